// ORACLE — TEST INFRASTRUCTURE ONLY.  CPU restatement of the reference's hot path
// (evoldoers/historian), operating on the same flattened descriptors as the
// product's C ABI (include/historian_b200.h) so tests can feed both sides the
// same bytes.  Plain FP64, single thread, the reference's own table-driven
// log_sum_exp and the reference's order of operations, so results are
// bit-comparable with oracle/_ref (the unmodified reference compiled here).
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may load
// this library.  Nothing under historian_b200/ links or imports it.
//
// Parity pinned: tests/test_oracle_golden.py checks this file against fixtures
// dumped from oracle/_ref (tests/golden/*.npz, made by tools/make_golden.py) and
// against the reference's own goldens (data/testforward.*.out etc.).
//
// Each function cites the reference file:line it follows.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <map>
#include <random>
#include <set>
#include <vector>

#include "historian_b200.h"

namespace {

const double NEG_INF = -std::numeric_limits<double>::infinity();

// ---------------------------------------------------------------- N1: logsumexp.h:22-64
const int LSE_MAX = 10;
const double LSE_PRECISION = .0001;
const int LSE_ENTRIES = ((int)(LSE_MAX / LSE_PRECISION)) + 1;

struct LseTable {
  std::vector<double> lookup;
  LseTable() : lookup(LSE_ENTRIES) {  // logsumexp.cpp:5-15
    int n;
    double x;
    for (n = 0; n < LSE_ENTRIES; n++) {
      x = n * LSE_PRECISION;
      lookup[n] = std::log(1. + std::exp(-x));  // log_sum_exp_unary_slow, logsumexp.cpp:53
    }
  }
};
const LseTable& lseTable() {
  static LseTable t;
  return t;
}

bool g_exact = false;  // switch to the reference's *_slow variants (logsumexp.cpp:22-53)

inline double lse_unary(double x) {  // logsumexp.h:42-64
  if (g_exact) return std::log(1. + std::exp(-x));
  if (x >= LSE_MAX || std::isnan(x) || std::isinf(x)) return 0;
  if (x < 0) return -x;
  const int n = (int)(x / LSE_PRECISION);
  const double* lookup = lseTable().lookup.data();
  const double f0 = lookup[n];
  const double dx = x - (n * LSE_PRECISION);
  const double f1 = lookup[n + 1];
  const double df = f1 - f0;
  return f0 + df * (dx / LSE_PRECISION);
}

inline double lse(double a, double b) {  // logsumexp.h:66-85
  double max, diff;
  if (a == b) { max = a; diff = 0; }
  else if (a < b) { max = b; diff = b - a; }
  else { max = a; diff = a - b; }
  if (g_exact) {  // log_sum_exp_slow, logsumexp.cpp:22-33: guards infinities the same way
    if (std::isinf(diff) || std::isnan(diff)) return max;
  }
  return max + lse_unary(diff);
}
inline double lse(double a, double b, double c) { return lse(lse(a, b), c); }
inline double lse(double a, double b, double c, double d) { return lse(lse(lse(a, b), c), d); }
inline double lse(double a, double b, double c, double d, double e) { return lse(lse(lse(lse(a, b), c), d), e); }
inline void lacc(double& a, double b) { a = lse(a, b); }

inline double logInner(const double* v1, const double* v2, int n) {  // logsumexp.h:132-137
  double lip = NEG_INF;
  for (int i = 0; i < n; ++i) lip = lse(lip, v1[i] + v2[i]);
  return lip;
}
inline double logInnerCA(const double* v1, const double* v2, int C, int A) {  // logsumexp.h:146-151
  double lip = NEG_INF;
  for (int c = 0; c < C; ++c) lip = lse(lip, logInner(v1 + c * A, v2 + c * A, A));
  return lip;
}

// MT19937 with an importable/exportable state image (624 words + position, the
// layout std::mt19937 streams out).  Same recurrence and tempering as
// std::mt19937, so libstdc++'s uniform_real_distribution driven by it consumes
// the stream exactly as the reference's generator does (forward.cpp:231-232).
struct Mt19937 {
  typedef uint32_t result_type;
  uint32_t mt[624];
  uint32_t p;
  static constexpr result_type min() { return 0; }
  static constexpr result_type max() { return 0xffffffffu; }
  void load(const uint32_t* st) { std::memcpy(mt, st, sizeof(mt)); p = st[624]; }
  void store(uint32_t* st) const { std::memcpy(st, mt, sizeof(mt)); st[624] = p; }
  void twist() {
    for (int k = 0; k < 624; ++k) {
      const uint32_t yv = (mt[k] & 0x80000000u) | (mt[(k + 1) % 624] & 0x7fffffffu);
      mt[k] = mt[(k + 397) % 624] ^ (yv >> 1) ^ ((yv & 1u) ? 0x9908b0dfu : 0u);
    }
    p = 0;
  }
  result_type operator()() {
    if (p >= 624) twist();
    uint32_t z = mt[p++];
    z ^= (z >> 11);
    z ^= (z << 7) & 0x9d2c5680u;
    z ^= (z << 15) & 0xefc60000u;
    z ^= (z >> 18);
    return z;
  }
};

// ---------------------------------------------------------------- T1: profile.h:13-76
struct Prof {
  int n, nt, A, C;
  std::vector<uint8_t> emit;
  std::vector<int> src, dest;
  std::vector<double> lp;
  std::vector<std::vector<int>> in, nullOut, absorbOut;
  const double* lpAbsorb;
  const int32_t* leafPos;
  bool empty;
  explicit Prof(const hist_profile_desc* d)
      : n(d->n_states), nt(d->n_trans), A(d->alph_size), C(d->components),
        emit(d->is_emit, d->is_emit + d->n_states),
        src(d->trans_src, d->trans_src + d->n_trans), dest(d->trans_dest, d->trans_dest + d->n_trans),
        lp(d->trans_lp, d->trans_lp + d->n_trans), in(d->n_states), nullOut(d->n_states),
        absorbOut(d->n_states), lpAbsorb(d->lp_absorb), leafPos(d->closest_leaf_pos) {
    for (int t = 0; t < nt; ++t) {
      in[dest[t]].push_back(t);
      (emit[dest[t]] ? absorbOut : nullOut)[src[t]].push_back(t);  // forward.cpp:828
    }
    empty = true;  // Profile::isEmpty, profile.cpp:356-361
    for (int i = 0; i < n; ++i)
      if (emit[i]) empty = false;
  }
  bool isNull(int i) const { return !emit[i]; }                       // profile.h:32
  bool isReady(int i) const { return nullOut[i].empty(); }             // profile.h:36
  bool isEmitOrStart(int i) const { return emit[i] || in[i].empty(); } // profile.h:34-35
  const double* absorb(int i) const { return lpAbsorb + (size_t)i * C * A; }
};

struct Hmm {
  int A, C;
  double w[6][6];  // w[src][dest], -inf where absent (pairhmm.cpp:46-115)
  const hist_pairhmm_desc* d;
  explicit Hmm(const hist_pairhmm_desc* d) : A(d->alph_size), C(d->components), d(d) {
    for (int a = 0; a < 6; ++a)
      for (int b = 0; b < 6; ++b) w[a][b] = NEG_INF;
    const double* l = d->lp_trans;
    w[0][0] = l[HIST_W_IMM_IMM]; w[0][1] = l[HIST_W_IMM_IMD]; w[0][2] = l[HIST_W_IMM_IDM];
    w[0][3] = l[HIST_W_IMM_IMI]; w[0][4] = l[HIST_W_IMM_IIW]; w[0][5] = l[HIST_W_IMM_EEE];
    w[1][0] = l[HIST_W_IMD_IMM]; w[1][1] = l[HIST_W_IMD_IMD]; w[1][2] = l[HIST_W_IMD_IDM]; w[1][5] = l[HIST_W_IMD_EEE];
    w[2][0] = l[HIST_W_IDM_IMM]; w[2][1] = l[HIST_W_IDM_IMD]; w[2][2] = l[HIST_W_IDM_IDM]; w[2][5] = l[HIST_W_IDM_EEE];
    w[3][0] = l[HIST_W_IMI_IMM]; w[3][1] = l[HIST_W_IMI_IMD]; w[3][3] = l[HIST_W_IMI_IMI];
    w[3][4] = l[HIST_W_IMI_IIW]; w[3][5] = l[HIST_W_IMI_EEE];
    w[4][0] = l[HIST_W_IIW_IMM]; w[4][2] = l[HIST_W_IIW_IDM]; w[4][4] = l[HIST_W_IIW_IIW]; w[4][5] = l[HIST_W_IIW_EEE];
  }
};

// PairHMM::sources (pairhmm.cpp:117-140)
const std::vector<int>& hmmSources(int dest) {
  static const std::vector<int> all = {0, 1, 2, 3, 4}, imd = {0, 1, 2, 3}, idm = {0, 1, 2, 4},
                                imi = {0, 3}, iiw = {0, 4, 3}, none;
  switch (dest) {
    case HIST_IMM: case HIST_EEE: return all;
    case HIST_IMD: return imd;
    case HIST_IDM: return idm;
    case HIST_IMI: return imi;
    case HIST_IIW: return iiw;
  }
  return none;
}

// ---------------------------------------------------------------- DPMatrix (forward.cpp:11-66, forward.h)
struct CellCoords {
  int xpos, ypos, state;
  bool operator<(const CellCoords& c) const {  // forward.h:36-37
    return xpos == c.xpos ? ypos == c.ypos ? state < c.state : ypos < c.ypos : xpos < c.xpos;
  }
  bool operator==(const CellCoords& c) const { return xpos == c.xpos && ypos == c.ypos && state == c.state; }
};

struct DP {
  Prof x, y;
  Hmm hmm;
  int xSize, ySize, A, C;
  std::vector<double> subx, suby, insx, insy, rootsubx, rootsuby;
  std::vector<int> xKey, yKey;
  std::vector<uint8_t> xNearStart, yNearEnd;
  int maxDistance;
  bool envInit;
  double* cells;  // dense [(xSize-1)][(ySize-1)][5]; absent cells read as -inf
  double lpEnd;

  DP(const hist_profile_desc* xd, const hist_profile_desc* yd, const hist_pairhmm_desc* hd,
     const hist_env_desc* env)
      : x(xd), y(yd), hmm(hd), xSize(xd->n_states), ySize(yd->n_states), A(hd->alph_size),
        C(hd->components), cells(nullptr), lpEnd(NEG_INF) {
    leftMultiply(x, hd->l_sub, subx);  // forward.cpp:20
    leftMultiply(y, hd->r_sub, suby);  // forward.cpp:21
    insx.assign(xSize, NEG_INF); insy.assign(ySize, NEG_INF);
    rootsubx.assign(xSize, NEG_INF); rootsuby.assign(ySize, NEG_INF);
    for (int i = 1; i < xSize - 1; ++i)  // forward.cpp:44-49
      if (!x.isNull(i))
        for (int cpt = 0; cpt < C; ++cpt) {
          lacc(insx[i], hd->l_log_cpt_weight[cpt] + logInner(hd->l_log_ins + cpt * A, x.absorb(i) + cpt * A, A));
          lacc(rootsubx[i], logInner(hd->log_root + cpt * A, &subx[((size_t)i * C + cpt) * A], A));
        }
    for (int j = 1; j < ySize - 1; ++j)  // forward.cpp:51-56
      if (!y.isNull(j))
        for (int cpt = 0; cpt < C; ++cpt) {
          lacc(insy[j], hd->r_log_cpt_weight[cpt] + logInner(hd->r_log_ins + cpt * A, y.absorb(j) + cpt * A, A));
          lacc(rootsuby[j], logInner(hd->log_root + cpt * A, &suby[((size_t)j * C + cpt) * A], A));
        }
    // envelope keys: forward.cpp:36-42 + alignpath.h:56-61
    maxDistance = env ? env->max_distance : -1;
    envInit = maxDistance >= 0;
    xKey.assign(xSize, 0); yKey.assign(ySize, 0);
    if (envInit) {
      for (int i = 0; i < xSize; ++i) {
        const int pos = i == 0 ? 0 : x.leafPos[i];
        xKey[i] = env->cumulative_matches[env->row1_pos_to_col[pos]];
      }
      for (int j = 0; j < ySize; ++j) {
        const int pos = j == 0 ? 0 : y.leafPos[j];
        yKey[j] = env->cumulative_matches[env->row2_pos_to_col[pos]];
      }
    }
    xNearStart.assign(xSize, 0); yNearEnd.assign(ySize, 0);
    xNearStart[0] = 1;  // forward.cpp:58-62
    for (int i = 0; i < xSize; ++i)
      if (xNearStart[i])
        for (int t : x.nullOut[i]) xNearStart[x.dest[t]] = 1;
    for (int yt : y.in[ySize - 1]) yNearEnd[y.src[yt]] = 1;  // forward.cpp:64-65
  }

  // Profile::leftMultiply, profile.cpp:78-91
  void leftMultiply(const Prof& p, const double* sub, std::vector<double>& out) {
    out.assign((size_t)p.n * C * A, NEG_INF);
    for (int i = 0; i < p.n; ++i)
      if (!p.isNull(i))
        for (int cpt = 0; cpt < C; ++cpt)
          for (int c = 0; c < A; ++c) {
            double lp = NEG_INF;
            for (int d = 0; d < A; ++d)
              lp = lse(lp, std::log(sub[((size_t)cpt * A + c) * A + d]) + p.absorb(i)[cpt * A + d]);
            out[((size_t)i * C + cpt) * A + c] = lp;
          }
  }

  bool atEdge(int i, int j) const { return xNearStart[i] || yNearEnd[j]; }  // forward.h:92-94
  bool inEnvelope(int i, int j) const {                                        // forward.h:96-98
    if (atEdge(i, j)) return true;
    if (!envInit) return true;
    const int d = xKey[i] - yKey[j];
    return std::abs(d) <= maxDistance;
  }

  size_t idx(int i, int j, int s) const { return ((size_t)i * (ySize - 1) + j) * 5 + s; }
  double cell(int i, int j, int s) const {  // forward.h:75-81: absent cells are -inf
    if (i < 0 || j < 0 || i >= xSize - 1 || j >= ySize - 1) return NEG_INF;
    return cells[idx(i, j, s)];
  }
  double cell(const CellCoords& c) const { return cell(c.xpos, c.ypos, c.state); }

  double absorb(int i, int j) const {  // computeLogProbAbsorb, forward.h:112-124
    std::vector<double> scratch((size_t)C * A);
    for (int cpt = 0; cpt < C; ++cpt)
      for (int n = 0; n < A; ++n)
        scratch[cpt * A + n] = subx[((size_t)i * C + cpt) * A + n] + suby[((size_t)j * C + cpt) * A + n];
    return logInnerCA(hmm.d->log_root, scratch.data(), C, A);
  }

  // ForwardMatrix::ForwardMatrix, forward.cpp:68-223
  void fillForward(double* out) {
    cells = out;
    const size_t total = (size_t)(xSize - 1) * (ySize - 1) * 5;
    for (size_t k = 0; k < total; ++k) cells[k] = NEG_INF;
    const double (*w)[6] = hmm.w;
    cells[idx(0, 0, HIST_IMM)] = 0;  // lpStart() = 0
    for (int i = 0; i < xSize - 1; ++i)
      for (int j = 0; j < ySize - 1; ++j)
        if (inEnvelope(i, j)) {
          double& imm = cells[idx(i, j, 0)];
          double& imd = cells[idx(i, j, 1)];
          double& idm = cells[idx(i, j, 2)];
          double& imi = cells[idx(i, j, 3)];
          double& iiw = cells[idx(i, j, 4)];
          const bool yReadyOrEmpty = y.isReady(j) || y.empty;
          const bool xReadyOrEmpty = x.isReady(i) || x.empty;
          if (!x.isNull(i)) {  // forward.cpp:95-118
            if (yReadyOrEmpty) {
              for (int xt : x.in[i]) {
                const int s = x.src[xt];
                lacc(imd, lse(cell(s, j, 0) + w[0][1], cell(s, j, 1) + w[1][1], cell(s, j, 2) + w[2][1],
                              cell(s, j, 3) + w[3][1]) + x.lp[xt]);
                lacc(iiw, lse(cell(s, j, 0) + w[0][4], cell(s, j, 3) + w[3][4], cell(s, j, 4) + w[4][4]) + x.lp[xt]);
              }
              imd += rootsubx[i];
              iiw += insx[i];
            }
          } else {  // forward.cpp:120-129
            if (yReadyOrEmpty)
              for (int xt : x.in[i]) {
                const int s = x.src[xt];
                lacc(imd, cell(s, j, 1) + x.lp[xt]);
                lacc(iiw, cell(s, j, 4) + x.lp[xt]);
              }
          }
          if (!y.isNull(j)) {  // forward.cpp:131-153
            if (xReadyOrEmpty) {
              for (int yt : y.in[j]) {
                const int s = y.src[yt];
                lacc(idm, lse(cell(i, s, 0) + w[0][2], cell(i, s, 1) + w[1][2], cell(i, s, 2) + w[2][2],
                              cell(i, s, 4) + w[4][2]) + y.lp[yt]);
                lacc(imi, lse(cell(i, s, 0) + w[0][3], cell(i, s, 3) + w[3][3]) + y.lp[yt]);
              }
              idm += rootsuby[j];
              imi += insy[j];
            }
          } else {  // forward.cpp:155-163
            for (int yt : y.in[j]) {
              const int s = y.src[yt];
              lacc(idm, cell(i, s, 2) + y.lp[yt]);
              lacc(imi, cell(i, s, 3) + y.lp[yt]);
            }
          }
          if (!x.isNull(i) && !y.isNull(j)) {  // forward.cpp:165-184
            for (int xt : x.in[i])
              for (int yt : y.in[j]) {
                const int sx = x.src[xt], sy = y.src[yt];
                lacc(imm, lse(cell(sx, sy, 0) + w[0][0], cell(sx, sy, 1) + w[1][0], cell(sx, sy, 2) + w[2][0],
                              cell(sx, sy, 3) + w[3][0], cell(sx, sy, 4) + w[4][0]) + x.lp[xt] + y.lp[yt]);
              }
            imm += absorb(i, j);
          } else if (y.isNull(j) && x.isEmitOrStart(i)) {  // forward.cpp:186-191
            for (int yt : y.in[j]) lacc(imm, cell(i, y.src[yt], 0) + y.lp[yt]);
          } else {  // forward.cpp:193-199
            if (yReadyOrEmpty)
              for (int xt : x.in[i]) lacc(imm, cell(x.src[xt], j, 0) + x.lp[xt]);
          }
        }
    lpEnd = NEG_INF;  // forward.cpp:205-220
    for (int xt : x.in[xSize - 1])
      for (int yt : y.in[ySize - 1]) {
        const int sx = x.src[xt], sy = y.src[yt];
        lacc(lpEnd, lse(cell(sx, sy, 0) + w[0][5], cell(sx, sy, 1) + w[1][5], cell(sx, sy, 2) + w[2][5],
                        cell(sx, sy, 3) + w[3][5], cell(sx, sy, 4) + w[4][5]) + x.lp[xt] + y.lp[yt]);
      }
  }

  // BackwardMatrix::BackwardMatrix, forward.cpp:975-1088
  void fillBackward(double* out) {
    cells = out;
    const size_t total = (size_t)(xSize - 1) * (ySize - 1) * 5;
    for (size_t k = 0; k < total; ++k) cells[k] = NEG_INF;
    const double (*w)[6] = hmm.w;
    for (int xt : x.in[xSize - 1])  // forward.cpp:981-995
      for (int yt : y.in[ySize - 1]) {
        const int sx = x.src[xt], sy = y.src[yt];
        if (inEnvelope(sx, sy))
          for (int s = 0; s < 5; ++s) cells[idx(sx, sy, s)] = x.lp[xt] + y.lp[yt] + w[s][5];
      }
    for (int i = xSize - 2; i >= 0; --i)
      for (int j = ySize - 2; j >= 0; --j)
        if (inEnvelope(i, j)) {
          double& imm = cells[idx(i, j, 0)];
          double& imd = cells[idx(i, j, 1)];
          double& idm = cells[idx(i, j, 2)];
          double& imi = cells[idx(i, j, 3)];
          double& iiw = cells[idx(i, j, 4)];
          const bool yReadyOrEmpty = y.isReady(j) || y.empty;
          const bool xReadyOrEmpty = x.isReady(i) || x.empty;
          for (int xt : x.absorbOut[i])  // forward.cpp:1017-1030
            for (int yt : y.absorbOut[j]) {
              const int dx = x.dest[xt], dy = y.dest[yt];
              const double dest_imm = x.lp[xt] + y.lp[yt] + absorb(dx, dy) + cell(dx, dy, 0);
              lacc(imm, w[0][0] + dest_imm);
              lacc(imd, w[1][0] + dest_imm);
              lacc(idm, w[2][0] + dest_imm);
              lacc(imi, w[3][0] + dest_imm);
              lacc(iiw, w[4][0] + dest_imm);
            }
          if (yReadyOrEmpty)  // forward.cpp:1032-1048
            for (int xt : x.absorbOut[i]) {
              const int dx = x.dest[xt];
              const double dest_imd = x.lp[xt] + rootsubx[dx] + cell(dx, j, 1);
              const double dest_iiw = x.lp[xt] + insx[dx] + cell(dx, j, 4);
              lacc(imm, w[0][1] + dest_imd);
              lacc(imd, w[1][1] + dest_imd);
              lacc(idm, w[2][1] + dest_imd);
              lacc(imi, w[3][1] + dest_imd);
              lacc(imm, w[0][4] + dest_iiw);
              lacc(imi, w[3][4] + dest_iiw);
              lacc(iiw, w[4][4] + dest_iiw);
            }
          if (xReadyOrEmpty)  // forward.cpp:1050-1065
            for (int yt : y.absorbOut[j]) {
              const int dy = y.dest[yt];
              const double dest_idm = y.lp[yt] + rootsuby[dy] + cell(i, dy, 2);
              const double dest_imi = y.lp[yt] + insy[dy] + cell(i, dy, 3);
              lacc(imm, w[0][2] + dest_idm);
              lacc(imd, w[1][2] + dest_idm);
              lacc(idm, w[2][2] + dest_idm);
              lacc(iiw, w[4][2] + dest_idm);
              lacc(imm, w[0][3] + dest_imi);
              lacc(imi, w[3][3] + dest_imi);
            }
          if (yReadyOrEmpty)  // forward.cpp:1067-1075
            for (int xt : x.nullOut[i]) {
              const int dx = x.dest[xt];
              lacc(imd, x.lp[xt] + cell(dx, j, 1));
              lacc(iiw, x.lp[xt] + cell(dx, j, 4));
              lacc(imm, x.lp[xt] + cell(dx, j, 0));
            }
          for (int yt : y.nullOut[j]) {  // forward.cpp:1077-1085
            const int dy = y.dest[yt];
            lacc(idm, y.lp[yt] + cell(i, dy, 2));
            lacc(imi, y.lp[yt] + cell(i, dy, 3));
            if (x.isEmitOrStart(i)) lacc(imm, y.lp[yt] + cell(i, dy, 0));
          }
        }
  }

  // lpCellEmitOrAbsorb, forward.cpp:404-440
  double lpCellEmitOrAbsorb(const CellCoords& c) const {
    double lp = 0;
    switch (c.state) {
      case HIST_IMD: if (!x.isNull(c.xpos)) lp = rootsubx[c.xpos]; break;
      case HIST_IIW: if (!x.isNull(c.xpos)) lp = insx[c.xpos]; break;
      case HIST_IDM: if (!y.isNull(c.ypos)) lp = rootsuby[c.ypos]; break;
      case HIST_IMI: if (!y.isNull(c.ypos)) lp = insy[c.ypos]; break;
      case HIST_IMM: if (!x.isNull(c.xpos) && !y.isNull(c.ypos)) lp = absorb(c.xpos, c.ypos); break;
      default: break;
    }
    return lp;
  }

  // sourceTransitionsWithoutEmitOrAbsorb, forward.cpp:326-398
  std::map<CellCoords, double> sourceTransNoEmit(const CellCoords& d) const {
    std::map<CellCoords, double> clp;
    const int xi = d.xpos, yj = d.ypos;
    const bool yReadyOrEmpty = y.isReady(yj) || y.empty;
    const bool xReadyOrEmpty = x.isReady(xi) || x.empty;
    switch (d.state) {
      case HIST_IMD: case HIST_IIW:
        if (x.isNull(xi)) {
          if (yReadyOrEmpty)
            if (xi < xSize - 1)
              for (int xt : x.in[xi]) clp[{x.src[xt], yj, d.state}] = x.lp[xt];
        } else if (yReadyOrEmpty)
          for (int xt : x.in[xi])
            for (int s : hmmSources(d.state)) clp[{x.src[xt], yj, s}] = hmm.w[s][d.state] + x.lp[xt];
        break;
      case HIST_IDM: case HIST_IMI:
        if (y.isNull(yj)) {
          if (yj < ySize - 1)
            for (int yt : y.in[yj]) clp[{xi, y.src[yt], d.state}] = y.lp[yt];
        } else if (xReadyOrEmpty)
          for (int yt : y.in[yj])
            for (int s : hmmSources(d.state)) clp[{xi, y.src[yt], s}] = hmm.w[s][d.state] + y.lp[yt];
        break;
      case HIST_IMM:
        if (y.isNull(yj) && x.isEmitOrStart(xi)) {
          if (yj < ySize - 1)
            for (int yt : y.in[yj]) clp[{xi, y.src[yt], d.state}] = y.lp[yt];
        } else if (x.isNull(xi)) {
          if (yReadyOrEmpty)
            if (xi < xSize - 1)
              for (int xt : x.in[xi]) clp[{x.src[xt], yj, d.state}] = x.lp[xt];
        } else if (!x.isNull(xi) && !y.isNull(yj))
          for (int xt : x.in[xi])
            for (int yt : y.in[yj])
              for (int s : hmmSources(d.state))
                clp[{x.src[xt], y.src[yt], s}] = hmm.w[s][d.state] + x.lp[xt] + y.lp[yt];
        break;
      case HIST_EEE:
        if (xi == xSize - 1 && yj == ySize - 1)
          for (int xt : x.in[xSize - 1])
            for (int yt : y.in[ySize - 1])
              for (int s : hmmSources(d.state))
                clp[{x.src[xt], y.src[yt], s}] = hmm.w[s][5] + x.lp[xt] + y.lp[yt];
        break;
      default: break;
    }
    return clp;
  }

  std::map<CellCoords, double> sourceTransitions(const CellCoords& d) const {  // forward.cpp:316-324
    auto clp = sourceTransNoEmit(d);
    const double lpAbs = lpCellEmitOrAbsorb(d);
    for (auto& kv : clp) kv.second += lpAbs;
    return clp;
  }
  std::map<CellCoords, double> sourceCells(const CellCoords& d) const {  // forward.cpp:309-314
    auto sc = sourceTransitions(d);
    for (auto& kv : sc) kv.second += cell(kv.first);
    return sc;
  }

  // BackwardMatrix::destTransitions, forward.cpp:1224-1285
  std::map<CellCoords, double> destTransitions(const CellCoords& sc) const {
    std::map<CellCoords, double> clp;
    const int xi = sc.xpos, yj = sc.ypos;
    const bool yReadyOrEmpty = y.isReady(yj) || y.empty;
    const bool xReadyOrEmpty = x.isReady(xi) || x.empty;
    for (int xt : x.absorbOut[xi])
      for (int yt : y.absorbOut[yj])
        clp[{x.dest[xt], y.dest[yt], HIST_IMM}] = hmm.w[sc.state][0] + x.lp[xt] + y.lp[yt];
    if (yReadyOrEmpty)
      for (int xt : x.absorbOut[xi]) {
        clp[{x.dest[xt], yj, HIST_IMD}] = hmm.w[sc.state][1] + x.lp[xt];
        clp[{x.dest[xt], yj, HIST_IIW}] = hmm.w[sc.state][4] + x.lp[xt];
      }
    if (xReadyOrEmpty)
      for (int yt : y.absorbOut[yj]) {
        clp[{xi, y.dest[yt], HIST_IDM}] = hmm.w[sc.state][2] + y.lp[yt];
        clp[{xi, y.dest[yt], HIST_IMI}] = hmm.w[sc.state][3] + y.lp[yt];
      }
    if (yReadyOrEmpty && (sc.state == HIST_IMD || sc.state == HIST_IIW || sc.state == HIST_IMM))
      for (int xt : x.nullOut[xi])
        if (x.dest[xt] != xSize - 1) clp[{x.dest[xt], yj, sc.state}] = x.lp[xt];
    if (sc.state == HIST_IDM || sc.state == HIST_IMI || (x.isEmitOrStart(xi) && sc.state == HIST_IMM))
      for (int yt : y.nullOut[yj])
        if (y.dest[yt] != ySize - 1) clp[{xi, y.dest[yt], sc.state}] = y.lp[yt];
    for (int xt : x.nullOut[xi])
      if (x.dest[xt] == xSize - 1)
        for (int yt : y.nullOut[yj])
          if (y.dest[yt] == ySize - 1)
            clp[{x.dest[xt], y.dest[yt], HIST_EEE}] = x.lp[xt] + y.lp[yt] + hmm.w[sc.state][5];
    for (auto& kv : clp) kv.second += lpCellEmitOrAbsorb(kv.first);
    return clp;
  }

  static CellCoords bestCell(const std::map<CellCoords, double>& clp) {  // forward.cpp:245-255
    CellCoords best{0, 0, HIST_EEE};
    double pBest = NEG_INF;
    for (auto& kv : clp)
      if (kv.second > pBest) { pBest = kv.second; best = kv.first; }
    return best;
  }

  // ForwardMatrix::bestTrace(end), forward.cpp:283-302
  std::vector<CellCoords> bestTrace(const CellCoords& end) const {
    std::vector<CellCoords> rev;
    rev.push_back(end);
    if (end.xpos > 0 || end.ypos > 0) {
      auto clp = sourceCells(end);
      while (true) {
        if (clp.empty()) break;
        const CellCoords cur = bestCell(clp);
        rev.push_back(cur);
        if (cur.xpos == 0 && cur.ypos == 0) break;
        clp = sourceCells(cur);
      }
    }
    return std::vector<CellCoords>(rev.rbegin(), rev.rend());
  }

  // sampleCell, forward.cpp:225-243 (libstdc++ uniform_real_distribution on mt19937)
  static bool sampleCell(const std::map<CellCoords, double>& clp, Mt19937& gen, CellCoords& out) {
    double ptot = 0, lpmax = NEG_INF;
    for (auto& kv : clp) lpmax = std::max(lpmax, kv.second);
    for (auto& kv : clp) ptot += std::exp(kv.second - lpmax);
    std::uniform_real_distribution<double> dist(0, ptot);
    const double p0 = dist(gen);
    double p = p0;
    for (auto& kv : clp)
      if ((p -= std::exp(kv.second - lpmax)) <= 0) { out = kv.first; return true; }
    return false;
  }

  // ForwardMatrix::sampleTrace, forward.cpp:257-276
  bool sampleTrace(Mt19937& gen, std::vector<CellCoords>& path) const {
    std::vector<CellCoords> rev;
    const CellCoords endCell{xSize - 1, ySize - 1, HIST_EEE};
    rev.push_back(endCell);
    auto clp = sourceCells(endCell);
    while (true) {
      CellCoords cur;
      if (!sampleCell(clp, gen, cur)) return false;
      rev.push_back(cur);
      if (cur.xpos == 0 && cur.ypos == 0) break;
      clp = sourceCells(cur);
    }
    path.assign(rev.rbegin(), rev.rend());
    return true;
  }

  bool isAbsorbing(const CellCoords& c) const {  // forward.cpp:470-474
    return (c.state == HIST_IMM && !x.isNull(c.xpos) && !y.isNull(c.ypos)) ||
           (c.state == HIST_IMD && !x.isNull(c.xpos)) || (c.state == HIST_IDM && !y.isNull(c.ypos));
  }
  double eliminatedLogProbInsert(const CellCoords& c) const {  // forward.cpp:503-526
    switch (c.state) {
      case HIST_IIW: return x.isNull(c.xpos) ? 0 : insx[c.xpos];
      case HIST_IMI: return y.isNull(c.ypos) ? 0 : insy[c.ypos];
      default: return 0;
    }
  }
};


}  // namespace

#include "gslshim/linalg.h"

extern "C" {

// gsl_cdf_gamma_Pinv as the GSL shim of oracle/_ref implements it (gamma.cpp:15 of the reference calls it for the
// discretised-gamma rate classes; no reference golden pins it, so tests/test_housekeeping.py checks it against scipy)
double oracle_gamma_pinv(double p, double a, double b) { return hb::linalg::gamma_pinv(p, a, b); }


void oracle_set_exact_lse(int exact) { g_exact = exact != 0; }

double oracle_log_sum_exp(double a, double b) { return lse(a, b); }

// DPMatrix constructor tables + envelope.  in_env: [(nx-1)*(ny-1)] bytes or NULL.
int oracle_pair_tables(const hist_profile_desc* x, const hist_profile_desc* y, const hist_pairhmm_desc* hmm,
                       const hist_env_desc* env, double* subx, double* suby, double* insx, double* insy,
                       double* rootsubx, double* rootsuby, int32_t* x_key, int32_t* y_key,
                       uint8_t* x_near_start, uint8_t* y_near_end, uint8_t* in_env, int64_t* n_cells) {
  DP dp(x, y, hmm, env);
  if (subx) std::copy(dp.subx.begin(), dp.subx.end(), subx);
  if (suby) std::copy(dp.suby.begin(), dp.suby.end(), suby);
  if (insx) std::copy(dp.insx.begin(), dp.insx.end(), insx);
  if (insy) std::copy(dp.insy.begin(), dp.insy.end(), insy);
  if (rootsubx) std::copy(dp.rootsubx.begin(), dp.rootsubx.end(), rootsubx);
  if (rootsuby) std::copy(dp.rootsuby.begin(), dp.rootsuby.end(), rootsuby);
  if (x_key) std::copy(dp.xKey.begin(), dp.xKey.end(), x_key);
  if (y_key) std::copy(dp.yKey.begin(), dp.yKey.end(), y_key);
  if (x_near_start) std::copy(dp.xNearStart.begin(), dp.xNearStart.end(), x_near_start);
  if (y_near_end) std::copy(dp.yNearEnd.begin(), dp.yNearEnd.end(), y_near_end);
  int64_t n = 0;
  for (int i = 0; i < dp.xSize - 1; ++i)
    for (int j = 0; j < dp.ySize - 1; ++j) {
      const bool e = dp.inEnvelope(i, j);
      if (in_env) in_env[(size_t)i * (dp.ySize - 1) + j] = e;
      n += e;
    }
  if (n_cells) *n_cells = n;
  return 0;
}

// Forward fill into a dense [(nx-1)][(ny-1)][5] array (-inf where no cell exists).
int oracle_forward(const hist_profile_desc* x, const hist_profile_desc* y, const hist_pairhmm_desc* hmm,
                   const hist_env_desc* env, double* cells, double* lp_end) {
  DP dp(x, y, hmm, env);
  dp.fillForward(cells);
  *lp_end = dp.lpEnd;
  return 0;
}

int oracle_backward(const hist_profile_desc* x, const hist_profile_desc* y, const hist_pairhmm_desc* hmm,
                    const hist_env_desc* env, double* cells, double* lp_start) {
  DP dp(x, y, hmm, env);
  dp.fillBackward(cells);
  *lp_start = dp.cell(0, 0, HIST_IMM);
  return 0;
}

// bestTrace from `end` through a filled forward matrix (dense cells as above).
int oracle_trace_best(const hist_profile_desc* x, const hist_profile_desc* y, const hist_pairhmm_desc* hmm,
                      const hist_env_desc* env, double* fwd_cells, hist_cell end, hist_cell* path,
                      int64_t capacity, int64_t* n) {
  DP dp(x, y, hmm, env);
  dp.cells = fwd_cells;
  const auto p = dp.bestTrace({end.xpos, end.ypos, end.state});
  *n = (int64_t)p.size();
  if ((int64_t)p.size() > capacity) return HIST_ERR_CAPACITY;
  for (size_t k = 0; k < p.size(); ++k) path[k] = {p[k].xpos, p[k].ypos, p[k].state};
  return 0;
}

int oracle_trace_sample(const hist_profile_desc* x, const hist_profile_desc* y, const hist_pairhmm_desc* hmm,
                        const hist_env_desc* env, double* fwd_cells, uint32_t* mt_state, int32_t n_traces,
                        hist_cell* paths, int64_t capacity, int64_t* path_off) {
  DP dp(x, y, hmm, env);
  dp.cells = fwd_cells;
  Mt19937 gen;
  gen.load(mt_state);
  int64_t off = 0;
  path_off[0] = 0;
  for (int t = 0; t < n_traces; ++t) {
    std::vector<CellCoords> p;
    if (!dp.sampleTrace(gen, p)) return HIST_ERR_NUMERIC;
    if (off + (int64_t)p.size() > capacity) return HIST_ERR_CAPACITY;
    for (size_t k = 0; k < p.size(); ++k) paths[off + k] = {p[k].xpos, p[k].ypos, p[k].state};
    off += (int64_t)p.size();
    path_off[t + 1] = off;
  }
  gen.store(mt_state);
  return 0;
}

// BackwardMatrix::bestTrace(start), forward.cpp:1287-1300, through dense backward cells.
int oracle_trace_best_forward(const hist_profile_desc* x, const hist_profile_desc* y,
                              const hist_pairhmm_desc* hmm, const hist_env_desc* env, double* bwd_cells,
                              hist_cell start, hist_cell* path, int64_t capacity, int64_t* n) {
  DP dp(x, y, hmm, env);
  dp.cells = bwd_cells;
  std::vector<CellCoords> p;
  CellCoords cur{start.xpos, start.ypos, start.state};
  while (cur.xpos < dp.xSize - 1 && cur.ypos < dp.ySize - 1) {
    auto clp = dp.destTransitions(cur);  // destCells, forward.cpp:1216-1222
    for (auto& kv : clp)
      if (kv.first.state != HIST_EEE) kv.second += dp.cell(kv.first);
    if (clp.empty()) break;
    cur = DP::bestCell(clp);
    p.push_back(cur);
  }
  p.push_back({dp.xSize - 1, dp.ySize - 1, HIST_EEE});
  *n = (int64_t)p.size();
  if ((int64_t)p.size() > capacity) return HIST_ERR_CAPACITY;
  for (size_t k = 0; k < p.size(); ++k) path[k] = {p[k].xpos, p[k].ypos, p[k].state};
  return 0;
}

// cellsAbovePostProbThreshold, forward.cpp:1302-1319.  Output in priority_queue pop order made
// deterministic: decreasing log posterior, ties by descending (xpos, ypos, state).
int oracle_posterior_cells(const hist_profile_desc* x, const hist_profile_desc* y, const hist_pairhmm_desc* hmm,
                           const hist_env_desc* env, double* fwd_cells, double* bwd_cells, double lp_end,
                           double min_post_prob, hist_cell* cells, double* log_post, int64_t capacity,
                           int64_t* n) {
  DP dp(x, y, hmm, env);
  const double thr = std::log(min_post_prob);
  std::vector<std::pair<double, CellCoords>> v;
  for (int i = dp.xSize - 2; i >= 0; --i)
    for (int j = dp.ySize - 2; j >= 0; --j)
      if (dp.inEnvelope(i, j))
        for (int s = 0; s < 5; ++s) {
          const double lpp = bwd_cells[dp.idx(i, j, s)] + fwd_cells[dp.idx(i, j, s)] - lp_end;
          if (lpp >= thr) v.push_back({lpp, {i, j, s}});
        }
  std::stable_sort(v.begin(), v.end(), [](const std::pair<double, CellCoords>& a, const std::pair<double, CellCoords>& b) {
    if (a.first != b.first) return a.first > b.first;
    return b.second < a.second;
  });
  *n = (int64_t)v.size();
  if ((int64_t)v.size() > capacity) return HIST_ERR_CAPACITY;
  for (size_t k = 0; k < v.size(); ++k) {
    cells[k] = {v[k].second.xpos, v[k].second.ypos, v[k].second.state};
    log_post[k] = v[k].first;
  }
  return 0;
}

// Indel part of ForwardMatrix::transitionEigenCounts (forward.cpp:579-652): c[6] = ins, del, insExt, delExt, insTime, delTime.
// x_tc / y_tc: [n_trans][6] counts of the children's transitions or NULL; waits = l.t, r.t, l.insWait, l.delWait, r.insWait, r.delWait.
static void indel_transition_counts(const DP& dp, const CellCoords& src, const CellCoords& dest, const double* x_tc, const double* y_tc,
                                    const double* waits, double* c) {
  enum { INS, DEL, INSEXT, DELEXT, INSTIME, DELTIME };
  const double lt = waits[0], rt = waits[1], lInsWait = waits[2], lDelWait = waits[3], rInsWait = waits[4], rDelWait = waits[5];
  for (int q = 0; q < 6; ++q) c[q] = 0;
  auto getTrans = [](const Prof& p, int s, int d) {
    for (int t : p.in[d])
      if (p.src[t] == s) return t;
    return -1;
  };
  if (src.xpos != dest.xpos && x_tc) {
    const int t = getTrans(dp.x, src.xpos, dest.xpos);
    if (t >= 0) for (int q = 0; q < 6; ++q) c[q] += x_tc[(size_t)t * 6 + q];
  }
  if (src.ypos != dest.ypos && y_tc) {
    const int t = getTrans(dp.y, src.ypos, dest.ypos);
    if (t >= 0) for (int q = 0; q < 6; ++q) c[q] += y_tc[(size_t)t * 6 + q];
  }
  if (dest.state == HIST_EEE) return;   // x.state[dest.xpos] is END: no case of the switch applies
  const bool xNull = dp.x.isNull(dest.xpos), yNull = dp.y.isNull(dest.ypos);
  const int s = dest.state;
  switch (s) {   // forward.cpp:586-650
    case HIST_IMM:
      if (!xNull && !yNull) {
        if (src.state == HIST_IMM || src.state == HIST_IMD) { c[INSTIME] += lt; c[DELTIME] += lt; }
        if (src.state == HIST_IMM || src.state == HIST_IDM) { c[INSTIME] += rt; c[DELTIME] += rt; }
      }
      break;
    case HIST_IMD:
      if (!xNull) {
        if (src.state == HIST_IMM || src.state == HIST_IMD) { c[INSTIME] += lt; c[DELTIME] += lt; }
        if (src.state == s) c[DELEXT] += 1;
        else { c[DEL] += 1; c[DELTIME] += rDelWait; }
      }
      break;
    case HIST_IIW:
      if (!xNull) {
        if (src.state == s) c[INSEXT] += 1;
        else { c[INS] += 1; c[INSTIME] += lInsWait; }
      }
      break;
    case HIST_IDM:
      if (!yNull) {
        if (src.state == HIST_IMM || src.state == HIST_IDM) { c[INSTIME] += rt; c[DELTIME] += rt; }
        if (src.state == s) c[DELEXT] += 1;
        else { c[DEL] += 1; c[DELTIME] += lDelWait; }
      }
      break;
    case HIST_IMI:
      if (!yNull) {
        if (src.state == s) c[INSEXT] += 1;
        else { c[INS] += 1; c[INSTIME] += rInsWait; }
      }
      break;
    default: break;
  }
}

// The indel counts of the effective transitions of makeProfile (forward.cpp:753-790): the elimination of
// oracle_profile_eliminate with an indel count vector per effective transition -- transitionEigenCounts for a direct
// transition, and for every path through an eliminated cell the reference's running posterior mixture
// (counts *= 1 - pp; counts += (src counts + cell->dest counts) * pp).  out [n_trans][6], transitions in the order
// oracle_profile_eliminate returns them (the first n_trans transitions of the profile after addReadyStates).
int oracle_eliminate_indel_counts(const hist_profile_desc* x, const hist_profile_desc* y, const hist_pairhmm_desc* hmm,
                                  const hist_env_desc* env, const hist_cell* cells_in, int64_t n_cells, int32_t strategy,
                                  const double* x_tc, const double* y_tc, const double* waits, double* out, int64_t capacity,
                                  int64_t* n_trans) {
  DP dp(x, y, hmm, env);
  std::set<CellCoords> cells;
  for (int64_t k = 0; k < n_cells; ++k) cells.insert({cells_in[k].xpos, cells_in[k].ypos, cells_in[k].state});
  const CellCoords startCell{0, 0, HIST_IMM}, endCell{dp.xSize - 1, dp.ySize - 1, HIST_EEE};
  if (!cells.count(startCell) || !cells.count(endCell)) return HIST_ERR_ARG;
  std::map<CellCoords, int> profStateIndex, outgoing;
  for (const auto& dest : cells)
    for (const auto& kv : dp.sourceTransitions(dest)) ++outgoing[kv.first];
  int n_states = 0;
  for (const auto& c : cells)
    if (dp.isAbsorbing(c) || c == startCell || c == endCell || outgoing[c] > 1 || (strategy & HIST_KEEP_GAPS_OPEN) != 0 ||
        (strategy & HIST_COLLAPSE_CHAINS) == 0)
      profStateIndex[c] = n_states++;
  struct Eff { double lpPath = NEG_INF; double cnt[6] = {0, 0, 0, 0, 0, 0}; };
  std::map<CellCoords, std::map<int, Eff>> effTrans;
  for (auto it = cells.crbegin(); it != cells.crend(); ++it) {
    const CellCoords& iterCell = *it;
    const auto slp = dp.sourceTransNoEmit(iterCell);
    const double ins = dp.eliminatedLogProbInsert(iterCell);
    auto found = profStateIndex.find(iterCell);
    if (found != profStateIndex.end()) {
      for (const auto& kv : slp) {
        Eff& eff = effTrans[kv.first][found->second];
        eff.lpPath = kv.second + ins;
        indel_transition_counts(dp, kv.first, iterCell, x_tc, y_tc, waits, eff.cnt);
      }
    } else {
      const auto cellEff = effTrans[iterCell];
      for (const auto& kv : slp) {
        double srcCell[6];
        indel_transition_counts(dp, kv.first, iterCell, x_tc, y_tc, waits, srcCell);
        auto& srcEff = effTrans[kv.first];
        for (const auto& ce : cellEff) {
          Eff& sd = srcEff[ce.first];
          const double lpPath = kv.second + ins + ce.second.lpPath;
          lacc(sd.lpPath, lpPath);
          const double pp = std::exp(lpPath - sd.lpPath);
          for (int q = 0; q < 6; ++q) sd.cnt[q] = sd.cnt[q] * (1 - pp) + (srcCell[q] + ce.second.cnt[q]) * pp;
        }
      }
    }
  }
  int64_t nt = 0;
  for (const auto& ps : profStateIndex)
    for (const auto& et : effTrans[ps.first]) {
      if (nt < capacity)
        for (int q = 0; q < 6; ++q) out[nt * 6 + q] = et.second.cnt[q];
      ++nt;
    }
  *n_trans = nt;
  return nt > capacity ? HIST_ERR_CAPACITY : 0;
}

// The indel half of BackwardMatrix::getCounts (forward.cpp:1183-1214) WITH counts carried by the children's transitions
// (forward.cpp:579-583): for every cell and every source transition, posterior weight x transitionEigenCounts' indel part
// (forward.cpp:579-652).  x_tc / y_tc: [n_trans][6] = ins, del, insExt, delExt, insTime, delTime of every transition of x / y
// (NULL: none); waits = hmm.l.t, hmm.r.t, l.insWait, l.delWait, r.insWait, r.delWait.  out[7]: the six sums and lp = lpEnd.
// The substitution half needs the reference's per-position cache of column counts (forward.cpp:903-917) and is not restated.
int oracle_fb_indel_counts(const hist_profile_desc* x, const hist_profile_desc* y, const hist_pairhmm_desc* hmm,
                           const hist_env_desc* env, const double* x_tc, const double* y_tc, const double* waits, double* out) {
  DP dp(x, y, hmm, env);
  const size_t n = (size_t)(dp.xSize - 1) * (dp.ySize - 1) * 5;
  std::vector<double> fwd(n), bwd(n);
  DP db(x, y, hmm, env);
  db.fillBackward(bwd.data());
  dp.fillForward(fwd.data());   // dp.cells = fwd: dp.cell() reads the Forward matrix
  const double lpEnd = dp.lpEnd;
  double tot[6] = {0, 0, 0, 0, 0, 0};
  for (int i = 0; i < dp.xSize - 1; ++i)
    for (int j = 0; j < dp.ySize - 1; ++j) {
      if (!dp.inEnvelope(i, j)) continue;
      for (int s = 0; s < 5; ++s) {
        const CellCoords dest{i, j, s};
        const double lpDest = bwd[dp.idx(i, j, s)];
        if (!(lpDest > NEG_INF)) continue;
        for (const auto& kv : dp.sourceTransitions(dest)) {
          const CellCoords& src = kv.first;
          const double w = std::exp(dp.cell(src) + kv.second + lpDest - lpEnd);
          if (!(w > 0)) continue;
          double c[6];
          indel_transition_counts(dp, src, dest, x_tc, y_tc, waits, c);
          for (int q = 0; q < 6; ++q) tot[q] += c[q] * w;
        }
      }
    }
  for (int q = 0; q < 6; ++q) out[q] = tot[q];
  out[6] = lpEnd;
  return 0;
}

// Numeric core of makeProfile, forward.cpp:686-843 (states, effective transitions).
int oracle_profile_eliminate(const hist_profile_desc* x, const hist_profile_desc* y, const hist_pairhmm_desc* hmm,
                             const hist_env_desc* env, double* fwd_cells, double lp_end, const hist_cell* cells_in,
                             int64_t n_cells, int32_t strategy, hist_elim_result* r) {
  DP dp(x, y, hmm, env);
  dp.cells = fwd_cells;
  const int C = dp.C, A = dp.A;
  std::set<CellCoords> cells;
  for (int64_t k = 0; k < n_cells; ++k) cells.insert({cells_in[k].xpos, cells_in[k].ypos, cells_in[k].state});
  const CellCoords startCell{0, 0, HIST_IMM}, endCell{dp.xSize - 1, dp.ySize - 1, HIST_EEE};
  if (!cells.count(startCell) || !cells.count(endCell)) return HIST_ERR_ARG;
  std::map<CellCoords, int> profStateIndex, outgoing;
  for (const auto& dest : cells)  // forward.cpp:699-701
    for (const auto& kv : dp.sourceTransitions(dest)) ++outgoing[kv.first];
  std::vector<CellCoords> states;
  for (const auto& c : cells)  // forward.cpp:703-732
    if (dp.isAbsorbing(c) || c == startCell || c == endCell || outgoing[c] > 1 ||
        (strategy & HIST_KEEP_GAPS_OPEN) != 0 || (strategy & HIST_COLLAPSE_CHAINS) == 0) {
      profStateIndex[c] = (int)states.size();
      states.push_back(c);
    }
  struct Eff { double lpPath = NEG_INF, lpBest = NEG_INF; std::vector<CellCoords> via; };
  std::map<CellCoords, std::map<int, Eff>> effTrans;
  for (auto it = cells.crbegin(); it != cells.crend(); ++it) {  // forward.cpp:751-807
    const CellCoords& iterCell = *it;
    const auto slp = dp.sourceTransNoEmit(iterCell);
    const double ins = dp.eliminatedLogProbInsert(iterCell);
    auto found = profStateIndex.find(iterCell);
    if (found != profStateIndex.end()) {
      const int cellIdx = found->second;
      for (const auto& kv : slp) {
        Eff& eff = effTrans[kv.first][cellIdx];
        eff.lpPath = eff.lpBest = kv.second + ins;
        eff.via.clear();
      }
    } else {
      const auto cellEff = effTrans[iterCell];  // copy: effTrans may grow below
      for (const auto& kv : slp) {
        auto& srcEff = effTrans[kv.first];
        for (const auto& ce : cellEff) {
          Eff& sd = srcEff[ce.first];
          const double lpPath = kv.second + ins + ce.second.lpPath;
          lacc(sd.lpPath, lpPath);
          const double lpBest = kv.second + ins + ce.second.lpBest;
          if (lpBest > sd.lpBest) {
            sd.lpBest = lpBest;
            sd.via.clear();
            sd.via.push_back(iterCell);
            sd.via.insert(sd.via.end(), ce.second.via.begin(), ce.second.via.end());
          }
        }
      }
    }
  }
  struct T { int s, d; double lp, best; std::vector<CellCoords> via; };
  std::vector<T> trans;
  for (const auto& ps : profStateIndex)  // forward.cpp:810-831
    for (const auto& et : effTrans[ps.first]) trans.push_back({ps.second, et.first, et.second.lpPath, et.second.lpBest, et.second.via});
  int64_t nvia = 0;
  for (auto& t : trans) nvia += (int64_t)t.via.size();
  r->n_states = (int)states.size();
  r->n_trans = (int)trans.size();
  r->n_via = nvia;
  if (r->n_states > r->state_capacity || r->n_trans > r->trans_capacity || nvia > r->via_capacity) return HIST_ERR_CAPACITY;
  for (size_t k = 0; k < states.size(); ++k) {
    const CellCoords& c = states[k];
    r->state_cell[k] = {c.xpos, c.ypos, c.state};
    double* la = r->state_lp_absorb + k * C * A;
    for (int q = 0; q < C * A; ++q) la[q] = NEG_INF;
    if (dp.isAbsorbing(c)) {  // forward.cpp:713-727
      for (int q = 0; q < C * A; ++q) {
        if (c.state == HIST_IMM) la[q] = dp.subx[(size_t)c.xpos * C * A + q] + dp.suby[(size_t)c.ypos * C * A + q];
        else if (c.state == HIST_IMD) la[q] = dp.subx[(size_t)c.xpos * C * A + q];
        else la[q] = dp.suby[(size_t)c.ypos * C * A + q];
      }
    }
    r->state_fwd_lp[k] = c.state == HIST_EEE ? lp_end : dp.cell(c);  // forward.cpp:731
  }
  int64_t off = 0;
  for (size_t k = 0; k < trans.size(); ++k) {
    r->trans_src[k] = trans[k].s;
    r->trans_dest[k] = trans[k].d;
    r->trans_lp[k] = trans[k].lp;
    r->trans_best_lp[k] = trans[k].best;
    r->trans_best_via_off[k] = off;
    for (auto& c : trans[k].via) r->trans_best_via[off++] = {c.xpos, c.ypos, c.state};
  }
  r->trans_best_via_off[trans.size()] = off;
  return 0;
}

// Profile::calcSumPathAbsorbProbs, profile.cpp:112-131; cum: [n_states] per-state cumLogProb.
double oracle_sum_path_absorb(const hist_profile_desc* p, const double* log_cpt_weight, const double* log_ins_prob,
                              double* cum) {
  Prof prof(p);
  std::vector<double> lpCum(prof.n, NEG_INF);
  lpCum[0] = 0;
  for (int pos = 1; pos < prof.n; ++pos) {
    double lpAbs = 0;
    if (!prof.isNull(pos)) {
      lpAbs = NEG_INF;
      for (int cpt = 0; cpt < prof.C; ++cpt)
        lacc(lpAbs, log_cpt_weight[cpt] + logInner(log_ins_prob + cpt * prof.A, prof.absorb(pos) + cpt * prof.A, prof.A));
    }
    for (int ti : prof.in[pos]) lacc(lpCum[pos], lpCum[prof.src[ti]] + prof.lp[ti] + lpAbs);
  }
  if (cum) std::copy(lpCum.begin(), lpCum.end(), cum);
  return lpCum.back();
}

}  // extern "C"
