// historian_b200: the DP graph of a pair (device side), shared by traceback, posterior decoding and
// the makeProfile elimination.
//
// Mirrors ForwardMatrix::sourceTransitionsWithoutEmitOrAbsorb / sourceTransitions / sourceCells
// (reference forward.cpp:309-398), lpCellEmitOrAbsorb (forward.cpp:404-440),
// BackwardMatrix::destTransitions (forward.cpp:1224-1285) and PairHMM::sources
// (pairhmm.cpp:117-140) as allocation-free enumerators: `count` + `get(c)` instead of a std::map.
// Values are natural-log doubles, like the reference's.
//
// Ordering contract: the reference iterates candidates in CellCoords order (xpos, ypos, state;
// forward.h:36-37).  source_get(c) enumerates in exactly that order PROVIDED every in-list is
// ascending in source state with no repeated source; hist_profile_upload checks this
// (hist_profile::sorted_in) and the trace entry points refuse profiles that violate it.
#pragma once
#include "hist_internal.cuh"

namespace hb {

struct DCell {
  int x, y, s;
};

__device__ __forceinline__ bool dcell_less(const DCell& a, const DCell& b) {
  return a.x == b.x ? (a.y == b.y ? a.s < b.s : a.y < b.y) : a.x < b.x;
}

__device__ __forceinline__ double lse2(double a, double b) {
  if (a == -INFINITY) return b;
  if (b == -INFINITY) return a;
  const double m = fmax(a, b);
  return m + log1p(exp(-fabs(a - b)));
}

__device__ __forceinline__ int dp_cell_slot(const PairDev& P, const DiagInfo& di, int i, int j) {
  const int r = i - di.ilo;
  if (r >= 0 && r < di.width) return P.nS + r;
  int s = P.x.slot[i];
  if (s >= 0) return s;
  s = P.y.slot[j];
  if (s >= 0) return P.nS + di.width + s;
  return -1;
}

// DPMatrix::cell (forward.h:75-81): natural log of state s of cell (i,j); -inf for absent cells.
// FAST: ln of the stored float by cell_log_fast (2e-10): for the traceback, which only compares and samples these values.
template <bool FAST = false>
__device__ __forceinline__ double dp_cell(const PairDev& P, bool backward, int i, int j, int s) {
  if (i < 0 || j < 0 || i > P.x.n - 2 || j > P.y.n - 2 || s < 0 || s > 4) return -INFINITY;
  const DiagInfo di = P.dinfo[i + j];
  const int q = dp_cell_slot(P, di, i, j);
  if (q < 0) return -INFINITY;
  const float4 a = (backward ? P.bA : P.fA)[di.off + q];
  const float2 b = (backward ? P.bB : P.fB)[di.off + q];
  const int E = __float_as_int(b.y);
  if (E <= HB_EXP_ZERO / 2) return -INFINITY;
  const float v = s == 0 ? a.x : s == 1 ? a.y : s == 2 ? a.z : s == 3 ? a.w : b.x;
  return FAST ? cell_log_fast(v, E) : cell_log(v, E);
}

__device__ __forceinline__ bool dp_in_envelope(const PairDev& P, int i, int j) {
  return P.x.near_[i] || P.y.near_[j] || P.max_dist < 0 || hb_iabs(P.x.key[i] - P.y.key[j]) <= P.max_dist;
}

// computeLogProbAbsorb (forward.h:112-124), serial exact log-sum-exp (one thread).
__device__ inline double dp_absorb(const PairDev& P, int i, int j) {
  const double* sx = P.x.sub + (size_t)i * P.K;
  const double* sy = P.y.sub + (size_t)j * P.K;
  double m = -INFINITY;
  for (int k = 0; k < P.K; ++k) m = fmax(m, P.log_root[k] + sx[k] + sy[k]);
  if (m == -INFINITY) return -INFINITY;
  double acc = 0;
  for (int k = 0; k < P.K; ++k) acc += exp(P.log_root[k] + sx[k] + sy[k] - m);
  return m + log(acc);
}

// lpCellEmitOrAbsorb (forward.cpp:404-440)
__device__ inline double dp_emit(const PairDev& P, const DCell& c) {
  const bool xe = c.x < P.x.n && (P.x.flags[c.x] & F_EMIT), ye = c.y < P.y.n && (P.y.flags[c.y] & F_EMIT);
  switch (c.s) {
    case HIST_IMD: return xe ? P.x.rootsub[c.x] : 0.0;
    case HIST_IIW: return xe ? P.x.insl[c.x] : 0.0;
    case HIST_IDM: return ye ? P.y.rootsub[c.y] : 0.0;
    case HIST_IMI: return ye ? P.y.insl[c.y] : 0.0;
    case HIST_IMM: return xe && ye ? dp_absorb(P, c.x, c.y) : 0.0;
    default: return 0.0;
  }
}

// PairHMM::sources (pairhmm.cpp:117-140), ascending in state
__device__ __forceinline__ int hmm_nsources(int dest) {
  return dest == HIST_IMM || dest == HIST_EEE ? 5 : dest == HIST_IMD || dest == HIST_IDM ? 4 : dest == HIST_IMI ? 2 : 3;
}
__device__ __forceinline__ int hmm_source(int dest, int k) {
  switch (dest) {
    case HIST_IMD: return k;                       // IMM IMD IDM IMI
    case HIST_IDM: return k < 3 ? k : 4;           // IMM IMD IDM IIW
    case HIST_IMI: return k == 0 ? 0 : 3;          // IMM IMI
    case HIST_IIW: return k == 0 ? 0 : k == 1 ? 3 : 4;  // IMM IMI IIW
    default: return k;
  }
}

// How the sources of a destination cell are laid out (sourceTransitionsWithoutEmitOrAbsorb, forward.cpp:326-398)
struct SrcEnum {
  int n;       // candidates
  int mode;    // 0: none, 1: x-list same state, 2: x-list x hmm sources, 3: y-list same state, 4: y-list x hmm sources,
               // 5: x-list x y-list x 5 states
  int xb, xn, yb, yn, ns;
};

__device__ inline SrcEnum source_enum(const PairDev& P, const DCell& d, double (*w_log)[6]) {
  (void)w_log;
  SrcEnum e;
  e.n = 0; e.mode = 0; e.ns = 1;
  const int xi = d.x, yj = d.y;
  const unsigned fx = P.x.flags[xi], fy = P.y.flags[yj];
  const bool xNull = !(fx & F_EMIT), yNull = !(fy & F_EMIT);
  const bool xReady = fx & F_READY, yReady = fy & F_READY;
  const bool xEmitOrStart = fx & F_EMIT_OR_START;
  e.xb = P.x.in_off[xi]; e.xn = P.x.in_off[xi + 1] - e.xb;
  e.yb = P.y.in_off[yj]; e.yn = P.y.in_off[yj + 1] - e.yb;
  const int xEnd = P.x.n - 1, yEnd = P.y.n - 1;
  switch (d.s) {
    case HIST_IMD: case HIST_IIW:
      if (xNull) {
        if (yReady && xi < xEnd) { e.mode = 1; e.n = e.xn; }
      } else if (yReady) { e.mode = 2; e.ns = hmm_nsources(d.s); e.n = e.xn * e.ns; }
      break;
    case HIST_IDM: case HIST_IMI:
      if (yNull) {
        if (yj < yEnd) { e.mode = 3; e.n = e.yn; }
      } else if (xReady) { e.mode = 4; e.ns = hmm_nsources(d.s); e.n = e.yn * e.ns; }
      break;
    case HIST_IMM:
      if (yNull && xEmitOrStart) {
        if (yj < yEnd) { e.mode = 3; e.n = e.yn; }
      } else if (xNull) {
        if (yReady && xi < xEnd) { e.mode = 1; e.n = e.xn; }
      } else if (!yNull) { e.mode = 5; e.ns = 5; e.n = e.xn * e.yn * 5; }
      break;
    case HIST_EEE:
      if (xi == xEnd && yj == yEnd) { e.mode = 5; e.ns = 5; e.n = e.xn * e.yn * 5; }
      break;
    default: break;
  }
  return e;
}

// c-th source (in CellCoords order) of destination d with its transition log-weight (no emission)
__device__ inline void source_get(const PairDev& P, const double (*w)[6], const DCell& d, const SrcEnum& e, int c, DCell& src, double& lp) {
  switch (e.mode) {
    case 1: {
      const int t = P.x.in_idx[e.xb + c];
      src = {P.x.t_src[t], d.y, d.s};
      lp = P.x.t_lp[t];
    } break;
    case 2: {
      const int t = P.x.in_idx[e.xb + c / e.ns];
      const int s = hmm_source(d.s, c % e.ns);
      src = {P.x.t_src[t], d.y, s};
      lp = w[s][d.s] + P.x.t_lp[t];
    } break;
    case 3: {
      const int t = P.y.in_idx[e.yb + c];
      src = {d.x, P.y.t_src[t], d.s};
      lp = P.y.t_lp[t];
    } break;
    case 4: {
      const int t = P.y.in_idx[e.yb + c / e.ns];
      const int s = hmm_source(d.s, c % e.ns);
      src = {d.x, P.y.t_src[t], s};
      lp = w[s][d.s] + P.y.t_lp[t];
    } break;
    default: {
      const int tx = P.x.in_idx[e.xb + c / (e.yn * 5)];
      const int ty = P.y.in_idx[e.yb + (c / 5) % e.yn];
      const int s = c % 5;
      src = {P.x.t_src[tx], P.y.t_src[ty], s};
      lp = w[s][d.s == HIST_EEE ? 5 : 0] + P.x.t_lp[tx] + P.y.t_lp[ty];
    } break;
  }
}

}  // namespace hb
