// historian_b200 INTEGRATION: the reference-side binding, compiled.
//
// This translation unit is what a maintainer of evoldoers/historian adds to bind the Forward/Backward hot path to
// libhistorian_b200.so: it DEFINES the three constructors of the reference's own classes -- DPMatrix::DPMatrix
// (forward.cpp:11-66), ForwardMatrix::ForwardMatrix (forward.cpp:68-223) and BackwardMatrix::BackwardMatrix
// (forward.cpp:975-1097) -- on top of the C ABI in include/historian_b200.h, against the reference's UNMODIFIED headers.
// Everything else of those classes (tracebacks, makeProfile, posteriors, getCounts, write ...) is the reference's own
// code working on the matrices the device filled.  It is linked AHEAD of the reference's forward.o with
// -Wl,--allow-multiple-definition (see integration/Makefile), so the reference's t/*.cpp test mains and its `historian`
// main build unchanged and run their golden tests through the GPU path (tests/test_integration_gpu.py).
//
// A production binding would keep the hist_pair alive inside the matrix object and bind the tracebacks and makeProfile
// as well (hist_recon_level does that for whole guide-tree levels); this one copies the filled cells back into the
// reference's cellStorage so that every reference method keeps working unchanged.
#include <cstdio>
#include <cstdlib>
#include <limits>
#include <vector>

#include "forward.h"
#include "util.h"
#include "logger.h"

#include "historian_b200.h"

namespace {

hist_ctx* b200_ctx() {
  static hist_ctx* ctx = nullptr;
  if (!ctx) {
    const char* dev = getenv("HIST_B200_DEVICE");
    if (hist_ctx_create(dev ? atoi(dev) : 0, &ctx) != HIST_OK) Abort("historian_b200: %s", hist_last_error(nullptr));
  }
  return ctx;
}

void check(int rc, const char* what) {
  if (rc != HIST_OK) Abort("historian_b200: %s failed: %s", what, hist_last_error(b200_ctx()));
}

// host arrays behind a hist_profile_desc
struct FlatProfile {
  std::vector<uint8_t> is_emit;
  std::vector<int32_t> src, dest, leaf;
  std::vector<double> lp, absorb;
  hist_profile_desc d;
  FlatProfile(const Profile& p, const GuideAlignmentEnvelope& env, AlignRowIndex row) {
    const int n = (int)p.size(), C = (int)p.components, A = (int)p.alphSize;
    is_emit.resize(n);
    absorb.assign((size_t)n * C * A, -std::numeric_limits<double>::infinity());
    for (int i = 0; i < n; ++i) {
      is_emit[i] = !p.state[i].isNull();
      if (is_emit[i])
        for (int c = 0; c < C; ++c)
          for (int a = 0; a < A; ++a) absorb[((size_t)i * C + c) * A + a] = p.state[i].lpAbsorb[c][a];
    }
    for (const auto& t : p.trans) {
      src.push_back((int32_t)t.src);
      dest.push_back((int32_t)t.dest);
      lp.push_back(t.lpTrans);
    }
    if (env.initialized()) {
      leaf.assign(n, 0);
      for (int i = 1; i < n; ++i) leaf[i] = (int32_t)p.state[i].seqCoords.at(row);   // forward.cpp:36-42
    }
    d.n_states = n;
    d.n_trans = (int32_t)src.size();
    d.alph_size = A;
    d.components = C;
    d.is_emit = is_emit.data();
    d.trans_src = src.data();
    d.trans_dest = dest.data();
    d.trans_lp = lp.data();
    d.lp_absorb = absorb.data();
    d.closest_leaf_pos = leaf.empty() ? nullptr : leaf.data();
  }
};

struct FlatHmm {
  std::vector<double> root, lsub, rsub, lins, rins, lcw, rcw;
  hist_pairhmm_desc d;
  explicit FlatHmm(const PairHMM& h) {
    const int C = h.components(), A = (int)h.alphabetSize();
    const double w[HIST_NWEIGHTS] = {h.imm_imm, h.imm_imd, h.imm_idm, h.imm_imi, h.imm_iiw, h.imm_eee, h.imd_imm, h.imd_imd,
                                     h.imd_idm, h.imd_eee, h.idm_imm, h.idm_imd, h.idm_idm, h.idm_eee, h.imi_imm, h.imi_imd,
                                     h.imi_imi, h.imi_iiw, h.imi_eee, h.iiw_imm, h.iiw_idm, h.iiw_iiw, h.iiw_eee};
    for (int k = 0; k < HIST_NWEIGHTS; ++k) d.lp_trans[k] = w[k];
    for (int c = 0; c < C; ++c)
      for (int a = 0; a < A; ++a) {
        root.push_back(h.logRoot[c][a]);
        lins.push_back(h.logl.logInsProb[c][a]);
        rins.push_back(h.logr.logInsProb[c][a]);
        for (int b = 0; b < A; ++b) {
          lsub.push_back(gsl_matrix_get(h.l.subMat[c], a, b));
          rsub.push_back(gsl_matrix_get(h.r.subMat[c], a, b));
        }
      }
    lcw.assign(h.logl.logCptWeight.begin(), h.logl.logCptWeight.end());
    rcw.assign(h.logr.logCptWeight.begin(), h.logr.logCptWeight.end());
    d.alph_size = A;
    d.components = C;
    d.log_root = root.data();
    d.l_sub = lsub.data();
    d.r_sub = rsub.data();
    d.l_log_ins = lins.data();
    d.r_log_ins = rins.data();
    d.l_log_cpt_weight = lcw.data();
    d.r_log_cpt_weight = rcw.data();
    d.l_t = h.l.t; d.r_t = h.r.t;
    d.l_ins_wait = h.l.insWait; d.l_del_wait = h.l.delWait;
    d.r_ins_wait = h.r.insWait; d.r_del_wait = h.r.delWait;
  }
};

struct FlatEnv {
  std::vector<int32_t> cum, r1, r2;
  hist_env_desc d;
  explicit FlatEnv(const GuideAlignmentEnvelope& e) {
    d.max_distance = e.initialized() ? e.maxDistance : -1;
    cum.assign(e.cumulativeMatches.begin(), e.cumulativeMatches.end());
    r1.assign(e.row1PosToCol.begin(), e.row1PosToCol.end());
    r2.assign(e.row2PosToCol.begin(), e.row2PosToCol.end());
    d.n_cols = cum.empty() ? 0 : (int32_t)cum.size() - 1;
    d.cumulative_matches = cum.data();
    d.row1_len = r1.empty() ? 0 : (int32_t)r1.size() - 1;
    d.row2_len = r2.empty() ? 0 : (int32_t)r2.size() - 1;
    d.row1_pos_to_col = r1.data();
    d.row2_pos_to_col = r2.data();
  }
};

// The device objects of the DPMatrix under construction: created while its first computed member (subx) is
// initialised, consumed by the constructor bodies below, released when the derived constructor is done.
struct Build {
  hist_profile* dx = nullptr;
  hist_profile* dy = nullptr;
  hist_pair* pair = nullptr;
  std::vector<double> subx, suby, insx, insy, rootsubx, rootsuby;
  std::vector<uint8_t> near_start, near_end;
  void release() {
    if (pair) hist_pair_destroy(pair);
    if (dx) hist_profile_destroy(dx);
    if (dy) hist_profile_destroy(dy);
    pair = nullptr; dx = dy = nullptr;
  }
};
thread_local Build g_build;

Profile with_absorb(const Profile& p, const std::vector<double>& sub) {   // what Profile::leftMultiply returns (profile.cpp:78-91)
  Profile q(p);
  const size_t C = p.components, A = p.alphSize;
  for (size_t i = 0; i < q.state.size(); ++i)
    if (!q.state[i].isNull())
      for (size_t c = 0; c < C; ++c)
        for (size_t a = 0; a < A; ++a) q.state[i].lpAbsorb[c][a] = sub[(i * C + c) * A + a];
  return q;
}

// Runs the DPMatrix constructor on the device.  Called from the initialiser of cellStorage, the FIRST member of DPMatrix
// in declaration order (members are initialised in that order, whatever the order of the initialiser list): every other
// member below then reads the finished tables.  Returns the size cellStorage is constructed with.
size_t prepare(const Profile& x, const Profile& y, const PairHMM& hmm, const GuideAlignmentEnvelope& env) {
  hist_ctx* ctx = b200_ctx();
  Build& b = g_build;
  b.release();
  FlatProfile fx(x, env, env.initialized() ? env.row1 : 0), fy(y, env, env.initialized() ? env.row2 : 0);
  FlatHmm fh(hmm);
  FlatEnv fe(env);
  check(hist_profile_upload(ctx, &fx.d, &b.dx), "hist_profile_upload(x)");
  check(hist_profile_upload(ctx, &fy.d, &b.dy), "hist_profile_upload(y)");
  check(hist_pair_prepare(ctx, b.dx, b.dy, &fh.d, &fe.d, &b.pair), "hist_pair_prepare");
  const size_t nx = x.size(), ny = y.size(), K = x.components * x.alphSize;
  b.subx.resize(nx * K); b.suby.resize(ny * K);
  b.insx.resize(nx); b.insy.resize(ny); b.rootsubx.resize(nx); b.rootsuby.resize(ny);
  check(hist_pair_fetch_tables(b.pair, b.subx.data(), b.suby.data(), b.insx.data(), b.insy.data(), b.rootsubx.data(), b.rootsuby.data()),
        "hist_pair_fetch_tables");
  b.near_start.resize(nx); b.near_end.resize(ny);
  check(hist_pair_fetch_envelope(b.pair, nullptr, nullptr, b.near_start.data(), b.near_end.data(), nullptr), "hist_pair_fetch_envelope");
  return x.size();
}

}  // namespace

// ------------------------------------------------------------------ DPMatrix::DPMatrix (forward.cpp:11-66)
DPMatrix::DPMatrix (const Profile& x, const Profile& y, const PairHMM& hmm, const GuideAlignmentEnvelope& env)
  : x(x),
    y(y),
    hmm(hmm),
    alphSize ((AlphTok) hmm.alphabetSize()),
    xEmpty (x.isEmpty()),
    yEmpty (y.isEmpty()),
    xSize (x.size()),
    ySize (y.size()),
    subx (with_absorb (x, g_build.subx)),       // Profile::leftMultiply of both profiles, computed on the device
    suby (with_absorb (y, g_build.suby)),
    cellStorage (prepare (x, y, hmm, env)),      // initialised first (declaration order): runs the constructor kernels
    absorbScratch (hmm.components(), vguard<LogProb> (hmm.alphabetSize())),
    insx (g_build.insx.begin(), g_build.insx.end()),
    insy (g_build.insy.begin(), g_build.insy.end()),
    rootsubx (g_build.rootsubx.begin(), g_build.rootsubx.end()),
    rootsuby (g_build.rootsuby.begin(), g_build.rootsuby.end()),
    startCell (0, 0, PairHMM::SSS),
    endCell (xSize - 1, ySize - 1, PairHMM::EEE),
    envelope (env),
    xClosestLeafPos (xSize, 0),
    yClosestLeafPos (ySize, 0),
    xNearStart (xSize, false),
    yNearEnd (ySize, false)
{
  if (env.initialized()) {
    for (ProfileStateIndex i = 1; i < xSize; ++i)
      xClosestLeafPos[i] = x.state[i].seqCoords.at(env.row1);
    for (ProfileStateIndex j = 1; j < ySize; ++j)
      yClosestLeafPos[j] = y.state[j].seqCoords.at(env.row2);
  }
  for (ProfileStateIndex i = 0; i < xSize; ++i) xNearStart[i] = g_build.near_start[i] != 0;
  for (ProfileStateIndex j = 0; j < ySize; ++j) yNearEnd[j] = g_build.near_end[j] != 0;
}

namespace {
// DPMatrix::cell of every in-envelope cell from the device, into the reference's sparse storage
template <class Store>
void fetch_cells(int which, Store& store) {
  hist_pair_info info;
  check(hist_pair_get_info(g_build.pair, &info), "hist_pair_get_info");
  const size_t n = (size_t)info.n_cells;
  std::vector<double> lp(5 * n);
  std::vector<int32_t> xi(n), yj(n);
  int64_t got = 0;
  check(hist_cells_fetch(g_build.pair, which, lp.data(), xi.data(), yj.data(), (int64_t)n, &got), "hist_cells_fetch");
  for (size_t k = 0; k < (size_t)got; ++k) {
    auto& cell = store[xi[k]][yj[k]];
    for (int s = 0; s < 5; ++s) cell.lp[s] = lp[5 * k + s];
  }
}
}  // namespace

// ------------------------------------------------------------------ ForwardMatrix::ForwardMatrix (forward.cpp:68-223)
ForwardMatrix::ForwardMatrix (const Profile& x, const Profile& y, const PairHMM& hmm, AlignRowIndex parentRowIndex, const GuideAlignmentEnvelope& env, SumProduct* sumProd)
  : DPMatrix (x, y, hmm, env),
    parentRowIndex (parentRowIndex),
    sumProd (sumProd)
{
  double lp = 0;
  check(hist_fwd_fill(g_build.pair, &lp), "hist_fwd_fill");
  lpEnd = lp;
  fetch_cells(0, cellStorage);
  g_build.release();
  LogThisAt(6,"Forward log-likelihood is " << lpEnd << endl);
}

// ------------------------------------------------------------------ BackwardMatrix::BackwardMatrix (forward.cpp:975-1097)
BackwardMatrix::BackwardMatrix (ForwardMatrix& fwd)
  : DPMatrix (fwd.x, fwd.y, fwd.hmm, fwd.envelope),
    fwd (fwd)
{
  lpEnd = 0;
  double lpe = 0, lps = 0;
  check(hist_fwd_fill(g_build.pair, &lpe), "hist_fwd_fill");      // BackwardMatrix takes a filled ForwardMatrix
  check(hist_bwd_fill(g_build.pair, &lps), "hist_bwd_fill");
  fetch_cells(1, cellStorage);
  g_build.release();
  LogThisAt(6,"Backward log-likelihood is " << lpStart() << endl);
  if (gsl_fcmp (lpStart(), fwd.lpEnd, .01) != 0)
    Warn ("Forward log-likelihood is %g, Backward log-likelihood is %g", fwd.lpEnd, lpStart());
}
