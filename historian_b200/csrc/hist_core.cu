// historian_b200: context, profile upload and DPMatrix construction on device.
//
// hist_pair_prepare is the B200 counterpart of DPMatrix::DPMatrix
// (reference forward.cpp:11-66): Profile::leftMultiply for both children
// (profile.cpp:78-91), the insx/insy/rootsubx/rootsuby tables (forward.cpp:44-56),
// the closest-leaf band keys (forward.cpp:36-42, alignpath.h:56-61) and the
// xNearStart / yNearEnd edge flags (forward.cpp:58-65).  The numeric tables are
// computed by kernels in this file; the integer band bookkeeping (bit-exact with
// the reference) and the anti-diagonal HBM layout are built on the host in
// O(states) and uploaded in one copy.
#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <limits>

#include "hist_internal.cuh"

namespace hb {

static thread_local std::string g_create_err;

int set_err(hist_ctx* ctx, int code, const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  if (ctx) ctx->err = buf; else g_create_err = buf;
  return code;
}

int check_cuda(hist_ctx* ctx, cudaError_t e, const char* what) {
  if (e == cudaSuccess) return 0;
  return set_err(ctx, HIST_ERR_CUDA, "CUDA error %s: %s (%s)", cudaGetErrorName(e), cudaGetErrorString(e), what);
}

// ------------------------------------------------------------------ F0 / F1 kernels
// One CTA per profile state.  Thread k < K computes sub[i][c][a] for k = c*A + a:
//   sub = m_c + log( sum_d P_c(a->d) * exp(lpAbsorb[i][c][d] - m_c) )        (profile.cpp:78-91)
// in FP64 probability space (the reference does A log-sum-exps in log space; the
// results agree to the reference table's own error).  Then the CTA
// reduces rootsub = log sum_k exp(logRoot[k] + sub[k]) and
// ins = LSE_c(logCptWeight[c] + log sum_a exp(logIns[c][a] + lpAbsorb[i][c][a]))   (forward.cpp:44-56)
// and writes the FP32 tables the fill kernels use, scaled by 2^-ex[i].
__global__ void prepare_side_kernel(SideDev s, int C, int A, const double* __restrict__ subMat,
                                    const double* __restrict__ logRoot, const double* __restrict__ logIns,
                                    const double* __restrict__ logCptWeight, int foldRoot) {
  extern __shared__ double sm[];
  const int K = C * A;
  double* e_abs = sm;          // [K] exp(lpAbsorb - m_c)
  double* red = sm + K;        // [K] scratch
  double* mc = sm + 2 * K;     // [C] per-component max
  const int i = blockIdx.x;
  const int k = threadIdx.x;
  const bool emit = (s.flags[i] & F_EMIT) != 0;
  if (!emit || i == 0 || i == s.n - 1) {
    if (k < K) {
      s.sub[(size_t)i * K + k] = -INFINITY;
      s.H[(size_t)i * K + k] = 0.f;
    }
    if (k == 0) {
      s.ex[i] = 0;
      s.rs[i] = 1.f;
      s.ins[i] = 1.f;
      s.rootsub[i] = -INFINITY;
      s.insl[i] = -INFINITY;
    }
    return;
  }
  const double* la = s.lp_absorb + (size_t)i * K;
  if (k < C) {
    double m = -INFINITY;
    for (int d = 0; d < A; ++d) m = fmax(m, la[k * A + d]);
    mc[k] = m;
  }
  __syncthreads();
  if (k < K) {
    const int c = k / A;
    e_abs[k] = mc[c] == -INFINITY ? 0.0 : exp(la[k] - mc[c]);
  }
  __syncthreads();
  double subv = -INFINITY;
  if (k < K) {
    const int c = k / A, a = k - c * A;
    const double* P = subMat + ((size_t)c * A + a) * A;
    double acc = 0;
    for (int d = 0; d < A; ++d) acc = fma(P[d], e_abs[c * A + d], acc);
    subv = acc > 0 ? mc[c] + log(acc) : -INFINITY;
    s.sub[(size_t)i * K + k] = subv;
    red[k] = subv;
  }
  __syncthreads();
  // block max of sub -> ex
  __shared__ double s_max, s_root, s_ins;
  if (k == 0) {
    double m = -INFINITY;
    for (int q = 0; q < K; ++q) m = fmax(m, red[q]);
    s_max = m;
    // rootsub
    double lr = -INFINITY;
    for (int q = 0; q < K; ++q) lr = fmax(lr, logRoot[q] + red[q]);
    double acc = 0;
    if (lr > -INFINITY)
      for (int q = 0; q < K; ++q) acc += exp(logRoot[q] + red[q] - lr);
    s_root = lr > -INFINITY ? lr + log(acc) : -INFINITY;
    // ins
    double li = -INFINITY;
    for (int c = 0; c < C; ++c)
      for (int a = 0; a < A; ++a) li = fmax(li, logCptWeight[c] + logIns[c * A + a] + la[c * A + a]);
    double acc2 = 0;
    if (li > -INFINITY)
      for (int c = 0; c < C; ++c)
        for (int a = 0; a < A; ++a) acc2 += exp(logCptWeight[c] + logIns[c * A + a] + la[c * A + a] - li);
    s_ins = li > -INFINITY ? li + log(acc2) : -INFINITY;
  }
  __syncthreads();
  const double mx = s_max;
  const int ex = mx > -INFINITY ? (int)floor(mx / HB_LN2) : 0;
  if (k < K) {
    double h = subv > -INFINITY ? exp(subv - ex * HB_LN2) : 0.0;
    if (foldRoot) h *= exp(logRoot[k]);
    s.H[(size_t)i * K + k] = (float)h;
  }
  if (k == 0) {
    s.ex[i] = ex;
    s.rootsub[i] = s_root;
    s.insl[i] = s_ins;
    s.rs[i] = s_root > -INFINITY ? (float)exp(s_root - ex * HB_LN2) : 0.f;
    s.ins[i] = s_ins > -INFINITY ? (float)exp(s_ins - ex * HB_LN2) : 0.f;
  }
}

// Per transition: lp -> (mantissa, exponent) with the destination's emission exponent folded in.
__global__ void prepare_trans_kernel(SideDev s) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= s.nt) return;
  const double lp = s.t_lp[t];
  if (!(lp > -INFINITY)) {
    s.t_m[t] = 0.f;
    s.t_e[t] = 0;
    return;
  }
  const double l2 = lp / HB_LN2;
  const double fe = floor(l2);
  s.t_m[t] = (float)exp2(l2 - fe);
  s.t_e[t] = (int)fe + s.ex[s.t_dst[t]];
}

// Exact in-envelope cell count (the GCUPS unit, forward.cpp:78-86), one thread per stored slot.
__global__ void count_cells_kernel(const PairDev* __restrict__ Pp) {
  const PairDev& P = *Pp;
  __shared__ unsigned long long s_cnt;
  if (threadIdx.x == 0) s_cnt = 0;
  __syncthreads();
  unsigned long long local = 0;
  for (int d = blockIdx.x; d < P.n_diag; d += gridDim.x) {
    const DiagInfo di = P.dinfo[d];
    const int nthr = P.nS + di.width + P.nE;
    for (int t = threadIdx.x; t < nthr; t += blockDim.x) {
      int i, j;
      bool ok = true;
      if (t < P.nS) {
        i = P.start_rows[t];
        j = d - i;
        if (i >= di.ilo && i < di.ilo + di.width) ok = false;
      } else if (t < P.nS + di.width) {
        i = di.ilo + (t - P.nS);
        j = d - i;
      } else {
        j = P.end_cols[t - P.nS - di.width];
        i = d - j;
        if (i >= di.ilo && i < di.ilo + di.width) ok = false;
        if (ok && i >= 0 && i <= P.x.n - 2 && P.x.slot[i] >= 0) ok = false;
      }
      if (!ok || i < 0 || j < 0 || i > P.x.n - 2 || j > P.y.n - 2) continue;
      const bool env = P.x.near_[i] || P.y.near_[j] || P.max_dist < 0 || hb_iabs(P.x.key[i] - P.y.key[j]) <= P.max_dist;
      if (env) ++local;
    }
  }
  atomicAdd(&s_cnt, local);
  __syncthreads();
  if (threadIdx.x == 0 && s_cnt) atomicAdd(P.n_cells, s_cnt);
}

}  // namespace hb

using namespace hb;

// ------------------------------------------------------------------ context
extern "C" int hist_ctx_create(int device, hist_ctx** out) {
  if (!out) return set_err(nullptr, HIST_ERR_ARG, "hist_ctx_create: out is NULL");
  *out = nullptr;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0)
    return set_err(nullptr, HIST_ERR_NO_DEVICE, "no CUDA device available (%s); historian_b200 has no CPU fallback",
                   e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
  if (device < 0 || device >= ndev) return set_err(nullptr, HIST_ERR_ARG, "device %d out of range (have %d)", device, ndev);
  e = cudaSetDevice(device);
  if (e != cudaSuccess) return set_err(nullptr, HIST_ERR_CUDA, "cudaSetDevice: %s", cudaGetErrorString(e));
  hist_ctx* ctx = new hist_ctx();
  ctx->device = device;
  ctx->launches = 0;
  ctx->timing = false;
  e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking);
  if (e != cudaSuccess) {
    set_err(nullptr, HIST_ERR_CUDA, "cudaStreamCreate: %s", cudaGetErrorString(e));
    delete ctx;
    return HIST_ERR_CUDA;
  }
  cudaDeviceProp prop;
  cudaGetDeviceProperties(&prop, device);
  ctx->sm_count = prop.multiProcessorCount;
  ctx->smem_optin = prop.sharedMemPerBlockOptin;
  *out = ctx;
  return HIST_OK;
}

extern "C" void hist_ctx_destroy(hist_ctx* ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  for (auto& e : ctx->fill_events) { cudaEventDestroy(e.first); cudaEventDestroy(e.second); }
  cudaStreamDestroy(ctx->stream);
  delete ctx;
}

extern "C" const char* hist_last_error(const hist_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_err.c_str(); }
extern "C" void* hist_ctx_stream(hist_ctx* ctx) { return ctx ? (void*)ctx->stream : nullptr; }
extern "C" int hist_ctx_synchronize(hist_ctx* ctx) {
  if (!ctx) return HIST_ERR_ARG;
  HB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return HIST_OK;
}
extern "C" int hist_ctx_set_timing(hist_ctx* ctx, int enabled) {
  if (!ctx) return HIST_ERR_ARG;
  ctx->timing = enabled != 0;
  return HIST_OK;
}
extern "C" int hist_ctx_fill_time_ms(hist_ctx* ctx, double* total_ms, int64_t* n_launches) {
  if (!ctx || !total_ms) return HIST_ERR_ARG;
  cudaSetDevice(ctx->device);
  double tot = 0;
  for (auto& e : ctx->fill_events) {
    HB_CUDA(ctx, cudaEventSynchronize(e.second));
    float ms = 0;
    HB_CUDA(ctx, cudaEventElapsedTime(&ms, e.first, e.second));
    tot += ms;
    cudaEventDestroy(e.first);
    cudaEventDestroy(e.second);
  }
  *total_ms = tot;
  if (n_launches) *n_launches = (int64_t)ctx->fill_events.size();
  ctx->fill_events.clear();
  return HIST_OK;
}
extern "C" int64_t hist_ctx_launch_count(const hist_ctx* ctx) { return ctx ? ctx->launches : 0; }

// ------------------------------------------------------------------ profiles
extern "C" int hist_profile_upload(hist_ctx* ctx, const hist_profile_desc* d, hist_profile** out) {
  if (!ctx || !d || !out) return set_err(ctx, HIST_ERR_ARG, "hist_profile_upload: NULL argument");
  *out = nullptr;
  if (d->n_states < 2 || d->n_trans < 0 || d->alph_size < 1 || d->components < 1)
    return set_err(ctx, HIST_ERR_ARG, "hist_profile_upload: bad sizes (%d states, %d transitions, A=%d, C=%d)", d->n_states,
                   d->n_trans, d->alph_size, d->components);
  for (int t = 0; t < d->n_trans; ++t)
    if (d->trans_src[t] < 0 || d->trans_dest[t] >= d->n_states || d->trans_src[t] >= d->trans_dest[t])
      return set_err(ctx, HIST_ERR_ARG, "hist_profile_upload: transition %d (%d -> %d) is not toposorted", t, d->trans_src[t],
                     d->trans_dest[t]);
  if (d->is_emit[0] || d->is_emit[d->n_states - 1])
    return set_err(ctx, HIST_ERR_ARG, "hist_profile_upload: START and END must be null states");
  cudaSetDevice(ctx->device);
  hist_profile* p = new hist_profile();
  p->ctx = ctx;
  p->n = d->n_states;
  p->nt = d->n_trans;
  p->A = d->alph_size;
  p->C = d->components;
  p->is_emit.assign(d->is_emit, d->is_emit + p->n);
  p->t_src.assign(d->trans_src, d->trans_src + p->nt);
  p->t_dst.assign(d->trans_dest, d->trans_dest + p->nt);
  p->t_lp.assign(d->trans_lp, d->trans_lp + p->nt);
  if (d->closest_leaf_pos) p->leaf_pos.assign(d->closest_leaf_pos, d->closest_leaf_pos + p->n);
  p->empty = true;
  for (int i = 0; i < p->n; ++i)
    if (p->is_emit[i]) p->empty = false;
  const size_t bytes = (size_t)p->n * p->C * p->A * sizeof(double);
  p->d_lp_absorb = nullptr;
  int rc = check_cuda(ctx, cudaMallocAsync((void**)&p->d_lp_absorb, bytes, ctx->stream), "cudaMallocAsync(lp_absorb)");
  if (!rc) rc = check_cuda(ctx, cudaMemcpyAsync(p->d_lp_absorb, d->lp_absorb, bytes, cudaMemcpyHostToDevice, ctx->stream), "H2D lp_absorb");
  // the descriptor is only borrowed for the duration of the call
  if (!rc) rc = check_cuda(ctx, cudaStreamSynchronize(ctx->stream), "sync after profile upload");
  if (rc) {
    if (p->d_lp_absorb) cudaFreeAsync(p->d_lp_absorb, ctx->stream);
    delete p;
    return rc;
  }
  *out = p;
  return HIST_OK;
}

extern "C" void hist_profile_destroy(hist_profile* p) {
  if (!p) return;
  cudaSetDevice(p->ctx->device);
  if (p->d_lp_absorb) cudaFreeAsync(p->d_lp_absorb, p->ctx->stream);
  delete p;
}

// ------------------------------------------------------------------ pair preparation
namespace {

struct SlabBuilder {
  size_t bytes = 0;
  std::vector<std::pair<size_t, std::pair<const void*, size_t>>> copies;  // (offset, (src, nbytes))
  size_t reserve(size_t n, size_t align = 16) {
    bytes = (bytes + align - 1) / align * align;
    const size_t off = bytes;
    bytes += n;
    return off;
  }
  size_t put(const void* src, size_t n, size_t align = 16) {
    const size_t off = reserve(n, align);
    copies.push_back({off, {src, n}});
    return off;
  }
};

struct SideHost {
  std::vector<uint8_t> flags, near_;
  std::vector<int> in_off, in_idx, ao_off, ao_idx, no_off, no_idx, key, slot;
};

void build_side_lists(const hist_profile* p, SideHost& s) {
  const int n = p->n, nt = p->nt;
  std::vector<int> in_cnt(n + 1, 0), ao_cnt(n + 1, 0), no_cnt(n + 1, 0);
  for (int t = 0; t < nt; ++t) {
    ++in_cnt[p->t_dst[t] + 1];
    if (p->is_emit[p->t_dst[t]]) ++ao_cnt[p->t_src[t] + 1]; else ++no_cnt[p->t_src[t] + 1];
  }
  for (int i = 0; i < n; ++i) {
    in_cnt[i + 1] += in_cnt[i];
    ao_cnt[i + 1] += ao_cnt[i];
    no_cnt[i + 1] += no_cnt[i];
  }
  s.in_off = in_cnt; s.ao_off = ao_cnt; s.no_off = no_cnt;
  s.in_idx.assign(std::max(1, in_cnt[n]), 0);
  s.ao_idx.assign(std::max(1, ao_cnt[n]), 0);
  s.no_idx.assign(std::max(1, no_cnt[n]), 0);
  std::vector<int> a(in_cnt.begin(), in_cnt.end() - 1), b(ao_cnt.begin(), ao_cnt.end() - 1), c(no_cnt.begin(), no_cnt.end() - 1);
  for (int t = 0; t < nt; ++t) {  // ascending transition index = the reference's list order
    s.in_idx[a[p->t_dst[t]]++] = t;
    if (p->is_emit[p->t_dst[t]]) s.ao_idx[b[p->t_src[t]]++] = t; else s.no_idx[c[p->t_src[t]]++] = t;
  }
  s.flags.assign(n, 0);
  for (int i = 0; i < n; ++i) {
    uint8_t f = 0;
    if (p->is_emit[i]) f |= F_EMIT;
    if (s.no_off[i + 1] == s.no_off[i] || p->empty) f |= F_READY;            // isReady() || xEmpty (forward.cpp:97)
    if (p->is_emit[i] || s.in_off[i + 1] == s.in_off[i]) f |= F_EMIT_OR_START;  // profile.h:34-35
    s.flags[i] = f;
  }
}

}  // namespace

extern "C" int hist_pair_prepare(hist_ctx* ctx, const hist_profile* x, const hist_profile* y, const hist_pairhmm_desc* hmm,
                                 const hist_env_desc* env, hist_pair** out) {
  if (!ctx || !x || !y || !hmm || !out) return set_err(ctx, HIST_ERR_ARG, "hist_pair_prepare: NULL argument");
  *out = nullptr;
  if (x->A != hmm->alph_size || y->A != hmm->alph_size || x->C != hmm->components || y->C != hmm->components)
    return set_err(ctx, HIST_ERR_ARG, "hist_pair_prepare: alphabet/component mismatch between profiles and PairHMM");
  const int nx = x->n, ny = y->n, A = x->A, C = x->C, K = A * C;
  if (K > 1024) return set_err(ctx, HIST_ERR_ARG, "hist_pair_prepare: C*A = %d exceeds 1024", K);
  const int maxDist = env ? env->max_distance : -1;
  if (maxDist >= 0) {
    if (x->leaf_pos.empty() || y->leaf_pos.empty())
      return set_err(ctx, HIST_ERR_ARG, "hist_pair_prepare: envelope given but profiles carry no closest_leaf_pos");
    if (!env->cumulative_matches || !env->row1_pos_to_col || !env->row2_pos_to_col)
      return set_err(ctx, HIST_ERR_ARG, "hist_pair_prepare: envelope arrays are NULL");
  }
  cudaSetDevice(ctx->device);

  SideHost sx, sy;
  build_side_lists(x, sx);
  build_side_lists(y, sy);

  // ---- envelope keys (forward.cpp:36-42, alignpath.h:56-61) and edge flags (forward.cpp:58-65)
  sx.key.assign(nx, 0);
  sy.key.assign(ny, 0);
  if (maxDist >= 0) {
    for (int i = 0; i < nx; ++i) {
      const int pos = i == 0 ? 0 : x->leaf_pos[i];
      if (pos < 0 || pos > env->row1_len) return set_err(ctx, HIST_ERR_ARG, "x state %d: closest-leaf position %d outside guide row (len %d)", i, pos, env->row1_len);
      sx.key[i] = env->cumulative_matches[env->row1_pos_to_col[pos]];
    }
    for (int j = 0; j < ny; ++j) {
      const int pos = j == 0 ? 0 : y->leaf_pos[j];
      if (pos < 0 || pos > env->row2_len) return set_err(ctx, HIST_ERR_ARG, "y state %d: closest-leaf position %d outside guide row (len %d)", j, pos, env->row2_len);
      sy.key[j] = env->cumulative_matches[env->row2_pos_to_col[pos]];
    }
  }
  sx.near_.assign(nx, 0);
  sy.near_.assign(ny, 0);
  sx.near_[0] = 1;
  for (int i = 0; i < nx; ++i)
    if (sx.near_[i])
      for (int q = sx.no_off[i]; q < sx.no_off[i + 1]; ++q) sx.near_[x->t_dst[sx.no_idx[q]]] = 1;
  for (int q = sy.in_off[ny - 1]; q < sy.in_off[ny]; ++q) sy.near_[y->t_src[sy.in_idx[q]]] = 1;

  // ---- anti-diagonal layout.  Cells exist for 0<=i<=nx-2, 0<=j<=ny-2.
  const int NI = nx - 1, NJ = ny - 1;  // number of rows / columns
  std::vector<int> jlo(NI), jhi(NI);   // bounding column interval of the in-range cells of each row (jlo>jhi: empty)
  if (maxDist < 0) {
    for (int i = 0; i < NI; ++i) { jlo[i] = 0; jhi[i] = NJ - 1; }
  } else {
    // prefix max / suffix min of the y keys make "first j with key >= a" / "last j with key <= b" binary-searchable
    std::vector<int> pmax(NJ), smin(NJ);
    for (int j = 0; j < NJ; ++j) pmax[j] = j ? std::max(pmax[j - 1], sy.key[j]) : sy.key[0];
    for (int j = NJ - 1; j >= 0; --j) smin[j] = j < NJ - 1 ? std::min(smin[j + 1], sy.key[j]) : sy.key[NJ - 1];
    for (int i = 0; i < NI; ++i) {
      const int a = sx.key[i] - maxDist, b = sx.key[i] + maxDist;
      const int lo = (int)(std::lower_bound(pmax.begin(), pmax.end(), a) - pmax.begin());        // first j: key[j] >= a
      const int hi = (int)(std::upper_bound(smin.begin(), smin.end(), b) - smin.begin()) - 1;    // last j: key[j] <= b
      jlo[i] = lo;
      jhi[i] = hi;
    }
  }
  // monotone hull so that every anti-diagonal meets the band in ONE interval of i
  {
    int run = -1;
    for (int i = 0; i < NI; ++i) {  // jhi non-decreasing
      if (jlo[i] <= jhi[i]) run = std::max(run, jhi[i]);
      if (run >= 0 && jlo[i] <= jhi[i]) jhi[i] = run;
    }
    run = NJ;
    for (int i = NI - 1; i >= 0; --i) {  // jlo non-decreasing
      if (jlo[i] <= jhi[i]) run = std::min(run, jlo[i]);
      if (jlo[i] <= jhi[i]) jlo[i] = run;
    }
  }
  const int nDiag = NI + NJ - 1;
  std::vector<int> dlo(nDiag, std::numeric_limits<int>::max()), dhi(nDiag, -1);
  for (int i = 0; i < NI; ++i) {
    if (jlo[i] > jhi[i]) continue;
    // row i covers diagonals i+jlo .. i+jhi; with monotone jlo/jhi the per-diagonal set is an interval,
    // so only its end points need updating: sweep the two boundary chains
    const int d0 = i + jlo[i], d1 = i + jhi[i];
    // dlo[d] = min i with d in [d0,d1]; dhi[d] = max i.  (rows are visited in ascending i)
    for (int d = d0; d <= d1; ++d) {
      if (dlo[d] == std::numeric_limits<int>::max()) dlo[d] = i;
      dhi[d] = i;
    }
  }
  // (the loop above is O(stored cells); good enough on the host for banded merges)

  hist_pair* pr = new hist_pair();
  pr->ctx = ctx;
  pr->x = x;
  pr->y = y;
  pr->fwd_done = pr->bwd_done = false;
  pr->lp_end = pr->lp_start = -INFINITY;
  // edge slots
  sx.slot.assign(nx, -1);
  sy.slot.assign(ny, -1);
  for (int i = 0; i < NI; ++i)
    if (sx.near_[i]) { sx.slot[i] = (int)pr->start_rows.size(); pr->start_rows.push_back(i); }
  for (int j = 0; j < NJ; ++j)
    if (sy.near_[j]) { sy.slot[j] = (int)pr->end_cols.size(); pr->end_cols.push_back(j); }
  const int nS = (int)pr->start_rows.size(), nE = (int)pr->end_cols.size();
  pr->dinfo.resize(nDiag);
  unsigned long long off = 0;
  int maxThreads = 1, maxWidth = 0;
  for (int d = 0; d < nDiag; ++d) {
    DiagInfo di;
    if (dhi[d] >= 0) { di.ilo = dlo[d]; di.width = dhi[d] - dlo[d] + 1; } else { di.ilo = 0; di.width = 0; }
    di.off = (unsigned)off;
    di.pad = 0;
    pr->dinfo[d] = di;
    off += (unsigned long long)(nS + di.width + nE);
    maxThreads = std::max(maxThreads, nS + di.width + nE);
    maxWidth = std::max(maxWidth, di.width);
  }
  if (off >= (1ull << 31)) {
    delete pr;
    return set_err(ctx, HIST_ERR_ARG, "hist_pair_prepare: %llu stored cells exceed the 2^31 limit of one pair", off);
  }
  const unsigned nStored = (unsigned)off;

  // ---- slab: every per-pair array in one allocation, host-built parts in one staging buffer
  SlabBuilder sb;
  auto putv = [&](const auto& v) { return sb.put(v.data(), v.size() * sizeof(v[0])); };
  struct SideOff { size_t flags, t_src, t_dst, t_lp, t_m, t_e, in_off, in_idx, ao_off, ao_idx, no_off, no_idx, ex, rs, ins, H, sub, rootsub, insl, key, near_, slot; };
  auto layout_side = [&](const hist_profile* p, const SideHost& s) {
    SideOff o;
    o.flags = putv(s.flags);
    o.t_src = sb.put(p->t_src.data(), std::max<size_t>(1, p->nt) * sizeof(int));
    o.t_dst = sb.put(p->t_dst.data(), std::max<size_t>(1, p->nt) * sizeof(int));
    o.t_lp = sb.put(p->t_lp.data(), std::max<size_t>(1, p->nt) * sizeof(double));
    o.t_m = sb.reserve(std::max<size_t>(1, p->nt) * sizeof(float));
    o.t_e = sb.reserve(std::max<size_t>(1, p->nt) * sizeof(int));
    o.in_off = putv(s.in_off); o.in_idx = putv(s.in_idx);
    o.ao_off = putv(s.ao_off); o.ao_idx = putv(s.ao_idx);
    o.no_off = putv(s.no_off); o.no_idx = putv(s.no_idx);
    o.ex = sb.reserve((size_t)p->n * sizeof(int));
    o.rs = sb.reserve((size_t)p->n * sizeof(float));
    o.ins = sb.reserve((size_t)p->n * sizeof(float));
    o.H = sb.reserve((size_t)p->n * K * sizeof(float));
    o.sub = sb.reserve((size_t)p->n * K * sizeof(double));
    o.rootsub = sb.reserve((size_t)p->n * sizeof(double));
    o.insl = sb.reserve((size_t)p->n * sizeof(double));
    o.key = putv(s.key); o.near_ = putv(s.near_); o.slot = putv(s.slot);
    return o;
  };
  // guard against zero-length vectors handed to put(): t_src etc. of a profile always has >= 1 entry
  // (START -> ... -> END needs at least one transition), so .data() is valid.
  const SideOff ox = layout_side(x, sx), oy = layout_side(y, sy);
  std::vector<int> one0(1, 0);
  const size_t o_start = sb.put(nS ? (const void*)pr->start_rows.data() : (const void*)one0.data(), std::max(1, nS) * sizeof(int));
  const size_t o_end = sb.put(nE ? (const void*)pr->end_cols.data() : (const void*)one0.data(), std::max(1, nE) * sizeof(int));
  const size_t o_dinfo = sb.put(pr->dinfo.data(), pr->dinfo.size() * sizeof(DiagInfo));
  // model tables
  const size_t o_lsub = sb.put(hmm->l_sub, (size_t)K * A * sizeof(double));
  const size_t o_rsub = sb.put(hmm->r_sub, (size_t)K * A * sizeof(double));
  const size_t o_root = sb.put(hmm->log_root, (size_t)K * sizeof(double));
  const size_t o_lins = sb.put(hmm->l_log_ins, (size_t)K * sizeof(double));
  const size_t o_rins = sb.put(hmm->r_log_ins, (size_t)K * sizeof(double));
  const size_t o_lcw = sb.put(hmm->l_log_cpt_weight, (size_t)C * sizeof(double));
  const size_t o_rcw = sb.put(hmm->r_log_cpt_weight, (size_t)C * sizeof(double));
  const size_t o_host_end = sb.bytes;  // everything before this offset is (partly) host-initialised
  const size_t o_scal = sb.reserve(4 * sizeof(double));  // lp_end, lp_start, n_cells, spare
  const size_t o_dev = sb.reserve(sizeof(PairDev), 64);
  const size_t o_fA = sb.reserve((size_t)nStored * sizeof(float4), 256);
  const size_t o_fB = sb.reserve((size_t)nStored * sizeof(float2), 256);
  // backward planes are allocated lazily by hist_bwd_fill (most merges never need them)

  char* slab = nullptr;
  int rc = check_cuda(ctx, cudaMallocAsync((void**)&slab, sb.bytes, ctx->stream), "cudaMallocAsync(pair slab)");
  if (rc) { delete pr; return rc; }
  pr->slab = slab;
  pr->slab_bytes = sb.bytes;
  std::vector<char> staging(o_host_end, 0);
  for (auto& c : sb.copies) memcpy(staging.data() + c.first, c.second.first, c.second.second);
  rc = check_cuda(ctx, cudaMemcpyAsync(slab, staging.data(), o_host_end, cudaMemcpyHostToDevice, ctx->stream), "H2D pair slab");
  if (!rc) rc = check_cuda(ctx, cudaMemsetAsync(slab + o_scal, 0, 4 * sizeof(double), ctx->stream), "memset scalars");

  auto side_dev = [&](const hist_profile* p, const SideOff& o) {
    SideDev s;
    s.n = p->n; s.nt = p->nt;
    s.flags = (const uint8_t*)(slab + o.flags);
    s.t_src = (const int*)(slab + o.t_src); s.t_dst = (const int*)(slab + o.t_dst);
    s.t_lp = (const double*)(slab + o.t_lp);
    s.t_m = (float*)(slab + o.t_m); s.t_e = (int*)(slab + o.t_e);
    s.in_off = (const int*)(slab + o.in_off); s.in_idx = (const int*)(slab + o.in_idx);
    s.ao_off = (const int*)(slab + o.ao_off); s.ao_idx = (const int*)(slab + o.ao_idx);
    s.no_off = (const int*)(slab + o.no_off); s.no_idx = (const int*)(slab + o.no_idx);
    s.ex = (int*)(slab + o.ex); s.rs = (float*)(slab + o.rs); s.ins = (float*)(slab + o.ins);
    s.H = (float*)(slab + o.H);
    s.lp_absorb = p->d_lp_absorb;
    s.sub = (double*)(slab + o.sub); s.rootsub = (double*)(slab + o.rootsub); s.insl = (double*)(slab + o.insl);
    s.key = (const int*)(slab + o.key); s.near_ = (const uint8_t*)(slab + o.near_); s.slot = (const int*)(slab + o.slot);
    return s;
  };
  PairDev& D = pr->dev;
  D.x = side_dev(x, ox);
  D.y = side_dev(y, oy);
  D.K = K; D.C = C; D.A = A;
  D.max_dist = maxDist;
  // HMM weights, w[src][dest] (pairhmm.cpp:46-115); absent pairs are -inf / 0
  for (int a = 0; a < 6; ++a)
    for (int b = 0; b < 6; ++b) pr->w_log[a][b] = -INFINITY;
  {
    const double* l = hmm->lp_trans;
    double (*w)[6] = pr->w_log;
    w[0][0] = l[HIST_W_IMM_IMM]; w[0][1] = l[HIST_W_IMM_IMD]; w[0][2] = l[HIST_W_IMM_IDM];
    w[0][3] = l[HIST_W_IMM_IMI]; w[0][4] = l[HIST_W_IMM_IIW]; w[0][5] = l[HIST_W_IMM_EEE];
    w[1][0] = l[HIST_W_IMD_IMM]; w[1][1] = l[HIST_W_IMD_IMD]; w[1][2] = l[HIST_W_IMD_IDM]; w[1][5] = l[HIST_W_IMD_EEE];
    w[2][0] = l[HIST_W_IDM_IMM]; w[2][1] = l[HIST_W_IDM_IMD]; w[2][2] = l[HIST_W_IDM_IDM]; w[2][5] = l[HIST_W_IDM_EEE];
    w[3][0] = l[HIST_W_IMI_IMM]; w[3][1] = l[HIST_W_IMI_IMD]; w[3][3] = l[HIST_W_IMI_IMI];
    w[3][4] = l[HIST_W_IMI_IIW]; w[3][5] = l[HIST_W_IMI_EEE];
    w[4][0] = l[HIST_W_IIW_IMM]; w[4][2] = l[HIST_W_IIW_IDM]; w[4][4] = l[HIST_W_IIW_IIW]; w[4][5] = l[HIST_W_IIW_EEE];
    for (int a = 0; a < 5; ++a)
      for (int b = 0; b < 6; ++b) D.W[a][b] = w[a][b] > -INFINITY ? (float)exp(w[a][b]) : 0.f;
  }
  pr->hmm_scalars = *hmm;
  pr->hmm_scalars.log_root = pr->hmm_scalars.l_sub = pr->hmm_scalars.r_sub = nullptr;
  pr->hmm_scalars.l_log_ins = pr->hmm_scalars.r_log_ins = nullptr;
  pr->hmm_scalars.l_log_cpt_weight = pr->hmm_scalars.r_log_cpt_weight = nullptr;
  D.n_diag = nDiag;
  D.nS = nS; D.nE = nE;
  D.start_rows = (const int*)(slab + o_start);
  D.end_cols = (const int*)(slab + o_end);
  D.dinfo = (const DiagInfo*)(slab + o_dinfo);
  D.max_threads = maxThreads;
  D.n_stored = nStored;
  D.fA = (float4*)(slab + o_fA); D.fB = (float2*)(slab + o_fB);
  D.bA = nullptr; D.bB = nullptr;
  D.lp_end = (double*)(slab + o_scal);
  D.lp_start = (double*)(slab + o_scal) + 1;
  D.n_cells = (unsigned long long*)((double*)(slab + o_scal) + 2);
  pr->d_dev = (PairDev*)(slab + o_dev);
  if (!rc) rc = check_cuda(ctx, cudaMemcpyAsync(pr->d_dev, &D, sizeof(PairDev), cudaMemcpyHostToDevice, ctx->stream), "H2D PairDev");

  // ---- F0/F1 on device
  if (!rc) {
    const int threads = std::max(32, (K + 31) / 32 * 32);
    const size_t smem = (size_t)(2 * K + C) * sizeof(double);
    prepare_side_kernel<<<nx, threads, smem, ctx->stream>>>(D.x, C, A, (const double*)(slab + o_lsub), (const double*)(slab + o_root),
                                                           (const double*)(slab + o_lins), (const double*)(slab + o_lcw), 0);
    prepare_side_kernel<<<ny, threads, smem, ctx->stream>>>(D.y, C, A, (const double*)(slab + o_rsub), (const double*)(slab + o_root),
                                                           (const double*)(slab + o_rins), (const double*)(slab + o_rcw), 1);
    prepare_trans_kernel<<<(x->nt + 127) / 128, 128, 0, ctx->stream>>>(D.x);
    prepare_trans_kernel<<<(y->nt + 127) / 128, 128, 0, ctx->stream>>>(D.y);
    count_cells_kernel<<<std::min(nDiag, 4 * ctx->sm_count), 128, 0, ctx->stream>>>(pr->d_dev);
    ctx->launches += 5;
    rc = check_cuda(ctx, cudaGetLastError(), "prepare kernels");
  }
  unsigned long long ncells = 0;
  if (!rc) rc = check_cuda(ctx, cudaMemcpyAsync(&ncells, D.n_cells, sizeof(ncells), cudaMemcpyDeviceToHost, ctx->stream), "D2H n_cells");
  // staging buffers are host-owned: wait before they go out of scope
  if (!rc) rc = check_cuda(ctx, cudaStreamSynchronize(ctx->stream), "sync after pair prepare");
  if (rc) {
    cudaFreeAsync(slab, ctx->stream);
    delete pr;
    return rc;
  }
  pr->x_key = sx.key; pr->y_key = sy.key;
  pr->x_near = sx.near_; pr->y_near = sy.near_;
  pr->x_slot = sx.slot; pr->y_slot = sy.slot;
  pr->info.x_size = nx;
  pr->info.y_size = ny;
  pr->info.n_cells = (int64_t)ncells;
  pr->info.n_stored = nStored;
  pr->info.n_diagonals = nDiag;
  pr->info.max_diag_width = maxWidth;
  pr->info.h2d_bytes = (int64_t)(o_host_end + sizeof(PairDev));
  *out = pr;
  return HIST_OK;
}

extern "C" void hist_pair_destroy(hist_pair* pr) {
  if (!pr) return;
  cudaSetDevice(pr->ctx->device);
  if (pr->dev.bA) cudaFreeAsync(pr->dev.bA, pr->ctx->stream);
  if (pr->slab) cudaFreeAsync(pr->slab, pr->ctx->stream);
  delete pr;
}

extern "C" int hist_pair_get_info(const hist_pair* pr, hist_pair_info* info) {
  if (!pr || !info) return HIST_ERR_ARG;
  *info = pr->info;
  return HIST_OK;
}

extern "C" int hist_pair_fetch_tables(hist_pair* pr, double* subx, double* suby, double* insx, double* insy, double* rootsubx,
                                      double* rootsuby) {
  if (!pr) return HIST_ERR_ARG;
  hist_ctx* ctx = pr->ctx;
  cudaSetDevice(ctx->device);
  const PairDev& D = pr->dev;
  const size_t K = D.K;
  if (subx) HB_CUDA(ctx, cudaMemcpyAsync(subx, D.x.sub, D.x.n * K * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
  if (suby) HB_CUDA(ctx, cudaMemcpyAsync(suby, D.y.sub, D.y.n * K * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
  if (insx) HB_CUDA(ctx, cudaMemcpyAsync(insx, D.x.insl, D.x.n * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
  if (insy) HB_CUDA(ctx, cudaMemcpyAsync(insy, D.y.insl, D.y.n * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
  if (rootsubx) HB_CUDA(ctx, cudaMemcpyAsync(rootsubx, D.x.rootsub, D.x.n * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
  if (rootsuby) HB_CUDA(ctx, cudaMemcpyAsync(rootsuby, D.y.rootsub, D.y.n * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
  HB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  return HIST_OK;
}

extern "C" int hist_pair_fetch_envelope(hist_pair* pr, int32_t* x_key, int32_t* y_key, uint8_t* x_near_start, uint8_t* y_near_end,
                                        uint8_t* in_envelope) {
  if (!pr) return HIST_ERR_ARG;
  const int nx = pr->info.x_size, ny = pr->info.y_size;
  if (x_key) memcpy(x_key, pr->x_key.data(), nx * sizeof(int));
  if (y_key) memcpy(y_key, pr->y_key.data(), ny * sizeof(int));
  if (x_near_start) memcpy(x_near_start, pr->x_near.data(), nx);
  if (y_near_end) memcpy(y_near_end, pr->y_near.data(), ny);
  if (in_envelope) {
    const int md = pr->dev.max_dist;
    for (int i = 0; i < nx - 1; ++i)
      for (int j = 0; j < ny - 1; ++j)
        in_envelope[(size_t)i * (ny - 1) + j] =
            pr->x_near[i] || pr->y_near[j] || md < 0 || hb_iabs(pr->x_key[i] - pr->y_key[j]) <= md;
  }
  return HIST_OK;
}
