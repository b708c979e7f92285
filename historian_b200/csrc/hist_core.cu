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
#include <cctype>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <climits>

#include <atomic>
#include <condition_variable>
#include <functional>
#include <chrono>
#include <mutex>
#include <thread>
#include "hist_internal.cuh"

namespace hb {

static thread_local std::string g_create_err;

int set_err(hist_ctx* ctx, int code, const char* fmt, ...) {
  char buf[1024];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof(buf), fmt, ap);
  va_end(ap);
  {
    static std::mutex mu;  // build_pair_host runs on several host threads
    std::lock_guard<std::mutex> lock(mu);
    if (ctx) ctx->err = buf; else g_create_err = buf;
  }
  return code;
}

int check_cuda(hist_ctx* ctx, cudaError_t e, const char* what) {
  if (e == cudaSuccess) return 0;
  return set_err(ctx, HIST_ERR_CUDA, "CUDA error %s: %s (%s)", cudaGetErrorName(e), cudaGetErrorString(e), what);
}

// ------------------------------------------------------------------ F0 / F1 kernels
// grid = (states, pairs, 2 sides).  One CTA per profile state.  Thread k < K computes sub[i][c][a] for k = c*A + a:
//   sub = m_c + log( sum_d P_c(a->d) * exp(lpAbsorb[i][c][d] - m_c) )        (profile.cpp:78-91)
// in FP64 probability space (the reference does A log-sum-exps in log space; the
// results agree to the reference table's own error).  Then the CTA
// reduces rootsub = log sum_k exp(logRoot[k] + sub[k]) and
// ins = LSE_c(logCptWeight[c] + log sum_a exp(logIns[c][a] + lpAbsorb[i][c][a]))   (forward.cpp:44-56)
// and writes the FP32 tables the fill kernels use, scaled by 2^-ex[i].
__device__ __forceinline__ double warp_max(double v) {
  for (int o = 16; o; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ double warp_sum(double v) {
  for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
// block-wide reduction through shared scratch `red` (blockDim.x/32 doubles); every thread gets the result
template <bool MAX>
__device__ __forceinline__ double block_reduce(double v, double* red) {
  v = MAX ? warp_max(v) : warp_sum(v);
  const int w = threadIdx.x >> 5, nw = blockDim.x >> 5;
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[w] = v;
  __syncthreads();
  double r = red[0];
  for (int q = 1; q < nw; ++q) r = MAX ? fmax(r, red[q]) : r + red[q];
  return r;
}

// exp / log with the exact trivial cases taken out (one-hot leaf rows are all trivial cases): same results
__device__ __forceinline__ double exp_t(double x) { return x == 0.0 ? 1.0 : x == -INFINITY ? 0.0 : exp(x); }
__device__ __forceinline__ double log_t(double x) { return x == 1.0 ? 0.0 : log(x); }

// WARP: C*A <= 32, one warp per state and blockDim.y states per CTA (a CTA per state would be a million
// 32-thread CTAs for a level of merges); otherwise one CTA per state.  Same arithmetic either way.
// TOK: the rows are the A+1 token classes of a leaf side (index = token, A = wildcard) instead of its states.
template <bool WARP, bool TOK = false>
__global__ void prepare_side_kernel(PairDev* const* __restrict__ pairs, int sm_stride) {
  extern __shared__ double sm_all[];
  double* sm = WARP ? sm_all + (size_t)threadIdx.y * sm_stride : sm_all;
  const PairDev& P = *pairs[blockIdx.y];
  const bool isY = blockIdx.z == 1;
  const SideDev& s = isY ? P.y : P.x;
  const int i = WARP ? blockIdx.x * blockDim.y + threadIdx.y : blockIdx.x;
  if ((s.tokens != nullptr) != TOK) return;   // leaf sides go through the token-class pass and leaf_scatter_kernel
  if (i >= (TOK ? P.A + 1 : s.n)) return;
  double* const o_sub = TOK ? s.tk_sub : s.sub;
  float* const o_H = TOK ? s.tk_H : s.H;
  int* const o_ex = TOK ? s.tk_ex : s.ex;
  float* const o_rs = TOK ? s.tk_rs : s.rs;
  float* const o_ins = TOK ? s.tk_ins : s.ins;
  double* const o_rootsub = TOK ? s.tk_rootsub : s.rootsub;
  double* const o_insl = TOK ? s.tk_insl : s.insl;
  auto sync = [&]() { if (WARP) __syncwarp(); else __syncthreads(); };
  auto red_max = [&](double v, double* r) { return WARP ? warp_max(v) : block_reduce<true>(v, r); };
  auto red_sum = [&](double v, double* r) { return WARP ? warp_sum(v) : block_reduce<false>(v, r); };
  const int C = P.C, A = P.A, K = P.K, Kp = P.Kp;
  const double* __restrict__ subMat = isY ? P.r_sub : P.l_sub;
  const double* __restrict__ logRoot = P.log_root;
  const double* __restrict__ logIns = isY ? P.r_ins : P.l_ins;
  const double* __restrict__ logCptWeight = isY ? P.r_cw : P.l_cw;
  double* e_abs = sm;          // [K] exp(lpAbsorb - m_c)
  double* mc = sm + K;         // [C] per-component max
  double* red = sm + K + C;    // [32] reduction scratch
  const int k = threadIdx.x;
  const bool emit = TOK || (s.flags[i] & F_EMIT) != 0;
  if (!TOK && (!emit || i == 0 || i == s.n - 1)) {
    if (k < K) s.sub[(size_t)i * K + k] = -INFINITY;
    if (k < Kp) s.H[(size_t)i * Kp + k] = 0.f;
    if (k == 0) {
      s.ex[i] = 0;
      s.rs[i] = 1.f;
      s.ins[i] = 1.f;
      s.rootsub[i] = -INFINITY;
      s.insl[i] = -INFINITY;
    }
    return;
  }
  const double* la = (TOK ? s.tk_onehot : s.lp_absorb) + (size_t)i * K;
  if (k < C) {
    double m = -INFINITY;
    for (int d = 0; d < A; ++d) m = fmax(m, la[k * A + d]);
    mc[k] = m;
  }
  sync();
  const int c = k < K ? k / A : 0, a = k - c * A;
  const double lak = k < K ? la[k] : -INFINITY;
  if (k < K) e_abs[k] = mc[c] == -INFINITY ? 0.0 : exp_t(lak - mc[c]);
  sync();
  double subv = -INFINITY;
  if (k < K) {
    const double* Pm = subMat + ((size_t)c * A + a) * A;
    double acc = 0;
    for (int d = 0; d < A; ++d) acc = fma(Pm[d], e_abs[c * A + d], acc);
    subv = acc > 0 ? mc[c] + log(acc) : -INFINITY;
    o_sub[(size_t)i * K + k] = subv;
  }
  // block reductions: max of sub -> ex; rootsub; ins
  const double mx = red_max(subv, red);
  const double rterm = k < K ? logRoot[k] + subv : -INFINITY;
  const double lr = red_max(rterm, red);
  const double racc = red_sum(lr > -INFINITY && k < K ? exp(rterm - lr) : 0.0, red);
  const double s_root = lr > -INFINITY ? lr + log(racc) : -INFINITY;
  const double iterm = k < K ? logCptWeight[c] + logIns[k] + lak : -INFINITY;
  const double li = red_max(iterm, red);
  const double iacc = red_sum(li > -INFINITY && k < K ? exp_t(iterm - li) : 0.0, red);
  const double s_ins = li > -INFINITY ? li + log_t(iacc) : -INFINITY;
  const int ex = mx > -INFINITY ? (int)floor(mx / HB_LN2) : 0;
  if (k < Kp) {
    double h = subv > -INFINITY ? exp(subv - ex * HB_LN2) : 0.0;
    if (isY && k < K) h *= exp(logRoot[k]);
    o_H[(size_t)i * Kp + k] = (float)h;
  }
  if (k == 0) {
    o_ex[i] = ex;
    o_rootsub[i] = s_root;
    o_insl[i] = s_ins;
    o_rs[i] = s_root > -INFINITY ? (float)exp(s_root - ex * HB_LN2) : 0.f;
    o_ins[i] = s_ins > -INFINITY ? (float)exp(s_ins - ex * HB_LN2) : 0.f;
  }
}

// Leaf sides: every state takes the rows of its residue's token class.  grid = (chunks, pairs, 2 sides); thread e of a
// side writes element e of its state-major `sub` array, and of `H` and of the per-state scalars where e is in range, so
// that consecutive threads write consecutive addresses.
__global__ void leaf_scatter_kernel(PairDev* const* __restrict__ pairs) {
  const PairDev& P = *pairs[blockIdx.y];
  const SideDev& s = blockIdx.z == 1 ? P.y : P.x;
  if (!s.tokens) return;
  const int K = P.K, Kp = P.Kp, A = P.A, n = s.n;
  const unsigned total = (unsigned)n * (unsigned)max(K, Kp);   // the host keeps n * max(K, Kp) below 2^31 for token sides
  const unsigned nK = (unsigned)n * (unsigned)K, nKp = (unsigned)n * (unsigned)Kp;
  for (unsigned e = blockIdx.x * blockDim.x + threadIdx.x; e < total; e += gridDim.x * blockDim.x) {
    auto token_class = [&](int i) {   // -1: START / END (the null-state rows of prepare_side_kernel)
      if (i == 0 || i == n - 1) return -1;
      const int q = s.tokens[i - 1];
      return q < 0 || q >= A ? A : q;
    };
    if (e < nK) {
      const unsigned i = e / (unsigned)K, k = e - i * (unsigned)K;
      const int q = token_class((int)i);
      s.sub[e] = q < 0 ? -INFINITY : s.tk_sub[(size_t)q * K + k];
    }
    if (e < nKp) {
      const unsigned i = e / (unsigned)Kp, k = e - i * (unsigned)Kp;
      const int q = token_class((int)i);
      s.H[e] = q < 0 ? 0.f : s.tk_H[(size_t)q * Kp + k];
    }
    if (e < (unsigned)n) {
      const int i = (int)e, q = token_class(i);
      s.ex[i] = q < 0 ? 0 : s.tk_ex[q];
      s.rs[i] = q < 0 ? 1.f : s.tk_rs[q];
      s.ins[i] = q < 0 ? 1.f : s.tk_ins[q];
      s.rootsub[i] = q < 0 ? -INFINITY : s.tk_rootsub[q];
      s.insl[i] = q < 0 ? -INFINITY : s.tk_insl[q];
    }
  }
}

// Per transition: lp -> (mantissa, exponent) with the destination's emission exponent folded in.
// grid = (chunks, pairs, 2 sides).
__global__ void prepare_trans_kernel(PairDev* const* __restrict__ pairs) {
  const PairDev& P = *pairs[blockIdx.y];
  const SideDev& s = blockIdx.z == 1 ? P.y : P.x;
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

// Packed per-state and per-edge records of the fill kernels (layout: hist_internal.cuh).
// grid = (chunks, pairs, 2 sides), one thread per state.
__global__ void pack_meta_kernel(PairDev* const* __restrict__ pairs) {
  const PairDev& P = *pairs[blockIdx.y];
  const SideDev& s = blockIdx.z == 1 ? P.y : P.x;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= s.n) return;
  int flags = s.flags[i];
  if (s.near_[i]) flags |= F_NEAR;
  // in-edges
  const int ib = s.in_off[i], ic = s.in_off[i + 1] - ib;
  int4 first = make_int4(0, 0, 0, -1);
  for (int q = 0; q < ic; ++q) {
    const int t = s.in_idx[ib + q];
    const int src = s.t_src[t];
    const int4 e = make_int4(src, s.t_e[t], __float_as_int(s.t_m[t]), s.slot[src]);
    s.in_edges[ib + q] = e;
    if (q == 0) first = e;
  }
  // out-edges
  const int ab = s.ao_off[i], ac = s.ao_off[i + 1] - ab;
  int4 f0 = make_int4(0, 0, -1, 0), f1 = make_int4(0, 0, 0, 0);
  for (int q = 0; q < ac; ++q) {
    const int t = s.ao_idx[ab + q];
    const int dst = s.t_dst[t];
    const float m = s.t_m[t];
    const int4 e0 = make_int4(dst, s.t_e[t], s.slot[dst], 0);
    const int4 e1 = make_int4(__float_as_int(m), __float_as_int(m * s.rs[dst]), __float_as_int(m * s.ins[dst]), 0);
    s.ao_edges[2 * (ab + q)] = e0;
    s.ao_edges[2 * (ab + q) + 1] = e1;
    if (q == 0) { f0 = e0; f1 = e1; }
  }
  const int nb = s.no_off[i], nc = s.no_off[i + 1] - nb;
  // null transitions into END (summed if there are several): (mantissa, exponent) of their total weight
  double end_sum = 0;
  int end_e = 0;
  bool has_end = false;
  for (int q = 0; q < nc; ++q) {
    const int t = s.no_idx[nb + q];
    const int dst = s.t_dst[t];
    const bool toEnd = dst == s.n - 1;
    const float m = s.t_m[t];
    s.no_edges[2 * (nb + q)] = make_int4(dst, s.t_e[t], toEnd ? -1 : s.slot[dst], toEnd ? 1 : 0);
    s.no_edges[2 * (nb + q) + 1] = make_int4(__float_as_int(m), 0, 0, 0);
    if (toEnd && m > 0.f) {
      const int e = s.t_e[t];
      if (!has_end) { end_sum = m; end_e = e; has_end = true; }
      else if (e > end_e) { end_sum = ldexp(end_sum, end_e - e) + m; end_e = e; }
      else end_sum += ldexp((double)m, e - end_e);
    }
  }
  if (has_end) flags |= F_TO_END;
  s.fm_a[i] = make_int4(flags | (ic << 8), ib, s.key[i], s.slot[i]);
  s.fm_b[i] = first;
  s.emis[i] = make_float2(s.rs[i], s.ins[i]);
  s.bm[i] = make_int4(flags | (ac << 8), ab, s.key[i], s.slot[i]);
  s.bm[s.n + i] = make_int4(nb, nc, end_e, __float_as_int((float)end_sum));
  s.bm[2 * s.n + i] = f0;
  s.bm[3 * s.n + i] = f1;
  // records of the chain kernels (leaf x leaf merges): own emission factors, and those of the next state
  s.cf[i] = make_int4(s.ex[i], __float_as_int(s.rs[i]), __float_as_int(s.ins[i]), s.key[i]);
  const int nx = min(i + 1, s.n - 1);
  s.cb[i] = make_int4(s.ex[nx], __float_as_int(s.rs[nx]), __float_as_int(s.ins[nx]), s.key[i]);
}

// Exact in-envelope cell count (the GCUPS unit, forward.cpp:78-86), one thread per stored slot.
// grid = (chunks, pairs).
__global__ void count_cells_kernel(PairDev* const* __restrict__ pairs) {
  const PairDev& P = *pairs[blockIdx.y];
  __shared__ unsigned long long s_cnt;
  if (threadIdx.x == 0) s_cnt = 0;
  __syncthreads();
  unsigned long long local = 0;
  // one warp per anti-diagonal (banded strips hold a few dozen slots)
  const int warps = blockDim.x >> 5, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int d = blockIdx.x * warps + warp; d < P.n_diag; d += gridDim.x * warps) {
    const DiagInfo di = P.dinfo[d];
    const int nthr = P.nS + di.width + P.nE;
    for (int t = lane; t < nthr; t += 32) {
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


// ------------------------------------------------------------------ profile structure images
// Offsets of the arrays of a ProfImage inside its allocation: the same function places them on the host (images built
// from host lists) and in chain_image_kernel (leaf chains, generated on the device).
struct ImgOff {
  size_t flags, t_src, t_dst, t_lp, in_off, in_idx, ao_off, ao_idx, no_off, no_idx, leaf_pos, near_start, near_end, slot_start,
      slot_end, start_rows, end_cols, bytes;
};
__host__ __device__ inline size_t img_take(size_t& b, size_t bytes) {
  const size_t off = b;
  b = (b + bytes + 15) & ~(size_t)15;
  return off;
}
__host__ __device__ inline ImgOff image_offsets(int n, int nt) {
  ImgOff o;
  size_t b = 0;
  const size_t nt1 = nt > 0 ? nt : 1, ns = (size_t)n;
  o.t_lp = img_take(b, nt1 * sizeof(double));
  o.t_src = img_take(b, nt1 * sizeof(int));
  o.t_dst = img_take(b, nt1 * sizeof(int));
  o.in_off = img_take(b, (ns + 1) * sizeof(int));
  o.in_idx = img_take(b, nt1 * sizeof(int));
  o.ao_off = img_take(b, (ns + 1) * sizeof(int));
  o.ao_idx = img_take(b, nt1 * sizeof(int));
  o.no_off = img_take(b, (ns + 1) * sizeof(int));
  o.no_idx = img_take(b, nt1 * sizeof(int));
  o.leaf_pos = img_take(b, ns * sizeof(int));
  o.slot_start = img_take(b, ns * sizeof(int));
  o.slot_end = img_take(b, ns * sizeof(int));
  o.start_rows = img_take(b, ns * sizeof(int));
  o.end_cols = img_take(b, ns * sizeof(int));
  o.flags = img_take(b, ns);
  o.near_start = img_take(b, ns);
  o.near_end = img_take(b, ns);
  o.bytes = (b + 255) & ~(size_t)255;
  return o;
}
__host__ __device__ inline ProfImage image_at(char* base, const ImgOff& o) {
  ProfImage im;
  im.flags = (uint8_t*)(base + o.flags);
  im.t_src = (int*)(base + o.t_src); im.t_dst = (int*)(base + o.t_dst); im.t_lp = (double*)(base + o.t_lp);
  im.in_off = (int*)(base + o.in_off); im.in_idx = (int*)(base + o.in_idx);
  im.ao_off = (int*)(base + o.ao_off); im.ao_idx = (int*)(base + o.ao_idx);
  im.no_off = (int*)(base + o.no_off); im.no_idx = (int*)(base + o.no_idx);
  im.leaf_pos = (int*)(base + o.leaf_pos);
  im.near_start = (uint8_t*)(base + o.near_start); im.near_end = (uint8_t*)(base + o.near_end);
  im.slot_start = (int*)(base + o.slot_start); im.slot_end = (int*)(base + o.slot_end);
  im.start_rows = (int*)(base + o.start_rows); im.end_cols = (int*)(base + o.end_cols);
  return im;
}

// A leaf (Profile(components, alphabet, FastSeq, row), profile.cpp:23-76) is the chain
// START -> residue 1 -> ... -> residue L -> END with log-weight 0 transitions: transition t enters state t+1; it is an
// absorbing out-edge of state t unless it enters END (t == n-2), which makes it the one null out-edge.  Every list of
// the image is arithmetic on the index.  grid = (chunks, profiles); requires n >= 3.
struct ChainImgJob {
  long long off;
  int n, pad;
};
__global__ void chain_image_kernel(char* arena, const ChainImgJob* __restrict__ jobs) {
  const ChainImgJob J = jobs[blockIdx.y];
  const int n = J.n, nt = n - 1;
  const ProfImage im = image_at(arena + J.off, image_offsets(n, nt));
  for (int k = blockIdx.x * blockDim.x + threadIdx.x; k <= n; k += gridDim.x * blockDim.x) {
    im.in_off[k] = max(0, k - 1);
    im.ao_off[k] = min(k, n - 2);
    im.no_off[k] = k > n - 2 ? 1 : 0;
    if (k < n) {
      const bool emit = k >= 1 && k <= n - 2;
      int f = 0;
      if (emit) f |= F_EMIT;
      if (k != n - 2) f |= F_READY;               // state n-2 has the null transition into END
      if (emit || k == 0) f |= F_EMIT_OR_START;
      im.flags[k] = (uint8_t)f;
      im.leaf_pos[k] = min(k, n - 2);             // seqCoords of the leaf row (profile.cpp:33-36, 49)
      im.near_start[k] = k == 0;
      im.near_end[k] = k == n - 2;
      im.slot_start[k] = k == 0 ? 0 : -1;
      im.slot_end[k] = k == n - 2 ? 0 : -1;
    }
    if (k < nt) {
      im.t_src[k] = k;
      im.t_dst[k] = k + 1;
      im.t_lp[k] = 0.0;
      im.in_idx[k] = k;
    }
    if (k < n - 2) im.ao_idx[k] = k;
    if (k == 0) {
      im.no_idx[0] = n - 2;
      im.start_rows[0] = 0;
      im.end_cols[0] = n - 2;
    }
  }
}

// ------------------------------------------------------------------ band keys and the anti-diagonal layout
// Inclusive scan of v[0..n) in place by the whole CTA (forward, or from the end when REVERSE): every thread folds a
// contiguous chunk, the chunk totals are combined through shared memory, the chunk is folded again from its prefix.
struct OpMax { __device__ int operator()(int a, int b) const { return max(a, b); } };
struct OpMin { __device__ int operator()(int a, int b) const { return min(a, b); } };
struct OpAdd { __device__ int operator()(int a, int b) const { return a + b; } };
template <bool REVERSE, class Op>
__device__ void block_scan(int* v, int n, int identity, Op op, int* sm) {
  const int T = blockDim.x, t = threadIdx.x;
  const int chunk = (n + T - 1) / T, b = min(n, t * chunk), e = min(n, b + chunk);
  __syncthreads();   // v was written by other threads; sm may still be read by the previous scan
  int acc = identity;
  for (int q = b; q < e; ++q) acc = op(acc, v[REVERSE ? n - 1 - q : q]);
  sm[t] = acc;
  __syncthreads();
  acc = identity;
  for (int q = 0; q < t; ++q) acc = op(acc, sm[q]);
  for (int q = b; q < e; ++q) {
    const int idx = REVERSE ? n - 1 - q : q;
    acc = op(acc, v[idx]);
    v[idx] = acc;
  }
  __syncthreads();
}

// One CTA per pair.  Band keys (forward.cpp:36-42, alignpath.h:56-61), the bounding column interval of every row, its
// monotone hull (so that every anti-diagonal meets the band in ONE interval of i), and the per-diagonal layout records
// with their plane offsets; integer work, bit-exact with the reference's inEnvelope predicate.  The totals go to
// P.layout_out: the host sizes the cell planes from them.
__global__ void __launch_bounds__(256) pair_layout_kernel(PairDev* const* __restrict__ pairs) {
  PairDev& P = *pairs[blockIdx.x];
  __shared__ int sm[256];
  __shared__ unsigned long long smu[256];
  __shared__ int s_err;
  const int T = blockDim.x, t = threadIdx.x;
  const int nx = P.x.n, ny = P.y.n, NI = nx - 1, NJ = ny - 1, md = P.max_dist;
  int* xkey = P.x.key;
  int* ykey = P.y.key;
  int* jlo = P.scratch;
  int* jhi = jlo + NI;
  int* rows = jhi + NI;
  int* cnt = rows + NI;
  int* pmax = cnt + NI;
  int* smin = pmax + NJ;
  LayoutOut* out = P.layout_out;
  if (t == 0) s_err = 0;
  __syncthreads();
  for (int i = t; i < nx; i += T) {
    int key = 0;
    if (md >= 0) {
      const int pos = i == 0 ? 0 : P.x_leaf_pos[i];
      if (pos < 0 || pos > P.env_len1) {
        if (atomicCAS(&s_err, 0, 1) == 0) { out->err_state = i; out->err_pos = pos; }
      } else {
        key = P.env_k1[pos];
      }
    }
    xkey[i] = key;
  }
  for (int j = t; j < ny; j += T) {
    int key = 0;
    if (md >= 0) {
      const int pos = j == 0 ? 0 : P.y_leaf_pos[j];
      if (pos < 0 || pos > P.env_len2) {
        if (atomicCAS(&s_err, 0, 2) == 0) { out->err_state = j; out->err_pos = pos; }
      } else {
        key = P.env_k2[pos];
      }
    }
    ykey[j] = key;
  }
  __syncthreads();
  if (s_err) {
    if (t == 0) { out->err = s_err; out->n_stored = 0; out->max_threads = 1; out->max_width = 0; }
    return;
  }
  if (md < 0) {
    for (int i = t; i < NI; i += T) { jlo[i] = 0; jhi[i] = NJ - 1; }
  } else {
    // prefix max / suffix min of the y keys make "first j with key >= a" / "last j with key <= b" binary-searchable
    for (int j = t; j < NJ; j += T) { pmax[j] = ykey[j]; smin[j] = ykey[j]; }
    block_scan<false>(pmax, NJ, INT_MIN, OpMax(), sm);
    block_scan<true>(smin, NJ, INT_MAX, OpMin(), sm);
    for (int i = t; i < NI; i += T) {
      const int a = xkey[i] - md, b = xkey[i] + md;
      int lo = 0, hi = NJ;   // first j: pmax[j] >= a
      while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (pmax[mid] < a) lo = mid + 1; else hi = mid;
      }
      jlo[i] = lo;
      lo = 0; hi = NJ;       // first j: smin[j] > b
      while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        if (smin[mid] <= b) lo = mid + 1; else hi = mid;
      }
      jhi[i] = lo - 1;
    }
  }
  __syncthreads();
  // monotone hull over the non-empty rows: jhi becomes its running maximum, jlo its running minimum from the end
  for (int i = t; i < NI; i += T) rows[i] = jlo[i] <= jhi[i] ? jhi[i] : -1;
  block_scan<false>(rows, NI, -1, OpMax(), sm);
  for (int i = t; i < NI; i += T)
    if (jlo[i] <= jhi[i]) jhi[i] = rows[i];
  __syncthreads();
  for (int i = t; i < NI; i += T) rows[i] = jlo[i] <= jhi[i] ? jlo[i] : NJ;
  block_scan<true>(rows, NI, NJ, OpMin(), sm);
  for (int i = t; i < NI; i += T)
    if (jlo[i] <= jhi[i]) jlo[i] = rows[i];
  __syncthreads();
  // the non-empty rows, ascending
  for (int i = t; i < NI; i += T) cnt[i] = jlo[i] <= jhi[i] ? 1 : 0;
  block_scan<false>(cnt, NI, 0, OpAdd(), sm);
  const int nrows = cnt[NI - 1];
  for (int i = t; i < NI; i += T)
    if (jlo[i] <= jhi[i]) rows[cnt[i] - 1] = i;
  __syncthreads();
  // Over the non-empty rows i + jlo[i] and i + jhi[i] are strictly increasing, so the rows meeting diagonal d form a
  // contiguous run of `rows`, found by two binary searches.
  const int nDiag = NI + NJ - 1, nS = P.nS, nE = P.nE;
  DiagInfo* dinfo = P.dinfo;
  int maxW = 0;
  for (int d = t; d < nDiag; d += T) {
    int lo = 0, hi = nrows;   // first r: rows[r] + jhi[rows[r]] >= d
    while (lo < hi) {
      const int mid = (lo + hi) >> 1, r = rows[mid];
      if (r + jhi[r] < d) lo = mid + 1; else hi = mid;
    }
    const int first = lo;
    lo = 0; hi = nrows;       // first r: rows[r] + jlo[rows[r]] > d
    while (lo < hi) {
      const int mid = (lo + hi) >> 1, r = rows[mid];
      if (r + jlo[r] <= d) lo = mid + 1; else hi = mid;
    }
    DiagInfo di;
    di.ilo = 0; di.width = 0; di.off = 0; di.pad = 0;
    if (first < lo) { di.ilo = rows[first]; di.width = rows[lo - 1] - rows[first] + 1; }
    dinfo[d] = di;
    maxW = max(maxW, di.width);
  }
  __syncthreads();
  // plane offsets: exclusive sum of nS + width + nE over the diagonals
  {
    const int chunk = (nDiag + T - 1) / T, b = min(nDiag, t * chunk), e = min(nDiag, b + chunk);
    unsigned long long acc = 0;
    for (int d = b; d < e; ++d) acc += (unsigned long long)(nS + dinfo[d].width + nE);
    smu[t] = acc;
    sm[t] = maxW;
    __syncthreads();
    acc = 0;
    for (int q = 0; q < t; ++q) acc += smu[q];
    for (int d = b; d < e; ++d) {
      dinfo[d].off = (unsigned)acc;
      acc += (unsigned long long)(nS + dinfo[d].width + nE);
    }
    if (t == 0) {
      unsigned long long total = 0;
      int w = 0;
      for (int q = 0; q < T; ++q) { total += smu[q]; w = max(w, sm[q]); }
      out->n_stored = total;
      out->max_width = w;
      out->max_threads = max(1, nS + w + nE);
      out->err = 0;
      P.n_stored = (unsigned)total;
      P.max_threads = max(1, nS + w + nE);
    }
  }
}

// The cell planes are allocated once the layout totals are back on the host: patch their addresses into the descriptors.
struct PlanePlace {
  float4* fA; float2* fB; float* absorb; float4* bA; float2* bB;
};
__global__ void place_planes_kernel(PairDev* const* __restrict__ pairs, const PlanePlace* __restrict__ pl, int n) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  PairDev& P = *pairs[k];
  P.fA = pl[k].fA; P.fB = pl[k].fB; P.absorb = pl[k].absorb; P.bA = pl[k].bA; P.bB = pl[k].bB;
}

}  // namespace hb

using namespace hb;

namespace {
// Persistent host worker threads: a level of merges goes through half a dozen parallel loops per call, and spawning
// sixteen threads for each costs more than some of the loops themselves.
class HostPool {
 public:
  static HostPool& get() {
    static HostPool pool;
    return pool;
  }
  // fn(k) for k in [0, n) on `nthreads` threads (the caller is one of them); items are handed out dynamically
  void run(int n, int nthreads, const std::function<void(int)>& fn) {
    std::lock_guard<std::mutex> serial(run_mu_);
    grow(nthreads - 1);
    {
      std::lock_guard<std::mutex> lk(mu_);
      job_ = &fn;
      n_ = n;
      next_.store(0);
      want_ = nthreads - 1;
      active_ = want_;
      ++gen_;
    }
    cv_.notify_all();
    work(fn, n);
    std::unique_lock<std::mutex> lk(mu_);
    done_.wait(lk, [&] { return active_ == 0; });
    job_ = nullptr;
  }

 private:
  HostPool() {}
  ~HostPool() {
    {
      std::lock_guard<std::mutex> lk(mu_);
      stop_ = true;
    }
    cv_.notify_all();
    for (auto& t : threads_) t.join();
  }
  void work(const std::function<void(int)>& fn, int n) {
    for (int k = next_.fetch_add(1); k < n; k = next_.fetch_add(1)) fn(k);
  }
  void grow(int workers) {
    while ((int)threads_.size() < workers) {
      const int id = (int)threads_.size();
      threads_.emplace_back([this, id] {
        unsigned long long seen = 0;
        for (;;) {
          const std::function<void(int)>* job = nullptr;
          int n = 0;
          {
            std::unique_lock<std::mutex> lk(mu_);
            cv_.wait(lk, [&] { return stop_ || (gen_ != seen && id < want_); });
            if (stop_) return;
            seen = gen_;
            job = job_;
            n = n_;
          }
          work(*job, n);
          {
            std::lock_guard<std::mutex> lk(mu_);
            --active_;
          }
          done_.notify_one();
        }
      });
    }
  }
  std::vector<std::thread> threads_;
  std::mutex mu_, run_mu_;
  std::condition_variable cv_, done_;
  const std::function<void(int)>* job_ = nullptr;
  std::atomic<int> next_{0};
  int n_ = 0, want_ = 0, active_ = 0;
  unsigned long long gen_ = 0;
  bool stop_ = false;
};

// Runs fn(0..n-1) on up to hardware_concurrency host threads (at least `grain` items per thread).
template <class F>
void parallel_for(int n, int grain, F fn) {
  int nthreads = (int)std::min<unsigned>(std::max(1u, std::thread::hardware_concurrency()), 16u);
  if (const char* lw = getenv("LOCAL_WORLD_SIZE"))  // one process per GPU: share the host cores between the ranks of the box
    nthreads = std::max(2, (int)std::thread::hardware_concurrency() / std::max(1, atoi(lw)));
  nthreads = std::min(nthreads, 16);
  nthreads = std::max(1, std::min(nthreads, n / std::max(1, grain)));
  if (nthreads <= 1) {
    for (int k = 0; k < n; ++k) fn(k);
    return;
  }
  HostPool::get().run(n, nthreads, std::function<void(int)>(fn));
}

// HIST_B200_HOST_TIMING=1: wall-clock split of the host side of the batched constructors on stderr
struct HostTimer {
  bool on;
  const char* what;
  std::chrono::steady_clock::time_point t0;
  explicit HostTimer(const char* w) : on(getenv("HIST_B200_HOST_TIMING") != nullptr), what(w), t0(std::chrono::steady_clock::now()) {}
  void lap(const char* stage) {
    if (!on) return;
    const auto t1 = std::chrono::steady_clock::now();
    fprintf(stderr, "[%s] %-18s %.2f ms\n", what, stage, std::chrono::duration<double, std::milli>(t1 - t0).count());
    t0 = t1;
  }
};
}  // namespace

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
  ctx->fill_smem_configured = 0;
  ctx->absorb_smem_configured = 0;
  ctx->staging = nullptr;
  ctx->staging_bytes = 0;
  ctx->staging2 = nullptr;
  ctx->staging2_bytes = 0;
  cudaEventCreateWithFlags(&ctx->ev_stage[0], cudaEventDisableTiming);
  cudaEventCreateWithFlags(&ctx->ev_stage[1], cudaEventDisableTiming);
  ctx->staging_img = nullptr;
  ctx->staging_img_bytes = 0;
  cudaEventCreateWithFlags(&ctx->ev_img, cudaEventDisableTiming);
  ctx->h2d_bytes = 0;
  ctx->stream2 = nullptr;
  ctx->ev_fork = ctx->ev_join = nullptr;
  e = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking);
  if (e != cudaSuccess) {
    set_err(nullptr, HIST_ERR_CUDA, "cudaStreamCreate: %s", cudaGetErrorString(e));
    delete ctx;
    return HIST_ERR_CUDA;
  }
  {
    // A private stream-ordered pool (the device's default pool, which a host application may share with torch or its own
    // cudaMallocAsync calls, is left alone).  Freed arenas stay in it: every level of a reconstruction allocates a multi-GB
    // arena, and the default release threshold (0) would hand it back to the driver at each synchronisation.
    cudaMemPoolProps props = {};
    props.allocType = cudaMemAllocationTypePinned;
    props.handleTypes = cudaMemHandleTypeNone;
    props.location.type = cudaMemLocationTypeDevice;
    props.location.id = device;
    e = cudaMemPoolCreate(&ctx->pool, &props);
    if (e != cudaSuccess) {
      set_err(nullptr, HIST_ERR_CUDA, "cudaMemPoolCreate: %s", cudaGetErrorString(e));
      cudaStreamDestroy(ctx->stream);
      delete ctx;
      return HIST_ERR_CUDA;
    }
    unsigned long long keep = ~0ull;
    cudaMemPoolSetAttribute(ctx->pool, cudaMemPoolAttrReleaseThreshold, &keep);
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
  if (ctx->staging) cudaFreeHost(ctx->staging);
  if (ctx->staging2) cudaFreeHost(ctx->staging2);
  if (ctx->staging_img) cudaFreeHost(ctx->staging_img);
  for (auto& e : ctx->onehot) cudaFree(e.second);
  cudaEventDestroy(ctx->ev_img);
  cudaEventDestroy(ctx->ev_stage[0]);
  cudaEventDestroy(ctx->ev_stage[1]);
  if (ctx->stream2) { cudaStreamSynchronize(ctx->stream2); cudaStreamDestroy(ctx->stream2); }
  if (ctx->ev_fork) cudaEventDestroy(ctx->ev_fork);
  if (ctx->ev_join) cudaEventDestroy(ctx->ev_join);
  cudaStreamDestroy(ctx->stream);
  cudaMemPoolDestroy(ctx->pool);
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
extern "C" int64_t hist_ctx_h2d_bytes(const hist_ctx* ctx) { return ctx ? ctx->h2d_bytes : 0; }

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
  p->sorted_in = true;
  {
    std::vector<int> last(p->n, -1);
    for (int t = 0; t < p->nt; ++t) {
      if (p->t_src[t] <= last[p->t_dst[t]]) p->sorted_in = false;
      last[p->t_dst[t]] = p->t_src[t];
    }
  }
  p->empty = true;
  for (int i = 0; i < p->n; ++i)
    if (p->is_emit[i]) p->empty = false;
  // a leaf chain (Profile(components, alphabet, FastSeq, row), profile.cpp:23-76) uploaded by hand gets the same fast paths
  // as one built by hist_profile_from_sequences
  p->chain = p->nt == p->n - 1;
  for (int t = 0; t < p->nt && p->chain; ++t) p->chain = p->t_src[t] == t && p->t_dst[t] == t + 1 && p->t_lp[t] == 0.0;
  for (int i = 1; i + 1 < p->n && p->chain; ++i) p->chain = p->is_emit[i] != 0;
  p->leaf_std = p->chain;
  for (size_t i = 0; i < p->leaf_pos.size() && p->leaf_std; ++i) p->leaf_std = p->leaf_pos[i] == std::min((int)i, p->n - 2);
  const size_t bytes = (size_t)p->n * p->C * p->A * sizeof(double);
  p->d_lp_absorb = nullptr;
  p->arena = nullptr;
  int rc = check_cuda(ctx, cudaMallocFromPoolAsync((void**)&p->d_lp_absorb, bytes, ctx->pool, ctx->stream), "cudaMallocAsync(lp_absorb)");
  if (!rc) rc = check_cuda(ctx, cudaMemcpyAsync(p->d_lp_absorb, d->lp_absorb, bytes, cudaMemcpyHostToDevice, ctx->stream), "H2D lp_absorb");
  ctx->h2d_bytes += (long long)bytes;
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

namespace {
int ensure_staging(hist_ctx* ctx, size_t bytes, int slot = 0);
void arena_release(hist_ctx* ctx, hb_arena* a);
}  // namespace

namespace hb {
// lpAbsorb of leaf profiles (profile.cpp:52-66): state 0 and END are null (rows unused, kept -inf), state pos+1
// is 0 at the residue's token and -inf elsewhere, or all 0 for a wildcard / invalid character.
// grid = (chunks, sequences): a block row walks the [states][C][A] elements of ITS sequence in order (coalesced stores,
// 32-bit index arithmetic, no search for the owning sequence).
__global__ void leaf_absorb_kernel(const int8_t* tokens, const long long* state_off, int n_seq, int C, int A, double* lp) {
  const int K = C * A;
  for (int s = blockIdx.y; s < n_seq; s += gridDim.y) {
    const long long s0 = state_off[s];
    const long long n_states = state_off[s + 1] - s0, total = n_states * K;
    const int8_t* tk = tokens + (s0 - 2LL * s);   // token of local state i (1 .. n_states - 2) is tk[i - 1]: minus the START/END rows before it
    double* out = lp + s0 * K;
    if (total < (1LL << 31)) {
      const unsigned tot = (unsigned)total, uK = (unsigned)K, uA = (unsigned)A, last = (unsigned)n_states - 1;
      for (unsigned q = blockIdx.x * blockDim.x + threadIdx.x; q < tot; q += gridDim.x * blockDim.x) {
        const unsigned i = q / uK, a = (q - i * uK) % uA;
        double v = -INFINITY;
        if (i > 0 && i < last) {
          const int tok = tk[i - 1];
          v = (tok < 0 || tok == (int)a) ? 0.0 : -INFINITY;
        }
        out[q] = v;
      }
    } else {
      for (long long q = blockIdx.x * (long long)blockDim.x + threadIdx.x; q < total; q += (long long)gridDim.x * blockDim.x) {
        const long long i = q / K;
        const int a = (int)((q - i * K) % A);
        double v = -INFINITY;
        if (i > 0 && i < n_states - 1) {
          const int tok = tk[i - 1];
          v = (tok < 0 || tok == a) ? 0.0 : -INFINITY;
        }
        out[q] = v;
      }
    }
  }
}
}  // namespace hb

// lut == nullptr: `tokens` are token strings; else they are residue characters mapped through lut[256]
static int leaf_profiles(hist_ctx* ctx, int32_t n, const int8_t* const* tokens, const int32_t* lengths, int32_t alph_size,
                         int32_t components, const int8_t* lut, hist_profile** out) {
  if (!ctx || n <= 0 || !tokens || !lengths || !out) return set_err(ctx, HIST_ERR_ARG, "hist_profile_from_sequences: bad arguments");
  if (alph_size < 1 || alph_size > 127 || components < 1) return set_err(ctx, HIST_ERR_ARG, "hist_profile_from_sequences: bad A/C");
  for (int k = 0; k < n; ++k) out[k] = nullptr;
  cudaSetDevice(ctx->device);
  std::vector<long long> state_off(n + 1, 0);
  long long n_tok = 0;
  for (int k = 0; k < n; ++k) {
    if (lengths[k] < 0 || (lengths[k] > 0 && !tokens[k])) return set_err(ctx, HIST_ERR_ARG, "hist_profile_from_sequences: sequence %d", k);
    state_off[k + 1] = state_off[k] + lengths[k] + 2;
    n_tok += lengths[k];
  }
  const size_t K = (size_t)alph_size * components;
  const size_t lp_bytes = (size_t)state_off[n] * K * sizeof(double), off_bytes = (n + 1) * sizeof(long long);
  const size_t tok_off = (lp_bytes + 255) & ~(size_t)255, so_off = (tok_off + n_tok + 255) & ~(size_t)255, total = so_off + off_bytes;
  hb_arena* arena = new hb_arena{nullptr, total, 0};
  int rc = check_cuda(ctx, cudaMallocFromPoolAsync((void**)&arena->dev, total, ctx->pool, ctx->stream), "cudaMallocAsync(leaf profiles)");
  const size_t so_st = (n_tok + 255) & ~(size_t)255;
  if (!rc) rc = ensure_staging(ctx, so_st + off_bytes);
  if (!rc) rc = check_cuda(ctx, cudaEventSynchronize(ctx->ev_stage[0]), "wait for the staging buffer");   // its previous transfer
  if (rc) {
    if (arena->dev) cudaFreeAsync(arena->dev, ctx->stream);
    delete arena;
    return rc;
  }
  {
    char* st = ctx->staging;
    parallel_for(n, 64, [&](int k) {
      char* dst = st + (state_off[k] - 2 * (long long)k);
      const int8_t* src = tokens[k];
      if (lut)
        for (int q = 0; q < lengths[k]; ++q) dst[q] = lut[(uint8_t)src[q]];
      else
        for (int q = 0; q < lengths[k]; ++q) dst[q] = src[q] >= alph_size ? -1 : src[q];
    });
    memcpy(ctx->staging + so_st, state_off.data(), off_bytes);
    rc = check_cuda(ctx, cudaMemcpyAsync(arena->dev + tok_off, ctx->staging, n_tok, cudaMemcpyHostToDevice, ctx->stream), "H2D tokens");
    if (!rc) rc = check_cuda(ctx, cudaMemcpyAsync(arena->dev + so_off, ctx->staging + so_st, off_bytes, cudaMemcpyHostToDevice, ctx->stream), "H2D offsets");
    ctx->h2d_bytes += (long long)(n_tok + off_bytes);
    if (!rc) rc = check_cuda(ctx, cudaEventRecord(ctx->ev_stage[0], ctx->stream), "record staging event");
  }
  if (!rc) {
    long long longest = 1;
    for (int k = 0; k < n; ++k) longest = std::max(longest, state_off[k + 1] - state_off[k]);
    const int chunks = (int)std::max<long long>(1, std::min<long long>((longest * K + 1023) / 1024, 64));   // ~4 elements per thread
    hb::leaf_absorb_kernel<<<dim3(chunks, std::min(n, 65535)), 256, 0, ctx->stream>>>((const int8_t*)(arena->dev + tok_off), (const long long*)(arena->dev + so_off), n,
                                                           components, alph_size, (double*)arena->dev);
    ctx->launches += 1;
    rc = check_cuda(ctx, cudaGetLastError(), "leaf_absorb_kernel");
  }
  // the borrowed token strings have been copied into the staging buffer (guarded by its event): nothing to wait for;
  // the profiles' tables are complete for whatever is enqueued next on the context's stream
  if (rc) {
    cudaStreamSynchronize(ctx->stream);
    cudaFreeAsync(arena->dev, ctx->stream);
    delete arena;
    return rc;
  }
  HostTimer tm("from_sequences");
  parallel_for(n, 16, [&](int k) {
    hist_profile* p = new hist_profile();
    const int L = lengths[k];
    p->ctx = ctx;
    p->n = L + 2;
    p->nt = L + 1;
    p->A = alph_size;
    p->C = components;
    p->is_emit.assign(p->n, 1);
    p->is_emit.front() = p->is_emit.back() = 0;
    p->t_src.resize(p->nt);
    p->t_dst.resize(p->nt);
    p->t_lp.assign(p->nt, 0.0);
    p->leaf_pos.resize(p->n);
    for (int t = 0; t < p->nt; ++t) { p->t_src[t] = t; p->t_dst[t] = t + 1; }
    for (int i = 0; i < p->n; ++i) p->leaf_pos[i] = std::min(i, L);   // seqCoords of the leaf row (profile.cpp:33-36, 49)
    p->sorted_in = true;
    p->chain = true;
    p->leaf_std = true;
    p->empty = L == 0;
    p->d_lp_absorb = (double*)arena->dev + (size_t)state_off[k] * K;
    p->d_tokens = (const int8_t*)(arena->dev + tok_off) + (state_off[k] - 2 * (long long)k);
    p->arena = arena;
    out[k] = p;
  });
  arena->refs += n;
  tm.lap("host objects");
  return HIST_OK;
}

extern "C" int hist_profile_from_sequences(hist_ctx* ctx, int32_t n, const int8_t* const* tokens, const int32_t* lengths,
                                           int32_t alph_size, int32_t components, hist_profile** out) {
  return leaf_profiles(ctx, n, tokens, lengths, alph_size, components, nullptr, out);
}

extern "C" int hist_profile_from_residues(hist_ctx* ctx, int32_t n, const char* const* residues, const int32_t* lengths,
                                          const char* alphabet, int32_t components, hist_profile** out) {
  if (!alphabet) return set_err(ctx, HIST_ERR_ARG, "hist_profile_from_residues: no alphabet");
  const size_t A = strlen(alphabet);
  if (A < 1 || A > 127) return set_err(ctx, HIST_ERR_ARG, "hist_profile_from_residues: bad alphabet size");
  // tokenize() (fastseq.cpp:10-16): the exact character first, then the other case; anything else is a wildcard
  int8_t lut[256];
  for (int c = 0; c < 256; ++c) lut[c] = -1;
  for (size_t k = 0; k < A; ++k) {
    const unsigned char ch = (unsigned char)alphabet[k];
    const int other = isupper(ch) ? tolower(ch) : toupper(ch);
    if (lut[other] < 0) lut[other] = (int8_t)k;
  }
  for (size_t k = A; k-- > 0;) lut[(unsigned char)alphabet[k]] = (int8_t)k;   // first exact occurrence wins
  return leaf_profiles(ctx, n, (const int8_t* const*)residues, lengths, (int32_t)A, components, lut, out);
}

extern "C" void hist_profile_destroy(hist_profile* p) {
  if (!p) return;
  if (p->refs > 0) {   // pairs built on this profile still read its host lists and device table: released with the last of them
    p->doomed = true;
    return;
  }
  cudaSetDevice(p->ctx->device);
  if (p->arena) arena_release(p->ctx, p->arena);
  else if (p->d_lp_absorb) cudaFreeAsync(p->d_lp_absorb, p->ctx->stream);
  if (p->img_arena) arena_release(p->ctx, p->img_arena);
  delete p;
}

extern "C" void hist_profiles_destroy(hist_profile* const* profiles, int32_t n) {
  if (!profiles || n <= 0) return;
  std::vector<hist_profile*> gone;
  gone.reserve(n);
  for (int k = 0; k < n; ++k) {
    hist_profile* p = profiles[k];
    if (!p) continue;
    if (p->refs > 0) {
      p->doomed = true;
      continue;
    }
    cudaSetDevice(p->ctx->device);
    if (p->arena) arena_release(p->ctx, p->arena);
    else if (p->d_lp_absorb) cudaFreeAsync(p->d_lp_absorb, p->ctx->stream);
    if (p->img_arena) arena_release(p->ctx, p->img_arena);
    gone.push_back(p);
  }
  parallel_for((int)gone.size(), 64, [&](int k) { delete gone[k]; });
}

// ------------------------------------------------------------------ pair preparation
namespace {

// Byte layout of one pair inside the batch arena: a host-initialised part (copied in ONE H2D
// transfer for the whole batch) and a device-only part.
struct Layout {
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
    if (n) copies.push_back({off, {src, n}});
    return off;
  }
};

struct SideHost {
  std::vector<uint8_t> flags;
  std::vector<int> in_off, in_idx, ao_off, ao_idx, no_off, no_idx;
};

struct SideOff {  // offsets in the device-only part of the pair arena
  bool tok = false;   // leaf side: token-class rows
  size_t k_sub = 0, k_H = 0, k_ex = 0, k_rs = 0, k_ins = 0, k_rootsub = 0, k_insl = 0;
  size_t d_key, d_t_m, d_t_e, d_ex, d_rs, d_ins, d_H, d_sub, d_rootsub, d_insl, d_fm_a, d_fm_b, d_emis, d_in_edges, d_bm, d_ao_edges, d_no_edges, d_cf, d_cb;
};

struct PairBuild {
  hist_pair* pr = nullptr;
  Layout host, dev;
  SideOff ox, oy;
  size_t h_k1 = 0, h_k2 = 0, h_lsub = 0, h_rsub = 0, h_root = 0, h_lins = 0, h_rins = 0, h_lcw = 0, h_rcw = 0, h_pairdev = 0;
  size_t d_dinfo = 0, d_scratch = 0;
  size_t host_base = 0, dev_base = 0;  // position of this pair's two parts in the arena
  std::vector<int> k1, k2;             // cumulativeMatches[rowPosToCol[pos]] of the two guide rows
  size_t plane_base = 0;               // position of this pair's cell planes in the plane arena
  size_t p_fA = 0, p_fB = 0, p_absorb = 0, p_bA = 0, p_bB = 0;
};

// The transition lists of a profile (the reference's ProfileState::in / absorbOut / nullOut, profile.h:13-40)
// and its state flags, from the host transition arrays.
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

// Writes the structure image of a profile (host lists) into `base`, an image-sized block of host memory laid out by
// image_offsets; returns the edge counts.  The edge flags are the reference's xNearStart / yNearEnd (forward.cpp:58-65).
int build_image_host(hist_ctx* ctx, const hist_profile* p, char* base, int* nS_out, int* nE_out) {
  const int n = p->n, nt = p->nt;
  SideHost s;
  build_side_lists(p, s);
  for (int i = 0; i < n; ++i)
    if (s.in_off[i + 1] - s.in_off[i] > HB_MAX_DEGREE || s.ao_off[i + 1] - s.ao_off[i] > HB_MAX_DEGREE)
      return set_err(ctx, HIST_ERR_ARG, "hist_pair_prepare: state %d has more than %d transitions", i, HB_MAX_DEGREE);
  const ImgOff off = image_offsets(n, nt);
  memset(base, 0, off.bytes);
  const ProfImage im = image_at(base, off);
  memcpy(im.flags, s.flags.data(), n);
  if (nt) {
    memcpy(im.t_src, p->t_src.data(), (size_t)nt * sizeof(int));
    memcpy(im.t_dst, p->t_dst.data(), (size_t)nt * sizeof(int));
    memcpy(im.t_lp, p->t_lp.data(), (size_t)nt * sizeof(double));
  }
  memcpy(im.in_off, s.in_off.data(), (size_t)(n + 1) * sizeof(int));
  memcpy(im.ao_off, s.ao_off.data(), (size_t)(n + 1) * sizeof(int));
  memcpy(im.no_off, s.no_off.data(), (size_t)(n + 1) * sizeof(int));
  memcpy(im.in_idx, s.in_idx.data(), (size_t)s.in_off[n] * sizeof(int));
  memcpy(im.ao_idx, s.ao_idx.data(), (size_t)s.ao_off[n] * sizeof(int));
  memcpy(im.no_idx, s.no_idx.data(), (size_t)s.no_off[n] * sizeof(int));
  if (!p->leaf_pos.empty()) memcpy(im.leaf_pos, p->leaf_pos.data(), (size_t)n * sizeof(int));
  im.near_start[0] = 1;
  for (int i = 0; i < n; ++i)
    if (im.near_start[i])
      for (int q = s.no_off[i]; q < s.no_off[i + 1]; ++q) im.near_start[p->t_dst[s.no_idx[q]]] = 1;
  for (int q = s.in_off[n - 1]; q < s.in_off[n]; ++q) im.near_end[p->t_src[s.in_idx[q]]] = 1;
  int nS = 0, nE = 0;
  for (int i = 0; i < n; ++i) {   // rows / columns 0 .. n-2 hold cells
    im.slot_start[i] = -1;
    im.slot_end[i] = -1;
    if (i < n - 1 && im.near_start[i]) { im.slot_start[i] = nS; im.start_rows[nS++] = i; }
    if (i < n - 1 && im.near_end[i]) { im.slot_end[i] = nE; im.end_cols[nE++] = i; }
  }
  *nS_out = nS;
  *nE_out = nE;
  return HIST_OK;
}

int ensure_pinned(hist_ctx* ctx, char*& buf, size_t& have, size_t bytes) {
  if (have >= bytes) return HIST_OK;
  if (buf) {
    HB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    cudaFreeHost(buf);
    buf = nullptr;
    have = 0;
  }
  const size_t want = std::max(bytes + bytes / 4, (size_t)1 << 20);
  HB_CUDA(ctx, cudaMallocHost((void**)&buf, want));
  have = want;
  return HIST_OK;
}

void arena_release(hist_ctx* ctx, hb_arena* a);

// Structure images for every profile of the list that has none yet: one allocation, one transfer (the images built from
// host lists, and the job records of the leaf chains, whose images chain_image_kernel then generates on the device).
int ensure_images(hist_ctx* ctx, const hist_profile* const* xs, const hist_profile* const* ys, int n) {
  std::vector<hist_profile*> need;
  for (int k = 0; k < 2 * n; ++k) {
    hist_profile* p = const_cast<hist_profile*>(k < n ? xs[k] : ys[k - n]);
    if (p && p->ctx == ctx && !p->img_arena) need.push_back(p);
  }
  if (need.empty()) return HIST_OK;
  std::sort(need.begin(), need.end());
  need.erase(std::unique(need.begin(), need.end()), need.end());
  // a leaf chain whose closest-leaf positions are the leaf's own (or absent) gets its image from chain_image_kernel
  auto on_device = [](const hist_profile* p) { return p->chain && p->leaf_std && p->n >= 3; };
  const int m = (int)need.size();
  std::vector<size_t> base(m + 1, 0);
  const int n_host = (int)(std::stable_partition(need.begin(), need.end(), [&](hist_profile* p) { return !on_device(p); }) - need.begin());
  int max_chain_n = 0;
  for (int k = 0; k < m; ++k) {
    const hist_profile* p = need[k];
    base[k + 1] = base[k] + image_offsets(p->n, p->nt).bytes;
    if (k >= n_host) max_chain_n = std::max(max_chain_n, p->n);
  }
  const int n_chain = m - n_host;
  const size_t host_bytes = base[n_host], jobs_off = base[m], jobs_bytes = (size_t)n_chain * sizeof(ChainImgJob);
  const size_t total = jobs_off + (jobs_bytes + 255) / 256 * 256 + 256;
  hb_arena* arena = new hb_arena{nullptr, total, 0};
  int rc = check_cuda(ctx, cudaMallocFromPoolAsync((void**)&arena->dev, total, ctx->pool, ctx->stream), "cudaMallocAsync(profile images)");
  if (!rc) rc = ensure_pinned(ctx, ctx->staging_img, ctx->staging_img_bytes, host_bytes + jobs_bytes + 256);
  if (!rc) rc = check_cuda(ctx, cudaEventSynchronize(ctx->ev_img), "wait for the image staging buffer");
  std::vector<int> nS(m, 1), nE(m, 1);
  if (!rc) {
    std::vector<int> rcs(std::max(1, n_host), 0);
    char* st = ctx->staging_img;
    parallel_for(n_host, 8, [&](int k) { rcs[k] = build_image_host(ctx, need[k], st + base[k], &nS[k], &nE[k]); });
    for (int k = 0; k < n_host && !rc; ++k) rc = rcs[k];
    ChainImgJob* jobs = (ChainImgJob*)(st + host_bytes);
    for (int k = 0; k < n_chain; ++k) jobs[k] = ChainImgJob{(long long)base[n_host + k], need[n_host + k]->n, 0};
  }
  if (!rc && host_bytes) rc = check_cuda(ctx, cudaMemcpyAsync(arena->dev, ctx->staging_img, host_bytes, cudaMemcpyHostToDevice, ctx->stream), "H2D profile images");
  if (!rc && n_chain) {
    rc = check_cuda(ctx, cudaMemcpyAsync(arena->dev + jobs_off, ctx->staging_img + host_bytes, jobs_bytes, cudaMemcpyHostToDevice, ctx->stream), "H2D chain jobs");
    if (!rc) {
      chain_image_kernel<<<dim3((max_chain_n + 1 + 255) / 256, n_chain), 256, 0, ctx->stream>>>(arena->dev, (const ChainImgJob*)(arena->dev + jobs_off));
      ctx->launches += 1;
      rc = check_cuda(ctx, cudaGetLastError(), "chain_image_kernel");
    }
  }
  if (!rc) rc = check_cuda(ctx, cudaEventRecord(ctx->ev_img, ctx->stream), "record image staging event");
  if (rc) {
    if (arena->dev) cudaFreeAsync(arena->dev, ctx->stream);
    delete arena;
    return rc;
  }
  ctx->h2d_bytes += (long long)(host_bytes + jobs_bytes);
  for (int k = 0; k < m; ++k) {
    hist_profile* p = need[k];
    p->img = image_at(arena->dev + base[k], image_offsets(p->n, p->nt));
    p->img_arena = arena;
    p->nS = nS[k];
    p->nE = nE[k];
    ++arena->refs;
  }
  return HIST_OK;
}

// lpAbsorb rows of the A+1 token classes of a leaf (profile.cpp:52-66): row q < A is 0 at residue q of every component
// and -inf elsewhere, row A (wildcard) is all 0.  One device table per (A, C) and context, built on first use.
int ensure_onehot(hist_ctx* ctx, int A, int C, const double** out) {
  for (auto& e : ctx->onehot)
    if (e.first.first == A && e.first.second == C) {
      *out = e.second;
      return HIST_OK;
    }
  const size_t K = (size_t)A * C;
  std::vector<double> h((size_t)(A + 1) * K, -INFINITY);
  for (int q = 0; q < A; ++q)
    for (int c = 0; c < C; ++c) h[(size_t)q * K + (size_t)c * A + q] = 0.0;
  for (size_t k = 0; k < K; ++k) h[(size_t)A * K + k] = 0.0;
  double* d = nullptr;
  HB_CUDA(ctx, cudaMalloc((void**)&d, h.size() * sizeof(double)));
  HB_CUDA(ctx, cudaMemcpy(d, h.data(), h.size() * sizeof(double), cudaMemcpyHostToDevice));
  ctx->h2d_bytes += (long long)(h.size() * sizeof(double));
  ctx->onehot.push_back({{A, C}, d});
  *out = d;
  return HIST_OK;
}

// Host part of the DPMatrix constructor for one pair: argument checks, the composed key tables of the two guide rows and
// the byte layout of the pair's arrays.  Band keys, edge slots and the anti-diagonal layout are computed on the device
// (pair_layout_kernel) from the profiles' structure images.
int plan_pair(hist_ctx* ctx, const hist_profile* x, const hist_profile* y, const hist_pairhmm_desc* hmm,
              const hist_env_desc* env, PairBuild& b) {
  if (!x || !y || !hmm) return set_err(ctx, HIST_ERR_ARG, "hist_pair_prepare: NULL argument");
  if (x->ctx != ctx || y->ctx != ctx) return set_err(ctx, HIST_ERR_ARG, "hist_pair_prepare: profile belongs to another context");
  if (x->A != hmm->alph_size || y->A != hmm->alph_size || x->C != hmm->components || y->C != hmm->components)
    return set_err(ctx, HIST_ERR_ARG, "hist_pair_prepare: alphabet/component mismatch between profiles and PairHMM");
  const int nx = x->n, ny = y->n, A = x->A, C = x->C, K = A * C, Kp = (K + 3) & ~3;
  if (Kp > 1024) return set_err(ctx, HIST_ERR_ARG, "hist_pair_prepare: C*A = %d exceeds 1024", K);
  const int maxDist = env ? env->max_distance : -1;
  if (maxDist >= 0) {
    if (x->leaf_pos.empty() || y->leaf_pos.empty())
      return set_err(ctx, HIST_ERR_ARG, "hist_pair_prepare: envelope given but profiles carry no closest_leaf_pos");
    if (!env->cumulative_matches || !env->row1_pos_to_col || !env->row2_pos_to_col)
      return set_err(ctx, HIST_ERR_ARG, "hist_pair_prepare: envelope arrays are NULL");
    if (env->row1_len < 0 || env->row2_len < 0 || env->n_cols < 0) return set_err(ctx, HIST_ERR_ARG, "hist_pair_prepare: negative envelope sizes");
    b.k1.resize((size_t)env->row1_len + 1);
    b.k2.resize((size_t)env->row2_len + 1);
    for (int q = 0; q <= env->row1_len; ++q) {
      const int col = env->row1_pos_to_col[q];
      if (col < 0 || col > env->n_cols) return set_err(ctx, HIST_ERR_ARG, "hist_pair_prepare: row1_pos_to_col[%d] = %d outside the guide alignment", q, col);
      b.k1[q] = env->cumulative_matches[col];
    }
    for (int q = 0; q <= env->row2_len; ++q) {
      const int col = env->row2_pos_to_col[q];
      if (col < 0 || col > env->n_cols) return set_err(ctx, HIST_ERR_ARG, "hist_pair_prepare: row2_pos_to_col[%d] = %d outside the guide alignment", q, col);
      b.k2[q] = env->cumulative_matches[col];
    }
  } else {
    b.k1.assign(1, 0);
    b.k2.assign(1, 0);
  }
  const int NI = nx - 1, NJ = ny - 1, nDiag = NI + NJ - 1;

  hist_pair* pr = new hist_pair();
  b.pr = pr;
  pr->ctx = ctx;
  pr->x = x;
  pr->y = y;
  pr->arena = nullptr;
  pr->planes = nullptr;
  pr->bwd_planes = nullptr;
  pr->d_dev = nullptr;
  pr->fwd_done = pr->bwd_done = false;
  pr->lp_end = pr->lp_start = -INFINITY;

  // ---- layouts
  Layout& H = b.host;
  Layout& Dv = b.dev;
  auto layout_side = [&](const hist_profile* p) {
    SideOff o;
    const size_t nt1 = std::max<size_t>(1, p->nt);
    o.d_key = Dv.reserve((size_t)p->n * sizeof(int));
    o.d_t_m = Dv.reserve(nt1 * sizeof(float));
    o.d_t_e = Dv.reserve(nt1 * sizeof(int));
    o.d_ex = Dv.reserve((size_t)p->n * sizeof(int));
    o.d_rs = Dv.reserve((size_t)p->n * sizeof(float));
    o.d_ins = Dv.reserve((size_t)p->n * sizeof(float));
    o.d_H = Dv.reserve((size_t)p->n * Kp * sizeof(float));
    o.d_sub = Dv.reserve((size_t)p->n * K * sizeof(double));
    o.d_rootsub = Dv.reserve((size_t)p->n * sizeof(double));
    o.d_insl = Dv.reserve((size_t)p->n * sizeof(double));
    o.d_fm_a = Dv.reserve((size_t)p->n * sizeof(int4));
    o.d_fm_b = Dv.reserve((size_t)p->n * sizeof(int4));
    o.d_emis = Dv.reserve((size_t)p->n * sizeof(float2));
    o.d_in_edges = Dv.reserve(nt1 * sizeof(int4));
    o.d_bm = Dv.reserve((size_t)p->n * 4 * sizeof(int4));
    o.d_cf = Dv.reserve((size_t)p->n * sizeof(int4));
    o.d_cb = Dv.reserve((size_t)p->n * sizeof(int4));
    o.d_ao_edges = Dv.reserve(nt1 * 2 * sizeof(int4));
    o.d_no_edges = Dv.reserve(nt1 * 2 * sizeof(int4));
    // a leaf the library built from residues (and not a codon batch, whose leftMultiply runs on the tensor cores)
    o.tok = p->d_tokens != nullptr && p->chain && p->n >= 3 && !leftmul_tc_applicable(A, C) &&
            (long long)p->n * (C * A + 8) < (1LL << 31);   // leaf_scatter_kernel indexes a side's arrays with 32 bits
    if (o.tok) {
      const size_t rows = (size_t)A + 1;
      o.k_sub = Dv.reserve(rows * K * sizeof(double));
      o.k_H = Dv.reserve(rows * Kp * sizeof(float));
      o.k_ex = Dv.reserve(rows * sizeof(int));
      o.k_rs = Dv.reserve(rows * sizeof(float));
      o.k_ins = Dv.reserve(rows * sizeof(float));
      o.k_rootsub = Dv.reserve(rows * sizeof(double));
      o.k_insl = Dv.reserve(rows * sizeof(double));
    }
    return o;
  };
  b.ox = layout_side(x);
  b.oy = layout_side(y);
  b.d_dinfo = Dv.reserve(((size_t)nDiag + 64) * sizeof(DiagInfo));   // the fill kernels prefetch layout records a few dozen steps ahead
  b.d_scratch = Dv.reserve(((size_t)4 * NI + 2 * NJ + 8) * sizeof(int));
  Dv.reserve(0, 256);
  b.h_k1 = H.put(b.k1.data(), b.k1.size() * sizeof(int));
  b.h_k2 = H.put(b.k2.data(), b.k2.size() * sizeof(int));
  b.h_lsub = H.put(hmm->l_sub, (size_t)K * A * sizeof(double));
  b.h_rsub = H.put(hmm->r_sub, (size_t)K * A * sizeof(double));
  b.h_root = H.put(hmm->log_root, (size_t)K * sizeof(double));
  b.h_lins = H.put(hmm->l_log_ins, (size_t)K * sizeof(double));
  b.h_rins = H.put(hmm->r_log_ins, (size_t)K * sizeof(double));
  b.h_lcw = H.put(hmm->l_log_cpt_weight, (size_t)C * sizeof(double));
  b.h_rcw = H.put(hmm->r_log_cpt_weight, (size_t)C * sizeof(double));
  b.h_pairdev = H.reserve(sizeof(PairDev), 64);
  H.reserve(0, 256);

  PairDev& D = pr->dev;
  memset(&D, 0, sizeof(D));
  D.K = K; D.Kp = Kp; D.C = C; D.A = A;
  D.max_dist = maxDist;
  D.env_len1 = maxDist >= 0 ? env->row1_len : 0;
  D.env_len2 = maxDist >= 0 ? env->row2_len : 0;
  D.x.n = nx; D.x.nt = x->nt;
  D.y.n = ny; D.y.nt = y->nt;
  // HMM weights, w[src][dest] (pairhmm.cpp:46-115); absent pairs are -inf / 0
  for (int a = 0; a < 6; ++a)
    for (int c = 0; c < 6; ++c) pr->w_log[a][c] = -INFINITY;
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
      for (int c = 0; c < 6; ++c) D.W[a][c] = w[a][c] > -INFINITY ? (float)exp(w[a][c]) : 0.f;
    for (int a = 0; a < 6; ++a)
      for (int c = 0; c < 6; ++c) D.Wlog[a][c] = w[a][c];
  }
  pr->hmm_scalars = *hmm;
  pr->hmm_scalars.log_root = pr->hmm_scalars.l_sub = pr->hmm_scalars.r_sub = nullptr;
  pr->hmm_scalars.l_log_ins = pr->hmm_scalars.r_log_ins = nullptr;
  pr->hmm_scalars.l_log_cpt_weight = pr->hmm_scalars.r_log_cpt_weight = nullptr;
  D.n_diag = nDiag;
  D.nS = x->nS; D.nE = y->nE;
  D.max_threads = 1;
  D.chain = x->chain && y->chain && nx >= 3 && ny >= 3 && D.nS == 1 && D.nE == 1 ? 1 : 0;
  D.n_stored = 0;
  pr->info.x_size = nx;
  pr->info.y_size = ny;
  pr->info.n_cells = 0;
  pr->info.n_stored = 0;
  pr->info.n_diagonals = nDiag;
  pr->info.max_diag_width = 0;
  pr->info.h2d_bytes = (int64_t)H.bytes;
  return HIST_OK;
}

// Resolve the device pointers of one pair once its two arena parts are placed (the cell planes follow later).
void place_pair(PairBuild& b, char* arena, size_t scal_off, size_t lay_off, const double* onehot) {
  hist_pair* pr = b.pr;
  char* hb_ = arena + b.host_base;
  char* db = arena + b.dev_base;
  auto side = [&](SideDev& s, const hist_profile* p, const SideOff& o, bool isY) {
    const ProfImage& im = p->img;
    s.flags = im.flags;
    s.t_src = im.t_src; s.t_dst = im.t_dst; s.t_lp = im.t_lp;
    s.in_off = im.in_off; s.in_idx = im.in_idx;
    s.ao_off = im.ao_off; s.ao_idx = im.ao_idx;
    s.no_off = im.no_off; s.no_idx = im.no_idx;
    s.near_ = isY ? im.near_end : im.near_start;
    s.slot = isY ? im.slot_end : im.slot_start;
    s.key = (int*)(db + o.d_key);
    s.t_m = (float*)(db + o.d_t_m); s.t_e = (int*)(db + o.d_t_e);
    s.ex = (int*)(db + o.d_ex); s.rs = (float*)(db + o.d_rs); s.ins = (float*)(db + o.d_ins);
    s.H = (float*)(db + o.d_H);
    s.lp_absorb = p->d_lp_absorb;
    s.sub = (double*)(db + o.d_sub); s.rootsub = (double*)(db + o.d_rootsub); s.insl = (double*)(db + o.d_insl);
    s.fm_a = (int4*)(db + o.d_fm_a); s.fm_b = (int4*)(db + o.d_fm_b); s.emis = (float2*)(db + o.d_emis);
    s.in_edges = (int4*)(db + o.d_in_edges); s.bm = (int4*)(db + o.d_bm);
    s.cf = (int4*)(db + o.d_cf); s.cb = (int4*)(db + o.d_cb);
    s.ao_edges = (int4*)(db + o.d_ao_edges); s.no_edges = (int4*)(db + o.d_no_edges);
    if (o.tok) {
      s.tokens = p->d_tokens;
      s.tk_onehot = onehot;
      s.tk_sub = (double*)(db + o.k_sub); s.tk_H = (float*)(db + o.k_H); s.tk_ex = (int*)(db + o.k_ex);
      s.tk_rs = (float*)(db + o.k_rs); s.tk_ins = (float*)(db + o.k_ins);
      s.tk_rootsub = (double*)(db + o.k_rootsub); s.tk_insl = (double*)(db + o.k_insl);
    }
  };
  PairDev& D = pr->dev;
  side(D.x, pr->x, b.ox, false);
  side(D.y, pr->y, b.oy, true);
  D.start_rows = pr->x->img.start_rows;
  D.end_cols = pr->y->img.end_cols;
  D.dinfo = (DiagInfo*)(db + b.d_dinfo);
  D.scratch = (int*)(db + b.d_scratch);
  D.env_k1 = (const int*)(hb_ + b.h_k1);
  D.env_k2 = (const int*)(hb_ + b.h_k2);
  D.x_leaf_pos = pr->x->img.leaf_pos;
  D.y_leaf_pos = pr->y->img.leaf_pos;
  D.l_sub = (const double*)(hb_ + b.h_lsub); D.r_sub = (const double*)(hb_ + b.h_rsub);
  D.log_root = (const double*)(hb_ + b.h_root);
  D.l_ins = (const double*)(hb_ + b.h_lins); D.r_ins = (const double*)(hb_ + b.h_rins);
  D.l_cw = (const double*)(hb_ + b.h_lcw); D.r_cw = (const double*)(hb_ + b.h_rcw);
  double* scal = (double*)(arena + scal_off);
  D.lp_end = scal;
  D.lp_start = scal + 1;
  D.n_cells = (unsigned long long*)(scal + 2);
  D.layout_out = (LayoutOut*)(arena + lay_off);
  pr->d_dev = (PairDev*)(hb_ + b.h_pairdev);
}

int ensure_staging(hist_ctx* ctx, size_t bytes, int slot) {
  char*& buf = slot ? ctx->staging2 : ctx->staging;
  size_t& have = slot ? ctx->staging2_bytes : ctx->staging_bytes;
  return ensure_pinned(ctx, buf, have, bytes);
}

// A sub-batch whose transfer and table kernels are enqueued but not yet waited for.
struct PrepPending {
  std::vector<PairBuild> builds;
  hb_arena* arena = nullptr;
  hb_arena* planes = nullptr;
  size_t scal_off = 0;
  hist_pair** out = nullptr;
};

void drop_pending(hist_ctx* ctx, PrepPending& pd) {
  for (auto& b : pd.builds) delete b.pr;
  pd.builds.clear();
  for (hb_arena** a : {&pd.arena, &pd.planes})
    if (*a) {
      if ((*a)->dev) cudaFreeAsync((*a)->dev, ctx->stream);
      delete *a;
      *a = nullptr;
    }
}

// Pairs handed to the caller.  Nothing is waited for: the table kernels are still running, and the exact cell counts
// they produce are read back by pair_cells when somebody asks.
int finish_pending(hist_ctx* ctx, PrepPending& pd) {
  (void)ctx;
  const int n = (int)pd.builds.size();
  pd.arena->scal_off = pd.scal_off;
  pd.arena->n_pairs = n;
  for (int k = 0; k < n; ++k) {
    hist_pair* pr = pd.builds[k].pr;
    pr->batch_index = k;
    pr->info.n_cells = -1;
    pr->arena = pd.arena;
    pr->planes = pd.planes;
    ++pd.arena->refs;
    ++pd.planes->refs;
    ++const_cast<hist_profile*>(pr->x)->refs;
    ++const_cast<hist_profile*>(pr->y)->refs;
    pd.out[k] = pr;
  }
  pd.builds.clear();
  pd.arena = nullptr;
  pd.planes = nullptr;
  return HIST_OK;
}

void arena_release(hist_ctx* ctx, hb_arena* a) {
  if (!a) return;
  if (--a->refs == 0) {
    if (a->dev) cudaFreeAsync(a->dev, ctx->stream);
    delete a;
  }
}

}  // namespace

namespace hb {
void host_parallel_for(int n, int grain, const std::function<void(int)>& fn) { parallel_for(n, grain, fn); }
int ensure_staging_buf(hist_ctx* ctx, size_t bytes, int slot) { return ensure_staging(ctx, bytes, slot); }
void arena_drop(hist_ctx* ctx, hb_arena* a) { arena_release(ctx, a); }
int pair_cells(hist_pair* pr, long long* n_cells) {
  hist_ctx* ctx = pr->ctx;
  hb_arena* a = pr->arena;
  if (!a->counted) {
    // the count is the GCUPS unit, not something the reference's constructor computes: its kernel runs when asked for
    cudaSetDevice(ctx->device);
    const int n = a->n_pairs;
    const int chunks = std::max(1, std::min(a->max_diag / 16 + 1, std::max(32, (4 * ctx->sm_count + n - 1) / n)));  // 4 diagonals in flight per CTA
    count_cells_kernel<<<dim3(chunks, n), 128, 0, ctx->stream>>>((PairDev* const*)(a->dev + a->ptr_off));
    ctx->launches += 1;
    HB_CUDA(ctx, cudaGetLastError());
    std::vector<double> scal((size_t)a->n_pairs * 4);
    HB_CUDA(ctx, cudaMemcpyAsync(scal.data(), a->dev + a->scal_off, scal.size() * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    HB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    a->n_cells.resize(a->n_pairs);
    for (int k = 0; k < a->n_pairs; ++k) {
      unsigned long long nc;
      memcpy(&nc, &scal[(size_t)k * 4 + 2], sizeof(nc));
      a->n_cells[k] = (long long)nc;
    }
    a->counted = true;
  }
  pr->info.n_cells = a->n_cells[pr->batch_index];
  if (n_cells) *n_cells = pr->info.n_cells;
  return HIST_OK;
}
}  // namespace hb

namespace {
// HIST_B200_VERIFY_PREP=1: the band keys, edge flags and layout records the device produced are read back and compared
// with a plain host computation of the same integers (forward.cpp:36-42, 58-65); mismatches go to stderr.
int verify_pair_layout(hist_ctx* ctx, const hist_pair* pr, const hist_env_desc* env, int k) {
  const hist_profile* x = pr->x;
  const hist_profile* y = pr->y;
  const PairDev& D = pr->dev;
  const int nx = x->n, ny = y->n, NI = nx - 1, NJ = ny - 1, nDiag = NI + NJ - 1, maxDist = D.max_dist;
  SideHost sx, sy;
  build_side_lists(x, sx);
  build_side_lists(y, sy);
  std::vector<int> xk(nx, 0), yk(ny, 0);
  if (maxDist >= 0) {
    for (int i = 0; i < nx; ++i) xk[i] = env->cumulative_matches[env->row1_pos_to_col[i == 0 ? 0 : x->leaf_pos[i]]];
    for (int j = 0; j < ny; ++j) yk[j] = env->cumulative_matches[env->row2_pos_to_col[j == 0 ? 0 : y->leaf_pos[j]]];
  }
  std::vector<uint8_t> xn(nx, 0), yn(ny, 0);
  xn[0] = 1;
  for (int i = 0; i < nx; ++i)
    if (xn[i])
      for (int q = sx.no_off[i]; q < sx.no_off[i + 1]; ++q) xn[x->t_dst[sx.no_idx[q]]] = 1;
  for (int q = sy.in_off[ny - 1]; q < sy.in_off[ny]; ++q) yn[y->t_src[sy.in_idx[q]]] = 1;
  std::vector<int> jlo(NI), jhi(NI);
  if (maxDist < 0) {
    for (int i = 0; i < NI; ++i) { jlo[i] = 0; jhi[i] = NJ - 1; }
  } else {
    std::vector<int> pmax(NJ), smin(NJ);
    for (int j = 0; j < NJ; ++j) pmax[j] = j ? std::max(pmax[j - 1], yk[j]) : yk[0];
    for (int j = NJ - 1; j >= 0; --j) smin[j] = j < NJ - 1 ? std::min(smin[j + 1], yk[j]) : yk[NJ - 1];
    for (int i = 0; i < NI; ++i) {
      jlo[i] = (int)(std::lower_bound(pmax.begin(), pmax.end(), xk[i] - maxDist) - pmax.begin());
      jhi[i] = (int)(std::upper_bound(smin.begin(), smin.end(), xk[i] + maxDist) - smin.begin()) - 1;
    }
  }
  std::vector<int> rows;
  {
    int run = -1;
    for (int i = 0; i < NI; ++i)
      if (jlo[i] <= jhi[i]) { run = std::max(run, jhi[i]); jhi[i] = run; rows.push_back(i); }
    run = NJ;
    for (int i = NI - 1; i >= 0; --i)
      if (jlo[i] <= jhi[i]) { run = std::min(run, jlo[i]); jlo[i] = run; }
  }
  int nS = 0, nE = 0;
  for (int i = 0; i < NI; ++i) nS += xn[i];
  for (int j = 0; j < NJ; ++j) nE += yn[j];
  std::vector<DiagInfo> want(nDiag);
  unsigned long long off = 0;
  {
    size_t lo = 0, hi = 0;
    for (int d = 0; d < nDiag; ++d) {
      while (lo < rows.size() && rows[lo] + jhi[rows[lo]] < d) ++lo;
      while (hi < rows.size() && rows[hi] + jlo[rows[hi]] <= d) ++hi;
      DiagInfo di = {0, 0, (unsigned)off, 0};
      if (lo < hi) { di.ilo = rows[lo]; di.width = rows[hi - 1] - rows[lo] + 1; }
      want[d] = di;
      off += (unsigned long long)(nS + di.width + nE);
    }
  }
  std::vector<DiagInfo> got(nDiag);
  std::vector<int> gxk(nx), gyk(ny);
  std::vector<uint8_t> gxn(nx), gyn(ny), gxf(nx), gyf(ny);
  HB_CUDA(ctx, cudaMemcpy(got.data(), D.dinfo, nDiag * sizeof(DiagInfo), cudaMemcpyDeviceToHost));
  HB_CUDA(ctx, cudaMemcpy(gxk.data(), D.x.key, nx * sizeof(int), cudaMemcpyDeviceToHost));
  HB_CUDA(ctx, cudaMemcpy(gyk.data(), D.y.key, ny * sizeof(int), cudaMemcpyDeviceToHost));
  HB_CUDA(ctx, cudaMemcpy(gxn.data(), D.x.near_, nx, cudaMemcpyDeviceToHost));
  HB_CUDA(ctx, cudaMemcpy(gyn.data(), D.y.near_, ny, cudaMemcpyDeviceToHost));
  HB_CUDA(ctx, cudaMemcpy(gxf.data(), D.x.flags, nx, cudaMemcpyDeviceToHost));
  HB_CUDA(ctx, cudaMemcpy(gyf.data(), D.y.flags, ny, cudaMemcpyDeviceToHost));
  int bad = 0;
  auto report = [&](const char* what, int idx, long long g, long long w) {
    if (bad++ < 8) fprintf(stderr, "[verify_prep] pair %d %s[%d]: device %lld, host %lld\n", k, what, idx, g, w);
  };
  for (int i = 0; i < nx; ++i) {
    if (gxk[i] != xk[i]) report("x.key", i, gxk[i], xk[i]);
    if (gxn[i] != xn[i]) report("x.near", i, gxn[i], xn[i]);
    if (gxf[i] != sx.flags[i]) report("x.flags", i, gxf[i], sx.flags[i]);
  }
  for (int j = 0; j < ny; ++j) {
    if (gyk[j] != yk[j]) report("y.key", j, gyk[j], yk[j]);
    if (gyn[j] != yn[j]) report("y.near", j, gyn[j], yn[j]);
    if (gyf[j] != sy.flags[j]) report("y.flags", j, gyf[j], sy.flags[j]);
  }
  for (int d = 0; d < nDiag; ++d) {
    if (got[d].ilo != want[d].ilo) report("dinfo.ilo", d, got[d].ilo, want[d].ilo);
    if (got[d].width != want[d].width) report("dinfo.width", d, got[d].width, want[d].width);
    if (got[d].off != want[d].off) report("dinfo.off", d, got[d].off, want[d].off);
    if (got[d].pad != 0) report("dinfo.pad", d, got[d].pad, 0);
  }
  if (D.nS != nS) report("nS", 0, D.nS, nS);
  if (D.nE != nE) report("nE", 0, D.nE, nE);
  if ((unsigned long long)D.n_stored != off) report("n_stored", 0, D.n_stored, (long long)off);
  fprintf(stderr, "[verify_prep] pair %d: %d x %d states, %d diagonals, %llu stored: %s\n", k, nx, ny, nDiag, off, bad ? "MISMATCH" : "ok");
  return bad ? set_err(ctx, HIST_ERR_STATE, "verify_prep: device layout of pair %d differs from the host computation", k) : HIST_OK;
}
}  // namespace

namespace {
// The DPMatrix constructors of a sub-batch.  Host: argument checks and byte layouts; ONE transfer of the guide-row key
// tables, the PairHMM tables and the descriptors; the device builds band keys and the anti-diagonal layout
// (pair_layout_kernel); its totals come back (the one host synchronisation), the cell planes are allocated, and the
// table kernels are enqueued.  Does not wait for those.
int prepare_enqueue(hist_ctx* ctx, int32_t n, const hist_profile* const* xs, const hist_profile* const* ys,
                    const hist_pairhmm_desc* const* hmms, const hist_env_desc* const* envs, int32_t flags, int slot,
                    PrepPending& pd) {
  const bool want_bwd = (flags & HIST_PREP_BACKWARD) != 0;
  static const bool verify = getenv("HIST_B200_VERIFY_PREP") != nullptr;
  HostTimer tm("prepare_batch");
  for (int k = 0; k < n; ++k)
    if (!xs[k] || !ys[k] || !hmms[k]) return set_err(ctx, HIST_ERR_ARG, "hist_pair_prepare: NULL argument");
  int rc = ensure_images(ctx, xs, ys, n);
  if (rc) return rc;
  // lpAbsorb rows of the token classes of leaf sides, one table per (alphabet size, components) of the context
  std::vector<const double*> onehots(n, nullptr);
  for (int k = 0; k < n; ++k)
    if ((xs[k]->d_tokens || ys[k]->d_tokens) && xs[k]->A == ys[k]->A && xs[k]->C == ys[k]->C) {
      rc = ensure_onehot(ctx, xs[k]->A, xs[k]->C, &onehots[k]);
      if (rc) return rc;
    }
  tm.lap("profile images");
  pd.builds.clear();
  pd.builds.resize(n);
  std::vector<PairBuild>& builds = pd.builds;
  auto fail = [&](int rc_) {
    for (auto& b : builds) {
      delete b.pr;
      b.pr = nullptr;
    }
    return rc_;
  };
  {
    std::vector<int> rcs(n, 0);
    parallel_for(n, 32, [&](int k) { rcs[k] = plan_pair(ctx, xs[k], ys[k], hmms[k], envs ? envs[k] : nullptr, builds[k]); });
    for (int k = 0; k < n; ++k)
      if (rcs[k]) return fail(rcs[k]);
  }
  tm.lap("host plan");
  size_t host_total = 0;
  for (int k = 0; k < n; ++k) {
    builds[k].host_base = host_total;
    host_total += builds[k].host.bytes;
  }
  // arena = [host-initialised parts][pointer table][plane placements][scalars of all pairs][layout totals][device-only parts]
  auto up256 = [](size_t v) { return (v + 255) / 256 * 256; };
  const size_t ptr_off = host_total;
  const size_t place_off = ptr_off + up256((size_t)n * sizeof(PairDev*));
  const size_t scal_off = place_off + up256((size_t)n * sizeof(PlanePlace));
  const size_t lay_off = scal_off + up256((size_t)n * 4 * sizeof(double));
  size_t total = lay_off + up256((size_t)n * sizeof(LayoutOut));
  for (int k = 0; k < n; ++k) {
    builds[k].dev_base = total;
    total += builds[k].dev.bytes;
  }
  hb_arena* arena = new hb_arena{nullptr, total, 0};
  rc = check_cuda(ctx, cudaMallocFromPoolAsync((void**)&arena->dev, total, ctx->pool, ctx->stream), "cudaMallocAsync(pair arena)");
  if (!rc) rc = ensure_staging(ctx, scal_off, slot);
  if (!rc) rc = check_cuda(ctx, cudaEventSynchronize(ctx->ev_stage[slot]), "wait for the staging buffer");  // its previous transfer
  if (rc) {
    if (arena->dev) cudaFreeAsync(arena->dev, ctx->stream);
    delete arena;
    return fail(rc);
  }
  pd.arena = arena;
  auto fail_all = [&](int rc_) {
    cudaStreamSynchronize(ctx->stream);
    fail(rc_);
    builds.clear();
    drop_pending(ctx, pd);
    return rc_;
  };
  tm.lap("alloc");
  // fill the pinned staging image and ship it in one transfer
  char* st = slot ? ctx->staging2 : ctx->staging;
  int maxN = 1, maxNt = 1, maxDiag = 1, maxKp = 4, maxKC = 1;
  parallel_for(n, 32, [&](int k) {
    PairBuild& b = builds[k];
    place_pair(b, arena->dev, scal_off + (size_t)k * 4 * sizeof(double), lay_off + (size_t)k * sizeof(LayoutOut),
               (b.ox.tok || b.oy.tok) ? onehots[k] : nullptr);
    char* base = st + b.host_base;
    for (auto& c : b.host.copies) memcpy(base + c.first, c.second.first, c.second.second);
    memcpy(base + b.h_pairdev, &b.pr->dev, sizeof(PairDev));
    ((PairDev**)(st + ptr_off))[k] = b.pr->d_dev;
  });
  for (int k = 0; k < n; ++k) {
    const PairDev& D = builds[k].pr->dev;
    maxN = std::max(maxN, std::max(D.x.n, D.y.n));
    maxNt = std::max(maxNt, std::max(D.x.nt, D.y.nt));
    maxDiag = std::max(maxDiag, D.n_diag);
    maxKp = std::max(maxKp, D.Kp);
    maxKC = std::max(maxKC, D.K + D.C);
  }
  tm.lap("staging fill");
  PairDev* const* dptrs = (PairDev* const*)(arena->dev + ptr_off);
  std::vector<LayoutOut> lay(n);
  rc = check_cuda(ctx, cudaMemcpyAsync(arena->dev, st, place_off, cudaMemcpyHostToDevice, ctx->stream), "H2D pair arena");
  ctx->h2d_bytes += (long long)place_off;
  if (!rc) rc = check_cuda(ctx, cudaMemsetAsync(arena->dev + scal_off, 0, (size_t)n * 4 * sizeof(double), ctx->stream), "memset scalars");
  if (!rc) {
    pair_layout_kernel<<<n, 256, 0, ctx->stream>>>(dptrs);
    ctx->launches += 1;
    rc = check_cuda(ctx, cudaGetLastError(), "pair_layout_kernel");
  }
  if (!rc) rc = check_cuda(ctx, cudaMemcpyAsync(lay.data(), arena->dev + lay_off, (size_t)n * sizeof(LayoutOut), cudaMemcpyDeviceToHost, ctx->stream), "D2H layout totals");
  if (!rc) rc = check_cuda(ctx, cudaStreamSynchronize(ctx->stream), "sync after pair layout");
  if (rc) return fail_all(rc);
  tm.lap("layout on device");
  // ---- cell planes, sized by the layout
  size_t plane_total = 0;
  for (int k = 0; k < n; ++k) {
    PairBuild& b = builds[k];
    const LayoutOut& L = lay[k];
    if (L.err)
      return fail_all(set_err(ctx, HIST_ERR_ARG, "%s state %d: closest-leaf position %d outside guide row (len %d)", L.err == 1 ? "x" : "y",
                              L.err_state, L.err_pos, L.err == 1 ? b.pr->dev.env_len1 : b.pr->dev.env_len2));
    if (L.n_stored >= (1ull << 31))
      return fail_all(set_err(ctx, HIST_ERR_ARG, "hist_pair_prepare: %llu stored cells exceed the 2^31 limit of one pair", L.n_stored));
    const size_t ns = (size_t)L.n_stored;
    Layout P;
    b.p_fA = P.reserve(ns * sizeof(float4), 256);
    b.p_fB = P.reserve(ns * sizeof(float2), 256);
    b.p_absorb = P.reserve(ns * sizeof(float), 256);
    if (want_bwd) {
      b.p_bA = P.reserve(ns * sizeof(float4), 256);
      b.p_bB = P.reserve(ns * sizeof(float2), 256);
    }
    P.reserve(0, 256);
    b.plane_base = plane_total;
    plane_total += P.bytes;
    PairDev& D = b.pr->dev;
    D.n_stored = (unsigned)L.n_stored;
    D.max_threads = L.max_threads;
    b.pr->info.n_stored = (int64_t)L.n_stored;
    b.pr->info.max_diag_width = L.max_width;
  }
  hb_arena* planes = new hb_arena{nullptr, std::max<size_t>(plane_total, 256), 0};
  rc = check_cuda(ctx, cudaMallocFromPoolAsync((void**)&planes->dev, planes->bytes, ctx->pool, ctx->stream), "cudaMallocAsync(cell planes)");
  if (rc) {
    delete planes;
    return fail_all(rc);
  }
  pd.planes = planes;
  {
    PlanePlace* pl = (PlanePlace*)(st + place_off);   // the staging buffer's first transfer has completed (stream waited for above)
    for (int k = 0; k < n; ++k) {
      PairBuild& b = builds[k];
      char* pb = planes->dev + b.plane_base;
      PairDev& D = b.pr->dev;
      D.fA = (float4*)(pb + b.p_fA); D.fB = (float2*)(pb + b.p_fB); D.absorb = (float*)(pb + b.p_absorb);
      D.bA = want_bwd ? (float4*)(pb + b.p_bA) : nullptr;
      D.bB = want_bwd ? (float2*)(pb + b.p_bB) : nullptr;
      pl[k] = PlanePlace{D.fA, D.fB, D.absorb, D.bA, D.bB};
    }
    rc = check_cuda(ctx, cudaMemcpyAsync(arena->dev + place_off, pl, (size_t)n * sizeof(PlanePlace), cudaMemcpyHostToDevice, ctx->stream), "H2D plane placements");
    ctx->h2d_bytes += (long long)((size_t)n * sizeof(PlanePlace));
    if (!rc) rc = check_cuda(ctx, cudaEventRecord(ctx->ev_stage[slot], ctx->stream), "record staging event");
  }
  // ---- F0/F1 and the packed records on device, one launch each for the whole batch
  if (!rc) {
    place_planes_kernel<<<(n + 127) / 128, 128, 0, ctx->stream>>>(dptrs, (const PlanePlace*)(arena->dev + place_off), n);
    const int threads = std::max(32, (maxKp + 31) / 32 * 32);
    const int stride = maxKC + 32;
    bool tensor = true;   // every pair of the batch is a one-component codon merge: leftMultiply as a tcgen05 GEMM
    for (int k = 0; k < n && tensor; ++k) tensor = leftmul_tc_applicable(builds[k].pr->dev.A, builds[k].pr->dev.C);
    bool all_tok = true;   // every side of every pair is a leaf side: the per-state pass has nothing to do
    for (int k = 0; k < n && all_tok; ++k) all_tok = builds[k].ox.tok && builds[k].oy.tok;
    if (tensor) {
      rc = launch_leftmul_tc(ctx, dptrs, n, maxN);
    } else if (all_tok) {
      ctx->launches -= 1;
    } else if (threads == 32) {
      const int per_cta = 8;
      prepare_side_kernel<true><<<dim3((maxN + per_cta - 1) / per_cta, n, 2), dim3(32, per_cta), (size_t)per_cta * stride * sizeof(double), ctx->stream>>>(dptrs, stride);
    } else {
      prepare_side_kernel<false><<<dim3(maxN, n, 2), threads, (size_t)stride * sizeof(double), ctx->stream>>>(dptrs, stride);
    }
    bool any_tok = false;
    int maxA = 1;
    for (int k = 0; k < n; ++k) {
      any_tok = any_tok || builds[k].ox.tok || builds[k].oy.tok;
      maxA = std::max(maxA, builds[k].pr->dev.A);
    }
    if (any_tok) {   // (never in a tensor batch)
      if (threads == 32) {
        const int per_cta = 8;
        prepare_side_kernel<true, true><<<dim3((maxA + 1 + per_cta - 1) / per_cta, n, 2), dim3(32, per_cta), (size_t)per_cta * stride * sizeof(double), ctx->stream>>>(dptrs, stride);
      } else {
        prepare_side_kernel<false, true><<<dim3(maxA + 1, n, 2), threads, (size_t)stride * sizeof(double), ctx->stream>>>(dptrs, stride);
      }
      // ~8 elements per thread: the per-block cost (the PairDev record, the side's pointers) is paid 10 times per side, not 64
      leaf_scatter_kernel<<<dim3((unsigned)std::max<long long>(1, std::min<long long>(((long long)maxN * maxKp + 2047) / 2048, 64)), n, 2), 256, 0, ctx->stream>>>(dptrs);
      ctx->launches += 2;
    }
    prepare_trans_kernel<<<dim3((maxNt + 127) / 128, n, 2), 128, 0, ctx->stream>>>(dptrs);
    pack_meta_kernel<<<dim3((maxN + 127) / 128, n, 2), 128, 0, ctx->stream>>>(dptrs);
    arena->ptr_off = ptr_off;
    arena->max_diag = maxDiag;
    ctx->launches += 4;
    if (!rc) rc = check_cuda(ctx, cudaGetLastError(), "prepare kernels");
  }
  tm.lap("enqueue");
  if (!rc && verify) {
    rc = check_cuda(ctx, cudaStreamSynchronize(ctx->stream), "sync before verify");
    for (int k = 0; k < n && !rc; ++k) rc = verify_pair_layout(ctx, builds[k].pr, envs ? envs[k] : nullptr, k);
  }
  if (rc) return fail_all(rc);
  pd.scal_off = scal_off;
  return HIST_OK;
}
}  // namespace

extern "C" int hist_pair_prepare_batch(hist_ctx* ctx, int32_t n, const hist_profile* const* xs, const hist_profile* const* ys,
                                       const hist_pairhmm_desc* const* hmms, const hist_env_desc* const* envs, int32_t flags,
                                       hist_pair** out) {
  if (!ctx || n <= 0 || !xs || !ys || !hmms || !out) return set_err(ctx, HIST_ERR_ARG, "hist_pair_prepare_batch: bad arguments");
  for (int k = 0; k < n; ++k) out[k] = nullptr;
  cudaSetDevice(ctx->device);
  // a level is shipped in sub-batches: while the device receives and tabulates one, the host threads build the next
  // (every sub-batch costs one host synchronisation, the layout totals; measured on the 512-merge leaf level, end to
  // end: 1 piece 8.4 G cells/s, 2 pieces 8.1, 4 pieces 7.2 -- so one piece unless HIST_B200_PREP_PARTS says otherwise)
  int parts = 1;
  if (const char* e = getenv("HIST_B200_PREP_PARTS")) parts = std::max(1, std::min(n, atoi(e)));
  const int per = (n + parts - 1) / parts;
  std::vector<PrepPending> pend((n + per - 1) / per);
  int rc = HIST_OK;
  for (size_t c = 0; c < pend.size() && !rc; ++c) {
    const int first = (int)c * per, cnt = std::min(per, n - first);
    pend[c].out = out + first;
    rc = prepare_enqueue(ctx, cnt, xs + first, ys + first, hmms + first, envs ? envs + first : nullptr, flags, (int)(c & 1), pend[c]);
  }
  // The borrowed descriptors have been copied into the pinned staging buffers, which are guarded by events: nothing to
  // wait for.  The table kernels keep running behind this call; whatever is enqueued next on the context's stream
  // (the fills) follows them.
  for (size_t c = 0; c < pend.size() && !rc; ++c) rc = finish_pending(ctx, pend[c]);
  if (rc) {
    cudaStreamSynchronize(ctx->stream);
    for (auto& pd : pend) drop_pending(ctx, pd);
    for (int k = 0; k < n; ++k)
      if (out[k]) {
        hist_pair_destroy(out[k]);
        out[k] = nullptr;
      }
  }
  return rc;
}

extern "C" int hist_pair_prepare(hist_ctx* ctx, const hist_profile* x, const hist_profile* y, const hist_pairhmm_desc* hmm,
                                 const hist_env_desc* env, hist_pair** out) {
  if (!ctx || !x || !y || !hmm || !out) return set_err(ctx, HIST_ERR_ARG, "hist_pair_prepare: NULL argument");
  return hist_pair_prepare_batch(ctx, 1, &x, &y, &hmm, &env, 0, out);
}

namespace {
// a pair holds its two profiles (their host transition lists and the device lpAbsorb table are read after prepare)
void profile_unref(const hist_profile* cp) {
  hist_profile* p = const_cast<hist_profile*>(cp);
  if (p && --p->refs == 0 && p->doomed) hist_profile_destroy(p);
}
}  // namespace

extern "C" void hist_pair_destroy(hist_pair* pr) {
  if (!pr) return;
  cudaSetDevice(pr->ctx->device);
  if (pr->bwd_planes) cudaFreeAsync(pr->bwd_planes, pr->ctx->stream);
  arena_release(pr->ctx, pr->arena);
  arena_release(pr->ctx, pr->planes);
  profile_unref(pr->x);
  profile_unref(pr->y);
  delete pr;
}

extern "C" void hist_pairs_destroy(hist_pair* const* pairs, int32_t n) {
  if (!pairs || n <= 0) return;
  for (int k = 0; k < n; ++k) {   // device memory and reference counts on the calling thread, in order
    hist_pair* pr = pairs[k];
    if (!pr) continue;
    cudaSetDevice(pr->ctx->device);
    if (pr->bwd_planes) cudaFreeAsync(pr->bwd_planes, pr->ctx->stream);
    arena_release(pr->ctx, pr->arena);
    arena_release(pr->ctx, pr->planes);
    profile_unref(pr->x);
    profile_unref(pr->y);
  }
  for (int k = 0; k < n; ++k) delete pairs[k];
}

extern "C" int hist_pair_get_info(const hist_pair* pr, hist_pair_info* info) {
  if (!pr || !info) return HIST_ERR_ARG;
  int rc = pair_cells(const_cast<hist_pair*>(pr), nullptr);
  if (rc) return rc;
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
  hist_ctx* ctx = pr->ctx;
  cudaSetDevice(ctx->device);
  const int nx = pr->info.x_size, ny = pr->info.y_size;
  // keys and edge flags live on the device (pair_layout_kernel, the profiles' structure images)
  std::vector<int> xk(nx), yk(ny);
  std::vector<uint8_t> xn(nx), yn(ny);
  HB_CUDA(ctx, cudaMemcpyAsync(xk.data(), pr->dev.x.key, nx * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
  HB_CUDA(ctx, cudaMemcpyAsync(yk.data(), pr->dev.y.key, ny * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
  HB_CUDA(ctx, cudaMemcpyAsync(xn.data(), pr->dev.x.near_, nx, cudaMemcpyDeviceToHost, ctx->stream));
  HB_CUDA(ctx, cudaMemcpyAsync(yn.data(), pr->dev.y.near_, ny, cudaMemcpyDeviceToHost, ctx->stream));
  HB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  if (x_key) memcpy(x_key, xk.data(), nx * sizeof(int));
  if (y_key) memcpy(y_key, yk.data(), ny * sizeof(int));
  if (x_near_start) memcpy(x_near_start, xn.data(), nx);
  if (y_near_end) memcpy(y_near_end, yn.data(), ny);
  if (in_envelope) {
    const int md = pr->dev.max_dist;
    for (int i = 0; i < nx - 1; ++i)
      for (int j = 0; j < ny - 1; ++j)
        in_envelope[(size_t)i * (ny - 1) + j] = xn[i] || yn[j] || md < 0 || hb_iabs(xk[i] - yk[j]) <= md;
  }
  return HIST_OK;
}
