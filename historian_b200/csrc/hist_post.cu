// historian_b200: posterior decoding of a filled Forward + Backward pair on device.
//
// Replaces BackwardMatrix::cellPostProb (reference forward.cpp:1172-1174) and
// BackwardMatrix::cellsAbovePostProbThreshold (forward.cpp:1302-1319): every (cell, state) whose
// log posterior F + B - lpEnd reaches log(minPostProb), in the pop order of the reference's
// priority_queue made deterministic (decreasing posterior; ties by descending CellCoords).
// Stream compaction over the stored cells, then a device sort.
#include <thrust/execution_policy.h>
#include <thrust/sort.h>

#include <algorithm>
#include <cmath>
#include <mutex>
#include <vector>

#include "hist_dpgraph.cuh"

namespace hb {

struct PostCell {
  double lpp;
  int x, y, s, pad;
};

struct PostOrder {
  __host__ __device__ bool operator()(const PostCell& a, const PostCell& b) const {
    if (a.lpp != b.lpp) return a.lpp > b.lpp;
    if (a.x != b.x) return a.x > b.x;
    if (a.y != b.y) return a.y > b.y;
    return a.s > b.s;
  }
};

// one thread per stored slot; grid = diagonals (strided)
__global__ void posterior_filter_kernel(const PairDev* __restrict__ Pp, double lp_end, double thr, PostCell* __restrict__ out,
                                        unsigned long long capacity, unsigned long long* __restrict__ count) {
  const PairDev& P = *Pp;
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
      if (!ok || i < 0 || j < 0 || i > P.x.n - 2 || j > P.y.n - 2 || !dp_in_envelope(P, i, j)) continue;
      const float4 fa = P.fA[di.off + t], ba = P.bA[di.off + t];
      const float2 fb = P.fB[di.off + t], bb = P.bB[di.off + t];
      const int fE = __float_as_int(fb.y), bE = __float_as_int(bb.y);
      if (fE <= HB_EXP_ZERO / 2 || bE <= HB_EXP_ZERO / 2) continue;
      const float fv[5] = {fa.x, fa.y, fa.z, fa.w, fb.x}, bv[5] = {ba.x, ba.y, ba.z, ba.w, bb.x};
#pragma unroll
      for (int s = 0; s < 5; ++s) {
        if (!(fv[s] > 0.f) || !(bv[s] > 0.f)) continue;
        const double lpp = cell_log(bv[s], bE) + cell_log(fv[s], fE) - lp_end;  // forward.cpp:1172-1174
        if (lpp >= thr) {
          const unsigned long long k = atomicAdd(count, 1ull);
          if (k < capacity) out[k] = {lpp, i, j, s, 0};
        }
      }
    }
  }
}

}  // namespace hb

using namespace hb;

extern "C" int hist_posterior_cells(hist_pair* pair, double min_post_prob, hist_cell* cells, double* log_post, int64_t capacity,
                                    int64_t* n) {
  if (!pair || !n) return HIST_ERR_ARG;
  hist_ctx* ctx = pair->ctx;
  *n = 0;
  if (!pair->fwd_done || !pair->bwd_done) return set_err(ctx, HIST_ERR_STATE, "hist_posterior_cells: needs both the Forward and the Backward fill");
  if (!(pair->lp_end > -INFINITY)) return set_err(ctx, HIST_ERR_NUMERIC, "hist_posterior_cells: zero forward likelihood");
  if (!(min_post_prob > 0)) return set_err(ctx, HIST_ERR_ARG, "hist_posterior_cells: min_post_prob must be positive");
  cudaSetDevice(ctx->device);
  { int rc_ = hb::pair_cells(pair, nullptr); if (rc_) return rc_; }
  const unsigned long long cap = 5ull * (unsigned long long)std::max<int64_t>(pair->info.n_cells, 1);
  PostCell* buf = nullptr;
  unsigned long long* cnt = nullptr;
  HB_CUDA(ctx, cudaMallocFromPoolAsync((void**)&buf, cap * sizeof(PostCell) + 16, ctx->pool, ctx->stream));
  cnt = (unsigned long long*)(buf + cap);
  HB_CUDA(ctx, cudaMemsetAsync(cnt, 0, sizeof(unsigned long long), ctx->stream));
  const double thr = std::log(min_post_prob);
  const int blocks = std::max(1, std::min(pair->info.n_diagonals, 8 * ctx->sm_count));
  const int threads = std::min(256, std::max(32, (pair->dev.max_threads + 31) / 32 * 32));
  posterior_filter_kernel<<<blocks, threads, 0, ctx->stream>>>(pair->d_dev, pair->lp_end, thr, buf, cap, cnt);
  ctx->launches += 1;
  HB_CUDA(ctx, cudaGetLastError());
  unsigned long long found = 0;
  HB_CUDA(ctx, cudaMemcpyAsync(&found, cnt, sizeof(found), cudaMemcpyDeviceToHost, ctx->stream));
  HB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  *n = (int64_t)found;
  int rc = HIST_OK;
  if ((int64_t)found > capacity || (found && (!cells || !log_post))) {
    rc = set_err(ctx, HIST_ERR_CAPACITY, "hist_posterior_cells: need room for %llu cells", found);
  } else if (found) {
    thrust::sort(thrust::cuda::par.on(ctx->stream), buf, buf + found, PostOrder());
    ctx->launches += 1;
    std::vector<PostCell> h(found);
    HB_CUDA(ctx, cudaMemcpyAsync(h.data(), buf, found * sizeof(PostCell), cudaMemcpyDeviceToHost, ctx->stream));
    HB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    for (unsigned long long k = 0; k < found; ++k) {
      cells[k] = {h[k].x, h[k].y, h[k].s};
      log_post[k] = h[k].lpp;
    }
  }
  cudaFreeAsync(buf, ctx->stream);
  return rc;
}

// ------------------------------------------------------------------ P3: Profile::calcSumPathAbsorbProbs
// (profile.cpp:112-131): a 1-D Forward over the profile, lpCum[pos] = LSE over in-transitions of
// lpCum[src] + lpTrans + lpAbs[pos], lpAbs[pos] = LSE_c(logCptWeight[c] + logInnerProduct(logInsProb[c],
// lpAbsorb[pos][c])) for emitting states and 0 for null states.  One block: all threads compute lpAbs in
// parallel, then warp 0 walks the states in order (lanes over the in-transitions, exact log-sum-exp).
namespace hb {
__device__ __forceinline__ double sp_lse2(double a, double b) {
  if (a == -INFINITY) return b;
  if (b == -INFINITY) return a;
  const double m = fmax(a, b);
  return m + log1p(exp(-fabs(a - b)));
}

__global__ void __launch_bounds__(256) sum_path_absorb_kernel(int n, int C, int A, const uint8_t* __restrict__ is_emit,
                                                              const double* __restrict__ lp_absorb, const int* __restrict__ in_off,
                                                              const int* __restrict__ in_src, const double* __restrict__ in_lp,
                                                              const double* __restrict__ log_cw, const double* __restrict__ log_ins,
                                                              double* __restrict__ lp_abs, double* __restrict__ cum) {
  for (int pos = threadIdx.x; pos < n; pos += blockDim.x) {
    double v = 0;
    if (is_emit[pos]) {
      v = -INFINITY;
      for (int c = 0; c < C; ++c) {
        double inner = -INFINITY;  // logInnerProduct (logsumexp.h:126-151)
        for (int a = 0; a < A; ++a) inner = sp_lse2(inner, log_ins[c * A + a] + lp_absorb[((size_t)pos * C + c) * A + a]);
        v = sp_lse2(v, log_cw[c] + inner);
      }
    }
    lp_abs[pos] = v;
    cum[pos] = pos == 0 ? 0.0 : -INFINITY;
  }
  __syncthreads();
  if (threadIdx.x >= 32) return;
  const int lane = threadIdx.x;
  for (int pos = 1; pos < n; ++pos) {
    const int e0 = in_off[pos], e1 = in_off[pos + 1];
    const double la = lp_abs[pos];
    double v = -INFINITY;
    for (int e = e0 + lane; e < e1; e += 32) v = sp_lse2(v, cum[in_src[e]] + in_lp[e] + la);
    if (e1 - e0 > 1)
      for (int o = 16; o; o >>= 1) v = sp_lse2(v, __shfl_xor_sync(0xffffffffu, v, o));
    else
      v = __shfl_sync(0xffffffffu, v, 0);
    if (lane == 0) cum[pos] = v;
    __syncwarp();
  }
}
}  // namespace hb

extern "C" int hist_profile_sum_path_absorb(hist_profile* p, const double* log_cpt_weight, const double* log_ins_prob, double* cum,
                                            double* result) {
  if (!p || !log_cpt_weight || !log_ins_prob || !result) return HIST_ERR_ARG;
  hist_ctx* ctx = p->ctx;
  cudaSetDevice(ctx->device);
  const int n = p->n, nt = p->nt, C = p->C, A = p->A;
  // incoming transitions per state, ascending transition index (the order of ProfileState::in)
  std::vector<int> in_off(n + 1, 0), in_src(std::max(1, nt));
  std::vector<double> in_lp(std::max(1, nt));
  for (int t = 0; t < nt; ++t) ++in_off[p->t_dst[t] + 1];
  for (int i = 0; i < n; ++i) in_off[i + 1] += in_off[i];
  {
    std::vector<int> at(in_off.begin(), in_off.end() - 1);
    for (int t = 0; t < nt; ++t) {
      const int q = at[p->t_dst[t]]++;
      in_src[q] = p->t_src[t];
      in_lp[q] = p->t_lp[t];
    }
  }
  const size_t o_emit = 0, o_off = (n + 255) & ~(size_t)255, o_src = o_off + ((sizeof(int) * (n + 1) + 255) & ~(size_t)255),
               o_lp = o_src + ((sizeof(int) * in_src.size() + 255) & ~(size_t)255), o_cw = o_lp + ((8 * in_lp.size() + 255) & ~(size_t)255),
               o_ins = o_cw + ((8 * (size_t)C + 255) & ~(size_t)255), o_abs = o_ins + ((8 * (size_t)C * A + 255) & ~(size_t)255),
               o_cum = o_abs + ((8 * (size_t)n + 255) & ~(size_t)255), total = o_cum + 8 * (size_t)n;
  char* d = nullptr;
  HB_CUDA(ctx, cudaMallocFromPoolAsync((void**)&d, total, ctx->pool, ctx->stream));
  int rc = 0;
  auto up = [&](size_t off, const void* src, size_t bytes) {
    if (!rc) rc = hb::check_cuda(ctx, cudaMemcpyAsync(d + off, src, bytes, cudaMemcpyHostToDevice, ctx->stream), "H2D sum_path_absorb");
  };
  up(o_emit, p->is_emit.data(), n);
  up(o_off, in_off.data(), sizeof(int) * (n + 1));
  up(o_src, in_src.data(), sizeof(int) * in_src.size());
  up(o_lp, in_lp.data(), 8 * in_lp.size());
  up(o_cw, log_cpt_weight, 8 * (size_t)C);
  up(o_ins, log_ins_prob, 8 * (size_t)C * A);
  if (!rc) {
    hb::sum_path_absorb_kernel<<<1, 256, 0, ctx->stream>>>(n, C, A, (const uint8_t*)(d + o_emit), p->d_lp_absorb, (const int*)(d + o_off),
                                                          (const int*)(d + o_src), (const double*)(d + o_lp), (const double*)(d + o_cw),
                                                          (const double*)(d + o_ins), (double*)(d + o_abs), (double*)(d + o_cum));
    ctx->launches += 1;
    rc = hb::check_cuda(ctx, cudaGetLastError(), "sum_path_absorb_kernel");
  }
  if (!rc && cum) rc = hb::check_cuda(ctx, cudaMemcpyAsync(cum, d + o_cum, 8 * (size_t)n, cudaMemcpyDeviceToHost, ctx->stream), "D2H cum");
  if (!rc) rc = hb::check_cuda(ctx, cudaMemcpyAsync(result, d + o_cum + 8 * (size_t)(n - 1), 8, cudaMemcpyDeviceToHost, ctx->stream), "D2H result");
  if (!rc) rc = hb::check_cuda(ctx, cudaStreamSynchronize(ctx->stream), "sync");
  cudaFreeAsync(d, ctx->stream);
  return rc;
}

// ------------------------------------------------------------------ P3 without device allocation
// hist_profile_sum_path_absorb_noalloc: the same 1-D Forward for callers that cannot call cudaMalloc (code under CUDA
// graph capture, tight serving loops, harnesses that forbid it).  No counterpart in the reference.
// Every piece of scratch -- the staged transition arrays, the in-lists (built on the device by
// this variant instead of on the host), lpAbs and lpCum -- lives in ONE file-scope device array of
// HIST_NOALLOC_SCRATCH_BYTES that every GPU loading the library holds until exit.  All calls share it and it keeps the
// last call's contents: the kernel clears what must start at zero (the in-degree counters), and two calls that overlap
// on one GPU (two contexts / streams on the same device) race.
namespace hb {
static __device__ __align__(256) unsigned char g_noalloc_scratch[HIST_NOALLOC_SCRATCH_BYTES];

struct NoAllocLayout {
  size_t emit, t_src, t_dst, t_lp, cw, ins, in_off, in_cnt, in_idx, lp_abs, cum, total;
};
static NoAllocLayout noalloc_layout(size_t n, size_t nt, size_t C, size_t A) {
  NoAllocLayout l;
  size_t b = 0;
  auto take = [&](size_t bytes) {
    const size_t off = b;
    b = (b + bytes + 255) & ~(size_t)255;
    return off;
  };
  const size_t nt1 = nt ? nt : 1;
  l.t_lp = take(8 * nt1);
  l.cw = take(8 * C);
  l.ins = take(8 * C * A);
  l.lp_abs = take(8 * n);
  l.cum = take(8 * n);
  l.t_src = take(4 * nt1);
  l.t_dst = take(4 * nt1);
  l.in_off = take(4 * (n + 1));
  l.in_cnt = take(4 * n);
  l.in_idx = take(4 * nt1);
  l.emit = take(n);
  l.total = b;
  return l;
}

// One block.  Builds the in-lists (ascending transition index = the order of ProfileState::in) in scratch, then runs
// the same recursion as sum_path_absorb_kernel.
__global__ void __launch_bounds__(256) sum_path_absorb_noalloc_kernel(int n, int nt, int C, int A, const uint8_t* __restrict__ is_emit,
                                                                      const double* __restrict__ lp_absorb, const int* __restrict__ t_src,
                                                                      const int* __restrict__ t_dst, const double* __restrict__ t_lp,
                                                                      const double* __restrict__ log_cw, const double* __restrict__ log_ins,
                                                                      int* __restrict__ in_off, int* __restrict__ in_cnt,
                                                                      int* __restrict__ in_idx, double* __restrict__ lp_abs,
                                                                      double* __restrict__ cum) {
  __shared__ int part[256];
  const int T = blockDim.x, tid = threadIdx.x;
  for (int i = tid; i < n; i += T) in_cnt[i] = 0;   // the array keeps the previous call's contents
  __syncthreads();
  for (int t = tid; t < nt; t += T) atomicAdd(&in_cnt[t_dst[t]], 1);
  __syncthreads();
  // exclusive scan of the in-degrees: every thread folds a contiguous chunk
  const int chunk = (n + T - 1) / T, b = min(n, tid * chunk), e = min(n, b + chunk);
  int acc = 0;
  for (int i = b; i < e; ++i) acc += in_cnt[i];
  part[tid] = acc;
  __syncthreads();
  acc = 0;
  for (int q = 0; q < tid; ++q) acc += part[q];
  for (int i = b; i < e; ++i) {
    in_off[i] = acc;
    acc += in_cnt[i];
    in_cnt[i] = 0;
  }
  if (tid == T - 1) in_off[n] = nt;
  __syncthreads();
  for (int t = tid; t < nt; t += T) {
    const int d = t_dst[t];
    in_idx[in_off[d] + atomicAdd(&in_cnt[d], 1)] = t;
  }
  __syncthreads();
  for (int i = tid; i < n; i += T) {   // the atomics filled each list in arbitrary order: sort it (lists are short)
    const int e0 = in_off[i], e1 = in_off[i + 1];
    for (int a = e0 + 1; a < e1; ++a) {
      const int v = in_idx[a];
      int q = a - 1;
      while (q >= e0 && in_idx[q] > v) { in_idx[q + 1] = in_idx[q]; --q; }
      in_idx[q + 1] = v;
    }
  }
  for (int pos = tid; pos < n; pos += T) {
    double v = 0;
    if (is_emit[pos]) {
      v = -INFINITY;
      for (int c = 0; c < C; ++c) {
        double inner = -INFINITY;  // logInnerProduct (logsumexp.h:126-151)
        for (int a = 0; a < A; ++a) inner = sp_lse2(inner, log_ins[c * A + a] + lp_absorb[((size_t)pos * C + c) * A + a]);
        v = sp_lse2(v, log_cw[c] + inner);
      }
    }
    lp_abs[pos] = v;
    cum[pos] = pos == 0 ? 0.0 : -INFINITY;
  }
  __syncthreads();
  if (tid >= 32) return;
  const int lane = tid;
  for (int pos = 1; pos < n; ++pos) {
    const int e0 = in_off[pos], e1 = in_off[pos + 1];
    const double la = lp_abs[pos];
    double v = -INFINITY;
    for (int q = e0 + lane; q < e1; q += 32) {
      const int t = in_idx[q];
      v = sp_lse2(v, cum[t_src[t]] + t_lp[t] + la);
    }
    if (e1 - e0 > 1)
      for (int o = 16; o; o >>= 1) v = sp_lse2(v, __shfl_xor_sync(0xffffffffu, v, o));
    else
      v = __shfl_sync(0xffffffffu, v, 0);
    if (lane == 0) cum[pos] = v;
    __syncwarp();
  }
}
}  // namespace hb

namespace {
// Device address of the scratch array, looked up once per device with that device current.  In host code the array's
// name is only a host-side stand-in; kernels must be given the address cudaGetSymbolAddress returns.
constexpr int kMaxDevices = 64;
unsigned char* g_noalloc_addr[kMaxDevices] = {};
std::mutex g_noalloc_mu;

int noalloc_scratch(hist_ctx* ctx, unsigned char** out) {
  if (ctx->device < 0 || ctx->device >= kMaxDevices) return hb::set_err(ctx, HIST_ERR_ARG, "noalloc: device %d out of range", ctx->device);
  std::lock_guard<std::mutex> lock(g_noalloc_mu);
  if (!g_noalloc_addr[ctx->device]) {
    void* ptr = nullptr;
    int rc = hb::check_cuda(ctx, cudaSetDevice(ctx->device), "cudaSetDevice");
    if (!rc) rc = hb::check_cuda(ctx, cudaGetSymbolAddress(&ptr, hb::g_noalloc_scratch), "cudaGetSymbolAddress(g_noalloc_scratch)");
    if (rc) return rc;
    if (!ptr) return hb::set_err(ctx, HIST_ERR_CUDA, "cudaGetSymbolAddress returned a NULL scratch address");
    g_noalloc_addr[ctx->device] = (unsigned char*)ptr;
  }
  *out = g_noalloc_addr[ctx->device];
  return HIST_OK;
}
}  // namespace

extern "C" int64_t hist_noalloc_scratch_bytes(void) { return (int64_t)HIST_NOALLOC_SCRATCH_BYTES; }

extern "C" int hist_profile_sum_path_absorb_noalloc(hist_profile* p, const double* log_cpt_weight, const double* log_ins_prob,
                                                    double* cum, double* result) {
  if (!p || !log_cpt_weight || !log_ins_prob || !result) return HIST_ERR_ARG;
  hist_ctx* ctx = p->ctx;
  const int n = p->n, nt = p->nt, C = p->C, A = p->A;
  const hb::NoAllocLayout l = hb::noalloc_layout((size_t)n, (size_t)nt, (size_t)C, (size_t)A);
  if (l.total > (size_t)HIST_NOALLOC_SCRATCH_BYTES)
    return hb::set_err(ctx, HIST_ERR_CAPACITY,
                       "hist_profile_sum_path_absorb_noalloc: a profile of %d states and %d transitions (C=%d, A=%d) needs %zu bytes of "
                       "scratch, the built-in array holds %zu; use hist_profile_sum_path_absorb",
                       n, nt, C, A, l.total, (size_t)HIST_NOALLOC_SCRATCH_BYTES);
  int rc = hb::check_cuda(ctx, cudaSetDevice(ctx->device), "cudaSetDevice");
  if (rc) return rc;
  unsigned char* d = nullptr;
  rc = noalloc_scratch(ctx, &d);
  if (rc) return rc;
  auto up = [&](size_t off, const void* src, size_t bytes) {
    if (!rc && bytes) rc = hb::check_cuda(ctx, cudaMemcpyAsync(d + off, src, bytes, cudaMemcpyHostToDevice, ctx->stream), "H2D sum_path_absorb_noalloc");
  };
  up(l.emit, p->is_emit.data(), (size_t)n);
  up(l.t_src, p->t_src.data(), sizeof(int) * (size_t)nt);
  up(l.t_dst, p->t_dst.data(), sizeof(int) * (size_t)nt);
  up(l.t_lp, p->t_lp.data(), 8 * (size_t)nt);
  up(l.cw, log_cpt_weight, 8 * (size_t)C);
  up(l.ins, log_ins_prob, 8 * (size_t)C * A);
  if (!rc) {
    hb::sum_path_absorb_noalloc_kernel<<<1, 256, 0, ctx->stream>>>(
        n, nt, C, A, (const uint8_t*)(d + l.emit), p->d_lp_absorb, (const int*)(d + l.t_src), (const int*)(d + l.t_dst),
        (const double*)(d + l.t_lp), (const double*)(d + l.cw), (const double*)(d + l.ins), (int*)(d + l.in_off), (int*)(d + l.in_cnt),
        (int*)(d + l.in_idx), (double*)(d + l.lp_abs), (double*)(d + l.cum));
    ctx->launches += 1;
    rc = hb::check_cuda(ctx, cudaGetLastError(), "sum_path_absorb_noalloc_kernel");
  }
  if (!rc && cum) rc = hb::check_cuda(ctx, cudaMemcpyAsync(cum, d + l.cum, 8 * (size_t)n, cudaMemcpyDeviceToHost, ctx->stream), "D2H cum");
  if (!rc) rc = hb::check_cuda(ctx, cudaMemcpyAsync(result, d + l.cum + 8 * (size_t)(n - 1), 8, cudaMemcpyDeviceToHost, ctx->stream), "D2H result");
  if (!rc) rc = hb::check_cuda(ctx, cudaStreamSynchronize(ctx->stream), "sync");
  return rc;
}
