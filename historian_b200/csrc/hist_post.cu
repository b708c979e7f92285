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
  const unsigned long long cap = 5ull * (unsigned long long)std::max<int64_t>(pair->info.n_cells, 1);
  PostCell* buf = nullptr;
  unsigned long long* cnt = nullptr;
  HB_CUDA(ctx, cudaMallocAsync((void**)&buf, cap * sizeof(PostCell) + 16, ctx->stream));
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
