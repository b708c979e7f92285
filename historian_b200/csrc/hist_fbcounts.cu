// historian_b200: Forward-Backward expected counts of a filled pair on device.
//
// Replaces the numeric part of BackwardMatrix::getCounts (reference forward.cpp:1183-1214):
//   * posterior-weighted transitionEigenCounts (forward.cpp:579-652) summed over every
//     (destination cell, state, incoming transition) -> six indel counts;
//   * the posterior weight exp(F + B - lpEnd) of every (cell, state), reduced into the weights of the
//     alignment columns the reference runs a sum-product over (getAlignmentColumn, forward.cpp:940-973):
//       XI[i]  x.alignColumn(i)                    IIW cells, and IMD / IMM cells on a null x state
//       XD[i]  x.alignColumn(i) + wildcard parent  IMD cells on an emitting x state (ypos > 0)
//       YI[j], YD[j]                               the same for y (IMI / IDM)
//       one column per IMM cell with both states emitting (x column + y column + wildcard parent).
//     The host appends those columns to a weighted hist_sumprod_columns batch.
//
// The reference caches the insert-column counts per x (y) state in xInsertCounts (yInsertCounts),
// filled by the FIRST cell of that row (column) it visits (forward.cpp:903-918).  When that cell sits on
// ypos == 0 (xpos == 0) its column is empty (forward.cpp:942) and the cached - empty - counts are reused
// for the whole row (column).  XI[i] (YI[j]) is therefore zero whenever (i,0) ((0,j)) lies inside the
// envelope; this is reproduced, not repaired, so results stay identical to the reference's.
//
// Profile transitions are taken to carry no counts of their own (leaf profiles, and every profile
// built by hist_profile_eliminate, which does not propagate counts yet).
//
// One block per x state (pass 0) or y state (pass 1); threads stride the other coordinate; fixed-order
// shuffle + shared-memory reductions, so results are run-to-run deterministic.
#include <algorithm>
#include <cmath>
#include <vector>

#include "hist_dpgraph.cuh"

namespace hb {

constexpr int FB_THREADS = 128;

struct FbCell {
  double w;
  int x, y;
};

struct FbWaits {
  double lt, rt, l_ins, l_del, r_ins, r_del;
};

// transitionEigenCounts (forward.cpp:579-652) for transitions that carry no counts themselves:
// acc = {ins, del, insExt, delExt, insTime, delTime}
__device__ __forceinline__ void add_transition_counts(double* acc, const FbWaits& W, int src, int dest, bool xNull, bool yNull, double w) {
  switch (dest) {
    case HIST_IMM:
      if (!xNull && !yNull) {
        if (src == HIST_IMM || src == HIST_IMD) { acc[4] += w * W.lt; acc[5] += w * W.lt; }
        if (src == HIST_IMM || src == HIST_IDM) { acc[4] += w * W.rt; acc[5] += w * W.rt; }
      }
      break;
    case HIST_IMD:
      if (!xNull) {
        if (src == HIST_IMM || src == HIST_IMD) { acc[4] += w * W.lt; acc[5] += w * W.lt; }
        if (src == dest) acc[3] += w;
        else { acc[1] += w; acc[5] += w * W.r_del; }
      }
      break;
    case HIST_IIW:
      if (!xNull) {
        if (src == dest) acc[2] += w;
        else { acc[0] += w; acc[4] += w * W.l_ins; }
      }
      break;
    case HIST_IDM:
      if (!yNull) {
        if (src == HIST_IMM || src == HIST_IDM) { acc[4] += w * W.rt; acc[5] += w * W.rt; }
        if (src == dest) acc[3] += w;
        else { acc[1] += w; acc[5] += w * W.l_del; }
      }
      break;
    case HIST_IMI:
      if (!yNull) {
        if (src == dest) acc[2] += w;
        else { acc[0] += w; acc[4] += w * W.r_ins; }
      }
      break;
    default: break;
  }
}

template <int NV>
__device__ __forceinline__ void block_reduce(double* v, double (*sm)[NV]) {
#pragma unroll
  for (int k = 0; k < NV; ++k)
#pragma unroll
    for (int o = 16; o; o >>= 1) v[k] += __shfl_down_sync(0xffffffffu, v[k], o);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0)
#pragma unroll
    for (int k = 0; k < NV; ++k) sm[warp][k] = v[k];
  __syncthreads();
  if (threadIdx.x == 0)
    for (int w = 1; w < FB_THREADS / 32; ++w)
#pragma unroll
      for (int k = 0; k < NV; ++k) v[k] += sm[w][k];
}

// PASS 0: block = x state i; indel partials [i][6], w_x[0][i] = XI, w_x[1][i] = XD, IMM cell list.
// PASS 1: block = y state j; w_y[0][j] = YI, w_y[1][j] = YD.
template <int PASS>
__global__ void __launch_bounds__(FB_THREADS) fb_counts_kernel(const PairDev* __restrict__ Pp, double lp_end, FbWaits W,
                                                               double* __restrict__ indel_part, double* __restrict__ w_side,
                                                               FbCell* __restrict__ cells, unsigned long long capacity,
                                                               unsigned long long* __restrict__ n_cells) {
  const PairDev& P = *Pp;
  __shared__ double sm[FB_THREADS / 32][8];
  const int xs = P.x.n - 1, ys = P.y.n - 1;  // cells span [0, xSize-2] x [0, ySize-2]
  const int fixed = blockIdx.x, span = PASS == 0 ? ys : xs;
  double acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};  // 0-5 indel, 6 insert-column weight, 7 delete-column weight
  for (int o = threadIdx.x; o < span; o += FB_THREADS) {
    const int i = PASS == 0 ? fixed : o, j = PASS == 0 ? o : fixed;
    if (!dp_in_envelope(P, i, j)) continue;
    const bool xNull = !(P.x.flags[i] & F_EMIT), yNull = !(P.y.flags[j] & F_EMIT);
#pragma unroll 1
    for (int s = 0; s < 5; ++s) {
      const double B = dp_cell(P, true, i, j, s);
      if (B == -INFINITY) continue;
      const double F = dp_cell(P, false, i, j, s);
      const double post = F == -INFINITY ? 0.0 : exp(F + B - lp_end);  // forward.cpp:1204
      if (PASS == 0) {
        if (s == HIST_IIW || ((s == HIST_IMD || s == HIST_IMM) && xNull)) acc[6] += post;
        else if (s == HIST_IMD) { if (j > 0) acc[7] += post; }
        else if (s == HIST_IMM && !yNull && post > 0.0) {
          const unsigned long long k = atomicAdd(n_cells, 1ull);
          if (k < capacity) cells[k] = {post, i, j};
        }
        // posterior-weighted transition counts (forward.cpp:1205-1207)
        const DCell d = {i, j, s};
        const SrcEnum e = source_enum(P, d, nullptr);
        if (e.n) {
          const double base = dp_emit(P, d) + B - lp_end;
          for (int c = 0; c < e.n; ++c) {
            DCell src;
            double lp;
            source_get(P, P.Wlog, d, e, c, src, lp);
            if (lp == -INFINITY) continue;
            const double Fs = dp_cell(P, false, src.x, src.y, src.s);
            if (Fs == -INFINITY) continue;
            add_transition_counts(acc, W, src.s, s, xNull, yNull, exp(Fs + lp + base));
          }
        }
      } else {
        if (s == HIST_IMI || (s == HIST_IDM && yNull) || (s == HIST_IMM && yNull && !xNull)) acc[6] += post;
        else if (s == HIST_IDM) { if (i > 0) acc[7] += post; }
      }
    }
  }
  block_reduce<8>(acc, sm);
  if (threadIdx.x == 0) {
    // the reference's per-row cache holds an empty column when its first cell is on the matrix edge
    const bool edge = fixed == 0 || (PASS == 0 ? dp_in_envelope(P, fixed, 0) : dp_in_envelope(P, 0, fixed));
    const int n = PASS == 0 ? P.x.n : P.y.n;
    w_side[fixed] = edge ? 0.0 : acc[6];
    w_side[n + fixed] = fixed == 0 ? 0.0 : acc[7];
    if (PASS == 0)
      for (int k = 0; k < 6; ++k) indel_part[(size_t)fixed * 6 + k] = acc[k];
  }
}

}  // namespace hb

using namespace hb;

extern "C" int hist_fb_counts(hist_pair* pair, const double* waits, double* indel, double* w_x, double* w_y, hist_cell* cells,
                              double* cell_w, int64_t capacity, int64_t* n_cells) {
  if (!pair || !waits || !indel || !w_x || !w_y || !n_cells) return HIST_ERR_ARG;
  hist_ctx* ctx = pair->ctx;
  *n_cells = 0;
  if (!pair->fwd_done || !pair->bwd_done) return set_err(ctx, HIST_ERR_STATE, "hist_fb_counts: needs both the Forward and the Backward fill");
  if (!(pair->lp_end > -INFINITY)) return set_err(ctx, HIST_ERR_NUMERIC, "hist_fb_counts: zero forward likelihood");
  cudaSetDevice(ctx->device);
  const int xn = pair->dev.x.n, yn = pair->dev.y.n;
  { int rc_ = hb::pair_cells(pair, nullptr); if (rc_) return rc_; }
  const unsigned long long cap = (unsigned long long)std::max<int64_t>(pair->info.n_cells, 1);
  const size_t n_d = (size_t)(xn - 1) * 6 + 2 * (size_t)xn + 2 * (size_t)yn;
  double* d_buf = nullptr;
  HB_CUDA(ctx, cudaMallocFromPoolAsync((void**)&d_buf, n_d * sizeof(double) + cap * sizeof(FbCell) + 16, ctx->pool, ctx->stream));
  double* d_part = d_buf;
  double* d_wx = d_part + (size_t)(xn - 1) * 6;
  double* d_wy = d_wx + 2 * (size_t)xn;
  FbCell* d_cells = (FbCell*)(d_wy + 2 * (size_t)yn);
  unsigned long long* d_cnt = (unsigned long long*)(d_cells + cap);
  HB_CUDA(ctx, cudaMemsetAsync(d_buf, 0, n_d * sizeof(double), ctx->stream));
  HB_CUDA(ctx, cudaMemsetAsync(d_cnt, 0, sizeof(unsigned long long), ctx->stream));
  const FbWaits W = {waits[0], waits[1], waits[2], waits[3], waits[4], waits[5]};
  fb_counts_kernel<0><<<xn - 1, FB_THREADS, 0, ctx->stream>>>(pair->d_dev, pair->lp_end, W, d_part, d_wx, d_cells, cap, d_cnt);
  fb_counts_kernel<1><<<yn - 1, FB_THREADS, 0, ctx->stream>>>(pair->d_dev, pair->lp_end, W, nullptr, d_wy, nullptr, 0, nullptr);
  ctx->launches += 2;
  HB_CUDA(ctx, cudaGetLastError());
  std::vector<double> h(n_d);
  unsigned long long found = 0;
  HB_CUDA(ctx, cudaMemcpyAsync(h.data(), d_buf, n_d * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
  HB_CUDA(ctx, cudaMemcpyAsync(&found, d_cnt, sizeof(found), cudaMemcpyDeviceToHost, ctx->stream));
  HB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  for (int k = 0; k < 6; ++k) indel[k] = 0;
  for (int i = 0; i < xn - 1; ++i)
    for (int k = 0; k < 6; ++k) indel[k] += h[(size_t)i * 6 + k];
  std::copy(h.begin() + (size_t)(xn - 1) * 6, h.begin() + (size_t)(xn - 1) * 6 + 2 * (size_t)xn, w_x);
  std::copy(h.begin() + (size_t)(xn - 1) * 6 + 2 * (size_t)xn, h.end(), w_y);
  *n_cells = (int64_t)found;
  int rc = HIST_OK;
  if ((int64_t)found > capacity || (found && (!cells || !cell_w))) {
    rc = set_err(ctx, HIST_ERR_CAPACITY, "hist_fb_counts: need room for %llu match cells", found);
  } else if (found) {
    std::vector<FbCell> hc(found);
    HB_CUDA(ctx, cudaMemcpyAsync(hc.data(), d_cells, found * sizeof(FbCell), cudaMemcpyDeviceToHost, ctx->stream));
    HB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    // the compaction order is arbitrary; sort into the reference's visiting order (xpos, ypos)
    std::sort(hc.begin(), hc.end(), [](const FbCell& a, const FbCell& b) { return a.x != b.x ? a.x < b.x : a.y < b.y; });
    for (unsigned long long k = 0; k < found; ++k) {
      cells[k] = {hc[k].x, hc[k].y, HIST_IMM};
      cell_w[k] = hc[k].w;
    }
  }
  cudaFreeAsync(d_buf, ctx->stream);
  return rc;
}
