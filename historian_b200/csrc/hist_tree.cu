// historian_b200: guide-tree estimation from a distance matrix (second half).
//
// hist_build_tree is the B200 counterpart of Tree::buildByUPGMA (reference tree.cpp:362-454, the default of
// `historian recon`, recon.cpp:737-742) and Tree::buildByNeighborJoining (tree.cpp:240-352): n - 2 rounds of "find the
// closest (compensated) pair of active nodes, join them under a new node, compute the new node's distances", O(n^3)
// in the reference's nested loops over a std::set.
//
// Design.  The rounds are sequential, the work inside a round is not: one cooperative grid (a CTA per SM) stays resident for
// the whole job and meets at grid-wide barriers, two per round (three for neighbour-joining).  The active nodes are an
// ascending list every CTA keeps in its shared memory (a join removes two entries and appends the new, largest index).
// The rows of the pair scan are dealt out to the CTAs, a row's columns to the threads (consecutive active nodes are
// mostly consecutive indices: coalesced); every thread keeps its best (value, i, j), ties going to the lexicographically
// first pair -- exactly the pair the reference's nested loops with a strict "<" keep; warp shuffles, one shared-memory
// pass and one pass over the CTAs' candidates pick the winner.  Neighbour-joining's average distances are summed per
// node over the active nodes in ascending order, like the reference's loop.  All arithmetic is FP64 adds, subtractions, halvings and one division,
// in the reference's order: node indices, joins and branch lengths are bit-identical.
#include <algorithm>
#include <cmath>
#include <vector>

#include <cooperative_groups.h>

#include "hist_internal.cuh"

namespace cg = cooperative_groups;

namespace hb {

#define TR_THREADS 1024
#define TR_MAX_LEAVES 4096

struct TreeCand {
  double v;
  int i, j;
};
// (v, i, j) before (w, k, l): smaller value, ties to the lexicographically first pair -- the pair the reference's nested
// loops with a strict "<" keep (tree.cpp:276-293, 387-401)
__device__ __forceinline__ bool tr_before(double v, int i, int j, double w, int k, int l) {
  if (i < 0) return false;
  if (k < 0) return true;
  return v < w || (v == w && (i < k || (i == k && j < l)));
}

// Cooperative launch, one CTA per SM.  dist: [M][M], M = 2n - 1, the n x n input in its top-left corner; aux: [M] node
// heights (UPGMA) or average distances (NJ); cand: [gridDim.x]; child: [M][2]; blen: [M].  Dynamic shared memory: two
// copies of the ascending list of active nodes (2 M ints), kept by every CTA for itself.
template <bool NJ>
__global__ void __launch_bounds__(TR_THREADS) tree_kernel(double* __restrict__ dist, int n, int M, double min_len, double* __restrict__ aux,
                                                         TreeCand* __restrict__ cand, int* __restrict__ child, double* __restrict__ blen) {
  cg::grid_group grid = cg::this_grid();
  extern __shared__ int s_lists[];
  __shared__ double rV[TR_THREADS / 32];
  __shared__ int rI[TR_THREADS / 32], rJ[TR_THREADS / 32];
  __shared__ int s_i, s_j;
  const int tid = threadIdx.x, T = TR_THREADS, G = gridDim.x;
  const long long gtid = (long long)blockIdx.x * T + tid, GT = (long long)G * T;
  int* act = s_lists;
  int* act2 = s_lists + M;
  for (int k = tid; k < M; k += T) act[k] = k;
  for (long long k = gtid; k < M; k += GT) {
    aux[k] = 0.0;
    child[2 * k] = child[2 * k + 1] = -1;
    blen[k] = 0.0;
  }
  grid.sync();
  int nAct = n, next = n;   // next: index of the node the next join creates
  while (nAct > 2) {
    if (NJ) {
      // average distance of every active node, summed in ascending order like the reference (tree.cpp:265-274)
      for (long long p = gtid; p < nAct; p += GT) {
        const int ni = act[p];
        double a = 0;
        for (int q = 0; q < nAct; ++q) {
          const int nj = act[q];
          if (nj != ni) a += dist[(size_t)nj * M + ni];
        }
        aux[ni] = a / (double)(nAct - 2);
      }
      grid.sync();
    }
    // ---- the closest (compensated) pair: rows dealt out to the CTAs, a row's columns to the threads
    double bv = 0;
    int bi = -1, bj = -1;
    for (int p = blockIdx.x; p < nAct - 1; p += G) {
      const int ni = act[p];
      const double ai = NJ ? aux[ni] : 0.0;
      const double* row = dist + (size_t)ni * M;
      for (int q = p + 1 + tid; q < nAct; q += T) {
        const int nj = act[q];
        const double d = row[nj];
        const double v = NJ ? d - ai - aux[nj] : d;
        if (tr_before(v, ni, nj, bv, bi, bj)) { bv = v; bi = ni; bj = nj; }
      }
    }
    for (int o = 16; o; o >>= 1) {
      const double ov = __shfl_xor_sync(0xffffffffu, bv, o);
      const int oi = __shfl_xor_sync(0xffffffffu, bi, o), oj = __shfl_xor_sync(0xffffffffu, bj, o);
      if (tr_before(ov, oi, oj, bv, bi, bj)) { bv = ov; bi = oi; bj = oj; }
    }
    if ((tid & 31) == 0) { rV[tid >> 5] = bv; rI[tid >> 5] = bi; rJ[tid >> 5] = bj; }
    __syncthreads();
    if (tid == 0) {
      for (int w = 1; w < T / 32; ++w)
        if (tr_before(rV[w], rI[w], rJ[w], bv, bi, bj)) { bv = rV[w]; bi = rI[w]; bj = rJ[w]; }
      cand[blockIdx.x].v = bv; cand[blockIdx.x].i = bi; cand[blockIdx.x].j = bj;
    }
    grid.sync();
    if (tid == 0) {   // every CTA picks the winner for itself
      double mv = 0;
      int mi = -1, mj = -1;
      for (int g = 0; g < G; ++g) {
        const TreeCand c = cand[g];
        if (tr_before(c.v, c.i, c.j, mv, mi, mj)) { mv = c.v; mi = c.i; mj = c.j; }
      }
      s_i = mi; s_j = mj;
    }
    __syncthreads();
    const int mi = s_i, mj = s_j, k = next;
    const double d_ij = dist[(size_t)mi * M + mj];
    double d_ik, d_jk, h_k = 0;
    if (NJ) {
      d_ik = 0.5 * (d_ij + aux[mi] - aux[mj]);
      d_jk = d_ij - d_ik;
      // Kuhner-Felsenstein correction and minimum branch lengths (tree.cpp:305-316)
      if (d_ik < min_len) { d_jk -= d_ik - min_len; d_ik = min_len; }
      if (d_jk < 0) { d_ik -= d_jk - min_len; d_jk = min_len; }
    } else {
      const double hi = aux[mi], hj = aux[mj];
      const double a = hi + min_len, b = hj + min_len, c = (hi + hj + d_ij) / 2;
      const double bc = b < c ? c : b;           // std::max
      h_k = a < bc ? bc : a;
      d_ik = h_k - hi;
      d_jk = h_k - hj;
    }
    // distances of the new node (tree.cpp:298-300, 318-319, 410-413), for every existing node like the reference
    for (long long m = gtid; m <= k; m += GT) {
      double v;
      if (m == k) v = 0;
      else if (m == mi) v = d_ik;
      else if (m == mj) v = d_jk;
      else {
        const double a = dist[(size_t)mi * M + m], b = dist[(size_t)mj * M + m];
        v = NJ ? 0.5 * (a + b - d_ij) : (a + b) / 2;
      }
      dist[(size_t)k * M + m] = v;
      dist[(size_t)m * M + k] = v;
    }
    if (gtid == 0) {
      if (!NJ) aux[k] = h_k;
      child[2 * k] = mi; child[2 * k + 1] = mj;
      blen[mi] = 0. < d_ik ? d_ik : 0.;            // max (0., d_ik)
      blen[mj] = 0. < d_jk ? d_jk : 0.;
    }
    // the active list stays ascending: drop mi and mj (mi < mj), append k
    for (int p = tid; p < nAct; p += T) {
      const int v = act[p];
      if (v != mi && v != mj) act2[p - (v > mi ? 1 : 0) - (v > mj ? 1 : 0)] = v;
    }
    if (tid == 0) act2[nAct - 2] = k;
    { int* t = act; act = act2; act2 = t; }
    --nAct;
    ++next;
    grid.sync();
  }
  if (gtid == 0) {
    // the root joins the last two nodes (tree.cpp:331-345, 432-446)
    const int i = act[0], j = act[1], k = next;
    child[2 * k] = i; child[2 * k + 1] = j;
    if (NJ) {
      const double dij = dist[(size_t)i * M + j];
      const double d = dij < 0. ? 0. : dij;        // max (dist[i][j], 0.)
      blen[i] = 0. < d / 2 ? d / 2 : 0.;
      blen[j] = blen[i];
    } else {
      const double hi = aux[i], hj = aux[j];
      const double a = hi + min_len, b = hj + min_len, c = (hi + hj + dist[(size_t)i * M + j]) / 2;
      const double bc = b < c ? c : b, hk = a < bc ? bc : a;
      blen[i] = 0. < hk - hi ? hk - hi : 0.;
      blen[j] = 0. < hk - hj ? hk - hj : 0.;
    }
  }
}

}  // namespace hb

using namespace hb;

extern "C" int hist_build_tree(hist_ctx* ctx, int32_t n, const double* dist, int32_t method, double min_branch_length, int32_t* child,
                               double* branch_length) {
  if (!ctx || n < 2 || !dist || !child || !branch_length || (method != 0 && method != 1)) return set_err(ctx, HIST_ERR_ARG, "hist_build_tree: bad arguments");
  if (n > TR_MAX_LEAVES) return set_err(ctx, HIST_ERR_ARG, "hist_build_tree: %d leaves exceed the limit of %d", n, TR_MAX_LEAVES);
  cudaSetDevice(ctx->device);
  const int M = 2 * n - 1;
  const int G = ctx->sm_count;   // one CTA per SM: the cooperative launch needs every CTA resident
  auto up = [](size_t v) { return (v + 255) / 256 * 256; };
  const size_t o_dist = 0, o_aux = up((size_t)M * M * 8), o_cand = o_aux + up((size_t)M * 8), o_child = o_cand + up((size_t)G * sizeof(TreeCand)),
               o_blen = o_child + up((size_t)M * 8), total = o_blen + up((size_t)M * 8);
  const size_t smem = (size_t)2 * M * sizeof(int);
  const void* fn = method == 1 ? (const void*)tree_kernel<true> : (const void*)tree_kernel<false>;
  if (smem > 48 * 1024) HB_CUDA(ctx, cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  int per_sm = 0;
  HB_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, TR_THREADS, smem));
  if (per_sm < 1) return set_err(ctx, HIST_ERR_CUDA, "hist_build_tree: the kernel does not fit one CTA per SM");
  char* d = nullptr;
  HB_CUDA(ctx, cudaMallocFromPoolAsync((void**)&d, total, ctx->pool, ctx->stream));
  int rc = check_cuda(ctx, cudaMemsetAsync(d, 0, o_aux, ctx->stream), "memset tree distances");
  if (!rc) rc = check_cuda(ctx, cudaMemcpy2DAsync(d + o_dist, (size_t)M * 8, dist, (size_t)n * 8, (size_t)n * 8, n, cudaMemcpyHostToDevice, ctx->stream), "H2D distance matrix");
  ctx->h2d_bytes += (long long)n * n * 8;
  if (!rc) {
    double* a_dist = (double*)(d + o_dist);
    int a_n = n, a_M = M;
    double a_min = min_branch_length;
    double* a_aux = (double*)(d + o_aux);
    TreeCand* a_cand = (TreeCand*)(d + o_cand);
    int* a_child = (int*)(d + o_child);
    double* a_blen = (double*)(d + o_blen);
    void* args[] = {&a_dist, &a_n, &a_M, &a_min, &a_aux, &a_cand, &a_child, &a_blen};
    rc = check_cuda(ctx, cudaLaunchCooperativeKernel(fn, dim3(G), dim3(TR_THREADS), args, smem, ctx->stream), "tree_kernel (cooperative launch)");
    ctx->launches += 1;
  }
  if (!rc) rc = check_cuda(ctx, cudaMemcpyAsync(child, d + o_child, (size_t)M * 2 * 4, cudaMemcpyDeviceToHost, ctx->stream), "D2H tree children");
  if (!rc) rc = check_cuda(ctx, cudaMemcpyAsync(branch_length, d + o_blen, (size_t)M * 8, cudaMemcpyDeviceToHost, ctx->stream), "D2H branch lengths");
  if (!rc) rc = check_cuda(ctx, cudaStreamSynchronize(ctx->stream), "sync after tree building");
  cudaFreeAsync(d, ctx->stream);
  return rc;
}
