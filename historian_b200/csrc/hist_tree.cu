// historian_b200: guide-tree estimation from a distance matrix (second half).
//
// hist_build_tree is the B200 counterpart of Tree::buildByUPGMA (reference tree.cpp:362-454, the default of
// `historian recon`, recon.cpp:737-742) and Tree::buildByNeighborJoining (tree.cpp:240-352): n - 2 rounds of "find the
// closest (compensated) pair of active nodes, join them under a new node, compute the new node's distances", O(n^3)
// in the reference's nested loops over a std::set.
//
// Design.  The rounds are sequential, the work inside a round is not: ONE CTA of 1024 threads owns the whole job.  The
// active nodes are kept as an ascending list (joining removes two entries and appends the new, largest index).  Thread
// p takes the p-th active node i: (neighbour-joining only) its average distance, summed over the active nodes in
// ascending order like the reference's loop; then the first minimum over the active j > i of the (compensated) distance.
// A block-wide reduction picks the smallest value, ties going to the smallest i -- exactly the pair the reference's
// lexicographic scan with a strict "<" keeps.  The matrix is symmetric, so thread i reads column i (dist[j][i]):
// consecutive threads, consecutive addresses.  All arithmetic is FP64 adds, subtractions, halvings and one division,
// in the reference's order: node indices, joins and branch lengths are bit-identical.
#include <algorithm>
#include <cmath>
#include <vector>

#include "hist_internal.cuh"

namespace hb {

#define TR_THREADS 1024

// dist: [M][M], M = 2n - 1, the n x n input in its top-left corner; act: [M] scratch; aux: [M] node heights (UPGMA) or
// average distances (NJ); child: [M][2]; blen: [M]
template <bool NJ>
__global__ void __launch_bounds__(TR_THREADS) tree_kernel(double* __restrict__ dist, int n, int M, double min_len, int* __restrict__ act,
                                                         double* __restrict__ aux, int* __restrict__ child, double* __restrict__ blen) {
  __shared__ double rV[TR_THREADS];
  __shared__ int rI[TR_THREADS], rJ[TR_THREADS];
  __shared__ int s_i, s_j, s_n;
  const int tid = threadIdx.x;
  for (int k = tid; k < M; k += TR_THREADS) {
    act[k] = k;
    aux[k] = 0.0;
    child[2 * k] = child[2 * k + 1] = -1;
    blen[k] = 0.0;
  }
  if (tid == 0) s_n = n;
  __syncthreads();
  int next = n;   // index of the node the next join creates
  while (s_n > 2) {
    const int nAct = s_n;
    if (NJ) {
      // average distance of every active node (tree.cpp:265-274)
      for (int p = tid; p < nAct; p += TR_THREADS) {
        const int ni = act[p];
        double a = 0;
        for (int q = 0; q < nAct; ++q) {
          const int nj = act[q];
          if (nj != ni) a += dist[(size_t)nj * M + ni];
        }
        aux[ni] = a / (double)(nAct - 2);
      }
      __syncthreads();
    }
    // first minimum of the (compensated) distance over the pairs i < j in lexicographic order (tree.cpp:276-293, 387-401)
    double bv = 0;
    int bi = -1, bj = -1;
    for (int p = tid; p < nAct; p += TR_THREADS) {
      const int ni = act[p];
      const double ai = NJ ? aux[ni] : 0.0;
      for (int q = p + 1; q < nAct; ++q) {
        const int nj = act[q];
        const double d = dist[(size_t)nj * M + ni];
        const double v = NJ ? d - ai - aux[nj] : d;
        if (bi < 0 || v < bv) { bv = v; bi = ni; bj = nj; }   // strict "<": the first minimum of this thread's rows
      }
    }
    rV[tid] = bv; rI[tid] = bi; rJ[tid] = bj;
    __syncthreads();
    if (tid == 0) {
      double mv = 0;
      int mi = -1, mj = -1;
      for (int t = 0; t < TR_THREADS; ++t) {
        if (rI[t] < 0) continue;
        if (mi < 0 || rV[t] < mv || (rV[t] == mv && rI[t] < mi)) { mv = rV[t]; mi = rI[t]; mj = rJ[t]; }
      }
      s_i = mi; s_j = mj;
    }
    __syncthreads();
    const int mi = s_i, mj = s_j, k = next;
    const double d_ij = dist[(size_t)mi * M + mj];
    // distances of the new node (tree.cpp:298-300, 410-411), for every existing node like the reference
    for (int m = tid; m < k; m += TR_THREADS) {
      const double a = dist[(size_t)mi * M + m], b = dist[(size_t)mj * M + m];
      const double v = NJ ? 0.5 * (a + b - d_ij) : (a + b) / 2;
      dist[(size_t)k * M + m] = v;
      dist[(size_t)m * M + k] = v;
    }
    __syncthreads();
    if (tid == 0) {
      dist[(size_t)k * M + k] = 0;
      double d_ik, d_jk;
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
        aux[k] = a < bc ? bc : a;
        d_ik = aux[k] - hi;
        d_jk = aux[k] - hj;
      }
      dist[(size_t)mi * M + k] = dist[(size_t)k * M + mi] = d_ik;
      dist[(size_t)mj * M + k] = dist[(size_t)k * M + mj] = d_jk;
      child[2 * k] = mi; child[2 * k + 1] = mj;
      blen[mi] = 0. < d_ik ? d_ik : 0.;            // max (0., d_ik)
      blen[mj] = 0. < d_jk ? d_jk : 0.;
      // the active list stays ascending: drop mi and mj, append k
      int w = 0;
      for (int p = 0; p < nAct; ++p)
        if (act[p] != mi && act[p] != mj) act[w++] = act[p];
      act[w++] = k;
      s_n = w;
    }
    ++next;
    __syncthreads();
  }
  if (tid == 0) {
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
  cudaSetDevice(ctx->device);
  const int M = 2 * n - 1;
  auto up = [](size_t v) { return (v + 255) / 256 * 256; };
  const size_t o_dist = 0, o_act = up((size_t)M * M * 8), o_aux = o_act + up((size_t)M * 4), o_child = o_aux + up((size_t)M * 8),
               o_blen = o_child + up((size_t)M * 8), total = o_blen + up((size_t)M * 8);
  char* d = nullptr;
  HB_CUDA(ctx, cudaMallocAsync((void**)&d, total, ctx->stream));
  int rc = check_cuda(ctx, cudaMemsetAsync(d, 0, o_act, ctx->stream), "memset tree distances");
  if (!rc) rc = check_cuda(ctx, cudaMemcpy2DAsync(d + o_dist, (size_t)M * 8, dist, (size_t)n * 8, (size_t)n * 8, n, cudaMemcpyHostToDevice, ctx->stream), "H2D distance matrix");
  ctx->h2d_bytes += (long long)n * n * 8;
  if (!rc) {
    if (method == 1) tree_kernel<true><<<1, TR_THREADS, 0, ctx->stream>>>((double*)(d + o_dist), n, M, min_branch_length, (int*)(d + o_act), (double*)(d + o_aux), (int*)(d + o_child), (double*)(d + o_blen));
    else tree_kernel<false><<<1, TR_THREADS, 0, ctx->stream>>>((double*)(d + o_dist), n, M, min_branch_length, (int*)(d + o_act), (double*)(d + o_aux), (int*)(d + o_child), (double*)(d + o_blen));
    ctx->launches += 1;
    rc = check_cuda(ctx, cudaGetLastError(), "tree_kernel");
  }
  if (!rc) rc = check_cuda(ctx, cudaMemcpyAsync(child, d + o_child, (size_t)M * 2 * 4, cudaMemcpyDeviceToHost, ctx->stream), "D2H tree children");
  if (!rc) rc = check_cuda(ctx, cudaMemcpyAsync(branch_length, d + o_blen, (size_t)M * 8, cudaMemcpyDeviceToHost, ctx->stream), "D2H branch lengths");
  if (!rc) rc = check_cuda(ctx, cudaStreamSynchronize(ctx->stream), "sync after tree building");
  cudaFreeAsync(d, ctx->stream);
  return rc;
}
