// historian_b200: the parent-child (branch) alignment of the refinement stage.
//
// hist_branch_align is the B200 counterpart of Refiner::BranchMatrix (reference refiner.cpp:10-106) over
// TreeAlignFuncs::BranchMatrixBase / SparseDPMatrix<3> (sampler.h:64-210, sampler.cpp:452-476, 1005-1031, 1162-1184): a
// three-state (Match / Insert / Delete) Viterbi alignment of two position weight matrices -- the conditional
// residue distributions of a parent and a child node given the rest of the tree (hist_sumprod_excluded_posteriors) --
// under the branch's ProbModel, inside the envelope of the current alignment, and its traceback.
//
// Design.  Refinement is a serial chain of such alignments (every move starts from the alignment the previous one
// produced), so one alignment is one small problem: a preparation kernel premultiplies the child's matrix by the branch's
// substitution matrix (preMultiply) and sums its insertion scores (calcInsProbs), a second one computes the match scores of
// every cell of the envelope in parallel (they do not depend on the recurrence), one CTA then sweeps the anti-diagonals
// x + y of the (xLen + 1) x (yLen + 1) matrix -- Delete comes from (x-1, y) and Insert from (x, y-1) on the previous
// anti-diagonal, Match from (x-1, y-1) two back -- one thread per cell, one barrier per step, cells outside the envelope
// written as -inf, and thread 0 walks the traceback with the reference's own candidate order (Match, Insert, Delete; the
// first strict maximum).  The Viterbi recurrences are the reference's expressions on the same doubles; the emission
// scores are exact log-sum-exps where the reference uses its lookup table.
#include <cmath>
#include <cstring>
#include <vector>

#include "hist_internal.cuh"

namespace hb {

struct BrDev {
  int A, C, xS, yS, maxd;
  const double* x_pwm;     // [xS-1][C][A]
  const double* y_sub;     // [yS-1][C][A]
  const double* y_emit;    // [yS-1]
  const int* kx;           // [xS] envelope key of every parent position
  const int* ky;           // [yS]
  const double* lm;        // [xS][yS] logMatch of the cells inside the envelope (x, y > 0), filled by branch_match_kernel
  double lpt[3][4];        // log transProb(src, dest), dest 3 = End
};

__device__ __forceinline__ double br_lse(double a, double b) {
  if (a == -INFINITY) return b;
  if (b == -INFINITY) return a;
  const double m = a < b ? b : a, d = a < b ? a - b : b - a;
  return m + log1p(exp(d));
}
__device__ __forceinline__ double br_max3(double a, double b, double c) {   // max (max (a, b), c)
  const double ab = a < b ? b : a;
  return ab < c ? c : ab;
}

// preMultiply (sampler.cpp:452-463) and calcInsProbs (sampler.cpp:465-476); grid over child positions
__global__ void branch_prep_kernel(int yLen, int C, int A, const double* __restrict__ y_pwm, const double* __restrict__ log_sub,
                                   const double* __restrict__ log_ins, const double* __restrict__ log_cw, double* __restrict__ y_sub,
                                   double* __restrict__ y_emit) {
  const int y = blockIdx.x;
  if (y >= yLen) return;
  const double* lpp = y_pwm + (size_t)y * C * A;
  for (int q = threadIdx.x; q < C * A; q += blockDim.x) {
    const int c = q / A, i = q % A;
    double v = -INFINITY;
    for (int j = 0; j < A; ++j) v = br_lse(v, log_sub[((size_t)c * A + i) * A + j] + lpp[c * A + j]);
    y_sub[(size_t)y * C * A + q] = v;
  }
  if (threadIdx.x == 0) {
    double v = -INFINITY;
    for (int c = 0; c < C; ++c)
      for (int i = 0; i < A; ++i) v = br_lse(v, log_cw[c] + log_ins[c * A + i] + lpp[c * A + i]);
    y_emit[y] = v;
  }
}

// logMatch (sampler.h:203-205): logInnerProduct over components of the inner products over residues, as one
// max-shifted sum (independent exponentials instead of a dependent chain of two-term log-sum-exps)
__device__ double br_log_match(const BrDev& B, int x, int y) {
  const double* xv = B.x_pwm + (size_t)(x - 1) * B.C * B.A;
  const double* yv = B.y_sub + (size_t)(y - 1) * B.C * B.A;
  const int n = B.C * B.A;
  double m = -INFINITY;
  for (int q = 0; q < n; ++q) {
    const double v = xv[q] + yv[q];
    m = m < v ? v : m;
  }
  if (m == -INFINITY) return m;
  double s = 0;
  for (int q = 0; q < n; ++q) s += exp(xv[q] + yv[q] - m);
  return m + log(s);
}

__device__ __forceinline__ bool br_in_env(const BrDev& B, int x, int y) {   // SparseDPMatrix::inEnvelope, sampler.h:146-149
  if (x == 0 || y == 0 || x == B.xS - 1 || y == B.yS - 1 || B.maxd < 0) return true;
  const int d = B.kx[x] - B.ky[y];
  return (d < 0 ? -d : d) <= B.maxd;
}

// every cell of the envelope at once: the emission scores do not depend on the recurrence
__global__ void branch_match_kernel(BrDev B, double* __restrict__ lm) {
  const long long n = (long long)B.xS * B.yS;
  for (long long q = (long long)blockIdx.x * blockDim.x + threadIdx.x; q < n; q += (long long)gridDim.x * blockDim.x) {
    const int x = (int)(q / B.yS), y = (int)(q % B.yS);
    if (x > 0 && y > 0 && br_in_env(B, x, y)) lm[q] = br_log_match(B, x, y);
  }
}

// one CTA; cells [xS][yS][3]; out: lp_end, path (reversed into place by thread 0), path_len (-1: no traceback state found)
__global__ void __launch_bounds__(1024) branch_viterbi_kernel(BrDev B, double* cells, double* __restrict__ lp_end,
                                                             unsigned char* __restrict__ path, int* __restrict__ path_len) {
  const int xS = B.xS, yS = B.yS, tid = threadIdx.x, T = blockDim.x;
  const double mm = B.lpt[0][0], mi = B.lpt[0][1], md = B.lpt[0][2], me = B.lpt[0][3];
  const double im = B.lpt[1][0], ii = B.lpt[1][1], id = B.lpt[1][2], ie = B.lpt[1][3];
  const double dm = B.lpt[2][0], dd = B.lpt[2][2], de = B.lpt[2][3];
  for (int a = 0; a <= xS + yS - 2; ++a) {
    const int x0 = a < yS ? 0 : a - yS + 1, x1 = a < xS ? a : xS - 1;
    for (int x = x0 + tid; x <= x1; x += T) {
      const int y = a - x;
      double M = -INFINITY, I = -INFINITY, D = -INFINITY;
      if (br_in_env(B, x, y)) {
        if (x == 0 && y == 0) M = 0;   // lpStart
        if (x > 0) {
          const double* s = cells + ((size_t)(x - 1) * yS + y) * 3;
          D = br_max3(s[0] + md, s[1] + id, s[2] + dd);
        }
        if (y > 0) {
          const double* s = cells + ((size_t)x * yS + y - 1) * 3;
          const double a0 = s[0] + mi, a1 = s[1] + ii;
          I = B.y_emit[y - 1] + (a0 < a1 ? a1 : a0);
        }
        if (x > 0 && y > 0) {
          const double* s = cells + ((size_t)(x - 1) * yS + y - 1) * 3;
          M = B.lm[(size_t)x * yS + y] + br_max3(s[0] + mm, s[1] + im, s[2] + dm);
        }
      }
      double* d = cells + ((size_t)x * yS + y) * 3;
      d[0] = M; d[1] = I; d[2] = D;
    }
    __syncthreads();
  }
  if (tid != 0) return;
  const double* e = cells + ((size_t)(xS - 1) * yS + yS - 1) * 3;
  *lp_end = br_max3(e[0] + me, e[1] + ie, e[2] + de);
  // BranchMatrix::best (refiner.cpp:62-106)
  int x = xS - 1, y = yS - 1, state = 3, n = 0;
  while (x > 0 || y > 0) {
    bool bx = false, by = false;
    if (state == 0) { if (x > 0 && y > 0) bx = by = true; }
    else if (state == 1) by = true;
    else if (state == 2) bx = true;
    if (bx || by) path[n++] = (unsigned char)((bx ? 1 : 0) | (by ? 2 : 0));
    const int sx = x - (bx ? 1 : 0), sy = y - (by ? 1 : 0);
    if (sx < 0 || sy < 0) { *path_len = -1; return; }
    double em = 0;
    if (state == 0) em = (x > 0 && y > 0) ? B.lm[(size_t)x * yS + y] : -INFINITY;
    else if (state == 1) em = y > 0 ? B.y_emit[y - 1] : -INFINITY;
    const double* s = cells + ((size_t)sx * yS + sy) * 3;
    double best = -INFINITY;
    int bs = -1;
    for (int q = 0; q < 3; ++q) {
      const double v = s[q] + B.lpt[q][state] + em;
      if (v > best) { best = v; bs = q; }
    }
    if (bs < 0) { *path_len = -1; return; }
    x = sx; y = sy; state = bs;
  }
  for (int q = 0; q < n / 2; ++q) {
    const unsigned char t = path[q];
    path[q] = path[n - 1 - q];
    path[n - 1 - q] = t;
  }
  *path_len = n;
}

}  // namespace hb

using namespace hb;

extern "C" int hist_branch_align(hist_ctx* ctx, const hist_branch_model* model, int32_t x_len, const double* x_pwm, int32_t y_len,
                                 const double* y_pwm, int32_t max_distance, const int32_t* x_key, const int32_t* y_key, double* lp_end,
                                 uint8_t* path, int32_t* path_len) {
  if (!ctx || !model || x_len < 0 || y_len < 0 || (x_len && !x_pwm) || (y_len && !y_pwm) || !lp_end || !path || !path_len)
    return set_err(ctx, HIST_ERR_ARG, "hist_branch_align: bad arguments");
  if (max_distance >= 0 && (!x_key || !y_key)) return set_err(ctx, HIST_ERR_ARG, "hist_branch_align: an envelope needs x_key and y_key");
  const int A = model->alph_size, C = model->components;
  if (A < 1 || C < 1 || !model->log_sub || !model->log_ins || !model->log_cpt_weight) return set_err(ctx, HIST_ERR_ARG, "hist_branch_align: bad model");
  cudaSetDevice(ctx->device);
  const size_t CA = (size_t)C * A, xS = (size_t)x_len + 1, yS = (size_t)y_len + 1;
  // the matrix is stored whole (3 scores + 1 match score per cell): 32 bytes x (xLen + 1) x (yLen + 1)
  if ((double)xS * (double)yS * 32.0 > 64.0 * 1024 * 1024 * 1024)
    return set_err(ctx, HIST_ERR_CAPACITY, "hist_branch_align: a %d x %d matrix needs %.1f GB of device memory (limit 64 GB)", x_len, y_len,
                   (double)xS * (double)yS * 32.0 / (1024.0 * 1024 * 1024));
  auto up = [](size_t v) { return (v + 255) / 256 * 256; };
  // host blob: [x_pwm | y_pwm | log_sub | log_ins | log_cw | kx | ky]; device: blob + [y_sub | y_emit | cells | lp_end | path_len | path]
  const size_t o_x = 0, o_y = o_x + up(8 * x_len * CA), o_sub = o_y + up(8 * y_len * CA), o_ins = o_sub + up(8 * CA * A), o_cw = o_ins + up(8 * CA),
               o_kx = o_cw + up(8 * (size_t)C), o_ky = o_kx + up(4 * xS), blob_bytes = o_ky + up(4 * yS);
  const size_t o_ysub = blob_bytes, o_yemit = o_ysub + up(8 * y_len * CA), o_cells = o_yemit + up(8 * (size_t)y_len + 8), o_lm = o_cells + up(24 * xS * yS), o_lp = o_lm + up(8 * xS * yS),
               o_len = o_lp + 256, o_path = o_len + 256, total = o_path + up(xS + yS);
  std::vector<char> blob(blob_bytes, 0);
  if (x_len) memcpy(blob.data() + o_x, x_pwm, 8 * x_len * CA);
  if (y_len) memcpy(blob.data() + o_y, y_pwm, 8 * y_len * CA);
  memcpy(blob.data() + o_sub, model->log_sub, 8 * CA * A);
  memcpy(blob.data() + o_ins, model->log_ins, 8 * CA);
  memcpy(blob.data() + o_cw, model->log_cpt_weight, 8 * (size_t)C);
  if (max_distance >= 0) {
    memcpy(blob.data() + o_kx, x_key, 4 * xS);
    memcpy(blob.data() + o_ky, y_key, 4 * yS);
  }
  char* d = nullptr;
  HB_CUDA(ctx, cudaMallocFromPoolAsync((void**)&d, total, ctx->pool, ctx->stream));
  int rc = check_cuda(ctx, cudaMemcpyAsync(d, blob.data(), blob_bytes, cudaMemcpyHostToDevice, ctx->stream), "H2D branch alignment inputs");
  ctx->h2d_bytes += (long long)blob_bytes;
  if (!rc && y_len) {
    branch_prep_kernel<<<y_len, 128, 0, ctx->stream>>>(y_len, C, A, (const double*)(d + o_y), (const double*)(d + o_sub), (const double*)(d + o_ins),
                                                       (const double*)(d + o_cw), (double*)(d + o_ysub), (double*)(d + o_yemit));
    ctx->launches += 1;
  }
  if (!rc) {
    BrDev B;
    B.A = A; B.C = C; B.xS = (int)xS; B.yS = (int)yS; B.maxd = max_distance;
    B.x_pwm = (const double*)(d + o_x); B.y_sub = (const double*)(d + o_ysub); B.y_emit = (const double*)(d + o_yemit);
    B.kx = (const int*)(d + o_kx); B.ky = (const int*)(d + o_ky);
    for (int s = 0; s < 3; ++s)
      for (int t = 0; t < 4; ++t) B.lpt[s][t] = model->lp_trans[s][t];
    B.lm = (const double*)(d + o_lm);
    {
      const long long cells = (long long)xS * yS;
      const int blocks = (int)std::min<long long>((cells + 127) / 128, 148 * 16);
      branch_match_kernel<<<blocks, 128, 0, ctx->stream>>>(B, (double*)(d + o_lm));
      ctx->launches += 1;
    }
    const int threads = (int)std::min<size_t>(1024, (std::min(xS, yS) + 31) / 32 * 32);
    branch_viterbi_kernel<<<1, threads, 0, ctx->stream>>>(B, (double*)(d + o_cells), (double*)(d + o_lp), (unsigned char*)(d + o_path), (int*)(d + o_len));
    ctx->launches += 1;
    rc = check_cuda(ctx, cudaGetLastError(), "branch_viterbi_kernel");
  }
  int n = 0;
  std::vector<unsigned char> hp(xS + yS);
  if (!rc) rc = check_cuda(ctx, cudaMemcpyAsync(lp_end, d + o_lp, 8, cudaMemcpyDeviceToHost, ctx->stream), "D2H branch score");
  if (!rc) rc = check_cuda(ctx, cudaMemcpyAsync(&n, d + o_len, 4, cudaMemcpyDeviceToHost, ctx->stream), "D2H branch path length");
  if (!rc) rc = check_cuda(ctx, cudaMemcpyAsync(hp.data(), d + o_path, xS + yS, cudaMemcpyDeviceToHost, ctx->stream), "D2H branch path");
  if (!rc) rc = check_cuda(ctx, cudaStreamSynchronize(ctx->stream), "sync after branch alignment");
  cudaFreeAsync(d, ctx->stream);
  if (rc) return rc;
  if (n < 0) return set_err(ctx, HIST_ERR_NUMERIC, "hist_branch_align: could not find a traceback state (refiner.cpp:93)");
  memcpy(path, hp.data(), (size_t)n);
  *path_len = n;
  return HIST_OK;
}
