// historian_b200: the distance-matrix stage.
//
// hist_distance_matrix is the B200 counterpart of RateModel::distanceMatrix (reference model.cpp:533-553): for every
// pair of rows of the guide alignment, mlDistance (model.cpp:506-531) counts the aligned residue pairs and
// DistanceMatrixParams::tML (model.cpp:583-655) finds the branch length that maximises their likelihood -- a
// Jukes-Cantor first guess, a coarse scan for a bracket, then golden-section steps (GSL's
// gsl_min_fminimizer_goldensection) to 1 %.  The reference evaluates up to ~60 matrix exponentials per pair, n(n-1)/2
// pairs, serially: the second bottleneck of `historian recon -seqs` for large n.
//
// Design.  One CTA per pair (i < j).  Counting: the threads stride over the alignment columns (coalesced int8 tokens) and
// bump a shared A x A histogram.  Likelihood: the non-zero entries are listed in (i, j) order (the reference's std::map
// order); for a trial t the threads evaluate P_t(i,j) = sum_c w_c Re sum_k V_c[i][k] exp(lambda_ck t) V_c^-1[k][j] from the
// model's eigen-decomposition (the sum-product path's "eigen-decomposed matrix exponentials") for their entries, and one
// thread adds count * log p in list order, like the reference's loop.  The search itself (model.cpp:583-655) is control
// flow every thread follows identically.  FP64 throughout.
//
// The reference's P_t comes from GSL's scaling-and-squaring exponential instead; the two agree to ~1e-13, so a
// comparison inside the search can only flip when two trial likelihoods tie to that precision, and the result moves by
// less than the search's own 1 % tolerance.
#include <cctype>
#include <cmath>
#include <cstring>
#include <vector>

#include "hist_internal.cuh"

namespace hb {

#define DM_THREADS 128
#define DM_MAX_A 64      // codon alphabets included (A = 61)
#define DM_MAX_CA 256    // components * alphabet size

struct DmModel {
  int A, C;
  const double* w;        // [C]
  const double* ev_re; const double* ev_im;          // [C][A]
  const double* v_re; const double* v_im;            // [C][A][A]  V[a][k]
  const double* vi_re; const double* vi_im;          // [C][A][A]  V^-1[k][b]
  int is_real;            // every eigenvalue and eigenvector entry has a zero imaginary part (reversible models)
  int staged;             // is_real and the real parts fit the shared-memory budget: V and V^-1 transposed are staged per CTA
};

__device__ __forceinline__ double dm_max(double a, double b) { return a < b ? b : a; }   // std::max
__device__ __forceinline__ double dm_min(double a, double b) { return b < a ? b : a; }   // std::min

// dynamic shared memory: [term A*A doubles | e_re, e_im C*A doubles | staged V, V^-1 transposed (optional) | counts, ent_idx, ent_cnt A*A ints]
struct DmShared {
  double* term; double* e_re; double* e_im;
  double* sv; double* svit;   // staged V[c][i][k] and V^-1 transposed [c][j][k], row stride A | 1 (bank-conflict padding)
  int* counts; int* ent_idx; int* ent_cnt;
  int n_ent;
  double f;
};
__host__ __device__ inline size_t dm_staged_doubles(int A, int C) { return 2 * (size_t)C * A * (A | 1); }
__host__ __device__ inline size_t dm_shared_bytes(int A, int C, int staged) {
  return ((size_t)A * A + 2 * (size_t)C * A + (staged ? dm_staged_doubles(A, C) : 0)) * sizeof(double) + 3 * (size_t)A * A * sizeof(int);
}

// distanceMatrixNegLogLike (model.cpp:555-569) for the CTA's pair; every thread returns the value
__device__ double dm_neg_loglike(const DmModel& M, DmShared& S, double t) {
  const int A = M.A, C = M.C, tid = threadIdx.x;
  __syncthreads();
  if (M.is_real) {
    // Real spectrum: the complex expressions below reduce to these, bit for bit (cos 0 = 1, every imaginary term is a
    // signed zero), with the roundings written out so that no contraction can change them.
    for (int k = tid; k < A * C; k += DM_THREADS) S.e_re[k] = exp(M.ev_re[k] * t);
    __syncthreads();
    const int ld = A | 1;
    for (int q = tid; q < S.n_ent; q += DM_THREADS) {
      const int i = S.ent_idx[q] / A, j = S.ent_idx[q] % A;
      double p = 0;
      for (int c = 0; c < C; ++c) {
        double acc = 0;
        if (M.staged) {
          const double* vr = S.sv + ((size_t)c * A + i) * ld;
          const double* ir = S.svit + ((size_t)c * A + j) * ld;
          for (int k = 0; k < A; ++k) acc = __dadd_rn(acc, __dmul_rn(__dmul_rn(vr[k], S.e_re[c * A + k]), ir[k]));
        } else {
          const double* vr = M.v_re + ((size_t)c * A + i) * A;
          const double* ir = M.vi_re + (size_t)c * A * A + j;
          for (int k = 0; k < A; ++k) acc = __dadd_rn(acc, __dmul_rn(__dmul_rn(vr[k], S.e_re[c * A + k]), ir[(size_t)k * A]));
        }
        p = __fma_rn(M.w[c], acc, p);   // as the compiler contracts `p += w * acc` below
      }
      S.term[q] = log(p) * (double)S.ent_cnt[q];
    }
    __syncthreads();
    if (tid == 0) {
      double ll = 0;
      for (int q = 0; q < S.n_ent; ++q) ll += S.term[q];
      S.f = -ll;
    }
    __syncthreads();
    return S.f;
  }
  for (int k = tid; k < A * C; k += DM_THREADS) {
    const double m = exp(M.ev_re[k] * t), ph = M.ev_im[k] * t;
    S.e_re[k] = m * cos(ph);
    S.e_im[k] = m * sin(ph);
  }
  __syncthreads();
  for (int q = tid; q < S.n_ent; q += DM_THREADS) {
    const int i = S.ent_idx[q] / A, j = S.ent_idx[q] % A;
    double p = 0;
    for (int c = 0; c < C; ++c) {
      const double* vr = M.v_re + ((size_t)c * A + i) * A;
      const double* vm = M.v_im + ((size_t)c * A + i) * A;
      const double* ir = M.vi_re + (size_t)c * A * A + j;
      const double* im = M.vi_im + (size_t)c * A * A + j;
      double acc = 0;
      for (int k = 0; k < A; ++k) {
        // Re( V[i][k] * e_k * Vinv[k][j] )
        const double ar = vr[k] * S.e_re[c * A + k] - vm[k] * S.e_im[c * A + k];
        const double ai = vr[k] * S.e_im[c * A + k] + vm[k] * S.e_re[c * A + k];
        acc += ar * ir[(size_t)k * A] - ai * im[(size_t)k * A];
      }
      p += M.w[c] * acc;
    }
    S.term[q] = log(p) * (double)S.ent_cnt[q];
  }
  __syncthreads();
  if (tid == 0) {
    double ll = 0;
    for (int q = 0; q < S.n_ent; ++q) ll += S.term[q];
    S.f = -ll;
  }
  __syncthreads();
  return S.f;
}

// grid = pairs (i < j) in row-major order of the upper triangle
__global__ void __launch_bounds__(DM_THREADS) distmat_kernel(DmModel M, const signed char* __restrict__ tok, int n_rows, int n_cols,
                                                            double sub_rate, int max_iter, double* __restrict__ dist) {
  extern __shared__ double dm_smem[];
  __shared__ DmShared S;
  const int tid = threadIdx.x, A = M.A;
  if (tid == 0) {
    S.term = dm_smem;
    S.e_re = S.term + A * A;
    S.e_im = S.e_re + M.C * A;
    S.sv = S.e_im + M.C * A;
    S.svit = S.sv + (M.staged ? (size_t)M.C * A * (A | 1) : 0);
    S.counts = reinterpret_cast<int*>(S.svit + (M.staged ? (size_t)M.C * A * (A | 1) : 0));
    S.ent_idx = S.counts + A * A;
    S.ent_cnt = S.ent_idx + A * A;
  }
  __syncthreads();
  if (M.staged) {
    const int ld = A | 1;
    for (int q = tid; q < M.C * A * A; q += DM_THREADS) {
      const int c = q / (A * A), r = (q / A) % A, k = q % A;
      S.sv[((size_t)c * A + r) * ld + k] = M.v_re[q];                                   // V[c][r][k]
      S.svit[((size_t)c * A + r) * ld + k] = M.vi_re[((size_t)c * A + k) * A + r];      // V^-1[c][k][r]
    }
  }
  // pair index -> (i, j)
  long long p = blockIdx.x;
  int i = 0;
  while (p >= n_rows - 1 - i) { p -= n_rows - 1 - i; ++i; }
  const int j = i + 1 + (int)p;
  for (int k = tid; k < A * A; k += DM_THREADS) S.counts[k] = 0;
  __syncthreads();
  const signed char* ti = tok + (size_t)i * n_cols;
  const signed char* tj = tok + (size_t)j * n_cols;
  for (int c = tid; c < n_cols; c += DM_THREADS) {
    const int a = ti[c], b = tj[c];
    if (a >= 0 && b >= 0) atomicAdd(&S.counts[a * A + b], 1);
  }
  __syncthreads();
  if (tid == 0) {
    int n = 0;
    for (int k = 0; k < A * A; ++k)
      if (S.counts[k]) { S.ent_idx[n] = k; S.ent_cnt[n] = S.counts[k]; ++n; }
    S.n_ent = n;
  }
  __syncthreads();
  // ---- tJC (model.cpp:571-581)
  int same = 0, diff = 0;
  for (int q = 0; q < S.n_ent; ++q)
    if (S.ent_idx[q] / A == S.ent_idx[q] % A) same += S.ent_cnt[q]; else diff += S.ent_cnt[q];
  const double pDiff = (double)diff / (double)(same + diff);
  const double dA = (double)A;
  double tjc_raw;
  if (pDiff >= (dA - 1) / dA) tjc_raw = INFINITY;
  else tjc_raw = -((dA - 1) / dA) * log(1 - (dA / (dA - 1)) * pDiff) / sub_rate;
  // ---- tML (model.cpp:583-655)
  const double tMin = 1e-9, tMax = 10;
  const double tjc = dm_min(tMax, dm_max(tMin, tjc_raw));
  double result;
  if (max_iter <= 0) {
    result = tjc;
  } else {
    const double tLower = dm_min(tMin, tjc / 2), tUpper = dm_max(tMax, tjc * 2);
    const double llLower = dm_neg_loglike(M, S, tLower), llUpper = dm_neg_loglike(M, S, tUpper);
    const double lljc = dm_neg_loglike(M, S, tjc);
    double t = 0;
    bool done = false;
    if (lljc < llLower && lljc < llUpper) {
      t = tjc;
    } else {
      bool found = false;
      double lo = tLower, hi = tUpper;
      const double nScanSteps = 4;
      while (!found && hi - lo > tLower) {
        const double step = (hi - lo) / nScanSteps;
        for (double x = lo; x < hi && !found; x += step) {
          const double ll = dm_neg_loglike(M, S, x);
          if (ll < llLower && ll < llUpper) { found = true; t = x; }
        }
        if (!found) {
          if (llLower < llUpper) hi = (lo + hi) / 2; else lo = (lo + hi) / 2;
        }
      }
      if (!found) {
        result = llLower < llUpper ? tLower : tUpper;
        done = true;
      }
    }
    if (!done) {
      // gsl_min_fminimizer_set + golden-section iterations (GSL min/golden.c), convergence to 1 %
      double x_min = t, x_lo = tLower, x_hi = tUpper;
      double f_min = dm_neg_loglike(M, S, x_min);
      (void)dm_neg_loglike(M, S, x_lo);
      (void)dm_neg_loglike(M, S, x_hi);
      for (int iter = 0; iter < max_iter; ++iter) {
        const double golden = 0.3819660;
        const double w_lower = x_min - x_lo, w_upper = x_hi - x_min;
        const double x_new = x_min + golden * ((w_upper > w_lower) ? w_upper : -w_lower);
        const double f_new = dm_neg_loglike(M, S, x_new);
        if (f_new < f_min) { x_min = x_new; f_min = f_new; }
        else if (x_new < x_min && f_new > f_min) x_lo = x_new;
        else if (x_new > x_min && f_new > f_min) x_hi = x_new;
        t = x_min;
        // gsl_min_test_interval (a, b, 0, 0.01)
        const double al = fabs(x_lo), au = fabs(x_hi);
        double min_abs = 0;
        if ((x_lo > 0 && x_hi > 0) || (x_lo < 0 && x_hi < 0)) min_abs = al < au ? al : au;
        if (fabs(x_hi - x_lo) < 0.01 * min_abs) break;
      }
      result = t;
    }
  }
  if (tid == 0) {
    dist[(size_t)i * n_rows + j] = result;
    dist[(size_t)j * n_rows + i] = result;
  }
}

}  // namespace hb

using namespace hb;

extern "C" int hist_distance_matrix(hist_ctx* ctx, const hist_model_desc* model, double expected_sub_rate, int32_t n_rows, int32_t n_cols,
                                    const char* rows, int32_t max_iterations, double* dist) {
  if (!ctx || !model || !rows || !dist || n_rows < 1 || n_cols < 0) return set_err(ctx, HIST_ERR_ARG, "hist_distance_matrix: bad arguments");
  const int A = model->alph_size, C = model->components;
  if (A < 2 || A > DM_MAX_A || C < 1 || A * C > DM_MAX_CA) return set_err(ctx, HIST_ERR_ARG, "hist_distance_matrix: alphabet of %d x %d components not supported (A <= %d, A*C <= %d)", A, C, DM_MAX_A, DM_MAX_CA);
  if (!model->alphabet || !model->cpt_weight || !model->eval_re || !model->eval_im || !model->evec_re || !model->evec_im || !model->evec_inv_re || !model->evec_inv_im)
    return set_err(ctx, HIST_ERR_ARG, "hist_distance_matrix: the model needs its alphabet and eigen-decomposition");
  if (!(expected_sub_rate > 0)) return set_err(ctx, HIST_ERR_ARG, "hist_distance_matrix: expected substitution rate must be positive");
  cudaSetDevice(ctx->device);
  for (size_t k = 0; k < (size_t)n_rows * n_rows; ++k) dist[k] = 0.0;
  if (n_rows < 2) return HIST_OK;
  // tokenize() (fastseq.cpp:10-16): the exact character first, then the other case; gaps, wildcards and anything else: -1
  signed char lut[256];
  for (int c = 0; c < 256; ++c) lut[c] = -1;
  for (int k = 0; k < A; ++k) {
    const unsigned char ch = (unsigned char)model->alphabet[k];
    const int other = isupper(ch) ? tolower(ch) : toupper(ch);
    if (lut[other] < 0) lut[other] = (signed char)k;
  }
  for (int k = A; k-- > 0;) lut[(unsigned char)model->alphabet[k]] = (signed char)k;
  lut[(unsigned char)'-'] = lut[(unsigned char)'.'] = lut[(unsigned char)'*'] = -1;   // Alignment::isGap / isWildcard (model.cpp:513)
  const size_t n_tok = (size_t)n_rows * n_cols;
  auto up = [](size_t v) { return (v + 255) / 256 * 256; };
  const size_t AA = (size_t)C * A * A, CA = (size_t)C * A;
  const size_t o_tok = 0, o_w = up(n_tok), o_ev = o_w + up(C * 8), o_v = o_ev + up(2 * CA * 8), o_vi = o_v + up(2 * AA * 8), o_dist = o_vi + up(2 * AA * 8),
               blob_bytes = o_dist, total = o_dist + up((size_t)n_rows * n_rows * 8);
  std::vector<char> blob(blob_bytes, 0);
  for (size_t k = 0; k < n_tok; ++k) blob[o_tok + k] = (char)lut[(unsigned char)rows[k]];
  memcpy(blob.data() + o_w, model->cpt_weight, C * 8);
  memcpy(blob.data() + o_ev, model->eval_re, CA * 8);
  memcpy(blob.data() + o_ev + CA * 8, model->eval_im, CA * 8);
  memcpy(blob.data() + o_v, model->evec_re, AA * 8);
  memcpy(blob.data() + o_v + AA * 8, model->evec_im, AA * 8);
  memcpy(blob.data() + o_vi, model->evec_inv_re, AA * 8);
  memcpy(blob.data() + o_vi + AA * 8, model->evec_inv_im, AA * 8);
  char* d = nullptr;
  HB_CUDA(ctx, cudaMallocFromPoolAsync((void**)&d, total, ctx->pool, ctx->stream));
  int rc = check_cuda(ctx, cudaMemcpyAsync(d, blob.data(), blob_bytes, cudaMemcpyHostToDevice, ctx->stream), "H2D distance-matrix inputs");
  ctx->h2d_bytes += (long long)blob_bytes;
  if (!rc) rc = check_cuda(ctx, cudaMemsetAsync(d + o_dist, 0, (size_t)n_rows * n_rows * 8, ctx->stream), "memset distances");
  if (!rc) {
    DmModel M;
    M.A = A; M.C = C;
    M.w = (const double*)(d + o_w);
    M.ev_re = (const double*)(d + o_ev); M.ev_im = M.ev_re + CA;
    M.v_re = (const double*)(d + o_v); M.v_im = M.v_re + AA;
    M.vi_re = (const double*)(d + o_vi); M.vi_im = M.vi_re + AA;
    const long long n_pairs = (long long)n_rows * (n_rows - 1) / 2;
    auto all_zero = [](const double* v, size_t n) {
      for (size_t k = 0; k < n; ++k)
        if (v[k] != 0.0) return false;
      return true;
    };
    M.is_real = all_zero(model->eval_im, CA) && all_zero(model->evec_im, AA) && all_zero(model->evec_inv_im, AA);
    M.staged = M.is_real && dm_staged_doubles(A, C) * sizeof(double) <= 16 * 1024;
    const size_t smem = dm_shared_bytes(A, C, M.staged);
    if (smem > 48 * 1024) rc = check_cuda(ctx, cudaFuncSetAttribute(distmat_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), "distmat shared memory");
    if (!rc) distmat_kernel<<<(unsigned)n_pairs, DM_THREADS, smem, ctx->stream>>>(M, (const signed char*)(d + o_tok), n_rows, n_cols, expected_sub_rate, max_iterations,
                                                                       (double*)(d + o_dist));
    ctx->launches += 1;
    rc = check_cuda(ctx, cudaGetLastError(), "distmat_kernel");
  }
  if (!rc) rc = check_cuda(ctx, cudaMemcpyAsync(dist, d + o_dist, (size_t)n_rows * n_rows * 8, cudaMemcpyDeviceToHost, ctx->stream), "D2H distances");
  if (!rc) rc = check_cuda(ctx, cudaStreamSynchronize(ctx->stream), "sync after distance matrix");
  cudaFreeAsync(d, ctx->stream);
  return rc;
}
