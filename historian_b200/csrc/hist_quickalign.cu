// historian_b200: the guide-alignment stage.
//
// hist_quickalign_batch is the B200 counterpart of QuickAlignMatrix (reference quickalign.h:8-76,
// quickalign.cpp:7-214): a three-state (match / insert / delete) Viterbi alignment of two sequences restricted to
// the diagonals of a DiagonalEnvelope (diagenv.h:38-155), the pairwise step of the guide alignment that
// AlignGraph runs for ceil(n log2 n) or all pairs of sequences (span.cpp:57-158) before the progressive
// reconstruction starts.
//
// Design.  The reference walks row by row (j outer, the envelope's diagonals inner) over a map-free but serial
// recurrence in which del(i,j) depends on (i-1,j) of the same row.  Here one CTA takes one pair and sweeps the
// anti-diagonals a = i + j; a thread owns envelope diagonals (delta = i - j), and on a diagonal consecutive cells are
// two anti-diagonals apart, so at step a the diagonals of parity a are computed from their own previous cell (a-2) and
// from the latest cells of the neighbouring diagonals delta+1 / delta-1 (parity a-1): the three score arrays
// [n_diagonals] live in shared memory, one __syncthreads per step, no cell plane at all.  What is stored per cell is
// one byte of traceback choices, made with the reference traceback's own expressions and candidate order
// (quickalign.cpp:158-193), so the path is the reference's path.
//
// Arithmetic is FP64 with explicitly unfused adds and multiplies (__dadd_rn / __dmul_rn): Viterbi has no
// log-sum-exp, every value is a sum of the same doubles in the same order as on the CPU, so scores are BIT-EXACT and
// alignments identical, ties included.  The reference's unsigned wrap-around in endGapScore (xLen - i - 2 with
// 32-bit unsigned SeqIdx, quickalign.h:66-69) is reproduced.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>

#include "hist_internal.cuh"

namespace hb {

struct QaPair {
  int x_len, y_len, nd, pad;
  long long x_tok, y_tok;     // offsets (bytes) of the token strings in the blob
  long long diag;             // offset (int32 elements) of the diagonal list
  long long tb;               // offset (bytes) of the traceback plane [nd][y_len + 1]
  long long path;             // offset (bytes) of the path / scratch, capacity 2 * (x_len + y_len)
};

struct QaScores {
  int A;
  double no_gap, gap_open, gap_extend, m2m, m2i, m2d, i2i, i2m, i2d, d2d, d2m;
};

__device__ __forceinline__ double qadd(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ double qmul(double a, double b) { return __dmul_rn(a, b); }
// std::max(a, b): b if a < b, else a (no NaNs here)
__device__ __forceinline__ double qmax(double a, double b) { return a < b ? b : a; }

// startGapScore / endGapScore (quickalign.h:60-69); the products use the reference's 32-bit unsigned arithmetic
__device__ __forceinline__ double qa_start_gap(const QaScores& S, unsigned i, unsigned j) {
  const double gi = i == 1 ? S.no_gap : qadd(S.gap_open, qmul((double)(unsigned)(i - 2u), S.gap_extend));
  const double gj = j == 1 ? S.no_gap : qadd(S.gap_open, qmul((double)(unsigned)(j - 2u), S.gap_extend));
  return qadd(gi, gj);
}
__device__ __forceinline__ double qa_end_gap(const QaScores& S, unsigned i, unsigned j, unsigned xLen, unsigned yLen) {
  const double gi = i == xLen ? S.no_gap : qadd(S.gap_open, qmul((double)(unsigned)(xLen - i - 2u), S.gap_extend));
  const double gj = j == yLen ? S.no_gap : qadd(S.gap_open, qmul((double)(unsigned)(yLen - j - 2u), S.gap_extend));
  return qadd(gi, gj);
}

// grid = pairs; dynamic shared memory: 3 * max_nd doubles + blockDim (double + 2 int) for the end-cell reduction
__global__ void __launch_bounds__(1024, 2) quickalign_kernel(const QaPair* __restrict__ pairs, const char* __restrict__ blob, const int* __restrict__ diags,
                                                          const double* __restrict__ submat, QaScores S, unsigned char* __restrict__ tb_all,
                                                          unsigned char* __restrict__ path_all, double* __restrict__ result,
                                                          int* __restrict__ ends, int* __restrict__ path_len, int max_nd) {
  extern __shared__ double sm[];
  const QaPair P = pairs[blockIdx.x];
  const int T = blockDim.x, tid = threadIdx.x;
  const int xLen = P.x_len, yLen = P.y_len, nd = P.nd;
  double* sM = sm;
  double* sI = sm + max_nd;
  double* sD = sm + 2 * max_nd;
  double* rV = sm + 3 * max_nd;
  int* rJ = reinterpret_cast<int*>(rV + T);
  int* rI = rJ + T;
  const signed char* xt = reinterpret_cast<const signed char*>(blob + P.x_tok);
  const signed char* yt = reinterpret_cast<const signed char*>(blob + P.y_tok);
  const int* dgp = P.diag >= 0 ? diags + P.diag : nullptr;   // NULL: the full envelope, slot k holds diagonal 1 - yLen + k
  const int dlo = 1 - P.y_len;
#define dg(k) (dgp ? dgp[(k)] : dlo + (k))
  unsigned char* tb = tb_all + P.tb;
  const double NEG = -INFINITY;
  for (int k = tid; k < nd; k += T) { sM[k] = NEG; sI[k] = NEG; sD[k] = NEG; }
  double best = NEG;
  int bestI = 0, bestJ = 0;
  __syncthreads();
  const int A = S.A;
  for (int a = 2; a <= xLen + yLen; ++a) {
    // a thread owns PAIRS of consecutive slots: neighbouring diagonals have opposite parity, so in a contiguous band every
    // thread has exactly one cell per step (slot-per-thread would leave every other lane idle)
    for (int kk = 2 * tid; kk < nd + 1; kk += 2 * T)
    for (int k = kk; k < min(kk + 2, nd); ++k) {
      const int delta = dg(k);
      if ((a ^ delta) & 1) continue;
      const int i = (a + delta) / 2, j = a - i;
      if (i < 1 || i > xLen || j < 1 || j > yLen) continue;
      // (i-1, j-1): this diagonal's previous cell
      const double M0 = sM[k], I0 = sI[k], D0 = sD[k];
      // (i, j-1): diagonal delta + 1;  (i-1, j): diagonal delta - 1
      const bool up = k + 1 < nd && dg(k + 1) == delta + 1, lf = k > 0 && dg(k - 1) == delta - 1;
      const double Mu = up ? sM[k + 1] : NEG, Iu = up ? sI[k + 1] : NEG;
      const double Ml = lf ? sM[k - 1] : NEG, Il = lf ? sI[k - 1] : NEG, Dl = lf ? sD[k - 1] : NEG;
      const int tx = xt[i - 1], ty = yt[j - 1];
      const double emit = (tx < 0 || ty < 0) ? 0.0 : submat[tx * A + ty];
      const double sgs = qadd(0.0, qa_start_gap(S, (unsigned)i, (unsigned)j));   // start + startGapScore, start == 0
      const double fM = qadd(M0, S.m2m), fD = qadd(D0, S.d2m), fI = qadd(I0, S.i2m);
      // fill (quickalign.cpp:68-85)
      double m = qmax(qmax(fM, fD), fI);
      m = qmax(m, sgs);
      m = qadd(m, emit);
      const double ins = qmax(qadd(Iu, S.i2i), qadd(Mu, S.m2i));
      const double del = qmax(qmax(qadd(Il, S.i2d), qadd(Dl, S.d2d)), qadd(Ml, S.m2d));
      // the traceback's choices (quickalign.cpp:158-193): first strict maximum in the order Match, Insert, Delete, Start
      int pm = 0;
      {
        double sc = NEG;
        const double cM = qadd(fM, emit), cI = qadd(fI, emit), cD = qadd(fD, emit), cS = qadd(sgs, emit);
        if (cM > sc) { sc = cM; pm = 0; }
        if (cI > sc) { sc = cI; pm = 1; }
        if (cD > sc) { sc = cD; pm = 2; }
        if (cS > sc) { sc = cS; pm = 3; }
      }
      int pi = 0;
      {
        double sc = NEG;
        const double cM = qadd(Mu, S.m2i), cI = qadd(Iu, S.i2i);
        if (cM > sc) { sc = cM; pi = 0; }
        if (cI > sc) { sc = cI; pi = 1; }
      }
      int pd = 0;
      {
        double sc = NEG;
        const double cM = qadd(Ml, S.m2d), cI = qadd(Il, S.i2d), cD = qadd(Dl, S.d2d);
        if (cM > sc) { sc = cM; pd = 0; }
        if (cI > sc) { sc = cI; pd = 1; }
        if (cD > sc) { sc = cD; pd = 2; }
      }
      tb[(size_t)k * (yLen + 1) + j] = (unsigned char)(pm | (pi << 2) | (pd << 3));
      sM[k] = m; sI[k] = ins; sD[k] = del;
      // the end cell (quickalign.cpp:87-92): first strict maximum in (j, i) order
      const double e = qadd(m, qa_end_gap(S, (unsigned)i, (unsigned)j, (unsigned)xLen, (unsigned)yLen));
      if (e > best || (e == best && e > NEG && (j < bestJ || (j == bestJ && i < bestI)))) { best = e; bestI = i; bestJ = j; }
    }
    __syncthreads();
  }
  rV[tid] = best; rJ[tid] = bestJ; rI[tid] = bestI;
  __syncthreads();
  if (tid != 0) return;
  for (int q = 1; q < T; ++q) {
    const double e = rV[q];
    if (e > best || (e == best && e > NEG && (rJ[q] < bestJ || (rJ[q] == bestJ && rI[q] < bestI)))) { best = e; bestI = rI[q]; bestJ = rJ[q]; }
  }
  result[blockIdx.x] = best;
  ends[2 * blockIdx.x] = bestI;
  ends[2 * blockIdx.x + 1] = bestJ;
  if (!(best > NEG)) {
    path_len[blockIdx.x] = 0;
    return;
  }
  // ---- traceback (quickalign.cpp:144-207); ops: 3 both rows, 1 x only, 2 y only.  Written end -> start into the second
  // half of this pair's path area, then laid out start -> end in the first half.
  unsigned char* out = path_all + P.path;
  unsigned char* rev = out + (xLen + yLen);
  int i = bestI, j = bestJ, state = 0, n = 0;
  int k;
  {
    int lo = 0, hi = nd;   // slot of diagonal i - j
    const int delta = i - j;
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if (dg(mid) < delta) lo = mid + 1; else hi = mid;
    }
    k = lo;
  }
  bool bad = false;
  while (state != 3) {
    if (i < 1 || j < 1 || k < 0 || k >= nd || n >= xLen + yLen) { bad = true; break; }
    const int b = tb[(size_t)k * (yLen + 1) + j];
    if (state == 0) { rev[n++] = 3; --i; --j; state = b & 3; }
    else if (state == 1) { rev[n++] = 2; --j; ++k; state = (b >> 2) & 1; }
    else { rev[n++] = 1; --i; --k; state = (b >> 3) & 3; }
  }
  if (bad) {
    path_len[blockIdx.x] = -1;
    return;
  }
  int c = 0;
  for (int q = 0; q < j; ++q) out[c++] = 2;
  for (int q = 0; q < i; ++q) out[c++] = 1;
  for (int q = n - 1; q >= 0; --q) out[c++] = rev[q];
  for (int q = 0; q < xLen - bestI; ++q) out[c++] = 1;
  for (int q = 0; q < yLen - bestJ; ++q) out[c++] = 2;
  path_len[blockIdx.x] = c;
#undef dg
}

}  // namespace hb

using namespace hb;

extern "C" int hist_quickalign_batch(hist_ctx* ctx, int32_t n, const hist_quickalign_pair* pairs, const hist_quickalign_scores* scores,
                                     double* result, int32_t* x_end, int32_t* y_end, uint8_t* paths, const int64_t* path_off,
                                     int32_t* path_len) {
  if (!ctx || n <= 0 || !pairs || !scores || !result) return set_err(ctx, HIST_ERR_ARG, "hist_quickalign_batch: bad arguments");
  if (scores->alph_size < 1 || scores->alph_size > 127 || !scores->submat) return set_err(ctx, HIST_ERR_ARG, "hist_quickalign_batch: bad score table");
  if (paths && (!path_off || !path_len)) return set_err(ctx, HIST_ERR_ARG, "hist_quickalign_batch: paths without path_off / path_len");
  cudaSetDevice(ctx->device);
  const int A = scores->alph_size;
  std::vector<QaPair> qp(n);
  size_t tok_bytes = 0, n_diag = 0, tb_bytes = 0, path_bytes = 0;
  int max_nd = 1;
  for (int k = 0; k < n; ++k) {
    const hist_quickalign_pair& p = pairs[k];
    if (p.x_len < 1 || p.y_len < 1 || !p.x_tok || !p.y_tok || p.n_diagonals < 1)
      return set_err(ctx, HIST_ERR_ARG, "hist_quickalign_batch: pair %d: empty sequence or envelope", k);
    const bool full = !p.diagonals;   // initFull without the list
    if (full && p.n_diagonals != p.x_len + p.y_len - 1)
      return set_err(ctx, HIST_ERR_ARG, "hist_quickalign_batch: pair %d: a full envelope (diagonals == NULL) has xLen + yLen - 1 = %d diagonals, not %d", k,
                     p.x_len + p.y_len - 1, p.n_diagonals);
    for (int q = 0; !full && q < p.n_diagonals; ++q)
      if (p.diagonals[q] < 1 - p.y_len || p.diagonals[q] > p.x_len - 1 || (q && p.diagonals[q] <= p.diagonals[q - 1]))
        return set_err(ctx, HIST_ERR_ARG, "hist_quickalign_batch: pair %d: diagonals must be ascending and inside [1 - yLen, xLen - 1]", k);
    QaPair& q = qp[k];
    q.x_len = p.x_len; q.y_len = p.y_len; q.nd = p.n_diagonals; q.pad = 0;
    q.x_tok = (long long)tok_bytes; tok_bytes += (size_t)p.x_len;
    q.y_tok = (long long)tok_bytes; tok_bytes += (size_t)p.y_len;
    q.diag = full ? -1 : (long long)n_diag;
    if (!full) n_diag += (size_t)p.n_diagonals;
    q.tb = (long long)tb_bytes; tb_bytes += (size_t)p.n_diagonals * ((size_t)p.y_len + 1);
    q.path = (long long)path_bytes; path_bytes += 2 * ((size_t)p.x_len + p.y_len);
    max_nd = std::max(max_nd, p.n_diagonals);
  }
  const int threads = std::min(1024, ((max_nd + 1) / 2 + 31) / 32 * 32);   // one thread per pair of slots
  const size_t smem = (size_t)3 * max_nd * sizeof(double) + (size_t)threads * (sizeof(double) + 2 * sizeof(int));
  if (smem > ctx->smem_optin) return set_err(ctx, HIST_ERR_ARG, "hist_quickalign_batch: an envelope of %d diagonals needs %zu bytes of shared memory (limit %zu)", max_nd, smem, ctx->smem_optin);
  // one host blob: [pair records | tokens | diagonals | score table], one device allocation: [blob | results | traceback | paths]
  auto up = [](size_t v) { return (v + 255) / 256 * 256; };
  const size_t o_pairs = 0, o_tok = up((size_t)n * sizeof(QaPair)), o_diag = o_tok + up(tok_bytes), o_sub = o_diag + up(n_diag * sizeof(int)),
               blob_bytes = o_sub + up((size_t)A * A * sizeof(double));
  const size_t o_res = blob_bytes, o_end = o_res + up((size_t)n * sizeof(double)), o_len = o_end + up((size_t)n * 2 * sizeof(int)),
               o_tb = o_len + up((size_t)n * sizeof(int)), o_path = o_tb + up(tb_bytes), total = o_path + up(path_bytes);
  std::vector<char> blob(blob_bytes, 0);
  memcpy(blob.data() + o_pairs, qp.data(), (size_t)n * sizeof(QaPair));
  for (int k = 0; k < n; ++k) {
    memcpy(blob.data() + o_tok + qp[k].x_tok, pairs[k].x_tok, (size_t)pairs[k].x_len);
    memcpy(blob.data() + o_tok + qp[k].y_tok, pairs[k].y_tok, (size_t)pairs[k].y_len);
    if (qp[k].diag >= 0) memcpy(blob.data() + o_diag + (size_t)qp[k].diag * sizeof(int), pairs[k].diagonals, (size_t)pairs[k].n_diagonals * sizeof(int));
  }
  for (size_t q = 0; q < tok_bytes; ++q) {   // tokens outside the alphabet count as "not in the alphabet"
    signed char& t = reinterpret_cast<signed char&>(blob[o_tok + q]);
    if (t >= A) t = -1;
  }
  memcpy(blob.data() + o_sub, scores->submat, (size_t)A * A * sizeof(double));
  char* d = nullptr;
  HB_CUDA(ctx, cudaMallocFromPoolAsync((void**)&d, total, ctx->pool, ctx->stream));
  int rc = check_cuda(ctx, cudaMemcpyAsync(d, blob.data(), blob_bytes, cudaMemcpyHostToDevice, ctx->stream), "H2D quickalign inputs");
  ctx->h2d_bytes += (long long)blob_bytes;
  if (!rc && smem > 48 * 1024) rc = check_cuda(ctx, cudaFuncSetAttribute(quickalign_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem), "quickalign shared memory");
  if (!rc) {
    QaScores S;
    S.A = A;
    S.no_gap = scores->no_gap; S.gap_open = scores->gap_open; S.gap_extend = scores->gap_extend;
    S.m2m = scores->m2m; S.m2i = scores->m2i; S.m2d = scores->m2d; S.i2i = scores->i2i; S.i2m = scores->i2m; S.i2d = scores->i2d;
    S.d2d = scores->d2d; S.d2m = scores->d2m;
    quickalign_kernel<<<n, threads, smem, ctx->stream>>>((const QaPair*)(d + o_pairs), d + o_tok, (const int*)(d + o_diag), (const double*)(d + o_sub), S,
                                                         (unsigned char*)(d + o_tb), (unsigned char*)(d + o_path), (double*)(d + o_res),
                                                         (int*)(d + o_end), (int*)(d + o_len), max_nd);
    ctx->launches += 1;
    rc = check_cuda(ctx, cudaGetLastError(), "quickalign_kernel");
  }
  std::vector<int> ends((size_t)n * 2), lens(n);
  std::vector<unsigned char> hp;
  if (!rc) rc = check_cuda(ctx, cudaMemcpyAsync(result, d + o_res, (size_t)n * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream), "D2H scores");
  if (!rc) rc = check_cuda(ctx, cudaMemcpyAsync(ends.data(), d + o_end, ends.size() * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream), "D2H end cells");
  if (!rc) rc = check_cuda(ctx, cudaMemcpyAsync(lens.data(), d + o_len, lens.size() * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream), "D2H path lengths");
  if (!rc && paths) {
    hp.resize(path_bytes);
    rc = check_cuda(ctx, cudaMemcpyAsync(hp.data(), d + o_path, path_bytes, cudaMemcpyDeviceToHost, ctx->stream), "D2H paths");
  }
  if (!rc) rc = check_cuda(ctx, cudaStreamSynchronize(ctx->stream), "sync after quickalign");
  cudaFreeAsync(d, ctx->stream);
  if (rc) return rc;
  for (int k = 0; k < n; ++k) {
    if (x_end) x_end[k] = ends[2 * (size_t)k];
    if (y_end) y_end[k] = ends[2 * (size_t)k + 1];
    if (lens[k] < 0) return set_err(ctx, HIST_ERR_NUMERIC, "hist_quickalign_batch: pair %d: traceback left the envelope", k);
    if (paths) {
      path_len[k] = lens[k];
      memcpy(paths + path_off[k], hp.data() + qp[k].path, (size_t)lens[k]);
    }
  }
  return HIST_OK;
}
