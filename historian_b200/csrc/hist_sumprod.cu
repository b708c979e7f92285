// Column-wise Felsenstein sum-product with EM substitution counts on sm_100a.
//
//   SumProduct::SumProduct branch tables      sumprod.cpp:21-50, model.cpp:1267-1297, 1329-1350  -> branch_tables_kernel
//   initColumn / fillUp / fillDown            sumprod.cpp:58-87, 99-161, 163-198                 -> sumprod_columns_kernel
//   logNodePostProb / maxPostState            sumprod.cpp:208-218, 259-262                       -> (same kernel, epilogue)
//   accumulateRootCounts/accumulateEigenCounts sumprod.cpp:264-271, 294-372                      -> (same kernel, count phase)
//   EigenCounts::accumulateSubstitutionCounts  model.cpp:900-915 (the column loop)               -> the grid
//   EigenModel::getSubCounts                  model.cpp:1352-1377                                -> transform_counts_kernel
//
// One thread block per alignment column (persistent: block b takes columns b, b+grid, ...).  The tree is
// renumbered breadth-first so that the nodes of one depth are contiguous: the upward pass walks the
// depths deepest-first, the downward pass root-first, threads spread over (node of the level, component,
// residue).  All arithmetic is FP64 in probability space with log scale factors, the reference's own
// representation, operations in the reference's order, so messages agree to rounding.  Messages live in
// a per-block workspace ([N][C][A] doubles x 3) that stays in L2; eigen-counts are accumulated per
// block in shared memory, written as per-block partials and summed in a fixed order (deterministic).
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <vector>

#include "hist_internal.cuh"

namespace hb {

#define SP_RESCALE 1e-30   // SUMPROD_RESCALE_THRESHOLD, sumprod.cpp:8
#define SP_EIGEN_EPS 1e-6  // EIGENMODEL_EPSILON, model.cpp:22
#define SP_THREADS 256
#define SP_TOK_GAP (-2)
#define SP_TOK_WILD (-1)

struct SpDev {
  int A, C, N, n_levels, is_real;
  const int* lvl_off;    // [n_levels+1] internal node ranges by depth (level 0 = the tree root)
  const int* parent;     // [N] internal index of the parent, -1 for the root
  const int* kid_off;    // [N+1]
  const int* kid_idx;    // children in the reference's order (ascending original index)
  const int* orig;       // [N] internal -> original (postorder) index
  const int* lvl_of;     // [N] depth level of each internal index
  const double* P;       // [N][C][A][A] branch matrices P(i->j)
  const double* PT;      // [N][C][A][A] transposed (j major)
  const double2* eet;    // [N][C][A] exp(lambda_k t)
  const double* t;       // [N] branch lengths
  const double2* V;      // [C][A][A] right eigenvectors V[a][k]
  const double2* Vinv;   // [C][A][A] V^-1[k][b]
  const double2* invdl;  // [C][A][A] 1 / (lambda_k - lambda_l)
  const uint8_t* eveq;   // [C][A][A] eigenvalues equal within EIGENMODEL_EPSILON (or k == l)
  const double* ins;     // [C][A]
  const double* logcw;   // [C]
  const double* sub_rate;  // [C][A][A]
  const int8_t* tokmap;  // [256] char -> token, SP_TOK_WILD if not in the alphabet
  const char* alphabet;  // [A]
};

__device__ __forceinline__ double2 cmul(double2 a, double2 b) { return make_double2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x); }
__device__ __forceinline__ double2 cadd(double2 a, double2 b) { return make_double2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ double2 csub(double2 a, double2 b) { return make_double2(a.x - b.x, a.y - b.y); }
__device__ __forceinline__ double2 cscale(double2 a, double s) { return make_double2(a.x * s, a.y * s); }

__device__ __forceinline__ double lse2_exact(double a, double b) {  // log_sum_exp_slow, logsumexp.cpp:22-32
  if (a == -INFINITY) return b;
  if (b == -INFINITY) return a;
  const double m = fmax(a, b);
  return m + log1p(exp(-fabs(a - b)));
}

// eigenSubCount (model.cpp:1329-1350) of one branch from its exp(lambda t) row
__device__ __forceinline__ double2 eigen_sub_count(const SpDev& S, int node, int c, int k, int l) {
  const double2* e = S.eet + ((size_t)node * S.C + c) * S.A;
  const size_t q = ((size_t)c * S.A + k) * S.A + l;
  if (S.eveq[q]) return cscale(e[k], S.t[node]);
  return cmul(csub(e[k], e[l]), S.invdl[q]);
}

// ------------------------------------------------------------------ S0: per-branch tables
__global__ void branch_tables_kernel(SpDev S, double* P, double* PT, const double2* eet, const double2* eval,
                                     const double* given /* [N][C][A][A] internal or NULL */) {
  const int A = S.A, C = S.C;
  const size_t total = (size_t)S.N * C * A * A;
  for (size_t q = blockIdx.x * (size_t)blockDim.x + threadIdx.x; q < total; q += (size_t)gridDim.x * blockDim.x) {
    const int j = q % A, i = (q / A) % A, c = (q / ((size_t)A * A)) % C, node = q / ((size_t)A * A * C);
    double p = 0;
    if (node != 0) {  // the root's slices stay zero (sumprod.cpp:36-49)
      const double t = S.t[node];
      if (given) p = given[q];
      else if (S.is_real) {  // getSubProbInner, model.cpp:1267-1273
        for (int k = 0; k < A; ++k)
          p += S.V[((size_t)c * A + i) * A + k].x * S.Vinv[((size_t)c * A + k) * A + j].x * exp(eval[c * A + k].x * t);
        p = fmin(1., fmax(0., p));
      } else {  // model.cpp:1274-1282
        double2 acc = make_double2(0, 0);
        for (int k = 0; k < A; ++k)
          acc = cadd(acc, cmul(cmul(S.V[((size_t)c * A + i) * A + k], S.Vinv[((size_t)c * A + k) * A + j]), eet[((size_t)node * C + c) * A + k]));
        p = fmin(1., fmax(0., acc.x));
      }
    }
    P[q] = p;
    PT[(((size_t)node * C + c) * A + j) * A + i] = p;
  }
}

// exp(lambda_k t) per branch (compute_exp_ev_t, model.cpp:1213-1229)
__global__ void exp_eval_kernel(int N, int C, int A, const double* t, const double2* eval, int is_real, double2* eet) {
  const size_t total = (size_t)N * C * A;
  for (size_t q = blockIdx.x * (size_t)blockDim.x + threadIdx.x; q < total; q += (size_t)gridDim.x * blockDim.x) {
    const int k = q % A, c = (q / A) % C, node = q / ((size_t)A * C);
    const double2 lam = eval[c * A + k];
    const double tt = t[node];
    const double m = exp(lam.x * tt);
    double s, co;
    sincos(lam.y * tt, &s, &co);
    eet[q] = lam.y == 0 ? make_double2(m, 0) : make_double2(m * co, m * s);
  }
}

// ------------------------------------------------------------------ load balance
// The cost of a column grows with its number of ungapped rows, and full-tree columns sit next to insertions
// that touch a handful of rows.  Columns are therefore processed in order of decreasing occupancy class
// (floor(log2(ungapped rows))), round-robin over the resident blocks: every block gets its share of the dense
// columns first and fills up with cheap ones.  The order is computed on device and is deterministic (stable
// within a class), so the summation order of the counts is reproducible.
__global__ void column_occupancy_kernel(const char* __restrict__ cols, long long n_cols, int N, int* __restrict__ occ) {
  const long long col = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
  if (col >= n_cols) return;
  const int lane = threadIdx.x & 31;
  const char* c = cols + (size_t)col * N;
  int n = 0;
  for (int r = lane; r < N; r += 32) n += c[r] != '-' && c[r] != '.';
  for (int o = 16; o; o >>= 1) n += __shfl_down_sync(0xffffffffu, n, o);
  if (lane == 0) occ[col] = n;
}

#define SP_BINS 32
// one block of 1024 threads: stable multi-split of the columns by class, densest class first
__global__ void __launch_bounds__(1024) column_order_kernel(const int* __restrict__ occ, int n_cols, int* __restrict__ order) {
  __shared__ int s_cnt[32][SP_BINS];   // per warp, per class: members in this chunk
  __shared__ int s_base[SP_BINS];      // next output position of each class
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  auto cls = [](int o) { return o <= 0 ? SP_BINS - 1 : min(SP_BINS - 2, __clz(o)); };  // fewer leading zeros = denser = earlier
  // class totals
  if (tid < SP_BINS) s_base[tid] = 0;
  __syncthreads();
  for (int c = tid; c < n_cols; c += 1024) atomicAdd(&s_base[cls(occ[c])], 1);
  __syncthreads();
  if (tid == 0) {
    int run = 0;
    for (int b = 0; b < SP_BINS; ++b) { const int n = s_base[b]; s_base[b] = run; run += n; }
  }
  __syncthreads();
  for (int c0 = 0; c0 < n_cols; c0 += 1024) {
    const int c = c0 + tid;
    const int b = c < n_cols ? cls(occ[c]) : -1;
    if (lane < SP_BINS) s_cnt[warp][lane] = 0;
    __syncwarp();
    const unsigned peers = __match_any_sync(0xffffffffu, b);
    const int rank = __popc(peers & ((1u << lane) - 1));
    if (b >= 0 && rank == 0) s_cnt[warp][b] = __popc(peers);
    __syncthreads();
    // exclusive prefix over the warps, per class (thread = class)
    if (tid < SP_BINS) {
      int run = s_base[tid];
      for (int w = 0; w < 32; ++w) { const int n = s_cnt[w][tid]; s_cnt[w][tid] = run; run += n; }
      s_base[tid] = run;
    }
    __syncthreads();
    if (b >= 0) order[s_cnt[warp][b] + rank] = c;
    __syncthreads();
  }
}

// ------------------------------------------------------------------ S1-S4: the column kernel
struct SpArgs {
  const char* columns;     // [n_cols][N] original row order
  const double* weights;   // [n_cols] or NULL
  long long n_cols;
  int flags;
  double* col_loglike;     // [n_cols] or NULL
  char* map_state;         // [n_cols][N] or NULL
  double* node_post;       // [n_cols][N][A] or NULL
  double* work;            // [grid] blocks of workspace
  size_t work_stride;      // doubles per block
  double* partials;        // [grid][counts_size] or NULL
  int counts_size;
  int chunk;               // nodes per staging chunk of the count phase
  int n_slices;            // node slices per (c,k,l) in the count phase
  int* err;                // device error flag
  const int* order;        // [n_cols] processing order (densest columns first: static load balance) or NULL
  int full;                // process every node of every level (debug fetch of all messages); else only the
                           // per-level index ranges that hold ungapped nodes
  int n_excl;              // logNodeExcludedPostProb queries per column
  const int* excl_q;       // [n_excl][2] internal indices (node, excluded neighbour or -1)
  double* excl_out;        // [n_excl][n_cols][C][A] or NULL
  int excl_norm;           // normalise every [C][A] block
};

// (forcing more resident blocks through a register cap was measured: 48 / 40 / 32 registers all spill and lose)
__global__ void __launch_bounds__(SP_THREADS) sumprod_columns_kernel(SpDev S, SpArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int A = S.A, C = S.C, N = S.N, CA = C * A, CAA = C * A * A;
  const int tid = threadIdx.x, nt = blockDim.x;
  // ---- shared layout
  double2* s_acc = (double2*)smem_raw;                       // [n_slices][CAA]
  double2* s_Ub = s_acc + (size_t)a.n_slices * CAA;          // [chunk][C][A]
  double2* s_Db = s_Ub + (size_t)a.chunk * CA;               // [chunk][C][A]
  double* s_U = (double*)(s_Db + (size_t)a.chunk * CA);      // [chunk][C][A]
  double* s_D = s_U + (size_t)a.chunk * CA;                  // [chunk][C][A]
  double* s_scale = s_D + (size_t)a.chunk * CA;              // [chunk][C]
  double* s_root_acc = s_scale + (size_t)a.chunk * C;        // [C][A]
  double* s_cptll = s_root_acc + CA;                         // [C]
  double* s_red = s_cptll + C;                               // [SP_THREADS / 32][C] warp partials of the root terms
  double* s_misc = s_red + (SP_THREADS / 32) * C;            // [4]: colLL, lp accumulator
  int* s_int = (int*)(s_misc + 4);                           // [4]: n_ungapped, n_roots, root (internal), -
  signed char* s_tok = (signed char*)(s_int + 4);            // [N]
  unsigned char* s_valid = (unsigned char*)(s_tok + ((N + 15) & ~15));  // [chunk][C]
  int* s_lo = (int*)(s_valid + (((size_t)a.chunk * C + 15) & ~(size_t)15));  // [n_levels] first / last internal index of
  int* s_hi = s_lo + S.n_levels;                                             // the ungapped nodes of each level

  // ---- per-block workspace (stays in L2)
  double* W = a.work + (size_t)blockIdx.x * a.work_stride;
  double* E = W;
  double* F = E + (size_t)N * CA;
  double* G = F + (size_t)N * CA;
  double* logE = G + (size_t)N * CA;
  double* logF = logE + (size_t)N * C;
  double* logG = logF + (size_t)N * C;
  double* rootterm = logG + (size_t)N * C;  // [N][C]

  for (int q = tid; q < a.n_slices * CAA; q += nt) s_acc[q] = make_double2(0, 0);
  for (int q = tid; q < CA; q += nt) s_root_acc[q] = 0;
  if (tid == 0) s_misc[1] = 0;
  __syncthreads();

  for (long long slot = blockIdx.x; slot < a.n_cols; slot += gridDim.x) {
    const long long col = a.order ? a.order[slot] : slot;
    const char* cs = a.columns + (size_t)col * N;
    // ---- initColumn (sumprod.cpp:58-87)
    if (tid < 4) s_int[tid] = tid == 2 ? 0x7fffffff : 0;
    for (int lev = tid; lev < S.n_levels; lev += nt) {
      s_lo[lev] = a.full ? S.lvl_off[lev] : 0x7fffffff;
      s_hi[lev] = a.full ? S.lvl_off[lev + 1] - 1 : -1;
    }
    __syncthreads();
    for (int node = tid; node < N; node += nt) {
      const unsigned char ch = (unsigned char)cs[S.orig[node]];
      signed char tk;
      if (ch == '-' || ch == '.') tk = SP_TOK_GAP;
      else tk = ch == '*' ? SP_TOK_WILD : S.tokmap[ch];
      s_tok[node] = tk;
    }
    __syncthreads();
    for (int node = tid; node < N; node += nt) {
      const bool gap = s_tok[node] == SP_TOK_GAP;
      bool is_root = false;
      if (!gap) {
        atomicAdd(&s_int[0], 1);
        if (!a.full) {
          // the ungapped nodes of a column form (a few) subtrees, which are contiguous index ranges per level
          const int lev = S.lvl_of[node];
          atomicMin(&s_lo[lev], node);
          atomicMax(&s_hi[lev], node);
        }
        const int p = S.parent[node];
        is_root = p < 0 || s_tok[p] == SP_TOK_GAP;
        if (is_root) {
          s_int[3] = node;  // internal index of the column root (unique whenever it is used)
          atomicAdd(&s_int[1], 1);
          atomicMin(&s_int[2], S.orig[node]);  // roots.front(): the first root in postorder
        }
      }
      if (gap && a.full)  // E = 1, logE = 0 of gapped rows (sumprod.cpp:76-80); the passes below skip gapped children instead
        for (int c = 0; c < C; ++c) {
          logE[(size_t)node * C + c] = 0;
          for (int i = 0; i < A; ++i) E[((size_t)node * C + c) * A + i] = 1;
        }
    }
    __syncthreads();
    const int n_ungapped = s_int[0], n_roots = s_int[1];

    // ---- fillUp (sumprod.cpp:99-161), deepest level first
    for (int lev = S.n_levels - 1; lev >= 0; --lev) {
      const int l0 = s_lo[lev], ln = s_hi[lev] - l0 + 1;
      if (ln <= 0) continue;  // uniform over the block
      // F, logF, root terms: one thread per (node, component)
      for (int q = tid; q < ln * C; q += nt) {
        const int node = l0 + q / C, c = q % C;
        const int k0 = S.kid_off[node], k1 = S.kid_off[node + 1];
        double lF = 0;
        for (int k = k0; k < k1; ++k)
          if (s_tok[S.kid_idx[k]] != SP_TOK_GAP) lF += logE[(size_t)S.kid_idx[k] * C + c];  // gapped children: logE = 0, E = 1
        const int tk = s_tok[node];
        double rterm = 0;
        if (tk != SP_TOK_GAP) {
          double* Fn = F + ((size_t)node * C + c) * A;
          if (tk == SP_TOK_WILD) {
            double Fmax = 0;
            for (int i = 0; i < A; ++i) {
              double Fi = 1;
              for (int k = k0; k < k1; ++k)
                if (s_tok[S.kid_idx[k]] != SP_TOK_GAP) Fi *= E[((size_t)S.kid_idx[k] * C + c) * A + i];
              Fn[i] = Fi;
              if (Fi > Fmax) Fmax = Fi;
            }
            if (Fmax < SP_RESCALE) {
              for (int i = 0; i < A; ++i) Fn[i] /= Fmax;
              lF += log(Fmax);
            }
          } else {
            double Ftok = 1;
            for (int k = k0; k < k1; ++k)
              if (s_tok[S.kid_idx[k]] != SP_TOK_GAP) Ftok *= E[((size_t)S.kid_idx[k] * C + c) * A + tk];
            if (Ftok < SP_RESCALE) {
              lF += log(Ftok);
              Ftok = 1;
            }
            for (int i = 0; i < A; ++i) Fn[i] = 0;
            Fn[tk] = Ftok;
          }
          const int p = S.parent[node];
          if (p < 0 || s_tok[p] == SP_TOK_GAP) {
            double ip = 0;
            for (int i = 0; i < A; ++i) ip += Fn[i] * S.ins[c * A + i];
            rterm = lF + log(ip);
            logE[(size_t)node * C + c] = 0;  // never read as a message (the parent is a gap); defined for the debug fetch
          } else {
            logE[(size_t)node * C + c] = lF;
          }
        }
        rootterm[(size_t)node * C + c] = rterm;
        logF[(size_t)node * C + c] = lF;
      }
      __syncthreads();
      // E = P F: one thread per (node, component, residue)
      for (int q = tid; q < ln * CA; q += nt) {
        const int node = l0 + q / CA, c = (q / A) % C, i = q % A;
        const int tk = s_tok[node];
        if (tk == SP_TOK_GAP) continue;
        const int p = S.parent[node];
        if (p < 0 || s_tok[p] == SP_TOK_GAP) continue;
        const double* Fn = F + ((size_t)node * C + c) * A;
        const double* Pt = S.PT + ((size_t)node * C + c) * A * A;  // Pt[j * A + i] = P(i -> j)
        double Ei;
        if (tk >= 0) Ei = Pt[tk * A + i] * Fn[tk];  // one-hot F: the other terms of the sum are exact zeros
        else {
          Ei = 0;
          for (int j = 0; j < A; ++j) Ei += Pt[j * A + i] * Fn[j];
        }
        E[((size_t)node * C + c) * A + i] = Ei;
      }
      __syncthreads();
    }
    // cptLogLike[c] = sum of the root terms; colLogLike = LSE_c(logCptWeight + cptLogLike)
    {
      const int warp = tid >> 5, lane = tid & 31;
      for (int c = 0; c < C; ++c) {
        double v = 0;
        for (int lev = 0; lev < S.n_levels; ++lev)
          if (s_hi[lev] >= 0)  // an empty level has lo = INT_MAX
            for (int node = s_lo[lev] + tid; node <= s_hi[lev]; node += nt) v += rootterm[(size_t)node * C + c];
        for (int o = 16; o; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
        if (lane == 0) s_red[warp * C + c] = v;
      }
      __syncthreads();
      if (tid == 0) {
        double ll = -INFINITY;
        for (int c = 0; c < C; ++c) {
          double v = 0;
          for (int w = 0; w < nt / 32; ++w) v += s_red[w * C + c];
          s_cptll[c] = v;
          ll = lse2_exact(ll, S.logcw[c] + v);
        }
        s_misc[0] = ll;
        if (a.col_loglike) a.col_loglike[col] = ll;
      }
      __syncthreads();
    }
    const double colLL = s_misc[0];
    if (n_ungapped == 0) {  // uniform over the block
      if (a.excl_out)
        for (int q = tid; q < a.n_excl * CA; q += nt) a.excl_out[((size_t)(q / CA) * a.n_cols + col) * CA + q % CA] = -INFINITY;
      continue;
    }

    // ---- fillDown (sumprod.cpp:163-198), root first
    for (int lev = 0; lev < S.n_levels; ++lev) {
      const int l0 = s_lo[lev], ln = s_hi[lev] - l0 + 1;
      if (ln <= 0) continue;
      for (int q = tid; q < ln * CA; q += nt) {
        const int node = l0 + q / CA, c = (q / A) % C, j = q % A;
        if (s_tok[node] == SP_TOK_GAP) continue;
        const int p = S.parent[node];
        if (p < 0 || s_tok[p] == SP_TOK_GAP) {
          G[((size_t)node * C + c) * A + j] = S.ins[c * A + j];
          if (j == 0) logG[(size_t)node * C + c] = 0;
        } else {
          const int k0 = S.kid_off[p], k1 = S.kid_off[p + 1];
          if (j == 0) {
            double lG = logG[(size_t)p * C + c];
            for (int k = k0; k < k1; ++k)
              if (S.kid_idx[k] != node && s_tok[S.kid_idx[k]] != SP_TOK_GAP) lG += logE[(size_t)S.kid_idx[k] * C + c];
            logG[(size_t)node * C + c] = lG;
          }
          const double* Gp = G + ((size_t)p * C + c) * A;
          const double* Pn = S.P + ((size_t)node * C + c) * A * A;
          double Gj = 0;
          for (int i = 0; i < A; ++i) {
            double pr = Gp[i] * Pn[i * A + j];
            for (int k = k0; k < k1; ++k) {
              const int s = S.kid_idx[k];
              if (s != node && s_tok[s] != SP_TOK_GAP) pr *= E[((size_t)s * C + c) * A + i];
            }
            Gj += pr;
          }
          G[((size_t)node * C + c) * A + j] = Gj;
        }
      }
      __syncthreads();
    }

    // ---- logNodeExcludedPostProb without normalisation (sumprod.cpp:219-250): the messages a node receives from all its
    // neighbours but one, restricted to its own residue; rows of -inf where the node is gapped
    if (a.excl_out)
      for (int q = tid; q < a.n_excl * CA; q += nt) {
        const int e = q / CA, c = (q / A) % C, i = q % A;
        const int node = a.excl_q[2 * e], ex = a.excl_q[2 * e + 1];
        const int tk = s_tok[node];
        double v = -INFINITY;
        if (tk != SP_TOK_GAP) {
          v = (tk == SP_TOK_WILD || tk == i) ? 0. : -INFINITY;
          v += S.logcw[c];
          for (int k = S.kid_off[node]; k < S.kid_off[node + 1]; ++k) {
            const int ch = S.kid_idx[k];
            if (ch != ex && s_tok[ch] != SP_TOK_GAP) v += log(E[((size_t)ch * C + c) * A + i]) + logE[(size_t)ch * C + c];   // gapped children: E = 1, logE = 0
          }
          if (S.parent[node] != ex || ex < 0) v += log(G[((size_t)node * C + c) * A + i]) + logG[(size_t)node * C + c];
        }
        a.excl_out[((size_t)e * a.n_cols + col) * CA + c * A + i] = v;
      }
    if (a.excl_out && a.excl_norm) {   // lp -= norm, norm = log-sum-exp over components and residues (sumprod.cpp:242, 245-248)
      __syncthreads();
      for (int e = tid; e < a.n_excl; e += nt) {
        double* o = a.excl_out + ((size_t)e * a.n_cols + col) * CA;
        double norm = -INFINITY;
        for (int q = 0; q < CA; ++q) norm = lse2_exact(norm, o[q]);
        if (norm > -INFINITY)
          for (int q = 0; q < CA; ++q) o[q] -= norm;
      }
    }

    const bool need_root = (a.flags & (HIST_SP_COUNTS | HIST_SP_MAP_STATES | HIST_SP_NODE_POST)) != 0;
    if (need_root && n_roots != 1) {  // assertSingleRoot
      if (tid == 0) atomicExch(a.err, 1);
      continue;
    }

    // ---- accumulateRootCounts + accumulateEigenCounts (sumprod.cpp:264-271, 294-372)
    if ((a.flags & HIST_SP_COUNTS) && a.partials) {
      const double weight = a.weights ? a.weights[col] : 1.0;
      // internal index of the column root
      int root = -1;
      {
        root = s_int[3];  // recorded by initColumn above
      }
      for (int q = tid; q < CA; q += nt) {
        const int c = q / A;
        const double norm = exp(S.logcw[c] + logF[(size_t)root * C + c] - colLL);
        s_root_acc[q] += weight * S.ins[q] * F[(size_t)root * CA + q] * norm;
      }
      if (tid == 0) s_misc[1] += colLL;  // indelCounts.lp += columnLogLikelihood (model.cpp:909)
      for (int lev = 0; lev < S.n_levels; ++lev)
      for (int n0 = s_lo[lev]; n0 <= s_hi[lev]; n0 += a.chunk) {
        const int nn = min(a.chunk, s_hi[lev] + 1 - n0);
        // U, D, scale: one thread per (node, component)
        for (int q = tid; q < nn * C; q += nt) {
          const int node = n0 + q / C, c = q % C;
          const bool ok = s_tok[node] != SP_TOK_GAP && node != root;
          s_valid[q] = ok;
          if (!ok) continue;
          const int p = S.parent[node];
          const int k0 = S.kid_off[p], k1 = S.kid_off[p + 1];
          const double* U0 = F + ((size_t)node * C + c) * A;
          double* U = s_U + (size_t)q * A;
          double* D = s_D + (size_t)q * A;
          double maxU = -INFINITY, maxD = -INFINITY, lsib = 0;
          for (int k = k0; k < k1; ++k)
            if (S.kid_idx[k] != node && s_tok[S.kid_idx[k]] != SP_TOK_GAP) lsib += logE[(size_t)S.kid_idx[k] * C + c];
          for (int i = 0; i < A; ++i) {
            double d = G[((size_t)p * C + c) * A + i];
            for (int k = k0; k < k1; ++k)
              if (S.kid_idx[k] != node && s_tok[S.kid_idx[k]] != SP_TOK_GAP) d *= E[((size_t)S.kid_idx[k] * C + c) * A + i];
            D[i] = d;
            maxD = fmax(maxD, d);
            maxU = fmax(maxU, U0[i]);
          }
          const double norm = exp(colLL - S.logcw[c] - logF[(size_t)node * C + c] - logG[(size_t)p * C + c] - lsib) / (maxU * maxD);
          for (int i = 0; i < A; ++i) {
            U[i] = U0[i] / maxU;
            D[i] = D[i] / maxD;
          }
          s_scale[q] = weight / norm;
        }
        __syncthreads();
        // Ubasis = V^-1 U, Dbasis = D^T V: one thread per (node, component, k)
        for (int q = tid; q < nn * CA; q += nt) {
          const int nc = q / A, k = q % A, c = nc % C;
          if (!s_valid[nc]) continue;
          const double* U = s_U + (size_t)nc * A;
          const double* D = s_D + (size_t)nc * A;
          double2 ub = make_double2(0, 0), db = make_double2(0, 0);
          if (S.is_real) {
            for (int b = 0; b < A; ++b) {
              ub.x += S.Vinv[((size_t)c * A + k) * A + b].x * U[b];
              db.x += S.V[((size_t)c * A + b) * A + k].x * D[b];
            }
          } else {
            for (int b = 0; b < A; ++b) {
              ub = cadd(ub, cscale(S.Vinv[((size_t)c * A + k) * A + b], U[b]));
              db = cadd(db, cscale(S.V[((size_t)c * A + b) * A + k], D[b]));
            }
          }
          s_Ub[q] = ub;
          s_Db[q] = db;
        }
        __syncthreads();
        // eigenCounts[c][k][l] += Dbasis[k] * eigenSubCount[k][l] * Ubasis[l] * weight / norm
        for (int q = tid; q < a.n_slices * CAA; q += nt) {
          const int slice = q / CAA, ckl = q % CAA, c = ckl / (A * A), k = (ckl / A) % A, l = ckl % A;
          double2 acc = s_acc[q];
          for (int n = slice; n < nn; n += a.n_slices) {
            const int nc = n * C + c;
            if (!s_valid[nc]) continue;
            const double2 esc = eigen_sub_count(S, n0 + n, c, k, l);
            acc = cadd(acc, cscale(cmul(s_Db[(size_t)nc * A + k], cmul(esc, s_Ub[(size_t)nc * A + l])), s_scale[nc]));
          }
          s_acc[q] = acc;
        }
        __syncthreads();
      }
    }

    // ---- logNodePostProb / maxPostState (sumprod.cpp:208-218, 259-262, 398-412)
    if (((a.flags & HIST_SP_NODE_POST) && a.node_post) || ((a.flags & HIST_SP_MAP_STATES) && a.map_state)) {
      const bool want_post = (a.flags & HIST_SP_NODE_POST) && a.node_post;
      const bool want_map = (a.flags & HIST_SP_MAP_STATES) && a.map_state;
      for (int node = tid; node < N; node += nt) {
        const int o = S.orig[node];
        const int tk = s_tok[node];
        char ch = cs[o];
        double* out = want_post ? a.node_post + ((size_t)col * N + o) * A : nullptr;
        if (tk == SP_TOK_GAP) {
          if (out)
            for (int i = 0; i < A; ++i) out[i] = -INFINITY;
        } else if (out || (want_map && ch == '*')) {
          double best = -INFINITY;
          int besti = 0;
          for (int i = 0; i < A; ++i) {
            double v = -INFINITY;
            for (int c = 0; c < C; ++c)
              v = lse2_exact(v, S.logcw[c] + logF[(size_t)node * C + c] + log(F[((size_t)node * C + c) * A + i]) + logG[(size_t)node * C + c] +
                                    log(G[((size_t)node * C + c) * A + i]) - colLL);
            v = fmin(v, 0.);
            if (out) out[i] = v;
            if (v > best) {
              best = v;
              besti = i;
            }
          }
          if (ch == '*') ch = S.alphabet[besti];
        }
        if (want_map) a.map_state[(size_t)col * N + o] = ch;
      }
    }
    __syncthreads();
  }

  // ---- per-block partial of the flat EigenCounts buffer
  if (a.partials) {
    __syncthreads();
    double* out = a.partials + (size_t)blockIdx.x * a.counts_size;
    for (int q = tid; q < 7; q += nt) out[q] = q == 6 ? s_misc[1] : 0;
    for (int q = tid; q < CA; q += nt) out[7 + q] = s_root_acc[q];
    for (int q = tid; q < CAA; q += nt) {
      double2 v = make_double2(0, 0);
      for (int s = 0; s < a.n_slices; ++s) v = cadd(v, s_acc[(size_t)s * CAA + q]);
      out[7 + CA + 2 * q] = v.x;
      out[7 + CA + 2 * q + 1] = v.y;
    }
  }
}


// ------------------------------------------------------------------ dense columns: node-major, level-synchronous
// A column in which most of the tree is ungapped (an alignment column that descends from the root sequence) costs
// a whole-tree pruning pass, and the one-block-per-column kernel above walks its levels with 256 threads whose every
// message is a dependent L2 access: ~3.7 ms per 8191-node column however many SMs are idle.  The dense columns of a
// batch are therefore taken out and processed TOGETHER, node-major: messages are laid out [node][component][residue]
// [column] so that the threads of a warp work on 32 columns of the same node and every load and store is coalesced;
// one launch per tree level (deepest first for fillUp, root first for fillDown) over all dense columns, then one
// count launch whose threads keep the A x A eigen-count accumulators of one (component, column) in registers over a
// chunk of nodes and reduce over columns per block.  Same arithmetic, same operation order inside a message as above.
// Alphabets up to 8 residues (the accumulators live in registers); larger alphabets keep the kernel above.
#define DN_COLS 128     // columns per block (thread = column)
#define DN_NODES 64     // nodes per block of the count kernel
#define DN_MAXA 8

struct DnArgs {
  const char* columns;     // [n_cols][N]
  const int* order;        // dense column b is columns[order[b]]
  const double* weights;   // indexed by original column, or NULL
  int nb;                  // dense columns in this batch (= stride of the column axis)
  signed char* tok;        // [N][nb]
  double* E; double* F; double* G;        // [N][C][A][nb]
  double* logE; double* logF; double* logG;   // [N][C][nb]
  double* colll;           // [nb]
  double* cpt;             // [C][nb] sum of the root terms of each component (zeroed per batch)
  int* n_roots;            // [nb] (zeroed per batch)
  int* root_node;          // [nb] internal index of the (last) root (zeroed per batch)
  double* col_loglike;     // user output indexed by original column, or NULL
  double* partials;        // [n_partials][counts_size]
  int counts_size;
  int* err;
};

__global__ void dn_init_kernel(SpDev S, DnArgs a) {
  const int col = blockIdx.x * blockDim.x + threadIdx.x, node = blockIdx.y;
  if (col >= a.nb) return;
  const unsigned char ch = (unsigned char)a.columns[(size_t)a.order[col] * S.N + S.orig[node]];
  signed char tk;
  if (ch == '-' || ch == '.') tk = SP_TOK_GAP;
  else tk = ch == '*' ? SP_TOK_WILD : S.tokmap[ch];
  a.tok[(size_t)node * a.nb + col] = tk;
}

// fillUp of one level (sumprod.cpp:99-161): thread = (column, node of the level, component)
__global__ void __launch_bounds__(DN_COLS) dn_up_kernel(SpDev S, DnArgs a, int lev) {
  const int col = blockIdx.x * blockDim.x + threadIdx.x;
  if (col >= a.nb) return;
  const int A = S.A, C = S.C, nb = a.nb;
  const int node = S.lvl_off[lev] + blockIdx.y, c = blockIdx.z;
  const size_t nc = (size_t)node * C + c;
  const int tk = a.tok[(size_t)node * nb + col];
  if (tk == SP_TOK_GAP) return;
  const int k0 = S.kid_off[node], k1 = S.kid_off[node + 1];
  double lF = 0;
  for (int k = k0; k < k1; ++k) {
    const int kid = S.kid_idx[k];
    if (a.tok[(size_t)kid * nb + col] != SP_TOK_GAP) lF += a.logE[((size_t)kid * C + c) * nb + col];
  }
  double* Fn = a.F + nc * A * nb + col;   // Fn[i * nb]
  if (tk == SP_TOK_WILD) {
    double Fmax = 0;
    for (int i = 0; i < A; ++i) {
      double Fi = 1;
      for (int k = k0; k < k1; ++k) {
        const int kid = S.kid_idx[k];
        if (a.tok[(size_t)kid * nb + col] != SP_TOK_GAP) Fi *= a.E[(((size_t)kid * C + c) * A + i) * nb + col];
      }
      Fn[(size_t)i * nb] = Fi;
      if (Fi > Fmax) Fmax = Fi;
    }
    if (Fmax < SP_RESCALE) {
      for (int i = 0; i < A; ++i) Fn[(size_t)i * nb] /= Fmax;
      lF += log(Fmax);
    }
  } else {
    double Ftok = 1;
    for (int k = k0; k < k1; ++k) {
      const int kid = S.kid_idx[k];
      if (a.tok[(size_t)kid * nb + col] != SP_TOK_GAP) Ftok *= a.E[(((size_t)kid * C + c) * A + tk) * nb + col];
    }
    if (Ftok < SP_RESCALE) {
      lF += log(Ftok);
      Ftok = 1;
    }
    for (int i = 0; i < A; ++i) Fn[(size_t)i * nb] = i == tk ? Ftok : 0;
  }
  const int p = S.parent[node];
  const bool is_root = p < 0 || a.tok[(size_t)p * nb + col] == SP_TOK_GAP;
  if (is_root) {
    // a column has one root (several only where no counts are asked for: their terms are then summed in arrival order)
    double ip = 0;
    for (int i = 0; i < A; ++i) ip += Fn[(size_t)i * nb] * S.ins[c * A + i];
    atomicAdd(a.cpt + (size_t)c * nb + col, lF + log(ip));
    if (c == 0) {
      atomicAdd(a.n_roots + col, 1);
      atomicMax(a.root_node + col, node);
    }
    a.logE[nc * nb + col] = 0;
  } else {
    a.logE[nc * nb + col] = lF;
    // E = P F
    const double* Pt = S.PT + nc * A * A;  // Pt[j * A + i] = P(i -> j)
    double* En = a.E + nc * A * nb + col;
    for (int i = 0; i < A; ++i) {
      double Ei;
      if (tk >= 0) Ei = Pt[tk * A + i] * Fn[(size_t)tk * nb];
      else {
        Ei = 0;
        for (int j = 0; j < A; ++j) Ei += Pt[j * A + i] * Fn[(size_t)j * nb];
      }
      En[(size_t)i * nb] = Ei;
    }
  }
  a.logF[nc * nb + col] = lF;
}

// column log-likelihoods, the root's counts and the likelihood term of the count buffer: thread = column
__global__ void __launch_bounds__(DN_COLS) dn_root_kernel(SpDev S, DnArgs a, int want_counts, int first_partial) {
  const int col = blockIdx.x * blockDim.x + threadIdx.x;
  if (col >= a.nb) return;
  const int A = S.A, C = S.C, CA = C * A, nb = a.nb;
  const int n_roots = a.n_roots[col], root = n_roots ? a.root_node[col] : -1;
  double ll = -INFINITY;
  for (int c = 0; c < C; ++c) ll = lse2_exact(ll, S.logcw[c] + a.cpt[(size_t)c * nb + col]);
  a.colll[col] = ll;
  const int ocol = a.order[col];
  if (a.col_loglike) a.col_loglike[ocol] = ll;
  if (!want_counts) return;
  double* out = a.partials + (size_t)(first_partial + col) * a.counts_size;
  for (int q = 0; q < a.counts_size; ++q) out[q] = 0;
  if (n_roots != 1) {  // assertSingleRoot
    if (root >= 0) atomicExch(a.err, 1);
    return;
  }
  const double weight = a.weights ? a.weights[ocol] : 1.0;
  out[6] = ll;
  for (int q = 0; q < CA; ++q) {
    const int c = q / A;
    const double norm = exp(S.logcw[c] + a.logF[((size_t)root * C + c) * nb + col] - ll);
    out[7 + q] = weight * S.ins[q] * a.F[((size_t)root * CA + q) * nb + col] * norm;
  }
}

// fillDown of one level (sumprod.cpp:163-198): thread = (column, node of the level, component)
__global__ void __launch_bounds__(DN_COLS) dn_down_kernel(SpDev S, DnArgs a, int lev) {
  const int col = blockIdx.x * blockDim.x + threadIdx.x;
  if (col >= a.nb) return;
  const int A = S.A, C = S.C, nb = a.nb;
  const int node = S.lvl_off[lev] + blockIdx.y, c = blockIdx.z;
  if (a.tok[(size_t)node * nb + col] == SP_TOK_GAP) return;
  const size_t nc = (size_t)node * C + c;
  const int p = S.parent[node];
  double* Gn = a.G + nc * A * nb + col;
  if (p < 0 || a.tok[(size_t)p * nb + col] == SP_TOK_GAP) {
    for (int j = 0; j < A; ++j) Gn[(size_t)j * nb] = S.ins[c * A + j];
    a.logG[nc * nb + col] = 0;
    return;
  }
  const int k0 = S.kid_off[p], k1 = S.kid_off[p + 1];
  double lG = a.logG[((size_t)p * C + c) * nb + col];
  for (int k = k0; k < k1; ++k) {
    const int s = S.kid_idx[k];
    if (s != node && a.tok[(size_t)s * nb + col] != SP_TOK_GAP) lG += a.logE[((size_t)s * C + c) * nb + col];
  }
  a.logG[nc * nb + col] = lG;
  const double* Gp = a.G + ((size_t)p * C + c) * A * nb + col;
  const double* Pn = S.P + nc * A * A;
  // G_p[i] and the product of the siblings' E[i], loaded once per residue.  With one ungapped sibling (binary trees)
  // (G_p[i] * P) * E_s is the reference's own operation order; with more, the siblings are multiplied together first.
  double gp[DN_MAXA], sib[DN_MAXA];
  for (int i = 0; i < A; ++i) {
    gp[i] = Gp[(size_t)i * nb];
    double v = 1;
    bool any = false;
    for (int k = k0; k < k1; ++k) {
      const int s = S.kid_idx[k];
      if (s != node && a.tok[(size_t)s * nb + col] != SP_TOK_GAP) {
        const double e = a.E[(((size_t)s * C + c) * A + i) * nb + col];
        v = any ? v * e : e;
        any = true;
      }
    }
    sib[i] = any ? v : 1.0;
  }
  for (int j = 0; j < A; ++j) {
    double Gj = 0;
    for (int i = 0; i < A; ++i) Gj += (gp[i] * Pn[i * A + j]) * sib[i];
    Gn[(size_t)j * nb] = Gj;
  }
}

// accumulateEigenCounts (sumprod.cpp:294-372): block = DN_COLS columns x DN_NODES nodes x one component; every thread
// accumulates its column's A x A complex contributions over the block's nodes, then the block reduces over columns.
// REAL: the eigen-system is real (every reversible model): the imaginary halves are exact zeros and are not computed.
template <int A, bool REAL>
__global__ void __launch_bounds__(DN_COLS) dn_counts_kernel(SpDev S, DnArgs a, int first_partial) {
  const int C = S.C, nb = a.nb, CA = C * A, CAA = CA * A;
  const int col = blockIdx.x * blockDim.x + threadIdx.x, c = blockIdx.z;
  const int n0 = blockIdx.y * DN_NODES, n1 = min(n0 + DN_NODES, S.N);
  double2 acc[A * A];
#pragma unroll
  for (int q = 0; q < A * A; ++q) acc[q] = make_double2(0, 0);
  // eigenSubCount of the block's branches (model.cpp:1329-1350), shared by every column
  __shared__ double2 s_esc[DN_NODES][A * A];
  for (int q = threadIdx.x; q < (n1 - n0) * A * A; q += blockDim.x) {
    const int n = q / (A * A), kl = q % (A * A);
    s_esc[n][kl] = eigen_sub_count(S, n0 + n, c, kl / A, kl % A);
  }
  __syncthreads();
  if (col < nb) {
    const double colLL = a.colll[col];
    const double weight = a.weights ? a.weights[a.order[col]] : 1.0;
    for (int node = n0; node < n1; ++node) {
      if (a.tok[(size_t)node * nb + col] == SP_TOK_GAP) continue;
      const int p = S.parent[node];
      if (p < 0 || a.tok[(size_t)p * nb + col] == SP_TOK_GAP) continue;   // the column root
      const size_t nc = (size_t)node * C + c;
      const int k0 = S.kid_off[p], k1 = S.kid_off[p + 1];
      double U[A], D[A];
      double maxU = -INFINITY, maxD = -INFINITY, lsib = 0;
      for (int k = k0; k < k1; ++k) {
        const int s = S.kid_idx[k];
        if (s != node && a.tok[(size_t)s * nb + col] != SP_TOK_GAP) lsib += a.logE[((size_t)s * C + c) * nb + col];
      }
#pragma unroll
      for (int i = 0; i < A; ++i) {
        double d = a.G[(((size_t)p * C + c) * A + i) * nb + col];
        for (int k = k0; k < k1; ++k) {
          const int s = S.kid_idx[k];
          if (s != node && a.tok[(size_t)s * nb + col] != SP_TOK_GAP) d *= a.E[(((size_t)s * C + c) * A + i) * nb + col];
        }
        D[i] = d;
        U[i] = a.F[(nc * A + i) * nb + col];
        maxD = fmax(maxD, d);
        maxU = fmax(maxU, U[i]);
      }
      const double norm = exp(colLL - S.logcw[c] - a.logF[nc * nb + col] - a.logG[((size_t)p * C + c) * nb + col] - lsib) / (maxU * maxD);
      const double scale = weight / norm;
#pragma unroll
      for (int i = 0; i < A; ++i) {
        U[i] = U[i] / maxU;
        D[i] = D[i] / maxD;
      }
      double2 ub[A], db[A];
#pragma unroll
      for (int k = 0; k < A; ++k) {
        double2 u = make_double2(0, 0), d = make_double2(0, 0);
        if (REAL) {
#pragma unroll
          for (int b = 0; b < A; ++b) {
            u.x += S.Vinv[((size_t)c * A + k) * A + b].x * U[b];
            d.x += S.V[((size_t)c * A + b) * A + k].x * D[b];
          }
        } else {
#pragma unroll
          for (int b = 0; b < A; ++b) {
            u = cadd(u, cscale(S.Vinv[((size_t)c * A + k) * A + b], U[b]));
            d = cadd(d, cscale(S.V[((size_t)c * A + b) * A + k], D[b]));
          }
        }
        ub[k] = u;
        db[k] = d;
      }
#pragma unroll
      for (int k = 0; k < A; ++k)
#pragma unroll
        for (int l = 0; l < A; ++l) {
          const double2 esc = s_esc[node - n0][k * A + l];
          if (REAL) acc[k * A + l].x += (db[k].x * (esc.x * ub[l].x)) * scale;   // the same products, imaginary parts all zero
          else acc[k * A + l] = cadd(acc[k * A + l], cscale(cmul(db[k], cmul(esc, ub[l])), scale));
        }
    }
  }
  // ---- block reduction over the columns, fixed order: lanes by xor-shuffle, warps in order
  __shared__ double2 s_part[DN_COLS / 32][A * A];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int q = 0; q < A * A; ++q) {
    double2 v = acc[q];
    for (int o = 16; o; o >>= 1) {
      v.x += __shfl_xor_sync(0xffffffffu, v.x, o);
      v.y += __shfl_xor_sync(0xffffffffu, v.y, o);
    }
    if (lane == 0) s_part[warp][q] = v;
  }
  __syncthreads();
  const int part = first_partial + (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
  double* out = a.partials + (size_t)part * a.counts_size;
  for (int q = threadIdx.x; q < a.counts_size; q += blockDim.x) out[q] = 0;
  __syncthreads();
  if (threadIdx.x < A * A) {
    double2 v = make_double2(0, 0);
    for (int w = 0; w < DN_COLS / 32; ++w) v = cadd(v, s_part[w][threadIdx.x]);
    out[7 + CA + 2 * (c * A * A + threadIdx.x)] = v.x;
    out[7 + CA + 2 * (c * A * A + threadIdx.x) + 1] = v.y;
  }
  (void)CAA;
}

// number of columns whose occupancy reaches `threshold` (they are the first entries of the occupancy order)
__global__ void count_dense_kernel(const int* __restrict__ occ, int n_cols, int threshold, int* __restrict__ out) {
  int n = 0;
  for (int c = blockIdx.x * blockDim.x + threadIdx.x; c < n_cols; c += gridDim.x * blockDim.x) n += occ[c] >= threshold;
  for (int o = 16; o; o >>= 1) n += __shfl_down_sync(0xffffffffu, n, o);
  if ((threadIdx.x & 31) == 0 && n) atomicAdd(out, n);
}

// counts[q] += sum over blocks, in block order
__global__ void reduce_partials_kernel(const double* partials, int n_blocks, int counts_size, double* counts) {
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= counts_size) return;
  double v = 0;
  for (int b = 0; b < n_blocks; ++b) v += partials[(size_t)b * counts_size + q];
  counts[q] += v;
}

// the same for thousands of partials: one block per count, fixed summation order (thread t takes partials t, t+256, ...;
// then lanes by shuffle, warps in order)
__global__ void __launch_bounds__(256) reduce_partials_wide_kernel(const double* partials, int n_blocks, int counts_size, double* counts) {
  const int q = blockIdx.x;
  double v = 0;
  for (int b = threadIdx.x; b < n_blocks; b += 256) v += partials[(size_t)b * counts_size + q];
  for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __shared__ double s[8];
  if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = v;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0;
    for (int w = 0; w < 8; ++w) t += s[w];
    counts[q] += t;
  }
}

// EigenModel::getSubCounts (model.cpp:1352-1377): one thread per (c, i, j)
__global__ void transform_counts_kernel(SpDev S, const double* counts, double* sub_count) {
  const int A = S.A, C = S.C;
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= C * A * A) return;
  const int j = q % A, i = (q / A) % A, c = q / (A * A);
  const double* ec = counts + 7 + C * A;  // (re, im) pairs; only 8-byte aligned
  double2 acc = make_double2(0, 0);
  for (int k = 0; k < A; ++k) {
    double2 ck = make_double2(0, 0);
    for (int l = 0; l < A; ++l) {
      const size_t e = 2 * (((size_t)c * A + k) * A + l);
      ck = cadd(ck, cmul(make_double2(ec[e], ec[e + 1]), S.V[((size_t)c * A + j) * A + l]));
    }
    acc = cadd(acc, cmul(S.Vinv[((size_t)c * A + k) * A + i], ck));
  }
  sub_count[q] = acc.x * (i == j ? 1. : S.sub_rate[q]);
}

}  // namespace hb

using namespace hb;

// ------------------------------------------------------------------ host object
struct hist_sumprod {
  hist_ctx* ctx;
  SpDev dev;
  int A, C, N;
  std::vector<int> orig;      // internal -> original
  std::vector<int> internal;  // original -> internal
  char* arena;                // all device tables
  double* work;               // workspace of the column kernel
  size_t work_stride;         // doubles per block
  int work_blocks;
  double* partials;
  double* d_counts;           // staging of a host counts buffer
  int* d_err;
  int* d_order;               // [order_cap] occupancy + processing order scratch (2 * order_cap ints)
  int64_t order_cap;
  int chunk, n_slices;
  size_t smem;
  int max_blocks;
  // dense (node-major) path
  std::vector<int> lvl_off_h;   // host copy of the level ranges
  char* dn_work = nullptr;      // workspace of a batch of dense columns
  size_t dn_work_bytes = 0;
  int* d_ndense = nullptr;
};

namespace {

int gsl_fcmp_host(double x1, double x2, double eps) {  // GSL sys/fcmp.c (used by model.cpp:1334-1335)
  int exponent;
  const double mx = fabs(x1) > fabs(x2) ? x1 : x2;
  frexp(mx, &exponent);
  const double delta = ldexp(eps, exponent), diff = x1 - x2;
  return diff > delta ? 1 : diff < -delta ? -1 : 0;
}

struct ArenaPlan {
  size_t bytes = 0;
  size_t add(size_t n) {
    const size_t at = bytes;
    bytes += (n + 255) & ~(size_t)255;
    return at;
  }
};

size_t sp_smem_bytes(int A, int C, int N, int chunk, int n_slices, int n_levels) {
  const size_t CA = (size_t)C * A, CAA = CA * A;
  size_t b = 16 * n_slices * CAA;           // s_acc
  b += 2 * 16 * chunk * CA;                 // s_Ub, s_Db
  b += 2 * 8 * chunk * CA;                  // s_U, s_D
  b += 8 * (size_t)chunk * C;               // s_scale
  b += 8 * CA + 8 * C + 8 * (SP_THREADS / 32) * C + 8 * 4 + 4 * 4;
  b += (N + 15) & ~15;                      // s_tok
  b += ((size_t)chunk * C + 15) & ~(size_t)15;  // s_valid
  b += 8 * (size_t)n_levels;                // s_lo, s_hi
  return (b + 15) & ~(size_t)15;
}

}  // namespace

extern "C" int hist_sumprod_create(hist_ctx* ctx, const hist_model_desc* m, const hist_tree_desc* tr, hist_sumprod** out) {
  if (!ctx || !m || !tr || !out) return set_err(ctx, HIST_ERR_ARG, "hist_sumprod_create: NULL argument");
  *out = nullptr;
  const int A = m->alph_size, C = m->components, N = tr->n_nodes;
  if (A < 1 || A > 127 || C < 1 || N < 1) return set_err(ctx, HIST_ERR_ARG, "hist_sumprod_create: bad dimensions A=%d C=%d N=%d", A, C, N);
  if (!m->alphabet || !m->cpt_weight || !m->ins_prob || !m->eval_re || !m->eval_im || !m->evec_re || !m->evec_im || !m->evec_inv_re ||
      !m->evec_inv_im || !tr->parent || !tr->branch_length)
    return set_err(ctx, HIST_ERR_ARG, "hist_sumprod_create: NULL table in descriptor");
  // ---- tree: children < parent, root last (tree.cpp:141)
  std::vector<std::vector<int>> kids(N);
  for (int r = 0; r < N; ++r) {
    const int p = tr->parent[r];
    if (r == N - 1 ? p != -1 : (p <= r || p >= N))
      return set_err(ctx, HIST_ERR_ARG, "hist_sumprod_create: tree is not in postorder (node %d has parent %d)", r, p);
    if (p >= 0) kids[p].push_back(r);
  }
  // breadth-first renumbering: depth levels are contiguous, siblings adjacent
  std::vector<int> orig, internal(N, -1), lvl_off;
  orig.reserve(N);
  orig.push_back(N - 1);
  lvl_off.push_back(0);
  size_t head = 0;
  while (head < orig.size()) {
    const size_t end = orig.size();
    lvl_off.push_back((int)end);
    for (; head < end; ++head)
      for (int k : kids[orig[head]]) orig.push_back(k);
  }
  // lvl_off now holds 0, then the end of every level (the last one repeated once): trim
  while (lvl_off.size() >= 2 && lvl_off[lvl_off.size() - 1] == lvl_off[lvl_off.size() - 2]) lvl_off.pop_back();
  if ((int)orig.size() != N) return set_err(ctx, HIST_ERR_ARG, "hist_sumprod_create: tree is not connected");
  const int n_levels = (int)lvl_off.size() - 1;
  for (int q = 0; q < N; ++q) internal[orig[q]] = q;
  std::vector<int> parent(N), kid_off(N + 1, 0), kid_idx;
  std::vector<double> t(N);
  for (int q = 0; q < N; ++q) {
    const int o = orig[q];
    parent[q] = tr->parent[o] < 0 ? -1 : internal[tr->parent[o]];
    t[q] = tr->branch_length[o];
    for (int k : kids[o]) kid_idx.push_back(internal[k]);
    kid_off[q + 1] = (int)kid_idx.size();
  }
  if (kid_idx.empty()) kid_idx.push_back(0);
  std::vector<int> lvl_of(N);
  for (int lev = 0; lev < n_levels; ++lev)
    for (int q = lvl_off[lev]; q < lvl_off[lev + 1]; ++q) lvl_of[q] = lev;
  // ---- model tables
  const size_t CA = (size_t)C * A, CAA = CA * A;
  bool is_real = true;  // EigenModel ctor, model.cpp:1175-1181
  for (size_t q = 0; q < CA; ++q) is_real = is_real && fabs(m->eval_im[q]) < SP_EIGEN_EPS;
  for (size_t q = 0; q < CAA; ++q) is_real = is_real && fabs(m->evec_im[q]) < SP_EIGEN_EPS && fabs(m->evec_inv_im[q]) < SP_EIGEN_EPS;
  std::vector<double2> eval(CA), V(CAA), Vinv(CAA), invdl(CAA);
  std::vector<uint8_t> eveq(CAA);
  std::vector<double> logcw(C);
  for (size_t q = 0; q < CA; ++q) eval[q] = make_double2(m->eval_re[q], m->eval_im[q]);
  for (size_t q = 0; q < CAA; ++q) {
    V[q] = make_double2(m->evec_re[q], m->evec_im[q]);
    Vinv[q] = make_double2(m->evec_inv_re[q], m->evec_inv_im[q]);
  }
  for (int c = 0; c < C; ++c) {
    logcw[c] = log(m->cpt_weight[c]);
    for (int k = 0; k < A; ++k)
      for (int l = 0; l < A; ++l) {
        const double2 a = eval[c * A + k], b = eval[c * A + l];
        const bool eq = k == l || (gsl_fcmp_host(a.x, b.x, SP_EIGEN_EPS) == 0 && gsl_fcmp_host(a.y, b.y, SP_EIGEN_EPS) == 0);
        eveq[((size_t)c * A + k) * A + l] = eq;
        const double dx = a.x - b.x, dy = a.y - b.y, den = dx * dx + dy * dy;
        invdl[((size_t)c * A + k) * A + l] = eq || den == 0 ? make_double2(0, 0) : make_double2(dx / den, -dy / den);
      }
  }
  std::vector<int8_t> tokmap(256, SP_TOK_WILD);
  for (int ch = 1; ch < 256; ++ch) {  // Alphabet tokenize with case folding (fastseq.cpp:10-16)
    const char* p = (const char*)memchr(m->alphabet, ch, A);
    if (!p) {
      const int sw = isupper(ch) ? tolower(ch) : toupper(ch);
      p = sw ? (const char*)memchr(m->alphabet, sw, A) : nullptr;
    }
    if (p) tokmap[ch] = (int8_t)(p - m->alphabet);
  }
  std::vector<double> given;
  if (m->branch_sub) {  // [C][N][A][A] original order -> [N][C][A][A] internal
    given.resize((size_t)N * CAA);
    for (int q = 0; q < N; ++q)
      for (int c = 0; c < C; ++c)
        memcpy(&given[((size_t)q * C + c) * A * A], m->branch_sub + ((size_t)c * N + orig[q]) * A * A, sizeof(double) * A * A);
  }

  hist_sumprod* sp = new hist_sumprod();
  sp->ctx = ctx;
  sp->A = A; sp->C = C; sp->N = N;
  sp->orig = orig;
  sp->internal = internal;
  sp->arena = nullptr; sp->work = nullptr; sp->partials = nullptr; sp->d_counts = nullptr; sp->d_err = nullptr; sp->d_order = nullptr; sp->order_cap = 0;
  sp->work_blocks = 0;
  cudaSetDevice(ctx->device);

  ArenaPlan plan;
  const size_t o_lvl = plan.add(sizeof(int) * (n_levels + 1)), o_par = plan.add(sizeof(int) * N), o_koff = plan.add(sizeof(int) * (N + 1)),
               o_kidx = plan.add(sizeof(int) * kid_idx.size()), o_orig = plan.add(sizeof(int) * N), o_lvlof = plan.add(sizeof(int) * N), o_P = plan.add(8 * (size_t)N * CAA),
               o_PT = plan.add(8 * (size_t)N * CAA), o_eet = plan.add(16 * (size_t)N * CA), o_t = plan.add(8 * (size_t)N),
               o_V = plan.add(16 * CAA), o_Vinv = plan.add(16 * CAA), o_invdl = plan.add(16 * CAA), o_eveq = plan.add(CAA),
               o_ins = plan.add(8 * CA), o_logcw = plan.add(8 * C), o_rate = plan.add(8 * CAA), o_tok = plan.add(256),
               o_alpha = plan.add(A), o_eval = plan.add(16 * CA), o_given = plan.add(given.empty() ? 8 : 8 * given.size());
  cudaError_t e = cudaMalloc(&sp->arena, plan.bytes);
  if (e != cudaSuccess) {
    delete sp;
    return set_err(ctx, HIST_ERR_CUDA, "hist_sumprod_create: cudaMalloc(%zu): %s", plan.bytes, cudaGetErrorString(e));
  }
  char* base = sp->arena;
  auto up = [&](size_t off, const void* src, size_t bytes) { return cudaMemcpyAsync(base + off, src, bytes, cudaMemcpyHostToDevice, ctx->stream); };
  std::vector<double> zero_rate;
  const double* rate = m->sub_rate;
  if (!rate) {
    zero_rate.assign(CAA, 0.0);
    rate = zero_rate.data();
  }
  cudaError_t es[] = {up(o_lvl, lvl_off.data(), sizeof(int) * (n_levels + 1)), up(o_par, parent.data(), sizeof(int) * N),
                      up(o_koff, kid_off.data(), sizeof(int) * (N + 1)), up(o_kidx, kid_idx.data(), sizeof(int) * kid_idx.size()),
                      up(o_orig, orig.data(), sizeof(int) * N), up(o_lvlof, lvl_of.data(), sizeof(int) * N), up(o_t, t.data(), 8 * (size_t)N), up(o_V, V.data(), 16 * CAA),
                      up(o_Vinv, Vinv.data(), 16 * CAA), up(o_invdl, invdl.data(), 16 * CAA), up(o_eveq, eveq.data(), CAA),
                      up(o_ins, m->ins_prob, 8 * CA), up(o_logcw, logcw.data(), 8 * C), up(o_rate, rate, 8 * CAA),
                      up(o_tok, tokmap.data(), 256), up(o_alpha, m->alphabet, A), up(o_eval, eval.data(), 16 * CA),
                      given.empty() ? cudaSuccess : up(o_given, given.data(), 8 * given.size())};
  for (cudaError_t x : es)
    if (x != cudaSuccess) {
      cudaFree(sp->arena);
      delete sp;
      return set_err(ctx, HIST_ERR_CUDA, "hist_sumprod_create: upload: %s", cudaGetErrorString(x));
    }
  SpDev& S = sp->dev;
  S.A = A; S.C = C; S.N = N; S.n_levels = n_levels; S.is_real = is_real;
  S.lvl_off = (const int*)(base + o_lvl); S.parent = (const int*)(base + o_par); S.kid_off = (const int*)(base + o_koff);
  S.kid_idx = (const int*)(base + o_kidx); S.orig = (const int*)(base + o_orig); S.lvl_of = (const int*)(base + o_lvlof); S.P = (const double*)(base + o_P);
  S.PT = (const double*)(base + o_PT); S.eet = (const double2*)(base + o_eet); S.t = (const double*)(base + o_t);
  S.V = (const double2*)(base + o_V); S.Vinv = (const double2*)(base + o_Vinv); S.invdl = (const double2*)(base + o_invdl);
  S.eveq = (const uint8_t*)(base + o_eveq); S.ins = (const double*)(base + o_ins); S.logcw = (const double*)(base + o_logcw);
  S.sub_rate = (const double*)(base + o_rate); S.tokmap = (const int8_t*)(base + o_tok); S.alphabet = base + o_alpha;

  // ---- S0 on device
  const int gb = std::max(1, std::min(ctx->sm_count * 8, (int)(((size_t)N * CAA + 255) / 256)));
  exp_eval_kernel<<<gb, 256, 0, ctx->stream>>>(N, C, A, S.t, (const double2*)(base + o_eval), is_real, (double2*)(base + o_eet));
  branch_tables_kernel<<<gb, 256, 0, ctx->stream>>>(S, (double*)(base + o_P), (double*)(base + o_PT), (const double2*)(base + o_eet),
                                                    (const double2*)(base + o_eval),
                                                    given.empty() ? nullptr : (const double*)(base + o_given));
  ctx->launches += 2;
  e = cudaStreamSynchronize(ctx->stream);
  if (e == cudaSuccess) e = cudaGetLastError();
  if (e != cudaSuccess) {
    cudaFree(sp->arena);
    delete sp;
    return set_err(ctx, HIST_ERR_CUDA, "hist_sumprod_create: branch tables: %s", cudaGetErrorString(e));
  }

  // ---- launch shape of the column kernel
  // aim for several resident blocks per SM (56 KB each); larger alphabets take what they need
  int n_slices = std::max(1, SP_THREADS / (int)CAA);
  int chunk = 256;
  while (chunk > 32 && sp_smem_bytes(A, C, N, chunk, n_slices, n_levels) > 56 * 1024) chunk /= 2;
  while (chunk > 8 && sp_smem_bytes(A, C, N, chunk, n_slices, n_levels) > std::min<size_t>(ctx->smem_optin, 160 * 1024)) chunk /= 2;
  if (sp_smem_bytes(A, C, N, chunk, n_slices, n_levels) > ctx->smem_optin) {
    cudaFree(sp->arena);
    delete sp;
    return set_err(ctx, HIST_ERR_CAPACITY, "hist_sumprod_create: tree/alphabet too large for shared memory (N=%d A=%d C=%d)", N, A, C);
  }
  sp->chunk = chunk;
  sp->n_slices = n_slices;
  sp->smem = sp_smem_bytes(A, C, N, chunk, n_slices, n_levels);
  sp->lvl_off_h = lvl_off;
  e = cudaFuncSetAttribute(sumprod_columns_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sp->smem);
  int per_sm = 1;
  if (e == cudaSuccess) e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, sumprod_columns_kernel, SP_THREADS, sp->smem);
  if (e != cudaSuccess) {
    cudaFree(sp->arena);
    delete sp;
    return set_err(ctx, HIST_ERR_CUDA, "hist_sumprod_create: kernel attributes: %s", cudaGetErrorString(e));
  }
  if (const char* e = getenv("HIST_B200_SP_BLOCKS_PER_SM")) per_sm = std::max(1, std::min(per_sm, atoi(e)));  // tuning knob
  sp->max_blocks = ctx->sm_count * std::max(1, per_sm);
  sp->work_stride = ((size_t)N * (3 * CA + 4 * C) + 31) & ~(size_t)31;
  e = cudaMalloc(&sp->d_err, sizeof(int));
  if (e == cudaSuccess) e = cudaMalloc(&sp->d_counts, 8 * (7 + CA + 2 * CAA));
  if (e != cudaSuccess) {
    cudaFree(sp->arena);
    delete sp;
    return set_err(ctx, HIST_ERR_CUDA, "hist_sumprod_create: cudaMalloc: %s", cudaGetErrorString(e));
  }
  *out = sp;
  return HIST_OK;
}

extern "C" void hist_sumprod_destroy(hist_sumprod* sp) {
  if (!sp) return;
  cudaSetDevice(sp->ctx->device);
  cudaStreamSynchronize(sp->ctx->stream);
  cudaFree(sp->arena);
  cudaFree(sp->work);
  cudaFree(sp->partials);
  cudaFree(sp->d_counts);
  cudaFree(sp->d_err);
  cudaFree(sp->d_order);
  cudaFree(sp->dn_work);
  cudaFree(sp->d_ndense);
  delete sp;
}

extern "C" size_t hist_sumprod_counts_size(const hist_sumprod* sp) {
  return sp ? (size_t)7 + (size_t)sp->C * sp->A + 2 * (size_t)sp->C * sp->A * sp->A : 0;
}

namespace {

bool is_device_ptr(const void* p) {
  if (!p) return false;
  cudaPointerAttributes at;
  if (cudaPointerGetAttributes(&at, p) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged;
}

int ensure_work(hist_sumprod* sp, int blocks) {
  hist_ctx* ctx = sp->ctx;
  if (blocks <= sp->work_blocks) return HIST_OK;
  cudaStreamSynchronize(ctx->stream);
  cudaFree(sp->work);
  cudaFree(sp->partials);
  sp->work = nullptr; sp->partials = nullptr; sp->work_blocks = 0;
  HB_CUDA(ctx, cudaMalloc(&sp->work, 8 * sp->work_stride * blocks));
  HB_CUDA(ctx, cudaMalloc(&sp->partials, 8 * hist_sumprod_counts_size(sp) * blocks));
  sp->work_blocks = blocks;
  return HIST_OK;
}

// Runs the column kernel on device-resident columns; counts_dev (device, accumulated into) may be NULL.
int run_columns(hist_sumprod* sp, const char* d_cols, const double* d_w, int64_t n_cols, int flags, double* d_ll, double* counts_dev,
                char* d_map, double* d_post, int force_blocks, int n_excl = 0, const int* d_excl_q = nullptr, double* d_excl_out = nullptr, int excl_norm = 0) {
  hist_ctx* ctx = sp->ctx;
  const int blocks = force_blocks > 0 ? force_blocks : (int)std::min<int64_t>(n_cols, sp->max_blocks);
  if (int rc = ensure_work(sp, blocks)) return rc;
  HB_CUDA(ctx, cudaMemsetAsync(sp->d_err, 0, sizeof(int), ctx->stream));
  SpArgs a;
  a.columns = d_cols; a.weights = d_w; a.n_cols = n_cols; a.flags = flags; a.col_loglike = d_ll; a.map_state = d_map; a.node_post = d_post;
  a.work = sp->work; a.work_stride = sp->work_stride;
  a.partials = ((flags & HIST_SP_COUNTS) && counts_dev) ? sp->partials : nullptr;
  a.counts_size = (int)hist_sumprod_counts_size(sp);
  a.chunk = sp->chunk; a.n_slices = sp->n_slices; a.err = sp->d_err;
  a.full = force_blocks > 0;  // the debug fetch reads every node's messages
  a.n_excl = n_excl; a.excl_q = d_excl_q; a.excl_out = d_excl_out; a.excl_norm = excl_norm;
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  if (ctx->timing) {
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0, ctx->stream);
  }
  a.order = nullptr;
  // ---- dense columns first, node-major over the whole batch (see dn_* kernels); the rest goes to the column kernel
  const int csz = a.counts_size;
  int widest = 1;
  for (size_t l = 0; l + 1 < sp->lvl_off_h.size(); ++l) widest = std::max(widest, sp->lvl_off_h[l + 1] - sp->lvl_off_h[l]);
  static const int dense_min = getenv("HIST_B200_SP_DENSE_MIN") ? atoi(getenv("HIST_B200_SP_DENSE_MIN")) : 64;
  const bool dense_ok = force_blocks == 0 && sp->A == 4 && sp->C <= 4 && !d_map && !d_post && !d_excl_out && widest <= 65535 && n_cols >= dense_min &&
                        n_cols < (1ll << 30) && dense_min >= 0;
  int64_t n_dense = 0;
  if ((n_cols > blocks || dense_ok) && n_cols < (1ll << 30)) {
    if (sp->order_cap < n_cols) {
      cudaStreamSynchronize(ctx->stream);
      cudaFree(sp->d_order);
      sp->d_order = nullptr; sp->order_cap = 0;
      HB_CUDA(ctx, cudaMalloc(&sp->d_order, 2 * sizeof(int) * (size_t)n_cols));
      sp->order_cap = n_cols;
    }
    int* occ = sp->d_order + sp->order_cap;
    column_occupancy_kernel<<<(int)((n_cols * 32 + 255) / 256), 256, 0, ctx->stream>>>(d_cols, n_cols, sp->N, occ);
    column_order_kernel<<<1, 1024, 0, ctx->stream>>>(occ, (int)n_cols, sp->d_order);
    ctx->launches += 2;
    a.order = sp->d_order;
    if (dense_ok) {
      // the four densest occupancy classes (at least 2^floor(log2 N) / 8 ungapped rows) are a prefix of the order: the
      // full-tree columns, and the few large-subtree insertions whose level walks would otherwise be the column kernel's tail
      int threshold = 1;
      while (threshold * 2 <= sp->N) threshold *= 2;
      threshold = std::max(1, threshold / 8);
      if (!sp->d_ndense) HB_CUDA(ctx, cudaMalloc(&sp->d_ndense, sizeof(int)));
      HB_CUDA(ctx, cudaMemsetAsync(sp->d_ndense, 0, sizeof(int), ctx->stream));
      count_dense_kernel<<<64, 256, 0, ctx->stream>>>(occ, (int)n_cols, threshold, sp->d_ndense);
      ctx->launches += 1;
      int nd = 0;
      HB_CUDA(ctx, cudaMemcpyAsync(&nd, sp->d_ndense, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
      HB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
      n_dense = nd >= dense_min ? nd : 0;
    }
  }
  const bool want_partials = (flags & HIST_SP_COUNTS) && counts_dev;
  for (int64_t b0 = 0; b0 < n_dense;) {
    const int A = sp->A, C = sp->C, N = sp->N, CA = C * A;
    const int nb = (int)std::min<int64_t>(n_dense - b0, 1024);
    const size_t nbs = (size_t)nb;
    const int col_blocks = (nb + DN_COLS - 1) / DN_COLS, node_blocks = (N + DN_NODES - 1) / DN_NODES;
    const int n_part = want_partials ? col_blocks * node_blocks * C + nb : 0;
    ArenaPlan plan;
    const size_t o_tok = plan.add((size_t)N * nbs), o_E = plan.add(8 * (size_t)N * CA * nbs), o_F = plan.add(8 * (size_t)N * CA * nbs),
                 o_G = plan.add(8 * (size_t)N * CA * nbs), o_lE = plan.add(8 * (size_t)N * C * nbs), o_lF = plan.add(8 * (size_t)N * C * nbs),
                 o_lG = plan.add(8 * (size_t)N * C * nbs), o_ll = plan.add(8 * nbs), o_cpt = plan.add(8 * (size_t)C * nbs + 8 * nbs),
                 o_part = plan.add(8 * (size_t)std::max(1, n_part) * csz);
    if (sp->dn_work_bytes < plan.bytes) {
      cudaStreamSynchronize(ctx->stream);
      cudaFree(sp->dn_work);
      sp->dn_work = nullptr; sp->dn_work_bytes = 0;
      HB_CUDA(ctx, cudaMalloc(&sp->dn_work, plan.bytes));
      sp->dn_work_bytes = plan.bytes;
    }
    char* w = sp->dn_work;
    DnArgs d;
    d.columns = d_cols; d.order = sp->d_order + b0; d.weights = d_w; d.nb = nb;
    d.tok = (signed char*)(w + o_tok);
    d.E = (double*)(w + o_E); d.F = (double*)(w + o_F); d.G = (double*)(w + o_G);
    d.logE = (double*)(w + o_lE); d.logF = (double*)(w + o_lF); d.logG = (double*)(w + o_lG);
    d.cpt = (double*)(w + o_cpt); d.n_roots = (int*)(w + o_cpt + 8 * (size_t)C * nbs); d.root_node = d.n_roots + nbs;
    HB_CUDA(ctx, cudaMemsetAsync(w + o_cpt, 0, 8 * (size_t)C * nbs + 8 * nbs, ctx->stream));
    d.colll = (double*)(w + o_ll); d.col_loglike = d_ll; d.partials = (double*)(w + o_part); d.counts_size = csz; d.err = sp->d_err;
    const int n_levels = (int)sp->lvl_off_h.size() - 1;
    dn_init_kernel<<<dim3(col_blocks, N), DN_COLS, 0, ctx->stream>>>(sp->dev, d);
    for (int lev = n_levels - 1; lev >= 0; --lev)
      dn_up_kernel<<<dim3(col_blocks, sp->lvl_off_h[lev + 1] - sp->lvl_off_h[lev], C), DN_COLS, 0, ctx->stream>>>(sp->dev, d, lev);
    dn_root_kernel<<<col_blocks, DN_COLS, 0, ctx->stream>>>(sp->dev, d, want_partials ? 1 : 0, col_blocks * node_blocks * C);
    ctx->launches += 2 + n_levels;
    if (want_partials) {
      for (int lev = 0; lev < n_levels; ++lev)
        dn_down_kernel<<<dim3(col_blocks, sp->lvl_off_h[lev + 1] - sp->lvl_off_h[lev], C), DN_COLS, 0, ctx->stream>>>(sp->dev, d, lev);
      if (sp->dev.is_real) dn_counts_kernel<4, true><<<dim3(col_blocks, node_blocks, C), DN_COLS, 0, ctx->stream>>>(sp->dev, d, 0);
      else dn_counts_kernel<4, false><<<dim3(col_blocks, node_blocks, C), DN_COLS, 0, ctx->stream>>>(sp->dev, d, 0);
      reduce_partials_wide_kernel<<<csz, 256, 0, ctx->stream>>>(d.partials, n_part, csz, counts_dev);
      ctx->launches += 2 + n_levels;
    }
    b0 += nb;
  }
  const int64_t n_rest = n_cols - n_dense;
  if (n_rest > 0) {
    if (a.order) a.order += n_dense;
    a.n_cols = n_rest;
    const int rblocks = (int)std::min<int64_t>(n_rest, blocks);
    sumprod_columns_kernel<<<rblocks, SP_THREADS, sp->smem, ctx->stream>>>(sp->dev, a);
    ctx->launches += 1;
    if (a.partials) {
      reduce_partials_kernel<<<(a.counts_size + 127) / 128, 128, 0, ctx->stream>>>(sp->partials, rblocks, a.counts_size, counts_dev);
      ctx->launches += 1;
    }
  }
  if (ctx->timing) {
    cudaEventRecord(e1, ctx->stream);
    ctx->fill_events.push_back(std::make_pair(e0, e1));
  }
  HB_CUDA(ctx, cudaGetLastError());
  return HIST_OK;
}

}  // namespace

extern "C" int hist_sumprod_columns(hist_sumprod* sp, const char* columns, const double* weights, int64_t n_cols, int32_t flags,
                                    double* col_loglike, double* counts, char* map_state, double* node_post) {
  if (!sp) return HIST_ERR_ARG;
  hist_ctx* ctx = sp->ctx;
  if (n_cols < 0 || (n_cols > 0 && !columns)) return set_err(ctx, HIST_ERR_ARG, "hist_sumprod_columns: bad columns argument");
  if (n_cols == 0) return HIST_OK;
  cudaSetDevice(ctx->device);
  const size_t N = sp->N, A = sp->A, csz = hist_sumprod_counts_size(sp);
  const bool cols_dev = is_device_ptr(columns), counts_dev = is_device_ptr(counts), ll_dev = is_device_ptr(col_loglike),
             w_dev = is_device_ptr(weights), map_dev = is_device_ptr(map_state), post_dev = is_device_ptr(node_post);
  const bool want_counts = (flags & HIST_SP_COUNTS) && counts, want_map = (flags & HIST_SP_MAP_STATES) && map_state,
             want_post = (flags & HIST_SP_NODE_POST) && node_post;
  // one scratch allocation for whatever lives on the host
  ArenaPlan plan;
  const size_t o_cols = cols_dev ? 0 : plan.add(n_cols * N), o_w = (weights && !w_dev) ? plan.add(8 * n_cols) : 0,
               o_ll = (col_loglike && !ll_dev) ? plan.add(8 * n_cols) : 0, o_map = (want_map && !map_dev) ? plan.add(n_cols * N) : 0,
               o_post = (want_post && !post_dev) ? plan.add(8 * n_cols * N * A) : 0;
  char* scratch = nullptr;
  if (plan.bytes) HB_CUDA(ctx, cudaMalloc(&scratch, plan.bytes));
  int rc = HIST_OK;
  auto fail = [&](int code) {
    cudaStreamSynchronize(ctx->stream);
    cudaFree(scratch);
    return code;
  };
#define SP_TRY(call)                                  \
  do {                                                \
    rc = hb::check_cuda(ctx, (call), #call);          \
    if (rc) return fail(rc);                          \
  } while (0)
  const char* d_cols = columns;
  if (!cols_dev) {
    SP_TRY(cudaMemcpyAsync(scratch + o_cols, columns, n_cols * N, cudaMemcpyHostToDevice, ctx->stream));
    d_cols = scratch + o_cols;
  }
  const double* d_w = weights;
  if (weights && !w_dev) {
    SP_TRY(cudaMemcpyAsync(scratch + o_w, weights, 8 * n_cols, cudaMemcpyHostToDevice, ctx->stream));
    d_w = (const double*)(scratch + o_w);
  }
  double* d_ll = col_loglike ? (ll_dev ? col_loglike : (double*)(scratch + o_ll)) : nullptr;
  char* d_map = want_map ? (map_dev ? map_state : scratch + o_map) : nullptr;
  double* d_post = want_post ? (post_dev ? node_post : (double*)(scratch + o_post)) : nullptr;
  double* d_counts = nullptr;
  if (want_counts) {
    d_counts = counts_dev ? counts : sp->d_counts;
    if (!counts_dev) SP_TRY(cudaMemsetAsync(d_counts, 0, 8 * csz, ctx->stream));
  }
  rc = run_columns(sp, d_cols, d_w, n_cols, flags, d_ll, d_counts, d_map, d_post, 0);
  if (rc) return fail(rc);
  int err = 0;
  SP_TRY(cudaMemcpyAsync(&err, sp->d_err, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
  if (col_loglike && !ll_dev) SP_TRY(cudaMemcpyAsync(col_loglike, d_ll, 8 * n_cols, cudaMemcpyDeviceToHost, ctx->stream));
  if (want_map && !map_dev) SP_TRY(cudaMemcpyAsync(map_state, d_map, n_cols * N, cudaMemcpyDeviceToHost, ctx->stream));
  if (want_post && !post_dev) SP_TRY(cudaMemcpyAsync(node_post, d_post, 8 * n_cols * N * A, cudaMemcpyDeviceToHost, ctx->stream));
  std::vector<double> host_counts;
  if (want_counts && !counts_dev) {
    host_counts.resize(csz);
    SP_TRY(cudaMemcpyAsync(host_counts.data(), d_counts, 8 * csz, cudaMemcpyDeviceToHost, ctx->stream));
  }
  // results on the host (or the error flag) need the stream drained; an all-device call stays asynchronous
  // except for the 4-byte flag
  SP_TRY(cudaStreamSynchronize(ctx->stream));
#undef SP_TRY
  cudaFree(scratch);
  if (err) return set_err(ctx, HIST_ERR_ARG, "hist_sumprod_columns: a column has more than one root (assertSingleRoot, sumprod.cpp:273)");
  for (size_t q = 0; q < host_counts.size(); ++q) counts[q] += host_counts[q];
  return HIST_OK;
}

extern "C" int hist_sumprod_excluded_posteriors(hist_sumprod* sp, const char* columns, int64_t n_cols, int32_t n_queries, const int32_t* nodes,
                                                const int32_t* excludes, int32_t normalize, double* out) {
  if (!sp) return HIST_ERR_ARG;
  hist_ctx* ctx = sp->ctx;
  if (n_cols < 0 || (n_cols > 0 && !columns) || n_queries < 1 || !nodes || !excludes || !out)
    return set_err(ctx, HIST_ERR_ARG, "hist_sumprod_excluded_posteriors: bad arguments");
  if (n_cols == 0) return HIST_OK;
  cudaSetDevice(ctx->device);
  const size_t N = sp->N, CA = (size_t)sp->C * sp->A;
  std::vector<int> q(2 * (size_t)n_queries);
  for (int k = 0; k < n_queries; ++k) {
    if (nodes[k] < 0 || nodes[k] >= (int)N || excludes[k] >= (int)N)
      return set_err(ctx, HIST_ERR_ARG, "hist_sumprod_excluded_posteriors: query %d: node %d / %d out of range", k, nodes[k], excludes[k]);
    q[2 * k] = sp->internal[nodes[k]];
    q[2 * k + 1] = excludes[k] < 0 ? -1 : sp->internal[excludes[k]];
  }
  ArenaPlan plan;
  const size_t o_cols = plan.add(n_cols * N), o_q = plan.add(q.size() * sizeof(int)), o_out = plan.add(8 * (size_t)n_queries * n_cols * CA);
  char* scratch = nullptr;
  HB_CUDA(ctx, cudaMalloc(&scratch, plan.bytes));
  int rc = hb::check_cuda(ctx, cudaMemcpyAsync(scratch + o_cols, columns, n_cols * N, cudaMemcpyHostToDevice, ctx->stream), "H2D columns");
  if (!rc) rc = hb::check_cuda(ctx, cudaMemcpyAsync(scratch + o_q, q.data(), q.size() * sizeof(int), cudaMemcpyHostToDevice, ctx->stream), "H2D queries");
  if (!rc) rc = run_columns(sp, scratch + o_cols, nullptr, n_cols, 0, nullptr, nullptr, nullptr, nullptr, 0, n_queries, (const int*)(scratch + o_q), (double*)(scratch + o_out), normalize ? 1 : 0);
  if (!rc) rc = hb::check_cuda(ctx, cudaMemcpyAsync(out, scratch + o_out, 8 * (size_t)n_queries * n_cols * CA, cudaMemcpyDeviceToHost, ctx->stream), "D2H excluded posteriors");
  const int rs = hb::check_cuda(ctx, cudaStreamSynchronize(ctx->stream), "sync after excluded posteriors");
  cudaFree(scratch);
  return rc ? rc : rs;
}

extern "C" int hist_sumprod_debug_column(hist_sumprod* sp, const char* column, double* E, double* F, double* G, double* logE, double* logF,
                                         double* logG, double* col_loglike) {
  if (!sp || !column) return HIST_ERR_ARG;
  hist_ctx* ctx = sp->ctx;
  cudaSetDevice(ctx->device);
  const size_t N = sp->N, A = sp->A, C = sp->C, CA = C * A;
  char* d_col = nullptr;
  double* d_ll = nullptr;
  HB_CUDA(ctx, cudaMalloc(&d_col, N));
  HB_CUDA(ctx, cudaMalloc(&d_ll, 8));
  HB_CUDA(ctx, cudaMemcpyAsync(d_col, column, N, cudaMemcpyHostToDevice, ctx->stream));
  int rc = run_columns(sp, d_col, nullptr, 1, 0, d_ll, nullptr, nullptr, nullptr, 1);
  std::vector<double> w(sp->work_stride);
  double ll = 0;
  if (!rc) rc = hb::check_cuda(ctx, cudaMemcpyAsync(w.data(), sp->work, 8 * sp->work_stride, cudaMemcpyDeviceToHost, ctx->stream), "fetch workspace");
  if (!rc) rc = hb::check_cuda(ctx, cudaMemcpyAsync(&ll, d_ll, 8, cudaMemcpyDeviceToHost, ctx->stream), "fetch colLogLike");
  if (!rc) rc = hb::check_cuda(ctx, cudaStreamSynchronize(ctx->stream), "sync");
  cudaFree(d_col);
  cudaFree(d_ll);
  if (rc) return rc;
  const double* srcs[3] = {w.data(), w.data() + N * CA, w.data() + 2 * N * CA};
  double* dsts[3] = {E, F, G};
  for (int m = 0; m < 3; ++m)
    if (dsts[m])
      for (size_t q = 0; q < N; ++q)
        for (size_t c = 0; c < C; ++c)
          memcpy(dsts[m] + (c * N + sp->orig[q]) * A, srcs[m] + (q * C + c) * A, 8 * A);
  const double* lsrc[3] = {w.data() + 3 * N * CA, w.data() + 3 * N * CA + N * C, w.data() + 3 * N * CA + 2 * N * C};
  double* ldst[3] = {logE, logF, logG};
  for (int m = 0; m < 3; ++m)
    if (ldst[m])
      for (size_t q = 0; q < N; ++q)
        for (size_t c = 0; c < C; ++c) ldst[m][c * N + sp->orig[q]] = lsrc[m][q * C + c];
  if (col_loglike) *col_loglike = ll;
  return HIST_OK;
}

extern "C" int hist_sumprod_branch_tables(hist_sumprod* sp, double* branch_sub, double* branch_esc) {
  if (!sp) return HIST_ERR_ARG;
  hist_ctx* ctx = sp->ctx;
  cudaSetDevice(ctx->device);
  const size_t N = sp->N, A = sp->A, C = sp->C, CA = C * A, CAA = CA * A;
  std::vector<double> P(N * CAA);
  std::vector<double2> eet(N * CA), invdl(CAA);
  std::vector<uint8_t> eveq(CAA);
  std::vector<double> t(N);
  HB_CUDA(ctx, cudaMemcpyAsync(P.data(), sp->dev.P, 8 * P.size(), cudaMemcpyDeviceToHost, ctx->stream));
  HB_CUDA(ctx, cudaMemcpyAsync(eet.data(), sp->dev.eet, 16 * eet.size(), cudaMemcpyDeviceToHost, ctx->stream));
  HB_CUDA(ctx, cudaMemcpyAsync(invdl.data(), sp->dev.invdl, 16 * CAA, cudaMemcpyDeviceToHost, ctx->stream));
  HB_CUDA(ctx, cudaMemcpyAsync(eveq.data(), sp->dev.eveq, CAA, cudaMemcpyDeviceToHost, ctx->stream));
  HB_CUDA(ctx, cudaMemcpyAsync(t.data(), sp->dev.t, 8 * N, cudaMemcpyDeviceToHost, ctx->stream));
  HB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  for (size_t q = 0; q < N; ++q)
    for (size_t c = 0; c < C; ++c) {
      const size_t o = (c * N + sp->orig[q]) * A * A;
      if (branch_sub) memcpy(branch_sub + o, &P[(q * C + c) * A * A], 8 * A * A);
      if (branch_esc)
        for (size_t k = 0; k < A; ++k)
          for (size_t l = 0; l < A; ++l) {
            // the same formula the count phase evaluates on device (eigen_sub_count)
            const double2 ek = eet[(q * C + c) * A + k], el = eet[(q * C + c) * A + l], d = invdl[(c * A + k) * A + l];
            double2 v;
            if (q == 0) v = make_double2(0, 0);
            else if (eveq[(c * A + k) * A + l]) v = make_double2(ek.x * t[q], ek.y * t[q]);
            else {
              const double rx = ek.x - el.x, ry = ek.y - el.y;
              v = make_double2(rx * d.x - ry * d.y, rx * d.y + ry * d.x);
            }
            branch_esc[2 * (o + k * A + l)] = v.x;
            branch_esc[2 * (o + k * A + l) + 1] = v.y;
          }
    }
  return HIST_OK;
}

extern "C" int hist_sumprod_transform_counts(hist_sumprod* sp, const double* counts, double* root_count, double* sub_count) {
  if (!sp || !counts || !sub_count) return HIST_ERR_ARG;
  hist_ctx* ctx = sp->ctx;
  cudaSetDevice(ctx->device);
  const size_t CA = (size_t)sp->C * sp->A, CAA = CA * sp->A, csz = hist_sumprod_counts_size(sp);
  const bool dev = is_device_ptr(counts);
  double* d_in = nullptr;
  double* d_out = nullptr;
  HB_CUDA(ctx, cudaMalloc(&d_out, 8 * CAA));
  const double* src = counts;
  if (!dev) {
    HB_CUDA(ctx, cudaMalloc(&d_in, 8 * csz));
    HB_CUDA(ctx, cudaMemcpyAsync(d_in, counts, 8 * csz, cudaMemcpyHostToDevice, ctx->stream));
    src = d_in;
  }
  transform_counts_kernel<<<(int)((CAA + 127) / 128), 128, 0, ctx->stream>>>(sp->dev, src, d_out);
  ctx->launches += 1;
  int rc = hb::check_cuda(ctx, cudaMemcpyAsync(sub_count, d_out, 8 * CAA, cudaMemcpyDeviceToHost, ctx->stream), "fetch sub counts");
  if (!rc && root_count) rc = hb::check_cuda(ctx, cudaMemcpyAsync(root_count, src + 7, 8 * CA, cudaMemcpyDeviceToHost, ctx->stream), "fetch root counts");
  if (!rc) rc = hb::check_cuda(ctx, cudaStreamSynchronize(ctx->stream), "sync");
  cudaFree(d_in);
  cudaFree(d_out);
  return rc;
}
