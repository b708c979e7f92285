// historian_b200: Forward / Backward fills of the pair-HMM DP matrix on device.
//
// Replaces ForwardMatrix::ForwardMatrix (reference forward.cpp:68-223) and
// BackwardMatrix::BackwardMatrix (forward.cpp:975-1097).
//
// Execution model: one CTA per (pair, direction).  The CTA sweeps the anti-diagonals
// d = i + j of the envelope in order; every transition of a profile has src < dest
// (profile.cpp:345,350), so all predecessors of a cell lie on earlier diagonals, for
// arbitrary state DAGs.  Cells are stored diagonal-major (hist_internal.cuh), so the
// threads of a warp read and write consecutive cells of a diagonal: coalesced 16-B +
// 8-B accesses.  The last RING diagonals are also kept in shared memory, which serves
// every chain-like predecessor (the overwhelming majority: profiles are near-linear
// DAGs) without touching L2; older predecessors come from HBM/L2.  A batch of sibling
// merges is ONE launch (grid = pairs x directions): the band of one merge is only
// ~20-40 cells per diagonal, so the machine is filled by running hundreds of
// wavefronts side by side.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <vector>

#include "hist_internal.cuh"

namespace hb {

#define HB_RING 4  // diagonals kept in shared memory (power of two)

struct Ring {
  float4* A;
  float2* B;
  int cap;   // slots per ring row; 0 = ring disabled
};

__device__ __forceinline__ int cell_slot(const PairDev& P, const DiagInfo& di, int i, int j) {
  const int r = i - di.ilo;
  if (r >= 0 && r < di.width) return P.nS + r;
  int s = P.x.slot[i];
  if (s >= 0) return s;
  s = P.y.slot[j];
  if (s >= 0) return P.nS + di.width + s;
  return -1;
}

// Fetch cell (i,j) written `dist` = |d - dcur| diagonals ago.  Returns false for cells that do
// not exist (outside the matrix / storage) or are all-zero ("-inf" in the reference).
__device__ __forceinline__ bool fetch_cell(const PairDev& P, const float4* __restrict__ gA, const float2* __restrict__ gB,
                                           const Ring& ring, int dcur, int i, int j, float4& a, float2& b) {
  if (i < 0 || j < 0 || i > P.x.n - 2 || j > P.y.n - 2) return false;
  const int d = i + j;
  const DiagInfo di = P.dinfo[d];
  const int s = cell_slot(P, di, i, j);
  if (s < 0) return false;
  const int dist = d > dcur ? d - dcur : dcur - d;
  if (ring.cap > 0 && dist < HB_RING) {
    const int r = (d & (HB_RING - 1)) * ring.cap + s;
    a = ring.A[r];
    b = ring.B[r];
  } else {
    a = gA[di.off + s];
    b = gB[di.off + s];
  }
  return __float_as_int(b.y) > HB_EXP_ZERO / 2;
}

__device__ __forceinline__ float absorb_hat(const PairDev& P, int i, int j) {
  // sum_k Xhat[i][k] * (root[k] * Yhat[j][k])  ==  exp(computeLogProbAbsorb(i,j)) * 2^-(ex[i]+ey[j])   (forward.h:112-124)
  const float* __restrict__ hx = P.x.H + (size_t)i * P.K;
  const float* __restrict__ hy = P.y.H + (size_t)j * P.K;
  float acc = 0.f;
  if ((P.K & 3) == 0) {
    const float4* x4 = reinterpret_cast<const float4*>(hx);
    const float4* y4 = reinterpret_cast<const float4*>(hy);
    for (int k = 0; k < P.K / 4; ++k) {
      const float4 u = __ldg(x4 + k), v = __ldg(y4 + k);
      acc = fmaf(u.x, v.x, acc);
      acc = fmaf(u.y, v.y, acc);
      acc = fmaf(u.z, v.z, acc);
      acc = fmaf(u.w, v.w, acc);
    }
  } else {
    for (int k = 0; k < P.K; ++k) acc = fmaf(__ldg(hx + k), __ldg(hy + k), acc);
  }
  return acc;
}

// ------------------------------------------------------------------ Forward cell (forward.cpp:86-200)
__device__ Cell forward_cell(const PairDev& P, const Ring& ring, int d, int i, int j) {
  const unsigned fx = P.x.flags[i], fy = P.y.flags[j];
  const bool xEmit = fx & F_EMIT, yEmit = fy & F_EMIT;
  const bool xReady = fx & F_READY, yReady = fy & F_READY;
  const bool xEmitOrStart = fx & F_EMIT_OR_START;
  Cell a = zero_cell();
  if (i == 0 && j == 0) {  // lpStart() = 0 (forward.cpp:73)
    a.v[0] = 1.f;
    a.E = 0;
  }
  const float (*W)[6] = P.W;
  float4 sa;
  float2 sb;
  // x-moves: sources (src_x, j)   (forward.cpp:95-129, 193-199)
  if (yReady) {
    const bool immByX = !xEmit && !(!yEmit && xEmitOrStart);  // third IMM branch
    for (int q = P.x.in_off[i]; q < P.x.in_off[i + 1]; ++q) {
      const int t = P.x.in_idx[q];
      if (!fetch_cell(P, P.fA, P.fB, ring, d, P.x.t_src[t], j, sa, sb)) continue;
      const float s = acc_align(a, __float_as_int(sb.y) + P.x.t_e[t]) * P.x.t_m[t];
      if (xEmit) {
        a.v[1] = fmaf(s, sa.x * W[0][1] + sa.y * W[1][1] + sa.z * W[2][1] + sa.w * W[3][1], a.v[1]);
        a.v[4] = fmaf(s, sa.x * W[0][4] + sa.w * W[3][4] + sb.x * W[4][4], a.v[4]);
      } else {
        a.v[1] = fmaf(s, sa.y, a.v[1]);
        a.v[4] = fmaf(s, sb.x, a.v[4]);
        if (immByX) a.v[0] = fmaf(s, sa.x, a.v[0]);
      }
    }
  }
  // y-moves: sources (i, src_y)   (forward.cpp:131-163, 186-191)
  if (yEmit ? xReady : true) {
    for (int q = P.y.in_off[j]; q < P.y.in_off[j + 1]; ++q) {
      const int t = P.y.in_idx[q];
      if (!fetch_cell(P, P.fA, P.fB, ring, d, i, P.y.t_src[t], sa, sb)) continue;
      const float s = acc_align(a, __float_as_int(sb.y) + P.y.t_e[t]) * P.y.t_m[t];
      if (yEmit) {
        a.v[2] = fmaf(s, sa.x * W[0][2] + sa.y * W[1][2] + sa.z * W[2][2] + sb.x * W[4][2], a.v[2]);
        a.v[3] = fmaf(s, sa.x * W[0][3] + sa.w * W[3][3], a.v[3]);
      } else {
        a.v[2] = fmaf(s, sa.z, a.v[2]);
        a.v[3] = fmaf(s, sa.w, a.v[3]);
        if (xEmitOrStart) a.v[0] = fmaf(s, sa.x, a.v[0]);
      }
    }
  }
  // xy-moves into IMM   (forward.cpp:165-184)
  if (xEmit && yEmit) {
    for (int qx = P.x.in_off[i]; qx < P.x.in_off[i + 1]; ++qx) {
      const int tx = P.x.in_idx[qx];
      const int sx = P.x.t_src[tx];
      const float mx = P.x.t_m[tx];
      const int ex = P.x.t_e[tx];
      for (int qy = P.y.in_off[j]; qy < P.y.in_off[j + 1]; ++qy) {
        const int ty = P.y.in_idx[qy];
        if (!fetch_cell(P, P.fA, P.fB, ring, d, sx, P.y.t_src[ty], sa, sb)) continue;
        const float s = acc_align(a, __float_as_int(sb.y) + ex + P.y.t_e[ty]) * mx * P.y.t_m[ty];
        a.v[0] = fmaf(s, sa.x * W[0][0] + sa.y * W[1][0] + sa.z * W[2][0] + sa.w * W[3][0] + sb.x * W[4][0], a.v[0]);
      }
    }
    a.v[0] *= absorb_hat(P, i, j);
  }
  // emissions are applied after the predecessor sums (forward.cpp:116-117, 151-152)
  if (xEmit && yReady) {
    a.v[1] *= P.x.rs[i];
    a.v[4] *= P.x.ins[i];
  }
  if (yEmit && xReady) {
    a.v[2] *= P.y.rs[j];
    a.v[3] *= P.y.ins[j];
  }
  cell_normalise(a);
  return a;
}

// ------------------------------------------------------------------ Backward cell (forward.cpp:1008-1086)
__device__ Cell backward_cell(const PairDev& P, const Ring& ring, int d, int i, int j) {
  const unsigned fx = P.x.flags[i], fy = P.y.flags[j];
  const bool xReady = fx & F_READY, yReady = fy & F_READY;
  const bool xEmitOrStart = fx & F_EMIT_OR_START;
  const int xEnd = P.x.n - 1, yEnd = P.y.n - 1;
  Cell a = zero_cell();
  const float (*W)[6] = P.W;
  float4 sa;
  float2 sb;
  // xy-absorbing transitions into IMM (forward.cpp:1017-1030)
  for (int qx = P.x.ao_off[i]; qx < P.x.ao_off[i + 1]; ++qx) {
    const int tx = P.x.ao_idx[qx];
    const int dx = P.x.t_dst[tx];
    for (int qy = P.y.ao_off[j]; qy < P.y.ao_off[j + 1]; ++qy) {
      const int ty = P.y.ao_idx[qy];
      const int dy = P.y.t_dst[ty];
      if (!fetch_cell(P, P.bA, P.bB, ring, d, dx, dy, sa, sb)) continue;
      const float s = acc_align(a, __float_as_int(sb.y) + P.x.t_e[tx] + P.y.t_e[ty]) * P.x.t_m[tx] * P.y.t_m[ty];
      const float dest = s * absorb_hat(P, dx, dy) * sa.x;
#pragma unroll
      for (int k = 0; k < 5; ++k) a.v[k] = fmaf(W[k][0], dest, a.v[k]);
    }
  }
  // x-absorbing transitions into IMD, IIW (forward.cpp:1032-1048)
  if (yReady)
    for (int qx = P.x.ao_off[i]; qx < P.x.ao_off[i + 1]; ++qx) {
      const int tx = P.x.ao_idx[qx];
      const int dx = P.x.t_dst[tx];
      if (!fetch_cell(P, P.bA, P.bB, ring, d, dx, j, sa, sb)) continue;
      const float s = acc_align(a, __float_as_int(sb.y) + P.x.t_e[tx]) * P.x.t_m[tx];
      const float dmd = s * P.x.rs[dx] * sa.y, diw = s * P.x.ins[dx] * sb.x;
      a.v[0] = fmaf(W[0][1], dmd, fmaf(W[0][4], diw, a.v[0]));
      a.v[1] = fmaf(W[1][1], dmd, a.v[1]);
      a.v[2] = fmaf(W[2][1], dmd, a.v[2]);
      a.v[3] = fmaf(W[3][1], dmd, fmaf(W[3][4], diw, a.v[3]));
      a.v[4] = fmaf(W[4][4], diw, a.v[4]);
    }
  // y-absorbing transitions into IDM, IMI (forward.cpp:1050-1065)
  if (xReady)
    for (int qy = P.y.ao_off[j]; qy < P.y.ao_off[j + 1]; ++qy) {
      const int ty = P.y.ao_idx[qy];
      const int dy = P.y.t_dst[ty];
      if (!fetch_cell(P, P.bA, P.bB, ring, d, i, dy, sa, sb)) continue;
      const float s = acc_align(a, __float_as_int(sb.y) + P.y.t_e[ty]) * P.y.t_m[ty];
      const float ddm = s * P.y.rs[dy] * sa.z, dmi = s * P.y.ins[dy] * sa.w;
      a.v[0] = fmaf(W[0][2], ddm, fmaf(W[0][3], dmi, a.v[0]));
      a.v[1] = fmaf(W[1][2], ddm, a.v[1]);
      a.v[2] = fmaf(W[2][2], ddm, a.v[2]);
      a.v[4] = fmaf(W[4][2], ddm, a.v[4]);
      a.v[3] = fmaf(W[3][3], dmi, a.v[3]);
    }
  // x-nonabsorbing transitions in IMD, IIW, IMM (forward.cpp:1067-1075)
  if (yReady)
    for (int qx = P.x.no_off[i]; qx < P.x.no_off[i + 1]; ++qx) {
      const int tx = P.x.no_idx[qx];
      if (!fetch_cell(P, P.bA, P.bB, ring, d, P.x.t_dst[tx], j, sa, sb)) continue;
      const float s = acc_align(a, __float_as_int(sb.y) + P.x.t_e[tx]) * P.x.t_m[tx];
      a.v[1] = fmaf(s, sa.y, a.v[1]);
      a.v[4] = fmaf(s, sb.x, a.v[4]);
      a.v[0] = fmaf(s, sa.x, a.v[0]);
    }
  // y-nonabsorbing transitions in IDM, IMI, IMM (forward.cpp:1077-1085)
  for (int qy = P.y.no_off[j]; qy < P.y.no_off[j + 1]; ++qy) {
    const int ty = P.y.no_idx[qy];
    if (!fetch_cell(P, P.bA, P.bB, ring, d, i, P.y.t_dst[ty], sa, sb)) continue;
    const float s = acc_align(a, __float_as_int(sb.y) + P.y.t_e[ty]) * P.y.t_m[ty];
    a.v[2] = fmaf(s, sa.z, a.v[2]);
    a.v[3] = fmaf(s, sa.w, a.v[3]);
    if (xEmitOrStart) a.v[0] = fmaf(s, sa.x, a.v[0]);
  }
  // transitions into EEE: the initial assignment of forward.cpp:981-995, which the sweep
  // then accumulates onto
  for (int qx = P.x.no_off[i]; qx < P.x.no_off[i + 1]; ++qx) {
    const int tx = P.x.no_idx[qx];
    if (P.x.t_dst[tx] != xEnd) continue;
    for (int qy = P.y.no_off[j]; qy < P.y.no_off[j + 1]; ++qy) {
      const int ty = P.y.no_idx[qy];
      if (P.y.t_dst[ty] != yEnd) continue;
      const float s = acc_align(a, P.x.t_e[tx] + P.y.t_e[ty]) * P.x.t_m[tx] * P.y.t_m[ty];
#pragma unroll
      for (int k = 0; k < 5; ++k) a.v[k] = fmaf(s, W[k][5], a.v[k]);
    }
  }
  cell_normalise(a);
  return a;
}

__device__ __forceinline__ double lse_exact(double a, double b) {
  if (a == -INFINITY) return b;
  if (b == -INFINITY) return a;
  const double m = fmax(a, b);
  return m + log1p(exp(-fabs(a - b)));
}

// ------------------------------------------------------------------ the wavefront kernel
// grid.x = n_pairs * n_dir; block b handles pair b / n_dir, direction (b % n_dir) (0 = Forward, 1 = Backward;
// with n_dir == 1 the direction is `only_dir`).
__global__ void __launch_bounds__(512) fill_kernel(PairDev* const* __restrict__ pairs, int n_dir, int only_dir, int ring_cap) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  __shared__ PairDev P;
  {
    const PairDev* src = pairs[blockIdx.x / n_dir];
    const int* s = reinterpret_cast<const int*>(src);
    int* dptr = reinterpret_cast<int*>(&P);
    for (int k = threadIdx.x; k < (int)(sizeof(PairDev) / sizeof(int)); k += blockDim.x) dptr[k] = s[k];
  }
  __syncthreads();
  const bool bwd = n_dir == 2 ? (blockIdx.x % n_dir) == 1 : only_dir == 1;
  Ring ring;
  ring.cap = ring_cap;
  ring.A = reinterpret_cast<float4*>(smem_raw);
  ring.B = reinterpret_cast<float2*>(smem_raw + (size_t)HB_RING * ring_cap * sizeof(float4));
  float4* gA = bwd ? P.bA : P.fA;
  float2* gB = bwd ? P.bB : P.fB;
  const int nx2 = P.x.n - 2, ny2 = P.y.n - 2;
  const bool one_warp = blockDim.x == 32;

  for (int step = 0; step < P.n_diag; ++step) {
    const int d = bwd ? P.n_diag - 1 - step : step;
    const DiagInfo di = P.dinfo[d];
    const int nthr = P.nS + di.width + P.nE;
    for (int t = threadIdx.x; t < nthr; t += blockDim.x) {
      int i, j;
      bool ok = true;
      if (t < P.nS) {
        i = P.start_rows[t];
        j = d - i;
        if (i >= di.ilo && i < di.ilo + di.width) ok = false;  // the main interval owns this cell
      } else if (t < P.nS + di.width) {
        i = di.ilo + (t - P.nS);
        j = d - i;
      } else {
        j = P.end_cols[t - P.nS - di.width];
        i = d - j;
        if (i >= di.ilo && i < di.ilo + di.width) ok = false;
        if (ok && i >= 0 && i <= nx2 && P.x.slot[i] >= 0) ok = false;  // a near-start slot owns it
      }
      ok = ok && i >= 0 && j >= 0 && i <= nx2 && j <= ny2;
      if (ok)  // DPMatrix::inEnvelope (forward.h:92-98)
        ok = P.x.near_[i] || P.y.near_[j] || P.max_dist < 0 || hb_iabs(P.x.key[i] - P.y.key[j]) <= P.max_dist;
      Cell c = ok ? (bwd ? backward_cell(P, ring, d, i, j) : forward_cell(P, ring, d, i, j)) : zero_cell();
      const float4 va = make_float4(c.v[0], c.v[1], c.v[2], c.v[3]);
      const float2 vb = make_float2(c.v[4], __int_as_float(c.E));
      gA[di.off + t] = va;
      gB[di.off + t] = vb;
      if (ring_cap > 0) {
        const int r = (d & (HB_RING - 1)) * ring_cap + t;
        ring.A[r] = va;
        ring.B[r] = vb;
      }
    }
    if (one_warp) __syncwarp(); else __syncthreads();
  }

  if (threadIdx.x == 0) {
    Ring none;
    none.cap = 0;
    none.A = nullptr;
    none.B = nullptr;
    float4 sa;
    float2 sb;
    if (!bwd) {
      // transitions into EEE (forward.cpp:205-220)
      double lp = -INFINITY;
      const int xe = P.x.n - 1, ye = P.y.n - 1;
      for (int qx = P.x.in_off[xe]; qx < P.x.in_off[xe + 1]; ++qx) {
        const int tx = P.x.in_idx[qx];
        for (int qy = P.y.in_off[ye]; qy < P.y.in_off[ye + 1]; ++qy) {
          const int ty = P.y.in_idx[qy];
          if (!fetch_cell(P, gA, gB, none, 1 << 30, P.x.t_src[tx], P.y.t_src[ty], sa, sb)) continue;
          const double s = (double)sa.x * P.W[0][5] + (double)sa.y * P.W[1][5] + (double)sa.z * P.W[2][5] +
                           (double)sa.w * P.W[3][5] + (double)sb.x * P.W[4][5];
          if (!(s > 0)) continue;
          const double term = log(s * (double)P.x.t_m[tx] * (double)P.y.t_m[ty]) +
                              (double)(__float_as_int(sb.y) + P.x.t_e[tx] + P.y.t_e[ty]) * HB_LN2;
          lp = lse_exact(lp, term);
        }
      }
      *P.lp_end = lp;
    } else {
      double lp = -INFINITY;
      if (fetch_cell(P, gA, gB, none, 1 << 30, 0, 0, sa, sb)) lp = cell_log(sa.x, __float_as_int(sb.y));
      *P.lp_start = lp;  // BackwardMatrix::lpStart() = cell(0,0,IMM)
    }
  }
}

// Convert stored cells to the reference's representation: natural-log doubles in the row-major
// order of DPMatrix::write (forward.cpp:448-464).  One thread per (i,j); a prefix count over the
// envelope predicate gives each cell its output rank.
__global__ void cells_rank_kernel(const PairDev* __restrict__ Pp, unsigned* __restrict__ row_count) {
  const PairDev& P = *Pp;
  const int i = blockIdx.x;
  const int NJ = P.y.n - 1;
  __shared__ unsigned s_cnt;
  if (threadIdx.x == 0) s_cnt = 0;
  __syncthreads();
  unsigned local = 0;
  for (int j = threadIdx.x; j < NJ; j += blockDim.x)
    if (P.x.near_[i] || P.y.near_[j] || P.max_dist < 0 || hb_iabs(P.x.key[i] - P.y.key[j]) <= P.max_dist) ++local;
  atomicAdd(&s_cnt, local);
  __syncthreads();
  if (threadIdx.x == 0) row_count[i] = s_cnt;
}

__global__ void cells_export_kernel(const PairDev* __restrict__ Pp, int which, const unsigned long long* __restrict__ row_start,
                                    double* __restrict__ lp, int* __restrict__ xpos, int* __restrict__ ypos) {
  const PairDev& P = *Pp;
  const int i = blockIdx.x;
  const int NJ = P.y.n - 1;
  const float4* gA = which ? P.bA : P.fA;
  const float2* gB = which ? P.bB : P.fB;
  // single warp per row: ordered compaction with ballot
  unsigned long long base = row_start[i];
  for (int j0 = 0; j0 < NJ; j0 += 32) {
    const int j = j0 + threadIdx.x;
    bool env = false;
    if (j < NJ) env = P.x.near_[i] || P.y.near_[j] || P.max_dist < 0 || hb_iabs(P.x.key[i] - P.y.key[j]) <= P.max_dist;
    const unsigned mask = __ballot_sync(0xffffffffu, env);
    if (env) {
      const unsigned long long k = base + __popc(mask & ((1u << threadIdx.x) - 1));
      const DiagInfo di = P.dinfo[i + j];
      const int s = cell_slot(P, di, i, j);
      double v[5] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY, -INFINITY};
      if (s >= 0) {
        const float4 a = gA[di.off + s];
        const float2 b = gB[di.off + s];
        const int E = __float_as_int(b.y);
        if (E > HB_EXP_ZERO / 2) {
          v[0] = cell_log(a.x, E); v[1] = cell_log(a.y, E); v[2] = cell_log(a.z, E);
          v[3] = cell_log(a.w, E); v[4] = cell_log(b.x, E);
        }
      }
      for (int q = 0; q < 5; ++q) lp[k * 5 + q] = v[q];
      if (xpos) xpos[k] = i;
      if (ypos) ypos[k] = j;
    }
    base += __popc(mask);
  }
}

}  // namespace hb

using namespace hb;

namespace {

int ensure_backward_planes(hist_pair* pr) {
  if (pr->dev.bA) return HIST_OK;
  hist_ctx* ctx = pr->ctx;
  const size_t n = pr->dev.n_stored;
  const size_t bytesA = (n * sizeof(float4) + 255) / 256 * 256;
  char* p = nullptr;
  HB_CUDA(ctx, cudaMallocAsync((void**)&p, bytesA + n * sizeof(float2), ctx->stream));
  pr->dev.bA = (float4*)p;
  pr->dev.bB = (float2*)(p + bytesA);
  HB_CUDA(ctx, cudaMemcpyAsync(pr->d_dev, &pr->dev, sizeof(PairDev), cudaMemcpyHostToDevice, ctx->stream));
  return HIST_OK;
}

int launch_fill(hist_ctx* ctx, hist_pair* const* pairs, int n, int n_dir, int only_dir) {
  int maxThreads = 1;
  for (int k = 0; k < n; ++k) maxThreads = std::max(maxThreads, pairs[k]->dev.max_threads);
  int block = std::min(512, (maxThreads + 31) / 32 * 32);
  // shared-memory ring of the last HB_RING diagonals, if it fits
  int ring_cap = maxThreads;
  size_t smem = (size_t)HB_RING * ring_cap * (sizeof(float4) + sizeof(float2));
  const size_t limit = ctx->smem_optin > 4096 ? ctx->smem_optin - 2048 : 0;
  if (smem > limit) {
    ring_cap = 0;
    smem = 0;
  }
  static size_t configured = 0;
  if (smem > 48 * 1024 && smem > configured) {
    HB_CUDA(ctx, cudaFuncSetAttribute(fill_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)limit));
    configured = limit;
  }
  std::vector<PairDev*> h(n);
  for (int k = 0; k < n; ++k) h[k] = pairs[k]->d_dev;
  PairDev** dptrs = nullptr;
  HB_CUDA(ctx, cudaMallocAsync((void**)&dptrs, n * sizeof(PairDev*), ctx->stream));
  HB_CUDA(ctx, cudaMemcpyAsync(dptrs, h.data(), n * sizeof(PairDev*), cudaMemcpyHostToDevice, ctx->stream));
  cudaEvent_t e0 = nullptr, e1 = nullptr;
  if (ctx->timing) {
    HB_CUDA(ctx, cudaEventCreate(&e0));
    HB_CUDA(ctx, cudaEventCreate(&e1));
    HB_CUDA(ctx, cudaEventRecord(e0, ctx->stream));
  }
  fill_kernel<<<n * n_dir, block, smem, ctx->stream>>>(dptrs, n_dir, only_dir, ring_cap);
  ctx->launches += 1;
  if (ctx->timing) {
    HB_CUDA(ctx, cudaEventRecord(e1, ctx->stream));
    ctx->fill_events.push_back({e0, e1});
  }
  HB_CUDA(ctx, cudaGetLastError());
  HB_CUDA(ctx, cudaFreeAsync(dptrs, ctx->stream));
  // h[] is pageable host memory: the copy above must finish before it goes out of scope
  return HIST_OK;
}

}  // namespace

extern "C" int hist_fill_batch(hist_ctx* ctx, hist_pair* const* pairs, int32_t n, int do_backward, double* lp_end, double* lp_start) {
  if (!ctx || !pairs || n <= 0 || !lp_end) return set_err(ctx, HIST_ERR_ARG, "hist_fill_batch: bad arguments");
  cudaSetDevice(ctx->device);
  for (int k = 0; k < n; ++k) {
    if (!pairs[k] || pairs[k]->ctx != ctx) return set_err(ctx, HIST_ERR_ARG, "hist_fill_batch: pair %d belongs to another context", k);
    if (do_backward) {
      int rc = ensure_backward_planes(pairs[k]);
      if (rc) return rc;
    }
  }
  int rc = launch_fill(ctx, pairs, n, do_backward ? 2 : 1, 0);
  if (rc) return rc;
  std::vector<double> tmp(2 * (size_t)n);
  for (int k = 0; k < n; ++k) {
    HB_CUDA(ctx, cudaMemcpyAsync(&tmp[2 * k], pairs[k]->dev.lp_end, (do_backward ? 2 : 1) * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
  }
  HB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  for (int k = 0; k < n; ++k) {
    pairs[k]->fwd_done = true;
    pairs[k]->lp_end = lp_end[k] = tmp[2 * k];
    if (do_backward) {
      pairs[k]->bwd_done = true;
      pairs[k]->lp_start = tmp[2 * k + 1];
      if (lp_start) lp_start[k] = tmp[2 * k + 1];
    }
  }
  return HIST_OK;
}

extern "C" int hist_fwd_fill(hist_pair* pair, double* lp_end) {
  if (!pair || !lp_end) return HIST_ERR_ARG;
  hist_pair* arr[1] = {pair};
  return hist_fill_batch(pair->ctx, arr, 1, 0, lp_end, nullptr);
}

extern "C" int hist_bwd_fill(hist_pair* pair, double* lp_start) {
  if (!pair || !lp_start) return HIST_ERR_ARG;
  hist_ctx* ctx = pair->ctx;
  if (!pair->fwd_done) return set_err(ctx, HIST_ERR_STATE, "hist_bwd_fill: BackwardMatrix needs a filled ForwardMatrix (call hist_fwd_fill first)");
  cudaSetDevice(ctx->device);
  int rc = ensure_backward_planes(pair);
  if (rc) return rc;
  hist_pair* arr[1] = {pair};
  rc = launch_fill(ctx, arr, 1, 1, 1);
  if (rc) return rc;
  HB_CUDA(ctx, cudaMemcpyAsync(lp_start, pair->dev.lp_start, sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
  HB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  pair->bwd_done = true;
  pair->lp_start = *lp_start;
  return HIST_OK;
}

extern "C" int hist_cells_fetch(hist_pair* pair, int which, double* lp, int32_t* xpos, int32_t* ypos, int64_t capacity, int64_t* n_cells) {
  if (!pair || !n_cells) return HIST_ERR_ARG;
  hist_ctx* ctx = pair->ctx;
  *n_cells = pair->info.n_cells;
  if (which ? !pair->bwd_done : !pair->fwd_done) return set_err(ctx, HIST_ERR_STATE, "hist_cells_fetch: matrix %d has not been filled", which);
  if (capacity < pair->info.n_cells || !lp) return set_err(ctx, HIST_ERR_CAPACITY, "hist_cells_fetch: need room for %lld cells", (long long)pair->info.n_cells);
  cudaSetDevice(ctx->device);
  const int NI = pair->info.x_size - 1;
  const size_t nc = (size_t)pair->info.n_cells;
  unsigned* d_cnt = nullptr;
  unsigned long long* d_start = nullptr;
  double* d_lp = nullptr;
  int* d_pos = nullptr;
  HB_CUDA(ctx, cudaMallocAsync((void**)&d_cnt, NI * sizeof(unsigned), ctx->stream));
  HB_CUDA(ctx, cudaMallocAsync((void**)&d_start, NI * sizeof(unsigned long long), ctx->stream));
  HB_CUDA(ctx, cudaMallocAsync((void**)&d_lp, std::max<size_t>(1, nc) * 5 * sizeof(double), ctx->stream));
  HB_CUDA(ctx, cudaMallocAsync((void**)&d_pos, std::max<size_t>(1, nc) * 2 * sizeof(int), ctx->stream));
  cells_rank_kernel<<<NI, 128, 0, ctx->stream>>>(pair->d_dev, d_cnt);
  std::vector<unsigned> cnt(NI);
  HB_CUDA(ctx, cudaMemcpyAsync(cnt.data(), d_cnt, NI * sizeof(unsigned), cudaMemcpyDeviceToHost, ctx->stream));
  HB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  std::vector<unsigned long long> start(NI);
  unsigned long long acc = 0;
  for (int i = 0; i < NI; ++i) { start[i] = acc; acc += cnt[i]; }
  HB_CUDA(ctx, cudaMemcpyAsync(d_start, start.data(), NI * sizeof(unsigned long long), cudaMemcpyHostToDevice, ctx->stream));
  cells_export_kernel<<<NI, 32, 0, ctx->stream>>>(pair->d_dev, which, d_start, d_lp, d_pos, d_pos + nc);
  ctx->launches += 2;
  HB_CUDA(ctx, cudaGetLastError());
  HB_CUDA(ctx, cudaMemcpyAsync(lp, d_lp, nc * 5 * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
  if (xpos) HB_CUDA(ctx, cudaMemcpyAsync(xpos, d_pos, nc * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
  if (ypos) HB_CUDA(ctx, cudaMemcpyAsync(ypos, d_pos + nc, nc * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
  HB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  cudaFreeAsync(d_cnt, ctx->stream);
  cudaFreeAsync(d_start, ctx->stream);
  cudaFreeAsync(d_lp, ctx->stream);
  cudaFreeAsync(d_pos, ctx->stream);
  if (acc != nc) return set_err(ctx, HIST_ERR_STATE, "hist_cells_fetch: envelope count mismatch (%llu vs %zu)", acc, nc);
  return HIST_OK;
}
