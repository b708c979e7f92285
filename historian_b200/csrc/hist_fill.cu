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
#include <cstdlib>
#include <cstring>
#include <vector>

#include "hist_internal.cuh"

namespace hb {

#define HB_RING_MAX 16  // most diagonals kept in shared memory (power of two)

// Shared-memory view of one strip: the last R diagonals of cells ("ring") plus their layout records.
struct Ring {
  float4* A;     // [R][cap] (IMM, IMD, IDM, IMI)
  float4* B;     // [R][cap] (IIW, exponent bits, Backward only: IMM * absorb, 0)
  int4* info;    // [R] {ilo, width, off, 0} of the diagonal held in each ring row
  int cap;       // slots per ring row
  int R;         // ring rows (0 = ring disabled), power of two
  const int* srows;  // near-start rows / near-end columns of the edge slots (shared-memory copies in the
  const int* ecols;  // narrow kernel, the global arrays otherwise)
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

// Layout of the two most recently finished diagonals (d -/+ 1, d -/+ 2 in sweep direction), kept in
// registers: they hold nearly every predecessor, so the common fetch needs no layout lookup.
struct Recent {
  int ilo1, w1, ilo2, w2;
};

// Fetch cell (i2,j2), whose edge slots (xs for row i2, ys for column j2) the caller already holds.
// Returns false when !en, for cells that are not stored, or all-zero cells ("-inf" in the reference).
template <bool BWD>
__device__ __forceinline__ bool fetch_cell(const PairDev& P, const Ring& ring, const Recent& rc, int dcur, bool en, int i2, int j2,
                                           int xs, int ys, float4& a, float4& b) {
  if (!en) return false;
  const int d2 = i2 + j2;
  const int dist = BWD ? d2 - dcur : dcur - d2;
  const bool near = dist < ring.R;
  int ilo, width;
  unsigned off = 0;
  if (near) {
    if (dist <= 2) {
      ilo = dist == 1 ? rc.ilo1 : rc.ilo2;
      width = dist == 1 ? rc.w1 : rc.w2;
    } else {
      const int4 ri = ring.info[d2 & (ring.R - 1)];
      ilo = ri.x; width = ri.y;
    }
  } else {
    const DiagInfo di = P.dinfo[d2];
    ilo = di.ilo; width = di.width; off = di.off;
  }
  const int r = i2 - ilo;
  int s;
  if ((unsigned)r < (unsigned)width) s = P.nS + r;
  else if (xs >= 0) s = xs;
  else if (ys >= 0) s = P.nS + width + ys;
  else return false;
  if (near) {
    const int q = (d2 & (ring.R - 1)) * ring.cap + s;
    a = ring.A[q];
    b = ring.B[q];
  } else {
    const float4* gA = BWD ? P.bA : P.fA;
    const float2* gB = BWD ? P.bB : P.fB;
    a = gA[off + s];
    const float2 t = gB[off + s];
    b = make_float4(t.x, t.y, BWD ? a.x * P.absorb[off + s] : 0.f, 0.f);
  }
  return __float_as_int(b.y) > HB_EXP_ZERO / 2;
}

// ------------------------------------------------------------------ Forward cell (forward.cpp:86-200)
// xa/xb/xe, ya/yb/ye: packed records of x state i and y state j; ab = absorb plane value of this cell.
// The contributions through the FIRST in-edge of each state (the only one for almost every state) are
// fetched together and combined under one common exponent, so the three cell reads and the FMA groups
// are independent instructions; further in-edges go through the generic aligned accumulation.
__device__ __forceinline__ Cell forward_cell(const PairDev& P, const Ring& ring, const Recent& rc, int d, int i, int j,
                                             const int4 xa, const int4 xb, const float2 xe, const int4 ya, const int4 yb,
                                             const float2 ye, const float ab) {
  const unsigned fx = xa.x & 0xff, fy = ya.x & 0xff;
  const int xc = (int)((unsigned)xa.x >> 8), yc = (int)((unsigned)ya.x >> 8);
  const bool xEmit = fx & F_EMIT, yEmit = fy & F_EMIT;
  const bool xReady = fx & F_READY, yReady = fy & F_READY;
  const bool xEmitOrStart = fx & F_EMIT_OR_START;
  const bool immByX = !xEmit && !(!yEmit && xEmitOrStart);  // third IMM branch (forward.cpp:193-199)
  const bool doX = yReady && xc > 0;                         // x-moves: sources (src_x, j)   (forward.cpp:95-129)
  const bool doY = (yEmit ? xReady : true) && yc > 0;        // y-moves: sources (i, src_y)   (forward.cpp:131-163)
  const bool doD = xEmit && yEmit && xc > 0 && yc > 0;       // xy-moves into IMM              (forward.cpp:165-184)
  const float (*W)[6] = P.W;
  float4 xA, xB, yA, yB, dA, dB;
  const bool vX = fetch_cell<false>(P, ring, rc, d, doX, xb.x, j, xb.w, ya.w, xA, xB);
  const bool vY = fetch_cell<false>(P, ring, rc, d, doY, i, yb.x, xa.w, yb.w, yA, yB);
  const bool vD = fetch_cell<false>(P, ring, rc, d, doD, xb.x, yb.x, xb.w, yb.w, dA, dB);
  const bool origin = i == 0 && j == 0;  // lpStart() = 0 (forward.cpp:73)
  const int eX = vX ? __float_as_int(xB.y) + xb.y : HB_EXP_ZERO;
  const int eY = vY ? __float_as_int(yB.y) + yb.y : HB_EXP_ZERO;
  const int eD = vD ? __float_as_int(dB.y) + xb.y + yb.y : HB_EXP_ZERO;
  Cell a;
  a.E = max(max(eX, eY), max(eD, origin ? 0 : HB_EXP_ZERO));
  const float sX = vX ? pow2_neg(eX - a.E) * __int_as_float(xb.z) : 0.f;
  const float sY = vY ? pow2_neg(eY - a.E) * __int_as_float(yb.z) : 0.f;
  const float sD = vD ? pow2_neg(eD - a.E) * __int_as_float(xb.z) * __int_as_float(yb.z) : 0.f;
  a.v[0] = origin ? pow2_neg(-a.E) : 0.f;
  a.v[1] = a.v[2] = a.v[3] = a.v[4] = 0.f;
  if (vX) {
    if (xEmit) {
      a.v[1] = sX * (xA.x * W[0][1] + xA.y * W[1][1] + xA.z * W[2][1] + xA.w * W[3][1]);
      a.v[4] = sX * (xA.x * W[0][4] + xA.w * W[3][4] + xB.x * W[4][4]);
    } else {
      a.v[1] = sX * xA.y;
      a.v[4] = sX * xB.x;
      if (immByX) a.v[0] = fmaf(sX, xA.x, a.v[0]);
    }
  }
  if (vY) {
    if (yEmit) {
      a.v[2] = sY * (yA.x * W[0][2] + yA.y * W[1][2] + yA.z * W[2][2] + yB.x * W[4][2]);
      a.v[3] = sY * (yA.x * W[0][3] + yA.w * W[3][3]);
    } else {
      a.v[2] = sY * yA.z;
      a.v[3] = sY * yA.w;
      if (xEmitOrStart) a.v[0] = fmaf(sY, yA.x, a.v[0]);
    }
  }
  if (vD) a.v[0] = fmaf(sD, dA.x * W[0][0] + dA.y * W[1][0] + dA.z * W[2][0] + dA.w * W[3][0] + dB.x * W[4][0], a.v[0]);
  // ---- further in-edges (state DAGs)
  float4 sa, sb;
  if (doX)
    for (int q = 1; q < xc; ++q) {
      const int4 e = P.x.in_edges[xa.y + q];
      if (!fetch_cell<false>(P, ring, rc, d, true, e.x, j, e.w, ya.w, sa, sb)) continue;
      const float s = acc_align(a, __float_as_int(sb.y) + e.y) * __int_as_float(e.z);
      if (xEmit) {
        a.v[1] = fmaf(s, sa.x * W[0][1] + sa.y * W[1][1] + sa.z * W[2][1] + sa.w * W[3][1], a.v[1]);
        a.v[4] = fmaf(s, sa.x * W[0][4] + sa.w * W[3][4] + sb.x * W[4][4], a.v[4]);
      } else {
        a.v[1] = fmaf(s, sa.y, a.v[1]);
        a.v[4] = fmaf(s, sb.x, a.v[4]);
        if (immByX) a.v[0] = fmaf(s, sa.x, a.v[0]);
      }
    }
  if (doY)
    for (int q = 1; q < yc; ++q) {
      const int4 e = P.y.in_edges[ya.y + q];
      if (!fetch_cell<false>(P, ring, rc, d, true, i, e.x, xa.w, e.w, sa, sb)) continue;
      const float s = acc_align(a, __float_as_int(sb.y) + e.y) * __int_as_float(e.z);
      if (yEmit) {
        a.v[2] = fmaf(s, sa.x * W[0][2] + sa.y * W[1][2] + sa.z * W[2][2] + sb.x * W[4][2], a.v[2]);
        a.v[3] = fmaf(s, sa.x * W[0][3] + sa.w * W[3][3], a.v[3]);
      } else {
        a.v[2] = fmaf(s, sa.z, a.v[2]);
        a.v[3] = fmaf(s, sa.w, a.v[3]);
        if (xEmitOrStart) a.v[0] = fmaf(s, sa.x, a.v[0]);
      }
    }
  if (doD && xc + yc > 2) {
    int4 ex = xb;
    for (int qx = 0; qx < xc;) {
      int4 ey = yb;
      for (int qy = 0; qy < yc;) {
        if ((qx | qy) && fetch_cell<false>(P, ring, rc, d, true, ex.x, ey.x, ex.w, ey.w, sa, sb)) {
          const float s = acc_align(a, __float_as_int(sb.y) + ex.y + ey.y) * __int_as_float(ex.z) * __int_as_float(ey.z);
          a.v[0] = fmaf(s, sa.x * W[0][0] + sa.y * W[1][0] + sa.z * W[2][0] + sa.w * W[3][0] + sb.x * W[4][0], a.v[0]);
        }
        if (++qy < yc) ey = P.y.in_edges[ya.y + qy];
      }
      if (++qx < xc) ex = P.x.in_edges[xa.y + qx];
    }
  }
  if (xEmit && yEmit) a.v[0] *= ab;
  // emissions are applied after the predecessor sums (forward.cpp:116-117, 151-152)
  if (xEmit && yReady) {
    a.v[1] *= xe.x;
    a.v[4] *= xe.y;
  }
  if (yEmit && xReady) {
    a.v[2] *= ye.x;
    a.v[3] *= ye.y;
  }
  cell_normalise(a);
  return a;
}

// ------------------------------------------------------------------ Backward cell (forward.cpp:1008-1086)
// xa,xn,x0,x1 / ya,yn,y0,y1: the four backward records of x state i / y state j.
__device__ __forceinline__ Cell backward_cell(const PairDev& P, const Ring& ring, const Recent& rc, int d, int i, int j, const int4 xa,
                                              const int4 xn, const int4 x0, const int4 x1, const int4 ya, const int4 yn,
                                              const int4 y0, const int4 y1) {
  const unsigned fx = xa.x & 0xff, fy = ya.x & 0xff;
  const int xac = (int)((unsigned)xa.x >> 8), yac = (int)((unsigned)ya.x >> 8);
  const bool xReady = fx & F_READY, yReady = fy & F_READY;
  const bool xEmitOrStart = fx & F_EMIT_OR_START;
  const float (*W)[6] = P.W;
  const bool doD = xac > 0 && yac > 0;   // xy-absorbing transitions into IMM (forward.cpp:1017-1030)
  const bool doX = yReady && xac > 0;    // x-absorbing transitions into IMD, IIW (forward.cpp:1032-1048)
  const bool doY = xReady && yac > 0;    // y-absorbing transitions into IDM, IMI (forward.cpp:1050-1065)
  // transitions into EEE: the initial assignment of forward.cpp:981-995, which the sweep then accumulates onto
  const bool toEnd = (fx & F_TO_END) && (fy & F_TO_END);
  float4 dA, dB, xA, xB, yA, yB;
  const bool vD = fetch_cell<true>(P, ring, rc, d, doD, x0.x, y0.x, x0.z, y0.z, dA, dB);
  const bool vX = fetch_cell<true>(P, ring, rc, d, doX, x0.x, j, x0.z, ya.w, xA, xB);
  const bool vY = fetch_cell<true>(P, ring, rc, d, doY, i, y0.x, xa.w, y0.z, yA, yB);
  const int eD = vD ? __float_as_int(dB.y) + x0.y + y0.y : HB_EXP_ZERO;
  const int eX = vX ? __float_as_int(xB.y) + x0.y : HB_EXP_ZERO;
  const int eY = vY ? __float_as_int(yB.y) + y0.y : HB_EXP_ZERO;
  const int eE = toEnd ? xn.z + yn.z : HB_EXP_ZERO;
  Cell a;
  a.E = max(max(eD, eX), max(eY, eE));
  const float dest = vD ? pow2_neg(eD - a.E) * __int_as_float(x1.x) * __int_as_float(y1.x) * dB.z : 0.f;  // absorb(dx,dy) * B(dx,dy,IMM)
  const float sX = vX ? pow2_neg(eX - a.E) : 0.f;
  const float sY = vY ? pow2_neg(eY - a.E) : 0.f;
  const float sE = toEnd ? pow2_neg(eE - a.E) * __int_as_float(xn.w) * __int_as_float(yn.w) : 0.f;
  const float dmd = vX ? sX * __int_as_float(x1.y) * xA.y : 0.f, diw = vX ? sX * __int_as_float(x1.z) * xB.x : 0.f;
  const float ddm = vY ? sY * __int_as_float(y1.y) * yA.z : 0.f, dmi = vY ? sY * __int_as_float(y1.z) * yA.w : 0.f;
  a.v[0] = W[0][0] * dest + W[0][1] * dmd + W[0][4] * diw + W[0][2] * ddm + W[0][3] * dmi + sE * W[0][5];
  a.v[1] = W[1][0] * dest + W[1][1] * dmd + W[1][2] * ddm + sE * W[1][5];
  a.v[2] = W[2][0] * dest + W[2][1] * dmd + W[2][2] * ddm + sE * W[2][5];
  a.v[3] = W[3][0] * dest + W[3][1] * dmd + W[3][4] * diw + W[3][3] * dmi + sE * W[3][5];
  a.v[4] = W[4][0] * dest + W[4][4] * diw + W[4][2] * ddm + sE * W[4][5];
  // ---- further absorbing out-edges (state DAGs)
  float4 sa, sb;
  if (doD && xac + yac > 2) {
    int4 ex0 = x0, ex1 = x1;
    for (int qx = 0; qx < xac;) {
      int4 ey0 = y0, ey1 = y1;
      for (int qy = 0; qy < yac;) {
        if ((qx | qy) && fetch_cell<true>(P, ring, rc, d, true, ex0.x, ey0.x, ex0.z, ey0.z, sa, sb)) {
          const float s = acc_align(a, __float_as_int(sb.y) + ex0.y + ey0.y) * __int_as_float(ex1.x) * __int_as_float(ey1.x);
          const float dst = s * sb.z;
#pragma unroll
          for (int k = 0; k < 5; ++k) a.v[k] = fmaf(W[k][0], dst, a.v[k]);
        }
        if (++qy < yac) { ey0 = P.y.ao_edges[2 * (ya.y + qy)]; ey1 = P.y.ao_edges[2 * (ya.y + qy) + 1]; }
      }
      if (++qx < xac) { ex0 = P.x.ao_edges[2 * (xa.y + qx)]; ex1 = P.x.ao_edges[2 * (xa.y + qx) + 1]; }
    }
  }
  if (doX)
    for (int q = 1; q < xac; ++q) {
      const int4 e0 = P.x.ao_edges[2 * (xa.y + q)], e1 = P.x.ao_edges[2 * (xa.y + q) + 1];
      if (!fetch_cell<true>(P, ring, rc, d, true, e0.x, j, e0.z, ya.w, sa, sb)) continue;
      const float s = acc_align(a, __float_as_int(sb.y) + e0.y);
      const float md = s * __int_as_float(e1.y) * sa.y, iw = s * __int_as_float(e1.z) * sb.x;
      a.v[0] = fmaf(W[0][1], md, fmaf(W[0][4], iw, a.v[0]));
      a.v[1] = fmaf(W[1][1], md, a.v[1]);
      a.v[2] = fmaf(W[2][1], md, a.v[2]);
      a.v[3] = fmaf(W[3][1], md, fmaf(W[3][4], iw, a.v[3]));
      a.v[4] = fmaf(W[4][4], iw, a.v[4]);
    }
  if (doY)
    for (int q = 1; q < yac; ++q) {
      const int4 e0 = P.y.ao_edges[2 * (ya.y + q)], e1 = P.y.ao_edges[2 * (ya.y + q) + 1];
      if (!fetch_cell<true>(P, ring, rc, d, true, i, e0.x, xa.w, e0.z, sa, sb)) continue;
      const float s = acc_align(a, __float_as_int(sb.y) + e0.y);
      const float dm = s * __int_as_float(e1.y) * sa.z, mi = s * __int_as_float(e1.z) * sa.w;
      a.v[0] = fmaf(W[0][2], dm, fmaf(W[0][3], mi, a.v[0]));
      a.v[1] = fmaf(W[1][2], dm, a.v[1]);
      a.v[2] = fmaf(W[2][2], dm, a.v[2]);
      a.v[4] = fmaf(W[4][2], dm, a.v[4]);
      a.v[3] = fmaf(W[3][3], mi, a.v[3]);
    }
  // x-nonabsorbing transitions in IMD, IIW, IMM (forward.cpp:1067-1075); transitions into END carry no cell
  if (yReady)
    for (int q = 0; q < xn.y; ++q) {
      const int4 e0 = P.x.no_edges[2 * (xn.x + q)], e1 = P.x.no_edges[2 * (xn.x + q) + 1];
      if (e0.w) continue;  // -> END
      if (!fetch_cell<true>(P, ring, rc, d, true, e0.x, j, e0.z, ya.w, sa, sb)) continue;
      const float s = acc_align(a, __float_as_int(sb.y) + e0.y) * __int_as_float(e1.x);
      a.v[1] = fmaf(s, sa.y, a.v[1]);
      a.v[4] = fmaf(s, sb.x, a.v[4]);
      a.v[0] = fmaf(s, sa.x, a.v[0]);
    }
  // y-nonabsorbing transitions in IDM, IMI, IMM (forward.cpp:1077-1085)
  for (int q = 0; q < yn.y; ++q) {
    const int4 e0 = P.y.no_edges[2 * (yn.x + q)], e1 = P.y.no_edges[2 * (yn.x + q) + 1];
    if (e0.w) continue;
    if (!fetch_cell<true>(P, ring, rc, d, true, i, e0.x, xa.w, e0.z, sa, sb)) continue;
    const float s = acc_align(a, __float_as_int(sb.y) + e0.y) * __int_as_float(e1.x);
    a.v[2] = fmaf(s, sa.z, a.v[2]);
    a.v[3] = fmaf(s, sa.w, a.v[3]);
    if (xEmitOrStart) a.v[0] = fmaf(s, sa.x, a.v[0]);
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

// Map thread slot t of diagonal d to its cell; false if the slot owns no in-envelope cell.
// xk / yk receive the first packed record of the two states (flags, band key, edge slot).
template <bool BWD>
__device__ __forceinline__ bool slot_cell(const PairDev& P, const DiagInfo& di, int d, int t, int& i, int& j, int4& xk, int4& yk) {
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
  }
  if (!ok || i < 0 || j < 0 || i > P.x.n - 2 || j > P.y.n - 2) return false;
  xk = BWD ? P.x.bm[i] : P.x.fm_a[i];
  yk = BWD ? P.y.bm[j] : P.y.fm_a[j];
  if (t >= P.nS + di.width && xk.w >= 0) return false;  // a near-start slot owns it
  // DPMatrix::inEnvelope (forward.h:92-98)
  return ((xk.x | yk.x) & F_NEAR) || P.max_dist < 0 || hb_iabs(xk.z - yk.z) <= P.max_dist;
}

// ------------------------------------------------------------------ absorb plane
// exp(computeLogProbAbsorb(i,j)) (forward.h:112-124) for every stored both-emit cell, as
// sum_k Xhat[i][k] * (root[k] Yhat[j][k]) relative to 2^(ex[i]+ey[j]).  A banded GEMM on CUDA cores:
// a CTA takes HB_ABS_DIAGS consecutive anti-diagonals, stages the rows of the two emission tables
// they touch in shared memory once (coalesced), then one warp per diagonal computes the dot products
// from shared memory (the row of a state is reused by every cell of its band row: ~40x) and writes
// consecutive floats.  Slots whose rows fall outside the staged tiles (the near-start / near-end
// edge slots, or diagonals wider than a tile) read their rows from global memory.
// grid = (diagonal blocks, pairs); dynamic smem = 2 * tile_rows * Kp floats.
#ifndef HB_ABS_DIAGS
#define HB_ABS_DIAGS 32
#endif

__device__ __forceinline__ float dot_rows(const float4* __restrict__ x4, const float4* __restrict__ y4, int K4) {
  float a0 = 0.f, a1 = 0.f;
#pragma unroll 5
  for (int k = 0; k < K4; ++k) {
    const float4 u = x4[k], v = y4[k];
    a0 = fmaf(u.x, v.x, a0);
    a1 = fmaf(u.y, v.y, a1);
    a0 = fmaf(u.z, v.z, a0);
    a1 = fmaf(u.w, v.w, a1);
  }
  return a0 + a1;
}

__global__ void __launch_bounds__(256) absorb_kernel(PairDev* const* __restrict__ pairs, int tile_rows) {
  extern __shared__ __align__(16) float tile[];
  __shared__ PairDev P;
  __shared__ int s_rng[4];  // rlo, rhi, clo, chi of the staged block
  {
    const int* s = reinterpret_cast<const int*>(pairs[blockIdx.y]);
    int* dptr = reinterpret_cast<int*>(&P);
    for (int k = threadIdx.x; k < (int)(sizeof(PairDev) / sizeof(int)); k += blockDim.x) dptr[k] = s[k];
  }
  __syncthreads();
  const int d0 = blockIdx.x * HB_ABS_DIAGS;
  if (d0 >= P.n_diag) return;
  const int d1 = min(d0 + HB_ABS_DIAGS, P.n_diag);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (warp == 0) {
    int rlo = 1 << 30, rhi = -1, clo = 1 << 30, chi = -1;
    if (d0 + lane < d1) {
      const DiagInfo di = P.dinfo[d0 + lane];
      if (di.width > 0) {
        rlo = di.ilo; rhi = di.ilo + di.width - 1;
        clo = d0 + lane - rhi; chi = d0 + lane - rlo;
      }
    }
    for (int o = 16; o; o >>= 1) {
      rlo = min(rlo, __shfl_xor_sync(0xffffffffu, rlo, o)); rhi = max(rhi, __shfl_xor_sync(0xffffffffu, rhi, o));
      clo = min(clo, __shfl_xor_sync(0xffffffffu, clo, o)); chi = max(chi, __shfl_xor_sync(0xffffffffu, chi, o));
    }
    if (lane == 0) {
      // keep what fits: rows beyond the tile are served from global memory
      s_rng[0] = rlo; s_rng[1] = min(rhi, rlo + tile_rows - 1);
      s_rng[2] = clo; s_rng[3] = min(chi, clo + tile_rows - 1);
    }
  }
  __syncthreads();
  const int rlo = s_rng[0], rhi = s_rng[1], clo = s_rng[2], chi = s_rng[3];
  const int Kp = P.Kp, K4 = Kp >> 2;
  float4* tx = reinterpret_cast<float4*>(tile);
  float4* ty = tx + (size_t)tile_rows * K4;
  if (rhi >= rlo) {
    const float4* gx = reinterpret_cast<const float4*>(P.x.H + (size_t)rlo * Kp);
    const float4* gy = reinterpret_cast<const float4*>(P.y.H + (size_t)clo * Kp);
    const int nx4 = (rhi - rlo + 1) * K4, ny4 = (chi - clo + 1) * K4;
    for (int k = threadIdx.x; k < nx4; k += blockDim.x) tx[k] = __ldg(gx + k);
    for (int k = threadIdx.x; k < ny4; k += blockDim.x) ty[k] = __ldg(gy + k);
  }
  __syncthreads();
  for (int d = d0 + warp; d < d1; d += (blockDim.x >> 5)) {
    const DiagInfo di = P.dinfo[d];
    const int nthr = P.nS + di.width + P.nE;
    for (int t = lane; t < nthr; t += 32) {
      int i, j;
      int4 xk, yk;
      float acc = 0.f;
      if (slot_cell<false>(P, di, d, t, i, j, xk, yk) && (xk.x & yk.x & F_EMIT)) {
        const float4* x4 = (i >= rlo && i <= rhi) ? tx + (size_t)(i - rlo) * K4 : reinterpret_cast<const float4*>(P.x.H + (size_t)i * Kp);
        const float4* y4 = (j >= clo && j <= chi) ? ty + (size_t)(j - clo) * K4 : reinterpret_cast<const float4*>(P.y.H + (size_t)j * Kp);
        acc = dot_rows(x4, y4, K4);
      }
      P.absorb[di.off + t] = acc;
    }
  }
}

// ------------------------------------------------------------------ the wavefront kernel
// Everything a lane needs for one cell that does NOT depend on earlier cells: coordinates, packed
// state records, absorb value.  Loaded one step ahead of its use in the narrow sweep.
template <bool BWD>
struct LaneMeta {
  int i, j;
  bool ok;       // slot owns an in-range cell (the envelope test still follows, it needs the records)
  bool tail;     // slot is a near-end column slot
  int4 xa, xb, ya, yb;   // Forward: fm_a, fm_b;  Backward: bm[0], bm[1]
  int4 x0, x1, y0, y1;   // Backward only: first absorbing out-edge
  float2 xe, ye;         // Forward only
  float ab;
};

template <bool BWD>
__device__ __forceinline__ LaneMeta<BWD> load_meta(const PairDev& P, const Ring& ring, const DiagInfo& di, int d, int t, bool with_ab = true) {
  LaneMeta<BWD> m;
  bool ok = t < P.nS + di.width + P.nE;
  int i = 0, j = 0;
  m.tail = false;
  if (t < P.nS) {
    i = ring.srows[t];
    j = d - i;
    if (i >= di.ilo && i < di.ilo + di.width) ok = false;  // the main interval owns this cell
  } else if (t < P.nS + di.width) {
    i = di.ilo + (t - P.nS);
    j = d - i;
  } else if (ok) {
    j = ring.ecols[t - P.nS - di.width];
    i = d - j;
    m.tail = true;
    if (i >= di.ilo && i < di.ilo + di.width) ok = false;
  }
  ok = ok && i >= 0 && j >= 0 && i <= P.x.n - 2 && j <= P.y.n - 2;
  if (!ok) i = j = 0;
  m.i = i; m.j = j; m.ok = ok;
  // whole 16-B records through the read-only path: one LDG.128 each (plain struct copies get split into
  // 4/8-byte loads of the fields in use, several times the L1 wavefronts)
  if (BWD) {
    const int4* xm = P.x.bm + i;   // four planes of n records: consecutive lanes read consecutive 16-B records
    const int4* ym = P.y.bm + j;
    const int nx = P.x.n, ny = P.y.n;
    m.xa = __ldg(xm); m.xb = __ldg(xm + nx); m.x0 = __ldg(xm + 2 * nx); m.x1 = __ldg(xm + 3 * nx);
    m.ya = __ldg(ym); m.yb = __ldg(ym + ny); m.y0 = __ldg(ym + 2 * ny); m.y1 = __ldg(ym + 3 * ny);
  } else {
    m.xa = __ldg(P.x.fm_a + i); m.xb = __ldg(P.x.fm_b + i); m.xe = __ldg(P.x.emis + i);
    m.ya = __ldg(P.y.fm_a + j); m.yb = __ldg(P.y.fm_b + j); m.ye = __ldg(P.y.emis + j);
  }
  m.ab = ok && with_ab ? __ldcs(P.absorb + di.off + t) : 0.f;
  return m;
}

template <bool BWD>
__device__ __forceinline__ void compute_store(const PairDev& P, const Ring& ring, const Recent& rc, const DiagInfo& di, int d, int t,
                                              const LaneMeta<BWD>& m, float4* gA, float2* gB) {
  Cell c = zero_cell();
  float aux = 0.f;
  // a near-start slot owns cells of near-end columns; then DPMatrix::inEnvelope (forward.h:92-98)
  const bool ok = m.ok && !(m.tail && m.xa.w >= 0) &&
                  (((m.xa.x | m.ya.x) & F_NEAR) || P.max_dist < 0 || hb_iabs(m.xa.z - m.ya.z) <= P.max_dist);
  if (ok) {
    if (BWD) {
      c = backward_cell(P, ring, rc, d, m.i, m.j, m.xa, m.xb, m.x0, m.x1, m.ya, m.yb, m.y0, m.y1);
      aux = c.v[0] * m.ab;  // absorb is 0 unless both states emit
    } else {
      c = forward_cell(P, ring, rc, d, m.i, m.j, m.xa, m.xb, m.xe, m.ya, m.yb, m.ye, m.ab);
    }
  }
  const float4 va = make_float4(c.v[0], c.v[1], c.v[2], c.v[3]);
  __stcs(gA + di.off + t, va);
  __stcs(gB + di.off + t, make_float2(c.v[4], __int_as_float(c.E)));
  if (ring.R > 0) {
    const int r = (d & (ring.R - 1)) * ring.cap + t;
    ring.A[r] = va;
    ring.B[r] = make_float4(c.v[4], __int_as_float(c.E), aux, 0.f);
  }
}

// ------------------------------------------------------------------ narrow strips: warp-specialised sweep
// Banded merges are thin strips: ~20-40 cells per anti-diagonal, thousands of dependent steps.  The
// time of a step is the latency of ONE warp's instruction stream, so the stream of the warp that
// carries the dependency is cut to the arithmetic: a CONSUMER warp (one cell per lane per step) reads
// its predecessors from the shared-memory ring, does the FMAs, normalises and stores.  Everything
// that does not depend on earlier cells -- slot -> (i,j), packed state records, envelope test, ring
// addresses of the predecessors, transition mantissas/exponents, emission factors and the absorb
// dot product (forward.h:112-124) -- is prepared by HB_NPROD PRODUCER warps, each working
// HB_NPROD steps ahead, and handed over as a 64-B per-lane record through a shared-memory queue
// guarded by named barriers (bar.arrive / bar.sync full/empty pairs).  Cells with more than one
// in-/out-edge, null out-edges or predecessors older than the ring fall back to the generic cell
// routine on the consumer ("slow" flag).
// measured on B200 (tools/ab_variants.sh): 1 producer + 128 registers (8 CTAs/SM: 1024 strips in ONE wave,
// no spills) 3.27 ms; 2 producers + 96 registers (6 CTAs/SM, two waves, spills) 3.6 ms
#ifndef HB_NPROD
#define HB_NPROD 1
#endif
#ifndef HB_DEPTH
#define HB_DEPTH 4
#endif
#ifndef HB_NARROW_REGS
#define HB_NARROW_REGS (HB_NPROD == 1 ? 128 : 96)
#endif
enum { R_OK = 1, R_XEMIT = 2, R_YEMIT = 4, R_IMM_BY_X = 8, R_IMM_BY_Y = 16, R_ORIGIN = 32, R_SLOW = 64, R_BOTH = 128, R_TO_END = 256 };

// shared-memory mbarriers (no per-SM limit, unlike named barriers): full[slot] counts the 32 lanes of
// the producing warp, empty[slot] the 32 lanes of the consumer
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"((unsigned)__cvta_generic_to_shared(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("{\n .reg .b64 st;\n mbarrier.arrive.shared::cta.b64 st, [%0];\n}" ::"r"((unsigned)__cvta_generic_to_shared(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, unsigned parity) {
  asm volatile(
      "{\n .reg .pred p;\n WAIT_LOOP:\n mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n @p bra WAIT_DONE;\n bra WAIT_LOOP;\n WAIT_DONE:\n}" ::
          "r"((unsigned)__cvta_generic_to_shared(bar)), "r"(parity)
      : "memory");
}

// Ring entry index of cell (i2,j2); -1: not stored; -2: older than the ring (generic path).
template <bool BWD>
__device__ __forceinline__ int pred_index(const PairDev& P, const Ring& ring, const Recent& rc, int dcur, bool en, int i2, int j2, int xs,
                                          int ys) {
  if (!en) return -1;
  const int d2 = i2 + j2;
  const int dist = BWD ? d2 - dcur : dcur - d2;
  if (dist >= ring.R) return -2;
  int ilo, width;
  if (dist <= 2) {
    ilo = dist == 1 ? rc.ilo1 : rc.ilo2;
    width = dist == 1 ? rc.w1 : rc.w2;
  } else {
    const DiagInfo di = P.dinfo[d2];
    ilo = di.ilo; width = di.width;
  }
  const int r = i2 - ilo;
  int s;
  if ((unsigned)r < (unsigned)width) s = P.nS + r;
  else if (xs >= 0) s = xs;
  else if (ys >= 0) s = P.nS + width + ys;
  else return -1;
  return (d2 & (ring.R - 1)) * ring.cap + s;
}

// Register wavefront: lane (= slot) of cell (i2,j2) in its diagonal if that diagonal lies exactly `want` steps back;
// -1: not stored; -2: elsewhere (generic path through HBM).
template <bool BWD>
__device__ __forceinline__ int pred_lane(const PairDev& P, const Recent& rc, int dcur, bool en, int i2, int j2, int xs, int ys, int want) {
  if (!en) return -1;
  const int d2 = i2 + j2;
  const int dist = BWD ? d2 - dcur : dcur - d2;
  if (dist != want) return -2;
  const int ilo = want == 1 ? rc.ilo1 : rc.ilo2, width = want == 1 ? rc.w1 : rc.w2;
  const int r = i2 - ilo;
  if ((unsigned)r < (unsigned)width) return P.nS + r;
  if (xs >= 0) return xs;
  if (ys >= 0) return P.nS + width + ys;
  return -1;
}

__device__ __forceinline__ DiagInfo load_dinfo_early(const DiagInfo* p) {
  // volatile: keeps the load where it is written (ahead of its use in the NEXT iteration)
  DiagInfo d;
  asm volatile("ld.global.nc.v4.s32 {%0,%1,%2,%3}, [%4];" : "=r"(d.ilo), "=r"(d.width), "=r"(d.off), "=r"(d.pad) : "l"(p));
  return d;
}

// Everything about one cell that does not depend on earlier cells, as the 64-B record (r0..r3) the consumer
// works from: ring addresses of its predecessors, flags, transition exponents / mantissas, emission factors,
// absorb value.  `m`: the per-lane state records of diagonal d.
// LANES: the predecessor fields of r0 are lanes of the previous (X, Y moves) / second previous (XY move) diagonal
// (register wavefront) instead of ring addresses.
template <bool BWD, bool LANES = false>
__device__ __forceinline__ void make_record(const PairDev& P, const Ring& ring, int d, const Recent rc, const LaneMeta<BWD>& m, int4& r0,
                                            int4& r1, int4& r2, int4& r3) {
  // a near-start slot owns cells of near-end columns; then DPMatrix::inEnvelope (forward.h:92-98)
  const bool ok = m.ok && !(m.tail && m.xa.w >= 0) &&
                  (((m.xa.x | m.ya.x) & F_NEAR) || P.max_dist < 0 || hb_iabs(m.xa.z - m.ya.z) <= P.max_dist);
  const int i = m.i, j = m.j;
  const unsigned fx = m.xa.x & 0xff, fy = m.ya.x & 0xff;
  const bool xEmit = fx & F_EMIT, yEmit = fy & F_EMIT;
  const bool xReady = fx & F_READY, yReady = fy & F_READY;
  const bool xEmitOrStart = fx & F_EMIT_OR_START;
  const bool both = xEmit && yEmit;
  const float ab = ok && both ? m.ab : 0.f;
  int flags = ok ? R_OK : 0;
  if (!BWD) {
    const int xc = (int)((unsigned)m.xa.x >> 8), yc = (int)((unsigned)m.ya.x >> 8);
    const bool doX = ok && yReady && xc > 0;
    const bool doY = ok && (yEmit ? xReady : true) && yc > 0;
    const bool doD = ok && both && xc > 0 && yc > 0;
    const int aX = LANES ? pred_lane<false>(P, rc, d, doX, m.xb.x, j, m.xb.w, m.ya.w, 1) : pred_index<false>(P, ring, rc, d, doX, m.xb.x, j, m.xb.w, m.ya.w);
    const int aY = LANES ? pred_lane<false>(P, rc, d, doY, i, m.yb.x, m.xa.w, m.yb.w, 1) : pred_index<false>(P, ring, rc, d, doY, i, m.yb.x, m.xa.w, m.yb.w);
    const int aD = LANES ? pred_lane<false>(P, rc, d, doD, m.xb.x, m.yb.x, m.xb.w, m.yb.w, 2) : pred_index<false>(P, ring, rc, d, doD, m.xb.x, m.yb.x, m.xb.w, m.yb.w);
    if (xEmit) flags |= R_XEMIT;
    if (yEmit) flags |= R_YEMIT;
    if (!xEmit && !(!yEmit && xEmitOrStart)) flags |= R_IMM_BY_X;
    if (!yEmit && xEmitOrStart) flags |= R_IMM_BY_Y;
    if (i == 0 && j == 0) flags |= R_ORIGIN;
    if (both) flags |= R_BOTH;
    if (xc > 1 || yc > 1 || aX == -2 || aY == -2 || aD == -2) flags |= R_SLOW;
    r0 = make_int4(aX, aY, aD, flags);
    r1 = make_int4(m.xb.y, m.yb.y, m.xb.y + m.yb.y, i);
    r2 = make_int4(m.xb.z, m.yb.z, __float_as_int(__int_as_float(m.xb.z) * __int_as_float(m.yb.z)), __float_as_int(ab));
    const bool ex = xEmit && yReady, ey = yEmit && xReady;
    r3 = make_int4(__float_as_int(ex ? m.xe.x : 1.f), __float_as_int(ex ? m.xe.y : 1.f), __float_as_int(ey ? m.ye.x : 1.f),
                   __float_as_int(ey ? m.ye.y : 1.f));
  } else {
    const int xac = (int)((unsigned)m.xa.x >> 8), yac = (int)((unsigned)m.ya.x >> 8);
    const bool doD = ok && xac > 0 && yac > 0;
    const bool doX = ok && yReady && xac > 0;
    const bool doY = ok && xReady && yac > 0;
    const int aD = LANES ? pred_lane<true>(P, rc, d, doD, m.x0.x, m.y0.x, m.x0.z, m.y0.z, 2) : pred_index<true>(P, ring, rc, d, doD, m.x0.x, m.y0.x, m.x0.z, m.y0.z);
    const int aX = LANES ? pred_lane<true>(P, rc, d, doX, m.x0.x, j, m.x0.z, m.ya.w, 1) : pred_index<true>(P, ring, rc, d, doX, m.x0.x, j, m.x0.z, m.ya.w);
    const int aY = LANES ? pred_lane<true>(P, rc, d, doY, i, m.y0.x, m.xa.w, m.y0.z, 1) : pred_index<true>(P, ring, rc, d, doY, i, m.y0.x, m.xa.w, m.y0.z);
    const bool toEnd = (fx & F_TO_END) && (fy & F_TO_END);
    if (toEnd) flags |= R_TO_END;
    const bool xNull = yReady && m.xb.y > ((fx & F_TO_END) ? 1 : 0);
    const bool yNull = m.yb.y > ((fy & F_TO_END) ? 1 : 0);
    if (xac > 1 || yac > 1 || xNull || yNull || aX == -2 || aY == -2 || aD == -2) flags |= R_SLOW;
    r0 = make_int4(aD, aX, aY, flags);
    r1 = make_int4(m.x0.y + m.y0.y, m.x0.y, m.y0.y, i);
    r2 = make_int4(__float_as_int(__int_as_float(m.x1.x) * __int_as_float(m.y1.x)), m.x1.y, m.x1.z, __float_as_int(ab));
    r3 = make_int4(m.y1.y, m.y1.z, m.xb.z + m.yb.z, __float_as_int(__int_as_float(m.xb.w) * __int_as_float(m.yb.w)));
  }
}

template <bool BWD>
__device__ __forceinline__ void produce_one(const PairDev& P, const Ring& ring, int4* rec, uint64_t* full, uint64_t* empty, int t, int q,
                                            int d, const Recent rc, const LaneMeta<BWD>& m) {
  const int slot = q % HB_DEPTH;
  int4 r0, r1, r2, r3;
  make_record<BWD>(P, ring, d, rc, m, r0, r1, r2, r3);
  if (q >= HB_DEPTH) mbar_wait(empty + slot, ((q / HB_DEPTH) - 1) & 1);
  int4* R = rec + slot * 128;
  R[t] = r0; R[32 + t] = r1; R[64 + t] = r2; R[96 + t] = r3;
  mbar_arrive(full + slot);
}

template <bool BWD>
__device__ __forceinline__ void producer(const PairDev& P, const Ring& ring, int4* rec, uint64_t* full, uint64_t* empty, int p) {
  const int t = threadIdx.x & 31;
  const int nD = P.n_diag;
  auto diag_of = [&](int q) { return min(max(BWD ? nD - 1 - q : q, 0), nD - 1); };
  auto dq = [&](int q) { return BWD ? nD - 1 - q : q; };
  // Software pipeline: the layout records (dinfo) are requested two iterations ahead and the per-lane state
  // records / absorb value one iteration ahead of their use, so their L2/HBM latency overlaps the record
  // construction and the wait for a free queue slot.  The loop is unrolled by two over two named buffers
  // (mA, mB): rotating one buffer through a register copy would make the copy wait for the loads.
  DiagInfo c0 = P.dinfo[diag_of(p)], c1 = P.dinfo[diag_of(p - 1)], c2 = P.dinfo[diag_of(p - 2)];
  DiagInfo n0 = P.dinfo[diag_of(p + HB_NPROD)], n1 = P.dinfo[diag_of(p + HB_NPROD - 1)], n2 = P.dinfo[diag_of(p + HB_NPROD - 2)];
  LaneMeta<BWD> mA = load_meta<BWD>(P, ring, c0, dq(p), t, true), mB;
  for (int q = p; q < nD; q += 2 * HB_NPROD) {
    {
      const Recent rc = {c1.ilo, c1.width, c2.ilo, c2.width};
      mB = load_meta<BWD>(P, ring, n0, diag_of(q + HB_NPROD), t, true);   // clamped past the end; then unused
      c0 = n0; c1 = n1; c2 = n2;
      n0 = load_dinfo_early(P.dinfo + diag_of(q + 2 * HB_NPROD));
      n1 = load_dinfo_early(P.dinfo + diag_of(q + 2 * HB_NPROD - 1));
      n2 = load_dinfo_early(P.dinfo + diag_of(q + 2 * HB_NPROD - 2));
      produce_one<BWD>(P, ring, rec, full, empty, t, q, dq(q), rc, mA);
    }
    if (q + HB_NPROD >= nD) break;
    {
      const Recent rc = {c1.ilo, c1.width, c2.ilo, c2.width};
      mA = load_meta<BWD>(P, ring, n0, diag_of(q + 2 * HB_NPROD), t, true);
      c0 = n0; c1 = n1; c2 = n2;
      n0 = load_dinfo_early(P.dinfo + diag_of(q + 3 * HB_NPROD));
      n1 = load_dinfo_early(P.dinfo + diag_of(q + 3 * HB_NPROD - 1));
      n2 = load_dinfo_early(P.dinfo + diag_of(q + 3 * HB_NPROD - 2));
      produce_one<BWD>(P, ring, rec, full, empty, t, q + HB_NPROD, dq(q + HB_NPROD), rc, mB);
    }
  }
}

__device__ __forceinline__ bool ring_read(const Ring& ring, int idx, float4& a, float4& b) {
  if (idx < 0) return false;
  a = ring.A[idx];
  b = ring.B[idx];
  return __float_as_int(b.y) > HB_EXP_ZERO / 2;
}

// Forward cell from a producer record (same arithmetic as the first-edge part of forward_cell)
// The 5x6 transition weights held in registers for the whole sweep (the shared-memory copy in PairDev would be
// re-read every step: the barrier waits are compiler memory barriers).
struct WReg { float w[5][6]; };
__device__ __forceinline__ WReg load_weights(const PairDev& P) {
  WReg r;
#pragma unroll
  for (int a = 0; a < 5; ++a)
#pragma unroll
    for (int b = 0; b < 6; ++b) r.w[a][b] = P.W[a][b];
  return r;
}

// Forward cell from its record and the three predecessor cells (A = IMM,IMD,IDM,IMI; B = IIW, exponent bits, aux)
__device__ __forceinline__ Cell forward_math(const WReg& WR, const int flags, const int4 r1, const int4 r2, const int4 r3, const bool vX,
                                             const float4 xA, const float4 xB, const bool vY, const float4 yA, const float4 yB, const bool vD,
                                             const float4 dA, const float4 dB) {
  const float (*W)[6] = WR.w;
  const bool origin = flags & R_ORIGIN;
  const int eX = vX ? __float_as_int(xB.y) + r1.x : HB_EXP_ZERO;
  const int eY = vY ? __float_as_int(yB.y) + r1.y : HB_EXP_ZERO;
  const int eD = vD ? __float_as_int(dB.y) + r1.z : HB_EXP_ZERO;
  Cell a;
  a.E = max(max(eX, eY), max(eD, origin ? 0 : HB_EXP_ZERO));
  const float sX = vX ? pow2_neg(eX - a.E) * __int_as_float(r2.x) : 0.f;
  const float sY = vY ? pow2_neg(eY - a.E) * __int_as_float(r2.y) : 0.f;
  const float sD = vD ? pow2_neg(eD - a.E) * __int_as_float(r2.z) : 0.f;
  a.v[0] = origin ? pow2_neg(-a.E) : 0.f;
  a.v[1] = a.v[2] = a.v[3] = a.v[4] = 0.f;
  if (vX) {
    if (flags & R_XEMIT) {
      a.v[1] = sX * (xA.x * W[0][1] + xA.y * W[1][1] + xA.z * W[2][1] + xA.w * W[3][1]);
      a.v[4] = sX * (xA.x * W[0][4] + xA.w * W[3][4] + xB.x * W[4][4]);
    } else {
      a.v[1] = sX * xA.y;
      a.v[4] = sX * xB.x;
      if (flags & R_IMM_BY_X) a.v[0] = fmaf(sX, xA.x, a.v[0]);
    }
  }
  if (vY) {
    if (flags & R_YEMIT) {
      a.v[2] = sY * (yA.x * W[0][2] + yA.y * W[1][2] + yA.z * W[2][2] + yB.x * W[4][2]);
      a.v[3] = sY * (yA.x * W[0][3] + yA.w * W[3][3]);
    } else {
      a.v[2] = sY * yA.z;
      a.v[3] = sY * yA.w;
      if (flags & R_IMM_BY_Y) a.v[0] = fmaf(sY, yA.x, a.v[0]);
    }
  }
  if (vD) a.v[0] = fmaf(sD, dA.x * W[0][0] + dA.y * W[1][0] + dA.z * W[2][0] + dA.w * W[3][0] + dB.x * W[4][0], a.v[0]);
  if (flags & R_BOTH) a.v[0] *= __int_as_float(r2.w);
  a.v[1] *= __int_as_float(r3.x);
  a.v[4] *= __int_as_float(r3.y);
  a.v[2] *= __int_as_float(r3.z);
  a.v[3] *= __int_as_float(r3.w);
  cell_normalise(a);
  return a;
}

__device__ __forceinline__ Cell forward_fast(const WReg& WR, const Ring& ring, const int4 r0, const int4 r1, const int4 r2, const int4 r3) {
  float4 xA, xB, yA, yB, dA, dB;
  const bool vX = ring_read(ring, r0.x, xA, xB);
  const bool vY = ring_read(ring, r0.y, yA, yB);
  const bool vD = ring_read(ring, r0.z, dA, dB);
  return forward_math(WR, r0.w, r1, r2, r3, vX, xA, xB, vY, yA, yB, vD, dA, dB);
}

template <bool NORMALISE = true>
__device__ __forceinline__ Cell backward_math(const WReg& WR, const int flags, const int4 r1, const int4 r2, const int4 r3, const bool vD,
                                              const float4 dA, const float4 dB, const bool vX, const float4 xA, const float4 xB, const bool vY,
                                              const float4 yA, const float4 yB) {
  const float (*W)[6] = WR.w;
  const bool toEnd = flags & R_TO_END;
  const int eD = vD ? __float_as_int(dB.y) + r1.x : HB_EXP_ZERO;
  const int eX = vX ? __float_as_int(xB.y) + r1.y : HB_EXP_ZERO;
  const int eY = vY ? __float_as_int(yB.y) + r1.z : HB_EXP_ZERO;
  const int eE = toEnd ? r3.z : HB_EXP_ZERO;
  Cell a;
  a.E = max(max(eD, eX), max(eY, eE));
  const float dest = vD ? pow2_neg(eD - a.E) * __int_as_float(r2.x) * dB.z : 0.f;  // absorb(dx,dy) * B(dx,dy,IMM)
  const float sX = vX ? pow2_neg(eX - a.E) : 0.f;
  const float sY = vY ? pow2_neg(eY - a.E) : 0.f;
  const float sE = toEnd ? pow2_neg(eE - a.E) * __int_as_float(r3.w) : 0.f;
  const float dmd = vX ? sX * __int_as_float(r2.y) * xA.y : 0.f, diw = vX ? sX * __int_as_float(r2.z) * xB.x : 0.f;
  const float ddm = vY ? sY * __int_as_float(r3.x) * yA.z : 0.f, dmi = vY ? sY * __int_as_float(r3.y) * yA.w : 0.f;
  a.v[0] = W[0][0] * dest + W[0][1] * dmd + W[0][4] * diw + W[0][2] * ddm + W[0][3] * dmi + sE * W[0][5];
  a.v[1] = W[1][0] * dest + W[1][1] * dmd + W[1][2] * ddm + sE * W[1][5];
  a.v[2] = W[2][0] * dest + W[2][1] * dmd + W[2][2] * ddm + sE * W[2][5];
  a.v[3] = W[3][0] * dest + W[3][1] * dmd + W[3][4] * diw + W[3][3] * dmi + sE * W[3][5];
  a.v[4] = W[4][0] * dest + W[4][4] * diw + W[4][2] * ddm + sE * W[4][5];
  if (NORMALISE) cell_normalise(a);
  return a;
}

__device__ __forceinline__ Cell backward_fast(const WReg& WR, const Ring& ring, const int4 r0, const int4 r1, const int4 r2, const int4 r3) {
  float4 dA, dB, xA, xB, yA, yB;
  const bool vD = ring_read(ring, r0.x, dA, dB);
  const bool vX = ring_read(ring, r0.y, xA, xB);
  const bool vY = ring_read(ring, r0.z, yA, yB);
  return backward_math(WR, r0.w, r1, r2, r3, vD, dA, dB, vX, xA, xB, vY, yA, yB);
}

template <bool BWD>
__device__ __noinline__ Cell generic_cell(const PairDev& P, const Ring& ring, const Recent& rc, const DiagInfo& di, int d, int t, float ab) {
  const LaneMeta<BWD> m = load_meta<BWD>(P, ring, di, d, t, false);
  if (BWD) return backward_cell(P, ring, rc, d, m.i, m.j, m.xa, m.xb, m.x0, m.x1, m.ya, m.yb, m.y0, m.y1);
  return forward_cell(P, ring, rc, d, m.i, m.j, m.xa, m.xb, m.xe, m.ya, m.yb, m.ye, ab);
}

// One cell of the consumer: arithmetic from the record, store to HBM (streaming) and to the ring.
template <bool BWD>
__device__ __forceinline__ void consume_store(const PairDev& P, const WReg& WR, const Ring& ring, const Recent& rc, const DiagInfo& di, int d, int t,
                                              const int4 r0, const int4 r1, const int4 r2, const int4 r3, float4* gA, float2* gB) {
  Cell c = zero_cell();
  float aux = 0.f;
  if (r0.w & R_OK) {
    const float ab = __int_as_float(r2.w);
    if (r0.w & R_SLOW) c = generic_cell<BWD>(P, ring, rc, di, d, t, ab);
    else c = BWD ? backward_fast(WR, ring, r0, r1, r2, r3) : forward_fast(WR, ring, r0, r1, r2, r3);
    if (BWD) aux = c.v[0] * ab;  // absorb is 0 unless both states emit
  }
  const float4 va = make_float4(c.v[0], c.v[1], c.v[2], c.v[3]);
  __stcs(gA + di.off + t, va);
  __stcs(gB + di.off + t, make_float2(c.v[4], __int_as_float(c.E)));
  const int r = (d & (ring.R - 1)) * ring.cap + t;
  ring.A[r] = va;
  ring.B[r] = make_float4(c.v[4], __int_as_float(c.E), aux, 0.f);
}

template <bool BWD>
__device__ __forceinline__ void consumer(const PairDev& P, const Ring& ring, const int4* rec, uint64_t* full, uint64_t* empty) {
  float4* gA = BWD ? P.bA : P.fA;
  float2* gB = BWD ? P.bB : P.fB;
  const int nD = P.n_diag;
  const int t = threadIdx.x;
  Recent rc = {0, 0, 0, 0};
  const WReg WR = load_weights(P);
  DiagInfo di_next = P.dinfo[BWD ? nD - 1 : 0];
  for (int q = 0; q < nD; ++q) {
    const int slot = q % HB_DEPTH;
    const int d = BWD ? nD - 1 - q : q;
    const DiagInfo di = di_next;
    // layout record of the next diagonal, requested a whole step ahead of its use
    di_next = load_dinfo_early(P.dinfo + (BWD ? max(d - 1, 0) : min(d + 1, nD - 1)));
    mbar_wait(full + slot, (q / HB_DEPTH) & 1);
    const int4* R = rec + slot * 128;
    const int4 r0 = R[t], r1 = R[32 + t], r2 = R[64 + t], r3 = R[96 + t];
    mbar_arrive(empty + slot);
    if (t == 0) ring.info[d & (ring.R - 1)] = make_int4(di.ilo, di.width, (int)di.off, 0);
    if (t < P.nS + di.width + P.nE) consume_store<BWD>(P, WR, ring, rc, di, d, t, r0, r1, r2, r3, gA, gB);
    __syncwarp();
    rc.ilo2 = rc.ilo1; rc.w2 = rc.w1;
    rc.ilo1 = di.ilo; rc.w1 = di.width;
  }
}

template <bool BWD>
__device__ __forceinline__ void sweep_narrow(const PairDev& P, const Ring& ring, int4* rec, uint64_t* full, uint64_t* empty) {
  if (threadIdx.x < 32) consumer<BWD>(P, ring, rec, full, empty);
  else producer<BWD>(P, ring, rec, full, empty, (threadIdx.x >> 5) - 1);
}

// ------------------------------------------------------------------ narrow strips, streamed records
// The producer's work depends on no DP value, so it does not have to ride along the wavefront at all:
// build_records_kernel computes the 64-B record of EVERY stored cell of a batch in one massively parallel,
// bandwidth-bound pass (one warp per anti-diagonal), laid out per diagonal as four planes of cnt = nS +
// width + nE int4's (one contiguous block of 64 * cnt bytes at 64 * off[d]).  fill_stream_kernel is then
// the consumer alone -- ONE warp per strip -- whose lane 0 streams the diagonals' blocks into a
// shared-memory queue with bulk async copies (cp.async.bulk + mbarrier complete_tx) HB_SDEPTH diagonals
// ahead of their use.  The dependent instruction stream of a step is the arithmetic and nothing else, and a
// strip costs 32 threads instead of 64-96, so a whole level of sibling merges is resident at once.
#define HB_SDEPTH 8
#ifndef HB_BUILD_WARPS
#define HB_BUILD_WARPS 4   // measured: 4 warps x 8 diagonals 1.04 ms, 8 x 8 1.14 ms, 8 x 16 1.09 ms, 16 x 4 1.79 ms
#endif
#ifndef HB_BUILD_DIAGS
#define HB_BUILD_DIAGS 16  // diagonals per warp (4 x 16: 1.00 ms)
#endif

__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, unsigned bytes) {
  asm volatile("{\n .reg .b64 st;\n mbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n}" ::"r"((unsigned)__cvta_generic_to_shared(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, unsigned bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   (unsigned)__cvta_generic_to_shared(dst)),
               "l"(src), "r"(bytes), "r"((unsigned)__cvta_generic_to_shared(bar))
               : "memory");
}

// Copies the PairDev descriptor and the edge-slot lists of this CTA's pair into shared memory.
__device__ __forceinline__ void stage_pair(const PairDev* src, PairDev& P, int* edge_rows, int* edge_cols) {
  const int* s = reinterpret_cast<const int*>(src);
  int* dptr = reinterpret_cast<int*>(&P);
  for (int k = threadIdx.x; k < (int)(sizeof(PairDev) / sizeof(int)); k += blockDim.x) dptr[k] = s[k];
  __syncthreads();
  for (int k = threadIdx.x; k < P.nS && k < 32; k += blockDim.x) edge_rows[k] = P.start_rows[k];
  for (int k = threadIdx.x; k < P.nE && k < 32; k += blockDim.x) edge_cols[k] = P.end_cols[k];
  __syncthreads();
}

template <bool BWD>
__device__ __forceinline__ void build_records(const PairDev& P, const Ring& ring, int4* __restrict__ out, int q0, int q1) {
  const int t = threadIdx.x & 31;
  const int nD = P.n_diag;
  auto diag_of = [&](int q) { return min(max(BWD ? nD - 1 - q : q, 0), nD - 1); };
  // layout records: the sweep-order neighbours are carried over from the previous iteration, the next diagonal's
  // and the "request ahead" diagonal's are loaded one iteration early
  DiagInfo di_n = P.dinfo[diag_of(q0)], c1 = P.dinfo[diag_of(q0 - 1)], c2 = P.dinfo[diag_of(q0 - 2)];
  DiagInfo dj_n = P.dinfo[diag_of(q0 + HB_SDEPTH)];
  for (int q = q0; q < q1; ++q) {
    const int d = BWD ? nD - 1 - q : q;
    const DiagInfo di = di_n, dj = dj_n;
    di_n = P.dinfo[diag_of(q + 1)];
    dj_n = P.dinfo[diag_of(q + 1 + HB_SDEPTH)];
    const Recent rc = {c1.ilo, c1.width, c2.ilo, c2.width};
    const int cnt = P.nS + di.width + P.nE;
    int4* blk = out + (size_t)di.off * 4 + 2 * (size_t)d;   // block of diagonal d: 2 header int4's + 4 planes of cnt
    if (t == 0) {
      // header: this diagonal's layout, and the layout of the diagonal the consumer requests at the end of this step
      __stcs(blk, make_int4(di.ilo, di.width, (int)di.off, 0));
      __stcs(blk + 1, make_int4((int)dj.off, dj.width, diag_of(q + HB_SDEPTH), 0));
    }
    for (int tt = t; tt < max(cnt, 1); tt += 32) {   // one pass for strips of up to 32 slots, more for wider ones
      const LaneMeta<BWD> m = load_meta<BWD>(P, ring, di, d, tt, true);
      int4 r0, r1, r2, r3;
      make_record<BWD>(P, ring, d, rc, m, r0, r1, r2, r3);
      if (tt < cnt) {
        int4* o = blk + 2 + tt;
        __stcs(o, r0);
        __stcs(o + cnt, r1);
        __stcs(o + 2 * cnt, r2);
        __stcs(o + 3 * cnt, r3);
      }
    }
    c2 = c1; c1 = di;
  }
}

// grid = (diagonal chunks, pairs * n_dir); recs[b] = record buffer of (pair, direction) b
__global__ void __launch_bounds__(32 * HB_BUILD_WARPS) build_records_kernel(PairDev* const* __restrict__ pairs, int4* const* __restrict__ recs,
                                                                           int n_dir, int only_dir, int ring_cap, int ring_rows) {
  __shared__ PairDev P;
  __shared__ int edge_rows[32], edge_cols[32];
  stage_pair(pairs[blockIdx.y / n_dir], P, edge_rows, edge_cols);
  const bool bwd = n_dir == 2 ? (blockIdx.y % n_dir) == 1 : only_dir == 1;
  Ring ring;
  ring.cap = ring_cap; ring.R = ring_rows;
  ring.A = nullptr; ring.B = nullptr; ring.info = nullptr;
  ring.srows = edge_rows; ring.ecols = edge_cols;
  const int q0 = (blockIdx.x * HB_BUILD_WARPS + (threadIdx.x >> 5)) * HB_BUILD_DIAGS;
  const int q1 = min(q0 + HB_BUILD_DIAGS, P.n_diag);
  if (q0 >= q1) return;
  if (bwd) build_records<true>(P, ring, recs[blockIdx.y], q0, q1); else build_records<false>(P, ring, recs[blockIdx.y], q0, q1);
}

#define HB_SMAX 96      // widest anti-diagonal (in slots) the streamed kernels take; wider strips use fill_kernel
// int4's per queue slot: 2 header + 4 planes of up to `cap` records
__host__ __device__ inline int hb_sslot(int cap) { return 2 + 4 * cap; }

template <bool BWD>
__device__ __forceinline__ void consumer_stream(const PairDev& P, const Ring& ring, const int4* __restrict__ recs, int4* queue, uint64_t* full) {
  const int sslot = hb_sslot(ring.cap);
  float4* gA = BWD ? P.bA : P.fA;
  float2* gB = BWD ? P.bB : P.fB;
  const int nD = P.n_diag;
  const int t = threadIdx.x;
  const int nEdge = P.nS + P.nE;
  // lane 0: request the block of diagonal dd (stored cells start at off, main interval `width` wide) into queue slot q % HB_SDEPTH
  auto issue = [&](int q, int dd, unsigned off, int width) {
    const int slot = q % HB_SDEPTH;
    const unsigned bytes = 32u + 64u * (unsigned)(nEdge + width);
    mbar_expect_tx(full + slot, bytes);
    bulk_g2s(queue + slot * sslot, recs + (size_t)off * 4 + 2 * (size_t)dd, bytes, full + slot);
  };
  if (t == 0)
    for (int q = 0; q < HB_SDEPTH && q < nD; ++q) {
      const int dd = BWD ? nD - 1 - q : q;
      const DiagInfo dj = P.dinfo[dd];
      issue(q, dd, dj.off, dj.width);
    }
  Recent rc = {0, 0, 0, 0};
  const WReg WR = load_weights(P);
  for (int q = 0; q < nD; ++q) {
    const int slot = q % HB_SDEPTH;
    const int d = BWD ? nD - 1 - q : q;
    mbar_wait(full + slot, (q / HB_SDEPTH) & 1);
    const int4* Q = queue + slot * sslot;
    const int4 h0 = Q[0], h1 = Q[1];   // the diagonal's layout travels with its records: no global load in this loop
    DiagInfo di;
    di.ilo = h0.x; di.width = h0.y; di.off = (unsigned)h0.z; di.pad = 0;
    const int cnt = nEdge + di.width;
    const int4* R = Q + 2;
    if (t == 0) ring.info[d & (ring.R - 1)] = h0;
    for (int tt = t; tt < cnt; tt += 32)   // one cell per lane; a second pass only where a diagonal exceeds 32 slots
      consume_store<BWD>(P, WR, ring, rc, di, d, tt, R[tt], R[cnt + tt], R[2 * cnt + tt], R[3 * cnt + tt], gA, gB);
    __syncwarp();
    if (t == 0 && q + HB_SDEPTH < nD) {
      // the slot has been read by every lane; order those generic-proxy reads before the async-proxy write
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      issue(q + HB_SDEPTH, h1.z, (unsigned)h1.x, h1.y);
    }
    rc.ilo2 = rc.ilo1; rc.w2 = rc.w1;
    rc.ilo1 = di.ilo; rc.w1 = di.width;
  }
}


// ------------------------------------------------------------------ narrow strips, register wavefront
// One warp per strip and NOTHING between the lanes but shuffles: lane t owns slot t of every anti-diagonal and keeps
// the cells it computed on the last two diagonals in registers; a cell's three predecessors (forward.cpp:95-200) sit
// on exactly those two diagonals for chain-like profiles, so they are fetched with warp shuffles from the lanes that
// own them.  There is no shared-memory ring, no record stream through HBM and no queue: everything that does not
// depend on DP values (state records, layout records, absorb value) is requested one step ahead of its use, and the
// dependent part of a step is shuffle -> FMAs -> normalise.  Predecessors that lie further back (long transitions of
// DAG profiles), extra in-edges and null out-edges take the generic cell routine, which reads earlier cells from HBM
// (every cell is stored as soon as it is computed).  Strips of up to 32 slots per anti-diagonal.
struct RegCell {
  float4 a;     // IMM, IMD, IDM, IMI
  float iiw;
  int e;        // exponent
  float aux;    // Backward: IMM * absorb
};

__device__ __forceinline__ bool shfl_cell(const RegCell& c, int src, float4& A, float4& B) {
  const int l = src & 31;
  A.x = __shfl_sync(0xffffffffu, c.a.x, l);
  A.y = __shfl_sync(0xffffffffu, c.a.y, l);
  A.z = __shfl_sync(0xffffffffu, c.a.z, l);
  A.w = __shfl_sync(0xffffffffu, c.a.w, l);
  B.x = __shfl_sync(0xffffffffu, c.iiw, l);
  const int e = __shfl_sync(0xffffffffu, c.e, l);
  B.y = __int_as_float(e);
  B.z = __shfl_sync(0xffffffffu, c.aux, l);
  B.w = 0.f;
  return src >= 0 && e > HB_EXP_ZERO / 2;
}

template <bool BWD>
__device__ __forceinline__ void reg_step(const PairDev& P, const WReg& WR, const Ring& ring, const Recent& rc, const DiagInfo& di, int d, int t,
                                         const LaneMeta<BWD>& m, RegCell& p1, RegCell& p2, float4* gA, float2* gB) {
  int4 r0, r1, r2, r3;
  make_record<BWD, true>(P, ring, d, rc, m, r0, r1, r2, r3);
  float4 xA, xB, yA, yB, dA, dB;
  // Forward: r0 = (X, Y, XY, flags); Backward: r0 = (XY, X, Y, flags).  X and Y moves come from the previous diagonal,
  // the XY move from the one before.
  const bool vX = shfl_cell(p1, BWD ? r0.y : r0.x, xA, xB);
  const bool vY = shfl_cell(p1, BWD ? r0.z : r0.y, yA, yB);
  const bool vD = shfl_cell(p2, BWD ? r0.x : r0.z, dA, dB);
  Cell c = zero_cell();
  float aux = 0.f;
  if (r0.w & R_OK) {
    const float ab = __int_as_float(r2.w);
    if (r0.w & R_SLOW) c = generic_cell<BWD>(P, ring, rc, di, d, t, ab);
    else c = BWD ? backward_math(WR, r0.w, r1, r2, r3, vD, dA, dB, vX, xA, xB, vY, yA, yB)
                 : forward_math(WR, r0.w, r1, r2, r3, vX, xA, xB, vY, yA, yB, vD, dA, dB);
    if (BWD) aux = c.v[0] * ab;  // absorb is 0 unless both states emit
  }
  const float4 va = make_float4(c.v[0], c.v[1], c.v[2], c.v[3]);
  if (t < P.nS + di.width + P.nE) {
    __stcs(gA + di.off + t, va);
    __stcs(gB + di.off + t, make_float2(c.v[4], __int_as_float(c.E)));
  }
  p2 = p1;
  p1.a = va; p1.iiw = c.v[4]; p1.e = c.E; p1.aux = aux;
}

template <bool BWD>
__device__ __forceinline__ void consumer_reg(const PairDev& P, const Ring& ring) {
  float4* gA = BWD ? P.bA : P.fA;
  float2* gB = BWD ? P.bB : P.fB;
  const int nD = P.n_diag;
  const int t = threadIdx.x;
  auto diag_of = [&](int q) { return min(max(BWD ? nD - 1 - q : q, 0), nD - 1); };
  const WReg WR = load_weights(P);
  RegCell p1, p2;
  p1.a = make_float4(0.f, 0.f, 0.f, 0.f); p1.iiw = 0.f; p1.e = HB_EXP_ZERO; p1.aux = 0.f;
  p2 = p1;
  Recent rc = {0, 0, 0, 0};
  // layout records two steps ahead, per-lane state records one step ahead; two named buffers (mA, mB) alternate so
  // that no register copy has to wait for the loads
  DiagInfo c0 = P.dinfo[diag_of(0)], n0 = P.dinfo[diag_of(1)], n1 = P.dinfo[diag_of(2)];
  LaneMeta<BWD> mA = load_meta<BWD>(P, ring, c0, diag_of(0), t, true), mB;
  for (int q = 0; q < nD; q += 2) {
    {
      const DiagInfo di = c0;
      mB = load_meta<BWD>(P, ring, n0, diag_of(q + 1), t, true);
      c0 = n0; n0 = n1;
      n1 = load_dinfo_early(P.dinfo + diag_of(q + 3));
      reg_step<BWD>(P, WR, ring, rc, di, diag_of(q), t, mA, p1, p2, gA, gB);
      __syncwarp();
      rc.ilo2 = rc.ilo1; rc.w2 = rc.w1;
      rc.ilo1 = di.ilo; rc.w1 = di.width;
    }
    if (q + 1 >= nD) break;
    {
      const DiagInfo di = c0;
      mA = load_meta<BWD>(P, ring, n0, diag_of(q + 2), t, true);
      c0 = n0; n0 = n1;
      n1 = load_dinfo_early(P.dinfo + diag_of(q + 4));
      reg_step<BWD>(P, WR, ring, rc, di, diag_of(q + 1), t, mB, p1, p2, gA, gB);
      __syncwarp();
      rc.ilo2 = rc.ilo1; rc.w2 = rc.w1;
      rc.ilo1 = di.ilo; rc.w1 = di.width;
    }
  }
}


// ------------------------------------------------------------------ leaf x leaf strips: chain kernel
// Both profiles are leaf chains (state t -> t+1 with log-weight 0, states 1..n-2 emit): every state has in-degree and
// out-degree 1 and the only null states are START and END, so everything make_record derives from packed state
// records follows from (i, j) by arithmetic: predecessor cells are (i-1,j), (i,j-1), (i-1,j-1), flags are comparisons
// against 0 and n-2, transition mantissas are 1.  What is left to load per step is ONE 16-B record per profile
// ({ex, rootsub, ins, band key} of the lane's state) and, because the lane keeps the emission rows of its two states
// in registers and only one of them changes from one anti-diagonal to the next, ONE row of K floats: the absorb term
// exp(computeLogProbAbsorb(i,j)) (forward.h:112-124) is a K-term dot product of two register-resident rows, computed
// in the wavefront warp itself instead of a separate pass through HBM.  No absorb plane, no record stream, no
// shared-memory ring: HBM sees the cell stores and nothing else.
struct ChainPos {
  int i, j;
  bool ok;     // the slot owns an in-range cell (envelope test not included)
};

__device__ __forceinline__ ChainPos chain_pos(const DiagInfo& di, int d, int t, int nx, int ny) {
  // slot layout with one near-start row (x START) and one near-end column (y state ny-2): [row 0 | main interval | column ny-2]
  ChainPos p;
  const bool head = t == 0, mid = t >= 1 && t <= di.width;
  const int i = head ? 0 : mid ? di.ilo + t - 1 : d - (ny - 2);
  const int j = d - i;
  const bool in_main = (unsigned)(i - di.ilo) < (unsigned)di.width;
  // row 0 / column ny-2 cells inside the main interval belong to it; (0, ny-2) belongs to the near-start slot
  bool ok = t < di.width + 2 && (mid || !in_main) && (head || mid || i != 0);
  ok = ok && i >= 0 && j >= 0 && i <= nx - 2 && j <= ny - 2;
  p.i = ok ? i : 0; p.j = ok ? j : 0; p.ok = ok;
  return p;
}

// slot of cell (i2,j2) on a diagonal with layout (ilo, width); -1 if it is not stored there
__device__ __forceinline__ int chain_slot(int ilo, int width, int i2, int j2, int ny) {
  const int r = i2 - ilo;
  const int edge = i2 == 0 ? 0 : (j2 == ny - 2 ? 1 + width : -1);
  return (unsigned)r < (unsigned)width ? 1 + r : edge;
}

// Predicated in-place load of one emission row: lanes whose predicate is false keep their registers (a plain
// `if (pred) h = load` makes the compiler load into temporaries and copy, 4*K4 moves per row and step).
template <int K4>
__device__ __forceinline__ void load_row(float4 (&h)[K4], const float* __restrict__ H, int row, int Kp, bool pred) {
  const float4* g = reinterpret_cast<const float4*>(H + (size_t)row * Kp);
#pragma unroll
  for (int k = 0; k < K4; ++k)
    asm volatile("{\n .reg .pred p;\n setp.ne.s32 p, %5, 0;\n @p ld.global.nc.v4.f32 {%0,%1,%2,%3}, [%4];\n}"
                 : "+f"(h[k].x), "+f"(h[k].y), "+f"(h[k].z), "+f"(h[k].w)
                 : "l"(g + k), "r"((int)pred));
}

// cell_normalise without a branch and with one scale factor: 2^-k is representable for every k a normal maximum can
// have here (k = 127 would need a cell value >= 2^127; predecessors are normalised to [1,2) and weights are <= 1).
__device__ __forceinline__ void cell_normalise_nb(Cell& a, bool ok) {
  const float m = fmaxf(fmaxf(fmaxf(a.v[0], a.v[1]), fmaxf(a.v[2], a.v[3])), a.v[4]);
  const int k = ((__float_as_int(m) >> 23) & 0xff) - 127;  // floor(log2 m) for normal m; -127 for 0
  const float s = __int_as_float((127 - k) << 23);
  const bool live = ok && m > 0.f;
#pragma unroll
  for (int q = 0; q < 5; ++q) a.v[q] = live ? a.v[q] * s : 0.f;
  a.E = live ? a.E + k : HB_EXP_ZERO;
}

// Forward cell of a chain x chain strip, branch-free (forward_math with the flag tests resolved: an x-move implies
// that x emits, a y-move that y emits, and IMM is fed by the xy-move alone; forward.cpp:95-200).
__device__ __forceinline__ Cell chain_forward_math(const WReg& WR, bool origin, int ex, int ey, float rsx, float insx, float rsy, float insy,
                                                   float ab, bool vX, const float4 xA, const float4 xB, bool vY, const float4 yA,
                                                   const float4 yB, bool vD, const float4 dA, const float4 dB) {
  const float (*W)[6] = WR.w;
  const int eX = vX ? __float_as_int(xB.y) + ex : HB_EXP_ZERO;
  const int eY = vY ? __float_as_int(yB.y) + ey : HB_EXP_ZERO;
  const int eD = vD ? __float_as_int(dB.y) + ex + ey : HB_EXP_ZERO;
  Cell a;
  a.E = max(max(eX, eY), max(eD, origin ? 0 : HB_EXP_ZERO));
  const float sX = vX ? pow2_neg(eX - a.E) : 0.f;
  const float sY = vY ? pow2_neg(eY - a.E) : 0.f;
  const float sD = vD ? pow2_neg(eD - a.E) : 0.f;
  a.v[1] = sX * rsx * (xA.x * W[0][1] + xA.y * W[1][1] + xA.z * W[2][1] + xA.w * W[3][1]);
  a.v[4] = sX * insx * (xA.x * W[0][4] + xA.w * W[3][4] + xB.x * W[4][4]);
  a.v[2] = sY * rsy * (yA.x * W[0][2] + yA.y * W[1][2] + yA.z * W[2][2] + yB.x * W[4][2]);
  a.v[3] = sY * insy * (yA.x * W[0][3] + yA.w * W[3][3]);
  const float imm = sD * ab * (dA.x * W[0][0] + dA.y * W[1][0] + dA.z * W[2][0] + dA.w * W[3][0] + dB.x * W[4][0]);
  a.v[0] = origin ? 1.f : imm;   // lpStart() = 0 at (0,0) (forward.cpp:73): no predecessor, exponent 0
  return a;
}

template <bool BWD, int K4>
__device__ __forceinline__ void consumer_chain(const PairDev& P) {
  float4* gA = BWD ? P.bA : P.fA;
  float2* gB = BWD ? P.bB : P.fB;
  const int nD = P.n_diag, nx = P.x.n, ny = P.y.n, Kp = P.Kp;
  const int t = threadIdx.x;
  auto diag_of = [&](int q) { return min(max(BWD ? nD - 1 - q : q, 0), nD - 1); };
  const WReg WR = load_weights(P);
  const int4* __restrict__ recx = BWD ? P.x.cb : P.x.cf;
  const int4* __restrict__ recy = BWD ? P.y.cb : P.y.cf;
  const float* __restrict__ Hx = P.x.H;
  const float* __restrict__ Hy = P.y.H;
  const int max_dist = P.max_dist;
  RegCell p1, p2;
  p1.a = make_float4(0.f, 0.f, 0.f, 0.f); p1.iiw = 0.f; p1.e = HB_EXP_ZERO; p1.aux = 0.f;
  p2 = p1;
  int ilo1 = 0, w1 = 0, ilo2 = 0, w2 = 0;
  DiagInfo c0 = P.dinfo[diag_of(0)], n0 = P.dinfo[diag_of(1)], n1 = P.dinfo[diag_of(2)];
  ChainPos cur = chain_pos(c0, diag_of(0), t, nx, ny);
  float4 hx[K4], hy[K4];
#pragma unroll
  for (int k = 0; k < K4; ++k) hx[k] = hy[k] = make_float4(0.f, 0.f, 0.f, 0.f);
  int hi = -1, hj = -1;   // states whose emission rows hx / hy hold
  {
    const bool px = cur.ok && cur.i >= 1, py = cur.ok && cur.j >= 1;
    load_row<K4>(hx, Hx, cur.i, Kp, px);
    load_row<K4>(hy, Hy, cur.j, Kp, py);
    hi = px ? cur.i : hi;
    hj = py ? cur.j : hj;
  }
  int4 rx = __ldg(recx + cur.i), ry = __ldg(recy + cur.j);
  for (int q = 0; q < nD; ++q) {
    const DiagInfo di = c0;
    const int i = cur.i, j = cur.j;
    const bool xEmit = i >= 1, yEmit = j >= 1;
    // ---- the absorb term of this cell from the register-resident rows, then the rows are free for the next step
    float ab = 0.f;
    {
      float a0 = 0.f, a1 = 0.f;
#pragma unroll
      for (int k = 0; k < K4; ++k) {
        a0 = fmaf(hx[k].x, hy[k].x, a0);
        a1 = fmaf(hx[k].y, hy[k].y, a1);
        a0 = fmaf(hx[k].z, hy[k].z, a0);
        a1 = fmaf(hx[k].w, hy[k].w, a1);
      }
      // the rows are overwritten in place below: pin the dot product ahead of those loads (both are volatile asm)
      asm volatile("" : "+f"(a0), "+f"(a1));
      if (cur.ok && xEmit && yEmit && hi == i && hj == j) ab = a0 + a1;
    }
    // ---- far ahead: pull the layout records, state records and emission rows the sweep will reach in ~24 steps into L1
    // (they come from HBM the first time; the loads below are then L1 hits instead of waiting out a DRAM round trip)
    if ((q & 3) == 0) {
      const int far = BWD ? -1 : 1;
      asm volatile("prefetch.global.L1 [%0];" ::"l"(P.dinfo + diag_of(q + 32 + (t & 3) * 8)));
      const int fi = min(max(i + far * 14, 0), nx - 1), fj = min(max(j + far * 14, 0), ny - 1);
      asm volatile("prefetch.global.L1 [%0];" ::"l"(recx + fi));
      asm volatile("prefetch.global.L1 [%0];" ::"l"(recy + fj));
      asm volatile("prefetch.global.L1 [%0];" ::"l"(Hx + (size_t)fi * Kp));
      asm volatile("prefetch.global.L1 [%0];" ::"l"(Hy + (size_t)fj * Kp));
    }
    // ---- requests for the next step (nothing here depends on DP values): at most one row changes for a main-interval lane
    const ChainPos nxt = chain_pos(n0, diag_of(q + 1), t, nx, ny);
    const int4 rxn = __ldg(recx + nxt.i), ryn = __ldg(recy + nxt.j);
    {
      const bool px = nxt.ok && nxt.i >= 1 && nxt.i != hi, py = nxt.ok && nxt.j >= 1 && nxt.j != hj;
      load_row<K4>(hx, Hx, nxt.i, Kp, px);
      load_row<K4>(hy, Hy, nxt.j, Kp, py);
      hi = px ? nxt.i : hi;
      hj = py ? nxt.j : hj;
    }
    c0 = n0; n0 = n1;
    n1 = load_dinfo_early(P.dinfo + diag_of(q + 3));
    // ---- this step
    const bool xLast = i == nx - 2, yLast = j == ny - 2;      // the states with the null transition into END: not ready
    const bool ok = cur.ok && (i == 0 || yLast || max_dist < 0 || hb_iabs(rx.w - ry.w) <= max_dist);   // DPMatrix::inEnvelope (forward.h:92-98)
    int sX, sY, sD;
    if (!BWD) {
      const bool doX = ok && !yLast && xEmit;             // x-move from (i-1, j)    (forward.cpp:95-129)
      const bool doY = ok && !xLast && yEmit;             // y-move from (i, j-1)    (forward.cpp:131-163)
      const bool doD = ok && xEmit && yEmit;              // xy-move from (i-1, j-1) (forward.cpp:165-184)
      sX = doX ? chain_slot(ilo1, w1, i - 1, j, ny) : -1;
      sY = doY ? chain_slot(ilo1, w1, i, j - 1, ny) : -1;
      sD = doD ? chain_slot(ilo2, w2, i - 1, j - 1, ny) : -1;
    } else {
      const bool xac = i + 1 <= nx - 2, yac = j + 1 <= ny - 2;   // the absorbing out-edge i -> i+1 exists iff i+1 emits
      sD = ok && xac && yac ? chain_slot(ilo2, w2, i + 1, j + 1, ny) : -1;     // forward.cpp:1017-1030
      sX = ok && !yLast && xac ? chain_slot(ilo1, w1, i + 1, j, ny) : -1;      // forward.cpp:1032-1048
      sY = ok && !xLast && yac ? chain_slot(ilo1, w1, i, j + 1, ny) : -1;      // forward.cpp:1050-1065
    }
    float4 xA, xB, yA, yB, dA, dB;
    const bool vX = shfl_cell(p1, sX, xA, xB);
    const bool vY = shfl_cell(p1, sY, yA, yB);
    const bool vD = shfl_cell(p2, sD, dA, dB);
    Cell c;
    if (!BWD) {
      c = chain_forward_math(WR, i == 0 && j == 0, rx.x, ry.x, __int_as_float(rx.y), __int_as_float(rx.z), __int_as_float(ry.y),
                             __int_as_float(ry.z), ab, vX, xA, xB, vY, yA, yB, vD, dA, dB);
    } else {
      const int4 r1 = make_int4(rx.x + ry.x, rx.x, ry.x, i);
      const int4 r2 = make_int4(__float_as_int(1.f), rx.y, rx.z, 0);
      const int4 r3 = make_int4(ry.y, ry.z, 0, __float_as_int(1.f));
      c = backward_math<false>(WR, xLast && yLast ? R_TO_END : 0, r1, r2, r3, vD, dA, dB, vX, xA, xB, vY, yA, yB);
    }
    cell_normalise_nb(c, ok);
    const float4 va = make_float4(c.v[0], c.v[1], c.v[2], c.v[3]);
    // (di.pad is 0: using it keeps the fourth register of the early layout load alive until here; dead, the compiler
    // reuses it as a scratch register and every write to it waits for the load issued three steps ahead)
    if (t < di.width + 2 + di.pad) {
      __stcs(gA + di.off + t, va);
      __stcs(gB + di.off + t, make_float2(c.v[4], __int_as_float(c.E)));
    }
    p2 = p1;
    p1.a = va; p1.iiw = c.v[4]; p1.e = c.E; p1.aux = BWD ? c.v[0] * ab : 0.f;
    ilo2 = ilo1; w2 = w1;
    ilo1 = di.ilo; w1 = di.width;
    cur = nxt;
    rx = rxn; ry = ryn;
  }
}

template <int K4>
__global__ void __launch_bounds__(32) fill_chain_kernel(PairDev* const* __restrict__ pairs, int n_dir, int only_dir) {
  __shared__ PairDev P;
  {
    const int* s = reinterpret_cast<const int*>(pairs[blockIdx.x / n_dir]);
    int* dptr = reinterpret_cast<int*>(&P);
    for (int k = threadIdx.x; k < (int)(sizeof(PairDev) / sizeof(int)); k += blockDim.x) dptr[k] = s[k];
  }
  __syncthreads();
  const bool bwd = n_dir == 2 ? (blockIdx.x % n_dir) == 1 : only_dir == 1;
  if (bwd) consumer_chain<true, K4>(P); else consumer_chain<false, K4>(P);
  __threadfence_block();
  __syncwarp();
  if (threadIdx.x == 0) {
    if (!bwd) {
      // transitions into EEE (forward.cpp:205-220): the one cell (nx-2, ny-2)
      const int i = P.x.n - 2, j = P.y.n - 2, d = i + j;
      const DiagInfo di = P.dinfo[d];
      const int s = chain_slot(di.ilo, di.width, i, j, P.y.n);
      double lp = -INFINITY;
      if (s >= 0) {
        const float4 a = P.fA[di.off + s];
        const float2 b = P.fB[di.off + s];
        const int E = __float_as_int(b.y);
        const double v = (double)a.x * P.W[0][5] + (double)a.y * P.W[1][5] + (double)a.z * P.W[2][5] + (double)a.w * P.W[3][5] + (double)b.x * P.W[4][5];
        if (E > HB_EXP_ZERO / 2 && v > 0) lp = log(v) + (double)E * HB_LN2;
      }
      *P.lp_end = lp;
    } else {
      const DiagInfo di = P.dinfo[0];
      const int s = chain_slot(di.ilo, di.width, 0, 0, P.y.n);
      double lp = -INFINITY;
      if (s >= 0) {
        const float4 a = P.bA[di.off + s];
        const float2 b = P.bB[di.off + s];
        if (__float_as_int(b.y) > HB_EXP_ZERO / 2) lp = cell_log(a.x, __float_as_int(b.y));
      }
      *P.lp_start = lp;  // BackwardMatrix::lpStart() = cell(0,0,IMM)
    }
  }
}


// ------------------------------------------------------------------ leaf x leaf strips, two warps per strip
// The step of consumer_chain has two halves: everything that does NOT depend on DP values (position of the lane's cell,
// its state records, the absorb dot product from the register-resident emission rows, the envelope test, the lanes its
// three predecessors sit on) and the part that does (shuffles, sums, normalisation, stores).  Here a HELPER warp runs the
// first half a few steps ahead and hands every lane a 32-byte record through shared memory; the MAIN warp carries the
// dependency chain with half the instructions.  The hand-over is lane-private (lane t of the main warp reads what lane t
// of the helper wrote), so it needs no barrier and no fence (a release would make the helper wait for its own prefetch
// loads at every step): every 16-byte chunk of a record carries the step number it belongs to -- a 16-byte shared
// store of one lane is a single transaction -- and the main lane re-reads until all three chunks show its step; a
// per-lane `consumed` counter, stored with a data dependency on the chunks just read, keeps the helper from lapping it.
#define CH_DEPTH 8
#define CH_SPIN_LIMIT (1 << 22)   // a hand-over that never arrives is a bug: give up (results invalid, flagged) instead of hanging the GPU

struct ChainShared {
  int4 rec[CH_DEPTH][3][32];   // per step, per lane: {ab, rs_x, ins_x, step} {rs_y, ins_y, ex | ey << 16, step} {store offset, control, step, 0}
  int consumed[32];
  int failed;
};

__device__ __forceinline__ int4 ld_volatile_shared4(const int4* p) {
  int4 v;
  asm volatile("ld.volatile.shared.v4.s32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"((unsigned)__cvta_generic_to_shared(p)) : "memory");
  return v;
}
__device__ __forceinline__ void st_volatile_shared4(int4* p, int4 v) {
  asm volatile("st.volatile.shared.v4.s32 [%0], {%1,%2,%3,%4};" ::"r"((unsigned)__cvta_generic_to_shared(p)), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ int ld_volatile_shared(const int* p) {
  int v;
  asm volatile("ld.volatile.shared.s32 %0, [%1];" : "=r"(v) : "r"((unsigned)__cvta_generic_to_shared(p)) : "memory");
  return v;
}
__device__ __forceinline__ void st_volatile_shared(int* p, int v) {
  asm volatile("st.volatile.shared.s32 [%0], %1;" ::"r"((unsigned)__cvta_generic_to_shared(p)), "r"(v) : "memory");
}

// control word: predecessor lanes (0xff = none) in bytes 0-2, flags in byte 3
enum { CH_OK = 1, CH_ORIGIN = 2, CH_TO_END = 4, CH_STORE = 8 };

template <bool BWD, int K4>
__device__ __forceinline__ void chain_helper(const PairDev& P, ChainShared& S) {
  const int nD = P.n_diag, nx = P.x.n, ny = P.y.n, Kp = P.Kp;
  const int t = threadIdx.x & 31;
  auto diag_of = [&](int q) { return min(max(BWD ? nD - 1 - q : q, 0), nD - 1); };
  const int4* __restrict__ recx = BWD ? P.x.cb : P.x.cf;
  const int4* __restrict__ recy = BWD ? P.y.cb : P.y.cf;
  const float* __restrict__ Hx = P.x.H;
  const float* __restrict__ Hy = P.y.H;
  const int max_dist = P.max_dist;
  int ilo1 = 0, w1 = 0, ilo2 = 0, w2 = 0;
  DiagInfo c0 = P.dinfo[diag_of(0)], n0 = P.dinfo[diag_of(1)], n1 = P.dinfo[diag_of(2)];
  ChainPos cur = chain_pos(c0, diag_of(0), t, nx, ny);
  float4 hx[K4], hy[K4];
#pragma unroll
  for (int k = 0; k < K4; ++k) hx[k] = hy[k] = make_float4(0.f, 0.f, 0.f, 0.f);
  int hi = -1, hj = -1;
  {
    const bool px = cur.ok && cur.i >= 1, py = cur.ok && cur.j >= 1;
    load_row<K4>(hx, Hx, cur.i, Kp, px);
    load_row<K4>(hy, Hy, cur.j, Kp, py);
    hi = px ? cur.i : hi;
    hj = py ? cur.j : hj;
  }
  int4 rx = __ldg(recx + cur.i), ry = __ldg(recy + cur.j);
  for (int q = 0; q < nD; ++q) {
    const DiagInfo di = c0;
    const int i = cur.i, j = cur.j;
    const bool xEmit = i >= 1, yEmit = j >= 1;
    float ab = 0.f;
    {
      float a0 = 0.f, a1 = 0.f;
#pragma unroll
      for (int k = 0; k < K4; ++k) {
        a0 = fmaf(hx[k].x, hy[k].x, a0);
        a1 = fmaf(hx[k].y, hy[k].y, a1);
        a0 = fmaf(hx[k].z, hy[k].z, a0);
        a1 = fmaf(hx[k].w, hy[k].w, a1);
      }
      asm volatile("" : "+f"(a0), "+f"(a1));   // the rows are overwritten in place below
      if (cur.ok && xEmit && yEmit && hi == i && hj == j) ab = a0 + a1;
    }
    if ((q & 3) == 0) {
      const int far = BWD ? -1 : 1;
      asm volatile("prefetch.global.L1 [%0];" ::"l"(P.dinfo + diag_of(q + 32 + (t & 3) * 8)));
      const int fi = min(max(i + far * 14, 0), nx - 1), fj = min(max(j + far * 14, 0), ny - 1);
      asm volatile("prefetch.global.L1 [%0];" ::"l"(recx + fi));
      asm volatile("prefetch.global.L1 [%0];" ::"l"(recy + fj));
      asm volatile("prefetch.global.L1 [%0];" ::"l"(Hx + (size_t)fi * Kp));
      asm volatile("prefetch.global.L1 [%0];" ::"l"(Hy + (size_t)fj * Kp));
    }
    const ChainPos nxt = chain_pos(n0, diag_of(q + 1), t, nx, ny);
    const int4 rxn = __ldg(recx + nxt.i), ryn = __ldg(recy + nxt.j);
    {
      const bool px = nxt.ok && nxt.i >= 1 && nxt.i != hi, py = nxt.ok && nxt.j >= 1 && nxt.j != hj;
      load_row<K4>(hx, Hx, nxt.i, Kp, px);
      load_row<K4>(hy, Hy, nxt.j, Kp, py);
      hi = px ? nxt.i : hi;
      hj = py ? nxt.j : hj;
    }
    c0 = n0; n0 = n1;
    n1 = load_dinfo_early(P.dinfo + diag_of(q + 3));
    const bool xLast = i == nx - 2, yLast = j == ny - 2;
    const bool ok = cur.ok && (i == 0 || yLast || max_dist < 0 || hb_iabs(rx.w - ry.w) <= max_dist);
    int sX, sY, sD;
    if (!BWD) {
      sX = ok && !yLast && xEmit ? chain_slot(ilo1, w1, i - 1, j, ny) : -1;
      sY = ok && !xLast && yEmit ? chain_slot(ilo1, w1, i, j - 1, ny) : -1;
      sD = ok && xEmit && yEmit ? chain_slot(ilo2, w2, i - 1, j - 1, ny) : -1;
    } else {
      const bool xac = i + 1 <= nx - 2, yac = j + 1 <= ny - 2;
      sD = ok && xac && yac ? chain_slot(ilo2, w2, i + 1, j + 1, ny) : -1;
      sX = ok && !yLast && xac ? chain_slot(ilo1, w1, i + 1, j, ny) : -1;
      sY = ok && !xLast && yac ? chain_slot(ilo1, w1, i, j + 1, ny) : -1;
    }
    const bool store = t < di.width + 2 + di.pad;
    int flags = (ok ? CH_OK : 0) | (store ? CH_STORE : 0);
    if (!BWD && i == 0 && j == 0) flags |= CH_ORIGIN;
    if (BWD && xLast && yLast) flags |= CH_TO_END;
    const int ctrl = (sX & 0xff) | ((sY & 0xff) << 8) | ((sD & 0xff) << 16) | (flags << 24);
    // ---- hand the record over: wait until the main warp has freed this ring slot
    const int slot = q & (CH_DEPTH - 1);
    if (q >= CH_DEPTH) {
      int spins = 0;
      while (ld_volatile_shared(&S.consumed[t]) < q - CH_DEPTH + 1)
        if (++spins > CH_SPIN_LIMIT) { S.failed = 1; break; }
    }
    st_volatile_shared4(&S.rec[slot][0][t], make_int4(__float_as_int(ab), rx.y, rx.z, q + 1));
    st_volatile_shared4(&S.rec[slot][1][t], make_int4(ry.y, ry.z, (rx.x & 0xffff) | (ry.x << 16), q + 1));
    st_volatile_shared4(&S.rec[slot][2][t], make_int4((int)(di.off + t), ctrl, q + 1, 0));
    ilo2 = ilo1; w2 = w1;
    ilo1 = di.ilo; w1 = di.width;
    cur = nxt;
    rx = rxn; ry = ryn;
  }
}

template <bool BWD>
__device__ __forceinline__ void chain_main(const PairDev& P, ChainShared& S) {
  float4* gA = BWD ? P.bA : P.fA;
  float2* gB = BWD ? P.bB : P.fB;
  const int nD = P.n_diag;
  const int t = threadIdx.x & 31;
  const WReg WR = load_weights(P);
  RegCell p1, p2;
  p1.a = make_float4(0.f, 0.f, 0.f, 0.f); p1.iiw = 0.f; p1.e = HB_EXP_ZERO; p1.aux = 0.f;
  p2 = p1;
  for (int q = 0; q < nD; ++q) {
    const int slot = q & (CH_DEPTH - 1);
    int4 ra, rb, rc;
    {
      int spins = 0;
      for (;;) {
        ra = ld_volatile_shared4(&S.rec[slot][0][t]);
        rb = ld_volatile_shared4(&S.rec[slot][1][t]);
        rc = ld_volatile_shared4(&S.rec[slot][2][t]);
        if (ra.w == q + 1 && rb.w == q + 1 && rc.z == q + 1) break;
        if (++spins > CH_SPIN_LIMIT) { S.failed = 1; break; }
      }
    }
    st_volatile_shared(&S.consumed[t], min(min(ra.w, rb.w), rc.z));   // = q + 1, and cannot be issued before the chunks have arrived
    const float ab = __int_as_float(ra.x);
    const int ctrl = rc.y, flags = (unsigned)ctrl >> 24;
    const int ex = (int)(short)(rb.z & 0xffff), ey = rb.z >> 16;
    const int bX = ctrl & 0xff, bY = (ctrl >> 8) & 0xff, bD = (ctrl >> 16) & 0xff;
    const int sX = bX == 0xff ? -1 : bX, sY = bY == 0xff ? -1 : bY, sD = bD == 0xff ? -1 : bD;
    const bool ok = flags & CH_OK;
    float4 xA, xB, yA, yB, dA, dB;
    const bool vX = shfl_cell(p1, sX, xA, xB);
    const bool vY = shfl_cell(p1, sY, yA, yB);
    const bool vD = shfl_cell(p2, sD, dA, dB);
    Cell c;
    if (!BWD) {
      c = chain_forward_math(WR, (flags & CH_ORIGIN) != 0, ex, ey, __int_as_float(ra.y), __int_as_float(ra.z), __int_as_float(rb.x),
                             __int_as_float(rb.y), ab, vX, xA, xB, vY, yA, yB, vD, dA, dB);
    } else {
      const int4 r1 = make_int4(ex + ey, ex, ey, 0);
      const int4 r2 = make_int4(__float_as_int(1.f), ra.y, ra.z, 0);
      const int4 r3 = make_int4(rb.x, rb.y, 0, __float_as_int(1.f));
      c = backward_math<false>(WR, (flags & CH_TO_END) ? R_TO_END : 0, r1, r2, r3, vD, dA, dB, vX, xA, xB, vY, yA, yB);
    }
    cell_normalise_nb(c, ok);
    const float4 va = make_float4(c.v[0], c.v[1], c.v[2], c.v[3]);
    if (flags & CH_STORE) {
      __stcs(gA + (unsigned)rc.x, va);
      __stcs(gB + (unsigned)rc.x, make_float2(c.v[4], __int_as_float(c.E)));
    }
    p2 = p1;
    p1.a = va; p1.iiw = c.v[4]; p1.e = c.E; p1.aux = BWD ? c.v[0] * ab : 0.f;
  }
}

template <int K4>
__global__ void __launch_bounds__(64) fill_chain2_kernel(PairDev* const* __restrict__ pairs, int n_dir, int only_dir, int* __restrict__ err) {
  __shared__ PairDev P;
  __shared__ ChainShared S;
  {
    const int* s = reinterpret_cast<const int*>(pairs[blockIdx.x / n_dir]);
    int* dptr = reinterpret_cast<int*>(&P);
    for (int k = threadIdx.x; k < (int)(sizeof(PairDev) / sizeof(int)); k += blockDim.x) dptr[k] = s[k];
  }
  if (threadIdx.x < 32) S.consumed[threadIdx.x] = 0;
  for (int k = threadIdx.x; k < CH_DEPTH * 3 * 32; k += blockDim.x) (&S.rec[0][0][0])[k] = make_int4(0, 0, 0, 0);   // step numbers start at 1
  if (threadIdx.x == 0) S.failed = 0;
  __syncthreads();
  const bool bwd = n_dir == 2 ? (blockIdx.x % n_dir) == 1 : only_dir == 1;
  if (threadIdx.x >= 32) {
    if (bwd) chain_helper<true, K4>(P, S); else chain_helper<false, K4>(P, S);
  } else {
    if (bwd) chain_main<true>(P, S); else chain_main<false>(P, S);
  }
  __threadfence_block();
  __syncthreads();
  if (threadIdx.x == 0) {
    if (S.failed && err) atomicExch(err, 1);
    if (!bwd) {
      // transitions into EEE (forward.cpp:205-220): the one cell (nx-2, ny-2)
      const int i = P.x.n - 2, j = P.y.n - 2, d = i + j;
      const DiagInfo di = P.dinfo[d];
      const int s = chain_slot(di.ilo, di.width, i, j, P.y.n);
      double lp = -INFINITY;
      if (s >= 0) {
        const float4 a = P.fA[di.off + s];
        const float2 b = P.fB[di.off + s];
        const int E = __float_as_int(b.y);
        const double v = (double)a.x * P.W[0][5] + (double)a.y * P.W[1][5] + (double)a.z * P.W[2][5] + (double)a.w * P.W[3][5] + (double)b.x * P.W[4][5];
        if (E > HB_EXP_ZERO / 2 && v > 0) lp = log(v) + (double)E * HB_LN2;
      }
      *P.lp_end = S.failed ? NAN : lp;
    } else {
      const DiagInfo di = P.dinfo[0];
      const int s = chain_slot(di.ilo, di.width, 0, 0, P.y.n);
      double lp = -INFINITY;
      if (s >= 0) {
        const float4 a = P.bA[di.off + s];
        const float2 b = P.bB[di.off + s];
        if (__float_as_int(b.y) > HB_EXP_ZERO / 2) lp = cell_log(a.x, __float_as_int(b.y));
      }
      *P.lp_start = S.failed ? NAN : lp;
    }
  }
}

// General strips: a CTA of several warps, each thread looping over the slots of the diagonal.
template <bool BWD>
__device__ __forceinline__ void sweep_wide(const PairDev& P, const Ring& ring) {
  float4* gA = BWD ? P.bA : P.fA;
  float2* gB = BWD ? P.bB : P.fB;
  Recent rc = {0, 0, 0, 0};
  for (int step = 0; step < P.n_diag; ++step) {
    const int d = BWD ? P.n_diag - 1 - step : step;
    const DiagInfo di = P.dinfo[d];
    if (ring.R > 0 && threadIdx.x == 0) ring.info[d & (ring.R - 1)] = make_int4(di.ilo, di.width, (int)di.off, 0);
    const int nthr = P.nS + di.width + P.nE;
    for (int t = threadIdx.x; t < nthr; t += blockDim.x) {
      const LaneMeta<BWD> m = load_meta<BWD>(P, ring, di, d, t);
      compute_store<BWD>(P, ring, rc, di, d, t, m, gA, gB);
    }
    __syncthreads();
    rc.ilo2 = rc.ilo1; rc.w2 = rc.w1;
    rc.ilo1 = di.ilo; rc.w1 = di.width;
  }
}

// grid.x = n_pairs * n_dir; block b handles pair b / n_dir, direction (b % n_dir) (0 = Forward, 1 = Backward;
// with n_dir == 1 the direction is `only_dir`).
// MODE 0: general strips (CTA-wide sweep); 1: narrow strips, producer/consumer warps; 2: narrow strips, streamed records;
// 3: narrow strips, register wavefront
template <int MODE>
__device__ __forceinline__ void fill_body(PairDev* const* __restrict__ pairs, int n_dir, int only_dir, int ring_cap, int ring_rows,
                                          int4* const* __restrict__ recs = nullptr) {
  constexpr bool NARROW = MODE != 0;
  extern __shared__ __align__(128) unsigned char smem_raw[];
  __shared__ PairDev P;
  __shared__ int4 ring_info[HB_RING_MAX];
  __shared__ uint64_t mbar[2 * HB_DEPTH > HB_SDEPTH ? 2 * HB_DEPTH : HB_SDEPTH];
  if (MODE == 1 && threadIdx.x < 2 * HB_DEPTH) mbar_init(mbar + threadIdx.x, 32);
  if (MODE == 2 && threadIdx.x < HB_SDEPTH) mbar_init(mbar + threadIdx.x, 1);
  if (MODE == 2) asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
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
  ring.R = ring_rows;
  ring.A = reinterpret_cast<float4*>(smem_raw);
  ring.B = ring.A + (size_t)ring_rows * ring_cap;
  ring.info = ring_info;
  ring.srows = P.start_rows;
  ring.ecols = P.end_cols;
  __shared__ int edge_rows[32], edge_cols[32];
  if (NARROW) {
    if (threadIdx.x < P.nS) edge_rows[threadIdx.x] = P.start_rows[threadIdx.x];
    if (threadIdx.x < P.nE) edge_cols[threadIdx.x] = P.end_cols[threadIdx.x];
    ring.srows = edge_rows;
    ring.ecols = edge_cols;
    __syncthreads();
  }
  if (MODE == 3) {
    ring.R = 0;   // no shared-memory ring: earlier cells are in registers or in HBM
    if (bwd) consumer_reg<true>(P, ring); else consumer_reg<false>(P, ring);
  } else if (MODE == 2) {
    int4* queue = reinterpret_cast<int4*>(ring.B + (size_t)ring_rows * ring_cap);
    if (bwd) consumer_stream<true>(P, ring, recs[blockIdx.x], queue, mbar); else consumer_stream<false>(P, ring, recs[blockIdx.x], queue, mbar);
  } else if (MODE == 1) {
    int4* rec = reinterpret_cast<int4*>(ring.B + (size_t)ring_rows * ring_cap);
    if (bwd) sweep_narrow<true>(P, ring, rec, mbar, mbar + HB_DEPTH); else sweep_narrow<false>(P, ring, rec, mbar, mbar + HB_DEPTH);
  } else {
    if (bwd) sweep_wide<true>(P, ring); else sweep_wide<false>(P, ring);
  }

  if (threadIdx.x == 0) {
    Ring none = ring;
    none.R = 0;
    const Recent rc = {0, 0, 0, 0};
    float4 sa, sb;
    if (!bwd) {
      // transitions into EEE (forward.cpp:205-220)
      double lp = -INFINITY;
      const int xe = P.x.n - 1, ye = P.y.n - 1;
      for (int qx = P.x.in_off[xe]; qx < P.x.in_off[xe + 1]; ++qx) {
        const int tx = P.x.in_idx[qx];
        const int sx = P.x.t_src[tx];
        for (int qy = P.y.in_off[ye]; qy < P.y.in_off[ye + 1]; ++qy) {
          const int ty = P.y.in_idx[qy];
          const int sy = P.y.t_src[ty];
          if (!fetch_cell<false>(P, none, rc, 1 << 30, true, sx, sy, P.x.slot[sx], P.y.slot[sy], sa, sb)) continue;
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
      if (fetch_cell<true>(P, none, rc, -(1 << 30), true, 0, 0, P.x.slot[0], P.y.slot[0], sa, sb)) lp = cell_log(sa.x, __float_as_int(sb.y));
      *P.lp_start = lp;  // BackwardMatrix::lpStart() = cell(0,0,IMM)
    }
  }
}

// one consumer + HB_NPROD producer warps per strip; requires blockDim.x == 32 * (1 + HB_NPROD),
// ring_rows >= 4 and max_threads <= 32 for every pair; dynamic smem = ring + HB_DEPTH * 2 KB of records
__global__ void __maxnreg__(HB_NARROW_REGS) fill_narrow_kernel(PairDev* const* __restrict__ pairs, int n_dir, int only_dir, int ring_cap, int ring_rows) {
  fill_body<1>(pairs, n_dir, only_dir, ring_cap, ring_rows);
}
// one warp per strip; dynamic smem = ring + HB_SDEPTH * 2 KB of record blocks; recs[blockIdx.x] built by build_records_kernel
__global__ void __maxnreg__(200) fill_stream_kernel(PairDev* const* __restrict__ pairs, int4* const* __restrict__ recs, int n_dir, int only_dir,
                                                         int ring_cap, int ring_rows) {
  fill_body<2>(pairs, n_dir, only_dir, ring_cap, ring_rows, recs);
}
// one warp per strip, no dynamic shared memory
__global__ void __launch_bounds__(32) fill_reg_kernel(PairDev* const* __restrict__ pairs, int n_dir, int only_dir) {
  fill_body<3>(pairs, n_dir, only_dir, 32, 0);
}
__global__ void __launch_bounds__(512) fill_kernel(PairDev* const* __restrict__ pairs, int n_dir, int only_dir, int ring_cap, int ring_rows) {
  fill_body<0>(pairs, n_dir, only_dir, ring_cap, ring_rows);
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
  HB_CUDA(ctx, cudaMallocFromPoolAsync((void**)&p, bytesA + n * sizeof(float2), ctx->pool, ctx->stream));
  pr->bwd_planes = p;
  pr->dev.bA = (float4*)p;
  pr->dev.bB = (float2*)(p + bytesA);
  HB_CUDA(ctx, cudaMemcpyAsync(pr->d_dev, &pr->dev, sizeof(PairDev), cudaMemcpyHostToDevice, ctx->stream));
  return HIST_OK;
}

int launch_fill_uniform(hist_ctx* ctx, hist_pair* const* pairs, int n, int n_dir, int only_dir) {
  int maxThreads = 1, maxDiag = 1;
  for (int k = 0; k < n; ++k) {
    maxThreads = std::max(maxThreads, pairs[k]->dev.max_threads);
    maxDiag = std::max(maxDiag, pairs[k]->dev.n_diag);
  }
  const int block = std::min(512, (maxThreads + 31) / 32 * 32);
  // shared-memory ring of the last R diagonals: as many rows as fit a modest budget (narrow strips
  // keep several CTAs per SM), at least 4 rows if the opt-in limit allows, none otherwise
  const int ring_cap = maxThreads;
  const size_t row = (size_t)ring_cap * 2 * sizeof(float4);
  const size_t limit = ctx->smem_optin > 4096 ? ctx->smem_optin - 4096 : 0;
  int rows = HB_RING_MAX;
  while (rows > 4 && rows * row > 32 * 1024) rows >>= 1;
  while (rows > 0 && rows * row > limit) rows >>= 1;
  if (rows < 4) rows = 0;
  const size_t smem = rows * row;
  if (smem + 4096 > 48 * 1024 && smem > ctx->fill_smem_configured) {
    HB_CUDA(ctx, cudaFuncSetAttribute(fill_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)limit));
    ctx->fill_smem_configured = limit;
  }
  std::vector<PairDev*> h(n);
  for (int k = 0; k < n; ++k) h[k] = pairs[k]->d_dev;
  PairDev** dptrs = nullptr;
  HB_CUDA(ctx, cudaMallocFromPoolAsync((void**)&dptrs, n * sizeof(PairDev*), ctx->pool, ctx->stream));
  HB_CUDA(ctx, cudaMemcpyAsync(dptrs, h.data(), n * sizeof(PairDev*), cudaMemcpyHostToDevice, ctx->stream));
  bool narrow = maxThreads <= HB_SMAX && rows >= 4;
  for (int k = 0; k < n && narrow; ++k) narrow = pairs[k]->dev.nS <= 32 && pairs[k]->dev.nE <= 32;   // edge-slot lists are staged in 32-entry arrays
  if (n_dir == 2 || only_dir == 0) {
    // the absorb plane is part of the Forward fill (computeLogProbAbsorb, forward.cpp:183); the
    // Backward fill reuses it (BackwardMatrix is built from a ForwardMatrix, forward.cpp:975)
    int maxKp = 4;
    for (int k = 0; k < n; ++k) maxKp = std::max(maxKp, pairs[k]->dev.Kp);
    int tile_rows = 96;
    while (tile_rows > 8 && (size_t)2 * tile_rows * maxKp * sizeof(float) > 96 * 1024) tile_rows >>= 1;
    const size_t abs_smem = (size_t)2 * tile_rows * maxKp * sizeof(float);
    if (abs_smem + 2048 > 48 * 1024 && abs_smem > ctx->absorb_smem_configured) {   // + the kernel's static shared memory
      HB_CUDA(ctx, cudaFuncSetAttribute(absorb_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
      ctx->absorb_smem_configured = 96 * 1024;
    }
    absorb_kernel<<<dim3((maxDiag + HB_ABS_DIAGS - 1) / HB_ABS_DIAGS, n), 256, abs_smem, ctx->stream>>>(dptrs, tile_rows);
    ctx->launches += 1;
  }
  static const bool use_queue = getenv("HIST_B200_FILL") && !strcmp(getenv("HIST_B200_FILL"), "queue");
  static const bool use_stream = getenv("HIST_B200_FILL") && !strcmp(getenv("HIST_B200_FILL"), "stream");
  static const bool use_reg = getenv("HIST_B200_FILL") && !strcmp(getenv("HIST_B200_FILL"), "reg");
  (void)use_stream;
  if (narrow && maxThreads <= 32 && use_reg) {
    // register wavefront: no record stream, no ring
    fill_reg_kernel<<<n * n_dir, 32, 0, ctx->stream>>>(dptrs, n_dir, only_dir);
    ctx->launches += 1;
    HB_CUDA(ctx, cudaGetLastError());
    HB_CUDA(ctx, cudaFreeAsync(dptrs, ctx->stream));
    return HIST_OK;
  }
  char* rec_buf = nullptr;
  const int n_cta = n * n_dir;
  std::vector<size_t> off(n_cta + 1, 0);
  const size_t tab = ((size_t)n_cta * sizeof(int4*) + 255) & ~(size_t)255;
  if (narrow && !use_queue) {
    // streamed records: 64 B per stored cell and direction (+ 32 B per diagonal), built by one parallel pass
    for (int b = 0; b < n_cta; ++b) off[b + 1] = off[b] + (((size_t)pairs[b / n_dir]->dev.n_stored * 64 + (size_t)pairs[b / n_dir]->dev.n_diag * 32 + 255) & ~(size_t)255);
    if (cudaMallocFromPoolAsync((void**)&rec_buf, tab + off[n_cta] + 256, ctx->pool, ctx->stream) != cudaSuccess) {
      cudaGetLastError();   // no room for the record buffer: the in-CTA producer/consumer kernel needs none
      rec_buf = nullptr;
    }
  }
  if (rec_buf) {
    std::vector<int4*> hrec(n_cta);
    for (int b = 0; b < n_cta; ++b) hrec[b] = (int4*)(rec_buf + tab + off[b]);
    HB_CUDA(ctx, cudaMemcpyAsync(rec_buf, hrec.data(), n_cta * sizeof(int4*), cudaMemcpyHostToDevice, ctx->stream));
    // ring rows: as many as still let the whole batch be resident at once (one wave)
    const int per_sm = (n_cta + ctx->sm_count - 1) / ctx->sm_count;
    int srows = HB_RING_MAX;
    const size_t qbytes = (size_t)HB_SDEPTH * hb_sslot(ring_cap) * 16;
    while (srows > 4 && (size_t)per_sm * (srows * row + qbytes + 2048) > 200 * 1024) srows >>= 1;
    const size_t ssmem = srows * row + qbytes;
    if (ssmem + 4096 > 48 * 1024) HB_CUDA(ctx, cudaFuncSetAttribute(fill_stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
    const int chunk = HB_BUILD_WARPS * HB_BUILD_DIAGS;
    build_records_kernel<<<dim3((maxDiag + chunk - 1) / chunk, n_cta), 32 * HB_BUILD_WARPS, 0, ctx->stream>>>(dptrs, (int4* const*)rec_buf, n_dir,
                                                                                                            only_dir, ring_cap, srows);
    fill_stream_kernel<<<n_cta, 32, ssmem, ctx->stream>>>(dptrs, (int4* const*)rec_buf, n_dir, only_dir, ring_cap, srows);
    ctx->launches += 1;
  } else if (narrow && maxThreads <= 32)
    fill_narrow_kernel<<<n * n_dir, 32 * (1 + HB_NPROD), smem + HB_DEPTH * 128 * sizeof(int4), ctx->stream>>>(dptrs, n_dir, only_dir, ring_cap, rows);
  else
    fill_kernel<<<n * n_dir, block, smem, ctx->stream>>>(dptrs, n_dir, only_dir, ring_cap, rows);
  ctx->launches += 1;
  HB_CUDA(ctx, cudaGetLastError());
  if (rec_buf) HB_CUDA(ctx, cudaFreeAsync(rec_buf, ctx->stream));
  HB_CUDA(ctx, cudaFreeAsync(dptrs, ctx->stream));
  // h[] is pageable host memory: the copy above has been staged by the time cudaMemcpyAsync returns
  return HIST_OK;
}

// leaf x leaf merges whose strips fit one warp and whose emission rows fit the register budget
bool chain_eligible(const hist_pair* pr) {
  static const bool off = getenv("HIST_B200_FILL") != nullptr;   // any explicit kernel choice switches the chain kernel off (A/B runs)
  const int K4 = pr->dev.Kp / 4;
  return !off && pr->dev.chain && pr->dev.max_threads <= 32 && (K4 == 1 || K4 == 2 || K4 == 5);
}

int launch_fill_chain(hist_ctx* ctx, hist_pair* const* pairs, int n, int n_dir, int only_dir) {
  // one launch per row length (a batch of leaf merges shares one alphabet: normally exactly one launch)
  for (int K4 : {1, 2, 5}) {
    std::vector<PairDev*> h;
    for (int k = 0; k < n; ++k)
      if (pairs[k]->dev.Kp / 4 == K4) h.push_back(pairs[k]->d_dev);
    if (h.empty()) continue;
    const int m = (int)h.size();
    PairDev** dptrs = nullptr;
    HB_CUDA(ctx, cudaMallocFromPoolAsync((void**)&dptrs, m * sizeof(PairDev*), ctx->pool, ctx->stream));
    HB_CUDA(ctx, cudaMemcpyAsync(dptrs, h.data(), m * sizeof(PairDev*), cudaMemcpyHostToDevice, ctx->stream));
    static const bool one_warp = getenv("HIST_B200_CHAIN") && !strcmp(getenv("HIST_B200_CHAIN"), "1warp");   // A/B runs
    if (one_warp) {
      if (K4 == 1) fill_chain_kernel<1><<<m * n_dir, 32, 0, ctx->stream>>>(dptrs, n_dir, only_dir);
      else if (K4 == 2) fill_chain_kernel<2><<<m * n_dir, 32, 0, ctx->stream>>>(dptrs, n_dir, only_dir);
      else fill_chain_kernel<5><<<m * n_dir, 32, 0, ctx->stream>>>(dptrs, n_dir, only_dir);
    } else {
      if (K4 == 1) fill_chain2_kernel<1><<<m * n_dir, 64, 0, ctx->stream>>>(dptrs, n_dir, only_dir, nullptr);
      else if (K4 == 2) fill_chain2_kernel<2><<<m * n_dir, 64, 0, ctx->stream>>>(dptrs, n_dir, only_dir, nullptr);
      else fill_chain2_kernel<5><<<m * n_dir, 64, 0, ctx->stream>>>(dptrs, n_dir, only_dir, nullptr);
    }
    ctx->launches += 1;
    HB_CUDA(ctx, cudaGetLastError());
    HB_CUDA(ctx, cudaFreeAsync(dptrs, ctx->stream));
  }
  return HIST_OK;
}

// Narrow strips (every anti-diagonal fits one warp) and general strips use different kernels; one wide merge
// must not push a whole level of narrow ones onto the general kernel, so a mixed batch is launched in two parts.
// Leaf x leaf merges go to the chain kernel first.
int launch_fill(hist_ctx* ctx, hist_pair* const* pairs_in, int n_in, int n_dir, int only_dir) {
  std::vector<hist_pair*> chain, rest;
  for (int k = 0; k < n_in; ++k) (chain_eligible(pairs_in[k]) ? chain : rest).push_back(pairs_in[k]);
  hist_pair* const* pairs = rest.data();
  const int n = (int)rest.size();
  std::vector<hist_pair*> narrow, wide;
  for (int k = 0; k < n; ++k) (pairs[k]->dev.max_threads <= 32 ? narrow : wide).push_back(pairs[k]);   // wide: streamed up to HB_SMAX slots, else general
  cudaEvent_t e0 = nullptr, e1 = nullptr;   // one event pair per fill call (bench.py's roofline leg)
  if (ctx->timing) {
    HB_CUDA(ctx, cudaEventCreate(&e0));
    HB_CUDA(ctx, cudaEventCreate(&e1));
    HB_CUDA(ctx, cudaEventRecord(e0, ctx->stream));
  }
  int rc = HIST_OK;
  auto ensure_stream2 = [&]() -> int {
    if (ctx->stream2) return HIST_OK;
    HB_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->stream2, cudaStreamNonBlocking));
    HB_CUDA(ctx, cudaEventCreateWithFlags(&ctx->ev_fork, cudaEventDisableTiming));
    HB_CUDA(ctx, cudaEventCreateWithFlags(&ctx->ev_join, cudaEventDisableTiming));
    return HIST_OK;
  };
  bool forked = false;
  if (!chain.empty()) {
    if (n == 0) {
      rc = launch_fill_chain(ctx, chain.data(), (int)chain.size(), n_dir, only_dir);
    } else {
      // a level of leaf merges with a few strips too wide for the chain kernel: those go through the three latency-bound
      // launches of the DAG path (~1 ms whatever their number); run them beside the chain kernel, not after it
      rc = ensure_stream2();
      if (!rc) {
        HB_CUDA(ctx, cudaEventRecord(ctx->ev_fork, ctx->stream));
        HB_CUDA(ctx, cudaStreamWaitEvent(ctx->stream2, ctx->ev_fork, 0));
        cudaStream_t main_stream = ctx->stream;
        ctx->stream = ctx->stream2;
        rc = launch_fill_chain(ctx, chain.data(), (int)chain.size(), n_dir, only_dir);
        ctx->stream = main_stream;
        forked = true;
      }
    }
  }
  if (rc || n == 0) {
    // nothing but leaf merges in this batch
  } else if (narrow.empty() || wide.empty()) {
    rc = launch_fill_uniform(ctx, pairs, n, n_dir, only_dir);
  } else {
    // the few general strips are latency-bound on their own (one CTA each): run them beside the narrow part
    { const int rc2 = ensure_stream2(); if (rc2) return rc2; }
    HB_CUDA(ctx, cudaEventRecord(ctx->ev_fork, ctx->stream));
    HB_CUDA(ctx, cudaStreamWaitEvent(ctx->stream2, ctx->ev_fork, 0));
    cudaStream_t main_stream = ctx->stream;
    ctx->stream = ctx->stream2;
    rc = launch_fill_uniform(ctx, wide.data(), (int)wide.size(), n_dir, only_dir);
    ctx->stream = main_stream;
    if (!rc) rc = launch_fill_uniform(ctx, narrow.data(), (int)narrow.size(), n_dir, only_dir);
    HB_CUDA(ctx, cudaEventRecord(ctx->ev_join, ctx->stream2));
    HB_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->ev_join, 0));
  }
  if (forked) {
    HB_CUDA(ctx, cudaEventRecord(ctx->ev_join, ctx->stream2));
    HB_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->ev_join, 0));
  }
  if (ctx->timing) {
    HB_CUDA(ctx, cudaEventRecord(e1, ctx->stream));
    ctx->fill_events.push_back({e0, e1});
  }
  return rc;
}

}  // namespace

// likelihoods of filled pairs: the scalar blocks (4 doubles) of pairs prepared in one batch are contiguous, so consecutive
// pairs come back in one transfer; waits for the context's stream
static int read_likelihoods(hist_ctx* ctx, hist_pair* const* pairs, int32_t n, int do_backward, double* lp_end, double* lp_start) {
  std::vector<double> tmp(4 * (size_t)n);
  for (int k = 0; k < n;) {
    int e = k + 1;
    while (e < n && pairs[e]->dev.lp_end == pairs[e - 1]->dev.lp_end + 4) ++e;
    HB_CUDA(ctx, cudaMemcpyAsync(&tmp[4 * (size_t)k], pairs[k]->dev.lp_end, (size_t)(e - k) * 4 * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    k = e;
  }
  HB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  for (int k = 0; k < n; ++k) {
    pairs[k]->fwd_done = true;
    pairs[k]->lp_end = lp_end[k] = tmp[4 * (size_t)k];
    if (do_backward) {
      pairs[k]->bwd_done = true;
      pairs[k]->lp_start = tmp[4 * (size_t)k + 1];
      if (lp_start) lp_start[k] = tmp[4 * (size_t)k + 1];
    }
  }
  return HIST_OK;
}

extern "C" int hist_fill_batch(hist_ctx* ctx, hist_pair* const* pairs, int32_t n, int do_backward, double* lp_end, double* lp_start) {
  if (!ctx || !pairs || n <= 0) return set_err(ctx, HIST_ERR_ARG, "hist_fill_batch: bad arguments");
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
  if (!lp_end) {   // enqueue only: the likelihoods stay on the device until a later call (or hist_ctx_synchronize) asks for them
    for (int k = 0; k < n; ++k) {
      pairs[k]->fwd_done = true;
      if (do_backward) pairs[k]->bwd_done = true;
    }
    return HIST_OK;
  }
  return read_likelihoods(ctx, pairs, n, do_backward, lp_end, lp_start);
}

extern "C" int hist_fill_batch_results(hist_ctx* ctx, hist_pair* const* pairs, int32_t n, int do_backward, double* lp_end, double* lp_start) {
  if (!ctx || !pairs || n <= 0 || !lp_end) return set_err(ctx, HIST_ERR_ARG, "hist_fill_batch_results: bad arguments");
  cudaSetDevice(ctx->device);
  for (int k = 0; k < n; ++k)
    if (!pairs[k] || pairs[k]->ctx != ctx || !pairs[k]->fwd_done || (do_backward && !pairs[k]->bwd_done))
      return set_err(ctx, HIST_ERR_STATE, "hist_fill_batch_results: pair %d has not been filled on this context", k);
  return read_likelihoods(ctx, pairs, n, do_backward, lp_end, lp_start);
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
  { int rc_ = hb::pair_cells(pair, nullptr); if (rc_) return rc_; }
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
  HB_CUDA(ctx, cudaMallocFromPoolAsync((void**)&d_cnt, NI * sizeof(unsigned), ctx->pool, ctx->stream));
  HB_CUDA(ctx, cudaMallocFromPoolAsync((void**)&d_start, NI * sizeof(unsigned long long), ctx->pool, ctx->stream));
  HB_CUDA(ctx, cudaMallocFromPoolAsync((void**)&d_lp, std::max<size_t>(1, nc) * 5 * sizeof(double), ctx->pool, ctx->stream));
  HB_CUDA(ctx, cudaMallocFromPoolAsync((void**)&d_pos, std::max<size_t>(1, nc) * 2 * sizeof(int), ctx->pool, ctx->stream));
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
