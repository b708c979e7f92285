// historian_b200: Profile::leftMultiply for codon alphabets on the 5th-generation tensor cores.
//
// Replaces, for 32 < A <= 64 (the 61 sense codons of ECMrest / ECMunrest) and one rate component, the CUDA-core
// prepare_side_kernel of hist_core.cu: Profile::leftMultiply (reference profile.cpp:78-91) is
//     sub[i][a] = m_i + log( sum_d P(a -> d) * exp(lpAbsorb[i][d] - m_i) ),
// a dense [states x 64] . [64 x 64]^T product per profile side -- the one place on the hot path that really is a GEMM
// (north_star (2)).  One CTA takes 128 consecutive states of one (pair, side):
//   * every thread owns one state: it forms the row exp(lpAbsorb - m) in FP32 and writes it to shared memory in the
//     canonical K-major UMMA layout (8-row x 16-byte core matrices), split into a TF32 head and an FP32 remainder
//     (3xTF32: a.b ~ ah.bh + ah.bl + al.bh, error ~2^-21 instead of TF32's 2^-11); the substitution matrix is staged
//     the same way (padded with zeros to 64 x 64);
//   * one thread issues 24 tcgen05.mma.kind::tf32 (M=128, N=64, K=8 each: 8 K-steps x 3 split terms) accumulating in
//     FP32 in tensor memory (64 TMEM columns), commits to an mbarrier;
//   * the four warps read their 32 accumulator lanes back with tcgen05.ld and finish the DPMatrix-constructor tables
//     of their state (sub, rootsub, ins, the FP32 emission row and its exponent; forward.cpp:44-56) in FP64.
// Numerics: the products are FP32-accurate (3xTF32 + FP32 accumulate): |sub - exact| <~ 2e-6, against the reference's
// own table log-sum-exp error of 4.6e-5 per term.
#include <cmath>
#include <cstdio>

#include "hist_internal.cuh"

namespace hb {

#define TC_M 128
#define TC_N 64
#define TC_K 64
#define TC_THREADS 512   // 16 warps: four read the accumulator back (one TMEM lane quarter each), all sixteen build operands and tables
#define TC_ROWS_PER_WARP (TC_M / (TC_THREADS / 32))

// smem byte offset of element (row r, column k) of a [rows x 64] FP32 operand in the canonical K-major, no-swizzle UMMA
// layout: core matrices of 8 rows x 16 bytes, the 16 core matrices of an 8-row group contiguous (LBO = 128 B), the
// 8-row groups 2048 B apart (SBO)
__device__ __forceinline__ int tc_off(int r, int k) { return (r >> 3) * 2048 + (k >> 2) * 128 + (r & 7) * 16 + (k & 3) * 4; }

__device__ __forceinline__ uint64_t tc_desc(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3fff);          // start address
  d |= (uint64_t)((128 >> 4) & 0x3fff) << 16;          // leading (K) byte offset between core matrices
  d |= (uint64_t)((2048 >> 4) & 0x3fff) << 32;         // stride (M/N) byte offset between 8-row groups
  d |= 1ull << 46;                                     // descriptor version (Blackwell)
  return d;                                            // base offset 0, layout type 0 = no swizzle
}

// kind::tf32, FP32 accumulate, A and B K-major, N = 64, M = 128
#define TC_IDESC ((1u << 4) | (2u << 7) | (2u << 10) | ((TC_N >> 3) << 17) | ((TC_M >> 4) << 24))

__device__ __forceinline__ void tc_split(float v, float& hi, float& lo) {
  hi = __uint_as_float(__float_as_uint(v) & 0xffffe000u);   // what the tensor core reads of v (10 mantissa bits)
  lo = v - hi;                                                // exact
}

// grid = (state tiles, pairs, 2 sides); TC_THREADS threads; dynamic smem = 2 * 32 KB (A hi/lo) + 2 * 16 KB (B hi/lo)
__global__ void __launch_bounds__(TC_THREADS) leftmul_tc_kernel(PairDev* const* __restrict__ pairs) {
  extern __shared__ __align__(1024) unsigned char tc_smem[];
  __shared__ uint64_t s_bar;
  __shared__ uint32_t s_tmem;
  const PairDev& P = *pairs[blockIdx.y];
  const bool isY = blockIdx.z == 1;
  const SideDev& s = isY ? P.y : P.x;
  const int i0 = blockIdx.x * TC_M;
  if (i0 >= s.n) return;   // uniform over the CTA
  const int A = P.A, K = P.K, Kp = P.Kp;
  const int tid = threadIdx.x, warp = tid >> 5;
  unsigned char* a_hi = tc_smem;
  unsigned char* a_lo = a_hi + TC_M * TC_K * 4;
  unsigned char* b_hi = a_lo + TC_M * TC_K * 4;
  unsigned char* b_lo = b_hi + TC_N * TC_K * 4;
  const double* __restrict__ subMat = isY ? P.r_sub : P.l_sub;
  const double* __restrict__ logRoot = P.log_root;
  const double* __restrict__ logIns = isY ? P.r_ins : P.l_ins;
  const double* __restrict__ logCptWeight = isY ? P.r_cw : P.l_cw;

  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"((uint32_t)__cvta_generic_to_shared(&s_bar)) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(&s_tmem)), "r"(TC_N)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  // ---- B operand: P(a -> d), rows a, zero-padded to 64 x 64
  for (int q = tid; q < TC_N * TC_K; q += TC_THREADS) {
    const int a = q >> 6, d = q & 63;
    const float v = (a < A && d < A) ? (float)subMat[(size_t)a * A + d] : 0.f;
    float hi, lo;
    tc_split(v, hi, lo);
    *reinterpret_cast<float*>(b_hi + tc_off(a, d)) = hi;
    *reinterpret_cast<float*>(b_lo + tc_off(a, d)) = lo;
  }
  // ---- A operand: rows exp(lpAbsorb[i] - m_i) of the tile's 128 states.  A warp takes a row at a time (lanes over the
  // residues: coalesced loads, the maximum by shuffle), 32 rows per warp.
  __shared__ double s_m[TC_M];
  __shared__ unsigned char s_live[TC_M];
  __shared__ double s_rootp[TC_N];   // exp(logRoot): the y side's emission rows carry the root distribution
  if (tid < TC_N) s_rootp[tid] = tid < A ? exp(logRoot[tid]) : 0.0;
  const int lane = tid & 31;
  for (int r = warp * TC_ROWS_PER_WARP; r < (warp + 1) * TC_ROWS_PER_WARP; ++r) {
    const int i = i0 + r;
    const bool live = i < s.n && i > 0 && i < s.n - 1 && (s.flags[i] & F_EMIT);
    const double* la = s.lp_absorb + (size_t)(live ? i : 0) * K;
    const double v0 = live && lane < A ? la[lane] : -INFINITY, v1 = live && lane + 32 < A ? la[lane + 32] : -INFINITY;
    double m = fmax(v0, v1);
    for (int o = 16; o; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
    // one-hot leaf rows are all trivial cases (exp(0), exp(-inf)): no FP64 exp for them
    const float e0 = v0 == -INFINITY ? 0.f : v0 == m ? 1.f : (float)exp(v0 - m);
    const float e1 = v1 == -INFINITY ? 0.f : v1 == m ? 1.f : (float)exp(v1 - m);
    float hi, lo;
    tc_split(e0, hi, lo);
    *reinterpret_cast<float*>(a_hi + tc_off(r, lane)) = hi;
    *reinterpret_cast<float*>(a_lo + tc_off(r, lane)) = lo;
    tc_split(e1, hi, lo);
    *reinterpret_cast<float*>(a_hi + tc_off(r, lane + 32)) = hi;
    *reinterpret_cast<float*>(a_lo + tc_off(r, lane + 32)) = lo;
    if (lane == 0) {
      s_m[r] = m;
      s_live[r] = live && m > -INFINITY;
    }
  }
  // generic-proxy writes -> visible to the tensor core (async proxy)
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = s_tmem;
  if (tid == 0) {
    const uint32_t ah = (uint32_t)__cvta_generic_to_shared(a_hi), al = (uint32_t)__cvta_generic_to_shared(a_lo);
    const uint32_t bh = (uint32_t)__cvta_generic_to_shared(b_hi), bl = (uint32_t)__cvta_generic_to_shared(b_lo);
    int first = 1;
    for (int term = 0; term < 3; ++term) {
      const uint32_t abase = term == 2 ? al : ah, bbase = term == 1 ? bl : bh;   // ah.bh + ah.bl + al.bh
      for (int kk = 0; kk < TC_K / 8; ++kk) {   // one instruction covers K = 8 tf32 = two 16-byte core-matrix columns
        const uint64_t da = tc_desc(abase + kk * 256), db = tc_desc(bbase + kk * 256);
        const uint32_t acc = first ? 0u : 1u;
        asm volatile(
            "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
            "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(tmem),
            "l"(da), "l"(db), "r"((uint32_t)TC_IDESC), "r"(acc)
            : "memory");
        first = 0;
      }
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"((uint32_t)__cvta_generic_to_shared(&s_bar))
                 : "memory");
  }
  // ---- wait for the accumulator
  {
    const uint32_t bar = (uint32_t)__cvta_generic_to_shared(&s_bar);
    asm volatile(
        "{\n .reg .pred p;\n TC_WAIT:\n mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n @p bra TC_DONE;\n bra TC_WAIT;\n TC_DONE:\n}" ::"r"(bar),
        "r"(0u)
        : "memory");
  }
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  // accumulator row of this thread: TMEM lane = row (warp w reads lanes 32w .. 32w+31), column = a.  The rows go back
  // to shared memory (row-major, over the A operand) so that the epilogue can again run with lanes over the residues.
  float* s_acc = reinterpret_cast<float*>(a_hi);   // [TC_M][TC_N]; the MMAs have completed (mbarrier), the operand is dead
  if (warp < 4) {
    const uint32_t taddr = tmem + ((uint32_t)(warp * 32) << 16);
#pragma unroll
    for (int c0 = 0; c0 < TC_N; c0 += 16) {
      uint32_t r[16];
      asm volatile(
          "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];\n"
          : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]),
            "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
          : "r"(taddr + c0));
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
      for (int q = 0; q < 16; ++q) s_acc[tid * TC_N + ((c0 + q + tid) & (TC_N - 1))] = __uint_as_float(r[q]);   // rotated: conflict-free
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(TC_N) : "memory");
  // ---- epilogue: the tables of the DPMatrix constructor (as prepare_side_kernel), a warp per state, lanes over the residues
  for (int r = warp * TC_ROWS_PER_WARP; r < (warp + 1) * TC_ROWS_PER_WARP; ++r) {
    const int i = i0 + r;
    if (i >= s.n) break;
    if (!s_live[r]) {
      for (int k = lane; k < K; k += 32) s.sub[(size_t)i * K + k] = -INFINITY;
      for (int k = lane; k < Kp; k += 32) s.H[(size_t)i * Kp + k] = 0.f;
      if (lane == 0) {
        s.ex[i] = 0; s.rs[i] = 1.f; s.ins[i] = 1.f;
        s.rootsub[i] = -INFINITY; s.insl[i] = -INFINITY;
      }
      continue;
    }
    const double m = s_m[r];
    const double* la = s.lp_absorb + (size_t)i * K;
    double sv[2], rt[2], it[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int a = lane + 32 * h;
      const float c = a < A ? s_acc[r * TC_N + ((a + r) & (TC_N - 1))] : 0.f;
      sv[h] = c > 0.f ? m + (c == 1.f ? 0.0 : log((double)c)) : -INFINITY;
      rt[h] = a < A ? logRoot[a] + sv[h] : -INFINITY;
      it[h] = a < A ? logCptWeight[0] + logIns[a] + la[a] : -INFINITY;
      if (a < A) s.sub[(size_t)i * K + a] = sv[h];
    }
    double mx = fmax(sv[0], sv[1]), lr = fmax(rt[0], rt[1]), li = fmax(it[0], it[1]);
    for (int o = 16; o; o >>= 1) {
      mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, o));
      lr = fmax(lr, __shfl_xor_sync(0xffffffffu, lr, o));
      li = fmax(li, __shfl_xor_sync(0xffffffffu, li, o));
    }
    double racc = 0, iacc = 0;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      if (rt[h] > -INFINITY) racc += exp(rt[h] - lr);
      if (it[h] > -INFINITY) iacc += it[h] == li ? 1.0 : exp(it[h] - li);
    }
    for (int o = 16; o; o >>= 1) {
      racc += __shfl_xor_sync(0xffffffffu, racc, o);
      iacc += __shfl_xor_sync(0xffffffffu, iacc, o);
    }
    const double s_root = lr > -INFINITY ? lr + log(racc) : -INFINITY;
    const double s_ins = li > -INFINITY ? li + (iacc == 1.0 ? 0.0 : log(iacc)) : -INFINITY;
    const int ex = mx > -INFINITY ? (int)floor(mx / HB_LN2) : 0;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int k = lane + 32 * h;
      if (k < Kp) {
        double hv = 0.0;
        if (k < A && sv[h] > -INFINITY) {
          hv = exp(sv[h] - ex * HB_LN2);
          if (isY) hv *= s_rootp[k];
        }
        s.H[(size_t)i * Kp + k] = (float)hv;
      }
    }
    if (lane == 0) {
      s.ex[i] = ex;
      s.rootsub[i] = s_root;
      s.insl[i] = s_ins;
      s.rs[i] = s_root > -INFINITY ? (float)exp(s_root - ex * HB_LN2) : 0.f;
      s.ins[i] = s_ins > -INFINITY ? (float)exp(s_ins - ex * HB_LN2) : 0.f;
    }
  }
}

// true if the tensor-core constructor applies to a batch of pairs with this alphabet
bool leftmul_tc_applicable(int A, int C) {
  static const bool off = getenv("HIST_B200_NO_TENSOR") != nullptr;   // A/B runs against the CUDA-core kernel
  return !off && C == 1 && A > 32 && A <= TC_N;
}

int launch_leftmul_tc(hist_ctx* ctx, PairDev* const* dptrs, int n_pairs, int max_states) {
  const size_t smem = 2 * TC_M * TC_K * 4 + 2 * TC_N * TC_K * 4;
  static bool configured = false;
  if (!configured) {
    HB_CUDA(ctx, cudaFuncSetAttribute(leftmul_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    configured = true;
  }
  leftmul_tc_kernel<<<dim3((max_states + TC_M - 1) / TC_M, n_pairs, 2), TC_THREADS, smem, ctx->stream>>>(dptrs);
  ctx->launches += 1;
  return check_cuda(ctx, cudaGetLastError(), "leftmul_tc_kernel");
}

}  // namespace hb
