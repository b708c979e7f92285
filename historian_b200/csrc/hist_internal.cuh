// Internal layouts shared by the historian_b200 CUDA translation units (sm_100a).
//
// Numeric representation of a DP cell (the B200-first design choice):
// the reference stores 5 natural-log doubles per cell and combines them with a
// table-driven log-sum-exp (forward.h:14-20, logsumexp.h:42-85).  Here a cell is
// five FP32 *linear-space* values sharing one int32 base-2 exponent E:
//     P(state s) = v[s] * 2^E,   natural log = ln(v[s]) + E * ln 2.
// Predecessor sums become FMAs; the per-cell exponent gives unbounded dynamic
// range (no underflow to "-inf" that the reference would not have), and the
// relative error per operation is 2^-24 instead of the reference table's 4.5e-5
// absolute.  Cells are 24 B (float4 + float2) instead of 40 B.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/historian_b200.h"

namespace hb {

// exponent of an all-zero cell ("-inf" in the reference)
#define HB_EXP_ZERO (-(1 << 28))
#define HB_LN2 0.69314718055994530942

// per-state flag bits
enum {
  F_EMIT = 1, F_READY = 2 /* isReady() || profile empty */, F_EMIT_OR_START = 4,
  F_NEAR = 8,    // xNearStart / yNearEnd (forward.cpp:58-65): the whole row / column is in the envelope
  F_TO_END = 16  // has a null transition into END
};

// Packed per-state / per-edge records read by the fill kernels (built on device by
// pack_meta_kernel).  One 16-B load per record, coalesced across the lanes of a diagonal.
//   forward  state: fm_a = {flags | in_cnt << 8, in_begin, band key, edge slot}
//                   fm_b = first in-edge (in-degree is 1 for almost every state)
//   in-edge       : {src, exponent (dest emission folded in), mantissa bits, edge slot of src}
//   backward state: bm[4] = {flags | ao_cnt << 8, ao_begin, band key, edge slot},
//                           {no_begin, no_cnt, END-edge exponent, END-edge mantissa bits},
//                           first absorbing out-edge (2 x int4)
//   out-edge      : {dst, exponent, edge slot of dst, 0}, {m, m*rs[dst], m*ins[dst], 0} as float bits
#define HB_MAX_DEGREE ((1 << 23) - 1)

// One side (x or y profile) of a pair, device pointers.
struct SideDev {
  int n;                 // states incl. START/END
  int nt;                // transitions
  const uint8_t* flags;  // [n]
  // transitions (profile order)
  const int* t_src;      // [nt]
  const int* t_dst;      // [nt]
  const double* t_lp;    // [nt] natural log (input)
  float* t_m;            // [nt] mantissa in [1,2): exp(lp - floor(lp/ln2) ln2)
  int* t_e;              // [nt] floor(lp/ln2) + ex[dst]  (emission exponent of the destination folded in)
  // CSR lists of transition ids, ascending
  const int* in_off;  const int* in_idx;    // incoming
  const int* ao_off;  const int* ao_idx;    // absorbing out (dest emits)
  const int* no_off;  const int* no_idx;    // null out
  // emission tables, all relative to 2^ex[i]
  int* ex;               // [n] floor(max_k sub[i][k] / ln2), 0 for null states
  float* rs;             // [n] exp(rootsub - ex ln2)   (1 for null states)
  float* ins;            // [n] exp(ins     - ex ln2)   (1 for null states)
  float* H;              // [n][Kp] exp(sub - ex ln2), rows zero-padded to Kp = 4*ceil(K/4); the y side is pre-multiplied by root[k]
  // packed records for the fill kernels
  int4* fm_a; int4* fm_b; float2* emis;   // [n] forward meta, (rs, ins)
  int4* in_edges;                         // [max(1,nt)]
  int4* bm;                               // [4][n] backward meta (plane-major)
  int4* cf; int4* cb;                     // [n] chain records: {ex[i], rs[i], ins[i], key[i]} and {ex[i+1], rs[i+1], ins[i+1], key[i]}
  int4* ao_edges; int4* no_edges;         // [max(1,n_ao)][2], [max(1,n_no)][2]
  // natural-log doubles kept for fetch / makeProfile
  const double* lp_absorb;  // [n][K] input
  double* sub;           // [n][K] leftMultiply result (profile.cpp:78-91), -inf rows for null states
  double* rootsub;       // [n]
  double* insl;          // [n]
  // Leaf sides (profiles the library built from residues): the constructor tables take only A+1 distinct rows, one per
  // residue and one for the wildcard.  They are computed once per (merge, side) and scattered by the residue of each state.
  const int8_t* tokens;       // [n-2] token of state i+1 (< 0: wildcard); NULL: not a leaf side, every row is computed
  const double* tk_onehot;    // [A+1][K] lpAbsorb rows of the token classes (profile.cpp:52-66)
  double* tk_sub; float* tk_H; int* tk_ex; float* tk_rs; float* tk_ins; double* tk_rootsub; double* tk_insl;   // [A+1] rows
  // envelope
  int* key;              // [n] cum[rowPosToCol[closestLeafPos]] (written by pair_layout_kernel)
  const uint8_t* near_;  // [n] xNearStart / yNearEnd
  const int* slot;       // [n] edge slot of this row/column, -1 if none
};

// Device image of a profile's integer structure: what the DPMatrix constructor derives from the profile alone,
// whichever side of a merge it ends up on.
struct ProfImage {
  uint8_t* flags;                     // [n] F_EMIT | F_READY | F_EMIT_OR_START
  int* t_src; int* t_dst; double* t_lp;
  int* in_off; int* in_idx; int* ao_off; int* ao_idx; int* no_off; int* no_idx;
  int* leaf_pos;                      // [n] closest-leaf sequence position (zeros if the profile carries none)
  uint8_t* near_start; uint8_t* near_end;   // [n] xNearStart when used as x / yNearEnd when used as y
  int* slot_start; int* slot_end;     // [n] rank among the near rows / columns, -1 elsewhere
  int* start_rows; int* end_cols;     // [nS], [nE]
};

// per anti-diagonal d = i + j
struct DiagInfo {
  int ilo;     // first x index of the main interval
  int width;   // cells in the main interval (may be 0)
  unsigned off;  // first stored cell of this diagonal
  int pad;
};

// What pair_layout_kernel reports back to the host for one pair (the plane allocation is sized from it).
struct LayoutOut {
  unsigned long long n_stored;   // stored cells over all diagonals
  int max_threads, max_width;
  int err;                       // 0 ok; 1/2: an x/y closest-leaf position lies outside its guide row
  int err_state, err_pos, pad;
};

struct PairDev {
  SideDev x, y;
  int K;                  // C * A
  int Kp;                 // K rounded up to a multiple of 4 (row stride of H)
  int C, A;
  int max_dist;           // <0: no envelope
  float W[5][6];          // exp(lpTrans[src][dest]), 0 where absent
  double Wlog[6][6];      // lpTrans[src][dest] (natural log), -inf where absent
  int n_diag;
  int nS, nE;             // edge slots per diagonal: near-start rows, near-end columns
  const int* start_rows;  // [nS]
  const int* end_cols;    // [nE]
  DiagInfo* dinfo;        // [n_diag] (written by pair_layout_kernel)
  int max_threads;        // max over d of nS + width + nE
  int chain;              // both profiles are leaf chains (START -> residue 1 -> ... -> residue L -> END, log-weight 0 transitions)
  unsigned n_stored;
  float4* fA; float2* fB; // forward cells: (IMM,IMD,IDM,IMI), (IIW, E)
  float4* bA; float2* bB; // backward cells
  float* absorb;          // [n_stored] exp(computeLogProbAbsorb(i,j)) * 2^-(ex[i]+ey[j]) (forward.h:112-124); 0 unless both states emit
  double* lp_end;         // [1]
  double* lp_start;       // [1]
  unsigned long long* n_cells;  // [1] in-envelope cell count
  // inputs, scratch and result slot of pair_layout_kernel (band keys and the anti-diagonal layout are built on the device)
  const int* env_k1; const int* env_k2;   // cumulativeMatches[rowPosToCol[pos]] of guide rows 1 and 2 (alignpath.cpp:282-310), [len+1]
  int env_len1, env_len2;
  const int* x_leaf_pos; const int* y_leaf_pos;
  int* scratch;                 // [4 (nx-1) + 2 (ny-1)] jlo, jhi, rows, cnt | pmax, smin
  LayoutOut* layout_out;
  // model tables of this merge (inputs of the DPMatrix-constructor kernels)
  const double* l_sub; const double* r_sub;      // [C][A][A]
  const double* log_root;                        // [C][A]
  const double* l_ins; const double* r_ins;      // [C][A]
  const double* l_cw; const double* r_cw;        // [C]
};

__host__ __device__ inline int hb_iabs(int a) { return a < 0 ? -a : a; }

// 2^sh as a float for sh <= 0 (0 when it would be subnormal)
__device__ __forceinline__ float pow2_neg(int sh) {
  return sh < -126 ? 0.f : __int_as_float((sh + 127) << 23);
}

struct Cell {
  float v[5];
  int E;
};

__device__ __forceinline__ Cell zero_cell() {
  Cell c;
#pragma unroll
  for (int k = 0; k < 5; ++k) c.v[k] = 0.f;
  c.E = HB_EXP_ZERO;
  return c;
}

// Bring the accumulator to exponent max(acc.E, ce); returns the factor by which a
// contribution expressed at exponent ce must be multiplied.
__device__ __forceinline__ float acc_align(Cell& a, int ce) {
  // branch-free: lanes of a diagonal take different sides, so a branch would only serialise
  const int ne = max(a.E, ce);
  const float s = pow2_neg(a.E - ne);
#pragma unroll
  for (int k = 0; k < 5; ++k) a.v[k] *= s;
  a.E = ne;
  return pow2_neg(ce - ne);
}

// Renormalise so that max_k v[k] is in [1,2) (E absorbs the shift).
__device__ __forceinline__ void cell_normalise(Cell& a) {
  float m = fmaxf(fmaxf(fmaxf(a.v[0], a.v[1]), fmaxf(a.v[2], a.v[3])), a.v[4]);
  if (!(m > 0.f)) {
    a = zero_cell();
    return;
  }
  const int k = ((__float_as_int(m) >> 23) & 0xff) - 127;  // floor(log2 m) for normal m
  if (k != 0) {
    // 2^-k may not be representable for k = 127/-127: split in two
    const int k1 = k / 2, k2 = k - k1;
    const float s1 = __int_as_float((127 - k1) << 23), s2 = __int_as_float((127 - k2) << 23);
#pragma unroll
    for (int q = 0; q < 5; ++q) a.v[q] = a.v[q] * s1 * s2;
    a.E += k;
  }
}

__device__ __forceinline__ double cell_log(float v, int E) {
  return v > 0.f ? log((double)v) + (double)E * HB_LN2 : -INFINITY;
}

// The same for the DP-graph walkers (traceback, elimination, posteriors), where it is evaluated for every candidate of
// every step: ln of a FLOAT in FP64 without the library's full-precision log.  v = m 2^e, m in [1,2) split by its top five
// mantissa bits into c_k (1 + r), c_k = 1 + (k + 1/2) / 32, |r| <= 1/64: ln v = e ln 2 + ln c_k + r - r^2/2 + r^3/3 - r^4/4,
// truncation error <= r^5/5 < 2e-10 -- three hundred times below the 2^-24 relative error the stored FP32 value itself carries.
static __device__ const double hb_log_inv[32] = {0.9846153846153847, 0.9552238805970149, 0.927536231884058, 0.9014084507042254, 0.8767123287671232, 0.8533333333333334, 0.8311688311688312, 0.810126582278481, 0.7901234567901234, 0.7710843373493976, 0.7529411764705882, 0.735632183908046, 0.7191011235955056, 0.7032967032967034, 0.6881720430107527, 0.6736842105263158, 0.6597938144329897, 0.6464646464646465, 0.6336633663366337, 0.6213592233009708, 0.6095238095238096, 0.5981308411214953, 0.5871559633027523, 0.5765765765765766, 0.5663716814159292, 0.5565217391304348, 0.5470085470085471, 0.5378151260504201, 0.5289256198347108, 0.5203252032520326, 0.512, 0.5039370078740157};
static __device__ const double hb_log_ck[32] = {0.015504186535965254, 0.0458095360312942, 0.07522342123758753, 0.10379679368164356, 0.13157635778871926, 0.15860503017663857, 0.184922338494012, 0.21056476910734964, 0.2355660713127669, 0.25995752443692605, 0.2837681731306446, 0.3070250352949119, 0.329753286372468, 0.3519764231571782, 0.37371640979358406, 0.394993808240869, 0.415827895143711, 0.43623676677491807, 0.4562374334815876, 0.4758459048699639, 0.4950772667978515, 0.5139457511022343, 0.5324647988694718, 0.5506471179526623, 0.5685047353526688, 0.5860490450035782, 0.6032908514380843, 0.6202404097518576, 0.6369074622370692, 0.6533012720127457, 0.6694306539426292, 0.6853040030989194};
__device__ __forceinline__ double cell_log_fast(float v, int E) {
  if (!(v > 0.f)) return -INFINITY;
  const int bits = __float_as_int(v);
  const int e = (bits >> 23) & 0xff;
  if (e == 0 || e == 0xff) return log((double)v) + (double)E * HB_LN2;   // subnormal / inf: the plain route
  const int k = (bits >> 18) & 31;
  const double m = (double)__int_as_float((bits & 0x007fffff) | 0x3f800000);
  const double r = fma(m, hb_log_inv[k], -1.0);
  double p = fma(r, -0.25, 1.0 / 3.0);
  p = fma(p, r, -0.5);
  p = fma(p, r, 1.0);
  return fma(p, r, hb_log_ck[k]) + (double)(e - 127 + E) * HB_LN2;
}

}  // namespace hb

// ------------------------------------------------------------------ host-side objects
#include <string>
#include <vector>

// One device allocation shared by the pairs (or profiles) of a batch; freed with the last of them.
struct hb_arena {
  char* dev;
  size_t bytes;
  int refs;
  // pair arenas: the exact in-envelope cell counts (the GCUPS unit) stay on the device until somebody asks (hb::pair_cells)
  size_t scal_off = 0, ptr_off = 0;
  int n_pairs = 0, max_diag = 1;
  bool counted = false;
  std::vector<long long> n_cells;
};

struct hist_ctx {
  int device;
  cudaStream_t stream;
  cudaMemPool_t pool;   // the context's own stream-ordered pool: every device allocation of the library comes from it
  std::string err;
  long long launches;
  int sm_count;
  size_t smem_optin;
  size_t fill_smem_configured;
  size_t absorb_smem_configured;
  // optional per-kernel timing of the fill launches (bench.py's roofline leg)
  bool timing;
  std::vector<std::pair<cudaEvent_t, cudaEvent_t>> fill_events;
  // second stream + fork/join events: the general-strip part of a mixed fill batch runs beside the narrow part
  cudaStream_t stream2;
  cudaEvent_t ev_fork, ev_join;
  // pinned staging buffer for host-built layouts (grown on demand)
  char* staging;
  size_t staging_bytes;
  // second staging buffer + one event per buffer: hist_pair_prepare_batch ships a level in sub-batches so that the
  // host build of one overlaps the transfer and table kernels of the previous one
  char* staging2;
  size_t staging2_bytes;
  cudaEvent_t ev_stage[2];
  // pinned staging buffer of the profile structure images (host-built ones and the chain job records)
  char* staging_img;
  size_t staging_img_bytes;
  cudaEvent_t ev_img;
  std::vector<std::pair<std::pair<int, int>, double*>> onehot;   // (A, C) -> device table of token-class lpAbsorb rows
  long long h2d_bytes;   // bytes the library has copied host -> device for profiles and DPMatrix constructors (bench.py)
};

struct hist_profile {
  hist_ctx* ctx;
  int n, nt, A, C;
  // host copies (structure is needed on the host for CSR/layout building)
  std::vector<uint8_t> is_emit;
  std::vector<int> t_src, t_dst;
  std::vector<double> t_lp;
  std::vector<int> leaf_pos;  // empty if not given
  bool empty;                 // Profile::isEmpty
  bool sorted_in;             // every in-list ascending in source state, no repeated (src,dest)
  bool chain = false;         // leaf profile built by the library: transition t is t -> t+1, states 1..n-2 emit
  const int8_t* d_tokens = nullptr;   // leaf built by the library: its tokens on the device ([n-2], inside `arena`)
  bool leaf_std = false;      // chain whose closest-leaf positions are its own residues' (or absent): image generated on the device
  int refs = 0;               // pairs currently built on this profile (hist_profile_destroy defers while > 0)
  bool doomed = false;        // destroy requested while pairs still referred to it
  // device copy of lp_absorb
  double* d_lp_absorb;
  hb_arena* arena;            // owner of d_lp_absorb
  // device image of the integer structure (flags, transitions, CSR lists, edge rows/columns), built once per profile --
  // by a kernel for leaf chains, from the host lists otherwise -- the first time a DPMatrix constructor needs it
  hb::ProfImage img = {};
  hb_arena* img_arena = nullptr;
  int nS = 0, nE = 0;         // states near START when this profile is x / near END when it is y (forward.cpp:58-65)
};

struct hist_pair {
  hist_ctx* ctx;
  const hist_profile* x;
  const hist_profile* y;
  hb::PairDev dev;      // host copy of the device descriptor (pointers are device pointers)
  hb::PairDev* d_dev;   // device copy
  hb_arena* arena;      // the batch allocation holding every per-pair array but the cell planes
  hb_arena* planes;     // the batch allocation of the cell planes (sized by the layout the device computed)
  void* bwd_planes;     // separately allocated Backward planes (when not part of the arena)
  hist_pair_info info;   // n_cells is valid once hb::pair_cells has run (hist_pair_get_info does that)
  int batch_index = 0;   // position in its prepare batch (arena->n_cells)
  bool fwd_done, bwd_done;
  double lp_end, lp_start;
  double w_log[6][6];   // natural-log HMM weights (host, for traceback etc.)
  hist_pairhmm_desc hmm_scalars;  // scalar fields only (pointers nulled)
};

#include <functional>

namespace hb {
int set_err(hist_ctx* ctx, int code, const char* fmt, ...);
int check_cuda(hist_ctx* ctx, cudaError_t e, const char* what);

// fn(0..n-1) on the library's persistent host worker pool (hist_core.cu); at least `grain` items per thread
void host_parallel_for(int n, int grain, const std::function<void(int)>& fn);
// pinned staging buffer `slot` (0/1) of the context, grown on demand (hist_core.cu)
int ensure_staging_buf(hist_ctx* ctx, size_t bytes, int slot);
void arena_drop(hist_ctx* ctx, hb_arena* a);
// exact in-envelope cell count of a pair (forward.cpp:78-86): read back from the device, for the whole batch, on first use
int pair_cells(hist_pair* pair, long long* n_cells);

// ---- codon alphabets: Profile::leftMultiply on the tensor cores (hist_codon.cu)
bool leftmul_tc_applicable(int A, int C);
int launch_leftmul_tc(hist_ctx* ctx, PairDev* const* dptrs, int n_pairs, int max_states);

// ---- batched traceback (hist_trace.cu)
// Best traces from EEE of n filled pairs in ONE launch (one warp per pair).  The paths stay on the device:
// *d_paths is one allocation (release with cudaFreeAsync on the context's stream) in which pair k's path
// (start -> end) occupies [path_off[k], path_off[k] + len[k]).
int trace_best_batch(hist_ctx* ctx, hist_pair* const* pairs, int n, hist_cell** d_paths, std::vector<long long>& path_off,
                     std::vector<long long>& len);
// n_traces sampled traces of one pair in one launch; mt_snap (host, [n_traces][625]) receives the generator state
// after each trace, so the caller can rewind to the state after the traces it actually consumes.
int trace_sample_snap(hist_pair* pair, const uint32_t* mt_state, int n_traces, std::vector<hist_cell>& paths,
                      std::vector<long long>& path_off, std::vector<uint32_t>& mt_snap);

// ---- batched makeProfile elimination (hist_elim.cu)
struct ElimJob {
  hist_pair* pair;
  const hist_cell* d_cells;   // sorted cell set on the device, or nullptr ...
  const hist_cell* h_cells;   // ... then on the host (uploaded by the batch)
  int n;
  int strategy;
};
struct ElimOut {
  int n_states = 0, n_trans = 0;
  long long n_via = 0;
  std::vector<hist_cell> state_cell;
  std::vector<int> t_src, t_dst;
  std::vector<double> t_lp, t_best;
  std::vector<long long> via_off;
  std::vector<hist_cell> via;
};
int eliminate_batch(hist_ctx* ctx, const std::vector<ElimJob>& jobs, std::vector<ElimOut>& out);
// One row of a parent profile's lpAbsorb table (forward.cpp:713-727): the absorb vector of cell (x, y, s) of pair P,
// or a row of -inf when s < 0.
struct GatherRow {
  const PairDev* P;
  int x, y, s;
};
int gather_absorb_rows(hist_ctx* ctx, const GatherRow* d_rows, long long n_rows, int K, double* d_out);
}  // namespace hb

#define HB_CUDA(ctx, call)                                         \
  do {                                                             \
    int _rc = hb::check_cuda((ctx), (call), #call);                \
    if (_rc) return _rc;                                           \
  } while (0)
