/* historian_b200 — C ABI of the B200-native hot path of Historian.
 *
 * The reference (evoldoers/historian) has no FFI layer: the hot path sits behind
 * ordinary C++ classes in one binary.  This header is the
 * thin C boundary UNDER those classes: every entry point names the reference
 * interface (file:line under the reference's src/) that it replaces.  Plain
 * pointers and sizes only; no C++ or torch types.  INTEGRATION.md shows how the
 * reference's classes bind to it.
 *
 * Conventions
 *   - every function returns 0 (HIST_OK) or a negative hist_status; the message
 *     is available from hist_last_error().  Nothing here throws, aborts or exits
 *     (the reference's Abort/Fail, util.cpp:28-55, are mapped by the C++ wrapper).
 *   - "zero forward likelihood" is NOT an error: lp_end = -inf is returned and the
 *     caller widens the band (recon.cpp:954-975).
 *   - descriptors (hist_*_desc) are borrowed for the duration of the call only.
 *   - handles (hist_profile, hist_pair, hist_sumprod) are library-owned device
 *     objects, released by the matching _destroy.  A hist_pair keeps its two
 *     profiles alive: hist_profile_destroy on a profile that pairs still use is
 *     deferred until the last of those pairs is destroyed.
 *   - a hist_ctx is bound to one device and one stream and is not thread-safe;
 *     multi-GPU = one ctx (one process) per device.
 *   - there is no CPU fallback: hist_ctx_create fails if no CUDA device is usable.
 *   - PairHMM state order everywhere: IMM=0, IMD=1, IDM=2, IMI=3, IIW=4, EEE=5
 *     (pairhmm.h:14-18).
 */
#ifndef HISTORIAN_B200_H
#define HISTORIAN_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
  HIST_OK = 0,
  HIST_ERR_ARG = -1,      /* malformed descriptor / argument */
  HIST_ERR_CUDA = -2,     /* CUDA runtime failure (message holds cudaGetErrorString) */
  HIST_ERR_NO_DEVICE = -3,/* no usable CUDA device: there is no CPU fallback */
  HIST_ERR_STATE = -4,    /* call order violated (e.g. trace before fill) */
  HIST_ERR_CAPACITY = -5, /* caller buffer too small; required size is reported */
  HIST_ERR_NUMERIC = -6   /* zero likelihood where the reference would Abort */
} hist_status;

enum { HIST_IMM = 0, HIST_IMD = 1, HIST_IDM = 2, HIST_IMI = 3, HIST_IIW = 4, HIST_EEE = 5, HIST_NSTATES = 5 };

/* Order of the 23 PairHMM log transition weights (pairhmm.h:50-54). */
enum {
  HIST_W_IMM_IMM = 0, HIST_W_IMM_IMD, HIST_W_IMM_IDM, HIST_W_IMM_IMI, HIST_W_IMM_IIW, HIST_W_IMM_EEE,
  HIST_W_IMD_IMM, HIST_W_IMD_IMD, HIST_W_IMD_IDM, HIST_W_IMD_EEE,
  HIST_W_IDM_IMM, HIST_W_IDM_IMD, HIST_W_IDM_IDM, HIST_W_IDM_EEE,
  HIST_W_IMI_IMM, HIST_W_IMI_IMD, HIST_W_IMI_IMI, HIST_W_IMI_IIW, HIST_W_IMI_EEE,
  HIST_W_IIW_IMM, HIST_W_IIW_IDM, HIST_W_IIW_IIW, HIST_W_IIW_EEE,
  HIST_NWEIGHTS = 23
};

/* ProfilingStrategy bits (forward.h:42-46). */
enum {
  HIST_COLLAPSE_CHAINS = 1, HIST_COUNT_SUBST = 2, HIST_COUNT_INDEL = 4,
  HIST_INCLUDE_BEST_TRACE = 8, HIST_KEEP_GAPS_OPEN = 16
};

typedef struct hist_ctx hist_ctx;
typedef struct hist_profile hist_profile;
typedef struct hist_pair hist_pair;
typedef struct hist_sumprod hist_sumprod;

/* A DP cell coordinate: DPMatrix::CellCoords (forward.h:26-38). */
typedef struct { int32_t xpos, ypos, state; } hist_cell;

/* ------------------------------------------------------------------ T1
 * Flattened Profile / ProfileState / ProfileTransition (profile.h:13-76).
 * State 0 is START, state n_states-1 is END; transitions are toposorted
 * (src < dest, profile.cpp:345,350).  A state is "null" iff is_emit == 0
 * (profile.h:32); "ready" iff it has no null out-transition (profile.h:36);
 * a transition is an absorbOut of its source iff its destination emits, else a
 * nullOut (forward.cpp:828).  Per-state transition lists (state.in, nullOut,
 * absorbOut) are the transitions in ascending index, which is the order the
 * reference builds them in. */
typedef struct {
  int32_t n_states;
  int32_t n_trans;
  int32_t alph_size;                 /* A */
  int32_t components;                /* C: rate-mixture components */
  const uint8_t* is_emit;            /* [n_states] */
  const int32_t* trans_src;          /* [n_trans] */
  const int32_t* trans_dest;         /* [n_trans] */
  const double* trans_lp;            /* [n_trans] ProfileTransition::lpTrans */
  const double* lp_absorb;           /* [n_states][C][A] ProfileState::lpAbsorb; rows of null states are ignored */
  const int32_t* closest_leaf_pos;   /* [n_states] state.seqCoords.at(closest-leaf row), 0 for START (forward.cpp:36-42); may be NULL when no envelope is used */
} hist_profile_desc;

/* ------------------------------------------------------------------ T2
 * PairHMM (pairhmm.h:8-65, pairhmm.cpp:5-44) with the two branch ProbModels
 * (model.cpp:374-391) and their LogProbModels (model.cpp:492). */
typedef struct {
  int32_t alph_size;
  int32_t components;
  double lp_trans[HIST_NWEIGHTS];    /* log transition weights, order HIST_W_* */
  const double* log_root;            /* [C][A] PairHMM::logRoot (log cptWeight folded in, pairhmm.cpp:13-15) */
  const double* l_sub;               /* [C][A][A] hmm.l.subMat, P(parent a -> child d) */
  const double* r_sub;               /* [C][A][A] hmm.r.subMat */
  const double* l_log_ins;           /* [C][A] hmm.logl.logInsProb */
  const double* r_log_ins;           /* [C][A] hmm.logr.logInsProb */
  const double* l_log_cpt_weight;    /* [C] hmm.logl.logCptWeight */
  const double* r_log_cpt_weight;    /* [C] hmm.logr.logCptWeight */
  double l_t, r_t;                   /* branch lengths (ProbModel::t) */
  double l_ins_wait, l_del_wait;     /* ProbModel::insWait/delWait */
  double r_ins_wait, r_del_wait;
} hist_pairhmm_desc;

/* ------------------------------------------------------------------ T3
 * GuideAlignmentEnvelope (alignpath.h:43-62, alignpath.cpp:282-310).
 * max_distance < 0 means "no envelope" (every cell in range) and the arrays may
 * be NULL.  The band predicate itself is integer and reproduced bit-exactly:
 * inEnvelope(i,j) = xNearStart[i] || yNearEnd[j] ||
 *   abs(cum[row1PosToCol[xpos(i)]] - cum[row2PosToCol[ypos(j)]]) <= max_distance
 * (forward.h:92-98). */
typedef struct {
  int32_t max_distance;
  int32_t n_cols;                    /* guide alignment columns */
  const int32_t* cumulative_matches; /* [n_cols+1] */
  int32_t row1_len, row2_len;        /* residues in row1 / row2 */
  const int32_t* row1_pos_to_col;    /* [row1_len+1] */
  const int32_t* row2_pos_to_col;    /* [row2_len+1] */
} hist_env_desc;

typedef struct {
  int32_t x_size, y_size;            /* profile sizes incl. START and END */
  int64_t n_cells;                   /* in-envelope (i,j) pairs, i<x_size-1, j<y_size-1: exactly the
                                        cells visited by forward.cpp:78-86 */
  int64_t n_stored;                  /* cells allocated in HBM (>= n_cells: band rows are padded to intervals) */
  int32_t n_diagonals;               /* wavefront steps */
  int32_t max_diag_width;            /* widest anti-diagonal in stored cells */
  int64_t h2d_bytes;                 /* bytes hist_pair_prepare copied host->device for this pair */
} hist_pair_info;

/* ------------------------------------------------------------------ context */
int hist_ctx_create(int device, hist_ctx** out);
void hist_ctx_destroy(hist_ctx* ctx);
/* Last error message of this context (or of the failed create when ctx == NULL). */
const char* hist_last_error(const hist_ctx* ctx);
/* CUDA stream the context launches on (a cudaStream_t), for event timing by the caller. */
void* hist_ctx_stream(hist_ctx* ctx);
int hist_ctx_synchronize(hist_ctx* ctx);
/* Number of kernels this context has launched so far (bench.py's gpu_launches). */
int64_t hist_ctx_launch_count(const hist_ctx* ctx);
/* Bytes this context has copied host -> device so far for leaf profiles, profile structure and the
 * DPMatrix constructors (bench.py's e2e.h2d_bytes_per_step is a difference of two readings). */
int64_t hist_ctx_h2d_bytes(const hist_ctx* ctx);
/* Device-side timing of the Forward/Backward fill kernels: when enabled, every fill launch is
 * bracketed by CUDA events on the context's stream.  hist_ctx_fill_time_ms waits for them,
 * returns the summed kernel time and launch count since the last query, and resets. */
int hist_ctx_set_timing(hist_ctx* ctx, int enabled);
int hist_ctx_fill_time_ms(hist_ctx* ctx, double* total_ms, int64_t* n_launches);

/* ------------------------------------------------------------------ profiles */
/* Replaces holding a `Profile` value on the host (profile.h:41-76). */
int hist_profile_upload(hist_ctx* ctx, const hist_profile_desc* desc, hist_profile** out);
/* Leaf profiles, Profile(components, alphabet, FastSeq, rowIndex) (profile.cpp:23-76), for a batch of n
 * sequences, built on device: START -> residue 1 -> ... -> residue L -> END with log-weight 0 transitions;
 * lpAbsorb of residue state pos+1 is 0 at its token and -inf elsewhere, or all 0 for a wildcard / invalid
 * character (token < 0).  closest_leaf_pos of state i is min(i, L) (the seqCoords of the leaf row).
 * tokens[k] points at lengths[k] tokens (Alphabet tokenize, fastseq.cpp:10-16).  Only the residues cross
 * the bus, not the [L+2][C][A] table. */
int hist_profile_from_sequences(hist_ctx* ctx, int32_t n, const int8_t* const* tokens, const int32_t* lengths,
                                int32_t alph_size, int32_t components, hist_profile** out);

/* The same constructor from residue characters: tokenisation as tokenize() does it
 * (fastseq.cpp:10-16: the exact character's first occurrence in `alphabet`, else the other
 * case's; anything else, the wildcard included, is an invalid token = wildcard row of zeros,
 * profile.cpp:56-60).  residues[k] need not be NUL-terminated. */
int hist_profile_from_residues(hist_ctx* ctx, int32_t n, const char* const* residues, const int32_t* lengths,
                               const char* alphabet, int32_t components, hist_profile** out);
void hist_profile_destroy(hist_profile* p);

/* ------------------------------------------------------------------ DPMatrix
 * hist_pair_prepare = the DPMatrix constructor (forward.cpp:11-66): leftMultiply
 * of both profiles (profile.cpp:78-91), the insx/insy/rootsubx/rootsuby tables
 * (forward.cpp:44-56), xNearStart/yNearEnd (forward.cpp:58-65), closest-leaf band
 * keys (forward.cpp:36-42), plus the HBM band layout.  All on device. */
int hist_pair_prepare(hist_ctx* ctx, const hist_profile* x, const hist_profile* y,
                      const hist_pairhmm_desc* hmm, const hist_env_desc* env, hist_pair** out);
/* Batched form: the DPMatrix constructors of all sibling merges of one guide-tree depth
 * (independent iterations of the recon.cpp:938-1033 loop) in one host->device transfer and one
 * launch per table kernel.  xs/ys/hmms: [n]; envs: [n] or NULL (entries may be NULL = no envelope).
 * flags: HIST_PREP_BACKWARD reserves the Backward planes up front (otherwise hist_bwd_fill
 * allocates them on first use).  out: [n] handles, each released by hist_pair_destroy. */
enum { HIST_PREP_BACKWARD = 1 };
int hist_pair_prepare_batch(hist_ctx* ctx, int32_t n, const hist_profile* const* xs, const hist_profile* const* ys,
                            const hist_pairhmm_desc* const* hmms, const hist_env_desc* const* envs, int32_t flags,
                            hist_pair** out);
void hist_pair_destroy(hist_pair* pair);
/* Destructors of a whole level at once (one call instead of n). */
void hist_pairs_destroy(hist_pair* const* pairs, int32_t n);
void hist_profiles_destroy(hist_profile* const* profiles, int32_t n);
int hist_pair_get_info(const hist_pair* pair, hist_pair_info* info);

/* Device results of the constructor tables, as natural-log doubles.
 * subx/suby: [n_states][C][A] (rows of null states hold -inf); the four scalar
 * tables: [n_states].  Any pointer may be NULL. */
int hist_pair_fetch_tables(hist_pair* pair, double* subx, double* suby,
                           double* insx, double* insy, double* rootsubx, double* rootsuby);

/* Band bookkeeping (bit-exact integers): x_key/y_key are
 * cum[rowPosToCol[closestLeafPos]] per state, near_start/near_end the edge flags.
 * in_envelope, if non-NULL, is the full [x_size-1][y_size-1] predicate as bytes
 * (only sensible for small matrices). */
int hist_pair_fetch_envelope(hist_pair* pair, int32_t* x_key, int32_t* y_key,
                             uint8_t* x_near_start, uint8_t* y_near_end, uint8_t* in_envelope);

/* ForwardMatrix::ForwardMatrix fill (forward.cpp:68-223): anti-diagonal
 * wavefront over the envelope; *lp_end = ForwardMatrix::lpEnd. */
int hist_fwd_fill(hist_pair* pair, double* lp_end);
/* BackwardMatrix::BackwardMatrix fill (forward.cpp:975-1097); *lp_start =
 * BackwardMatrix::lpStart().  Requires hist_fwd_fill first (BackwardMatrix takes
 * a ForwardMatrix). */
int hist_bwd_fill(hist_pair* pair, double* lp_start);

/* Batched form used for all sibling merges of one guide-tree depth
 * (recon.cpp:938-1033 loop, independent iterations): one launch fills every
 * pair.  do_backward != 0 also runs the Backward fills (concurrently).
 * lp_end / lp_start: [n] outputs (lp_start may be NULL when !do_backward).  lp_end == NULL: the fills are only enqueued on
 * the context's stream (no read-back, no synchronisation); the matrices are usable by later calls on the same context. */
int hist_fill_batch(hist_ctx* ctx, hist_pair* const* pairs, int32_t n, int do_backward,
                    double* lp_end, double* lp_start);
/* The likelihoods of pairs whose fills were enqueued with lp_end == NULL: waits for the context's stream and reads them
 * back (ForwardMatrix::lpEnd / BackwardMatrix::lpStart()).  With two contexts on one device a caller can prepare the next
 * batch on one while the other's fills run. */
int hist_fill_batch_results(hist_ctx* ctx, hist_pair* const* pairs, int32_t n, int do_backward,
                            double* lp_end, double* lp_start);

/* DPMatrix::cell for every in-envelope cell, in the row-major order of
 * DPMatrix::write (forward.cpp:448-464): lp[k*5+s] = natural-log value of state s
 * of the k-th cell, (xpos[k], ypos[k]) its coordinates.  which: 0 = Forward,
 * 1 = Backward.  capacity in cells; *n_cells receives the count. */
int hist_cells_fetch(hist_pair* pair, int which, double* lp, int32_t* xpos, int32_t* ypos,
                     int64_t capacity, int64_t* n_cells);

/* ------------------------------------------------------------------ traceback
 * ForwardMatrix::bestTrace(end) (forward.cpp:278-302): greedy arg-max walk from
 * `end` back to (0,0); candidates ordered by CellCoords::operator<, first strict
 * maximum wins (forward.cpp:245-255).  The path is returned start-to-end and
 * includes `end`.  Pass end.state == HIST_EEE at (x_size-1, y_size-1) for the
 * full trace. */
int hist_trace_best(hist_pair* pair, hist_cell end, hist_cell* path, int64_t capacity, int64_t* n);

/* ForwardMatrix::sampleTrace (forward.cpp:257-276) with sampleCell
 * (forward.cpp:225-243).  mt_state is a std::mt19937 state image: 624 words
 * followed by the position index (625 uint32); it is advanced exactly as
 * libstdc++'s uniform_real_distribution<double>(0,ptot)(generator) would
 * (two 32-bit draws per step), so the caller's generator stays in lock-step with
 * the reference's stream.  n_traces traces are sampled in sequence; path_off
 * [n_traces+1] delimits them inside `paths`. */
int hist_trace_sample(hist_pair* pair, uint32_t* mt_state, int32_t n_traces,
                      hist_cell* paths, int64_t capacity, int64_t* path_off);

/* BackwardMatrix::bestTrace(start) (forward.cpp:1287-1300): greedy forward walk
 * through the Backward matrix from `start` to EEE (path excludes `start`; the
 * reference appends endCell after the walk, so EEE appears twice at the end). */
int hist_trace_best_forward(hist_pair* pair, hist_cell start, hist_cell* path, int64_t capacity, int64_t* n);

/* BackwardMatrix::cellsAbovePostProbThreshold (forward.cpp:1302-1319): all
 * (cell,state) with log posterior F+B-lpEnd >= log(min_post_prob), sorted by
 * decreasing posterior (the priority_queue pop order), ties by descending
 * (xpos, ypos, state).  Requires both fills and min_post_prob > 0. */
int hist_posterior_cells(hist_pair* pair, double min_post_prob, hist_cell* cells, double* log_post,
                         int64_t capacity, int64_t* n);

/* BackwardMatrix::getCounts (forward.cpp:1183-1214), numeric part.  Requires both fills.
 *   waits[6]  = { l.t, r.t, l.insWait, l.delWait, r.insWait, r.delWait } of the PairHMM's two
 *               ProbModels (model.cpp:374-391), the constants transitionEigenCounts adds
 *               (forward.cpp:579-652);
 *   indel[6]  = posterior-weighted ins, del, insExt, delExt, insTime, delTime summed over every
 *               (cell, state, incoming transition); profile transitions are taken to carry no
 *               counts of their own;
 *   w_x[2][x.n_states] = posterior weight of the alignment column of each x state, [0] without
 *               and [1] with a wildcard parent row (getAlignmentColumn, forward.cpp:940-973: IIW and
 *               null-state cells / IMD cells); w_y[2][y.n_states] likewise (IMI / IDM);
 *   cells, cell_w = the IMM cells whose two states both emit, in (xpos, ypos) order, with
 *               their posterior weights: one column (x column + y column + wildcard parent) each.
 * The caller appends those columns, weighted, to a hist_sumprod_columns batch; that is what
 * accumulateCachedEigenCounts does one column at a time (forward.cpp:903-938).  The
 * reference's per-row insert-column cache keeps an EMPTY column for a row whose first
 * envelope cell is on ypos == 0 (forward.cpp:905-918 with :942); w_x[0] / w_y[0] are zero for
 * such rows / columns so the counts stay identical to the reference's.
 * Returns HIST_ERR_CAPACITY (with *n_cells set) when `capacity` is too small. */
int hist_fb_counts(hist_pair* pair, const double* waits, double* indel, double* w_x, double* w_y,
                   hist_cell* cells, double* cell_w, int64_t capacity, int64_t* n_cells);

/* ------------------------------------------------------------------ makeProfile
 * Numeric core of ForwardMatrix::makeProfile (forward.cpp:686-843): given the
 * sorted cell set (must contain SSS=(0,0,IMM) and EEE), decide retained cells
 * (forward.cpp:703-709), run the reverse-order elimination
 * (forward.cpp:751-807: lpPath by log-sum-exp, lpBestAlignPath by max) and
 * gather the retained states' absorb tables (forward.cpp:714-727).
 * Outputs (sizes queried with the *_capacity arguments; HIST_ERR_CAPACITY
 * reports the needed sizes in n_states/n_trans):
 *   state_cell[n_states]              retained cells in CellCoords order
 *   state_lp_absorb[n_states][C][A]   -inf rows for non-absorbing states
 *   state_fwd_lp[n_states]            meta "fwdLogProb"
 *   trans_src/dest[n_trans]           indices into state_cell (before addReadyStates)
 *   trans_lp[n_trans], trans_best_lp[n_trans]
 *   trans_best_via_off[n_trans+1], trans_best_via[]  eliminated cells along the
 *                                     best path of each effective transition
 *                                     (for host-side alignPath assembly) */
typedef struct {
  int32_t n_states, n_trans;
  int64_t n_via;
  hist_cell* state_cell;
  double* state_lp_absorb;
  double* state_fwd_lp;
  int32_t* trans_src;
  int32_t* trans_dest;
  double* trans_lp;
  double* trans_best_lp;
  int64_t* trans_best_via_off;
  hist_cell* trans_best_via;
  int32_t state_capacity, trans_capacity;
  int64_t via_capacity;
} hist_elim_result;
int hist_profile_eliminate(hist_pair* pair, const hist_cell* cells, int64_t n_cells, int32_t strategy,
                           hist_elim_result* result);

/* ------------------------------------------------------------------ reconstruction
 * The node loop of Reconstructor::reconstruct (recon.cpp:917-1052) one guide-tree LEVEL at a time: all sibling
 * merges whose children are complete are independent iterations of that loop (its only dependencies are
 * child -> parent).  One hist_recon_level call runs, for every merge of the level: the DPMatrix constructor and
 * Forward fill with the band-widening retry (recon.cpp:954-975), ForwardMatrix::sampleProfile / bestProfile
 * (forward.cpp:845-895) -> traces on device, ForwardMatrix::makeProfile (forward.cpp:686-843) -> elimination on
 * device + state/transition/alignPath/seqCoords bookkeeping on the library's host threads, and
 * Profile::addReadyStates (profile.cpp:268-319).  The parent profiles stay on the device (their lpAbsorb tables
 * are gathered from the pairs' leftMultiply results, never copied to the host) and are the inputs of the next
 * level.  The children of a merge are released once their parent is built.
 * Sampled profiles draw from ONE std::mt19937 stream in node order in the reference (recon.cpp:38, 1010):
 * pass the merges of such a run in node order (the draws of merge k+1 start where those of merge k end). */
typedef struct hist_recon hist_recon;

typedef struct {
  int32_t node;                      /* row index of the parent (the tree node being built) */
  int32_t left, right;               /* its children: tree.getChild(node,0/1) */
  int32_t closest_left, closest_right, closest;  /* Dataset::closestLeaf of left, right and node (recon.cpp:890-911) */
  const hist_pairhmm_desc* hmm;      /* PairHMM(ProbModel(model,t_left), ProbModel(model,t_right), insProb) (recon.cpp:946-948) */
} hist_merge_desc;

typedef struct {
  int32_t band;                      /* maxDistanceFromGuide (-band); < 0 or no guide set: no envelope */
  int32_t strategy;                  /* ProfilingStrategy bits of the non-root nodes (recon.cpp:925-930); Count* not supported */
  int32_t profile_samples;           /* -profsamples */
  int32_t max_profile_states;        /* maxProfileStates() (0 = unlimited) */
  double min_post_prob;              /* must be <= 0 here: posterior profiles use hist_posterior_cells */
  int32_t root_node;                 /* the tree's root, or -1: that merge uses bestAlignPath + bestProfile (recon.cpp:1008-1011) */
  uint32_t* mt_state;                /* [625] std::mt19937 image, advanced by the sampled traces; may be NULL when nothing is sampled */
} hist_recon_params;

/* wall-clock split of hist_recon_level, accumulated over calls (milliseconds) */
enum { HIST_RECON_ENV = 0, HIST_RECON_PREPARE, HIST_RECON_FILL, HIST_RECON_RETRY, HIST_RECON_TRACE, HIST_RECON_SELECT,
       HIST_RECON_ELIM, HIST_RECON_ASSEMBLE, HIST_RECON_GATHER, HIST_RECON_CLEANUP, HIST_RECON_NSTAGES };

int hist_recon_create(hist_ctx* ctx, int32_t n_nodes, const char* alphabet, int32_t components, hist_recon** out);
void hist_recon_destroy(hist_recon* rec);
/* Dataset::guide restricted to the leaves: bits[n_rows][n_cols] (non-zero = residue), row_node[r] = tree node of row r. */
int hist_recon_set_guide(hist_recon* rec, int32_t n_rows, int32_t n_cols, const int32_t* row_node, const uint8_t* bits);
/* prof[node] = Profile(components, alphabet, seq, node) for the leaves (recon.cpp:940-941), built on device. */
int hist_recon_add_leaves(hist_recon* rec, int32_t n, const int32_t* nodes, const char* const* residues, const int32_t* lengths);
int hist_recon_level(hist_recon* rec, int32_t n, const hist_merge_desc* merges, const hist_recon_params* params, double* lp_end);
/* makeAlignment(path) as a dense 0/1 matrix bits[n_nodes][*n_cols] (forward->bestAlignPath() of the root, recon.cpp:1009,
 * 1040); HIST_ERR_CAPACITY with *n_cols set when `capacity` (bytes) is too small. */
int hist_recon_alignment(hist_recon* rec, int32_t* n_cols, uint8_t* bits, int64_t capacity);
int hist_recon_stage_ms(hist_recon* rec, double* ms /* [HIST_RECON_NSTAGES] */, int reset);
/* The device profile of a built node that has not been consumed by its parent yet (borrowed; NULL otherwise). */
hist_profile* hist_recon_profile(hist_recon* rec, int32_t node);

/* Sibling merges over several GPUs (one process and one hist_recon per GPU): a finished profile moves to the rank
 * that builds its parent as a host blob (integer structure, transition weights, coordinates, path ids) plus its
 * [n][C][A] lpAbsorb table, which stays in DEVICE memory on both sides: export copies it device-to-device into the
 * caller's device buffer dev_out (*dev_bytes long; call with host_buf == NULL first to learn both sizes), the caller
 * moves those bytes (ncclSend/Recv or a peer copy) and passes the received device buffer to import, which copies it
 * device-to-device again.  release drops the sender's copy.  The alignment-path stores that hist_recon_alignment expands
 * are gathered once, at the end, with export_stores / import_stores (host blobs). */
int hist_recon_export_node(hist_recon* rec, int32_t node, void* host_buf, int64_t capacity, int64_t* bytes,
                           void* dev_out, int64_t* dev_bytes);
int hist_recon_import_node(hist_recon* rec, int32_t node, const void* host_buf, int64_t bytes, const void* dev_table);
int hist_recon_release_node(hist_recon* rec, int32_t node);
int hist_recon_export_stores(hist_recon* rec, void* buf, int64_t capacity, int64_t* bytes);
int hist_recon_import_stores(hist_recon* rec, const void* buf, int64_t bytes);

/* Read a device profile back in the hist_profile_desc layout (any pointer may be NULL). */
int hist_profile_get_sizes(const hist_profile* p, int32_t* n_states, int32_t* n_trans, int32_t* alph_size, int32_t* components);
int hist_profile_fetch(hist_profile* p, uint8_t* is_emit, int32_t* trans_src, int32_t* trans_dest, double* trans_lp,
                       double* lp_absorb, int32_t* closest_leaf_pos);

/* ------------------------------------------------------------------ P3
 * Profile::calcSumPathAbsorbProbs (profile.cpp:112-131): 1-D Forward over the profile with the insertion
 * distribution as emission (the "cumLogProb" tag of t/testforward and the log line of recon.cpp:1022).
 * log_cpt_weight [C], log_ins_prob [C][A]; cum (optional) receives lpCumAbs per state [n_states];
 * *result = lpCumAbs of the END state. */
int hist_profile_sum_path_absorb(hist_profile* profile, const double* log_cpt_weight, const double* log_ins_prob,
                                 double* cum, double* result);

/* The same computation for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it).  This entry point has no counterpart in the reference.
 * It calls no cudaMalloc / cudaFree / cudaMallocAsync and no allocating library
 * routine: inputs are staged in, and the in-lists, lpAbs and lpCum are built in, ONE built-in device array of
 * HIST_NOALLOC_SCRATCH_BYTES that every GPU loading the library holds until exit.  A profile whose scratch
 * layout (about 20 B per transition + 25 B per state + 8 B per (component, residue)) exceeds that size is
 * rejected with HIST_ERR_CAPACITY and a message naming the sizes -- use hist_profile_sum_path_absorb for it.
 * All calls on a GPU share the array, which keeps the last call's contents (the kernel clears what must start
 * at zero): calls that overlap on one GPU -- two contexts, two host threads -- race; serialise them.
 * Arguments and results as hist_profile_sum_path_absorb. */
#define HIST_NOALLOC_SCRATCH_BYTES (4u << 20)
int hist_profile_sum_path_absorb_noalloc(hist_profile* profile, const double* log_cpt_weight, const double* log_ins_prob,
                                         double* cum, double* result);
/* Size of the built-in scratch array (HIST_NOALLOC_SCRATCH_BYTES of the loaded library). */
int64_t hist_noalloc_scratch_bytes(void);

/* ------------------------------------------------------------------ guide-alignment stage
 * QuickAlignMatrix (quickalign.h:8-76, quickalign.cpp:7-214): three-state Viterbi alignment of two sequences inside
 * the diagonals of a DiagonalEnvelope (diagenv.h:38-155) -- the pairwise step AlignGraph runs for every selected pair
 * of sequences (span.cpp:57-158) before the progressive reconstruction.  The constructor does the fill; alignPath()
 * the traceback.  Scores are bit-exact FP64 and paths identical to the reference's (ties included).
 *
 * hist_quickalign_scores: what the QuickAlignMatrix constructor derives from (model, time), quickalign.cpp:22-50:
 *   submat [A][A] = logSubProb[i][j] - logInsProb[j]; the public members noGap, gapOpen, gapExtend, m2m .. d2m.
 * hist_quickalign_pair: tokens as FastSeq::unvalidatedTokens gives them (< 0 or >= A: not in the alphabet) and
 *   DiagonalEnvelope::diagonals (ascending values of i - j, within [1 - y_len, x_len - 1]; initFull or initSparse);
 *   diagonals == NULL with n_diagonals = x_len + y_len - 1 is the full envelope without the list.
 * Outputs per pair k: result[k] = mx.result (-inf: no alignment inside the envelope), x_end / y_end = mx.xEnd / yEnd,
 *   and, if paths != NULL, the columns of alignPath() from paths[path_off[k]] on (capacity x_len + y_len, path_len[k]
 *   columns): 3 = both rows, 1 = row 0 (x) only, 2 = row 1 (y) only; path_len[k] = 0 when result is -inf. */
typedef struct {
  int32_t alph_size;
  const double* submat;
  double no_gap, gap_open, gap_extend;
  double m2m, m2i, m2d, i2i, i2m, i2d, d2d, d2m;
} hist_quickalign_scores;
typedef struct {
  int32_t x_len, y_len;
  const int8_t* x_tok;
  const int8_t* y_tok;
  int32_t n_diagonals;
  const int32_t* diagonals;
} hist_quickalign_pair;
int hist_quickalign_batch(hist_ctx* ctx, int32_t n, const hist_quickalign_pair* pairs, const hist_quickalign_scores* scores,
                          double* result, int32_t* x_end, int32_t* y_end, uint8_t* paths, const int64_t* path_off,
                          int32_t* path_len);

/* AlignGraph (span.h:9-52, span.cpp:35-158) as `historian recon -seqs` runs it (recon.cpp:806-821): trial edges, one
 * DiagonalEnvelope (diagenv.cpp:93-190) and one QuickAlignMatrix per edge (on the device, batched), the spanning tree of
 * the best-scoring edges grown from sequence 0, and alignPathMerge of its n - 1 pairwise alignments.
 *
 * hist_guide_params: all_pairs = AlignGraph's all-vs-all constructor (`-allspan`, guideAlignTryAllPairs, recon.cpp:41,
 *   811-812), else the sparse random graph of ceil(n log2 n) edges, continued until connected (span.cpp:66-91), drawn from
 *   std::mt19937 seeded with `seed` (Reconstructor::rndSeed, recon.cpp:38, 745-748; 5489 by default).  The rest is
 *   DiagEnvParams (diagenv.h:24-35): sparse = 0 is -kmatchoff; kmer_threshold >= 0 is -kmatchn; kmer_threshold < 0 takes the
 *   full envelope when xLen * yLen * cell_size < max_size (the reference's default, with max_size = the machine's memory
 *   and cell_size = ForwardMatrix::cellSize() = 40) and otherwise the largest set of seeded diagonals that fits max_size.
 * tokens: FastSeq::unvalidatedTokens of every sequence, concatenated; tok_off [n_seqs + 1].
 * Outputs: residue_col [tok_off[n_seqs]] = the guide alignment's column of every residue (row s of Alignment::path has its
 *   k-th `true` at column residue_col[tok_off[s] + k]), *n_cols its number of columns; optionally the trial edges in the
 *   order they were drawn (*n_edges of them; the first edge_capacity written to edge_rows [2 * e], edge_lp [e] = mx.end) and
 *   the indices of the spanning tree's edges in joining order (mst_edges [n_seqs - 1]). */
typedef struct {
  int32_t all_pairs;
  uint32_t seed;
  int32_t sparse;
  int32_t kmer_len;
  int32_t kmer_threshold;
  int32_t band_size;
  int32_t cell_size;
  int32_t reserved;
  int64_t max_size;
} hist_guide_params;
int hist_guide_align(hist_ctx* ctx, int32_t n_seqs, const int8_t* tokens, const int64_t* tok_off, const hist_quickalign_scores* scores,
                     const hist_guide_params* params, int32_t* residue_col, int32_t* n_cols, int32_t* n_edges, int32_t* edge_rows,
                     double* edge_lp, int64_t edge_capacity, int32_t* mst_edges);

/* DiagonalEnvelope::initFull / initSparse (diagenv.cpp:93-190) of one pair, host only: the ascending diagonals i - j that
 * hist_guide_align hands to the pairwise alignment (params: the DiagEnvParams part of hist_guide_params).  capacity
 * x_len + y_len - 1 always suffices.  ctx may be NULL. */
int hist_diagonal_envelope(hist_ctx* ctx, const int8_t* x_tok, int32_t x_len, const int8_t* y_tok, int32_t y_len, int32_t alph_size,
                           const hist_guide_params* params, int32_t* diagonals, int32_t capacity, int32_t* n_diagonals);

/* alignPathMerge (alignpath.cpp:153-207) on the host: merges alignments that share rows into one.  Alignment k has
 * n_rows[k] rows (their global indices consecutive in `rows`) and n_cols[k] columns, its 0/1 matrix row-major and
 * consecutive in `bits`; row r of the result has row_off[r + 1] - row_off[r] residues (every alignment mentioning the row
 * must agree, alignpath.cpp:113).  residue_col [row_off[n_total_rows]]: output column of every residue (-1 for rows no
 * alignment mentions); *n_out_cols: number of columns.  HIST_ERR_STATE for inconsistent alignments (the reference aborts).
 * ctx may be NULL (no device work; used for the error text only). */
int hist_align_path_merge(hist_ctx* ctx, int32_t n_alignments, const int32_t* n_rows, const int32_t* n_cols, const int32_t* rows,
                          const uint8_t* bits, int32_t n_total_rows, const int64_t* row_off, int32_t* residue_col, int32_t* n_out_cols);

/* ------------------------------------------------------------------ sum-product
 * SumProduct / AlignColSumProduct (sumprod.h:32-104, sumprod.cpp).
 * The model is passed in eigen-decomposed form (EigenModel, model.cpp:1135-1199):
 * P(t) = V exp(Lambda t) V^-1 clamped to [0,1] (model.cpp:1267-1283) and
 * eigenSubCount(t) (model.cpp:1329-1350) are built on device per branch. */
typedef struct {
  int32_t alph_size;
  int32_t components;
  const char* alphabet;              /* [A] residue characters (token order) */
  const double* cpt_weight;          /* [C] */
  const double* ins_prob;            /* [C][A] model.insProb */
  const double* sub_rate;            /* [C][A][A] rate matrices (used for sub counts' R_ij) */
  const double* eval_re;             /* [C][A] eigenvalues */
  const double* eval_im;
  const double* evec_re;             /* [C][A][A] right eigenvectors V[a][k] */
  const double* evec_im;
  const double* evec_inv_re;         /* [C][A][A] V^-1[k][b] */
  const double* evec_inv_im;
  const double* branch_sub;          /* optional [C][N][A][A]: branch matrices to use verbatim
                                        (e.g. the reference's expm results); NULL = build from eigen */
} hist_model_desc;

typedef struct {
  int32_t n_nodes;                   /* nodes indexed in postorder, children < parent, root last (tree.cpp:141) */
  const int32_t* parent;             /* [N] parent index, -1 for the root */
  const double* branch_length;       /* [N] length of the branch above each node */
} hist_tree_desc;

int hist_sumprod_create(hist_ctx* ctx, const hist_model_desc* model, const hist_tree_desc* tree,
                        hist_sumprod** out);
void hist_sumprod_destroy(hist_sumprod* sp);

/* Flat layout of the count buffer: 7 indel slots (unused here, kept so the
 * buffer is the whole EigenCounts object: ins, del, insExt, delExt, insTime,
 * delTime, lp), then rootCount[C][A], then eigenCount[C][A][A] as (re,im) pairs. */
size_t hist_sumprod_counts_size(const hist_sumprod* sp);

enum {
  HIST_SP_COUNTS = 1,      /* accumulateEigenCounts + accumulateRootCounts (sumprod.cpp:264-271, 294-372) */
  HIST_SP_MAP_STATES = 2,  /* maxPostState for wildcard rows (sumprod.cpp:259, 398-412) */
  HIST_SP_NODE_POST = 4    /* logNodePostProb for every ungapped row (sumprod.cpp:208-218) */
};

/* Processes n_cols independent alignment columns (model.cpp:900-915 loop body):
 * initColumn, fillUp, fillDown, then whatever `flags` asks for.
 *   columns      [n_cols][N] bytes: residue char, '-' or '.' for gap, '*' wildcard
 *                (any char outside the alphabet is a wildcard, sumprod.cpp:67)
 *   weights      [n_cols] or NULL (=1): weight argument of accumulateEigenCounts
 *   col_loglike  [n_cols] columnLogLikelihood()
 *   counts       flat count buffer, ACCUMULATED into (+=); may be a DEVICE pointer
 *                (then it stays on device for an NCCL all-reduce by the caller)
 *   map_state    [n_cols][N] bytes: input char with wildcards replaced by the MAP residue
 *   node_post    [n_cols][N][A] log posteriors (-inf rows for gaps) */
int hist_sumprod_columns(hist_sumprod* sp, const char* columns, const double* weights, int64_t n_cols,
                         int32_t flags, double* col_loglike, double* counts, char* map_state,
                         double* node_post);

/* SumProduct::logNodeExcludedPostProb (sumprod.cpp:219-250), for every column and every query
 * (nodes[q], excludes[q]): log of cptWeight[c] x the messages node receives from all its tree neighbours except `exclude`
 * (a child, the parent, or -1 for none), restricted to the node's own residue -- the conditional position weight
 * matrices of TreeAlignFuncs::getConditionalPWMs (sampler.cpp:356-372), whose restricted fillUp / fillDown orders yield
 * the same messages as the full passes.  normalize != 0 (getConditionalPWMs' default): every [C][A] block minus its
 * log-sum-exp.  out [n_queries][n_cols][C][A]; rows of -inf in columns where the node is gapped. */
int hist_sumprod_excluded_posteriors(hist_sumprod* sp, const char* columns, int64_t n_cols, int32_t n_queries, const int32_t* nodes,
                                     const int32_t* excludes, int32_t normalize, double* out);

/* ------------------------------------------------------------------ refinement stage
 * Refiner::BranchMatrix (refiner.cpp:10-106) over TreeAlignFuncs::BranchMatrixBase (sampler.h:183-210, sampler.cpp:1005-1031):
 * Viterbi alignment of a parent's and a child's conditional position weight matrices (hist_sumprod_excluded_posteriors:
 * parent excluding the child, child excluding the parent) under the branch's ProbModel, and its traceback best().
 *   model: logProbModel.logSubProb [C][A][A] (parent residue i -> child residue j), logInsProb [C][A], logCptWeight [C] of
 *          ProbModel (model, max (Tree::minBranchLength, dist)), and lp_trans[src][dest] = log transProb (src, dest) for
 *          src in Match, Insert, Delete and dest in Match, Insert, Delete, End (model.cpp:400-447).
 *   x_pwm [x_len][C][A], y_pwm [y_len][C][A]: log weights of the parent's / child's residues.
 *   max_distance, x_key [x_len + 1], y_key [y_len + 1]: the envelope, cell (x, y) is inside iff it lies on the border of
 *          the matrix or |x_key[x] - y_key[y]| <= max_distance (sampler.h:146-149 with GuideAlignmentEnvelope::inRange:
 *          x_key[x] = cumulativeMatches[row1PosToCol[xEnvPos[x]]]); max_distance < 0: no envelope, keys may be NULL.
 * Outputs: *lp_end (the Viterbi log-likelihood), path [capacity x_len + y_len] / *path_len: the columns of best(),
 *          3 = both rows, 1 = parent only, 2 = child only. */
typedef struct {
  int32_t alph_size;
  int32_t components;
  const double* log_sub;
  const double* log_ins;
  const double* log_cpt_weight;
  double lp_trans[3][4];
} hist_branch_model;
int hist_branch_align(hist_ctx* ctx, const hist_branch_model* model, int32_t x_len, const double* x_pwm, int32_t y_len, const double* y_pwm,
                      int32_t max_distance, const int32_t* x_key, const int32_t* y_key, double* lp_end, uint8_t* path, int32_t* path_len);

/* Fetch E/F/G messages and their log scale factors of ONE column for parity tests
 * (SumProductStorage, sumprod.h:14-30): each [C][N][A] resp. [C][N]. */
int hist_sumprod_debug_column(hist_sumprod* sp, const char* column, double* E, double* F, double* G,
                              double* logE, double* logF, double* logG, double* col_loglike);

/* EigenModel::getSubCounts (model.cpp:1352-1377): eigenbasis counts -> (i,j)
 * substitution counts and wait times, sub_count[C][A][A] (diagonal = wait time). */
int hist_sumprod_transform_counts(hist_sumprod* sp, const double* counts, double* root_count,
                                  double* sub_count);

/* The per-branch tables built on device by hist_sumprod_create (SumProduct ctor,
 * sumprod.cpp:21-50): branch_sub [C][N][A][A] = branchSubProb, branch_esc
 * [C][N][A][A] (re,im) = branchEigenSubCount (model.cpp:1329-1350).  Either may
 * be NULL.  For parity tests. */
int hist_sumprod_branch_tables(hist_sumprod* sp, double* branch_sub, double* branch_esc);

/* ------------------------------------------------------------------ distance-matrix stage
 * RateModel::distanceMatrix (model.cpp:533-553): dist[i][j] = mlDistance of rows i and j of a gapped alignment --
 * aligned residue-pair counts (model.cpp:506-531), then DistanceMatrixParams::tML (model.cpp:583-655): Jukes-Cantor
 * first guess, bracket scan, golden-section steps to 1 % (max_iterations <= 0: the Jukes-Cantor estimate itself).
 * model: alphabet, cpt_weight and the eigen-decomposition are used (P_t from eigenvalues, where the reference calls
 * GSL's matrix exponential: agreement ~1e-13 in P, so results equal the reference's unless two trial likelihoods tie
 * to that precision, and then differ by less than the search's 1 %).  expected_sub_rate =
 * RateModel::expectedSubstitutionRate (model.cpp:336-347).  rows: [n_rows][n_cols] characters, gaps '-' '.', wildcard '*'.
 * dist: [n_rows][n_rows], symmetric, zero diagonal. */
int hist_distance_matrix(hist_ctx* ctx, const hist_model_desc* model, double expected_sub_rate, int32_t n_rows, int32_t n_cols,
                         const char* rows, int32_t max_iterations, double* dist);

/* Guide tree from a distance matrix: Tree::buildByUPGMA (tree.cpp:362-454; method 0, the default of `historian recon`) or
 * Tree::buildByNeighborJoining (tree.cpp:240-352; method 1).  dist: [n][n] symmetric; min_branch_length:
 * Tree::minBranchLength (TREE_MIN_BRANCH_LEN = 1e-9, tree.h:24).  Nodes 0 .. n-1 are the rows of the matrix, node n + r is
 * created by the r-th join, the root is node 2n - 2.  child: [2n-1][2] (-1, -1 for leaves), branch_length: [2n-1] length
 * of the branch above each node (0 for the root).  Joins and lengths are bit-identical to the reference's before it
 * serialises the tree with 6 significant digits and parses it back (tree.cpp:347-350). */
int hist_build_tree(hist_ctx* ctx, int32_t n, const double* dist, int32_t method, double min_branch_length, int32_t* child,
                    double* branch_length);

#ifdef __cplusplus
}
#endif
#endif /* HISTORIAN_B200_H */
