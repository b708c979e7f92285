// historian_b200: numeric core of ForwardMatrix::makeProfile on device.
//
// Replaces reference forward.cpp:686-843 (with isAbsorbing :470-474, eliminatedLogProbInsert
// :503-526 and the DP-graph enumerator :326-398): from a sorted set of DP cells, decide which cells
// become states of the parent profile, sum out the others in reverse order (lpPath by log-sum-exp,
// lpBestAlignPath by max, remembering the eliminated cells along the best path) and gather the
// states' absorb tables.  The alignPath / seqCoords / name assembly and addReadyStates
// (profile.cpp:268-319) are host bookkeeping on top of this result.
//
// Kernels: (1) one thread per selected cell lists its in-set predecessors (binary search in the
// sorted cell list) -> CSR of the selected sub-DAG + out-degree counts; (2) retained flags and
// state numbering (CellCoords order, forward.cpp:703-732); (3) the reverse-order elimination, a
// dependent chain over the sub-DAG, run by one thread over pooled sorted lists (the sub-DAG of a
// merge is a few thousand cells; batching across merges is the parallel axis); (4) per-state
// table gather and per-transition via-path expansion, fully parallel.
#include <thrust/execution_policy.h>
#include <thrust/scan.h>

#include <algorithm>
#include <cmath>
#include <vector>

#include "hist_dpgraph.cuh"

namespace hb {

struct ElimDev {
  const PairDev* P;
  const DCell* cells;   // [n] sorted
  int n;
  int strategy;
  // sub-DAG
  int* in_cnt;          // [n+1] -> offsets after the scan
  int* out_cnt;         // [n]
  int* e_src;           // [n_edges] index of the source cell
  double* e_lp;         // [n_edges] transition log-weight without emission
  // states
  int* retained;        // [n+1] flag -> state index after the scan
  // effective-transition pool (sorted singly linked list per cell, keyed by destination state)
  int* head;            // [n]
  int pool_cap;
  int* p_k; int* p_next; int* p_via_cell; int* p_via_next;
  double* p_path; double* p_best;
  // results
  int* counts;          // [4]: n_states, n_trans, n_via (low/high words unused), pool overflow flag
  long long* n_via;     // [1]
  int* t_src; int* t_dst; int* t_entry;   // [pool_cap]
  double* t_lp; double* t_best;
  long long* t_via_off; // [pool_cap + 1]
};

__device__ __forceinline__ int find_cell(const DCell* cells, int n, const DCell& c) {
  int lo = 0, hi = n;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (dcell_less(cells[mid], c)) lo = mid + 1; else hi = mid;
  }
  return (lo < n && cells[lo].x == c.x && cells[lo].y == c.y && cells[lo].s == c.s) ? lo : -1;
}

// pass 0: count in-set sources of every cell and out-degrees; pass 1: fill the CSR
__global__ void elim_edges_kernel(ElimDev E, int pass) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= E.n) return;
  const PairDev& P = *E.P;
  const DCell d = E.cells[c];
  const SrcEnum e = source_enum(P, d, nullptr);
  int cnt = 0;
  const int base = pass ? E.in_cnt[c] : 0;
  for (int q = 0; q < e.n; ++q) {
    DCell s;
    double lp;
    source_get(P, P.Wlog, d, e, q, s, lp);
    const int si = find_cell(E.cells, E.n, s);
    if (si < 0) continue;
    if (pass) {
      E.e_src[base + cnt] = si;
      E.e_lp[base + cnt] = lp;
    } else {
      atomicAdd(E.out_cnt + si, 1);
    }
    ++cnt;
  }
  if (!pass) E.in_cnt[c] = cnt;
}

// isAbsorbing (forward.cpp:470-474)
__device__ __forceinline__ bool is_absorbing(const PairDev& P, const DCell& c) {
  const bool xe = P.x.flags[c.x] & F_EMIT, ye = P.y.flags[c.y] & F_EMIT;
  return (c.s == HIST_IMM && xe && ye) || (c.s == HIST_IMD && xe) || (c.s == HIST_IDM && ye);
}

__global__ void elim_retain_kernel(ElimDev E) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= E.n) return;
  const PairDev& P = *E.P;
  const DCell d = E.cells[c];
  const bool isStart = d.x == 0 && d.y == 0 && d.s == HIST_IMM;
  const bool isEnd = d.x == P.x.n - 1 && d.y == P.y.n - 1 && d.s == HIST_EEE;
  // forward.cpp:703-709
  E.retained[c] = (is_absorbing(P, d) || isStart || isEnd || E.out_cnt[c] > 1 || (E.strategy & HIST_KEEP_GAPS_OPEN) ||
                   !(E.strategy & HIST_COLLAPSE_CHAINS)) ? 1 : 0;
  E.head[c] = -1;
}

// find-or-insert the entry of destination state k in the sorted list of cell s; -1 on pool overflow
__device__ int pool_entry(ElimDev& E, int& pool_n, int s, int k) {
  int prev = -1, cur = E.head[s];
  while (cur >= 0 && E.p_k[cur] < k) { prev = cur; cur = E.p_next[cur]; }
  if (cur >= 0 && E.p_k[cur] == k) return cur;
  if (pool_n >= E.pool_cap) return -1;
  const int e = pool_n++;
  E.p_k[e] = k;
  E.p_path[e] = -INFINITY;
  E.p_best[e] = -INFINITY;
  E.p_via_cell[e] = -1;
  E.p_via_next[e] = -1;
  E.p_next[e] = cur;
  if (prev < 0) E.head[s] = e; else E.p_next[prev] = e;
  return e;
}

// forward.cpp:751-807, then the transition list of forward.cpp:810-831.  `retained` holds state indices
// (exclusive scan) with retained[n] = number of states; is_ret = retained[c+1] != retained[c].
__global__ void elim_reduce_kernel(ElimDev E) {
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  const PairDev& P = *E.P;
  int pool_n = 0;
  bool overflow = false;
  for (int c = E.n - 1; c >= 0 && !overflow; --c) {
    const DCell d = E.cells[c];
    // eliminatedLogProbInsert (forward.cpp:503-526)
    double ins = 0;
    if (d.s == HIST_IIW && (P.x.flags[d.x] & F_EMIT)) ins = P.x.insl[d.x];
    else if (d.s == HIST_IMI && (P.y.flags[d.y] & F_EMIT)) ins = P.y.insl[d.y];
    const bool is_ret = E.retained[c + 1] != E.retained[c];
    for (int q = E.in_cnt[c]; q < E.in_cnt[c + 1] && !overflow; ++q) {
      const int s = E.e_src[q];
      const double lp = E.e_lp[q] + ins;
      if (is_ret) {
        const int e = pool_entry(E, pool_n, s, E.retained[c]);
        if (e < 0) { overflow = true; break; }
        E.p_path[e] = lp;
        E.p_best[e] = lp;
        E.p_via_cell[e] = -1;
        E.p_via_next[e] = -1;
      } else {
        for (int ce = E.head[c]; ce >= 0; ce = E.p_next[ce]) {
          const int e = pool_entry(E, pool_n, s, E.p_k[ce]);
          if (e < 0) { overflow = true; break; }
          E.p_path[e] = lse2(E.p_path[e], lp + E.p_path[ce]);
          const double best = lp + E.p_best[ce];
          if (best > E.p_best[e]) {
            E.p_best[e] = best;
            E.p_via_cell[e] = c;
            E.p_via_next[e] = ce;
          }
        }
      }
    }
  }
  // transitions of the retained cells, ascending destination state, and their via-path lengths
  int nt = 0;
  long long nv = 0;
  if (!overflow)
    for (int c = 0; c < E.n; ++c) {
      if (E.retained[c + 1] == E.retained[c]) continue;
      for (int e = E.head[c]; e >= 0; e = E.p_next[e]) {
        E.t_src[nt] = E.retained[c];
        E.t_dst[nt] = E.p_k[e];
        E.t_lp[nt] = E.p_path[e];
        E.t_best[nt] = E.p_best[e];
        E.t_entry[nt] = e;
        E.t_via_off[nt] = nv;
        for (int v = e; v >= 0 && E.p_via_cell[v] >= 0; v = E.p_via_next[v]) ++nv;
        ++nt;
      }
    }
  E.t_via_off[nt] = nv;
  E.counts[0] = E.retained[E.n];
  E.counts[1] = nt;
  E.counts[3] = overflow ? 1 : 0;
  *E.n_via = nv;
}

// state cells, absorb tables (forward.cpp:713-727) and fwdLogProb (forward.cpp:731)
__global__ void elim_states_kernel(ElimDev E, double lp_end, hist_cell* st_cell, double* st_abs, double* st_fwd) {
  const int c = blockIdx.x;
  if (E.retained[c + 1] == E.retained[c]) return;
  const PairDev& P = *E.P;
  const int k = E.retained[c];
  const DCell d = E.cells[c];
  const bool ab = is_absorbing(P, d);
  for (int q = threadIdx.x; q < P.K; q += blockDim.x) {
    double v = -INFINITY;
    if (ab) {
      if (d.s == HIST_IMM) v = P.x.sub[(size_t)d.x * P.K + q] + P.y.sub[(size_t)d.y * P.K + q];
      else if (d.s == HIST_IMD) v = P.x.sub[(size_t)d.x * P.K + q];
      else v = P.y.sub[(size_t)d.y * P.K + q];
    }
    st_abs[(size_t)k * P.K + q] = v;
  }
  if (threadIdx.x == 0) {
    st_cell[k] = {d.x, d.y, d.s};
    st_fwd[k] = d.s == HIST_EEE ? lp_end : dp_cell(P, false, d.x, d.y, d.s);
  }
}

__global__ void elim_via_kernel(ElimDev E, int n_trans, hist_cell* via) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n_trans) return;
  long long o = E.t_via_off[t];
  for (int v = E.t_entry[t]; v >= 0 && E.p_via_cell[v] >= 0; v = E.p_via_next[v]) {
    const DCell c = E.cells[E.p_via_cell[v]];
    via[o++] = {c.x, c.y, c.s};
  }
}

}  // namespace hb

using namespace hb;

extern "C" int hist_profile_eliminate(hist_pair* pair, const hist_cell* cells, int64_t n_cells, int32_t strategy, hist_elim_result* r) {
  if (!pair || !cells || !r || n_cells < 2) return HIST_ERR_ARG;
  hist_ctx* ctx = pair->ctx;
  if (!pair->fwd_done) return set_err(ctx, HIST_ERR_STATE, "hist_profile_eliminate: the Forward matrix has not been filled");
  if (!pair->x->sorted_in || !pair->y->sorted_in)
    return set_err(ctx, HIST_ERR_ARG, "hist_profile_eliminate: profile transitions must be sorted by (source, destination) without repeats");
  if (n_cells > (1ll << 28)) return set_err(ctx, HIST_ERR_ARG, "hist_profile_eliminate: too many cells");
  const int n = (int)n_cells, nx = pair->info.x_size, ny = pair->info.y_size, K = pair->dev.K;
  for (int k = 0; k < n; ++k) {
    const hist_cell& c = cells[k];
    const bool isEnd = c.xpos == nx - 1 && c.ypos == ny - 1 && c.state == HIST_EEE;
    if (!isEnd && (c.xpos < 0 || c.ypos < 0 || c.xpos > nx - 2 || c.ypos > ny - 2 || c.state < 0 || c.state > 4))
      return set_err(ctx, HIST_ERR_ARG, "hist_profile_eliminate: cell %d (%d,%d,%d) outside the matrix", k, c.xpos, c.ypos, c.state);
    if (k) {
      const hist_cell& p = cells[k - 1];
      const bool less = p.xpos != c.xpos ? p.xpos < c.xpos : p.ypos != c.ypos ? p.ypos < c.ypos : p.state < c.state;
      if (!less) return set_err(ctx, HIST_ERR_ARG, "hist_profile_eliminate: cells must be strictly ascending in (xpos, ypos, state)");
    }
  }
  if (!(cells[0].xpos == 0 && cells[0].ypos == 0 && cells[0].state == HIST_IMM) ||
      !(cells[n - 1].xpos == nx - 1 && cells[n - 1].ypos == ny - 1 && cells[n - 1].state == HIST_EEE))
    return set_err(ctx, HIST_ERR_ARG, "hist_profile_eliminate: the cell set must contain the start cell (0,0,IMM) and the end cell");
  cudaSetDevice(ctx->device);

  int rc = HIST_OK;
  for (int pool_cap = 8 * n + 1024;; pool_cap *= 4) {
    // ---- scratch
    size_t off = 0;
    auto take = [&](size_t bytes) { off = (off + 15) / 16 * 16; const size_t o = off; off += bytes; return o; };
    const size_t o_cells = take((size_t)n * sizeof(DCell)), o_in = take((size_t)(n + 1) * 4), o_out = take((size_t)n * 4),
                 o_ret = take((size_t)(n + 1) * 4), o_head = take((size_t)n * 4), o_pk = take((size_t)pool_cap * 4),
                 o_pn = take((size_t)pool_cap * 4), o_pvc = take((size_t)pool_cap * 4), o_pvn = take((size_t)pool_cap * 4),
                 o_pp = take((size_t)pool_cap * 8), o_pb = take((size_t)pool_cap * 8), o_cnt = take(16), o_nv = take(8),
                 o_ts = take((size_t)pool_cap * 4), o_td = take((size_t)pool_cap * 4), o_te = take((size_t)pool_cap * 4),
                 o_tl = take((size_t)pool_cap * 8), o_tb = take((size_t)pool_cap * 8), o_tv = take((size_t)(pool_cap + 1) * 8);
    char* buf = nullptr;
    HB_CUDA(ctx, cudaMallocAsync((void**)&buf, off, ctx->stream));
    ElimDev E;
    E.P = pair->d_dev;
    E.cells = (const DCell*)(buf + o_cells);
    E.n = n;
    E.strategy = strategy;
    E.in_cnt = (int*)(buf + o_in); E.out_cnt = (int*)(buf + o_out);
    E.retained = (int*)(buf + o_ret); E.head = (int*)(buf + o_head);
    E.pool_cap = pool_cap;
    E.p_k = (int*)(buf + o_pk); E.p_next = (int*)(buf + o_pn); E.p_via_cell = (int*)(buf + o_pvc); E.p_via_next = (int*)(buf + o_pvn);
    E.p_path = (double*)(buf + o_pp); E.p_best = (double*)(buf + o_pb);
    E.counts = (int*)(buf + o_cnt); E.n_via = (long long*)(buf + o_nv);
    E.t_src = (int*)(buf + o_ts); E.t_dst = (int*)(buf + o_td); E.t_entry = (int*)(buf + o_te);
    E.t_lp = (double*)(buf + o_tl); E.t_best = (double*)(buf + o_tb); E.t_via_off = (long long*)(buf + o_tv);
    E.e_src = nullptr; E.e_lp = nullptr;
    auto cleanup = [&](char* extra) { if (extra) cudaFreeAsync(extra, ctx->stream); cudaFreeAsync(buf, ctx->stream); };
    static_assert(sizeof(DCell) == sizeof(hist_cell), "cell layouts must agree");
    rc = check_cuda(ctx, cudaMemcpyAsync(buf + o_cells, cells, (size_t)n * sizeof(DCell), cudaMemcpyHostToDevice, ctx->stream), "H2D cells");
    if (!rc) rc = check_cuda(ctx, cudaMemsetAsync(buf + o_in, 0, o_head - o_in, ctx->stream), "memset counts");
    if (rc) { cleanup(nullptr); return rc; }
    const int tb = 128, gb = (n + tb - 1) / tb;
    elim_edges_kernel<<<gb, tb, 0, ctx->stream>>>(E, 0);
    thrust::exclusive_scan(thrust::cuda::par.on(ctx->stream), E.in_cnt, E.in_cnt + n + 1, E.in_cnt);
    int n_edges = 0;
    rc = check_cuda(ctx, cudaMemcpyAsync(&n_edges, E.in_cnt + n, 4, cudaMemcpyDeviceToHost, ctx->stream), "D2H edge count");
    if (!rc) rc = check_cuda(ctx, cudaStreamSynchronize(ctx->stream), "sync");
    if (rc) { cleanup(nullptr); return rc; }
    char* ebuf = nullptr;
    const size_t eb = (size_t)std::max(n_edges, 1);
    rc = check_cuda(ctx, cudaMallocAsync((void**)&ebuf, eb * 16, ctx->stream), "cudaMallocAsync(edges)");
    if (rc) { cleanup(nullptr); return rc; }
    E.e_lp = (double*)ebuf;
    E.e_src = (int*)(ebuf + eb * 8);
    elim_edges_kernel<<<gb, tb, 0, ctx->stream>>>(E, 1);
    elim_retain_kernel<<<gb, tb, 0, ctx->stream>>>(E);
    thrust::exclusive_scan(thrust::cuda::par.on(ctx->stream), E.retained, E.retained + n + 1, E.retained);
    elim_reduce_kernel<<<1, 32, 0, ctx->stream>>>(E);
    ctx->launches += 6;
    int counts[4] = {0, 0, 0, 0};
    long long nvia = 0;
    rc = check_cuda(ctx, cudaGetLastError(), "elimination kernels");
    if (!rc) rc = check_cuda(ctx, cudaMemcpyAsync(counts, E.counts, 16, cudaMemcpyDeviceToHost, ctx->stream), "D2H counts");
    if (!rc) rc = check_cuda(ctx, cudaMemcpyAsync(&nvia, E.n_via, 8, cudaMemcpyDeviceToHost, ctx->stream), "D2H via count");
    if (!rc) rc = check_cuda(ctx, cudaStreamSynchronize(ctx->stream), "sync");
    if (rc) { cleanup(ebuf); return rc; }
    if (counts[3]) {  // pool too small: retry with a larger one
      cleanup(ebuf);
      if (pool_cap > (1 << 28)) return set_err(ctx, HIST_ERR_CAPACITY, "hist_profile_eliminate: effective-transition pool exhausted");
      continue;
    }
    const int ns = counts[0], nt = counts[1];
    r->n_states = ns;
    r->n_trans = nt;
    r->n_via = nvia;
    if (ns > r->state_capacity || nt > r->trans_capacity || nvia > r->via_capacity) {
      cleanup(ebuf);
      return set_err(ctx, HIST_ERR_CAPACITY, "hist_profile_eliminate: need %d states, %d transitions, %lld via cells", ns, nt, nvia);
    }
    // ---- results
    char* obuf = nullptr;
    const size_t b_cell = ((size_t)ns * sizeof(hist_cell) + 15) / 16 * 16, b_abs = (size_t)ns * K * 8, b_fwd = (size_t)ns * 8,
                 b_via = (size_t)std::max<long long>(nvia, 1) * sizeof(hist_cell);
    rc = check_cuda(ctx, cudaMallocAsync((void**)&obuf, b_cell + b_abs + b_fwd + b_via, ctx->stream), "cudaMallocAsync(results)");
    if (rc) { cleanup(ebuf); return rc; }
    hist_cell* d_cell = (hist_cell*)obuf;
    double* d_abs = (double*)(obuf + b_cell);
    double* d_fwd = (double*)(obuf + b_cell + b_abs);
    hist_cell* d_via = (hist_cell*)(obuf + b_cell + b_abs + b_fwd);
    elim_states_kernel<<<n, 64, 0, ctx->stream>>>(E, pair->lp_end, d_cell, d_abs, d_fwd);
    if (nt) elim_via_kernel<<<(nt + 127) / 128, 128, 0, ctx->stream>>>(E, nt, d_via);
    ctx->launches += 2;
    rc = check_cuda(ctx, cudaGetLastError(), "result kernels");
    auto d2h = [&](void* dst, const void* src, size_t bytes) {
      if (!rc && bytes) rc = check_cuda(ctx, cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, ctx->stream), "D2H results");
    };
    d2h(r->state_cell, d_cell, (size_t)ns * sizeof(hist_cell));
    d2h(r->state_lp_absorb, d_abs, b_abs);
    d2h(r->state_fwd_lp, d_fwd, b_fwd);
    d2h(r->trans_src, E.t_src, (size_t)nt * 4);
    d2h(r->trans_dest, E.t_dst, (size_t)nt * 4);
    d2h(r->trans_lp, E.t_lp, (size_t)nt * 8);
    d2h(r->trans_best_lp, E.t_best, (size_t)nt * 8);
    d2h(r->trans_best_via_off, E.t_via_off, (size_t)(nt + 1) * 8);
    d2h(r->trans_best_via, d_via, (size_t)nvia * sizeof(hist_cell));
    if (!rc) rc = check_cuda(ctx, cudaStreamSynchronize(ctx->stream), "sync");
    cudaFreeAsync(obuf, ctx->stream);
    cleanup(ebuf);
    return rc;
  }
}
