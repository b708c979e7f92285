// historian_b200: numeric core of ForwardMatrix::makeProfile on device.
//
// Replaces reference forward.cpp:686-843 (with isAbsorbing :470-474, eliminatedLogProbInsert
// :503-526 and the DP-graph enumerator :326-398): from a sorted set of DP cells, decide which cells
// become states of the parent profile, sum out the others in reverse order (lpPath by log-sum-exp,
// lpBestAlignPath by max, remembering the eliminated cells along the best path) and gather the
// states' absorb tables.  The alignPath / seqCoords / name assembly and addReadyStates
// (profile.cpp:268-319) are host bookkeeping on top of this result (hist_recon.cu).
//
// Every kernel takes a BATCH of jobs (one per sibling merge of a guide-tree level; blockIdx.y = job):
// (1) one thread per selected cell lists its in-set predecessors (binary search in the sorted cell
// list) -> CSR of the selected sub-DAG + out-degree counts; (2) retained flags and state numbering
// (CellCoords order, forward.cpp:703-732); (3) the reverse-order elimination, a dependent chain over
// the sub-DAG of one merge, run by one thread per job over pooled sorted lists (the sub-DAG of a
// merge is a few thousand cells; the jobs of a level are the parallel axis); (4) a packing pass that
// lays the results of every job out compactly for ONE device->host copy; (5) the states' absorb rows
// are gathered straight into the parent profiles' device tables (gather_absorb_rows), they never
// visit the host.
#include <algorithm>
#include <cmath>
#include <cstring>
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
  int* counts;          // [4]: n_states, n_trans, unused, pool overflow flag
  long long* n_via;     // [1]
  int* t_src; int* t_dst; int* t_entry;   // [pool_cap]
  double* t_lp; double* t_best;
  long long* t_via_off; // [pool_cap + 1]
  // packed output of this job (set once the counts are known): state cells, transitions, via lists
  hist_cell* o_state; int* o_src; int* o_dst; double* o_lp; double* o_best; long long* o_via_off; hist_cell* o_via;
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
__global__ void elim_edges_kernel(const ElimDev* __restrict__ jobs, int pass) {
  const ElimDev E = jobs[blockIdx.y];
  const PairDev& P = *E.P;
  for (int c = blockIdx.x * blockDim.x + threadIdx.x; c < E.n; c += gridDim.x * blockDim.x) {
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
}

// Exclusive scan of in_cnt (which == 0) or retained (which == 1) over n+1 entries, one CTA per job;
// totals[job] receives the sum.
__global__ void __launch_bounds__(1024) elim_scan_kernel(const ElimDev* __restrict__ jobs, int which, int* __restrict__ totals) {
  const ElimDev E = jobs[blockIdx.x];
  int* a = which ? E.retained : E.in_cnt;
  const int n = E.n + 1;
  __shared__ int wsum[32];
  __shared__ int carry;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = blockDim.x >> 5;
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (int base = 0; base < n; base += blockDim.x) {
    const int i = base + threadIdx.x;
    const int v = i < n ? a[i] : 0;
    int x = v;
    for (int o = 1; o < 32; o <<= 1) {
      const int y = __shfl_up_sync(0xffffffffu, x, o);
      if (lane >= o) x += y;
    }
    if (lane == 31) wsum[w] = x;
    __syncthreads();
    if (w == 0) {
      int s = lane < nw ? wsum[lane] : 0;
      for (int o = 1; o < 32; o <<= 1) {
        const int y = __shfl_up_sync(0xffffffffu, s, o);
        if (lane >= o) s += y;
      }
      wsum[lane] = s;
    }
    __syncthreads();
    const int c0 = carry;
    const int incl = c0 + (w ? wsum[w - 1] : 0) + x;
    if (i < n) a[i] = incl - v;
    __syncthreads();
    if (threadIdx.x == blockDim.x - 1) carry = incl;
    __syncthreads();
  }
  if (threadIdx.x == 0 && totals) totals[blockIdx.x] = carry;
}

// isAbsorbing (forward.cpp:470-474)
__device__ __forceinline__ bool is_absorbing(const PairDev& P, const DCell& c) {
  const bool xe = P.x.flags[c.x] & F_EMIT, ye = P.y.flags[c.y] & F_EMIT;
  return (c.s == HIST_IMM && xe && ye) || (c.s == HIST_IMD && xe) || (c.s == HIST_IDM && ye);
}

__global__ void elim_retain_kernel(const ElimDev* __restrict__ jobs) {
  const ElimDev E = jobs[blockIdx.y];
  const PairDev& P = *E.P;
  for (int c = blockIdx.x * blockDim.x + threadIdx.x; c < E.n; c += gridDim.x * blockDim.x) {
    const DCell d = E.cells[c];
    const bool isStart = d.x == 0 && d.y == 0 && d.s == HIST_IMM;
    const bool isEnd = d.x == P.x.n - 1 && d.y == P.y.n - 1 && d.s == HIST_EEE;
    // forward.cpp:703-709
    E.retained[c] = (is_absorbing(P, d) || isStart || isEnd || E.out_cnt[c] > 1 || (E.strategy & HIST_KEEP_GAPS_OPEN) ||
                     !(E.strategy & HIST_COLLAPSE_CHAINS)) ? 1 : 0;
    E.head[c] = -1;
  }
}

// find-or-insert the entry of destination state k in the sorted list of cell s; -1 on pool overflow
__device__ int pool_entry(const ElimDev& E, int& pool_n, int s, int k) {
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

__device__ __forceinline__ void prefetch_l1(const void* p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }

// forward.cpp:751-807, then the transition list of forward.cpp:810-831.  `retained` holds state indices
// (exclusive scan) with retained[n] = number of states; is_ret = retained[c+1] != retained[c].
// One warp per job: the elimination itself is a dependent chain over the cells (lane 0 walks it), and what it
// costs is the latency of its loads; the other lanes run ahead of it and pull the next ELIM_CHUNK cells' records
// (cells, edge lists, list heads) into L1, so that the chain sees L1 hits instead of L2 round trips.
#define ELIM_CHUNK 128
__global__ void __launch_bounds__(32) elim_reduce_kernel(const ElimDev* __restrict__ jobs) {
  const ElimDev E = jobs[blockIdx.x];
  const PairDev& P = *E.P;
  const int lane = threadIdx.x;
  int pool_n = 0;
  bool overflow = false;
  for (int hi = E.n - 1; hi >= 0; hi -= ELIM_CHUNK) {
    const int lo = max(hi - ELIM_CHUNK + 1, 0);
    {
      // one 128-B line holds 10 cells / 32 ints / 16 doubles
      const int nxt_hi = lo - 1, nxt_lo = max(lo - ELIM_CHUNK, 0);
      const int first = hi == E.n - 1 ? lo : nxt_lo, last = hi == E.n - 1 ? hi : nxt_hi;   // the first round also covers its own chunk
      if (last >= first) {
        for (int c = first + lane * 8; c <= last; c += 32 * 8) prefetch_l1(E.cells + c);
        for (int c = first + lane * 32; c <= last + 1; c += 32 * 32) {
          prefetch_l1(E.in_cnt + c);
          prefetch_l1(E.retained + c);
          prefetch_l1(E.head + min(c, E.n - 1));
        }
        const int q0 = E.in_cnt[first], q1 = E.in_cnt[last + 1];
        for (int q = q0 + lane * 16; q < q1; q += 32 * 16) prefetch_l1(E.e_lp + q);
        for (int q = q0 + lane * 32; q < q1; q += 32 * 32) prefetch_l1(E.e_src + q);
      }
    }
    __syncwarp();
    if (lane == 0)
      for (int c = hi; c >= lo && !overflow; --c) {
        const DCell d = E.cells[c];
        // eliminatedLogProbInsert (forward.cpp:503-526)
        double ins = 0;
        if (d.s == HIST_IIW) { if (P.x.flags[d.x] & F_EMIT) ins = P.x.insl[d.x]; }
        else if (d.s == HIST_IMI) { if (P.y.flags[d.y] & F_EMIT) ins = P.y.insl[d.y]; }
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
    __syncwarp();
  }
  if (lane != 0) return;
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
  E.counts[2] = 0;
  E.counts[3] = overflow ? 1 : 0;
  *E.n_via = nv;
}

// compact results of every job: state cells in state order, transitions, via lists
__global__ void elim_pack_kernel(const ElimDev* __restrict__ jobs) {
  const ElimDev E = jobs[blockIdx.y];
  if (!E.o_state) return;
  const int nt = E.counts[1];
  const int stride = gridDim.x * blockDim.x, t0 = blockIdx.x * blockDim.x + threadIdx.x;
  for (int c = t0; c < E.n; c += stride)
    if (E.retained[c + 1] != E.retained[c]) {
      const DCell d = E.cells[c];
      E.o_state[E.retained[c]] = {d.x, d.y, d.s};
    }
  for (int t = t0; t <= nt; t += stride) {
    E.o_via_off[t] = E.t_via_off[t];
    if (t == nt) break;
    E.o_src[t] = E.t_src[t];
    E.o_dst[t] = E.t_dst[t];
    E.o_lp[t] = E.t_lp[t];
    E.o_best[t] = E.t_best[t];
    long long o = E.t_via_off[t];
    for (int v = E.t_entry[t]; v >= 0 && E.p_via_cell[v] >= 0; v = E.p_via_next[v]) {
      const DCell c = E.cells[E.p_via_cell[v]];
      E.o_via[o++] = {c.x, c.y, c.s};
    }
  }
}

// absorb tables of profile states (forward.cpp:713-727): IMM both-emit -> subx[i] + suby[j], IMD -> subx[i], IDM -> suby[j];
// anything else (and s < 0) is a null state: a row of -inf.  One warp per row.
__global__ void gather_rows_kernel(const GatherRow* __restrict__ rows, long long n_rows, int K, double* __restrict__ out) {
  const long long r = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (r >= n_rows) return;
  const GatherRow g = rows[r];
  const int lane = threadIdx.x & 31;
  int mode = 0;   // 0 null, 1 x + y, 2 x, 3 y
  if (g.s >= 0) {
    const PairDev& P = *g.P;
    const bool xe = P.x.flags[g.x] & F_EMIT, ye = P.y.flags[g.y] & F_EMIT;
    if (g.s == HIST_IMM && xe && ye) mode = 1;
    else if (g.s == HIST_IMD && xe) mode = 2;
    else if (g.s == HIST_IDM && ye) mode = 3;
  }
  double* o = out + r * K;
  if (mode == 0) {
    for (int q = lane; q < K; q += 32) o[q] = -INFINITY;
    return;
  }
  const PairDev& P = *g.P;
  const double* sx = P.x.sub + (size_t)g.x * K;
  const double* sy = P.y.sub + (size_t)g.y * K;
  for (int q = lane; q < K; q += 32) o[q] = mode == 1 ? sx[q] + sy[q] : mode == 2 ? sx[q] : sy[q];
}

// fwdLogProb tag of each state (forward.cpp:731)
__global__ void elim_fwd_kernel(const PairDev* __restrict__ Pp, const hist_cell* __restrict__ st, int ns, double lp_end, double* __restrict__ out) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= ns) return;
  const hist_cell d = st[k];
  out[k] = d.state == HIST_EEE ? lp_end : dp_cell(*Pp, false, d.xpos, d.ypos, d.state);
}

int gather_absorb_rows(hist_ctx* ctx, const GatherRow* d_rows, long long n_rows, int K, double* d_out) {
  if (n_rows <= 0) return HIST_OK;
  const int warps = 8;
  gather_rows_kernel<<<(unsigned)((n_rows + warps - 1) / warps), 32 * warps, 0, ctx->stream>>>(d_rows, n_rows, K, d_out);
  ctx->launches += 1;
  return check_cuda(ctx, cudaGetLastError(), "gather_rows_kernel");
}

namespace {
struct JobLayout {
  size_t cells, in, out, ret, head, small_end;                    // small per-cell arrays (zeroed by ONE memset for all jobs)
  size_t pk, pn, pvc, pvn, pp, pb, ts, td, te, tl, tb, tv, end;   // effective-transition pool and transition lists
  long long pool_cap;
};

void layout_small(JobLayout& L, size_t base, int n, bool own_cells) {
  size_t off = base;
  auto take = [&](size_t bytes) { off = (off + 15) / 16 * 16; const size_t o = off; off += bytes; return o; };
  L.cells = take(own_cells ? (size_t)n * sizeof(DCell) : 0);
  L.in = take((size_t)(n + 1) * 4); L.out = take((size_t)n * 4); L.ret = take((size_t)(n + 1) * 4); L.head = take((size_t)n * 4);
  L.small_end = (off + 255) & ~(size_t)255;
}

void layout_pool(JobLayout& L, size_t base, long long pool_cap) {
  size_t off = base;
  auto take = [&](size_t bytes) { off = (off + 15) / 16 * 16; const size_t o = off; off += bytes; return o; };
  L.pool_cap = pool_cap;
  L.pk = take((size_t)pool_cap * 4); L.pn = take((size_t)pool_cap * 4); L.pvc = take((size_t)pool_cap * 4); L.pvn = take((size_t)pool_cap * 4);
  L.pp = take((size_t)pool_cap * 8); L.pb = take((size_t)pool_cap * 8);
  L.ts = take((size_t)pool_cap * 4); L.td = take((size_t)pool_cap * 4); L.te = take((size_t)pool_cap * 4);
  L.tl = take((size_t)pool_cap * 8); L.tb = take((size_t)pool_cap * 8); L.tv = take((size_t)(pool_cap + 1) * 8);
  L.end = (off + 255) & ~(size_t)255;
}

// One attempt over a set of jobs with the given pool capacities; overflowed[k] is set for jobs whose pool ran out.
int eliminate_attempt(hist_ctx* ctx, const std::vector<ElimJob>& jobs, const std::vector<int>& which, const std::vector<long long>& pool_cap,
                      std::vector<ElimOut>& out, std::vector<char>& overflowed) {
  static_assert(sizeof(DCell) == sizeof(hist_cell), "cell layouts must agree");
  const int nj = (int)which.size();
  std::vector<JobLayout> L(nj);
  size_t total = ((size_t)nj * sizeof(ElimDev) + 255) & ~(size_t)255;
  const size_t tot_off = total;
  total += ((size_t)nj * sizeof(int) + 255) & ~(size_t)255;
  const size_t cnt_off = total;   // counts [nj][4] and via counts [nj] of all jobs, contiguous: one read-back
  total += ((size_t)nj * 16 + 255) & ~(size_t)255;
  const size_t nv_off = total;
  total += ((size_t)nj * 8 + 255) & ~(size_t)255;
  int max_n = 1;
  const size_t small_off = total;
  for (int q = 0; q < nj; ++q) {
    const ElimJob& J = jobs[which[q]];
    layout_small(L[q], total, J.n, J.d_cells == nullptr);
    total = L[q].small_end;
    max_n = std::max(max_n, J.n);
  }
  const size_t small_end = total;
  for (int q = 0; q < nj; ++q) {
    layout_pool(L[q], total, pool_cap[q]);
    total = L[q].end;
  }
  char* buf = nullptr;
  HB_CUDA(ctx, cudaMallocFromPoolAsync((void**)&buf, total, ctx->pool, ctx->stream));
  char* ebuf = nullptr;
  char* obuf = nullptr;
  auto cleanup = [&]() {
    if (obuf) cudaFreeAsync(obuf, ctx->stream);
    if (ebuf) cudaFreeAsync(ebuf, ctx->stream);
    cudaFreeAsync(buf, ctx->stream);
  };
  std::vector<ElimDev> E(nj);
  int rc = check_cuda(ctx, cudaMemsetAsync(buf + small_off, 0, small_end - small_off, ctx->stream), "memset counts");
  for (int q = 0; q < nj && !rc; ++q) {
    const ElimJob& J = jobs[which[q]];
    const JobLayout& l = L[q];
    ElimDev& e = E[q];
    memset(&e, 0, sizeof(e));
    e.P = J.pair->d_dev;
    e.n = J.n;
    e.strategy = J.strategy;
    if (J.d_cells) e.cells = (const DCell*)J.d_cells;
    else {
      e.cells = (const DCell*)(buf + l.cells);
      rc = check_cuda(ctx, cudaMemcpyAsync(buf + l.cells, J.h_cells, (size_t)J.n * sizeof(DCell), cudaMemcpyHostToDevice, ctx->stream), "H2D cells");
    }
    e.in_cnt = (int*)(buf + l.in); e.out_cnt = (int*)(buf + l.out); e.retained = (int*)(buf + l.ret); e.head = (int*)(buf + l.head);
    e.pool_cap = (int)l.pool_cap;
    e.p_k = (int*)(buf + l.pk); e.p_next = (int*)(buf + l.pn); e.p_via_cell = (int*)(buf + l.pvc); e.p_via_next = (int*)(buf + l.pvn);
    e.p_path = (double*)(buf + l.pp); e.p_best = (double*)(buf + l.pb);
    e.counts = (int*)(buf + cnt_off) + 4 * q; e.n_via = (long long*)(buf + nv_off) + q;
    e.t_src = (int*)(buf + l.ts); e.t_dst = (int*)(buf + l.td); e.t_entry = (int*)(buf + l.te);
    e.t_lp = (double*)(buf + l.tl); e.t_best = (double*)(buf + l.tb); e.t_via_off = (long long*)(buf + l.tv);
  }
  if (rc) { cleanup(); return rc; }
  ElimDev* d_jobs = (ElimDev*)buf;
  int* d_tot = (int*)(buf + tot_off);
  auto upload_jobs = [&]() { return check_cuda(ctx, cudaMemcpyAsync(d_jobs, E.data(), (size_t)nj * sizeof(ElimDev), cudaMemcpyHostToDevice, ctx->stream), "H2D jobs"); };
  rc = upload_jobs();
  if (rc) { cleanup(); return rc; }
  const int tb = 128;
  const dim3 grid((unsigned)std::min(64, (max_n + tb - 1) / tb), (unsigned)nj);
  elim_edges_kernel<<<grid, tb, 0, ctx->stream>>>(d_jobs, 0);
  elim_scan_kernel<<<nj, 1024, 0, ctx->stream>>>(d_jobs, 0, d_tot);
  ctx->launches += 2;
  std::vector<int> n_edges(nj);
  rc = check_cuda(ctx, cudaGetLastError(), "elimination kernels (edges)");
  if (!rc) rc = check_cuda(ctx, cudaMemcpyAsync(n_edges.data(), d_tot, (size_t)nj * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream), "D2H edge counts");
  if (!rc) rc = check_cuda(ctx, cudaStreamSynchronize(ctx->stream), "sync");
  if (rc) { cleanup(); return rc; }
  size_t etot = 0;
  std::vector<size_t> eoff(nj);
  for (int q = 0; q < nj; ++q) {
    eoff[q] = etot;
    etot += ((size_t)std::max(n_edges[q], 1) * 12 + 255) & ~(size_t)255;
  }
  rc = check_cuda(ctx, cudaMallocFromPoolAsync((void**)&ebuf, etot, ctx->pool, ctx->stream), "cudaMallocAsync(edges)");
  if (rc) { ebuf = nullptr; cleanup(); return rc; }
  for (int q = 0; q < nj; ++q) {
    E[q].e_lp = (double*)(ebuf + eoff[q]);
    E[q].e_src = (int*)(ebuf + eoff[q] + (size_t)std::max(n_edges[q], 1) * 8);
  }
  rc = upload_jobs();
  if (rc) { cleanup(); return rc; }
  elim_edges_kernel<<<grid, tb, 0, ctx->stream>>>(d_jobs, 1);
  elim_retain_kernel<<<grid, tb, 0, ctx->stream>>>(d_jobs);
  elim_scan_kernel<<<nj, 1024, 0, ctx->stream>>>(d_jobs, 1, nullptr);
  elim_reduce_kernel<<<nj, 32, 0, ctx->stream>>>(d_jobs);
  ctx->launches += 4;
  rc = check_cuda(ctx, cudaGetLastError(), "elimination kernels");
  std::vector<int> counts((size_t)nj * 4);
  std::vector<long long> nvia(nj);
  if (!rc) rc = check_cuda(ctx, cudaMemcpyAsync(counts.data(), buf + cnt_off, (size_t)nj * 16, cudaMemcpyDeviceToHost, ctx->stream), "D2H counts");
  if (!rc) rc = check_cuda(ctx, cudaMemcpyAsync(nvia.data(), buf + nv_off, (size_t)nj * 8, cudaMemcpyDeviceToHost, ctx->stream), "D2H via counts");
  if (!rc) rc = check_cuda(ctx, cudaStreamSynchronize(ctx->stream), "sync");
  if (rc) { cleanup(); return rc; }
  // ---- packed results, one device->host copy
  size_t otot = 0;
  std::vector<size_t> o_base(nj);
  struct OOff { size_t st, src, dst, lp, best, voff, via, end; };
  std::vector<OOff> oo(nj);
  int max_work = 1;
  for (int q = 0; q < nj; ++q) {
    overflowed[q] = counts[(size_t)q * 4 + 3] != 0;
    if (overflowed[q]) continue;
    const size_t ns = counts[(size_t)q * 4], nt = counts[(size_t)q * 4 + 1], nv = (size_t)nvia[q];
    size_t off = otot;
    auto take = [&](size_t bytes) { off = (off + 15) / 16 * 16; const size_t o = off; off += bytes; return o; };
    oo[q].voff = take((nt + 1) * 8); oo[q].lp = take(nt * 8); oo[q].best = take(nt * 8);
    oo[q].st = take(ns * sizeof(hist_cell)); oo[q].via = take(nv * sizeof(hist_cell));
    oo[q].src = take(nt * 4); oo[q].dst = take(nt * 4);
    oo[q].end = (off + 15) / 16 * 16;
    otot = oo[q].end;
    max_work = std::max(max_work, std::max(jobs[which[q]].n, (int)nt + 1));
  }
  if (otot) {
    rc = check_cuda(ctx, cudaMallocFromPoolAsync((void**)&obuf, otot, ctx->pool, ctx->stream), "cudaMallocAsync(results)");
    if (rc) { obuf = nullptr; cleanup(); return rc; }
    for (int q = 0; q < nj; ++q) {
      if (overflowed[q]) continue;
      E[q].o_state = (hist_cell*)(obuf + oo[q].st); E[q].o_src = (int*)(obuf + oo[q].src); E[q].o_dst = (int*)(obuf + oo[q].dst);
      E[q].o_lp = (double*)(obuf + oo[q].lp); E[q].o_best = (double*)(obuf + oo[q].best);
      E[q].o_via_off = (long long*)(obuf + oo[q].voff); E[q].o_via = (hist_cell*)(obuf + oo[q].via);
    }
    rc = upload_jobs();
    if (!rc) {
      elim_pack_kernel<<<dim3((unsigned)std::min(64, (max_work + tb - 1) / tb), (unsigned)nj), tb, 0, ctx->stream>>>(d_jobs);
      ctx->launches += 1;
      rc = check_cuda(ctx, cudaGetLastError(), "elim_pack_kernel");
    }
    std::vector<char> host;
    if (!rc) {
      host.resize(otot);
      rc = check_cuda(ctx, cudaMemcpyAsync(host.data(), obuf, otot, cudaMemcpyDeviceToHost, ctx->stream), "D2H elimination results");
    }
    if (!rc) rc = check_cuda(ctx, cudaStreamSynchronize(ctx->stream), "sync");
    if (rc) { cleanup(); return rc; }
    host_parallel_for(nj, 16, [&](int q) {
      if (overflowed[q]) return;
      ElimOut& o = out[which[q]];
      const size_t ns = counts[(size_t)q * 4], nt = counts[(size_t)q * 4 + 1], nv = (size_t)nvia[q];
      o.n_states = (int)ns; o.n_trans = (int)nt; o.n_via = (long long)nv;
      const char* h = host.data();
      o.state_cell.assign((const hist_cell*)(h + oo[q].st), (const hist_cell*)(h + oo[q].st) + ns);
      o.t_src.assign((const int*)(h + oo[q].src), (const int*)(h + oo[q].src) + nt);
      o.t_dst.assign((const int*)(h + oo[q].dst), (const int*)(h + oo[q].dst) + nt);
      o.t_lp.assign((const double*)(h + oo[q].lp), (const double*)(h + oo[q].lp) + nt);
      o.t_best.assign((const double*)(h + oo[q].best), (const double*)(h + oo[q].best) + nt);
      o.via_off.assign((const long long*)(h + oo[q].voff), (const long long*)(h + oo[q].voff) + nt + 1);
      o.via.assign((const hist_cell*)(h + oo[q].via), (const hist_cell*)(h + oo[q].via) + nv);
    });
  }
  cleanup();
  return HIST_OK;
}
}  // namespace

int eliminate_batch(hist_ctx* ctx, const std::vector<ElimJob>& jobs, std::vector<ElimOut>& out) {
  const int nj = (int)jobs.size();
  out.assign(nj, ElimOut());
  if (!nj) return HIST_OK;
  cudaSetDevice(ctx->device);
  std::vector<int> which(nj);
  std::vector<long long> cap(nj);
  for (int k = 0; k < nj; ++k) {
    which[k] = k;
    cap[k] = 8ll * jobs[k].n + 1024;
  }
  while (!which.empty()) {
    std::vector<char> over(which.size(), 0);
    const int rc = eliminate_attempt(ctx, jobs, which, cap, out, over);
    if (rc) return rc;
    std::vector<int> again;
    std::vector<long long> cap2;
    for (size_t q = 0; q < which.size(); ++q)
      if (over[q]) {   // pool too small: retry this job with a larger one
        if (cap[q] > (1ll << 28)) return set_err(ctx, HIST_ERR_CAPACITY, "hist_profile_eliminate: effective-transition pool exhausted");
        again.push_back(which[q]);
        cap2.push_back(cap[q] * 4);
      }
    which.swap(again);
    cap.swap(cap2);
  }
  return HIST_OK;
}

}  // namespace hb

using namespace hb;

extern "C" int hist_profile_eliminate(hist_pair* pair, const hist_cell* cells, int64_t n_cells, int32_t strategy, hist_elim_result* r) {
  if (!pair || !cells || !r || n_cells < 2) return HIST_ERR_ARG;
  hist_ctx* ctx = pair->ctx;
  if (!pair->fwd_done) return set_err(ctx, HIST_ERR_STATE, "hist_profile_eliminate: the Forward matrix has not been filled");
  if (!pair->x->sorted_in || !pair->y->sorted_in)
    return set_err(ctx, HIST_ERR_ARG, "hist_profile_eliminate: profile transitions must be sorted by (source, destination) without repeats");
  if (n_cells > (1ll << 24)) return set_err(ctx, HIST_ERR_ARG, "hist_profile_eliminate: too many cells");
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

  std::vector<ElimJob> jobs(1);
  jobs[0] = {pair, nullptr, cells, n, strategy};
  std::vector<ElimOut> outs;
  int rc = eliminate_batch(ctx, jobs, outs);
  if (rc) return rc;
  const ElimOut& o = outs[0];
  const int ns = o.n_states, nt = o.n_trans;
  r->n_states = ns;
  r->n_trans = nt;
  r->n_via = o.n_via;
  if (ns > r->state_capacity || nt > r->trans_capacity || o.n_via > r->via_capacity)
    return set_err(ctx, HIST_ERR_CAPACITY, "hist_profile_eliminate: need %d states, %d transitions, %lld via cells", ns, nt, o.n_via);
  memcpy(r->state_cell, o.state_cell.data(), (size_t)ns * sizeof(hist_cell));
  memcpy(r->trans_src, o.t_src.data(), (size_t)nt * 4);
  memcpy(r->trans_dest, o.t_dst.data(), (size_t)nt * 4);
  memcpy(r->trans_lp, o.t_lp.data(), (size_t)nt * 8);
  memcpy(r->trans_best_lp, o.t_best.data(), (size_t)nt * 8);
  memcpy(r->trans_best_via_off, o.via_off.data(), (size_t)(nt + 1) * 8);
  if (o.n_via) memcpy(r->trans_best_via, o.via.data(), (size_t)o.n_via * sizeof(hist_cell));
  // ---- absorb tables and fwdLogProb of the states
  const size_t b_rows = ((size_t)ns * sizeof(GatherRow) + 255) & ~(size_t)255, b_abs = ((size_t)ns * K * 8 + 255) & ~(size_t)255,
               b_cell = ((size_t)ns * sizeof(hist_cell) + 255) & ~(size_t)255;
  char* obuf = nullptr;
  HB_CUDA(ctx, cudaMallocFromPoolAsync((void**)&obuf, b_rows + b_abs + b_cell + (size_t)ns * 8, ctx->pool, ctx->stream));
  std::vector<GatherRow> rows(ns);
  for (int k = 0; k < ns; ++k) rows[k] = {pair->d_dev, o.state_cell[k].xpos, o.state_cell[k].ypos, o.state_cell[k].state == HIST_EEE ? -1 : o.state_cell[k].state};
  double* d_abs = (double*)(obuf + b_rows);
  hist_cell* d_cell = (hist_cell*)(obuf + b_rows + b_abs);
  double* d_fwd = (double*)(obuf + b_rows + b_abs + b_cell);
  rc = check_cuda(ctx, cudaMemcpyAsync(obuf, rows.data(), (size_t)ns * sizeof(GatherRow), cudaMemcpyHostToDevice, ctx->stream), "H2D rows");
  if (!rc) rc = check_cuda(ctx, cudaMemcpyAsync(d_cell, o.state_cell.data(), (size_t)ns * sizeof(hist_cell), cudaMemcpyHostToDevice, ctx->stream), "H2D state cells");
  if (!rc) rc = gather_absorb_rows(ctx, (const GatherRow*)obuf, ns, K, d_abs);
  if (!rc) {
    elim_fwd_kernel<<<(ns + 127) / 128, 128, 0, ctx->stream>>>(pair->d_dev, d_cell, ns, pair->lp_end, d_fwd);
    ctx->launches += 1;
    rc = check_cuda(ctx, cudaGetLastError(), "elim_fwd_kernel");
  }
  if (!rc) rc = check_cuda(ctx, cudaMemcpyAsync(r->state_lp_absorb, d_abs, (size_t)ns * K * 8, cudaMemcpyDeviceToHost, ctx->stream), "D2H absorb tables");
  if (!rc) rc = check_cuda(ctx, cudaMemcpyAsync(r->state_fwd_lp, d_fwd, (size_t)ns * 8, cudaMemcpyDeviceToHost, ctx->stream), "D2H fwd lp");
  if (!rc) rc = check_cuda(ctx, cudaStreamSynchronize(ctx->stream), "sync");
  cudaFreeAsync(obuf, ctx->stream);
  return rc;
}
