// historian_b200: traceback through the Forward / Backward matrices on device.
//
// Replaces ForwardMatrix::bestTrace / sampleTrace with bestCell / sampleCell
// (reference forward.cpp:225-302) and BackwardMatrix::bestTrace(start) (forward.cpp:1287-1300).
//
// One warp walks one trace.  At every step the lanes evaluate the candidate predecessor cells in
// parallel (transition weight + emission + stored Forward value, all natural-log doubles), then the
// warp picks the arg-max (first strict maximum in CellCoords order, forward.cpp:245-255) or lane 0
// draws from the candidates exactly as sampleCell does: p0 ~ U(0, sum exp(lp - max)), subtract in
// CellCoords order (forward.cpp:225-243).  The uniform draw replicates libstdc++'s
// uniform_real_distribution<double> on std::mt19937 (two 32-bit outputs per draw), and the generator
// state is passed in and out, so the caller's stream stays in lock-step with the reference's.
#include <algorithm>
#include <vector>

#include "hist_dpgraph.cuh"

namespace hb {

struct TraceArgs {
  const PairDev* P;
  int mode;            // 0: best backwards from `origin` through Forward; 1: sample from EEE; 2: best forwards from `origin` through Backward
  DCell origin;
  int n_traces;
  uint32_t* mt;        // [625] device copy of the generator state (mode 1)
  hist_cell* paths;    // output
  long long capacity;
  long long* offs;     // [n_traces + 1]
  int* status;         // 0 ok, 1 capacity, 2 numeric (no finite candidate while sampling), 3 candidate buffer too small
  int cand_cap;
  uint32_t* mt_snap;   // optional [n_traces][625]: generator state after each sampled trace
};

struct Mt {
  uint32_t* mt;  // [624] + position at [624]
  __device__ void twist() {
    for (int k = 0; k < 624; ++k) {
      const uint32_t y = (mt[k] & 0x80000000u) | (mt[(k + 1) % 624] & 0x7fffffffu);
      mt[k] = mt[(k + 397) % 624] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
    }
    mt[624] = 0;
  }
  __device__ uint32_t next() {
    if (mt[624] >= 624) twist();
    uint32_t z = mt[mt[624]++];
    z ^= (z >> 11);
    z ^= (z << 7) & 0x9d2c5680u;
    z ^= (z << 15) & 0xefc60000u;
    z ^= (z >> 18);
    return z;
  }
  // std::generate_canonical<double, 53>(mt19937) as libstdc++ computes it
  __device__ double canonical() {
    const double a = (double)next();
    const double b = (double)next();
    double r = (a + b * 4294967296.0) / 18446744073709551616.0;
    if (r >= 1.0) r = 0.99999999999999988897769753748;  // nextafter(1, 0)
    return r;
  }
};

// lpCellEmitOrAbsorb (forward.cpp:404-440) of one cell by the whole warp: the lanes split the C*A terms of
// computeLogProbAbsorb (forward.h:112-124) instead of every lane repeating the serial sum (FP64 exp is the
// expensive part of a traceback step).  Every lane returns the same value (xor-shuffle reductions).
__device__ __forceinline__ double dp_emit_warp(const PairDev& P, const DCell& c, int lane) {
  const bool xe = c.x < P.x.n && (P.x.flags[c.x] & F_EMIT), ye = c.y < P.y.n && (P.y.flags[c.y] & F_EMIT);
  if (c.s != HIST_IMM || !(xe && ye)) return dp_emit(P, c);
  const double* sx = P.x.sub + (size_t)c.x * P.K;
  const double* sy = P.y.sub + (size_t)c.y * P.K;
  double t0 = -INFINITY, t1 = -INFINITY, t2 = -INFINITY;   // K <= 96 in registers, more through a loop
  if (lane < P.K) t0 = P.log_root[lane] + sx[lane] + sy[lane];
  if (lane + 32 < P.K) t1 = P.log_root[lane + 32] + sx[lane + 32] + sy[lane + 32];
  if (lane + 64 < P.K) t2 = P.log_root[lane + 64] + sx[lane + 64] + sy[lane + 64];
  double m = fmax(fmax(t0, t1), t2);
  for (int k = lane + 96; k < P.K; k += 32) m = fmax(m, P.log_root[k] + sx[k] + sy[k]);
  for (int o = 16; o; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
  if (m == -INFINITY) return -INFINITY;
  double acc = 0;
  if (t0 > -INFINITY) acc += exp(t0 - m);
  if (t1 > -INFINITY) acc += exp(t1 - m);
  if (t2 > -INFINITY) acc += exp(t2 - m);
  for (int k = lane + 96; k < P.K; k += 32) acc += exp(P.log_root[k] + sx[k] + sy[k] - m);
  for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  return m + log(acc);
}

// destination candidates of the Backward walk (destTransitions + destCells, forward.cpp:1216-1285);
// lane 0 lists them (the lists are short), all lanes then evaluate them
__device__ int list_dests(const PairDev& P, const DCell& sc, DCell* cand, double* clp, int cap) {
  const double (*w)[6] = P.Wlog;
  const int xi = sc.x, yj = sc.y;
  const unsigned fx = P.x.flags[xi], fy = P.y.flags[yj];
  const bool xReady = fx & F_READY, yReady = fy & F_READY, xEmitOrStart = fx & F_EMIT_OR_START;
  const int xEnd = P.x.n - 1, yEnd = P.y.n - 1;
  int n = 0;
  auto push = [&](int a, int b, int s, double lp) {
    if (n < cap) { cand[n] = {a, b, s}; clp[n] = lp; }
    ++n;
  };
  for (int qx = P.x.ao_off[xi]; qx < P.x.ao_off[xi + 1]; ++qx)
    for (int qy = P.y.ao_off[yj]; qy < P.y.ao_off[yj + 1]; ++qy) {
      const int tx = P.x.ao_idx[qx], ty = P.y.ao_idx[qy];
      push(P.x.t_dst[tx], P.y.t_dst[ty], HIST_IMM, w[sc.s][0] + P.x.t_lp[tx] + P.y.t_lp[ty]);
    }
  if (yReady)
    for (int qx = P.x.ao_off[xi]; qx < P.x.ao_off[xi + 1]; ++qx) {
      const int tx = P.x.ao_idx[qx];
      push(P.x.t_dst[tx], yj, HIST_IMD, w[sc.s][1] + P.x.t_lp[tx]);
      push(P.x.t_dst[tx], yj, HIST_IIW, w[sc.s][4] + P.x.t_lp[tx]);
    }
  if (xReady)
    for (int qy = P.y.ao_off[yj]; qy < P.y.ao_off[yj + 1]; ++qy) {
      const int ty = P.y.ao_idx[qy];
      push(xi, P.y.t_dst[ty], HIST_IDM, w[sc.s][2] + P.y.t_lp[ty]);
      push(xi, P.y.t_dst[ty], HIST_IMI, w[sc.s][3] + P.y.t_lp[ty]);
    }
  if (yReady && (sc.s == HIST_IMD || sc.s == HIST_IIW || sc.s == HIST_IMM))
    for (int qx = P.x.no_off[xi]; qx < P.x.no_off[xi + 1]; ++qx) {
      const int tx = P.x.no_idx[qx];
      if (P.x.t_dst[tx] != xEnd) push(P.x.t_dst[tx], yj, sc.s, P.x.t_lp[tx]);
    }
  if (sc.s == HIST_IDM || sc.s == HIST_IMI || (xEmitOrStart && sc.s == HIST_IMM))
    for (int qy = P.y.no_off[yj]; qy < P.y.no_off[yj + 1]; ++qy) {
      const int ty = P.y.no_idx[qy];
      if (P.y.t_dst[ty] != yEnd) push(xi, P.y.t_dst[ty], sc.s, P.y.t_lp[ty]);
    }
  for (int qx = P.x.no_off[xi]; qx < P.x.no_off[xi + 1]; ++qx) {
    const int tx = P.x.no_idx[qx];
    if (P.x.t_dst[tx] != xEnd) continue;
    for (int qy = P.y.no_off[yj]; qy < P.y.no_off[yj + 1]; ++qy) {
      const int ty = P.y.no_idx[qy];
      if (P.y.t_dst[ty] == yEnd) push(xEnd, yEnd, HIST_EEE, P.x.t_lp[tx] + P.y.t_lp[ty] + w[sc.s][5]);
    }
  }
  return n;
}

// grid = traces batch: block b walks the trace(s) described by args[b]
__global__ void __launch_bounds__(32) trace_kernel(const TraceArgs* __restrict__ args) {
  extern __shared__ __align__(16) unsigned char sm_raw[];
  const TraceArgs a = args[blockIdx.x];
  const PairDev& P = *a.P;
  const int lane = threadIdx.x;
  double* clp = reinterpret_cast<double*>(sm_raw);                 // [cand_cap]
  DCell* cand = reinterpret_cast<DCell*>(clp + a.cand_cap);        // [cand_cap]
  __shared__ uint32_t s_mt[625];
  __shared__ DCell s_cur;
  __shared__ int s_flag;
  const double (*w)[6] = P.Wlog;
  Mt gen;
  gen.mt = s_mt;
  if (a.mode == 1)
    for (int k = lane; k < 625; k += 32) s_mt[k] = a.mt[k];
  if (lane == 0) { a.offs[0] = 0; s_flag = 0; }
  __syncwarp();
  long long off = 0;
  const DCell endCell = {P.x.n - 1, P.y.n - 1, HIST_EEE};
  for (int tr = 0; tr < a.n_traces; ++tr) {
    DCell cur = a.mode == 1 ? endCell : a.origin;
    long long len = 0;
    if (a.mode != 2) {  // the path includes its first cell
      if (off + len < a.capacity) { if (lane == 0) a.paths[off + len] = {cur.x, cur.y, cur.s}; } else if (lane == 0) s_flag = 1;
      ++len;
    }
    bool walking = a.mode == 2 ? (cur.x < P.x.n - 1 && cur.y < P.y.n - 1) : (cur.x > 0 || cur.y > 0);
    while (walking) {
      int n;
      if (a.mode == 2) {
        if (lane == 0) s_cur.x = list_dests(P, cur, cand, clp, a.cand_cap);
        __syncwarp();
        n = s_cur.x;
        if (n > a.cand_cap) { if (lane == 0) s_flag = 3; break; }
        for (int c = lane; c < n; c += 32) {
          const DCell d = cand[c];
          double lp = clp[c] + dp_emit(P, d);
          if (d.s != HIST_EEE) lp += dp_cell<true>(P, true, d.x, d.y, d.s);
          clp[c] = lp;
        }
      } else {
        const SrcEnum e = source_enum(P, cur, nullptr);
        n = e.n;
        if (n > a.cand_cap) { if (lane == 0) s_flag = 3; break; }
        // The reference adds lpCellEmitOrAbsorb(cur) to every candidate (forward.cpp:225-255).  It is the same term for all of
        // them: it cannot move the maximum and it cancels in exp(lp - max), so neither the best nor the sampled source depends
        // on it -- and it was a quarter of the step (an A-term log-sum-exp in FP64 behind two emission-row loads).  Left out.
        for (int c = lane; c < n; c += 32) {
          DCell s;
          double lp;
          source_get(P, w, cur, e, c, s, lp);
          cand[c] = s;
          clp[c] = lp + dp_cell<true>(P, false, s.x, s.y, s.s);
        }
      }
      __syncwarp();
      if (n == 0) break;
      // ---- choose
      if (a.mode == 1) {
        double m = -INFINITY;
        for (int c = lane; c < n; c += 32) m = fmax(m, clp[c]);
        for (int o = 16; o; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
        if (m == -INFINITY) { if (lane == 0) s_flag = 2; break; }
        for (int c = lane; c < n; c += 32) clp[c] = exp(clp[c] - m);
        __syncwarp();
        if (lane == 0) {
          double ptot = 0;
          for (int c = 0; c < n; ++c) ptot += clp[c];
          double p = gen.canonical() * ptot;  // uniform_real_distribution<double>(0, ptot)
          int pick = -1;
          for (int c = 0; c < n; ++c)
            if ((p -= clp[c]) <= 0) { pick = c; break; }
          if (pick < 0) s_flag = 2; else s_cur = cand[pick];
        }
        __syncwarp();
        if (s_flag == 2) break;
      } else {
        // first strict maximum in CellCoords order; all -inf -> {0,0,EEE} (bestCell's initial value)
        double bm = -INFINITY;
        DCell bc = {0, 0, HIST_EEE};
        for (int c = lane; c < n; c += 32) {
          const double v = clp[c];
          const DCell k = cand[c];
          if (v > bm || (v == bm && v > -INFINITY && dcell_less(k, bc))) { bm = v; bc = k; }
        }
        for (int o = 16; o; o >>= 1) {
          const double om = __shfl_xor_sync(0xffffffffu, bm, o);
          DCell oc;
          oc.x = __shfl_xor_sync(0xffffffffu, bc.x, o);
          oc.y = __shfl_xor_sync(0xffffffffu, bc.y, o);
          oc.s = __shfl_xor_sync(0xffffffffu, bc.s, o);
          if (om > bm || (om == bm && om > -INFINITY && dcell_less(oc, bc))) { bm = om; bc = oc; }
        }
        if (lane == 0) s_cur = bc;
        __syncwarp();
      }
      cur = s_cur;
      __syncwarp();
      if (off + len < a.capacity) { if (lane == 0) a.paths[off + len] = {cur.x, cur.y, cur.s}; } else if (lane == 0) s_flag = 1;
      ++len;
      walking = a.mode == 2 ? (cur.x < P.x.n - 1 && cur.y < P.y.n - 1) : !(cur.x == 0 && cur.y == 0);
      if (len > (long long)P.x.n + P.y.n + 4) break;  // cannot happen in a toposorted graph
    }
    __syncwarp();
    if (s_flag > 1) break;
    if (a.mode == 2) {  // BackwardMatrix::bestTrace always appends endCell after the walk (forward.cpp:1298)
      if (off + len < a.capacity) { if (lane == 0) a.paths[off + len] = {endCell.x, endCell.y, endCell.s}; } else if (lane == 0) s_flag = 1;
      ++len;
    } else if (s_flag == 0) {
      // walked end -> start: reverse in place so the path runs start -> end
      for (long long k = lane; k < len / 2; k += 32) {
        const hist_cell t = a.paths[off + k];
        a.paths[off + k] = a.paths[off + len - 1 - k];
        a.paths[off + len - 1 - k] = t;
      }
    }
    off += len;
    if (lane == 0) a.offs[tr + 1] = off;
    __syncwarp();
    if (a.mode == 1 && a.mt_snap)
      for (int k = lane; k < 625; k += 32) a.mt_snap[(size_t)tr * 625 + k] = s_mt[k];
  }
  __syncwarp();
  if (a.mode == 1)
    for (int k = lane; k < 625; k += 32) a.mt[k] = s_mt[k];
  if (lane == 0) *a.status = s_flag;
}

}  // namespace hb

using namespace hb;

namespace {

int max_degree(const std::vector<int>& key, int n) {
  std::vector<int> cnt(n, 0);
  int m = 0;
  for (int k : key) m = std::max(m, ++cnt[k]);
  return m;
}

int run_trace(hist_pair* pair, int mode, hist_cell origin, uint32_t* mt_state, int n_traces, hist_cell* paths, int64_t capacity,
              int64_t* path_off, const char* what) {
  hist_ctx* ctx = pair->ctx;
  if (mode == 2 ? !pair->bwd_done : !pair->fwd_done) return set_err(ctx, HIST_ERR_STATE, "%s: the matrix has not been filled", what);
  if (!pair->x->sorted_in || !pair->y->sorted_in)
    return set_err(ctx, HIST_ERR_ARG, "%s: profile transitions must be sorted by (source, destination) without repeats", what);
  if (capacity < 0 || (capacity > 0 && !paths) || n_traces < 1) return set_err(ctx, HIST_ERR_ARG, "%s: bad arguments", what);
  const int nx = pair->info.x_size, ny = pair->info.y_size;
  if (mode != 1) {
    const bool isEnd = origin.xpos == nx - 1 && origin.ypos == ny - 1 && origin.state == HIST_EEE;
    if (!isEnd && (origin.xpos < 0 || origin.ypos < 0 || origin.xpos > nx - 2 || origin.ypos > ny - 2 || origin.state < 0 || origin.state > 4))
      return set_err(ctx, HIST_ERR_ARG, "%s: cell (%d,%d,%d) outside the matrix", what, origin.xpos, origin.ypos, origin.state);
  }
  cudaSetDevice(ctx->device);
  // candidate buffer: 5 * max in-degree(x) * max in-degree(y) (sources) or the out-degree analogue (dests)
  const int dx = std::max(max_degree(pair->x->t_dst, nx), max_degree(pair->x->t_src, nx));
  const int dy = std::max(max_degree(pair->y->t_dst, ny), max_degree(pair->y->t_src, ny));
  long long cand = 5ll * dx * dy + 4ll * (dx + dy) + 8;
  if (cand * 24 > 200 * 1024) return set_err(ctx, HIST_ERR_CAPACITY, "%s: %lld candidate cells per step exceed the device buffer", what, cand);
  const size_t smem = (size_t)cand * (sizeof(double) + sizeof(DCell));
  if (smem + 4096 > 48 * 1024) HB_CUDA(ctx, cudaFuncSetAttribute(trace_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const int64_t cap = std::max<int64_t>(capacity, 1);
  char* buf = nullptr;
  const size_t b_paths = (size_t)cap * sizeof(hist_cell), b_offs = (size_t)(n_traces + 1) * sizeof(long long);
  const size_t o_offs = (b_paths + 15) / 16 * 16, o_mt = o_offs + (b_offs + 15) / 16 * 16, o_status = o_mt + 625 * 4 + 12;
  const size_t o_args = o_status + 16;
  HB_CUDA(ctx, cudaMallocFromPoolAsync((void**)&buf, o_args + sizeof(TraceArgs), ctx->pool, ctx->stream));
  TraceArgs a;
  a.P = pair->d_dev;
  a.mode = mode;
  a.origin = {origin.xpos, origin.ypos, origin.state};
  a.n_traces = n_traces;
  a.mt = (uint32_t*)(buf + o_mt);
  a.paths = (hist_cell*)buf;
  a.capacity = capacity;
  a.offs = (long long*)(buf + o_offs);
  a.status = (int*)(buf + o_status);
  a.cand_cap = (int)cand;
  a.mt_snap = nullptr;
  if (mode == 1) HB_CUDA(ctx, cudaMemcpyAsync(a.mt, mt_state, 625 * 4, cudaMemcpyHostToDevice, ctx->stream));
  HB_CUDA(ctx, cudaMemcpyAsync(buf + o_args, &a, sizeof(TraceArgs), cudaMemcpyHostToDevice, ctx->stream));   // pageable: staged before return
  trace_kernel<<<1, 32, smem, ctx->stream>>>((const TraceArgs*)(buf + o_args));
  ctx->launches += 1;
  HB_CUDA(ctx, cudaGetLastError());
  std::vector<long long> offs(n_traces + 1);
  int status = 0;
  HB_CUDA(ctx, cudaMemcpyAsync(offs.data(), a.offs, b_offs, cudaMemcpyDeviceToHost, ctx->stream));
  HB_CUDA(ctx, cudaMemcpyAsync(&status, a.status, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
  HB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  int rc = HIST_OK;
  if (status == 1) rc = set_err(ctx, HIST_ERR_CAPACITY, "%s: path buffer too small (capacity %lld cells)", what, (long long)capacity);
  else if (status == 2) rc = set_err(ctx, HIST_ERR_NUMERIC, "%s: no finite predecessor (zero forward likelihood)", what);
  else if (status == 3) rc = set_err(ctx, HIST_ERR_CAPACITY, "%s: candidate buffer too small", what);
  if (!rc) {
    const long long total = offs[n_traces];
    if (total > 0) HB_CUDA(ctx, cudaMemcpyAsync(paths, a.paths, (size_t)total * sizeof(hist_cell), cudaMemcpyDeviceToHost, ctx->stream));
    if (mode == 1) HB_CUDA(ctx, cudaMemcpyAsync(mt_state, a.mt, 625 * 4, cudaMemcpyDeviceToHost, ctx->stream));
    HB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    for (int k = 0; k <= n_traces; ++k) path_off[k] = offs[k];
  }
  cudaFreeAsync(buf, ctx->stream);
  return rc;
}

}  // namespace

namespace {
long long cand_bound(const hist_pair* pair) {
  const int nx = pair->info.x_size, ny = pair->info.y_size;
  const int dx = std::max(max_degree(pair->x->t_dst, nx), max_degree(pair->x->t_src, nx));
  const int dy = std::max(max_degree(pair->y->t_dst, ny), max_degree(pair->y->t_src, ny));
  return 5ll * dx * dy + 4ll * (dx + dy) + 8;
}
}  // namespace

namespace hb {

int trace_best_batch(hist_ctx* ctx, hist_pair* const* pairs, int n, hist_cell** d_paths, std::vector<long long>& path_off,
                     std::vector<long long>& len) {
  const char* what = "trace_best_batch";
  *d_paths = nullptr;
  if (n <= 0) return HIST_OK;
  cudaSetDevice(ctx->device);
  std::vector<long long> cands(n);
  host_parallel_for(n, 16, [&](int k) { cands[k] = pairs[k]->x->chain && pairs[k]->y->chain ? 21 : cand_bound(pairs[k]); });
  long long cand = 8;
  path_off.assign(n + 1, 0);
  for (int k = 0; k < n; ++k) {
    hist_pair* pr = pairs[k];
    if (!pr->fwd_done) return set_err(ctx, HIST_ERR_STATE, "%s: the matrix has not been filled", what);
    if (!pr->x->sorted_in || !pr->y->sorted_in)
      return set_err(ctx, HIST_ERR_ARG, "%s: profile transitions must be sorted by (source, destination) without repeats", what);
    cand = std::max(cand, cands[k]);
    path_off[k + 1] = path_off[k] + pr->info.x_size + pr->info.y_size + 4;
  }
  if (cand * 24 > 200 * 1024) return set_err(ctx, HIST_ERR_CAPACITY, "%s: %lld candidate cells per step exceed the device buffer", what, cand);
  const size_t smem = (size_t)cand * (sizeof(double) + sizeof(DCell));
  if (smem + 4096 > 48 * 1024) HB_CUDA(ctx, cudaFuncSetAttribute(trace_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const size_t b_paths = ((size_t)path_off[n] * sizeof(hist_cell) + 255) & ~(size_t)255;
  const size_t b_offs = ((size_t)n * 2 * sizeof(long long) + 255) & ~(size_t)255, b_stat = ((size_t)n * sizeof(int) + 255) & ~(size_t)255;
  char* buf = nullptr;
  HB_CUDA(ctx, cudaMallocFromPoolAsync((void**)&buf, b_paths + b_offs + b_stat + (size_t)n * sizeof(TraceArgs), ctx->pool, ctx->stream));
  std::vector<TraceArgs> args(n);
  for (int k = 0; k < n; ++k) {
    TraceArgs& a = args[k];
    a.P = pairs[k]->d_dev;
    a.mode = 0;
    a.origin = {pairs[k]->info.x_size - 1, pairs[k]->info.y_size - 1, HIST_EEE};
    a.n_traces = 1;
    a.mt = nullptr;
    a.mt_snap = nullptr;
    a.paths = (hist_cell*)buf + path_off[k];
    a.capacity = path_off[k + 1] - path_off[k];
    a.offs = (long long*)(buf + b_paths) + 2 * k;
    a.status = (int*)(buf + b_paths + b_offs) + k;
    a.cand_cap = (int)cand;
  }
  TraceArgs* d_args = (TraceArgs*)(buf + b_paths + b_offs + b_stat);
  int rc = check_cuda(ctx, cudaMemcpyAsync(d_args, args.data(), (size_t)n * sizeof(TraceArgs), cudaMemcpyHostToDevice, ctx->stream), "H2D trace args");
  if (!rc) {
    trace_kernel<<<n, 32, smem, ctx->stream>>>(d_args);
    ctx->launches += 1;
    rc = check_cuda(ctx, cudaGetLastError(), "trace_kernel");
  }
  std::vector<long long> offs((size_t)2 * n);
  std::vector<int> status(n);
  if (!rc) rc = check_cuda(ctx, cudaMemcpyAsync(offs.data(), buf + b_paths, offs.size() * sizeof(long long), cudaMemcpyDeviceToHost, ctx->stream), "D2H trace lengths");
  if (!rc) rc = check_cuda(ctx, cudaMemcpyAsync(status.data(), buf + b_paths + b_offs, (size_t)n * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream), "D2H trace status");
  if (!rc) rc = check_cuda(ctx, cudaStreamSynchronize(ctx->stream), "sync after traces");
  len.assign(n, 0);
  for (int k = 0; k < n && !rc; ++k) {
    if (status[k] == 1) rc = set_err(ctx, HIST_ERR_CAPACITY, "%s: path buffer too small (pair %d)", what, k);
    else if (status[k] == 2) rc = set_err(ctx, HIST_ERR_NUMERIC, "%s: no finite predecessor (pair %d)", what, k);
    else if (status[k] == 3) rc = set_err(ctx, HIST_ERR_CAPACITY, "%s: candidate buffer too small (pair %d)", what, k);
    len[k] = offs[(size_t)2 * k + 1];
  }
  if (rc) {
    cudaFreeAsync(buf, ctx->stream);
    return rc;
  }
  *d_paths = (hist_cell*)buf;
  return HIST_OK;
}

int trace_sample_snap(hist_pair* pair, const uint32_t* mt_state, int n_traces, std::vector<hist_cell>& paths,
                      std::vector<long long>& path_off, std::vector<uint32_t>& mt_snap) {
  const char* what = "trace_sample_snap";
  hist_ctx* ctx = pair->ctx;
  if (!pair->fwd_done) return set_err(ctx, HIST_ERR_STATE, "%s: the matrix has not been filled", what);
  if (!pair->x->sorted_in || !pair->y->sorted_in)
    return set_err(ctx, HIST_ERR_ARG, "%s: profile transitions must be sorted by (source, destination) without repeats", what);
  cudaSetDevice(ctx->device);
  const long long cand = cand_bound(pair);
  if (cand * 24 > 200 * 1024) return set_err(ctx, HIST_ERR_CAPACITY, "%s: %lld candidate cells per step exceed the device buffer", what, cand);
  const size_t smem = (size_t)cand * (sizeof(double) + sizeof(DCell));
  if (smem + 4096 > 48 * 1024) HB_CUDA(ctx, cudaFuncSetAttribute(trace_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const long long per = pair->info.x_size + pair->info.y_size + 4, cap = per * n_traces;
  const size_t b_paths = ((size_t)cap * sizeof(hist_cell) + 255) & ~(size_t)255, b_offs = ((size_t)(n_traces + 1) * 8 + 255) & ~(size_t)255;
  const size_t b_mt = (625 * 4 + 255) & ~(size_t)255, b_snap = ((size_t)n_traces * 625 * 4 + 255) & ~(size_t)255;
  char* buf = nullptr;
  HB_CUDA(ctx, cudaMallocFromPoolAsync((void**)&buf, b_paths + b_offs + b_mt + b_snap + 256 + sizeof(TraceArgs), ctx->pool, ctx->stream));
  TraceArgs a;
  a.P = pair->d_dev;
  a.mode = 1;
  a.origin = {0, 0, 0};
  a.n_traces = n_traces;
  a.paths = (hist_cell*)buf;
  a.capacity = cap;
  a.offs = (long long*)(buf + b_paths);
  a.mt = (uint32_t*)(buf + b_paths + b_offs);
  a.mt_snap = (uint32_t*)(buf + b_paths + b_offs + b_mt);
  a.status = (int*)(buf + b_paths + b_offs + b_mt + b_snap);
  a.cand_cap = (int)cand;
  TraceArgs* d_args = (TraceArgs*)(buf + b_paths + b_offs + b_mt + b_snap + 256);
  int rc = check_cuda(ctx, cudaMemcpyAsync(a.mt, mt_state, 625 * 4, cudaMemcpyHostToDevice, ctx->stream), "H2D generator");
  if (!rc) rc = check_cuda(ctx, cudaMemcpyAsync(d_args, &a, sizeof(TraceArgs), cudaMemcpyHostToDevice, ctx->stream), "H2D trace args");
  if (!rc) {
    trace_kernel<<<1, 32, smem, ctx->stream>>>(d_args);
    ctx->launches += 1;
    rc = check_cuda(ctx, cudaGetLastError(), "trace_kernel");
  }
  path_off.assign(n_traces + 1, 0);
  mt_snap.assign((size_t)n_traces * 625, 0);
  int status = 0;
  if (!rc) rc = check_cuda(ctx, cudaMemcpyAsync(path_off.data(), a.offs, (size_t)(n_traces + 1) * 8, cudaMemcpyDeviceToHost, ctx->stream), "D2H offsets");
  if (!rc) rc = check_cuda(ctx, cudaMemcpyAsync(mt_snap.data(), a.mt_snap, mt_snap.size() * 4, cudaMemcpyDeviceToHost, ctx->stream), "D2H generator snapshots");
  if (!rc) rc = check_cuda(ctx, cudaMemcpyAsync(&status, a.status, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream), "D2H status");
  if (!rc) rc = check_cuda(ctx, cudaStreamSynchronize(ctx->stream), "sync after sampled traces");
  if (!rc && status == 1) rc = set_err(ctx, HIST_ERR_CAPACITY, "%s: path buffer too small", what);
  if (!rc && status == 2) rc = set_err(ctx, HIST_ERR_NUMERIC, "%s: no finite predecessor (zero forward likelihood)", what);
  if (!rc && status == 3) rc = set_err(ctx, HIST_ERR_CAPACITY, "%s: candidate buffer too small", what);
  if (!rc) {
    paths.resize((size_t)path_off[n_traces]);
    if (!paths.empty()) rc = check_cuda(ctx, cudaMemcpyAsync(paths.data(), a.paths, paths.size() * sizeof(hist_cell), cudaMemcpyDeviceToHost, ctx->stream), "D2H paths");
    if (!rc) rc = check_cuda(ctx, cudaStreamSynchronize(ctx->stream), "sync after paths");
  }
  cudaFreeAsync(buf, ctx->stream);
  return rc;
}

}  // namespace hb

extern "C" int hist_trace_best(hist_pair* pair, hist_cell end, hist_cell* path, int64_t capacity, int64_t* n) {
  if (!pair || !n) return HIST_ERR_ARG;
  int64_t offs[2] = {0, 0};
  const int rc = run_trace(pair, 0, end, nullptr, 1, path, capacity, offs, "hist_trace_best");
  *n = offs[1];
  return rc;
}

extern "C" int hist_trace_best_forward(hist_pair* pair, hist_cell start, hist_cell* path, int64_t capacity, int64_t* n) {
  if (!pair || !n) return HIST_ERR_ARG;
  int64_t offs[2] = {0, 0};
  const int rc = run_trace(pair, 2, start, nullptr, 1, path, capacity, offs, "hist_trace_best_forward");
  *n = offs[1];
  return rc;
}

extern "C" int hist_trace_sample(hist_pair* pair, uint32_t* mt_state, int32_t n_traces, hist_cell* paths, int64_t capacity,
                                 int64_t* path_off) {
  if (!pair || !mt_state || !path_off) return HIST_ERR_ARG;
  hist_cell none = {0, 0, 0};
  return run_trace(pair, 1, none, mt_state, n_traces, paths, capacity, path_off, "hist_trace_sample");
}
