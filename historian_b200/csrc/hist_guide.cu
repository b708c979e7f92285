// historian_b200: the guide-alignment stage as a whole (the callers of the pairwise step).
//
// hist_guide_align is the B200 counterpart of AlignGraph (reference span.h:9-52, span.cpp:35-158) as `historian recon
// -seqs` uses it (recon.cpp:806-821): choose the pairs of sequences to align (all of them, or ceil(n log2 n) random
// ones, continued until the graph is connected), build a DiagonalEnvelope per pair (diagenv.cpp:93-190), align every
// pair with QuickAlignMatrix, take the maximum spanning tree of the scores from sequence 0 (span.cpp:119-144) and merge
// its n - 1 pairwise alignments into one multiple alignment (alignPathMerge, alignpath.cpp:9-21, 98-207).
//
// Design.  The O(n log n) pairwise Viterbi alignments are the work: they go to the device in batches
// (hist_quickalign_batch, csrc/hist_quickalign.cu: one CTA per pair), sized by their traceback planes.  Everything else is
// bookkeeping on the host and follows the reference's rules exactly where the outcome depends on them:
//  * the random pairs come from std::mt19937 through std::uniform_int_distribution<size_t>, two draws per trial, like the
//    reference (same standard library, same stream);
//  * the spanning tree grows from sequence 0's set with one std::priority_queue of edges per sequence, stale tops popped
//    while scanning the set's members in ascending order, a strictly better top replacing the current best;
//  * the merge emits, at every step, the linked column of the LOWEST-numbered alignment whose next column is ready (all
//    the columns it is linked to through shared residues are the next ones of their alignments).  The reference
//    recomputes the linked sets of every alignment for every output column; here residues are grouped into their final
//    columns once (union-find), a column counts how many of its alignments have reached it, and an ordered set holds the
//    alignments whose next column is ready: O(total columns x log n) instead of O(columns x alignments x rows).
//
// hist_align_path_merge exposes the merge on its own (host only, no device work): alignPathMerge is also what the
// reference's simulator, sampler and refiner call.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <numeric>
#include <queue>
#include <random>
#include <set>
#include <vector>

#include "hist_internal.cuh"

namespace hb {
namespace {

// One alignment to be merged: columns in order, each the list of residues (global ids) it holds.
struct MergeAlign {
  std::vector<int> col_off;        // [cols + 1]
  std::vector<long long> res;      // residue ids
};

struct Dsu {
  std::vector<long long> p;
  explicit Dsu(size_t n) : p(n) { std::iota(p.begin(), p.end(), 0LL); }
  long long find(long long a) {
    while (p[a] != a) { p[a] = p[p[a]]; a = p[a]; }
    return a;
  }
  void join(long long a, long long b) {
    a = find(a); b = find(b);
    if (a != b) p[std::max(a, b)] = std::min(a, b);
  }
};

// alignPathMerge (alignpath.cpp:153-207) over residue ids [0, n_res): residue_col[r] = output column of residue r, or -1
// for residues no alignment mentions.  Returns the number of columns, or -1 when the alignments are inconsistent (the
// reference's "Inconsistent alignments" / "no alignments ready" aborts).
long long merge_alignments(const std::vector<MergeAlign>& al, long long n_res, std::vector<int>& residue_col) {
  const int m = (int)al.size();
  Dsu dsu((size_t)n_res);
  std::vector<uint8_t> seen((size_t)n_res, 0);
  for (const MergeAlign& a : al)
    for (size_t c = 0; c + 1 < a.col_off.size(); ++c)
      for (int q = a.col_off[c]; q < a.col_off[c + 1]; ++q) {
        seen[(size_t)a.res[q]] = 1;
        if (q > a.col_off[c]) dsu.join(a.res[a.col_off[c]], a.res[q]);
      }
  // classes (final columns) numbered by their smallest residue; members = the (alignment, column) pairs linked into them
  std::vector<int> cls_of_res((size_t)n_res, -1);
  int n_cls = 0;
  for (long long r = 0; r < n_res; ++r)
    if (seen[(size_t)r] && dsu.find(r) == r) cls_of_res[(size_t)r] = n_cls++;
  for (long long r = 0; r < n_res; ++r)
    if (seen[(size_t)r]) cls_of_res[(size_t)r] = cls_of_res[(size_t)dsu.find(r)];
  std::vector<std::vector<int>> col_cls(m);       // class of every (alignment, column); -1: empty column
  std::vector<int> size(n_cls, 0);
  for (int k = 0; k < m; ++k) {
    const int cols = (int)al[k].col_off.size() - 1;
    col_cls[k].assign(cols, -1);
    for (int c = 0; c < cols; ++c)
      if (al[k].col_off[c + 1] > al[k].col_off[c]) {
        const int cl = cls_of_res[(size_t)al[k].res[al[k].col_off[c]]];
        col_cls[k][c] = cl;
        ++size[cl];
      }
  }
  std::vector<int> mem_off(n_cls + 1, 0);
  for (int cl = 0; cl < n_cls; ++cl) mem_off[cl + 1] = mem_off[cl] + size[cl];
  std::vector<int> mem(mem_off[n_cls]), fill(mem_off.begin(), mem_off.end() - 1);
  for (int k = 0; k < m; ++k)   // ascending alignment number inside a class; two columns of one alignment: inconsistent
    for (int cl : col_cls[k])
      if (cl >= 0) {
        if (fill[cl] > mem_off[cl] && mem[fill[cl] - 1] == k) return -1;
        mem[fill[cl]++] = k;
      }
  std::vector<int> next(m, 0), at_front(n_cls, 0), out_col(n_cls, -1);
  std::set<int> ready;   // alignments whose next column is ready
  auto arrive = [&](int k) {   // alignment k now points at column next[k]
    while (next[k] < (int)col_cls[k].size() && col_cls[k][next[k]] < 0) ++next[k];   // empty columns are skipped (alignpath.cpp:186-187)
    if (next[k] >= (int)col_cls[k].size()) return;
    const int cl = col_cls[k][next[k]];
    if (++at_front[cl] == size[cl])
      for (int q = mem_off[cl]; q < mem_off[cl + 1]; ++q) ready.insert(mem[q]);
  };
  for (int k = 0; k < m; ++k) arrive(k);
  long long n_out = 0;
  while (!ready.empty()) {
    const int k = *ready.begin();
    const int cl = col_cls[k][next[k]];
    out_col[cl] = (int)n_out++;
    for (int q = mem_off[cl]; q < mem_off[cl + 1]; ++q) ready.erase(mem[q]);
    for (int q = mem_off[cl]; q < mem_off[cl + 1]; ++q) {
      ++next[mem[q]];
      arrive(mem[q]);
    }
  }
  for (int k = 0; k < m; ++k)
    if (next[k] < (int)col_cls[k].size()) return -1;   // "no alignments ready"
  residue_col.assign((size_t)n_res, -1);
  for (long long r = 0; r < n_res; ++r)
    if (seen[(size_t)r]) residue_col[(size_t)r] = out_col[cls_of_res[(size_t)r]];
  return n_out;
}

// DiagonalEnvelope::initFull / initSparse (diagenv.cpp:93-190): the ascending list of diagonals i - j of a pair.
// tokens < 0: not in the alphabet.  P: hist_guide_params.
// An empty list stands for the full envelope (all xLen + yLen - 1 diagonals).
void envelope_diagonals(const int8_t* x, int xLen, const int8_t* y, int yLen, int A, const hist_guide_params& P, std::vector<int>& diags) {
  const int lo = 1 - yLen, hi = xLen - 1;
  auto full = [&]() { diags.clear(); };
  const int K = P.kmer_len;
  if (!P.sparse) return full();
  if (P.kmer_threshold >= 0) {
    const long long min_len = 2LL * ((long long)K + P.kmer_threshold);   // MIN_KMERS_FOR_SPARSE_ENVELOPE
    if (xLen < min_len || yLen < min_len) return full();
  } else if ((unsigned long long)((unsigned)xLen * (unsigned)yLen) * (unsigned long long)P.cell_size < (unsigned long long)P.max_size)
    return full();   // xLen * yLen in the reference's 32-bit unsigned SeqIdx (diagenv.cpp:113)
  if (xLen < K || yLen < K) return full();   // (the reference's unsigned xLen - kmerLen would wrap here)
  // KmerIndex of y (fastseq.cpp:255-266) as a sorted list; matches per diagonal
  auto kmer_at = [&](const int8_t* t, int at, unsigned long long& code) {
    code = 0;
    for (int q = 0; q < K; ++q) {
      if (t[at + q] < 0 || t[at + q] >= A) return false;
      code = code * (unsigned long long)A + (unsigned long long)t[at + q];   // makeKmer: the first token is the most significant
    }
    return true;
  };
  std::vector<std::pair<unsigned long long, int>> yk;
  yk.reserve((size_t)yLen);
  unsigned long long code;
  for (int j = 0; j + K <= yLen; ++j)
    if (kmer_at(y, j, code)) yk.push_back({code, j});
  std::sort(yk.begin(), yk.end());
  std::vector<unsigned> count((size_t)xLen + yLen, 0);   // index: diagonal + yLen
  for (int i = 0; i + K <= xLen; ++i)
    if (kmer_at(x, i, code)) {
      auto it = std::lower_bound(yk.begin(), yk.end(), std::make_pair(code, -1));
      for (; it != yk.end() && it->first == code; ++it) ++count[(size_t)(i - it->second + yLen)];
    }
  // diagonals by descending match count (countDistrib, diagenv.cpp:131-133, 158-186)
  std::vector<int> order;
  for (int d = lo - 1; d <= hi; ++d)   // get_diag of 0-based k-mer starts spans [1 - yLen + K - 1 .., xLen - K]: inside the table
    if (d + yLen >= 0 && count[(size_t)(d + yLen)]) order.push_back(d);
  std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return count[(size_t)(a + yLen)] > count[(size_t)(b + yLen)]; });
  const int half = P.band_size / 2;
  std::vector<uint8_t> in_env((size_t)xLen + yLen + 2, 0), in_store((size_t)xLen + yLen + 4, 0);   // index: d - lo, d - lo + 1
  size_t n_store = 0;
  auto add_store = [&](int d) {
    uint8_t& f = in_store[(size_t)(d - lo + 1)];
    if (!f) { f = 1; ++n_store; }
  };
  in_env[(size_t)(0 - lo)] = 1;   // the zeroth diagonal is always there
  add_store(0);
  const unsigned long long diag_size = (unsigned long long)std::min(xLen, yLen) * (unsigned long long)P.cell_size;
  for (size_t g = 0; g < order.size();) {
    const unsigned c = count[(size_t)(order[g] + yLen)];
    size_t g1 = g;
    while (g1 < order.size() && count[(size_t)(order[g1] + yLen)] == c) ++g1;
    if (P.kmer_threshold >= 0 && c < (unsigned)P.kmer_threshold) break;
    // the group's storage diagonals, counted before they are committed (memory mode stops BEFORE the group that would not fit)
    std::vector<int> new_store;
    if (P.kmer_threshold < 0) {
      size_t extra = 0;
      for (size_t q = g; q < g1; ++q) {
        const int dMin = std::max(lo, order[q] - half), dMax = std::min(hi, order[q] + half);
        for (int d = dMin - 1; d <= dMax + 1; ++d) {
          uint8_t& f = in_store[(size_t)(d - lo + 1)];
          if (!f) { f = 2; ++extra; new_store.push_back(d); }
        }
      }
      for (int d : new_store) in_store[(size_t)(d - lo + 1)] = 0;
      if ((unsigned long long)(n_store + extra) * diag_size >= (unsigned long long)P.max_size) break;
    }
    for (size_t q = g; q < g1; ++q) {
      const int dMin = std::max(lo, order[q] - half), dMax = std::min(hi, order[q] + half);
      for (int d = dMin; d <= dMax; ++d) in_env[(size_t)(d - lo)] = 1;
      for (int d = dMin - 1; d <= dMax + 1; ++d) add_store(d);
    }
    g = g1;
  }
  diags.clear();
  for (int d = lo; d <= hi; ++d)
    if (in_env[(size_t)(d - lo)]) diags.push_back(d);
}

struct GuideEdge {
  int row1, row2;
  double lp;
  int path;   // index of the pair
  bool operator<(const GuideEdge& e) const { return lp < e.lp; }   // span.h:21
};

// AlignGraph::Partition (span.cpp:4-33): the set with the smaller index absorbs the other
struct GuidePartition {
  std::vector<size_t> idx;
  std::vector<std::set<size_t>> sets;
  size_t n_sets;
  explicit GuidePartition(size_t n) : idx(n), sets(n), n_sets(n) {
    for (size_t i = 0; i < n; ++i) { idx[i] = i; sets[i].insert(i); }
  }
  bool same(size_t a, size_t b) const { return idx[a] == idx[b]; }
  void merge(size_t a, size_t b) {
    if (same(a, b)) return;
    size_t i1 = idx[a], i2 = idx[b];
    if (i1 > i2) std::swap(i1, i2);
    for (size_t v : sets[i2]) idx[v] = i1;
    sets[i1].insert(sets[i2].begin(), sets[i2].end());
    sets[i2].clear();
    --n_sets;
  }
};

}  // namespace
}  // namespace hb

using namespace hb;

extern "C" int hist_align_path_merge(hist_ctx* ctx, int32_t n_alignments, const int32_t* n_rows, const int32_t* n_cols, const int32_t* rows,
                                     const uint8_t* bits, int32_t n_total_rows, const int64_t* row_off, int32_t* residue_col,
                                     int32_t* n_out_cols) {
  if (n_alignments < 0 || (n_alignments && (!n_rows || !n_cols || !rows || !bits)) || n_total_rows < 0 || !row_off || !residue_col || !n_out_cols)
    return set_err(ctx, HIST_ERR_ARG, "hist_align_path_merge: bad arguments");
  const long long n_res = row_off[n_total_rows];
  std::vector<MergeAlign> al((size_t)n_alignments);
  size_t r_at = 0, b_at = 0;
  for (int k = 0; k < n_alignments; ++k) {
    MergeAlign& a = al[k];
    const int R = n_rows[k], Cn = n_cols[k];
    if (R < 0 || Cn < 0) return set_err(ctx, HIST_ERR_ARG, "hist_align_path_merge: alignment %d: negative size", k);
    std::vector<long long> pos(R, 0);
    a.col_off.assign(1, 0);
    for (int r = 0; r < R; ++r)
      if (rows[r_at + r] < 0 || rows[r_at + r] >= n_total_rows) return set_err(ctx, HIST_ERR_ARG, "hist_align_path_merge: alignment %d: row %d out of range", k, rows[r_at + r]);
    for (int c = 0; c < Cn; ++c) {
      for (int r = 0; r < R; ++r)
        if (bits[b_at + (size_t)r * Cn + c]) {
          const int row = rows[r_at + r];
          const long long id = row_off[row] + pos[r]++;
          if (id >= row_off[row + 1]) return set_err(ctx, HIST_ERR_ARG, "hist_align_path_merge: alignment %d: row %d has more residues than row_off allows", k, row);
          a.res.push_back(id);
        }
      if ((int)a.res.size() == a.col_off.back()) return set_err(ctx, HIST_ERR_ARG, "hist_align_path_merge: column %d of alignment %d is empty", c, k);   // alignpath.cpp:136
      a.col_off.push_back((int)a.res.size());
    }
    for (int r = 0; r < R; ++r) {
      const int row = rows[r_at + r];
      if (pos[r] != row_off[row + 1] - row_off[row])   // alignpath.cpp:113
        return set_err(ctx, HIST_ERR_ARG, "hist_align_path_merge: incompatible number of residues for row %d of alignment %d (%lld != %lld)", row, k, pos[r],
                       (long long)(row_off[row + 1] - row_off[row]));
    }
    r_at += (size_t)R;
    b_at += (size_t)R * Cn;
  }
  std::vector<int> rc;
  const long long cols = merge_alignments(al, n_res, rc);
  if (cols < 0) return set_err(ctx, HIST_ERR_STATE, "hist_align_path_merge: inconsistent alignments");
  for (long long r = 0; r < n_res; ++r) residue_col[r] = rc[(size_t)r];
  *n_out_cols = (int32_t)cols;
  return HIST_OK;
}

extern "C" int hist_guide_align(hist_ctx* ctx, int32_t n_seqs, const int8_t* tokens, const int64_t* tok_off, const hist_quickalign_scores* scores,
                                const hist_guide_params* params, int32_t* residue_col, int32_t* n_cols, int32_t* n_edges, int32_t* edge_rows,
                                double* edge_lp, int64_t edge_capacity, int32_t* mst_edges) {
  if (!ctx || n_seqs < 1 || !tokens || !tok_off || !scores || !params || !residue_col || !n_cols)
    return set_err(ctx, HIST_ERR_ARG, "hist_guide_align: bad arguments");
  const hist_guide_params& P = *params;
  if (P.sparse && (P.kmer_len < 1 || P.kmer_len > 32)) return set_err(ctx, HIST_ERR_ARG, "hist_guide_align: kmer_len %d out of range", P.kmer_len);
  const size_t n = (size_t)n_seqs;
  for (size_t s = 0; s < n; ++s)
    if (tok_off[s + 1] - tok_off[s] < 1 || tok_off[s + 1] - tok_off[s] > (1 << 30)) return set_err(ctx, HIST_ERR_ARG, "hist_guide_align: sequence %zu is empty or too long", s);
  if (n_edges) *n_edges = 0;
  if (n == 1) {   // nothing to align: the sequence is its own guide
    for (int64_t q = 0; q < tok_off[1]; ++q) residue_col[q] = (int32_t)q;
    *n_cols = (int32_t)tok_off[1];
    return HIST_OK;
  }
  // ---- the trial edges (span.cpp:57-91)
  std::vector<std::pair<int, int>> trial;
  if (P.all_pairs) {
    for (size_t a = 0; a + 1 < n; ++a)
      for (size_t b = a + 1; b < n; ++b) trial.push_back({(int)a, (int)b});
  } else {
    std::mt19937 generator(P.seed);
    std::vector<std::set<size_t>> targets(n);
    GuidePartition part(n);
    const size_t want = std::min((size_t)(n * (n - 1) / 2), (size_t)std::ceil(std::log((double)n) * (double)n / std::log(2.0)));
    std::uniform_int_distribution<size_t> dist(0, n - 1);
    for (size_t e = 0; e < want || part.n_sets > 1; ++e) {
      size_t src, dest;
      do {
        src = dist(generator);
        dest = dist(generator);
        if (dest < src) std::swap(src, dest);
      } while (src == dest || targets[src].count(dest));
      targets[src].insert(dest);
      trial.push_back({(int)src, (int)dest});
      part.merge(src, dest);
    }
  }
  const size_t E = trial.size();
  // ---- envelopes + pairwise alignments on the device, in batches bounded by their traceback planes (span.cpp:93-117)
  std::vector<double> lp(E);
  std::vector<std::vector<uint8_t>> path(E);
  const size_t tb_budget = (size_t)3 << 30;
  size_t at = 0;
  while (at < E) {
    size_t end = at, tb = 0;
    std::vector<std::vector<int>> diags;
    while (end < E) {
      const int a = trial[end].first, b = trial[end].second;
      const int xl = (int)(tok_off[a + 1] - tok_off[a]), yl = (int)(tok_off[b + 1] - tok_off[b]);
      std::vector<int> d;
      envelope_diagonals(tokens + tok_off[a], xl, tokens + tok_off[b], yl, scores->alph_size, P, d);
      const size_t need = (d.empty() ? (size_t)xl + yl - 1 : d.size()) * ((size_t)yl + 1);
      if (end > at && (tb + need > tb_budget || end - at >= 8192)) break;
      tb += need;
      diags.push_back(std::move(d));
      ++end;
    }
    const int nb = (int)(end - at);
    std::vector<hist_quickalign_pair> qp(nb);
    std::vector<int64_t> poff(nb);
    int64_t ptot = 0;
    for (int k = 0; k < nb; ++k) {
      const int a = trial[at + k].first, b = trial[at + k].second;
      qp[k].x_len = (int32_t)(tok_off[a + 1] - tok_off[a]);
      qp[k].y_len = (int32_t)(tok_off[b + 1] - tok_off[b]);
      qp[k].x_tok = tokens + tok_off[a];
      qp[k].y_tok = tokens + tok_off[b];
      qp[k].n_diagonals = diags[k].empty() ? qp[k].x_len + qp[k].y_len - 1 : (int32_t)diags[k].size();
      qp[k].diagonals = diags[k].empty() ? nullptr : diags[k].data();
      poff[k] = ptot;
      ptot += (int64_t)qp[k].x_len + qp[k].y_len;
    }
    std::vector<uint8_t> pbuf((size_t)ptot);
    std::vector<int32_t> plen(nb);
    const int rc = hist_quickalign_batch(ctx, nb, qp.data(), scores, lp.data() + at, nullptr, nullptr, pbuf.data(), poff.data(), plen.data());
    if (rc) return rc;
    for (int k = 0; k < nb; ++k) {
      if (!(lp[at + k] > -INFINITY))   // QuickAlignMatrix::alignPath requires a finite score (quickalign.cpp:148)
        return set_err(ctx, HIST_ERR_NUMERIC, "hist_guide_align: sequences %d and %d have no alignment inside their envelope", trial[at + k].first, trial[at + k].second);
      path[at + k].assign(pbuf.begin() + poff[k], pbuf.begin() + poff[k] + plen[k]);
    }
    at = end;
  }
  if (n_edges) *n_edges = (int32_t)E;
  for (size_t e = 0; e < E && (int64_t)e < edge_capacity; ++e) {
    if (edge_rows) { edge_rows[2 * e] = trial[e].first; edge_rows[2 * e + 1] = trial[e].second; }
    if (edge_lp) edge_lp[e] = lp[e];
  }
  // ---- the spanning tree (span.cpp:119-144)
  std::vector<std::priority_queue<GuideEdge>> edges(n);
  for (size_t e = 0; e < E; ++e) {
    const GuideEdge g{trial[e].first, trial[e].second, lp[e], (int)e};
    edges[(size_t)g.row1].push(g);
    edges[(size_t)g.row2].push(g);
  }
  std::vector<int> tree;
  GuidePartition part(n);
  while (part.n_sets > 1) {
    GuideEdge best{0, 0, -INFINITY, -1};
    bool found = false;
    for (size_t src : part.sets.front()) {
      auto& q = edges[src];
      while (!q.empty() && part.same((size_t)q.top().row1, (size_t)q.top().row2)) q.pop();
      if (!q.empty() && (!found || best < q.top())) { best = q.top(); found = true; }
    }
    if (!found) return set_err(ctx, HIST_ERR_STATE, "hist_guide_align: found no valid edge");
    tree.push_back(best.path);
    part.merge((size_t)best.row1, (size_t)best.row2);
  }
  if (mst_edges)
    for (size_t k = 0; k < tree.size(); ++k) mst_edges[k] = tree[k];
  // ---- merge (span.cpp:146-150)
  std::vector<MergeAlign> al(tree.size());
  for (size_t k = 0; k < tree.size(); ++k) {
    const int e = tree[k];
    const int64_t xo = tok_off[trial[e].first], yo = tok_off[trial[e].second];
    int64_t i = 0, j = 0;
    MergeAlign& a = al[k];
    a.col_off.reserve(path[e].size() + 1);
    a.col_off.push_back(0);
    for (uint8_t c : path[e]) {
      if (c & 1) a.res.push_back(xo + i++);
      if (c & 2) a.res.push_back(yo + j++);
      a.col_off.push_back((int)a.res.size());
    }
  }
  std::vector<int> rc;
  const long long cols = merge_alignments(al, tok_off[n], rc);
  if (cols < 0) return set_err(ctx, HIST_ERR_STATE, "hist_guide_align: inconsistent pairwise alignments");
  for (int64_t r = 0; r < tok_off[n]; ++r) residue_col[r] = rc[(size_t)r];
  *n_cols = (int32_t)cols;
  return HIST_OK;
}

extern "C" int hist_diagonal_envelope(hist_ctx* ctx, const int8_t* x_tok, int32_t x_len, const int8_t* y_tok, int32_t y_len, int32_t alph_size,
                                      const hist_guide_params* params, int32_t* diagonals, int32_t capacity, int32_t* n_diagonals) {
  if (!x_tok || !y_tok || x_len < 1 || y_len < 1 || alph_size < 1 || !params || !n_diagonals)
    return set_err(ctx, HIST_ERR_ARG, "hist_diagonal_envelope: bad arguments");
  if (params->sparse && (params->kmer_len < 1 || params->kmer_len > 32)) return set_err(ctx, HIST_ERR_ARG, "hist_diagonal_envelope: kmer_len %d out of range", params->kmer_len);
  std::vector<int> d;
  envelope_diagonals(x_tok, x_len, y_tok, y_len, alph_size, *params, d);
  if (d.empty()) {   // the full envelope
    d.resize((size_t)x_len + y_len - 1);
    std::iota(d.begin(), d.end(), 1 - y_len);
  }
  *n_diagonals = (int32_t)d.size();
  if (!diagonals || capacity < *n_diagonals) return set_err(ctx, HIST_ERR_CAPACITY, "hist_diagonal_envelope: need room for %d diagonals", *n_diagonals);
  for (size_t k = 0; k < d.size(); ++k) diagonals[k] = d[k];
  return HIST_OK;
}
