// historian_b200: level-batched, device-resident progressive reconstruction with a native host.
//
// Replaces the node loop of Reconstructor::reconstruct (reference recon.cpp:917-1052) together with the
// bookkeeping half of ForwardMatrix::makeProfile (forward.cpp:686-843: state numbering, seqCoords :533-538,
// alignPaths :540-578, transitions :808-831), Profile::addReadyStates (profile.cpp:268-319),
// ForwardMatrix::sampleProfile / bestProfile (forward.cpp:845-895), GuideAlignmentEnvelope
// (alignpath.cpp:282-310) and traceAlignPath (forward.cpp:654-684).
//
// All sibling merges of one guide-tree level go through ONE call: batched DPMatrix constructors, one fill
// launch, one traceback launch (one warp per merge), one batched elimination, host assembly of the parents on
// the worker pool, and ONE gather kernel that writes every parent's lpAbsorb table straight into a new device
// arena from the pairs' leftMultiply results -- profile tables never visit the host, and a parent is handed to
// the next level's hist_pair_prepare_batch as it is.  What does cross the bus per merge is the integer
// structure of the parent (cells, transitions: a few KB).
//
// AlignPaths are not materialised per state as the reference does (a map of vector<bool> per state and
// transition, concatenated at every elimination step): every profile keeps a small expression store (one
// column node per state, one item list per transition, referring to the children's stores), and only the
// root's best trace is ever expanded into alignment columns.
#include <algorithm>
#include <array>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <string>
#include <vector>

#include "hist_internal.cuh"

using namespace hb;

namespace {

// one alignment column = rows of child column a (store of child 0) + rows of child column b (child 1) + own row
struct ColNode {
  int a, b, row;
};
struct SeqNode {
  int first, count;   // items [first, first + count)
};
// item = (id << 2) | tag;  tag 0: column `id` of this store, 1: sequence `id` of child 0's store, 2: of child 1's
enum { IT_COL = 0, IT_SEQ0 = 1, IT_SEQ1 = 2 };

struct PathStore {
  int child_node[2] = {-1, -1};
  bool full = true;             // false: imported with a profile from another rank (state_col / trans_seq only)
  std::vector<ColNode> cols;
  std::vector<SeqNode> seqs;
  std::vector<int> items;
  std::vector<int> state_col;   // [n_states] column id or -1 (empty path)
  std::vector<int> trans_seq;   // [n_trans] sequence id or -1
};

struct NodeRec {
  hist_profile* prof = nullptr;       // owned until its parent has been built
  PathStore* store = nullptr;         // kept until the reconstruction is destroyed
  std::vector<int> in_off, in_idx;    // transitions by destination (Profile::getTrans)
};

struct GuideRow {
  std::vector<uint8_t> bits;      // [cols]
  std::vector<int> pos_to_col;    // [len + 1]
};

void build_in_lists(const hist_profile* p, std::vector<int>& in_off, std::vector<int>& in_idx) {
  in_off.assign(p->n + 1, 0);
  for (int t = 0; t < p->nt; ++t) ++in_off[p->t_dst[t] + 1];
  for (int i = 0; i < p->n; ++i) in_off[i + 1] += in_off[i];
  in_idx.assign(std::max(1, p->nt), 0);
  std::vector<int> cur(in_off.begin(), in_off.end() - 1);
  for (int t = 0; t < p->nt; ++t) in_idx[cur[p->t_dst[t]]++] = t;
}

struct Timer {
  std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
  double lap() {
    const auto t1 = std::chrono::steady_clock::now();
    const double ms = std::chrono::duration<double, std::milli>(t1 - t0).count();
    t0 = t1;
    return ms;
  }
};

}  // namespace

struct hist_recon {
  hist_ctx* ctx = nullptr;
  int n_nodes = 0;
  std::string alphabet;
  int C = 1;
  std::vector<NodeRec> node;
  // guide alignment (leaf rows only)
  int g_cols = 0;
  std::vector<int> g_row_of_node;   // [n_nodes] guide row or -1
  std::vector<GuideRow> g_rows;
  // result
  int root = -1;
  std::vector<int> root_items;      // items of the root's best alignment path (tags as seen from the root's store)
  int aln_cols = -1;
  std::vector<std::pair<const PathStore*, int>> aln_roots;   // [aln_cols] the column trees of the root's alignment path
  double stage_ms[HIST_RECON_NSTAGES] = {0};
};

namespace {

// ---------------------------------------------------------------- assembly of one parent (host)
inline bool emit_or_start(const hist_profile* p, const NodeRec* r, int i) {   // profile.h:34-35
  return p->is_emit[i] || r->in_off[i + 1] == r->in_off[i];
}

inline bool cell_absorbing(const hist_profile* x, const hist_profile* y, const hist_cell& c) {   // forward.cpp:470-474
  if (c.state == HIST_EEE) return false;
  const bool xe = x->is_emit[c.xpos], ye = y->is_emit[c.ypos];
  return (c.state == HIST_IMM && xe && ye) || (c.state == HIST_IMD && xe) || (c.state == HIST_IDM && ye);
}

// Profile::getTrans(src, dest): index of the transition, -1 if absent
inline int trans_between(const hist_profile* p, const NodeRec* r, int s, int d) {
  for (int q = r->in_off[d]; q < r->in_off[d + 1]; ++q)
    if (p->t_src[r->in_idx[q]] == s) return r->in_idx[q];
  return -1;
}

struct Assembler {
  const hist_profile* x;
  const hist_profile* y;
  const NodeRec* xr;
  const NodeRec* yr;
  int parent_row;
  PathStore* st;

  // cellAlignPath (forward.cpp:540-568) as a column node of the parent's store; -1 for an empty path
  int cell_col(const hist_cell& c) {
    int a = -1, b = -1;
    const int xs = xr->store->state_col[c.xpos], ys = c.state == HIST_EEE ? -1 : yr->store->state_col[c.ypos];
    switch (c.state) {
      case HIST_IMM:
        if (x->is_emit[c.xpos] && y->is_emit[c.ypos]) { a = xs; b = ys; }
        else if (emit_or_start(x, xr, c.xpos)) b = ys;
        else a = xs;
        break;
      case HIST_IMD: case HIST_IIW: a = xs; break;
      case HIST_IDM: case HIST_IMI: b = ys; break;
      default: break;
    }
    if (c.state == HIST_EEE) a = -1;
    const int row = cell_absorbing(x, y, c) ? parent_row : -1;
    if (a < 0 && b < 0 && row < 0) return -1;
    st->cols.push_back({a, b, row});
    return (int)st->cols.size() - 1;
  }
  // transitionAlignPath (forward.cpp:570-577): appends the items; false if a transition is missing
  bool trans_items(const hist_cell& s, const hist_cell& d, std::vector<int>& items) {
    if (s.xpos != d.xpos) {
      const int t = trans_between(x, xr, s.xpos, d.xpos);
      if (t < 0) return false;
      const int q = xr->store->trans_seq[t];
      if (q >= 0) items.push_back((q << 2) | IT_SEQ0);
    }
    if (s.ypos != d.ypos) {
      const int t = trans_between(y, yr, s.ypos, d.ypos);
      if (t < 0) return false;
      const int q = yr->store->trans_seq[t];
      if (q >= 0) items.push_back((q << 2) | IT_SEQ1);
    }
    return true;
  }
};

struct Assembled {
  hist_profile* prof = nullptr;
  PathStore* store = nullptr;
  std::vector<hist_cell> row_cell;    // [n_new] cell whose absorb vector is the state's row; state < 0 for null states
  std::vector<int> root_items;        // root only
  int rc = 0;
};

// ForwardMatrix::makeProfile's assembly from the device elimination, then Profile::addReadyStates.
void assemble_parent(hist_ctx* ctx, const hist_pair* pair, const NodeRec* xr, const NodeRec* yr, int parent_row, bool closest_in_x,
                     const ElimOut& e, int x_node, int y_node, Assembled& out) {
  const hist_profile* x = pair->x;
  const hist_profile* y = pair->y;
  PathStore* st = new PathStore();
  st->child_node[0] = x_node;
  st->child_node[1] = y_node;
  Assembler as{x, y, xr, yr, parent_row, st};
  const int ns = e.n_states, nt = e.n_trans;
  std::vector<uint8_t> is_emit(ns);
  std::vector<int> col_old(ns);
  st->cols.reserve(ns + 16);
  for (int k = 0; k < ns; ++k) {
    is_emit[k] = cell_absorbing(x, y, e.state_cell[k]) ? 1 : 0;
    col_old[k] = as.cell_col(e.state_cell[k]);
  }
  // transitions: tap(src,c1) cap(c1) tap(c1,c2) ... tap(c_last,dest) along the best path (forward.cpp:764, 797)
  std::vector<int> seq_old(nt, -1);
  std::vector<int> items;
  for (int t = 0; t < nt; ++t) {
    items.clear();
    hist_cell prev = e.state_cell[e.t_src[t]];
    for (long long v = e.via_off[t]; v < e.via_off[t + 1]; ++v) {
      const hist_cell& c = e.via[v];
      if (!as.trans_items(prev, c, items)) out.rc = HIST_ERR_STATE;
      const int col = as.cell_col(c);
      if (col >= 0) items.push_back((col << 2) | IT_COL);
      prev = c;
    }
    if (!as.trans_items(prev, e.state_cell[e.t_dst[t]], items)) out.rc = HIST_ERR_STATE;
    if (!items.empty()) {
      st->seqs.push_back({(int)st->items.size(), (int)items.size()});
      st->items.insert(st->items.end(), items.begin(), items.end());
      seq_old[t] = (int)st->seqs.size() - 1;
    }
  }
  // ---- addReadyStates (profile.cpp:268-319)
  std::vector<uint8_t> has_null(ns, 0), has_abs(ns, 0);
  for (int t = 0; t < nt; ++t) (is_emit[e.t_dst[t]] ? has_abs : has_null)[e.t_src[t]] = 1;
  std::vector<int> old2new(ns), ready_of(ns, -1);
  int n_new = 0;
  for (int s = 0; s < ns; ++s) {
    old2new[s] = n_new++;
    if (has_null[s] && has_abs[s]) ready_of[s] = n_new++;
  }
  const int n_ready = n_new - ns;
  hist_profile* p = new hist_profile();
  p->ctx = ctx;
  p->n = n_new;
  p->nt = nt + n_ready;
  p->A = x->A;
  p->C = x->C;
  p->is_emit.assign(n_new, 0);
  p->t_src.resize(p->nt);
  p->t_dst.resize(p->nt);
  p->t_lp.resize(p->nt);
  const bool have_leaf = !x->leaf_pos.empty() && !y->leaf_pos.empty();
  if (have_leaf) p->leaf_pos.assign(n_new, 0);
  st->state_col.assign(n_new, -1);
  st->trans_seq.assign(p->nt, -1);
  out.row_cell.assign(n_new, hist_cell{0, 0, -1});
  for (int s = 0; s < ns; ++s) {
    const int k = old2new[s];
    const hist_cell& c = e.state_cell[s];
    p->is_emit[k] = is_emit[s];
    st->state_col[k] = col_old[s];
    if (is_emit[s]) out.row_cell[k] = c;
    if (have_leaf) {
      // cellSeqCoords (forward.cpp:533-538), restricted to the closest leaf of the parent (forward.cpp:36-42)
      const int pos = closest_in_x ? x->leaf_pos[c.xpos] : y->leaf_pos[c.ypos];
      p->leaf_pos[k] = pos;
      if (ready_of[s] >= 0) p->leaf_pos[ready_of[s]] = pos;
    }
  }
  for (int t = 0; t < nt; ++t) {
    const int s = e.t_src[t], d = e.t_dst[t];
    p->t_src[t] = (ready_of[s] >= 0 && is_emit[d]) ? ready_of[s] : old2new[s];   // absorbing out-transitions move to the ready state
    p->t_dst[t] = old2new[d];
    p->t_lp[t] = e.t_lp[t];
    st->trans_seq[t] = seq_old[t];
  }
  {
    int t = nt;
    for (int s = 0; s < ns; ++s)   // ready transitions are appended in state order (profile.cpp:283, 299)
      if (ready_of[s] >= 0) {
        p->t_src[t] = old2new[s];
        p->t_dst[t] = ready_of[s];
        p->t_lp[t] = 0.0;
        ++t;
      }
  }
  p->sorted_in = true;
  {
    std::vector<int> last(p->n, -1);
    for (int t = 0; t < p->nt; ++t) {
      if (p->t_src[t] <= last[p->t_dst[t]]) p->sorted_in = false;
      last[p->t_dst[t]] = p->t_src[t];
    }
  }
  p->empty = true;
  for (int i = 0; i < p->n; ++i)
    if (p->is_emit[i]) p->empty = false;
  p->chain = false;
  p->d_lp_absorb = nullptr;
  p->arena = nullptr;
  out.prof = p;
  out.store = st;
}

// traceAlignPath (forward.cpp:654-684) as an item list relative to the parent's store
int trace_items(const hist_pair* pair, const NodeRec* xr, const NodeRec* yr, int parent_row, PathStore* st, const std::vector<hist_cell>& tr,
                std::vector<int>& items) {
  Assembler as{pair->x, pair->y, xr, yr, parent_row, st};
  for (size_t n = 0; n + 1 < tr.size(); ++n) {
    const int col = as.cell_col(tr[n]);
    if (col >= 0) items.push_back((col << 2) | IT_COL);
    if (!as.trans_items(tr[n], tr[n + 1], items)) return HIST_ERR_STATE;
  }
  const int col = as.cell_col(tr.back());
  if (col >= 0) items.push_back((col << 2) | IT_COL);
  return HIST_OK;
}

// The columns of the root's alignment path, in order: (store, column node) roots of the trees whose rows a column holds.
// Sequential (the item lists nest), but short: one entry per column.
void collect_columns(hist_recon* R) {
  struct Frame { const PathStore* st; int first, count, pos; };   // walking items [first, first+count) of store st
  const PathStore* root = R->node[R->root].store;
  std::vector<Frame> stack;
  std::vector<std::pair<const PathStore*, int>>& cols = R->aln_roots;
  cols.clear();
  // the root's items live outside its store: treat them as a virtual sequence
  const std::vector<int>& top = R->root_items;
  auto child_store = [&](const PathStore* st, int which) -> const PathStore* {
    const int cn = st->child_node[which];
    return cn >= 0 ? R->node[cn].store : nullptr;
  };
  for (size_t q = 0; q < top.size(); ++q) {
    const int it = top[q], id = it >> 2, tag = it & 3;
    if (tag == IT_COL) { cols.push_back({root, id}); continue; }
    const PathStore* cs = child_store(root, tag == IT_SEQ0 ? 0 : 1);
    stack.clear();
    stack.push_back({cs, cs->seqs[id].first, cs->seqs[id].count, 0});
    while (!stack.empty()) {
      Frame& f = stack.back();
      if (f.pos == f.count) { stack.pop_back(); continue; }
      const int it2 = f.st->items[f.first + f.pos++];
      const int id2 = it2 >> 2, tag2 = it2 & 3;
      if (tag2 == IT_COL) cols.push_back({f.st, id2});
      else {
        const PathStore* c2 = child_store(f.st, tag2 == IT_SEQ0 ? 0 : 1);
        stack.push_back({c2, c2->seqs[id2].first, c2->seqs[id2].count, 0});
      }
    }
  }
  R->aln_cols = (int)cols.size();
}

// Expands the columns into the caller's dense [n_nodes][cols] 0/1 matrix: rows are cleared, then blocks of 64 columns (a
// cache line of every row) walk their column trees, both over the host threads.
void expand_alignment(const hist_recon* R, uint8_t* bits) {
  const int n_cols = R->aln_cols, n_nodes = R->n_nodes;
  if (n_cols <= 0) return;
  const int row_grain = std::max(1, (int)((1 << 20) / n_cols));
  host_parallel_for((n_nodes + row_grain - 1) / row_grain, 1, [&](int b) {
    const int r0 = b * row_grain, r1 = std::min(n_nodes, r0 + row_grain);
    memset(bits + (size_t)r0 * n_cols, 0, (size_t)(r1 - r0) * n_cols);
  });
  const int blk = 64;
  host_parallel_for((n_cols + blk - 1) / blk, 1, [&](int b) {
    std::vector<std::pair<const PathStore*, int>> cstack;
    const int c1 = std::min(n_cols, (b + 1) * blk);
    for (int c = b * blk; c < c1; ++c) {
      cstack.clear();
      cstack.push_back(R->aln_roots[c]);
      while (!cstack.empty()) {
        const auto cur = cstack.back();
        cstack.pop_back();
        const ColNode& cn = cur.first->cols[cur.second];
        if (cn.row >= 0 && cn.row < n_nodes) bits[(size_t)cn.row * n_cols + c] = 1;
        for (int w = 0; w < 2; ++w) {
          const int id = w ? cn.b : cn.a;
          if (id < 0) continue;
          const int ch = cur.first->child_node[w];
          cstack.push_back({R->node[ch].store, id});
        }
      }
    }
  });
}

struct EnvHost {
  std::vector<int> cum;
  hist_env_desc d;
};

int make_env(hist_recon* R, int leaf1, int leaf2, int max_dist, EnvHost& e) {
  memset(&e.d, 0, sizeof(e.d));
  e.d.max_distance = max_dist;
  if (max_dist < 0) return HIST_OK;
  const int r1 = R->g_row_of_node[leaf1], r2 = R->g_row_of_node[leaf2];
  if (r1 < 0 || r2 < 0) return set_err(R->ctx, HIST_ERR_ARG, "hist_recon_level: closest leaf %d / %d has no guide row", leaf1, leaf2);
  const GuideRow& a = R->g_rows[r1];
  const GuideRow& b = R->g_rows[r2];
  // GuideAlignmentEnvelope (alignpath.cpp:282-310)
  e.cum.resize(R->g_cols + 1);
  e.cum[0] = 0;
  for (int c = 0; c < R->g_cols; ++c) e.cum[c + 1] = e.cum[c] + (a.bits[c] & b.bits[c]);
  e.d.n_cols = R->g_cols;
  e.d.cumulative_matches = e.cum.data();
  e.d.row1_len = (int)a.pos_to_col.size() - 1;
  e.d.row2_len = (int)b.pos_to_col.size() - 1;
  e.d.row1_pos_to_col = a.pos_to_col.data();
  e.d.row2_pos_to_col = b.pos_to_col.data();
  return HIST_OK;
}

inline bool cell_less(const hist_cell& a, const hist_cell& b) {
  return a.xpos != b.xpos ? a.xpos < b.xpos : a.ypos != b.ypos ? a.ypos < b.ypos : a.state < b.state;
}

}  // namespace

// ================================================================== C ABI
extern "C" int hist_recon_create(hist_ctx* ctx, int32_t n_nodes, const char* alphabet, int32_t components, hist_recon** out) {
  if (!ctx || !out || n_nodes < 1 || !alphabet || components < 1) return set_err(ctx, HIST_ERR_ARG, "hist_recon_create: bad arguments");
  hist_recon* R = new hist_recon();
  R->ctx = ctx;
  R->n_nodes = n_nodes;
  R->alphabet = alphabet;
  R->C = components;
  R->node.resize(n_nodes);
  R->g_row_of_node.assign(n_nodes, -1);
  *out = R;
  return HIST_OK;
}

extern "C" void hist_recon_destroy(hist_recon* R) {
  if (!R) return;
  std::vector<hist_profile*> profs;
  for (auto& nr : R->node) {
    if (nr.prof) profs.push_back(nr.prof);
    delete nr.store;
  }
  if (!profs.empty()) hist_profiles_destroy(profs.data(), (int32_t)profs.size());
  delete R;
}

extern "C" int hist_recon_set_guide(hist_recon* R, int32_t n_rows, int32_t n_cols, const int32_t* row_node, const uint8_t* bits) {
  if (!R || n_rows < 0 || n_cols < 0 || (n_rows && (!row_node || !bits))) return set_err(R ? R->ctx : nullptr, HIST_ERR_ARG, "hist_recon_set_guide: bad arguments");
  R->g_cols = n_cols;
  R->g_rows.assign(n_rows, GuideRow());
  std::fill(R->g_row_of_node.begin(), R->g_row_of_node.end(), -1);
  for (int r = 0; r < n_rows; ++r) {
    if (row_node[r] < 0 || row_node[r] >= R->n_nodes) return set_err(R->ctx, HIST_ERR_ARG, "hist_recon_set_guide: row %d names node %d", r, row_node[r]);
    R->g_row_of_node[row_node[r]] = r;
    GuideRow& g = R->g_rows[r];
    g.bits.assign(bits + (size_t)r * n_cols, bits + (size_t)(r + 1) * n_cols);
    g.pos_to_col.assign(1, 0);
    for (int c = 0; c < n_cols; ++c) {
      g.bits[c] = g.bits[c] ? 1 : 0;
      if (g.bits[c]) g.pos_to_col.push_back(c + 1);
    }
  }
  return HIST_OK;
}

extern "C" int hist_recon_add_leaves(hist_recon* R, int32_t n, const int32_t* nodes, const char* const* residues, const int32_t* lengths) {
  if (!R || n <= 0 || !nodes || !residues || !lengths) return set_err(R ? R->ctx : nullptr, HIST_ERR_ARG, "hist_recon_add_leaves: bad arguments");
  for (int k = 0; k < n; ++k)
    if (nodes[k] < 0 || nodes[k] >= R->n_nodes || R->node[nodes[k]].store)
      return set_err(R->ctx, HIST_ERR_ARG, "hist_recon_add_leaves: node %d out of range or already built", nodes[k]);
  std::vector<hist_profile*> profs(n, nullptr);
  int rc = hist_profile_from_residues(R->ctx, n, residues, lengths, R->alphabet.c_str(), R->C, profs.data());
  if (rc) return rc;
  host_parallel_for(n, 32, [&](int k) {
    NodeRec& nr = R->node[nodes[k]];
    nr.prof = profs[k];
    // Profile(components, alphabet, FastSeq, rowIndex) (profile.cpp:23-76): state pos+1 carries one column of the leaf row
    PathStore* st = new PathStore();
    const int L = lengths[k];
    st->cols.resize(L);
    st->state_col.assign(L + 2, -1);
    for (int q = 0; q < L; ++q) {
      st->cols[q] = {-1, -1, nodes[k]};
      st->state_col[q + 1] = q;
    }
    st->trans_seq.assign(L + 1, -1);
    nr.store = st;
    build_in_lists(nr.prof, nr.in_off, nr.in_idx);
  });
  return HIST_OK;
}

extern "C" hist_profile* hist_recon_profile(hist_recon* R, int32_t node) {
  if (!R || node < 0 || node >= R->n_nodes) return nullptr;
  return R->node[node].prof;
}

extern "C" int hist_recon_level(hist_recon* R, int32_t n, const hist_merge_desc* merges, const hist_recon_params* prm, double* lp_end_out) {
  if (!R || n <= 0 || !merges || !prm) return set_err(R ? R->ctx : nullptr, HIST_ERR_ARG, "hist_recon_level: bad arguments");
  hist_ctx* ctx = R->ctx;
  cudaSetDevice(ctx->device);
  Timer tm;
  const bool have_guide = prm->band >= 0 && !R->g_rows.empty();
  const bool sampling_mode = prm->min_post_prob <= 0;
  if (!sampling_mode) return set_err(ctx, HIST_ERR_ARG, "hist_recon_level: posterior profiles (-profminpost) go through hist_posterior_cells / hist_profile_eliminate");
  if (!(prm->strategy & HIST_INCLUDE_BEST_TRACE) && prm->profile_samples <= 0)
    return set_err(ctx, HIST_ERR_ARG, "Must allow at least one sample path in the profile");   // forward.cpp:848
  if (prm->strategy & (HIST_COUNT_SUBST | HIST_COUNT_INDEL))
    return set_err(ctx, HIST_ERR_ARG, "hist_recon_level: Count* strategies are not carried through makeProfile");
  std::vector<const hist_profile*> xs(n), ys(n);
  std::vector<const hist_pairhmm_desc*> hmms(n);
  for (int k = 0; k < n; ++k) {
    const hist_merge_desc& m = merges[k];
    if (m.node < 0 || m.node >= R->n_nodes || m.left < 0 || m.left >= R->n_nodes || m.right < 0 || m.right >= R->n_nodes || !m.hmm)
      return set_err(ctx, HIST_ERR_ARG, "hist_recon_level: merge %d names a node out of range", k);
    if (!R->node[m.left].prof || !R->node[m.right].prof) return set_err(ctx, HIST_ERR_STATE, "hist_recon_level: children of node %d are not built", m.node);
    if (R->node[m.node].store) return set_err(ctx, HIST_ERR_STATE, "hist_recon_level: node %d is already built", m.node);
    if (m.closest != m.closest_left && m.closest != m.closest_right)
      return set_err(ctx, HIST_ERR_ARG, "hist_recon_level: closest leaf of node %d is neither child's", m.node);
    xs[k] = R->node[m.left].prof;
    ys[k] = R->node[m.right].prof;
    hmms[k] = m.hmm;
  }
  // ---- DPMatrix constructors + Forward fills of the level, with the band-widening retry (recon.cpp:954-975)
  std::vector<EnvHost> envs(n);
  std::vector<const hist_env_desc*> envp(n, nullptr);
  std::vector<int> max_dist(n, have_guide ? prm->band : -1);
  std::vector<hist_pair*> pairs(n, nullptr);
  std::vector<double> lp_end(n, -INFINITY);
  auto fail = [&](int rc) {
    hist_pairs_destroy(pairs.data(), n);
    return rc;
  };
  {
    std::vector<int> rcs(n, 0);
    host_parallel_for(n, 16, [&](int k) {
      rcs[k] = make_env(R, merges[k].closest_left, merges[k].closest_right, max_dist[k], envs[k]);
      envp[k] = &envs[k].d;
    });
    for (int k = 0; k < n; ++k)
      if (rcs[k]) return rcs[k];
  }
  R->stage_ms[HIST_RECON_ENV] += tm.lap();
  int rc = hist_pair_prepare_batch(ctx, n, xs.data(), ys.data(), hmms.data(), have_guide ? envp.data() : nullptr, 0, pairs.data());
  if (rc) return rc;
  R->stage_ms[HIST_RECON_PREPARE] += tm.lap();
  rc = hist_fill_batch(ctx, pairs.data(), n, 0, lp_end.data(), nullptr);
  if (rc) return fail(rc);
  R->stage_ms[HIST_RECON_FILL] += tm.lap();
  for (int k = 0; k < n; ++k) {
    while (!(lp_end[k] > -INFINITY)) {
      if (max_dist[k] < 0) return fail(set_err(ctx, HIST_ERR_NUMERIC, "Zero forward likelihood even in the absence of guide alignment constraints (node %d)", merges[k].node));
      if (max_dist[k] * 2 > R->g_cols) max_dist[k] = -1;
      else if (max_dist[k] == 0) max_dist[k] = 1;
      else max_dist[k] *= 2;
      hist_pair_destroy(pairs[k]);
      pairs[k] = nullptr;
      rc = make_env(R, merges[k].closest_left, merges[k].closest_right, max_dist[k], envs[k]);
      if (!rc) rc = hist_pair_prepare(ctx, xs[k], ys[k], hmms[k], &envs[k].d, &pairs[k]);
      if (!rc) rc = hist_fwd_fill(pairs[k], &lp_end[k]);
      if (rc) return fail(rc);
    }
    if (lp_end_out) lp_end_out[k] = lp_end[k];
  }
  R->stage_ms[HIST_RECON_RETRY] += tm.lap();
  // ---- which cells make each parent: bestProfile / sampleProfile (forward.cpp:845-895)
  hist_cell* d_paths = nullptr;
  std::vector<long long> path_off, path_len;
  const bool want_best = (prm->strategy & HIST_INCLUDE_BEST_TRACE) != 0;
  std::vector<int> best_idx;   // merges whose best trace is needed: all with IncludeBestTrace, else only the root
  for (int k = 0; k < n; ++k)
    if (want_best || merges[k].node == prm->root_node) best_idx.push_back(k);
  std::vector<int> best_slot(n, -1);
  if (!best_idx.empty()) {
    std::vector<hist_pair*> bp(best_idx.size());
    for (size_t q = 0; q < best_idx.size(); ++q) { bp[q] = pairs[best_idx[q]]; best_slot[best_idx[q]] = (int)q; }
    rc = trace_best_batch(ctx, bp.data(), (int)bp.size(), &d_paths, path_off, path_len);
    if (rc) return fail(rc);
  }
  auto fail2 = [&](int rc2) {
    if (d_paths) cudaFreeAsync(d_paths, ctx->stream);
    return fail(rc2);
  };
  R->stage_ms[HIST_RECON_TRACE] += tm.lap();
  std::vector<ElimJob> jobs(n);
  std::vector<std::vector<hist_cell>> host_cells(n);
  std::vector<hist_cell> root_trace;
  for (int k = 0; k < n; ++k) {
    const hist_merge_desc& m = merges[k];
    const bool is_root = m.node == prm->root_node;
    const int slot = best_slot[k];
    ElimJob& J = jobs[k];
    J.pair = pairs[k];
    J.h_cells = nullptr;
    J.d_cells = nullptr;
    if (is_root) {
      // forward->bestAlignPath() and forward->bestProfile() with the default strategy (recon.cpp:1009-1010, forward.h:163)
      root_trace.resize((size_t)path_len[slot]);
      rc = check_cuda(ctx, cudaMemcpyAsync(root_trace.data(), d_paths + path_off[slot], root_trace.size() * sizeof(hist_cell), cudaMemcpyDeviceToHost, ctx->stream), "D2H root trace");
      if (!rc) rc = check_cuda(ctx, cudaStreamSynchronize(ctx->stream), "sync");
      if (rc) return fail2(rc);
      J.d_cells = d_paths + path_off[slot];
      J.n = (int)path_len[slot];
      J.strategy = HIST_COLLAPSE_CHAINS;
      continue;
    }
    J.strategy = prm->strategy;
    const long long best_len = want_best ? path_len[slot] : 0;
    const long long max_cells = prm->max_profile_states;
    const bool sample = prm->profile_samples > 0 && (max_cells == 0 || best_len < max_cells);
    if (!sample) {
      J.d_cells = d_paths + path_off[slot];
      J.n = (int)best_len;
      continue;
    }
    // sampleProfile: cell multiset of the best trace (count 2) and the sampled traces, drawn from the caller's generator
    if (!prm->mt_state) return fail2(set_err(ctx, HIST_ERR_ARG, "hist_recon_level: sampled profiles need mt_state"));
    std::map<std::array<int, 3>, int> count;
    auto key = [](const hist_cell& c) { return std::array<int, 3>{c.xpos, c.ypos, c.state}; };
    int n_traces = 0;
    if (want_best) {
      std::vector<hist_cell> best((size_t)best_len);
      rc = check_cuda(ctx, cudaMemcpyAsync(best.data(), d_paths + path_off[slot], best.size() * sizeof(hist_cell), cudaMemcpyDeviceToHost, ctx->stream), "D2H best trace");
      if (!rc) rc = check_cuda(ctx, cudaStreamSynchronize(ctx->stream), "sync");
      if (rc) return fail2(rc);
      for (auto& c : best) count[key(c)] = 2;   // avoid dropping these cells
      ++n_traces;
    }
    std::vector<hist_cell> paths;
    std::vector<long long> poff;
    std::vector<uint32_t> snap;
    rc = trace_sample_snap(pairs[k], prm->mt_state, prm->profile_samples, paths, poff, snap);
    if (rc) return fail2(rc);
    int accepted = 0, consumed = 0;
    while (accepted < prm->profile_samples && (max_cells == 0 || (long long)count.size() < max_cells)) {
      for (long long q = poff[consumed]; q < poff[consumed + 1]; ++q) ++count[key(paths[q])];
      ++consumed;
      ++n_traces;
      ++accepted;
    }
    if (consumed > 0) memcpy(prm->mt_state, &snap[(size_t)(consumed - 1) * 625], 625 * sizeof(uint32_t));
    const int threshold = (n_traces > 1 && max_cells > 0 && (long long)count.size() >= max_cells) ? 2 : 1;
    std::vector<hist_cell>& hc = host_cells[k];
    hc.reserve(count.size());
    for (auto& kv : count)
      if (kv.second >= threshold) hc.push_back({kv.first[0], kv.first[1], kv.first[2]});
    J.h_cells = hc.data();
    J.n = (int)hc.size();
  }
  R->stage_ms[HIST_RECON_SELECT] += tm.lap();
  // ---- numeric elimination of the level on device
  std::vector<ElimOut> elim;
  rc = eliminate_batch(ctx, jobs, elim);
  if (rc) return fail2(rc);
  R->stage_ms[HIST_RECON_ELIM] += tm.lap();
  // ---- parents: host assembly on the worker pool
  std::vector<Assembled> made(n);
  host_parallel_for(n, 4, [&](int k) {
    const hist_merge_desc& m = merges[k];
    assemble_parent(ctx, pairs[k], &R->node[m.left], &R->node[m.right], m.node, m.closest == m.closest_left, elim[k], m.left, m.right, made[k]);
    if (m.node == prm->root_node && !made[k].rc)
      made[k].rc = trace_items(pairs[k], &R->node[m.left], &R->node[m.right], m.node, made[k].store, root_trace, made[k].root_items);
  });
  auto drop_made = [&]() {
    for (auto& a : made) {
      delete a.prof;
      delete a.store;
    }
  };
  for (int k = 0; k < n; ++k)
    if (made[k].rc) {
      drop_made();
      return fail2(set_err(ctx, made[k].rc, "hist_recon_level: a child transition along a best path is missing (node %d)", merges[k].node));
    }
  R->stage_ms[HIST_RECON_ASSEMBLE] += tm.lap();
  // ---- lpAbsorb tables of all parents: one arena, one gather launch
  const int K = pairs[0]->dev.K;
  std::vector<long long> row_base(n + 1, 0);
  for (int k = 0; k < n; ++k) row_base[k + 1] = row_base[k] + made[k].prof->n;
  const long long total_rows = row_base[n];
  const size_t tab_bytes = ((size_t)total_rows * K * sizeof(double) + 255) & ~(size_t)255;
  const size_t rows_bytes = (size_t)total_rows * sizeof(GatherRow);
  hb_arena* arena = new hb_arena{nullptr, tab_bytes, 0};
  char* d_rows = nullptr;
  rc = check_cuda(ctx, cudaMallocFromPoolAsync((void**)&arena->dev, tab_bytes, ctx->pool, ctx->stream), "cudaMallocAsync(parent profiles)");
  if (!rc) rc = check_cuda(ctx, cudaMallocFromPoolAsync((void**)&d_rows, rows_bytes, ctx->pool, ctx->stream), "cudaMallocAsync(gather rows)");
  if (!rc) rc = ensure_staging_buf(ctx, rows_bytes, 0);
  if (!rc) {
    GatherRow* hr = (GatherRow*)ctx->staging;
    host_parallel_for(n, 8, [&](int k) {
      GatherRow* o = hr + row_base[k];
      const PairDev* P = pairs[k]->d_dev;
      const std::vector<hist_cell>& rc_ = made[k].row_cell;
      for (size_t q = 0; q < rc_.size(); ++q) o[q] = {P, rc_[q].xpos, rc_[q].ypos, rc_[q].state};
    });
    rc = check_cuda(ctx, cudaMemcpyAsync(d_rows, hr, rows_bytes, cudaMemcpyHostToDevice, ctx->stream), "H2D gather rows");
  }
  if (!rc) rc = gather_absorb_rows(ctx, (const GatherRow*)d_rows, total_rows, K, (double*)arena->dev);
  if (!rc) rc = check_cuda(ctx, cudaStreamSynchronize(ctx->stream), "sync after gather");   // the staging buffer is reused by the next call
  if (d_rows) cudaFreeAsync(d_rows, ctx->stream);
  if (rc) {
    if (arena->dev) cudaFreeAsync(arena->dev, ctx->stream);
    delete arena;
    drop_made();
    return fail2(rc);
  }
  R->stage_ms[HIST_RECON_GATHER] += tm.lap();
  for (int k = 0; k < n; ++k) {
    const hist_merge_desc& m = merges[k];
    hist_profile* p = made[k].prof;
    p->d_lp_absorb = (double*)arena->dev + (size_t)row_base[k] * K;
    p->arena = arena;
    ++arena->refs;
    NodeRec& nr = R->node[m.node];
    nr.prof = p;
    nr.store = made[k].store;
    if (m.node == prm->root_node) {
      R->root = m.node;
      R->root_items.swap(made[k].root_items);
      R->aln_cols = -1;
    }
  }
  host_parallel_for(n, 8, [&](int k) { build_in_lists(R->node[merges[k].node].prof, R->node[merges[k].node].in_off, R->node[merges[k].node].in_idx); });
  // ---- the level's matrices and children are done
  if (d_paths) cudaFreeAsync(d_paths, ctx->stream);
  hist_pairs_destroy(pairs.data(), n);
  std::vector<hist_profile*> kids;
  for (int k = 0; k < n; ++k)
    for (int c : {merges[k].left, merges[k].right}) {
      NodeRec& nr = R->node[c];
      if (nr.prof) kids.push_back(nr.prof);
      nr.prof = nullptr;
      std::vector<int>().swap(nr.in_off);
      std::vector<int>().swap(nr.in_idx);
    }
  hist_profiles_destroy(kids.data(), (int32_t)kids.size());
  R->stage_ms[HIST_RECON_CLEANUP] += tm.lap();
  return HIST_OK;
}

extern "C" int hist_recon_alignment(hist_recon* R, int32_t* n_cols, uint8_t* bits, int64_t capacity) {
  if (!R || !n_cols) return HIST_ERR_ARG;
  if (R->root < 0) return set_err(R->ctx, HIST_ERR_STATE, "hist_recon_alignment: the root has not been reconstructed");
  if (R->aln_cols < 0) collect_columns(R);
  *n_cols = R->aln_cols;
  const int64_t need = (int64_t)R->n_nodes * R->aln_cols;
  if (!bits || capacity < need) return set_err(R->ctx, HIST_ERR_CAPACITY, "hist_recon_alignment: need room for %d x %d bytes", R->n_nodes, R->aln_cols);
  if (need) expand_alignment(R, bits);
  return HIST_OK;
}

extern "C" int hist_recon_stage_ms(hist_recon* R, double* ms, int reset) {
  if (!R || !ms) return HIST_ERR_ARG;
  for (int k = 0; k < HIST_RECON_NSTAGES; ++k) {
    ms[k] = R->stage_ms[k];
    if (reset) R->stage_ms[k] = 0;
  }
  return HIST_OK;
}


// ------------------------------------------------------------------ exchange between ranks (one process per GPU)
// A profile moves as (host blob, device table): the blob holds the integer structure, the transition weights, the
// closest-leaf coordinates and the state/transition path ids its parent's assembly needs; the [n][C][A] lpAbsorb table
// stays in device memory (the caller moves it with ncclSend/Recv or a peer copy and hands the received copy to
// hist_recon_import_node).  The full path stores (needed only to expand the final alignment) travel separately, once.
namespace {
template <class T>
void put(std::vector<char>& b, const T* v, size_t n) {
  const char* p = (const char*)v;
  b.insert(b.end(), p, p + n * sizeof(T));
}
template <class T>
bool get(const char*& p, const char* end, T* v, size_t n) {
  if ((size_t)(end - p) < n * sizeof(T)) return false;
  memcpy(v, p, n * sizeof(T));
  p += n * sizeof(T);
  return true;
}
template <class T>
void put_vec(std::vector<char>& b, const std::vector<T>& v) {
  const long long n = (long long)v.size();
  put(b, &n, 1);
  put(b, v.data(), v.size());
}
template <class T>
bool get_vec(const char*& p, const char* end, std::vector<T>& v) {
  long long n = 0;
  if (!get(p, end, &n, 1) || n < 0 || (size_t)(end - p) < (size_t)n * sizeof(T)) return false;
  v.resize((size_t)n);
  return get(p, end, v.data(), (size_t)n);
}
int copy_out(hist_ctx* ctx, const std::vector<char>& b, void* buf, int64_t capacity, int64_t* bytes, const char* what) {
  *bytes = (int64_t)b.size();
  if (!buf || capacity < *bytes) return set_err(ctx, HIST_ERR_CAPACITY, "%s: need %lld bytes", what, (long long)b.size());
  memcpy(buf, b.data(), b.size());
  return HIST_OK;
}
}  // namespace

extern "C" int hist_recon_export_node(hist_recon* R, int32_t node, void* host_buf, int64_t capacity, int64_t* bytes, void* dev_out,
                                      int64_t* dev_bytes) {
  if (!R || !bytes || node < 0 || node >= R->n_nodes) return HIST_ERR_ARG;
  const NodeRec& nr = R->node[node];
  if (!nr.prof || !nr.store) return set_err(R->ctx, HIST_ERR_STATE, "hist_recon_export_node: node %d has no live profile", node);
  const hist_profile* p = nr.prof;
  std::vector<char> b;
  const int hdr[6] = {p->n, p->nt, p->A, p->C, p->leaf_pos.empty() ? 0 : 1, p->chain ? 1 : 0};
  put(b, hdr, 6);
  put(b, p->is_emit.data(), p->is_emit.size());
  put(b, p->t_src.data(), p->t_src.size());
  put(b, p->t_dst.data(), p->t_dst.size());
  put(b, p->t_lp.data(), p->t_lp.size());
  put(b, p->leaf_pos.data(), p->leaf_pos.size());
  put(b, nr.store->child_node, 2);
  put(b, nr.store->state_col.data(), nr.store->state_col.size());
  put(b, nr.store->trans_seq.data(), nr.store->trans_seq.size());
  const int64_t tb = (int64_t)p->n * p->A * p->C * (int64_t)sizeof(double);
  if (dev_bytes) *dev_bytes = tb;
  const int rc = copy_out(R->ctx, b, host_buf, capacity, bytes, "hist_recon_export_node");
  if (rc || !dev_out) return rc;
  cudaSetDevice(R->ctx->device);
  HB_CUDA(R->ctx, cudaMemcpyAsync(dev_out, p->d_lp_absorb, (size_t)tb, cudaMemcpyDeviceToDevice, R->ctx->stream));
  HB_CUDA(R->ctx, cudaStreamSynchronize(R->ctx->stream));   // the caller sends dev_out on a stream of its own
  return HIST_OK;
}

extern "C" int hist_recon_import_node(hist_recon* R, int32_t node, const void* host_buf, int64_t bytes, const void* dev_table) {
  if (!R || !host_buf || !dev_table || node < 0 || node >= R->n_nodes) return HIST_ERR_ARG;
  hist_ctx* ctx = R->ctx;
  NodeRec& nr = R->node[node];
  if (nr.prof) return set_err(ctx, HIST_ERR_STATE, "hist_recon_import_node: node %d already has a profile", node);
  const char* q = (const char*)host_buf;
  const char* end = q + bytes;
  int hdr[6];
  if (!get(q, end, hdr, 6) || hdr[0] < 2 || hdr[1] < 0) return set_err(ctx, HIST_ERR_ARG, "hist_recon_import_node: malformed blob");
  hist_profile* p = new hist_profile();
  PathStore* st = new PathStore();
  st->full = false;
  p->ctx = ctx;
  p->n = hdr[0]; p->nt = hdr[1]; p->A = hdr[2]; p->C = hdr[3];
  p->chain = hdr[5] != 0;
  p->is_emit.resize(p->n);
  p->t_src.resize(p->nt); p->t_dst.resize(p->nt); p->t_lp.resize(p->nt);
  if (hdr[4]) p->leaf_pos.resize(p->n);
  st->state_col.resize(p->n);
  st->trans_seq.resize(p->nt);
  bool ok = get(q, end, p->is_emit.data(), p->is_emit.size()) && get(q, end, p->t_src.data(), p->t_src.size()) &&
            get(q, end, p->t_dst.data(), p->t_dst.size()) && get(q, end, p->t_lp.data(), p->t_lp.size()) &&
            get(q, end, p->leaf_pos.data(), p->leaf_pos.size()) && get(q, end, st->child_node, 2) &&
            get(q, end, st->state_col.data(), st->state_col.size()) && get(q, end, st->trans_seq.data(), st->trans_seq.size());
  for (int t = 0; ok && t < p->nt; ++t) ok = p->t_src[t] >= 0 && p->t_src[t] < p->t_dst[t] && p->t_dst[t] < p->n;
  if (!ok) {
    delete p;
    delete st;
    return set_err(ctx, HIST_ERR_ARG, "hist_recon_import_node: malformed blob");
  }
  p->sorted_in = true;
  {
    std::vector<int> last(p->n, -1);
    for (int t = 0; t < p->nt; ++t) {
      if (p->t_src[t] <= last[p->t_dst[t]]) p->sorted_in = false;
      last[p->t_dst[t]] = p->t_src[t];
    }
  }
  p->empty = true;
  for (int i = 0; i < p->n; ++i)
    if (p->is_emit[i]) p->empty = false;
  cudaSetDevice(ctx->device);
  const size_t tb = (size_t)p->n * p->A * p->C * sizeof(double);
  p->arena = nullptr;
  int rc = check_cuda(ctx, cudaMallocFromPoolAsync((void**)&p->d_lp_absorb, tb, ctx->pool, ctx->stream), "cudaMallocAsync(imported profile)");
  if (!rc) rc = check_cuda(ctx, cudaMemcpyAsync(p->d_lp_absorb, dev_table, tb, cudaMemcpyDeviceToDevice, ctx->stream), "D2D imported table");
  if (!rc) rc = check_cuda(ctx, cudaStreamSynchronize(ctx->stream), "sync after import");   // the caller's buffer is only borrowed
  if (rc) {
    if (p->d_lp_absorb) cudaFreeAsync(p->d_lp_absorb, ctx->stream);
    delete p;
    delete st;
    return rc;
  }
  nr.prof = p;
  if (nr.store && nr.store->full) delete st;   // this rank built the node itself: its complete store stays
  else {
    delete nr.store;
    nr.store = st;
  }
  build_in_lists(p, nr.in_off, nr.in_idx);
  return HIST_OK;
}

// The profile of `node` has been handed to another rank: free its device table here (its path store stays).
extern "C" int hist_recon_release_node(hist_recon* R, int32_t node) {
  if (!R || node < 0 || node >= R->n_nodes) return HIST_ERR_ARG;
  NodeRec& nr = R->node[node];
  if (nr.prof) hist_profile_destroy(nr.prof);
  nr.prof = nullptr;
  std::vector<int>().swap(nr.in_off);
  std::vector<int>().swap(nr.in_idx);
  return HIST_OK;
}

extern "C" int hist_recon_export_stores(hist_recon* R, void* buf, int64_t capacity, int64_t* bytes) {
  if (!R || !bytes) return HIST_ERR_ARG;
  std::vector<char> b;
  int count = 0;
  for (auto& nr : R->node)
    if (nr.store && nr.store->full) ++count;
  put(b, &count, 1);
  for (int v = 0; v < R->n_nodes; ++v) {
    const PathStore* st = R->node[v].store;
    if (!st || !st->full) continue;
    put(b, &v, 1);
    put(b, st->child_node, 2);
    put_vec(b, st->cols);
    put_vec(b, st->seqs);
    put_vec(b, st->items);
    put_vec(b, st->state_col);
    put_vec(b, st->trans_seq);
  }
  return copy_out(R->ctx, b, buf, capacity, bytes, "hist_recon_export_stores");
}

extern "C" int hist_recon_import_stores(hist_recon* R, const void* buf, int64_t bytes) {
  if (!R || !buf) return HIST_ERR_ARG;
  const char* q = (const char*)buf;
  const char* end = q + bytes;
  int count = 0;
  if (!get(q, end, &count, 1)) return set_err(R->ctx, HIST_ERR_ARG, "hist_recon_import_stores: malformed blob");
  for (int k = 0; k < count; ++k) {
    int v = -1;
    PathStore* st = new PathStore();
    const bool ok = get(q, end, &v, 1) && v >= 0 && v < R->n_nodes && get(q, end, st->child_node, 2) && get_vec(q, end, st->cols) &&
                    get_vec(q, end, st->seqs) && get_vec(q, end, st->items) && get_vec(q, end, st->state_col) && get_vec(q, end, st->trans_seq);
    if (!ok) {
      delete st;
      return set_err(R->ctx, HIST_ERR_ARG, "hist_recon_import_stores: malformed blob");
    }
    if (R->node[v].store && R->node[v].store->full) {   // this rank built it itself
      delete st;
      continue;
    }
    delete R->node[v].store;
    R->node[v].store = st;
  }
  R->aln_cols = -1;
  return HIST_OK;
}

// ------------------------------------------------------------------ profile read-back (tests)
extern "C" int hist_profile_get_sizes(const hist_profile* p, int32_t* n_states, int32_t* n_trans, int32_t* alph_size, int32_t* components) {
  if (!p) return HIST_ERR_ARG;
  if (n_states) *n_states = p->n;
  if (n_trans) *n_trans = p->nt;
  if (alph_size) *alph_size = p->A;
  if (components) *components = p->C;
  return HIST_OK;
}

extern "C" int hist_profile_fetch(hist_profile* p, uint8_t* is_emit, int32_t* trans_src, int32_t* trans_dest, double* trans_lp, double* lp_absorb,
                                  int32_t* closest_leaf_pos) {
  if (!p) return HIST_ERR_ARG;
  hist_ctx* ctx = p->ctx;
  if (is_emit) memcpy(is_emit, p->is_emit.data(), p->n);
  if (trans_src) memcpy(trans_src, p->t_src.data(), (size_t)p->nt * 4);
  if (trans_dest) memcpy(trans_dest, p->t_dst.data(), (size_t)p->nt * 4);
  if (trans_lp) memcpy(trans_lp, p->t_lp.data(), (size_t)p->nt * 8);
  if (closest_leaf_pos) {
    if (p->leaf_pos.empty()) memset(closest_leaf_pos, 0, (size_t)p->n * 4);
    else memcpy(closest_leaf_pos, p->leaf_pos.data(), (size_t)p->n * 4);
  }
  if (lp_absorb) {
    cudaSetDevice(ctx->device);
    HB_CUDA(ctx, cudaMemcpyAsync(lp_absorb, p->d_lp_absorb, (size_t)p->n * p->A * p->C * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
    HB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
  }
  return HIST_OK;
}
