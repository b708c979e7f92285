"""Host-side driver of the progressive reconstruction: the caller of the hot path.

Mirrors, on top of the C ABI, the parts of the reference that are bookkeeping around the DP:
  * ForwardMatrix::bestProfile / sampleProfile (forward.cpp:845-895) and
    BackwardMatrix::postProbProfile / addTrace / addCells (forward.cpp:1321-1379): which cells go
    into the parent profile (the traces themselves come from the device);
  * ForwardMatrix::makeProfile's assembly (forward.cpp:686-843): states/transitions from the device
    elimination (hist_profile_eliminate), seqCoords (forward.cpp:533-538), alignPaths
    (forward.cpp:540-578), then Profile::addReadyStates (profile.cpp:268-319);
  * Dataset::prepareRecon's closest-leaf table (recon.cpp:864-915) and the node loop of
    Reconstructor::reconstruct with its band-widening retry (recon.cpp:917-1052);
  * traceAlignPath (forward.cpp:654-684) for the final alignment.
AlignPaths are kept as lazy expressions (only the cells of the final root path are ever expanded).
There is no CPU implementation of the DP here: every matrix is filled, traced and eliminated on the
device.
"""
import numpy as np

from . import descs as D
from .profile_ops import add_ready_states

IMM, IMD, IDM, IMI, IIW, EEE = D.IMM, D.IMD, D.IDM, D.IMI, D.IIW, D.EEE


# ------------------------------------------------------------------ lazy AlignPath (alignpath.cpp:56-83)
class Path:
    """AlignPath expression: ('empty',) | ('col', row) | ('union', a, b) | ('concat', [..]) | ('addrow', a, row)."""
    __slots__ = ("op", "args", "_cache")

    def __init__(self, op, *args):
        self.op, self.args, self._cache = op, args, None

    def expand(self):
        """dict row -> list of bool (the reference's AlignPath)."""
        if self._cache is None:
            op = self.op
            if op == "empty":
                r = {}
            elif op == "col":
                r = {self.args[0]: [True]}
            elif op == "union":  # alignPathUnion: rows of the second are added to the first
                r = dict(self.args[0].expand())
                for k, v in self.args[1].expand().items():
                    r.setdefault(k, v)
            elif op == "addrow":  # alignPath[parentRow].push_back(true)
                r = {k: list(v) for k, v in self.args[0].expand().items()}
                r.setdefault(self.args[1], []).append(True)
            else:
                r = concat_many([p.expand() for p in self.args[0]])
            self._cache = r
        return self._cache


EMPTY = Path("empty")


def path_columns(a):
    return max((len(v) for v in a.values()), default=0)


def concat_paths(a1, a2):
    """alignPathConcat (alignpath.cpp:62-78)."""
    c1, c2 = path_columns(a1), path_columns(a2)
    a = {k: list(v) for k, v in a1.items()}
    for k in a:
        if k not in a2:
            a[k].extend([False] * c2)
    for row, rp in a2.items():
        lp = a.setdefault(row, [])
        if not lp:
            lp.extend([False] * c1)
        lp.extend(rp)
    return a


def concat_many(paths):
    """Left fold of alignPathConcat over a list of expanded paths in linear time (the fold itself copies the
    accumulated path at every step).  Rows of one path all have that path's column count."""
    out, offset = {}, 0
    for a in paths:
        if not a:
            continue
        c = path_columns(a)
        for row, rp in a.items():
            lst = out.get(row)
            if lst is None:
                lst = out[row] = []
            if len(lst) < offset:
                lst.extend([False] * (offset - len(lst)))
            lst.extend(rp)
        offset += c
    for lst in out.values():
        if len(lst) < offset:
            lst.extend([False] * (offset - len(lst)))
    return out


# ------------------------------------------------------------------ host profile
class HostProfile:
    """A Profile (profile.h:41-76) on the host: the flattened arrays the C ABI takes plus the
    bookkeeping that never goes to the device."""

    def __init__(self, flat, state_path, trans_path, leaf_rows, root_row, name=""):
        self.flat = flat                  # descs.Profile (closest_leaf_pos set by select_leaf)
        self.state_path = state_path      # [n_states] Path
        self.trans_path = trans_path      # [n_trans] Path
        self.leaf_rows = leaf_rows        # {row: int32[n_states]} ProfileState::seqCoords of the rows still needed
        self.root_row = root_row
        self.name = name
        self._trans_index = None
        has_in = np.zeros(flat.n_states, bool)
        has_in[flat.trans_dest] = True
        self.emit_or_start = flat.is_emit.astype(bool) | ~has_in  # profile.h:34-35

    @staticmethod
    def leaf(seq, alphabet, components, row, name=""):
        """Profile(components, alphabet, FastSeq, row) (profile.cpp:23-76)."""
        flat = D.Profile.from_sequence(seq, alphabet, components)
        n = flat.n_states
        sp = [EMPTY] + [Path("col", row) for _ in range(n - 2)] + [EMPTY]
        coords = np.arange(n, dtype=np.int32).clip(0, len(seq))
        return HostProfile(flat, sp, [EMPTY] * flat.n_trans, {row: coords}, row, name)

    def with_leaf(self, row):
        """The flattened profile carrying seqCoords.at(row) as closest_leaf_pos (forward.cpp:36-42)."""
        f = self.flat
        return D.Profile(f.is_emit, f.trans_src, f.trans_dest, f.trans_lp, f.lp_absorb, self.leaf_rows[row], name=self.name)

    def trans_between(self, s, d):
        if self._trans_index is None:
            self._trans_index = {(int(a), int(b)): t for t, (a, b) in enumerate(zip(self.flat.trans_src, self.flat.trans_dest))}
        return self._trans_index[(int(s), int(d))]

    def is_null(self, i):
        return not self.flat.is_emit[i]

    def is_emit_or_start(self, i):
        return bool(self.emit_or_start[i])


class PairView:
    """What the assembly needs from one sibling merge: the two child profiles and the parent row."""

    def __init__(self, x, y, parent_row):
        self.x, self.y, self.parent_row = x, y, parent_row

    def is_absorbing(self, c):  # forward.cpp:470-474
        i, j, s = c
        return ((s == IMM and not self.x.is_null(i) and not self.y.is_null(j)) or (s == IMD and not self.x.is_null(i))
                or (s == IDM and not self.y.is_null(j)))

    def cell_path(self, c):  # cellAlignPath, forward.cpp:540-568
        i, j, s = c
        if s == IMM:
            if not self.x.is_null(i) and not self.y.is_null(j):
                p = Path("union", self.x.state_path[i], self.y.state_path[j])
            elif self.x.is_emit_or_start(i):
                p = self.y.state_path[j]
            else:
                p = self.x.state_path[i]
        elif s in (IMD, IIW):
            p = self.x.state_path[i]
        elif s in (IDM, IMI):
            p = self.y.state_path[j]
        else:
            p = EMPTY
        if self.is_absorbing(c):
            p = Path("addrow", p, self.parent_row)
        return p

    def trans_path(self, src, dest):  # transitionAlignPath, forward.cpp:570-577
        parts = []
        if src[0] != dest[0]:
            parts.append(self.x.trans_path[self.x.trans_between(src[0], dest[0])])
        if src[1] != dest[1]:
            parts.append(self.y.trans_path[self.y.trans_between(src[1], dest[1])])
        if not parts:
            return EMPTY
        return parts[0] if len(parts) == 1 else Path("concat", parts)

    def trace_path(self, trace):  # traceAlignPath, forward.cpp:654-684
        parts = []
        tr = [tuple(int(v) for v in c) for c in trace]
        for n in range(len(tr) - 1):
            parts += [self.cell_path(tr[n]), self.trans_path(tr[n], tr[n + 1])]
        parts.append(self.cell_path(tr[-1]))
        p = Path("concat", parts).expand()
        for row in (self.parent_row, self.x.root_row, self.y.root_row):  # ensureAlignPathHasRow
            p.setdefault(row, [])
        return p


# ------------------------------------------------------------------ which cells make the parent profile
def _cellset(cells):
    return np.array(sorted(cells), np.int32).reshape(-1, 3)


def best_cells(pair):
    """ForwardMatrix::bestProfile's cell set (forward.cpp:891-895)."""
    return _cellset({tuple(int(v) for v in c) for c in pair.trace_best()})


def sampled_cells(pair, mt_state, profile_samples, max_cells, strategy, min_len=0, max_len=None):
    """ForwardMatrix::sampleProfile's cell set (forward.cpp:845-889).  mt_state is advanced in place."""
    count = {}
    n_traces = 0
    if strategy & D.INCLUDE_BEST_TRACE:
        for c in pair.trace_best():
            count[tuple(int(v) for v in c)] = 2  # avoid dropping these cells
        n_traces += 1
    elif profile_samples <= 0:
        raise ValueError("Must allow at least one sample path in the profile")
    accepted = 0
    while accepted < profile_samples and (max_cells == 0 or len(count) < max_cells):
        tr = pair.trace_sample(mt_state, 1)[0]
        anc_len = int(np.sum((tr[:, 2] == IMM) | (tr[:, 2] == IDM) | (tr[:, 2] == IMD)))
        if anc_len < min_len or (max_len is not None and anc_len > max_len):
            break
        for c in tr:
            k = tuple(int(v) for v in c)
            count[k] = count.get(k, 0) + 1
        n_traces += 1
        accepted += 1
    threshold = 2 if (n_traces > 1 and max_cells > 0 and len(count) >= max_cells) else 1
    return _cellset({c for c, n in count.items() if n >= threshold})


def posterior_cells(pair, min_post_prob, max_cells, strategy):
    """BackwardMatrix::postProbProfile's cell set (forward.cpp:1327-1379), KeepGapsOpen excluded."""
    if strategy & D.KEEP_GAPS_OPEN:
        raise NotImplementedError("KeepGapsOpen posterior profiles (forward.cpp:1368-1375)")
    queue, _ = pair.posterior_cells(min_post_prob)
    cells = set()

    def add_cells(fwd_trace, back_trace):  # addCells, forward.cpp:1344-1377
        new = []
        for c in fwd_trace[::-1]:
            k = tuple(int(v) for v in c)
            if k in cells:
                break
            new.append(k)
        for c in back_trace:
            k = tuple(int(v) for v in c)
            if k in cells:
                break
            new.append(k)
        if max_cells > 0 and len(cells) > 0 and len(cells) + len(new) > max_cells:
            return False
        cells.update(new)
        return True

    if len(queue) == 0 or (strategy & D.INCLUDE_BEST_TRACE):
        add_cells(pair.trace_best(), [])
    k = 0
    while (max_cells == 0 or len(cells) < max_cells) and k < len(queue):
        best = tuple(int(v) for v in queue[k])
        if best in cells:
            k += 1
        elif not add_cells(pair.trace_best(best), pair.trace_best_forward(best)):  # addTrace, forward.cpp:1379-1384
            break
    return _cellset(cells)


# ------------------------------------------------------------------ makeProfile assembly
def make_profile(pair, view, cells, strategy, name=""):
    """ForwardMatrix::makeProfile (forward.cpp:686-843) from the device elimination, then
    Profile::addReadyStates (profile.cpp:268-319).  Returns (HostProfile, elimination dict)."""
    e = pair.eliminate(cells, strategy)
    st = [tuple(int(v) for v in c) for c in e["state_cell"]]
    ns = len(st)
    is_emit = np.array([view.is_absorbing(c) for c in st], np.uint8)
    state_path = [view.cell_path(c) for c in st]
    trans_path = []
    via, off = e["via"], e["via_off"]
    for t in range(len(e["trans_src"])):
        chain = [st[e["trans_src"][t]]] + [tuple(int(v) for v in c) for c in via[off[t]:off[t + 1]]] + [st[e["trans_dest"][t]]]
        parts = []
        for k in range(len(chain) - 1):  # tap(src,c1), cap(c1), tap(c1,c2), ... , tap(c_last,dest)   (forward.cpp:764, 797)
            parts.append(view.trans_path(chain[k], chain[k + 1]))
            if k + 1 < len(chain) - 1:
                parts.append(view.cell_path(chain[k + 1]))
        trans_path.append(parts[0] if len(parts) == 1 else Path("concat", parts))
    xi = np.array([c[0] for c in st]), np.array([c[1] for c in st])
    leaf_rows = {}
    for row, arr in view.x.leaf_rows.items():  # cellSeqCoords, forward.cpp:533-538 (y entries override x)
        leaf_rows[row] = arr[xi[0]]
    for row, arr in view.y.leaf_rows.items():
        leaf_rows[row] = arr[xi[1]]
    flat = D.Profile(is_emit, e["trans_src"], e["trans_dest"], e["trans_lp"], e["state_lp_absorb"], None, name=name)
    q, old2new, extra = add_ready_states(flat, dict(leaf_rows))
    n_new = q.n_states
    sp = [EMPTY] * n_new
    for s in range(ns):
        sp[old2new[s]] = state_path[s]
    tp = trans_path + [EMPTY] * (q.n_trans - len(trans_path))
    prof = HostProfile(q, sp, tp, {r: np.ascontiguousarray(a, np.int32) for r, a in extra.items()}, view.parent_row, name)
    prof.state_cell = [None] * n_new
    prof.fwd_lp = np.full(n_new, np.nan)
    for s in range(ns):
        prof.state_cell[old2new[s]] = st[s]
        prof.fwd_lp[old2new[s]] = e["state_fwd_lp"][s]
    return prof, e


# ------------------------------------------------------------------ Reconstructor::reconstruct (recon.cpp:917-1052)
def closest_leaves(tree):
    """Dataset::prepareRecon's closestLeaf / closestLeafDistance (recon.cpp:890-911)."""
    n = tree.n_nodes
    cl, dcl = np.zeros(n, np.int64), np.zeros(n)
    for node in range(n):
        if tree.is_leaf(node):
            cl[node], dcl[node] = node, 0.0
        else:
            best, dbest = -1, 0.0
            for nc, c in enumerate(tree.children[node]):
                dc = dcl[c] + tree.branch_length[c]
                if nc == 0 or dc < dbest:
                    best, dbest = cl[c], dc
            cl[node], dcl[node] = best, dbest
    return cl, dcl


def mt19937_seed(seed):
    """std::mt19937(seed) as the 625-word image the C ABI passes in and out."""
    mt = np.zeros(625, np.uint32)
    x = seed & 0xffffffff
    mt[0] = x
    for k in range(1, 624):
        x = (1812433253 * (x ^ (x >> 30)) + k) & 0xffffffff
        mt[k] = x
    mt[624] = 624
    return mt


class Reconstructor:
    """The node loop of Reconstructor::reconstruct on the device context `ctx`.

    Defaults follow recon.cpp:32-75 / recon.h:13-27: band 20, 10 profile samples, seed 5489 (mt19937's
    default), CollapseChains | IncludeBestTrace.  `max_profile_states` must be given explicitly (the reference
    derives it from host RAM, recon.cpp:77-79)."""

    def __init__(self, ctx, model, band=20, profile_samples=10, max_profile_states=0, min_post_prob=None, seed=5489,
                 include_best_trace=True, keep_gaps_open=False):
        self.ctx, self.model = ctx, model
        self.band, self.profile_samples, self.max_profile_states = band, profile_samples, max_profile_states
        self.min_post_prob = min_post_prob       # not None: usePosteriorsForProfile (-profminpost)
        self.seed = seed
        self.strategy = (D.COLLAPSE_CHAINS | (D.INCLUDE_BEST_TRACE if include_best_trace else 0)
                         | (D.KEEP_GAPS_OPEN if keep_gaps_open else 0))
        self.log = []

    def merge(self, node, lprof, rprof, tl, tr, guide_rows, cl_l, cl_r, is_root, mt_state):
        """One internal node: DPMatrix + Forward (+ Backward), cell selection, makeProfile."""
        from . import model as M
        hmm = M.pair_hmm(self.model, tl, tr)
        have_guide = guide_rows is not None
        xf = lprof.with_leaf(cl_l) if have_guide else lprof.flat
        yf = rprof.with_leaf(cl_r) if have_guide else rprof.flat
        dx, dy = self.ctx.upload(xf), self.ctx.upload(yf)
        max_dist = self.band if have_guide else -1
        while True:  # band-widening retry (recon.cpp:954-975)
            env = D.Envelope.from_guide_rows(guide_rows[cl_l], guide_rows[cl_r], max_dist) if max_dist >= 0 else D.Envelope(-1)
            pair = self.ctx.prepare(dx, dy, hmm, env)
            lp_end = pair.forward()
            if lp_end > -np.inf:
                break
            if max_dist < 0:
                raise RuntimeError("Zero forward likelihood even in the absence of guide alignment constraints")
            n_cols = len(next(iter(guide_rows.values())))
            max_dist = -1 if max_dist * 2 > n_cols else (1 if max_dist == 0 else max_dist * 2)
            pair.close()
        view = PairView(lprof, rprof, node)
        use_post = self.min_post_prob is not None
        path = None
        if is_root:
            best = pair.trace_best()
            path = view.trace_path(best)                      # forward->bestAlignPath()
            cells = _cellset({tuple(int(v) for v in c) for c in best})
            prof, _ = make_profile(pair, view, cells, D.COLLAPSE_CHAINS)  # forward->bestProfile() (default strategy, forward.h:163)
        elif use_post:
            pair.backward()
            cells = posterior_cells(pair, self.min_post_prob, self.max_profile_states, self.strategy)
            prof, _ = make_profile(pair, view, cells, self.strategy)
        else:
            cells = sampled_cells(pair, mt_state, self.profile_samples, self.max_profile_states, self.strategy)
            prof, _ = make_profile(pair, view, cells, self.strategy)
        self.log.append(dict(node=node, lp_end=lp_end, cells=int(pair.n_cells), states=prof.flat.n_states, band=max_dist))
        pair.close(); dx.close(); dy.close()
        return prof, lp_end, path

    def reconstruct(self, tree, seqs, guide=None):
        """tree: simulator.Tree (postorder, root last); seqs: {leaf node: str}; guide: {leaf node: bool[cols]} or None.
        Returns (alignment path {row: [bool]}, final Forward log-likelihood, {node: HostProfile})."""
        cl, _ = closest_leaves(tree)
        mt_state = mt19937_seed(self.seed)
        prof, path, lp_final = {}, None, -np.inf
        for node in range(tree.n_nodes):
            if tree.is_leaf(node):
                prof[node] = HostProfile.leaf(seqs[node], self.model.alphabet, self.model.components, node, tree.names[node])
                continue
            l, r = tree.children[node]
            keep = int(cl[node])
            p, lp_end, pth = self.merge(node, prof[l], prof[r], float(tree.branch_length[l]), float(tree.branch_length[r]), guide,
                                        int(cl[l]), int(cl[r]), node == tree.root, mt_state)
            # only the closest-leaf coordinates of this node are needed further up (forward.cpp:36-42)
            p.leaf_rows = {keep: p.leaf_rows[keep]} if keep in p.leaf_rows else p.leaf_rows
            prof[node] = p
            if node == tree.root:
                path, lp_final = pth, lp_end
            else:
                # children are not needed again except for their alignPath expressions, which the parent references
                prof[l].leaf_rows, prof[r].leaf_rows = {}, {}
        return path, lp_final, prof


class NativeReconstructor:
    """Reconstructor::reconstruct on the native, level-batched driver (hist_recon_*, csrc/hist_recon.cu): the same
    parameters and results as Reconstructor above, but every guide-tree level is ONE library call, the parent
    profiles stay on the device and the makeProfile bookkeeping runs on the library's host threads.

    Deterministic profiles (best trace only: profile_samples == 0 or max_profile_states == 1) are batched by level;
    sampled profiles share one mt19937 stream in node order (recon.cpp:38, forward.cpp:257-276), so their merges
    go through the library one node at a time, in node order."""

    def __init__(self, ctx, model, band=20, profile_samples=10, max_profile_states=0, seed=5489, include_best_trace=True,
                 keep_gaps_open=False):
        self.ctx, self.model = ctx, model
        self.band, self.profile_samples, self.max_profile_states, self.seed = band, profile_samples, max_profile_states, seed
        self.strategy = (D.COLLAPSE_CHAINS | (D.INCLUDE_BEST_TRACE if include_best_trace else 0)
                         | (D.KEEP_GAPS_OPEN if keep_gaps_open else 0))
        self.lp_end = {}
        self.stage_ms = {}
        self._hmm_cache = {}

    def draws(self):
        return self.profile_samples > 0 and not (self.strategy & D.INCLUDE_BEST_TRACE and self.max_profile_states == 1)

    def _hmm(self, tl, tr):
        from . import model as M
        key = (float(tl), float(tr))
        h = self._hmm_cache.get(key)
        if h is None:
            hm = M.pair_hmm(self.model, key[0], key[1])
            h = self._hmm_cache[key] = (hm, hm.desc())
        return h[1]

    def schedule(self, tree):
        """Lists of internal nodes that go through the library together."""
        if self.draws():
            return [[v] for v in range(tree.n_nodes) if not tree.is_leaf(v)]   # one generator stream, node order
        return merge_levels(tree)

    def reconstruct(self, tree, seqs, guide=None, rec=None, groups=None, on_group=None, mt_state=None):
        """Same arguments as Reconstructor.reconstruct.  Returns (alignment bool [n_nodes][cols], final Forward
        log-likelihood, capi.Recon).  `rec`/`groups`: continue on an existing hist_recon with a given schedule;
        on_group(nodes, rec) is called after every library call (the new parents are still live then); mt_state: the
        generator image to continue from (a resumed sampled run), updated in place; kept in self.mt_state."""
        from . import capi
        cl, _ = closest_leaves(tree)
        if rec is None:
            rec = capi.Recon(self.ctx, tree.n_nodes, self.model.alphabet, self.model.components)
            leaves = tree.leaves()
            if guide is not None:
                rec.set_guide(leaves, np.array([guide[v] for v in leaves], np.uint8))
            rec.add_leaves(leaves, [seqs[v] for v in leaves])
        mt_state = mt19937_seed(self.seed) if mt_state is None else mt_state
        self.mt_state = mt_state
        band = self.band if guide is not None else -1
        lp_final = -np.inf
        for nodes in (groups if groups is not None else self.schedule(tree)):
            merges, hd = [], []
            for v in nodes:
                l, r = tree.children[v]
                merges.append((v, l, r, cl[l], cl[r], cl[v]))
                hd.append(self._hmm(tree.branch_length[l], tree.branch_length[r]))
            lp = rec.level(merges, hd, band, self.strategy, self.profile_samples, self.max_profile_states,
                           tree.root if tree.root in nodes else -1, mt_state)
            for v, x in zip(nodes, lp):
                self.lp_end[v] = float(x)
            if tree.root in nodes:
                lp_final = self.lp_end[tree.root]
            if on_group is not None:
                on_group(nodes, rec)
        self.stage_ms = rec.stage_ms()
        if tree.root not in self.lp_end and groups is not None:
            return None, lp_final, rec   # a partial schedule: no root yet
        return rec.alignment(), lp_final, rec


# ------------------------------------------------------------------ sibling merges over several GPUs
def plan_owners(tree, world):
    """Placement of the progressive reconstruction on `world` ranks (one process per GPU): the leaves are dealt out in
    contiguous blocks of the tree's leaf order, and an internal node is built where its larger child (by leaf count;
    ties: the left one) lives, so whole subtrees stay on one rank and only the smaller child's flattened profile moves
    (SURVEY.md 8(e)).  Returns owner[node]."""
    n = tree.n_nodes
    leaves = tree.leaves()
    owner = np.zeros(n, np.int64)
    size = np.zeros(n, np.int64)
    for k, v in enumerate(leaves):
        owner[v] = k * world // len(leaves)
        size[v] = 1
    for v in range(n):
        kids = tree.children[v]
        if kids:
            size[v] = sum(size[c] for c in kids)
            owner[v] = owner[max(kids, key=lambda c: (size[c], -kids.index(c)))]
    return owner


def level_moves(tree, nodes, owner):
    """Profiles that must change rank before the merges `nodes` can run: [(child node, from rank, to rank)]."""
    return [(c, int(owner[c]), int(owner[v])) for v in nodes for c in tree.children[v] if owner[c] != owner[v]]


def exchange_profiles(rec, moves, rank, dist, device):
    """Moves finished profiles between ranks with point-to-point transfers (torch.distributed send/recv: NCCL over NVLink
    for CUDA tensors).  A profile travels as a header (two sizes), the host blob and the lpAbsorb table; the table goes
    device to device.  Every rank walks the same list, so sends and receives pair up in order."""
    import torch
    for node, src, dst in moves:
        if rank == src:
            nb, db = rec.export_sizes(node)
            table = torch.empty(db, dtype=torch.uint8, device=device)
            blob = rec.export_node(node, table.data_ptr())
            dist.send(torch.tensor([nb, db], dtype=torch.int64, device=device), dst)
            dist.send(torch.from_numpy(blob).to(device), dst)
            dist.send(table, dst)
            rec.release_node(node)
        elif rank == dst:
            hdr = torch.zeros(2, dtype=torch.int64, device=device)
            dist.recv(hdr, src)
            nb, db = (int(v) for v in hdr.tolist())
            blob = torch.empty(nb, dtype=torch.uint8, device=device)
            table = torch.empty(db, dtype=torch.uint8, device=device)
            dist.recv(blob, src)
            dist.recv(table, src)
            if table.is_cuda:
                torch.cuda.synchronize()
            rec.import_node(node, blob.cpu().numpy(), table.data_ptr())


def gather_stores(rec, to_rank, rank, world, dist, device):
    """The alignment-path stores of every rank, sent once to the rank that expands the final alignment."""
    import torch
    if rank != to_rank:
        blob = rec.export_stores()
        dist.send(torch.tensor([blob.size], dtype=torch.int64, device=device), to_rank)
        dist.send(torch.from_numpy(blob).to(device), to_rank)
        return
    for src in range(world):
        if src == to_rank:
            continue
        hdr = torch.zeros(1, dtype=torch.int64, device=device)
        dist.recv(hdr, src)
        blob = torch.empty(int(hdr.item()), dtype=torch.uint8, device=device)
        dist.recv(blob, src)
        rec.import_stores(blob.cpu().numpy())


def reconstruct_partitioned(nat, tree, seqs, guide, rank, world, dist, device):
    """NativeReconstructor.reconstruct with the sibling merges of every guide-tree level spread over `world` ranks
    (plan_owners).  Only RNG-free profiles (deterministic mode) can be spread: sampled profiles share one mt19937 stream
    in node order.  Returns (alignment or None, final log-likelihood or None, capi.Recon); the alignment and the
    likelihood are returned on the rank that owns the root."""
    from . import capi
    if nat.draws() and world > 1:
        raise ValueError("sampled profiles share one RNG stream in node order; use max_profile_states=1 or profile_samples=0 for multi-GPU runs")
    owner = plan_owners(tree, world)
    cl, _ = closest_leaves(tree)
    rec = capi.Recon(nat.ctx, tree.n_nodes, nat.model.alphabet, nat.model.components)
    leaves = tree.leaves()
    if guide is not None:
        rec.set_guide(leaves, np.array([guide[v] for v in leaves], np.uint8))
    mine = [v for v in leaves if owner[v] == rank]
    if mine:
        rec.add_leaves(mine, [seqs[v] for v in mine])
    band = nat.band if guide is not None else -1
    mt_state = mt19937_seed(nat.seed)
    lp_final = None
    for nodes in merge_levels(tree):
        exchange_profiles(rec, level_moves(tree, nodes, owner), rank, dist, device)
        todo = [v for v in nodes if owner[v] == rank]
        if not todo:
            continue
        merges, hd = [], []
        for v in todo:
            l, r = tree.children[v]
            merges.append((v, l, r, cl[l], cl[r], cl[v]))
            hd.append(nat._hmm(tree.branch_length[l], tree.branch_length[r]))
        lp = rec.level(merges, hd, band, nat.strategy, nat.profile_samples, nat.max_profile_states,
                       tree.root if tree.root in todo else -1, mt_state)
        for v, x in zip(todo, lp):
            nat.lp_end[v] = float(x)
        if tree.root in todo:
            lp_final = nat.lp_end[tree.root]
    root_rank = int(owner[tree.root])
    if world > 1:
        gather_stores(rec, root_rank, rank, world, dist, device)
    nat.stage_ms = rec.stage_ms()
    return (rec.alignment() if rank == root_rank else None), lp_final, rec


def gapped_rows_dense(bits, tree, seqs, wildcard="*"):
    """gapped_rows for the dense alignment matrix NativeReconstructor returns."""
    out = {}
    for row in range(bits.shape[0]):
        rp = bits[row]
        s = seqs[row] if tree.is_leaf(row) else wildcard * int(rp.sum())
        it = iter(s)
        out[row] = "".join(next(it) if b else "-" for b in rp)
    return out


def merge_levels(tree):
    """Internal nodes grouped by height above the leaves (level 0 = cherries).  The merges of one level
    depend only on lower levels (recon.cpp:938-1033 is a postorder loop whose only dependencies are
    child -> parent), so they are independent and can run side by side / on different GPUs."""
    h = np.zeros(tree.n_nodes, int)
    for v in range(tree.n_nodes):
        if tree.children[v]:
            h[v] = 1 + max(h[c] for c in tree.children[v])
    levels = [[] for _ in range(int(h.max()))]
    for v in range(tree.n_nodes):
        if tree.children[v]:
            levels[h[v] - 1].append(v)
    return levels


def level_owner(level_nodes, world):
    """Round-robin placement of the merges of one level: {node: rank}."""
    return {node: k % world for k, node in enumerate(level_nodes)}


def run_levels(tree, leaf_value, merge_fn, rank=0, world=1, all_gather=None):
    """Level-synchronous driver of the progressive reconstruction over `world` ranks (one process per GPU).

    leaf_value(node) -> value of a leaf; merge_fn(node, left_value, right_value) -> value of an internal node.
    Every rank computes the merges level_owner() assigns to it, then the level's results are exchanged with
    all_gather(dict) -> list of dicts (torch.distributed.all_gather_object: the flattened profiles are host
    objects of <= ~1 MB), so that every rank holds every profile of the finished level.
    There is no collective on the data path of a merge.  Returns {node: value} (complete on every rank)."""
    val = {v: leaf_value(v) for v in range(tree.n_nodes) if tree.is_leaf(v)}
    for nodes in merge_levels(tree):
        owner = level_owner(nodes, world)
        mine = {}
        for node in nodes:
            if owner[node] == rank:
                l, r = tree.children[node]
                mine[node] = merge_fn(node, val[l], val[r])
        if world > 1:
            if all_gather is None:
                raise ValueError("world > 1 needs an all_gather")
            for part in all_gather(mine):
                val.update(part)
        else:
            val.update(mine)
    return val


def reconstruct_distributed(rec, tree, seqs, guide=None, rank=0, world=1, all_gather=None):
    """Reconstructor.reconstruct with the sibling merges of each guide-tree level spread over the ranks.

    Sampled profiles draw from ONE mt19937 stream in node order in the reference (recon.cpp:38, forward.cpp:
    257-276), which serialises the tree; here a multi-rank run is only allowed for the stream-free modes
    (best-trace profiles: profile_samples == 0 or -profmaxstates 1; or posterior profiles, -profminpost), where the result is
    identical to the single-stream run.  Returns (path, final log-likelihood, {node: HostProfile})."""
    draws = rec.min_post_prob is None and rec.profile_samples > 0 and rec.max_profile_states != 1
    if world > 1 and draws:  # max_profile_states == 1: the best trace alone fills the cell budget, no draw is made
        raise ValueError("sampled profiles share one RNG stream in node order; use profile_samples=0 or min_post_prob for multi-GPU runs")
    cl, _ = closest_leaves(tree)
    mt_state = mt19937_seed(rec.seed)

    def leaf_value(node):
        return (HostProfile.leaf(seqs[node], rec.model.alphabet, rec.model.components, node, tree.names[node]), -np.inf, None)

    def merge_fn(node, lv, rv):
        l, r = tree.children[node]
        keep = int(cl[node])
        p, lp_end, pth = rec.merge(node, lv[0], rv[0], float(tree.branch_length[l]), float(tree.branch_length[r]), guide,
                                   int(cl[l]), int(cl[r]), node == tree.root, mt_state)
        p.leaf_rows = {keep: p.leaf_rows[keep]} if keep in p.leaf_rows else p.leaf_rows
        return (p, lp_end, pth)

    val = run_levels(tree, leaf_value, merge_fn, rank, world, all_gather)
    p, lp_final, path = val[tree.root]
    return path, lp_final, {node: v[0] for node, v in val.items()}


def gapped_rows(path, tree, seqs, wildcard="*"):
    """Alignment(ungapped, path).gapped() (alignpath.cpp:232-280) with wildcard ancestors (recon.cpp:1410-1421)."""
    out = {}
    for row, rp in path.items():
        s = seqs[row] if tree.is_leaf(row) else wildcard * int(np.sum(rp))
        it = iter(s)
        out[row] = "".join(next(it) if b else "-" for b in rp)
    return out


# ------------------------------------------------------------------ checkpoint / resume of a partial progressive run
def save_checkpoint(path, rec, live_nodes, done_groups, mt_state=None):
    """Writes a partial progressive reconstruction to `path` (the flattened-profile format of
    hist_recon_export_node / hist_recon_export_stores, one .npz): the profiles that still wait for their parent -- host blob
    (structure, transition weights, coordinates, path ids) and lpAbsorb table, read back from the device -- and the
    alignment-path stores of every node built so far.  `done_groups`: how many groups of the schedule are finished."""
    import torch
    out = {"stores": rec.export_stores(), "live": np.array(sorted(live_nodes), np.int32), "done_groups": np.array([done_groups], np.int64)}
    if mt_state is not None:
        out["mt_state"] = np.asarray(mt_state, np.uint32)
    dev = torch.device("cuda", rec.ctx.device) if hasattr(rec.ctx, "device") else torch.device("cuda")
    for v in out["live"]:
        nb, db = rec.export_sizes(int(v))
        table = torch.empty(db, dtype=torch.uint8, device=dev)
        out["blob%d" % v] = rec.export_node(int(v), table.data_ptr())
        torch.cuda.synchronize()
        out["table%d" % v] = table.cpu().numpy()
    np.savez(path, **out)


def load_checkpoint(path, rec):
    """Restores what save_checkpoint wrote into a fresh hist_recon (guide set, no leaves added).  Returns (live nodes,
    done_groups, mt_state or None)."""
    import torch
    d = np.load(path)
    dev = torch.device("cuda", rec.ctx.device) if hasattr(rec.ctx, "device") else torch.device("cuda")
    for v in d["live"]:
        table = torch.from_numpy(d["table%d" % v]).to(dev)
        torch.cuda.synchronize()
        rec.import_node(int(v), np.ascontiguousarray(d["blob%d" % v]), table.data_ptr())
    rec.import_stores(np.ascontiguousarray(d["stores"]))
    return [int(v) for v in d["live"]], int(d["done_groups"][0]), (d["mt_state"] if "mt_state" in d.files else None)
