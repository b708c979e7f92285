"""Synthetic sequence families for benchmarks and tests: the algorithm of the reference's
Simulator (simulator.cpp:5-146) restated on numpy — Gillespie indel events per branch
(simulateGapsByGillespie), then substitution by sampling the rows of exp(R t) down the tree
(simulateSubsByMatExp).

Why not the reference binary: its simulateTree merges the per-branch alignment paths with
alignPathMerge, whose cost grows ~14x per doubling of the leaf count (measured in this
container: 128 leaves 14 s, 256 leaves 191 s, 1024 leaves did not finish in 20 min, extrapolated
>10 h), and it uses libstdc++ distributions, so its stream is not reproducible elsewhere.
This module draws from numpy's PCG64 with a fixed seed, so the family is identical on every box.
"""
import numpy as np

from .model import RateModel


class Tree:
    """Postorder-indexed binary tree, children < parent, root last (tree.cpp:141)."""

    def __init__(self, parent, branch_length, names):
        self.parent = np.asarray(parent, np.int32)
        self.branch_length = np.asarray(branch_length, float)
        self.names = list(names)
        n = len(self.parent)
        self.children = [[] for _ in range(n)]
        for v, p in enumerate(self.parent):
            if p >= 0:
                self.children[p].append(v)

    @property
    def n_nodes(self):
        return len(self.parent)

    @property
    def root(self):
        return self.n_nodes - 1

    def is_leaf(self, v):
        return not self.children[v]

    def leaves(self):
        return [v for v in range(self.n_nodes) if self.is_leaf(v)]

    def depth_levels(self):
        """Internal nodes grouped by height above the leaves (level 0 = cherries): the
        independent sibling merges of recon.cpp:938-1033 that can run side by side."""
        h = np.zeros(self.n_nodes, int)
        for v in range(self.n_nodes):
            if self.children[v]:
                h[v] = 1 + max(h[c] for c in self.children[v])
        levels = [[] for _ in range(int(h.max()))]
        for v in range(self.n_nodes):
            if self.children[v]:
                levels[h[v] - 1].append(v)
        return levels

    @staticmethod
    def balanced(n_leaves, bl, prefix="s"):
        parent, blen, names = [], [], []

        def rec(lo, hi):
            if hi - lo == 1:
                parent.append(-1); blen.append(bl); names.append("%s%04d" % (prefix, lo))
                return len(parent) - 1
            mid = (lo + hi) // 2
            a = rec(lo, mid)
            b = rec(mid, hi)
            parent.append(-1); blen.append(bl); names.append("node%d" % len(parent))
            v = len(parent) - 1
            parent[a] = parent[b] = v
            return v

        rec(0, n_leaves)
        blen[-1] = 0.0
        return Tree(parent, blen, names)

    @staticmethod
    def from_newick(text):
        """Newick / NHX string -> Tree with the reference's node numbering (knhx.cpp:62-84: leaves are numbered as they
        are read, internal nodes as their ')' closes, i.e. postorder; NHX comments [...] are skipped)."""
        parent, blen, names = [], [], []
        stack, pos, n = [], 0, len(text)
        last = -1

        def label(p):
            q = p
            while q < n and text[q] not in ",():;[":
                q += 1
            return text[p:q].strip(), q

        while pos < n:
            ch = text[pos]
            if ch == "(":
                stack.append([])
                pos += 1
            elif ch == ",":
                pos += 1
            elif ch == ")" or ch not in " \t\r\n;[":
                if ch == ")":
                    kids = stack.pop()
                    name, pos = label(pos + 1)
                    parent.append(-1); blen.append(0.0); names.append(name)
                    v = len(parent) - 1
                    for c in kids:
                        parent[c] = v
                else:
                    name, pos = label(pos)
                    parent.append(-1); blen.append(0.0); names.append(name)
                    v = len(parent) - 1
                if pos < n and text[pos] == ":":
                    num, pos = label(pos + 1)
                    blen[v] = float(num)
                if pos < n and text[pos] == "[":
                    pos = text.index("]", pos) + 1
                if stack:
                    stack[-1].append(v)
                last = v
            elif ch == "[":
                pos = text.index("]", pos) + 1
            elif ch == ";":
                break
            else:
                pos += 1
        assert last == len(parent) - 1 and not stack, "malformed Newick string"
        return Tree(parent, blen, names)

    def newick(self):
        def rec(v):
            if self.is_leaf(v):
                s = self.names[v]
            else:
                s = "(" + ",".join(rec(c) for c in self.children[v]) + ")"
            return s + (":%g" % self.branch_length[v] if self.parent[v] >= 0 else "")
        return rec(self.root) + ";\n"


def simulate_family(model: RateModel, tree: Tree, root_len: int, seed: int):
    """Returns (gapped uint8 [n_nodes, n_cols] of residue characters / '-', component [n_nodes, n_cols])."""
    rng = np.random.default_rng(seed)
    N = tree.n_nodes
    next_col = root_len
    order = list(range(root_len))          # global left-to-right order of alignment columns
    cols = [None] * N                      # per node: column id of each residue
    cols[tree.root] = list(range(root_len))
    for node in range(tree.root - 1, -1, -1):
        c = list(cols[tree.parent[node]])
        t, T = 0.0, tree.branch_length[node]
        while True:
            ins_tot = model.insRate * (len(c) + 1)
            del_tot = model.delRate * len(c)
            tot = ins_tot + del_tot
            t += -np.log(rng.random()) / tot
            if t > T:
                break
            r = rng.random() * tot
            if r < ins_tot:
                pos = int(r / model.insRate)
                n_ins = 1 + int(rng.geometric(1.0 - model.insExtProb) - 1)
                new = list(range(next_col, next_col + n_ins))
                next_col += n_ins
                at = order.index(c[pos]) if pos < len(c) else len(order)
                order[at:at] = new
                c[pos:pos] = new
            else:
                pos = int((r - ins_tot) / model.delRate)
                n_del = 1 + int(rng.geometric(1.0 - model.delExtProb) - 1)
                del c[pos:min(pos + n_del, len(c))]
        cols[node] = c
    rank = np.empty(next_col, np.int64)
    rank[np.array(order)] = np.arange(next_col)
    A, C = model.alphabetSize, model.components
    tok = np.full((N, next_col), -1, np.int16)
    cpt = np.zeros((N, next_col), np.int8)
    cw = np.cumsum(model.cptWeight)
    ins_cdf = [np.cumsum(p / p.sum()) for p in model.insProb]
    for node in range(tree.root, -1, -1):  # parents before children
        mine = rank[np.array(cols[node], np.int64)] if cols[node] else np.zeros(0, np.int64)
        p = tree.parent[node]
        u, u2 = rng.random(len(mine)), rng.random(len(mine))
        if p < 0:
            inherited = np.zeros(len(mine), bool)
        else:
            inherited = tok[p, mine] >= 0
            sub_cdf = np.cumsum(model.getSubProbMatrix(tree.branch_length[node]), axis=2)
        c_new = np.minimum(np.searchsorted(cw, u2 * cw[-1]), C - 1)
        a_new = np.zeros(len(mine), np.int64)
        for c in range(C):
            sel = c_new == c
            a_new[sel] = np.searchsorted(ins_cdf[c], u[sel])
        if inherited.any():
            pc = cpt[p, mine[inherited]].astype(np.int64)
            rows = sub_cdf[pc, tok[p, mine[inherited]].astype(np.int64)]          # [n, A] cumulative rows
            a_new[inherited] = (rows < (u[inherited] * rows[:, -1])[:, None]).sum(axis=1)
            c_new[inherited] = pc
        tok[node, mine] = np.minimum(a_new, A - 1)
        cpt[node, mine] = c_new
    alph = np.frombuffer(model.alphabet.encode(), np.uint8)
    gapped = np.where(tok >= 0, alph[np.maximum(tok, 0)], ord("-")).astype(np.uint8)
    return gapped, cpt


def leaf_guide(tree: Tree, gapped):
    """Leaf rows only, with all-gap columns removed: (leaf node ids, uint8 [n_leaves, cols])."""
    leaves = tree.leaves()
    g = gapped[leaves]
    keep = (g != ord("-")).any(axis=0)
    return leaves, np.ascontiguousarray(g[:, keep])
