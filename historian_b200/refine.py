"""Iterative branch refinement: Refiner::refine (reference refiner.cpp:108-194) as `historian recon
-refine` runs it on a finished reconstruction (recon.cpp:1041-1042, 1054-1065).

One move re-aligns a node to its parent: the conditional position weight matrices of the two nodes given the rest of the
tree (TreeAlignFuncs::getConditionalPWMs, sampler.cpp:356-372 -> hist_sumprod_excluded_posteriors on the device), a
three-state Viterbi alignment of the two matrices inside the envelope of the current alignment (Refiner::BranchMatrix,
refiner.cpp:10-106 -> hist_branch_align on the device), and the merge of the new parent-child alignment with the two
clades' alignments (alignPathMerge -> hist_align_path_merge).  A move is kept when the log-likelihood of the whole history
(TreeAlignFuncs::logLikelihood, sampler.cpp:419-427 -> column likelihoods from hist_sumprod_columns) improves; the sweep
over the nodes stops after n - 1 moves in a row without improvement.  This module is the host loop; it has no numerics of its own
beyond the indel path likelihood (em.indel_counts)."""
import math

import numpy as np

from . import capi, descs as D, em
from .model import ProbModel

DEFAULT_MAX_DISTANCE_FROM_GUIDE = 20   # recon.h:16
TREE_MIN_BRANCH_LEN = 1e-9             # tree.h:24
GAPS = (ord("-"), ord("."))


def trans_logprob(pm):
    """log ProbModel::transProb (model.cpp:400-447): rows Match, Insert, Delete; columns Match, Insert, Delete, End."""
    p = np.array([[(1 - pm.ins) * (1 - pm.del_), pm.ins, (1 - pm.ins) * pm.del_, 1 - pm.ins],
                  [(1 - pm.insExt) * (1 - pm.del_), pm.insExt, (1 - pm.insExt) * pm.del_, 1 - pm.insExt],
                  [1 - pm.delExt, 0.0, pm.delExt, 1 - pm.delExt]])
    with np.errstate(divide="ignore"):
        return np.log(p)


class Refiner:
    """Refiner (refiner.h:8-35).  rows: uint8 [N][cols] gapped characters of every tree node (node order, root last)."""

    def __init__(self, ctx, model, tree, max_distance=DEFAULT_MAX_DISTANCE_FROM_GUIDE):
        self.ctx, self.model, self.tree, self.max_distance = ctx, model, tree, max_distance
        self.sp = ctx.sumprod(D.SumProdModel.from_rate_model(model), D.TreeTopology(tree.parent, tree.branch_length))
        self.moves = self.accepted = 0
        n = tree.n_nodes
        # nodes of every subtree (cladePath's row sets, sampler.cpp:136-148)
        self.subtree = [[v] for v in range(n)]
        for v in range(n):
            p = int(tree.parent[v])
            if p >= 0:
                self.subtree[p] = self.subtree[p] + self.subtree[v]   # children are numbered before their parents
        self._pm = {}

    def close(self):
        self.sp.close()

    def log_likelihood(self, rows):
        """TreeAlignFuncs::logLikelihood (model, history), sampler.cpp:419-427."""
        cols = np.ascontiguousarray(rows.T)
        cll = self.sp.columns(cols, flags=0)["col_loglike"]
        return em.tree_align_log_likelihood(self.model, self.tree, cols, cll)[0]

    def _branch_model(self, dist):
        key = float(dist)
        m = self._pm.get(key)
        if m is None:
            pm = ProbModel(self.model, max(TREE_MIN_BRANCH_LEN, key))
            with np.errstate(divide="ignore"):
                m = self._pm[key] = (np.log(np.asarray(pm.subMat, float)), np.log(np.asarray(pm.insVec, float)),
                                     np.log(np.asarray(pm.cptWeight, float)), trans_logprob(pm))
        return m

    def refine_node(self, rows, node):
        """Refiner::refine (history, node), refiner.cpp:127-171: the history after re-aligning `node` to its parent."""
        tree = self.tree
        parent = int(tree.parent[node])
        bits = ~np.isin(rows, GAPS)
        pb, nb = bits[parent], bits[node]
        # envelope of the old parent:node alignment (pairPath + GuideAlignmentEnvelope): the key of a position is the number of
        # matched columns up to and including its own
        match = pb & nb
        kx = np.concatenate([[0], np.cumsum(match)[pb]]).astype(np.int32)
        ky = np.concatenate([[0], np.cumsum(match)[nb]]).astype(np.int32)
        post = self.sp.excluded_posteriors(np.ascontiguousarray(rows.T), [(parent, node), (node, parent)])
        log_sub, log_ins, log_cw, lpt = self._branch_model(tree.branch_length[node])
        _, path = self.ctx.branch_align(log_sub, log_ins, log_cw, lpt, post[0][pb], post[1][nb], self.max_distance, kx, ky)
        # merge { parent's clade, new branch alignment, node's clade } (refiner.cpp:146-147)
        inside = np.zeros(tree.n_nodes, bool)
        inside[self.subtree[node]] = True
        comps = []
        for sel in (np.flatnonzero(~inside), None, np.flatnonzero(inside)):
            if sel is None:
                comps.append(([parent, node], np.stack([path & 1, (path >> 1) & 1]).astype(np.uint8)))
            else:
                sub = bits[sel]
                comps.append((sel.tolist(), sub[:, sub.any(axis=0)].astype(np.uint8)))   # alignPathRemoveEmptyColumns
        lens = bits.sum(axis=1)
        new_bits = capi.align_path_merge(comps, lens)
        new_rows = np.full(new_bits.shape, ord("-"), np.uint8)
        for r in range(tree.n_nodes):
            new_rows[r, np.flatnonzero(new_bits[r])] = rows[r][bits[r]]
        return new_rows

    def refine(self, rows):
        """Refiner::refine (history), refiner.cpp:173-194.  Returns (rows, log-likelihood)."""
        n = self.tree.n_nodes
        best = np.ascontiguousarray(rows, np.uint8)
        best_lp = self.log_likelihood(best)
        node = since = 0
        while since < n - 1:
            new = self.refine_node(best, node)
            new_lp = self.log_likelihood(new)
            self.moves += 1
            if new_lp > best_lp:
                best, best_lp, since = new, new_lp, 0
                self.accepted += 1
            else:
                since += 1
            node = (node + 1) % (n - 1)   # skip the root
        return best, best_lp


def refine(ctx, model, tree, rows, max_distance=DEFAULT_MAX_DISTANCE_FROM_GUIDE):
    r = Refiner(ctx, model, tree, max_distance)
    try:
        return r.refine(rows)
    finally:
        r.close()
