"""E-step driver of `historian count|fit -recon/-stockrecon`: the column loop of
EigenCounts::accumulateSubstitutionCounts (model.cpp:900-915) sharded over GPUs, plus
IndelCounts::accumulateIndelCounts (model.cpp:847-898) and EigenCounts::transform (model.cpp:924-931).

Partitioning: alignment columns are independent given the model, so rank r of W takes the
columns r, r+W, r+2W, ... (column_shard): interleaving balances the load, because the cost of a column grows
with its number of ungapped rows and dense (root-derived) and sparse (insertion) columns come in runs.  The
model, eigen-system and tree are replicated.  The only
exchange is ONE all-reduce (sum) of the flat EigenCounts buffer (7 + C*A + 2*C*A^2 doubles), issued on the
device-resident buffer the count kernel accumulated into (torch.distributed, NCCL over NVLink); the M-step
is replicated.  The per-branch indel counts are sharded by branch (each branch is a sequential scan over
the columns) and travel in the same buffer.
"""
import math

import numpy as np


def column_shard(n_cols, rank, world):
    """The columns of rank `rank` as a slice: rank, rank + world, ...  (a partition of range(n_cols))."""
    return slice(int(rank), int(n_cols), int(world))


def count_sharded(compute, n_cols, rank=0, world=1, all_reduce=None):
    """Runs `compute(shard)` (-> flat counts of this rank's columns, any array type the reducer understands)
    on this rank's shard (a slice of the column index) and sums over ranks with `all_reduce(buf)` (in place)."""
    buf = compute(column_shard(n_cols, rank, world))
    if world > 1:
        if all_reduce is None:
            raise ValueError("world > 1 needs an all_reduce")
        all_reduce(buf)
    return buf


def decay_wait_time(rate, t):  # decayWaitTime, model.cpp:24-26
    return 1.0 / rate - t / (math.exp(rate * t) - 1.0)


def indel_trans_logprob(ins_rate, del_rate, ins_ext, del_ext, t):
    """log ProbModel::transProb(src, dest) (model.cpp:400-447) for src in (Match, Insert, Delete) and
    dest in (Match, Insert, Delete, End), vectorised over branch lengths t -> [len(t), 3, 4]."""
    t = np.atleast_1d(np.asarray(t, float))
    ins, dele = 1 - np.exp(-ins_rate * t), 1 - np.exp(-del_rate * t)
    one, zero = np.ones_like(t), np.zeros_like(t)
    p = np.stack([
        np.stack([(1 - ins) * (1 - dele), ins, (1 - ins) * dele, 1 - ins], -1),
        np.stack([(1 - ins_ext) * (1 - dele), ins_ext * one, (1 - ins_ext) * dele, (1 - ins_ext) * one], -1),
        np.stack([(1 - del_ext) * one, zero, del_ext * one, (1 - del_ext) * one], -1)], 1)
    with np.errstate(divide="ignore"):
        return np.log(p)


def indel_counts(columns, parent, branch_length, ins_rate, del_rate, ins_ext, del_ext, weight=1.0, branches=None):
    """IndelCounts::accumulateIndelCounts over every branch of the tree (model.cpp:847-898): the seven
    slots (ins, del, insExt, delExt, insTime, delTime, lp) of the flat EigenCounts buffer.
    A per-branch three-state scan over the alignment columns, vectorised over the branches (host side:
    O(N * n_cols) booleans).  `branches`: subset of child nodes to process (multi-GPU shard)."""
    columns = np.asarray(columns)
    present = ~np.isin(columns, [ord("-"), ord(".")])           # [n_cols, N]
    parent = np.asarray(parent)
    kids = np.arange(len(parent) - 1) if branches is None else np.asarray(branches, np.int64)
    if len(kids) == 0:
        return np.zeros(7)
    t = np.asarray(branch_length, float)[kids]
    with np.errstate(divide="ignore", invalid="ignore"):
        ins_wait = 1.0 / ins_rate - t / np.expm1(ins_rate * t)
        del_wait = 1.0 / del_rate - t / np.expm1(del_rate * t)
    lpt = indel_trans_logprob(ins_rate, del_rate, ins_ext, del_ext, t)      # [B, 3, 4]
    state = np.zeros(len(kids), np.int64)                        # Match
    out = np.zeros((7, len(kids)))
    b = np.arange(len(kids))
    par_rows, kid_rows = present[:, parent[kids]], present[:, kids]
    for col in range(len(columns)):
        p, c = par_rows[col], kid_rows[col]
        nxt = np.where(p & c, 0, np.where(c, 1, 2))
        live = p | c
        same = live & (state == nxt)
        m, i, d = nxt == 0, nxt == 1, nxt == 2
        out[4] += np.where(same & m, t, 0) + np.where(live & i & ~same, ins_wait, 0)
        out[5] += np.where(same & m, t, 0) + np.where(live & d & ~same, del_wait, 0)
        out[2] += same & i
        out[3] += same & d
        out[0] += live & i & ~same
        out[1] += live & d & ~same
        out[6] += np.where(live, lpt[b, state, nxt], 0)
        state = np.where(live, nxt, state)
    out[6] += lpt[b, state, 3]
    return weight * out.sum(axis=1)


class ColumnCounter:
    """Substitution E-step on one GPU (one process per GPU); see module docstring."""

    def __init__(self, ctx, model, tree):
        self.ctx = ctx
        self.sp = ctx.sumprod(model, tree)
        self.model, self.tree = model, tree

    def close(self):
        self.sp.close()

    def count(self, columns, rank=0, world=1, weights=None):
        """Flat EigenCounts (numpy) of the whole alignment `columns` [n_cols, N], summed over ranks.
        With world > 1 the torch.distributed default group (backend nccl) must be initialised."""
        import torch
        columns = np.ascontiguousarray(columns, np.uint8)
        dev = torch.device("cuda", self.ctx.device)

        def compute(shard):
            buf = torch.zeros(self.sp.counts_size, dtype=torch.float64, device=dev)
            mine = np.ascontiguousarray(columns[shard])
            if len(mine):
                torch.cuda.synchronize(dev)
                w = None if weights is None else np.ascontiguousarray(np.asarray(weights, np.float64)[shard])
                self.sp.columns(mine, w, counts=buf, col_loglike=False)
            return buf

        def all_reduce(buf):
            import torch.distributed as dist
            self.ctx.synchronize()
            dist.all_reduce(buf)

        buf = count_sharded(compute, len(columns), rank, world, all_reduce)
        self.ctx.synchronize()
        return buf.cpu().numpy()
