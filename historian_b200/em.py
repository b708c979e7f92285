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


# ------------------------------------------------------------------ M-step and the EM loop (historian fit)
class EventCounts:
    """EventCounts (model.h): indel counts (ins, del, insExt, delExt, insTime, delTime, lp), rootCount [C][A] and
    subCount [C][A][A] (diagonal = wait time), as EigenCounts::transform returns them (model.cpp:924-931)."""

    def __init__(self, indel, root, sub):
        self.indel = np.array(indel, float)
        self.root = np.array(root, float)
        self.sub = np.array(sub, float)

    @property
    def lp(self):
        return float(self.indel[6])

    def __add__(self, other):
        return EventCounts(self.indel + other.indel, self.root + other.root, self.sub + other.sub)

    def optimize(self, model, fit_indel_rates=True, fit_subst_rates=True):
        """EventCounts::optimize (model.cpp:1022-1059): the closed-form M-step, in place on a model.RateModel."""
        C, A = self.root.shape
        if fit_subst_rates:
            cpt_count = []
            for c in range(C):
                ins_norm = self.root[c].sum()
                model.insProb[c] = self.root[c] / ins_norm
                r = np.zeros((A, A))
                for i in range(A):
                    for j in range(A):
                        if j != i:
                            r[i, j] = self.sub[c][i][j] / self.sub[c][i][i]
                    r[i, i] = -(r[i].sum() - r[i, i])
                model.subRate[c] = r
                cpt_count.append(ins_norm)
            tot = float(sum(cpt_count))
            model.cptWeight = [float(x) / tot for x in cpt_count]
        if fit_indel_rates:
            ins, dele, ins_ext, del_ext, ins_time, del_time = self.indel[:6]
            model.insRate = ins / ins_time
            model.delRate = dele / del_time
            model.insExtProb = ins_ext / (ins_ext + ins)
            model.delExtProb = del_ext / (del_ext + dele)
        return model

    def expected_log_likelihood(self, model):
        """EventCounts::expectedLogLikelihood (model.cpp:1083-1104)."""
        def xlogy(x, y):
            return 0.0 if x == 0 else x * math.log(y)
        ins, dele, ins_ext, del_ext, ins_time, del_time = self.indel[:6]
        lp = (-model.insRate * ins_time + xlogy(ins, model.insRate) - model.delRate * del_time + xlogy(dele, model.delRate)
              + xlogy(ins_ext, model.insExtProb) + xlogy(ins, 1 - model.insExtProb)
              + xlogy(del_ext, model.delExtProb) + xlogy(dele, 1 - model.delExtProb))
        C, A = self.root.shape
        for c in range(C):
            sr = np.asarray(model.subRate[c])
            for i in range(A):
                lp += xlogy(self.root[c][i], model.insProb[c][i]) - (-sr[i, i]) * self.sub[c][i][i]
                for j in range(A):
                    if i != j:
                        lp += xlogy(self.sub[c][i][j], sr[i, j])
        return lp


def laplace_prior(components, alph_size):
    """The default prior of `historian fit` (recon.cpp:1277-1280): one pseudocount on every event count and wait time."""
    return EventCounts([1, 1, 1, 1, 1, 1, 0], np.ones((components, alph_size)), np.ones((components, alph_size, alph_size)))


def log_prior(prior, model, fit_indel_rates=True, fit_subst_rates=True):
    """EventCounts::logPrior (model.cpp:1060-1078) with logGammaPdf / logBetaPdf / logDirichletPdf (logsumexp.cpp:101-115)."""
    from scipy import stats
    lp = 0.0
    ins, dele, ins_ext, del_ext, ins_time, del_time = prior.indel[:6]
    if fit_indel_rates:
        lp += stats.gamma.logpdf(model.insRate, ins + 1, scale=1.0 / ins_time) + stats.gamma.logpdf(model.delRate, dele + 1, scale=1.0 / del_time)
        lp += stats.beta.logpdf(model.insExtProb, ins_ext + 1, ins + 1) + stats.beta.logpdf(model.delExtProb, del_ext + 1, dele + 1)
    if fit_subst_rates:
        C, A = prior.root.shape
        for c in range(C):
            p = np.asarray(model.insProb[c], float)
            lp += stats.dirichlet.logpdf(p / p.sum(), prior.root[c] + 1)
            sr = np.asarray(model.subRate[c])
            for i in range(A):
                for j in range(A):
                    if i != j:
                        lp += stats.gamma.logpdf(sr[i, j], prior.sub[c][i][j] + 1, scale=1.0 / prior.sub[c][i][i])
    return float(lp)


def count_alignment(ctx, model, tree, columns, rank=0, world=1, fit_indel_rates=True, fit_subst_rates=True):
    """Reconstructor::count (recon.cpp:1284-1291) for one reconstructed alignment: EigenCounts::accumulateCounts
    (model.cpp:917-922) -- substitution counts on the device, columns sharded over the ranks with one all-reduce;
    indel counts by branch on the host -- then EigenCounts::transform.  tree: descs.TreeTopology."""
    from . import descs as D
    spm = D.SumProdModel.from_rate_model(model)
    cc = ColumnCounter(ctx, spm, tree)
    try:
        if fit_subst_rates:
            counts = cc.count(columns, rank, world)
        else:
            counts = np.zeros(cc.sp.counts_size)
        if fit_indel_rates:
            counts[:7] += indel_counts(columns, tree.parent, tree.branch_length, model.insRate, model.delRate, model.insExtProb, model.delExtProb)
        root, sub = cc.sp.transform_counts(counts)
    finally:
        cc.close()
    return EventCounts(counts[:7], root, sub)


def fit(ctx, model, tree, columns, max_iter=100, min_improvement=1e-3, rank=0, world=1, fit_indel_rates=True, fit_subst_rates=True,
        prior="laplace", counter=None, log=None):
    """Reconstructor::fit (recon.cpp:1385-1408) on one reconstructed alignment: E-step (count), stop when log-likelihood +
    log-prior improves by less than min_improvement (relative), else the closed-form M-step on data + prior counts.
    prior: "laplace" (the reference's default pseudocounts), None (-nolaplace) or an EventCounts.
    `counter(model) -> EventCounts` replaces the device E-step (tests run the loop against the CPU oracle).
    Returns (model, [(log-likelihood, log-prior) per iteration])."""
    if isinstance(prior, str):
        prior = laplace_prior(model.components, model.alphabetSize)
    lp_last, lps = -math.inf, []
    for it in range(max_iter):
        ev = counter(model) if counter is not None else count_alignment(ctx, model, tree, columns, rank, world, fit_indel_rates, fit_subst_rates)
        lp_prior = log_prior(prior, model, fit_indel_rates, fit_subst_rates) if prior is not None else 0.0
        lps.append((ev.lp, lp_prior))
        if log is not None:
            log("EM iteration #%d: log-likelihood (%.6f) + log-prior (%.6f) = %g" % (it + 1, ev.lp, lp_prior, ev.lp + lp_prior))
        if ev.lp + lp_prior <= lp_last + abs(lp_last) * min_improvement:
            break
        (ev + prior if prior is not None else ev).optimize(model, fit_indel_rates, fit_subst_rates)
        lp_last = ev.lp + lp_prior
    return model, lps


# ------------------------------------------------------------------ #=GF LP: TreeAlignFuncs::logLikelihood
def root_log_likelihood(model, root_row):
    """TreeAlignFuncs::rootLogLikelihood (sampler.cpp:372-380): geometric length of the root sequence with
    rootExtProb = insExtProb (sampler.h:52).  root_row: the root's characters over the alignment columns."""
    root_len = int(np.count_nonzero(~np.isin(np.asarray(root_row), [ord("-"), ord(".")])))
    ext = model.insExtProb
    return math.log(1.0 - ext) + math.log(ext) * root_len


def tree_align_log_likelihood(model, tree, columns, column_log_likelihoods):
    """TreeAlignFuncs::logLikelihood(model, tree, gapped) (sampler.cpp:404-411), the `#=GF LP` line of the Stockholm
    output (recon.cpp:1138): root + indels + substitutions.
      indels: logBranchPathLikelihood summed over the branches (sampler.cpp:382-392, 439-449) -- slot 6 of indel_counts;
      substitutions: the sum of the Felsenstein column likelihoods (fillUp, sampler.cpp:394-402), which the caller
      computes -- on the device with SumProd.columns(columns, flags=0)["col_loglike"] (hist_sumprod_columns).
    columns: uint8 [n_cols, N] characters in tree node order (root last).  Returns (lp, lp_root, lp_indels, lp_subst)."""
    columns = np.asarray(columns)
    lp_root = root_log_likelihood(model, columns[:, len(tree.parent) - 1])
    lp_gaps = float(indel_counts(columns, tree.parent, tree.branch_length, model.insRate, model.delRate, model.insExtProb,
                                 model.delExtProb)[6])
    lp_sub = float(np.sum(column_log_likelihoods))
    return lp_root + lp_gaps + lp_sub, lp_root, lp_gaps, lp_sub


def read_stockholm_recon(path):
    """A reconstruction as `historian recon -output stockholm` writes it: (names, rows, gf) with gf the #=GF tags
    (NH = tree, LP = log-likelihood).  Rows of blocks are concatenated."""
    rows, order, gf = {}, [], {}
    with open(path) as f:
        for line in f:
            line = line.rstrip("\n")
            if line.startswith("#=GF"):
                _, tag, val = (line.split(None, 2) + [""])[:3]
                gf[tag] = gf.get(tag, "") + val
            elif line and not line.startswith("#") and not line.startswith("//"):
                name, seq = line.split()
                if name not in rows:
                    order.append(name)
                    rows[name] = ""
                rows[name] += seq
    return order, [rows[n] for n in order], gf
