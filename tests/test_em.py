"""E-step driver (historian_b200/em.py): column sharding, the all-reduce of the flat EigenCounts buffer
(world_size-2 gloo on CPU), and the reference's own `testcount` golden (Makefile:317,
data/testcount.count.json: 6 significant digits) for indel counts, substitution counts, wait times,
root counts and the total log-likelihood."""
import json
import os

import numpy as np
import pytest

import helpers as H
import oracle_lib as O
from historian_b200 import descs as D, em, model as M

GOLD = json.load(open(os.path.join(H.GOLDEN_DIR, "testcount_recon.json")))


def golden_inputs():
    rm = M.RateModel.read(GOLD["model"])
    cols = np.array([list(r.encode()) for r in GOLD["rows"]], np.uint8).T.copy()   # [n_cols, N]
    tree = D.TreeTopology(GOLD["parent"], GOLD["branch_length"])
    return rm, D.SumProdModel.from_rate_model(rm), tree, cols


def sig6(a, b):
    return abs(a - b) <= 6e-6 * max(abs(a), abs(b)) + 1e-12


def check_against_golden(rm, counts, root, sub):
    e = GOLD["expect"]
    for k, slot in (("ins", 0), ("del", 1), ("insExt", 2), ("delExt", 3), ("insTime", 4), ("delTime", 5)):
        assert sig6(counts[slot], e["indel"][k]), (k, counts[slot], e["indel"][k])
    assert sig6(counts[6], e["logLikelihood"]), (counts[6], e["logLikelihood"])
    al = rm.alphabet
    for i, a in enumerate(al):
        assert sig6(root[0][i], e["sub"]["root"][a])
        assert sig6(sub[0][i][i], e["sub"]["wait"][a])
        for j, b in enumerate(al):
            if i != j:
                assert sig6(sub[0][i][j], e["sub"]["sub"][a][b]), (a, b, sub[0][i][j], e["sub"]["sub"][a][b])


def test_counts_match_the_reference_testcount_golden_cpu():
    rm, spm, tree, cols = golden_inputs()
    counts = O.sumprod_columns(spm, tree, cols, flags=O.SP_COUNTS)["counts"]
    counts[:7] += em.indel_counts(cols, tree.parent, tree.branch_length, rm.insRate, rm.delRate, rm.insExtProb, rm.delExtProb)
    root, sub = O.sumprod_transform_counts(spm, counts)
    check_against_golden(rm, counts, root, sub)


@pytest.mark.gpu
def test_counts_match_the_reference_testcount_golden_device():
    from historian_b200 import capi
    rm, spm, tree, cols = golden_inputs()
    ctx = capi.Context(0)
    cc = em.ColumnCounter(ctx, spm, tree)
    counts = cc.count(cols)
    counts[:7] += em.indel_counts(cols, tree.parent, tree.branch_length, rm.insRate, rm.delRate, rm.insExtProb, rm.delExtProb)
    root, sub = cc.sp.transform_counts(counts)
    check_against_golden(rm, counts, root, sub)
    cc.close()
    ctx.close()


def test_indel_counts_shard_by_branch():
    rm, spm, tree, cols = golden_inputs()
    args = (cols, tree.parent, tree.branch_length, rm.insRate, rm.delRate, rm.insExtProb, rm.delExtProb)
    full = em.indel_counts(*args)
    parts = sum(em.indel_counts(*args, branches=np.arange(r, tree.n_nodes - 1, 3)) for r in range(3))
    assert np.allclose(full, parts, rtol=1e-13)
    assert np.allclose(em.indel_counts(*args, weight=2.5), 2.5 * full)


def test_column_shard_is_a_partition():
    for n in (0, 1, 7, 8, 9, 12997):
        for w in (1, 2, 3, 8):
            idx = [np.arange(n)[em.column_shard(n, r, w)] for r in range(w)]
            assert sorted(np.concatenate(idx).tolist()) == list(range(n))
            sizes = [len(i) for i in idx]
            assert max(sizes) - min(sizes) <= 1


def _gloo_worker(rank, world, port, path, q):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    d = H.load_golden(path)
    model, tree = D.SumProdModel.from_golden(d), D.TreeTopology(d["parent"], d["branch_length"])
    cols = d["columns"]

    def compute(shard):  # the oracle stands in for the device kernel: this test covers sharding + reduction
        c = np.zeros(D.counts_size(model.components, model.alph_size))
        if len(cols[shard]):
            c = O.sumprod_columns(model, tree, np.ascontiguousarray(cols[shard]), flags=O.SP_COUNTS)["counts"]
        return torch.from_numpy(c)

    buf = em.count_sharded(compute, len(cols), rank, world, dist.all_reduce)
    if rank == 0:
        q.put(buf.numpy().copy())
    dist.destroy_process_group()


def test_column_sharded_counts_world2_gloo():
    import torch.multiprocessing as mp
    path = os.path.join(H.GOLDEN_DIR, "sumprod_dna12_gamma2_wild.npz")
    ctxm = mp.get_context("spawn")
    q = ctxm.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctxm.Process(target=_gloo_worker, args=(r, 2, port, path, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    d = H.load_golden(path)
    model, tree = D.SumProdModel.from_golden(d), D.TreeTopology(d["parent"], d["branch_length"])
    whole = O.sumprod_columns(model, tree, d["columns"], flags=O.SP_COUNTS)["counts"]
    assert np.allclose(got, whole, rtol=1e-12, atol=1e-14)
