"""Level-parallel driver of the progressive reconstruction (historian_b200/recon.py: merge_levels, run_levels,
reconstruct_distributed; SURVEY.md 8(e) "independent sibling merges at the same guide-tree depth").
CPU: scheduling + exchange with world_size 2 over gloo (a stub merge stands in for the device).  GPU: the
level-ordered driver gives the same reconstruction as the postorder loop."""
import os
import tempfile

import numpy as np
import pytest

from historian_b200 import recon as R, simulator as S


def test_levels_respect_dependencies():
    for n in (2, 5, 8, 13):
        tree = S.Tree.balanced(n, 0.1)
        seen = set(tree.leaves())
        levels = R.merge_levels(tree)
        assert sorted(v for lv in levels for v in lv) == [v for v in range(tree.n_nodes) if tree.children[v]]
        for lv in levels:
            for v in lv:
                assert all(c in seen for c in tree.children[v])
            seen.update(lv)
        assert levels[-1] == [tree.root]


def _stub_merge(node, a, b):
    return "(%s,%s)%d" % (a, b, node)


def _gloo_worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    tree = S.Tree.balanced(11, 0.1)
    calls = []

    def merge(node, a, b):
        calls.append(node)
        return _stub_merge(node, a, b)

    def gather(mine):
        out = [None] * world
        dist.all_gather_object(out, mine)
        return out

    val = R.run_levels(tree, lambda v: tree.names[v], merge, rank, world, gather)
    q.put((rank, calls, val[tree.root]))
    dist.destroy_process_group()


def test_run_levels_world2_gloo():
    import torch.multiprocessing as mp
    ctxm = mp.get_context("spawn")
    q = ctxm.Queue()
    port = 31500 + os.getpid() % 2000
    procs = [ctxm.Process(target=_gloo_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    tree = S.Tree.balanced(11, 0.1)
    serial = R.run_levels(tree, lambda v: tree.names[v], _stub_merge)[tree.root]
    calls = {}
    for rank, c, root in got:
        assert root == serial                      # every rank ends with the complete result
        calls[rank] = c
    internal = sorted(v for v in range(tree.n_nodes) if tree.children[v])
    assert sorted(calls[0] + calls[1]) == internal  # each merge ran exactly once
    assert calls[0] and calls[1]


@pytest.mark.gpu
def test_level_driver_matches_postorder_loop(ctx):
    from historian_b200 import model as M
    tree = S.Tree.balanced(12, 0.1)
    lg = M.RateModel.preset("lg")
    lg.insRate = lg.delRate = 0.05
    gapped, _ = S.simulate_family(lg, tree, 70, 4)
    lg = M.RateModel.preset("lg")
    leaves, guide = S.leaf_guide(tree, gapped)
    seqs = {v: bytes(row[row != ord("-")]).decode() for v, row in zip(leaves, guide)}
    grows = {v: row != ord("-") for v, row in zip(leaves, guide)}
    for kw in (dict(max_profile_states=1), dict(max_profile_states=300, min_post_prob=0.01)):
        a = R.Reconstructor(ctx, lg, band=10, **kw).reconstruct(tree, seqs, grows)
        b = R.reconstruct_distributed(R.Reconstructor(ctx, lg, band=10, **kw), tree, seqs, grows)
        assert a[1] == b[1]
        assert R.gapped_rows(a[0], tree, seqs) == R.gapped_rows(b[0], tree, seqs)
    with pytest.raises(ValueError):
        R.reconstruct_distributed(R.Reconstructor(ctx, lg, band=10, max_profile_states=300), tree, seqs, grows, 0, 2, lambda m: [m, {}])
