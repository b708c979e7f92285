"""Sibling merges over several GPUs (historian_b200/recon.py: plan_owners, level_moves, exchange_profiles, gather_stores,
reconstruct_partitioned).  CPU: the placement plan, and the point-to-point exchange with world_size 2
over gloo against a stand-in for the device object (the real transfers are NCCL send/recv of device buffers).
GPU (one device): the partitioned driver on a single rank equals the plain native driver."""
import ctypes
import os

import numpy as np
import pytest

from historian_b200 import recon as R, simulator as S


def test_plan_keeps_subtrees_together():
    tree = S.Tree.balanced(16, 0.1)
    owner = R.plan_owners(tree, 4)
    assert sorted(set(owner[tree.leaves()].tolist())) == [0, 1, 2, 3]
    moves = [m for lv in R.merge_levels(tree) for m in R.level_moves(tree, lv, owner)]
    assert len(moves) == 3                       # world - 1 profiles cross ranks in a balanced tree
    for v in range(tree.n_nodes):
        if tree.children[v]:
            assert owner[v] in [owner[c] for c in tree.children[v]]
    # a caterpillar: the growing spine never moves, only single leaves do
    n = 9
    parent, bl, names = [], [], []
    spine = None
    for k in range(n):
        parent.append(-1); bl.append(0.1); names.append("s%d" % k)
        leaf = len(parent) - 1
        if spine is None:
            spine = leaf
        else:
            parent.append(-1); bl.append(0.1); names.append("")
            v = len(parent) - 1
            parent[spine] = parent[leaf] = v
            spine = v
    cat = S.Tree(parent, bl, names)
    own = R.plan_owners(cat, 3)
    for lv in R.merge_levels(cat):
        for c, src, dst in R.level_moves(cat, lv, own):
            assert cat.is_leaf(c)
    assert R.plan_owners(cat, 1).max() == 0


class FakeRecon:
    """Stands in for capi.Recon in the exchange: a 'profile' is a (blob, table) pair of byte arrays."""

    def __init__(self, rank):
        self.rank, self.nodes, self.stores, self.released = rank, {}, {}, []

    def make(self, node):
        rng = np.random.default_rng(node)
        self.nodes[node] = (rng.integers(0, 255, 10 + node, dtype=np.uint8), rng.integers(0, 255, 64 * (1 + node), dtype=np.uint8))
        self.stores[node] = self.rank

    def export_sizes(self, node):
        return self.nodes[node][0].size, self.nodes[node][1].size

    def export_node(self, node, dev_out_ptr):
        blob, table = self.nodes[node]
        ctypes.memmove(dev_out_ptr, table.ctypes.data, table.size)
        return blob

    def import_node(self, node, blob, dev_table_ptr):
        n = 64 * (1 + node)
        table = np.empty(n, np.uint8)
        ctypes.memmove(table.ctypes.data, dev_table_ptr, n)
        self.nodes[node] = (np.array(blob), table)

    def release_node(self, node):
        self.released.append(node)
        del self.nodes[node]

    def export_stores(self):
        return np.array(sorted(self.stores), np.uint8)

    def import_stores(self, blob):
        for v in blob.tolist():
            self.stores.setdefault(v, "imported")


def _gloo_worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    tree = S.Tree.balanced(8, 0.1)
    owner = R.plan_owners(tree, world)
    rec = FakeRecon(rank)
    for v in tree.leaves():
        if owner[v] == rank:
            rec.make(v)
    built = []
    for nodes in R.merge_levels(tree):
        R.exchange_profiles(rec, R.level_moves(tree, nodes, owner), rank, dist, torch.device("cpu"))
        for v in nodes:
            if owner[v] == rank:
                assert all(c in rec.nodes for c in tree.children[v]), "children of %d not on rank %d" % (v, rank)
                for c in tree.children[v]:
                    ref = FakeRecon(-1)
                    ref.make(c)
                    assert np.array_equal(rec.nodes[c][0], ref.nodes[c][0]) and np.array_equal(rec.nodes[c][1], ref.nodes[c][1])
                    del rec.nodes[c]
                rec.make(v)
                built.append(v)
    root_rank = int(owner[tree.root])
    R.gather_stores(rec, root_rank, rank, world, dist, torch.device("cpu"))
    q.put((rank, built, rec.released, sorted(rec.stores) if rank == root_rank else None))
    dist.destroy_process_group()


def test_exchange_world2_gloo():
    import torch.multiprocessing as mp
    ctxm = mp.get_context("spawn")
    q = ctxm.Queue()
    port = 33500 + os.getpid() % 2000
    procs = [ctxm.Process(target=_gloo_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    tree = S.Tree.balanced(8, 0.1)
    built = sorted(v for _, b, _, _ in got for v in b)
    assert built == [v for v in range(tree.n_nodes) if tree.children[v]]      # every merge ran exactly once
    assert sum(len(rel) for _, _, rel, _ in got) == 1                           # one profile crossed (world - 1)
    stores = [s for _, _, _, s in got if s is not None]
    assert len(stores) == 1 and stores[0] == list(range(tree.n_nodes))         # the root's rank holds every path store


@pytest.mark.gpu
def test_partitioned_driver_single_rank(ctx):
    import tempfile
    import torch
    from test_recon_gpu import _simulate
    with tempfile.TemporaryDirectory() as tmp:
        tree, lg, seqs, grows, _, _ = _simulate(tmp, 12, 70, 4)
    a = R.NativeReconstructor(ctx, lg, band=10, max_profile_states=1)
    bits_a, lp_a, rec_a = a.reconstruct(tree, seqs, grows)
    b = R.NativeReconstructor(ctx, lg, band=10, max_profile_states=1)
    bits_b, lp_b, rec_b = R.reconstruct_partitioned(b, tree, seqs, grows, 0, 1, None, torch.device("cuda", 0))
    assert lp_a == lp_b and np.array_equal(bits_a, bits_b)
    # a profile survives export -> import on the same device unchanged
    c = R.NativeReconstructor(ctx, lg, band=10, max_profile_states=1)
    kept = {}

    def on_group(nodes, rec):
        if tree.root not in nodes and not kept:
            v = nodes[0]
            nb, db = rec.export_sizes(v)
            table = torch.empty(db, dtype=torch.uint8, device="cuda")
            blob = rec.export_node(v, table.data_ptr())
            before = rec.profile(v)
            rec.release_node(v)
            rec.import_node(v, blob, table.data_ptr())
            after = rec.profile(v)
            for f in ("is_emit", "trans_src", "trans_dest", "trans_lp", "lp_absorb", "closest_leaf_pos"):
                assert np.array_equal(getattr(before, f), getattr(after, f)), f
            kept[v] = True

    bits_c, lp_c, rec_c = c.reconstruct(tree, seqs, grows, on_group=on_group)
    assert kept and lp_c == lp_a and np.array_equal(bits_c, bits_a)
    # the stores of a finished reconstruction can be re-imported (no-ops for the ones this rank built)
    rec_c.import_stores(rec_c.export_stores())
    assert np.array_equal(rec_c.alignment(), bits_a)
    for r in (rec_a, rec_b, rec_c):
        r.close()
