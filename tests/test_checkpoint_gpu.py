"""Checkpoint / resume of a partial progressive reconstruction (second half): the flattened-profile
format of hist_recon_export_node / hist_recon_export_stores written to a file (recon.save_checkpoint) after some of the
schedule, read into a fresh hist_recon (recon.load_checkpoint) and continued -- the final alignment must equal the
uninterrupted run's, in deterministic mode and (with the generator's state saved) in sampled mode."""
import os

import numpy as np
import pytest

import tempfile

from historian_b200 import capi, recon as R
from test_recon_gpu import _simulate

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("samples,max_states,cut", [(10, 1, 0.5), (10, 1, 0.25), (4, 400, 0.6)])
def test_resume_equals_the_uninterrupted_run(ctx, tmp_path, samples, max_states, cut):
    with tempfile.TemporaryDirectory() as tmp:
        tree, lg, seqs, grows, _, _ = _simulate(tmp, 20, 60, 5)
    nat = R.NativeReconstructor(ctx, lg, band=10, profile_samples=samples, max_profile_states=max_states)
    want, lp_want, rec0 = nat.reconstruct(tree, seqs, grows)
    rec0.close()
    groups = nat.schedule(tree)
    k = max(1, int(len(groups) * cut))
    leaves = tree.leaves()

    def fresh(leaf_subset):
        rec = capi.Recon(ctx, tree.n_nodes, lg.alphabet, lg.components)
        rec.set_guide(leaves, np.array([grows[v] for v in leaves], np.uint8))
        if leaf_subset:
            rec.add_leaves(leaf_subset, [seqs[v] for v in leaf_subset])
        return rec

    rec = fresh(leaves)
    nat1 = R.NativeReconstructor(ctx, lg, band=10, profile_samples=samples, max_profile_states=max_states)
    bits, _, _ = nat1.reconstruct(tree, seqs, grows, rec=rec, groups=groups[:k])
    assert bits is None
    done = {v for g in groups[:k] for v in g}
    live = [v for v in done if int(tree.parent[v]) not in done]
    path = str(tmp_path / "ckpt.npz")
    R.save_checkpoint(path, rec, live, k, nat1.mt_state)
    rec.close()
    assert os.path.getsize(path) > 0

    rec2 = fresh([v for v in leaves if int(tree.parent[v]) not in done])
    live2, k2, mt = R.load_checkpoint(path, rec2)
    assert sorted(live2) == sorted(live) and k2 == k
    nat2 = R.NativeReconstructor(ctx, lg, band=10, profile_samples=samples, max_profile_states=max_states)
    got, lp_got, _ = nat2.reconstruct(tree, seqs, grows, rec=rec2, groups=groups[k:], mt_state=mt)
    rec2.close()
    assert lp_got == lp_want
    assert got.shape == want.shape and np.array_equal(got, want)
