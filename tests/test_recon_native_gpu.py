"""GPU tests of the native level-batched reconstruction (hist_recon_*, csrc/hist_recon.cu) against
  (a) the per-merge Python driver (historian_b200/recon.py: Reconstructor), itself pinned to the reference's dumps
      by tests/test_recon_gpu.py: every parent profile (states, transitions, closest-leaf coordinates: exact;
      numbers: the same device arithmetic, exact) and the final alignment must coincide, in deterministic AND in
      sampled mode (same mt19937 stream);
  (b) the reference BINARY end to end in deterministic mode: identical alignment, row by row.
"""
import os
import subprocess
import tempfile

import numpy as np
import pytest

from historian_b200 import descs as D, recon as R
from test_recon_gpu import _simulate, _read_fasta, REF_BIN

pytestmark = pytest.mark.gpu


def _dense_from_path(path, n_nodes):
    ncol = max((len(v) for v in path.values()), default=0)
    out = np.zeros((n_nodes, ncol), bool)
    for row, rp in path.items():
        out[row, :len(rp)] = rp
    return out


def _compare(ctx, tree, lg, seqs, grows, band, samples, max_states, seed=5489):
    guide = grows if band >= 0 else None
    py = R.Reconstructor(ctx, lg, band=band, profile_samples=samples, max_profile_states=max_states, seed=seed)
    path, lp_py, profs = py.reconstruct(tree, seqs, guide)
    nat = R.NativeReconstructor(ctx, lg, band=band, profile_samples=samples, max_profile_states=max_states, seed=seed)
    seen = {}

    def on_group(nodes, rec):
        for v in nodes:
            seen[v] = rec.profile(v)

    bits, lp_nat, rec = nat.reconstruct(tree, seqs, guide, on_group=on_group)
    assert lp_nat == lp_py
    for v, q in seen.items():
        p = profs[v].flat
        assert np.array_equal(q.is_emit, p.is_emit), "node %d: states" % v
        assert np.array_equal(q.trans_src, p.trans_src) and np.array_equal(q.trans_dest, p.trans_dest), "node %d: transitions" % v
        assert np.array_equal(q.trans_lp, p.trans_lp), "node %d: transition weights" % v
        em = p.is_emit.astype(bool)
        assert np.array_equal(q.lp_absorb[em], p.lp_absorb[em]), "node %d: absorb tables" % v
        if guide is not None and profs[v].leaf_rows:
            (row, coords), = profs[v].leaf_rows.items()
            assert np.array_equal(q.closest_leaf_pos[1:], coords[1:]), "node %d: closest-leaf coordinates" % v
        assert nat.lp_end[v] == [e for e in py.log if e["node"] == v][0]["lp_end"]
    assert len(seen) == tree.n_nodes - len(tree.leaves())
    want = _dense_from_path(path, tree.n_nodes)
    assert bits.shape == want.shape and np.array_equal(bits, want)
    rec.close()
    return bits


@pytest.mark.parametrize("n_leaves,root_len,seed,band", [(8, 60, 3, 20), (16, 80, 5, 5), (6, 50, 9, -1), (13, 70, 4, 10)])
def test_native_equals_python_driver_deterministic(ctx, n_leaves, root_len, seed, band):
    with tempfile.TemporaryDirectory() as tmp:
        tree, lg, seqs, grows, _, _ = _simulate(tmp, n_leaves, root_len, seed)
    _compare(ctx, tree, lg, seqs, grows, band, 10, 1)


@pytest.mark.parametrize("n_leaves,root_len,seed,band,samples,max_states", [(8, 50, 11, 10, 5, 300), (6, 40, 2, 20, 10, 0), (8, 60, 7, -1, 3, 150)])
def test_native_equals_python_driver_sampled(ctx, n_leaves, root_len, seed, band, samples, max_states):
    with tempfile.TemporaryDirectory() as tmp:
        tree, lg, seqs, grows, _, _ = _simulate(tmp, n_leaves, root_len, seed)
    _compare(ctx, tree, lg, seqs, grows, band, samples, max_states, seed=7)


@pytest.mark.parametrize("n_leaves,root_len,seed,band", [(8, 60, 3, 20), (32, 120, 6, 10)])
def test_native_matches_reference_binary(ctx, n_leaves, root_len, seed, band):
    if not os.path.exists(REF_BIN):
        pytest.skip("oracle/_ref/bin/historian not built")
    with tempfile.TemporaryDirectory() as tmp:
        tree, lg, seqs, grows, tf, gf = _simulate(tmp, n_leaves, root_len, seed)
        args = [REF_BIN, "recon", "-guide", gf, "-tree", tf, "-preset", "lg", "-profmaxstates", "1", "-output", "fasta", "-norefine",
                "-band", str(band)]
        out = subprocess.run(args, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, check=True, cwd=tmp).stdout.decode()
    ref_rows = _read_fasta(out)
    nat = R.NativeReconstructor(ctx, lg, band=band, profile_samples=10, max_profile_states=1)
    bits, lp_final, rec = nat.reconstruct(tree, seqs, grows)
    ours = R.gapped_rows_dense(bits, tree, seqs, wildcard="x")
    assert len(ref_rows) == tree.n_nodes
    for node, (name, row) in enumerate(ref_rows):
        assert ours[node] == row.lower() or ours[node] == row, "row %d (%s) differs:\n%s\n%s" % (node, name, ours[node], row)
    assert np.isfinite(lp_final)
    rec.close()


def test_level_call_rejects_bad_arguments(ctx):
    from historian_b200 import capi, model as M
    lg = M.RateModel.preset("lg")
    rec = capi.Recon(ctx, 3, lg.alphabet, 1)
    rec.add_leaves([0, 1], ["ARND", "ARD"])
    hm = M.pair_hmm(lg, 0.1, 0.1)
    hd = hm.desc()
    with pytest.raises(capi.HistError):   # node 2's closest leaf must be one of the children's
        rec.level([(2, 0, 1, 0, 1, 2)], [hd], -1, D.COLLAPSE_CHAINS | D.INCLUDE_BEST_TRACE, 0, 1, 2)
    with pytest.raises(capi.HistError):   # no best trace and no samples (forward.cpp:848)
        rec.level([(2, 0, 1, 0, 1, 0)], [hd], -1, D.COLLAPSE_CHAINS, 0, 1, 2)
    lp = rec.level([(2, 0, 1, 0, 1, 0)], [hd], -1, D.COLLAPSE_CHAINS | D.INCLUDE_BEST_TRACE, 0, 1, 2)
    assert np.isfinite(lp[0])
    bits = rec.alignment()
    assert bits.shape[0] == 3 and bits[0].sum() == 4 and bits[1].sum() == 3
    with pytest.raises(capi.HistError):   # already built
        rec.level([(2, 0, 1, 0, 1, 0)], [hd], -1, D.COLLAPSE_CHAINS | D.INCLUDE_BEST_TRACE, 0, 1, 2)
    rec.close()
