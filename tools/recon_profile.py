import os, sys, cProfile, pstats
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from historian_b200 import capi, model as M, recon as R, simulator as S
tree = S.Tree.balanced(32, 0.05)
lg = M.RateModel.preset("lg")
gapped, _ = S.simulate_family(lg, tree, 1000, 1)
leaves, guide = S.leaf_guide(tree, gapped)
seqs = {v: bytes(row[row != ord("-")]).decode() for v, row in zip(leaves, guide)}
grows = {v: row != ord("-") for v, row in zip(leaves, guide)}
ctx = capi.Context(0)
rec = R.Reconstructor(ctx, lg, band=20, max_profile_states=1)
rec.reconstruct(tree, seqs, grows)
rec = R.Reconstructor(ctx, lg, band=20, max_profile_states=1)
pr = cProfile.Profile(); pr.enable()
rec.reconstruct(tree, seqs, grows)
pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(28)
