"""2-GPU check of reconstruct_distributed (torchrun --nproc-per-node 2): the level-parallel reconstruction over
two ranks must equal the single-rank one.  Prints one line per rank."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch, torch.distributed as dist
from historian_b200 import capi, model as M, recon as R, simulator as S

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
tree = S.Tree.balanced(32, 0.05)
lg = M.RateModel.preset("lg")
gapped, _ = S.simulate_family(lg, tree, 300, 2)
leaves, guide = S.leaf_guide(tree, gapped)
seqs = {v: bytes(row[row != ord("-")]).decode() for v, row in zip(leaves, guide)}
grows = {v: row != ord("-") for v, row in zip(leaves, guide)}
ctx = capi.Context(local)

def gather(mine):
    out = [None] * world
    dist.all_gather_object(out, mine)
    return out

t0 = time.perf_counter()
serial = R.Reconstructor(ctx, lg, band=20, max_profile_states=1).reconstruct(tree, seqs, grows)
t1 = time.perf_counter()
dist.barrier()
t2 = time.perf_counter()
par = R.reconstruct_distributed(R.Reconstructor(ctx, lg, band=20, max_profile_states=1), tree, seqs, grows, rank, world, gather)
dist.barrier()
t3 = time.perf_counter()
same = serial[1] == par[1] and R.gapped_rows(serial[0], tree, seqs) == R.gapped_rows(par[0], tree, seqs)
print("rank %d/%d: identical=%s lp=%.4f serial %.2fs distributed %.2fs" % (rank, world, same, par[1], t1 - t0, t3 - t2), flush=True)
assert same
dist.destroy_process_group()
