"""Partitioned reconstruction of ONE family over the GPUs of a box (strong scaling), launched with torchrun:
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/recon_ngpu.py N_LEAVES ROOT_LEN
Every rank simulates the same family; rank 0 also runs the single-rank native driver and compares the alignments."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import torch.distributed as dist
from historian_b200 import capi, model as M, recon as R, simulator as S

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n_leaves, root_len = int(sys.argv[1]), int(sys.argv[2])
tree = S.Tree.balanced(n_leaves, 0.05)
lg = M.RateModel.preset("lg")
gapped, _ = S.simulate_family(lg, tree, root_len, 1)
leaves, guide = S.leaf_guide(tree, gapped)
seqs = {v: bytes(row[row != ord("-")]).decode() for v, row in zip(leaves, guide)}
grows = {v: row != ord("-") for v, row in zip(leaves, guide)}
ctx = capi.Context(local)
dev = torch.device("cuda", local)
owner = R.plan_owners(tree, world)
for rep in range(3):
    nat = R.NativeReconstructor(ctx, lg, band=20, max_profile_states=1)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    bits, lp, rec = R.reconstruct_partitioned(nat, tree, seqs, grows, rank, world, dist, dev)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t1 = time.perf_counter()
    if rank == int(owner[tree.root]):
        print("rank %d of %d: %d leaves x %d partitioned reconstruction %.3f s, %d columns, lp %.3f" % (rank, world, n_leaves, root_len, t1 - t0, bits.shape[1], lp), flush=True)
        print("  stages ms (this rank): " + ", ".join("%s %.1f" % kv for kv in nat.stage_ms.items()), flush=True)
        keep = bits
    rec.close()
if rank == int(owner[tree.root]):
    nat = R.NativeReconstructor(ctx, lg, band=20, max_profile_states=1)
    t0 = time.perf_counter()
    bits1, lp1, rec1 = nat.reconstruct(tree, seqs, grows)
    t1 = time.perf_counter()
    print("single rank: %.3f s; identical alignment: %s; lp equal: %s" % (t1 - t0, bits1.shape == keep.shape and bool((bits1 == keep).all()), lp1 == lp), flush=True)
    rec1.close()
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
