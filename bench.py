#!/usr/bin/env python3
"""Benchmark of the hot path: banded profile Forward/Backward fills (metric: cells/s, GCUPS).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Workload (BASELINE.json configs[2], the configuration the metric's target is quoted on): a
synthetic LG protein family, 1024 sequences x ~1000 residues on a balanced tree (branch length
0.05), simulated with a fixed seed (historian_b200/simulator.py: the reference simulator's
algorithm; the reference binary itself needs >10 h for 1024 leaves, see that file).  One STEP =
the Forward AND Backward fill (forward.cpp:68-223, 975-1097) of all 512 sibling merges of the
deepest guide-tree level (leaf x leaf, guide band 20: recon.cpp:954-984), one batched launch.
A cell = one in-envelope (i,j) of one matrix with all five states (forward.cpp:78-86); a step
updates 2 * sum(n_cells) cells.  N GPUs: weak scaling, every rank fills the 512 merges of its own
family (seed 1 + rank), no data-path collective (the reference loops over datasets the same way,
recon.cpp:1368-1372).

Prints ONE JSON line (rank 0).  `value`: inputs resident in HBM.  `e2e`: the same step through
the C ABI from HOST buffers (profile upload + DPMatrix preparation + fills + likelihood read-back
inside the timed region).  `roofline`: the fill kernel alone, CUDA events on its own stream.
`cpu_baseline`: the unmodified reference (oracle/_ref/bin/ref_time) on a bounded sample.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

BAND = 20
BRANCH = 0.05
N_LEAVES = 1024
ROOT_LEN = 1000
# dram__bytes_read.sum + dram__bytes_write.sum of the dominant kernel per launch, from the ncu --set full capture under profiles/ (None = not captured)
FILL_TRAFFIC_BYTES = 1.387e9   # profiles/r02_fill_chain2_ncu_details.txt: fill_chain2_kernel<5> 0.348 GB read + 1.039 GB written per launch (default family)
# FP32 work of one cell update in the chain kernel: K-term absorb dot product (K FMA) + 24 FMA/FMUL of the five state sums
# + 9 scale/emission multiplies + 5 normalisation multiplies, FMA counted as 2 flop
FILL_FLOP_PER_CELL = 2 * 20 + 2 * 24 + 9 + 5
ALG_BYTES_PER_CELL = 80.0   # 40 B written once + 40 B read once per DP cell (reference layout)
METRIC = "profile Forward/Backward cells/s (GCUPS)"
UNIT = "Gcells/s"


def make_family(seed, n_leaves=N_LEAVES, with_tree=False):
    from historian_b200 import model as M, simulator as S
    tree = S.Tree.balanced(n_leaves, BRANCH)
    lg = M.RateModel.preset("lg")
    gapped, _ = S.simulate_family(lg, tree, ROOT_LEN, seed)
    leaves, guide = S.leaf_guide(tree, gapped)
    if with_tree:
        return lg, guide, tree, leaves
    return lg, guide


def cherry_inputs(lg, guide, n_pairs):
    """Flattened inputs of the leaf x leaf merges (rows 2k, 2k+1 of the guide)."""
    from historian_b200 import descs as D, model as M
    hmm = M.pair_hmm(lg, BRANCH, BRANCH)
    out = []
    gap = ord("-")
    for k in range(n_pairs):
        r1, r2 = guide[2 * k], guide[2 * k + 1]
        x = D.Profile.from_sequence(bytes(r1[r1 != gap]).decode(), lg.alphabet, lg.components)
        y = D.Profile.from_sequence(bytes(r2[r2 != gap]).decode(), lg.alphabet, lg.components)
        env = D.Envelope.from_guide_rows(r1 != gap, r2 != gap, BAND)
        out.append((x, y, env))
    return hmm, out


def cherry_sequences(guide, n_pairs):
    """The ungapped residue strings of the leaf x leaf merges, as a reference user holds them after reading a
    FASTA file (one byte string per sequence, like FastSeq::seq)."""
    gap = ord("-")
    return [bytes(guide[r][guide[r] != gap]) for r in range(2 * n_pairs)]


def write_fasta(path, guide):
    with open(path, "w") as f:
        for k, row in enumerate(guide):
            f.write(">s%04d\n%s\n" % (k, bytes(row).decode()))


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.stop_flag, self.rows = index, threading.Event(), []

    def run(self):
        while not self.stop_flag.is_set():
            try:
                o = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits"],
                                   capture_output=True, text=True, timeout=5).stdout.strip()
                if o:
                    self.rows.append([c.strip() for c in o.split(",")])
            except Exception:
                pass
            self.stop_flag.wait(0.2)

    def summary(self):
        self.stop_flag.set()
        self.join(timeout=6)
        sm, mx, reasons = [], 0.0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx = max(mx, float(r[1]))
            except Exception:
                continue
            for n, v in zip(names, r[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons),
                "samples": len(sm)}


# ------------------------------------------------------------------ CPU side (reference arm and cpu_baseline)
REF_TIME = os.path.join(ROOT, "oracle", "_ref", "bin", "ref_time")


def run_ref_time(fasta, first, count, procs):
    """The unmodified reference's ForwardMatrix+BackwardMatrix on `count` cherries starting at
    `first`, spread over `procs` single-threaded processes (the reference has no threads).
    Returns (cells updated [both matrices], wall seconds of the slowest-finishing set)."""
    per = [count // procs + (1 if p < count % procs else 0) for p in range(procs)]
    jobs, start = [], first
    t0 = time.perf_counter()
    for p in range(procs):
        if per[p] == 0:
            continue
        jobs.append(subprocess.Popen([REF_TIME, "lg", fasta, str(BAND), str(BRANCH), str(BRANCH), str(start), str(per[p])],
                                     stdout=subprocess.PIPE, text=True))
        start += per[p]
    outs = [json.loads(j.communicate()[0]) for j in jobs]
    wall = time.perf_counter() - t0
    for j in jobs:
        if j.returncode != 0:
            raise RuntimeError("ref_time failed")
    # the time inside the reference's ForwardMatrix + BackwardMatrix constructors (the processes run side by side: the
    # slowest one bounds the set); process start, FASTA parsing and the cell-counting loop of ref_time are left out
    busy = max(o["fwd_s"] + o["bwd_s"] for o in outs)
    return 2 * sum(o["cells"] for o in outs), busy, outs, wall


def cpu_baseline(guide, sample_pairs, procs):
    if not os.path.exists(REF_TIME):
        return {"value": None, "unit": UNIT, "cores": procs, "kind": "reference",
                "sample": "oracle/_ref/bin/ref_time missing (build with __graft_entry__.build() where /root/reference exists)"}
    with tempfile.TemporaryDirectory() as tmp:
        fa = os.path.join(tmp, "leaves.fa")
        write_fasta(fa, guide[:2 * sample_pairs])
        cells, busy, _, wall = run_ref_time(fa, 0, sample_pairs, procs)
    return {"value": cells / busy / 1e9, "unit": UNIT, "cores": procs, "kind": "reference",
            "sample": "%d of the 512 leaf x leaf merges (Forward+Backward, band %d), unmodified reference ForwardMatrix/BackwardMatrix "
                      "via oracle/_ref/bin/ref_time, %d single-threaded process(es), %.1f s inside the two constructors (%.1f s wall)"
                      % (sample_pairs, BAND, procs, busy, wall)}


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    if not os.path.exists(REF_TIME):
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/bin/ref_time not built (needs /root/reference at build time)"}))
        return
    lg, guide = make_family(1)
    # one step = a bounded sample: 4 cherries per core (about 0.15 s of CPU work each)
    per_step = min(512, 4 * cores)
    with tempfile.TemporaryDirectory() as tmp:
        fa = os.path.join(tmp, "leaves.fa")
        write_fasta(fa, guide)
        for w in range(args.warmup):
            run_ref_time(fa, (w * per_step) % (512 - per_step + 1), per_step, cores)
        cells, wall = 0, 0.0
        for s in range(args.steps):
            c, t, _, _ = run_ref_time(fa, (s * per_step) % (512 - per_step + 1), per_step, cores)
            cells += c
            wall += t
    v = cells / wall / 1e9
    sample = ("%d leaf x leaf merges per step (Forward+Backward, band %d) on %d single-threaded processes; time = the slowest process's "
              "seconds inside the ForwardMatrix + BackwardMatrix constructors" % (per_step, BAND, cores))
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * wall / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic", "config": workload_config(per_step),
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "reference", "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}, "gpu_launches": 0}))


def fp32_roofline(cells_per_launch, per_launch_s):
    """The fill against the FP32 issue peak measured on this box (tools/microbench/peaks.cu, profiles/r02_peaks_microbench.json):
    SURVEY.md 8(d) names FP32/SFU issue, not HBM, as the binding roofline of the fills.  The cells hold linear-space
    mantissas, so there is no transcendental in the inner loop (MUFU is idle by design)."""
    peak = 69.8
    try:
        peak = float(json.load(open(os.path.join(ROOT, "profiles", "r02_peaks_microbench.json")))["fp32_fma_tflops"])
    except Exception:
        pass
    ach = cells_per_launch * FILL_FLOP_PER_CELL / per_launch_s / 1e12
    return {"achieved": ach, "peak": peak, "unit": "TFLOP/s", "frac": ach / peak, "flop_per_cell": FILL_FLOP_PER_CELL,
            "note": "one warp per strip walks 2000 dependent anti-diagonals: the kernel is bound by dependent-issue latency "
                    "(ncu: IPC 1.3 of 4, 0.36 eligible warps per scheduler), not by FP32 throughput"}


def workload_config(pairs):
    return {"workload": "lg_1024x1000_seed1, LEAF LEVEL ONLY: the level-0 sibling merges (leaf x leaf), Forward+Backward fill, guide band %d "
                        "(the whole tree, all levels, is the `recon` record)" % BAND,
            "family": "LG, %d leaves x %d root residues, balanced tree, branch %.2f, insrate=delrate=0.01" % (N_LEAVES, ROOT_LEN, BRANCH),
            "merges_per_step_per_gpu": pairs, "band": BAND, "alphabet": 20, "components": 1,
            "parallelism": "one copy of the family per GPU, no data-path collective (weak scaling)",
            "l2_policy": "cell planes written per step (>1 GB) exceed the 126 MB L2; no reuse between steps"}


# ------------------------------------------------------------------ EM count leg (BASELINE.json metric 3, configs[3])
EM_LEAVES, EM_ROOT_LEN = 4096, 1000


def em_family():
    """configs[3]: DNA, Jukes-Cantor with a 2-component discretised gamma, balanced 4096-leaf tree, full history
    (ancestral rows included) as `historian fit -stockrecon` reads it."""
    from historian_b200 import descs as D, model as M, simulator as S
    tree = S.Tree.balanced(EM_LEAVES, BRANCH)
    jc = M.RateModel.preset("jc").gamma(2, 1.0)
    gapped, _ = S.simulate_family(jc, tree, EM_ROOT_LEN, 1)
    cols = np.ascontiguousarray(gapped.T)                      # [n_cols, N]
    return D.SumProdModel.from_rate_model(jc), D.TreeTopology(tree.parent, tree.branch_length), cols


def em_leg(ctx, args, rank, world, local, barrier):
    """EM count columns/s: one step = the E-step substitution counts (sum-product up/down + eigen-counts,
    model.cpp:900-915) of the whole alignment, columns sharded over the ranks, one NCCL all-reduce of the
    count buffer per step (strong scaling: the alignment is fixed)."""
    import torch
    import torch.distributed as dist
    from historian_b200 import em
    model, tree, cols = em_family()
    n_cols, N = cols.shape
    mine = np.ascontiguousarray(cols[em.column_shard(n_cols, rank, world)])   # interleaved shard (load balance)
    lo, hi = 0, len(mine)
    sp = ctx.sumprod(model, tree)
    dev = torch.device("cuda", local)
    stream = torch.cuda.ExternalStream(ctx.stream, device=local)
    d_cols = torch.from_numpy(mine).to(dev)
    d_counts = torch.zeros(sp.counts_size, dtype=torch.float64, device=dev)
    d_ll = torch.zeros(max(1, hi - lo), dtype=torch.float64, device=dev)

    def step_resident():
        d_counts.zero_()
        torch.cuda.synchronize()
        sp.columns_device(d_cols, hi - lo, d_counts, d_ll)
        if world > 1:
            ctx.synchronize()
            dist.all_reduce(d_counts)

    for _ in range(args.warmup):
        step_resident()
    ctx.set_timing(True)
    ctx.fill_time_ms()
    barrier()
    steps = max(1, min(args.steps, 5))
    t0 = time.perf_counter()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(steps):
        step_resident()
    e1.record(stream)
    barrier()
    wall_ms = 1e3 * (time.perf_counter() - t0)
    ms = max(e0.elapsed_time(e1), 0.0)
    if world > 1:
        ms = wall_ms          # the all-reduce runs on torch's stream: bracket the whole step on the host clock between barriers
    kernel_ms, kernel_launches = ctx.fill_time_ms()
    ctx.set_timing(False)
    resident_counts = d_counts.cpu().numpy()

    pinned = torch.from_numpy(mine).pin_memory()
    host_cols = pinned.numpy()

    def step_e2e():
        buf = torch.zeros(sp.counts_size, dtype=torch.float64, device=dev)
        torch.cuda.synchronize()
        sp.columns(host_cols, counts=buf, col_loglike=False)
        if world > 1:
            dist.all_reduce(buf)
        return buf.cpu().numpy()

    step_e2e()
    barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        e2e_counts = step_e2e()
    barrier()
    e2e_ms = 1e3 * (time.perf_counter() - t0)
    assert np.allclose(e2e_counts, resident_counts, rtol=1e-9, atol=1e-9)

    t = torch.tensor([ms, e2e_ms, kernel_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, e2e_ms, kernel_ms = t.tolist()
    out = None
    if rank == 0:
        A, Cc = model.alph_size, model.components
        # E/F/G messages of the UNGAPPED nodes of every column (the kernels visit nothing else) + its characters
        ungapped = int(np.count_nonzero(~np.isin(mine, [ord("-"), ord(".")])))
        alg_total = 3.0 * Cc * A * 8 * ungapped + float(N) * (hi - lo)
        alg_bytes = alg_total / max(1, hi - lo)
        per_launch_s = kernel_ms / 1e3 / max(1, kernel_launches)
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        achieved = (hi - lo) * alg_bytes / per_launch_s / 1e9
        base = None
        if world == 1:
            sys.path.insert(0, os.path.join(ROOT, "tests"))
            import oracle_lib as O
            sample = cols[:: max(1, n_cols // args.cpu_cols)][: args.cpu_cols]
            c0 = time.perf_counter()
            oc = O.sumprod_columns(model, tree, sample, flags=O.SP_COUNTS)["counts"]
            cpu_s = time.perf_counter() - c0
            dc = sp.columns(sample)["counts"]
            assert np.allclose(dc, oc, rtol=1e-8, atol=1e-8), "device counts differ from the CPU oracle on the sample"
            base = {"value": len(sample) / cpu_s, "unit": "columns/s", "cores": 1, "kind": "port",
                    "sample": "%d evenly spaced columns of the same alignment through oracle/oracle_sumprod.cpp (FP64, 1 thread), %.1f s; "
                              "device counts checked against it" % (len(sample), cpu_s)}
        out = {"metric": "EM count columns/s", "value": n_cols * steps / (ms / 1e3), "unit": "columns/s", "steps": steps,
               "ms_per_step": ms / steps, "scaling": "strong", "dtype": "f64",
               "config": {"workload": "dna_jc_gamma2_4096x1000_seed1 full-history alignment, substitution E-step (sum-product + eigen-counts)",
                          "columns": int(n_cols), "nodes": int(N), "alphabet": int(A), "components": int(Cc),
                          "parallelism": "columns sharded over %d GPU(s), one all-reduce of %d doubles per step" % (world, sp.counts_size)},
               "e2e": {"value": n_cols * steps / (e2e_ms / 1e3), "unit": "columns/s", "h2d_bytes_per_step": int((hi - lo) * N),
                       "d2h_bytes_per_step": int(8 * sp.counts_size)},
               "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                            "kernel": "one hist_sumprod_columns call: hb::dn_up / dn_down / dn_counts / dn_root over the dense columns (node-major, one launch per tree level) "
                                      "+ hb::sumprod_columns_kernel over the sparse ones",
                            "avg_launch_ms": 1e3 * per_launch_s, "algorithmic_bytes_per_column": alg_bytes, "ungapped_nodes": ungapped,
                            "columns_per_launch": int(hi - lo),
                            "note": "algorithmic bytes = 3*C*A*8 per UNGAPPED node of each column + N characters per column (full-tree columns are "
                                    "996 of 12 997); the call is latency-bound on the level-by-level walk of the dense columns, not on HBM"},
               "cpu_baseline": base, "gpu_launches": int(2 * steps)}
    sp.close()
    return out



# ------------------------------------------------------------------ reconstruction leg (BASELINE.json metric 2, configs[2])
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "bin", "historian")
RECON_REF_LEAVES = 64


def _family_inputs(tree, leaves, guide):
    seqs = {v: bytes(row[row != ord("-")]).decode() for v, row in zip(leaves, guide)}
    grows = {v: row != ord("-") for v, row in zip(leaves, guide)}
    return seqs, grows


def _reference_recon(tree, leaves, guide, extra):
    """Wall seconds and FASTA rows of the unmodified reference binary (one core: it has no threads) on a family."""
    with tempfile.TemporaryDirectory() as tmp:
        tf, gf = os.path.join(tmp, "tree.nh"), os.path.join(tmp, "guide.fa")
        open(tf, "w").write(tree.newick())
        with open(gf, "w") as f:
            for v, row in zip(leaves, guide):
                f.write(">%s\n%s\n" % (tree.names[v], bytes(row).decode()))
        t0 = time.perf_counter()
        out = subprocess.run([REF_BIN, "recon", "-guide", gf, "-tree", tf, "-preset", "lg", "-output", "fasta", "-norefine", "-band", str(BAND)] + extra,
                             stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, check=True, cwd=tmp).stdout.decode()
        wall = time.perf_counter() - t0
    rows = []
    for line in out.splitlines():
        if line.startswith(">"):
            rows.append("")
        elif rows:
            rows[-1] += line.strip()
    return wall, rows


def recon_leg(ctx, args, rank, world, local, barrier):
    """recon wall-time: the whole progressive reconstruction (every level: DPMatrix constructors, Forward fills, traces,
    makeProfile, parent profiles kept on the device; recon.cpp:917-1052) of the 1024 x 1000 LG family, band 20,
    deterministic profiles (-profmaxstates 1).  N GPUs: ONE family, its sibling merges spread over the ranks
    (historian_b200.recon.plan_owners), profiles moved between ranks with NCCL send/recv: strong scaling.  At N = 1 also
    the sampled-profile mode (-profsamples 10 -profmaxstates 5000) and the reference binary, both on a 64-leaf family."""
    import torch
    import torch.distributed as dist
    from historian_b200 import recon as R
    dev = torch.device("cuda", local)
    lg, guide, tree, leaves = make_family(args.family_seed, with_tree=True)
    seqs, grows = _family_inputs(tree, leaves, guide)
    owner = R.plan_owners(tree, world)
    root_rank = int(owner[tree.root])
    walls, bits, lp, stage = [], None, None, None
    for rep in range(1 + args.recon_reps):      # the first run is the warm-up
        nat = R.NativeReconstructor(ctx, lg, band=BAND, max_profile_states=1)
        barrier()
        t0 = time.perf_counter()
        b, l, rec = R.reconstruct_partitioned(nat, tree, seqs, grows, rank, world, dist, dev)
        barrier()
        wall = time.perf_counter() - t0
        if rep:
            walls.append(wall)
        if rank == root_rank:
            bits, lp, stage = b, l, nat.stage_ms
        rec.close()
    t = torch.tensor([min(walls)], dtype=torch.float64, device=dev)
    flag = torch.tensor([1.0 if rank == root_rank else 0.0, float(bits.shape[1]) if bits is not None else 0.0, lp if lp is not None else 0.0],
                        dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(flag)      # only the root's rank contributes columns and likelihood
    out = None
    if rank == 0:
        levels = R.merge_levels(tree)
        moves = sum(len(R.level_moves(tree, lv, owner)) for lv in levels)
        out = {"metric": "recon wall-time", "value": t.item(), "unit": "s", "higher_is_better": False, "scaling": "strong", "n_gpus": world,
               "config": {"workload": "lg_1024x1000_seed%d whole-tree progressive reconstruction: %d merges on %d levels, band %d, -profmaxstates 1"
                                      % (args.family_seed, tree.n_nodes - len(leaves), len(levels), BAND),
                          "parallelism": "sibling merges of a level spread over %d GPU(s); %d profile(s) cross ranks (NCCL send/recv); "
                                         "levels with fewer merges than GPUs leave ranks idle (the last %d levels)" % (
                                             world, moves, sum(1 for lv in levels if len(lv) < world))},
               "reps": args.recon_reps, "alignment_columns": int(flag[1].item()), "lp_final": flag[2].item()}
        if world == 1:
            out["stage_ms"] = {k: round(v, 2) for k, v in stage.items()}
    if world == 1 and rank == 0:
        # ---- a 64-leaf family: sampled profiles, and the reference binary on the same inputs
        lg2, guide2, tree2, leaves2 = make_family(args.family_seed, n_leaves=RECON_REF_LEAVES, with_tree=True)
        seqs2, grows2 = _family_inputs(tree2, leaves2, guide2)
        small = {}
        for name, kw, extra in (("deterministic", dict(max_profile_states=1), ["-profmaxstates", "1"]),
                                ("sampled", dict(profile_samples=10, max_profile_states=5000, seed=1), ["-profsamples", "10", "-profmaxstates", "5000", "-seed", "1"])):
            nat = R.NativeReconstructor(ctx, lg2, band=BAND, **kw)
            nat.reconstruct(tree2, seqs2, grows2)[2].close()
            nat = R.NativeReconstructor(ctx, lg2, band=BAND, **kw)
            t0 = time.perf_counter()
            b2, l2, rec2 = nat.reconstruct(tree2, seqs2, grows2)
            w2 = time.perf_counter() - t0
            rec2.close()
            entry = {"ours_s": w2, "alignment_columns": int(b2.shape[1])}
            if os.path.exists(REF_BIN) and (name == "deterministic" or args.recon_ref_sampled):
                rw, rows = _reference_recon(tree2, leaves2, guide2, extra)
                ours = R.gapped_rows_dense(b2, tree2, seqs2, wildcard="x")
                same = len(rows) == tree2.n_nodes and all(ours[k] == r.lower() or ours[k] == r for k, r in enumerate(rows))
                entry.update({"reference_s": rw, "identical_alignment": bool(same)})
            small[name] = entry
        out["family_%d_leaves" % RECON_REF_LEAVES] = small
        det = small["deterministic"]
        if "reference_s" in det:
            out["cpu_baseline"] = {"value": det["reference_s"], "unit": "s", "cores": 1, "kind": "reference",
                                   "sample": "oracle/_ref/bin/historian recon on the %d-leaf x %d family of the same model and tree shape (deterministic "
                                             "profiles, band %d): %.2f s against %.3f s here, identical alignment: %s"
                                             % (RECON_REF_LEAVES, ROOT_LEN, BAND, det["reference_s"], det["ours_s"], det["identical_alignment"])}
    return out


# ------------------------------------------------------------------ codon leg (BASELINE.json configs[4])
CODON_LEAVES, CODON_ROOT_LEN = 256, 600
CODON_TENSOR_PIPE = 0.013   # profiles/r02_codon_sass_and_ncu.txt: tensor pipe active 1.3 % of leftmul_tc_kernel (24 UTCHMMA per CTA; the rest is the FP64 epilogue)


def codon_family(seed=1):
    from historian_b200 import model as M, simulator as S
    tree = S.Tree.balanced(CODON_LEAVES, BRANCH)
    rm = M.RateModel.preset("ECMrest")
    gapped, _ = S.simulate_family(rm, tree, CODON_ROOT_LEN, seed)
    leaves, guide = S.leaf_guide(tree, gapped)
    return rm, guide, tree, leaves


def codon_constructor_ms(reps=5):
    """Milliseconds of the batched DPMatrix constructors (hist_pair_prepare_batch) of the 128 leaf x leaf merges of the
    codon family; run in this process (tensor-core leftMultiply) and, by the caller, in a child with HIST_B200_NO_TENSOR=1."""
    from historian_b200 import capi, descs as D, model as M
    rm, guide, tree, leaves = codon_family()
    seqs, grows = _family_inputs(tree, leaves, guide)
    ctx = capi.Context(int(os.environ.get("LOCAL_RANK", "0")))
    hmm = M.pair_hmm(rm, BRANCH, BRANCH)
    lv = ctx.upload_sequences([seqs[v].encode("latin-1") for v in leaves], rm.alphabet, rm.components)
    envs = [D.Envelope.from_guide_rows(grows[a], grows[b], BAND) for a, b in zip(leaves[0::2], leaves[1::2])]
    plan = capi.PrepareBatch(ctx, lv[0::2], lv[1::2], hmm, envs, backward=False)
    best = 1e30
    for _ in range(reps + 1):
        ctx.synchronize()
        t0 = time.perf_counter()
        pr = plan.run()
        ctx.synchronize()
        best = min(best, 1e3 * (time.perf_counter() - t0))
        ctx.close_all(pr)
    ctx.close()
    return best


def codon_leg(ctx, args):
    """configs[4]: ECMrest (61 sense codons) family of 256 sequences x 600 codons: whole-tree reconstruction wall time, and the
    constructor level that holds the tensor-core GEMM against the CUDA-core kernel it replaces."""
    from historian_b200 import recon as R
    rm, guide, tree, leaves = codon_family()
    seqs, grows = _family_inputs(tree, leaves, guide)
    walls = []
    for rep in range(3):
        nat = R.NativeReconstructor(ctx, rm, band=BAND, max_profile_states=1)
        t0 = time.perf_counter()
        bits, lp, rec = nat.reconstruct(tree, seqs, grows)
        walls.append(time.perf_counter() - t0)
        stage = nat.stage_ms
        rec.close()
    tensor_ms = codon_constructor_ms()
    child = subprocess.run([sys.executable, os.path.abspath(__file__), "--codon-constructor-only"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL,
                           env=dict(os.environ, HIST_B200_NO_TENSOR="1"), text=True)
    cuda_ms = None
    try:
        cuda_ms = float(child.stdout.strip().splitlines()[-1])
    except Exception:
        pass
    return {"metric": "recon wall-time (codon)", "value": min(walls[1:]), "unit": "s", "higher_is_better": False,
            "config": {"workload": "ECMrest_256x600_seed1 whole-tree progressive reconstruction (61-state codon model, band %d, -profmaxstates 1)" % BAND,
                       "alphabet": 61, "components": 1},
            "alignment_columns": int(bits.shape[1]), "lp_final": lp, "stage_ms": {k: round(v, 2) for k, v in stage.items()},
            "constructors_128_leaf_merges_ms": {"tensor_core_leftmultiply": tensor_ms, "cuda_core_leftmultiply": cuda_ms,
                                                "kernel": "hb::leftmul_tc_kernel: tcgen05.mma kind::tf32 (3xTF32), M=128 N=64 K=64 per tile, FP32 accumulators in TMEM"},
            "tensor_pipe": CODON_TENSOR_PIPE}

# ------------------------------------------------------------------ guide-alignment leg
def guide_leg(ctx, args, lg, guide_rows):
    """The pairwise step of the guide-alignment stage on the bench family: QuickAlignMatrix (quickalign.cpp:7-214) for
    --guide-pairs random pairs of its 1024 leaves (AlignGraph takes ceil(n log2 n) such pairs, span.cpp:57-118), k-mer seeded
    envelopes (band 64, 6-mers, threshold 14), time 1.  End to end by construction: tokens and diagonals go up, scores, end
    cells and paths come back inside the timed call."""
    from historian_b200 import guide as G
    A = len(lg.alphabet)
    lut = {ch: k for k, ch in enumerate(lg.alphabet)}
    toks = [np.array([lut.get(chr(c), -1) for c in row[row != ord("-")]], np.int8) for row in np.asarray(guide_rows)]
    rng = np.random.default_rng(11)
    n = len(toks)
    picks = [(int(a), int(b)) for a, b in zip(rng.integers(0, n, args.guide_pairs), rng.integers(0, n, args.guide_pairs)) if a != b]
    sub, sc = G.quickalign_scores(lg, 1.0)
    # The reference's default envelope (kmerThreshold = -1, diagenv.cpp:109-116) is the FULL one whenever xLen * yLen * cellSize
    # fits in memory, as it does for 1000-residue sequences; --guide-envelope sparse takes the k-mer seeded bands instead.
    if args.guide_envelope == "sparse":
        pairs = [(toks[a], toks[b], G.sparse_envelope(toks[a], toks[b], A)) for a, b in picks]
    else:
        pairs = [(toks[a], toks[b], np.arange(1 - len(toks[b]), len(toks[a]), dtype=np.int32)) for a, b in picks]
    cells = sum(G.envelope_cells(dg, len(x), len(y)) for x, y, dg in pairs)
    h2d0 = ctx.h2d_bytes
    ctx.quickalign(pairs, sub, sc)
    h2d = ctx.h2d_bytes - h2d0
    launches0 = ctx.launch_count
    walls = []
    for _ in range(3):
        t0 = time.perf_counter()
        res, ends, paths = ctx.quickalign(pairs, sub, sc)
        walls.append(time.perf_counter() - t0)
    wall = min(walls)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as O
    sample = list(range(0, len(pairs), max(1, len(pairs) // args.guide_cpu_pairs)))[: args.guide_cpu_pairs]
    c0 = time.perf_counter()
    for k in sample:
        want, end, wpath = O.quickalign(*pairs[k], sub, sc)
        assert want == res[k] and tuple(ends[k]) == end and np.array_equal(paths[k], wpath), "device alignment differs from the CPU oracle"
    cpu_s = time.perf_counter() - c0
    cpu_cells = sum(G.envelope_cells(pairs[k][2], len(pairs[k][0]), len(pairs[k][1])) for k in sample)
    ref_base = None
    ref_bin = os.path.join(ROOT, "oracle", "_ref", "bin", "ref_dump_quickalign")
    if os.path.exists(ref_bin) and args.guide_envelope == "full":
        # the unmodified reference's QuickAlignMatrix constructor + alignPath() on all pairs of a few of the same sequences
        with tempfile.TemporaryDirectory() as tmp:
            fa = os.path.join(tmp, "s.fa")
            with open(fa, "w") as f:
                for k in range(args.guide_ref_seqs):
                    f.write(">s%d\n%s\n" % (k, "".join(lg.alphabet[t] if t >= 0 else "*" for t in toks[k])))
            out = subprocess.run([ref_bin, os.path.join(tmp, "d.bin"), "lg", fa, "1", "full"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
        try:
            f_ = out.split()
            rc_, rf_, rt_ = float(f_[f_.index("cells") + 1]), float(f_[f_.index("fill_s") + 1]), float(f_[f_.index("trace_s") + 1])
            ref_base = {"value": rc_ / (rf_ + rt_) / 1e9, "unit": "Gcells/s", "cores": 1, "kind": "reference",
                        "sample": "the %d pairs among %d of the family's sequences, unmodified reference QuickAlignMatrix constructor + alignPath() via "
                                  "oracle/_ref/bin/ref_dump_quickalign, %.1f s inside them" % (int(f_[f_.index("pairs") + 1]), args.guide_ref_seqs, rf_ + rt_)}
        except Exception:
            ref_base = None
    stage = guide_stage(ctx, args, lg, toks)
    return {"metric": "guide-alignment Viterbi cells/s", "value": cells / wall / 1e9, "unit": "Gcells/s", "higher_is_better": True, "ms": 1e3 * wall,
            "dtype": "f64 (bit-exact with the reference)", "pairs": len(pairs), "cells": int(cells),
            "diagonals_per_pair": [int(min(len(p[2]) for p in pairs)), int(max(len(p[2]) for p in pairs))],
            "config": {"workload": "lg_1024x1000_seed%d: %d random pairs of leaves, %s, time 1" % (
                args.family_seed, len(pairs), "full envelopes (the reference's default at this size)" if args.guide_envelope == "full"
                else "k-mer seeded envelopes (6-mers, >= 14 matches, band 64)")},
            "e2e": {"value": cells / wall / 1e9, "unit": "Gcells/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(sum(len(p) for p in paths) + 16 * len(pairs))},
            "gpu_launches": int((ctx.launch_count - launches0) // 3),
            "cpu_baseline": ref_base if ref_base is not None else
            {"value": cpu_cells / cpu_s / 1e9, "unit": "Gcells/s", "cores": 1, "kind": "port",
             "sample": "%d of the pairs through oracle/oracle_quickalign.cpp (dense cell array, 1 thread), %.1f s" % (len(sample), cpu_s)},
            "checked": "scores, end cells and paths of %d pairs equal the CPU port's (oracle/oracle_quickalign.cpp, %.3f Gcells/s on one core)" % (
                len(sample), cpu_cells / cpu_s / 1e9),
            "stage": stage}


def guide_stage(ctx, args, lg, toks):
    """The whole guide-alignment stage (AlignGraph, span.cpp:35-158) on the first --guide-stage-seqs leaves: random graph of
    ceil(n log2 n) pairs, full envelopes, Viterbi on the device, spanning tree and merge on the host (hist_guide_align); the
    unmodified reference's AlignGraph on the first --guide-stage-ref-seqs of them through oracle/_ref/bin/ref_dump_guide, its
    guide alignment compared with the device's for equality."""
    from historian_b200 import guide as G
    sub, sc = G.quickalign_scores(lg, 1.0)
    n = min(args.guide_stage_seqs, len(toks))
    ctx.guide_align(toks[:8], sub, sc, sparse=False)
    wall = float("inf")
    for _ in range(2):
        t0 = time.perf_counter()
        bits, info = ctx.guide_align(toks[:n], sub, sc, sparse=False)
        wall = min(wall, time.perf_counter() - t0)
    out = {"seqs": n, "pairs": int(len(info["edges"])), "columns": int(bits.shape[1]), "seconds": wall,
           "cells": int(sum(len(toks[a]) * len(toks[b]) for a, b in info["edges"]))}
    ref_bin = os.path.join(ROOT, "oracle", "_ref", "bin", "ref_dump_guide")
    m = min(args.guide_stage_ref_seqs, n)
    if os.path.exists(ref_bin) and m >= 2:
        with tempfile.TemporaryDirectory() as tmp:
            fa = os.path.join(tmp, "s.fa")
            with open(fa, "w") as f:
                for k in range(m):
                    f.write(">s%d\n%s\n" % (k, "".join(lg.alphabet[t] if t >= 0 else "*" for t in toks[k])))
            r = subprocess.run([ref_bin, "lg", fa, "1", "rnd", "5489", "full"], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
        try:
            tl = [ln.split() for ln in r.stderr.splitlines() if ln.startswith("timing")][0]
            rows = []
            for ln in r.stdout.splitlines():   # FASTA rows may be wrapped
                if ln.startswith(">"):
                    rows.append("")
                elif rows:
                    rows[-1] += ln.strip()
            mine, minfo = ctx.guide_align(toks[:m], sub, sc, sparse=False)
            same = mine.shape == (m, len(rows[0])) and all(np.array_equal(mine[k], np.frombuffer(rows[k].encode(), np.uint8) != ord("-")) for k in range(m))
            out["reference"] = {"seqs": m, "pairs": int(len(minfo["edges"])), "graph_seconds": float(tl[2]), "merge_seconds": float(tl[4]), "cores": 1,
                                "kind": "reference", "guide_alignment_equal": bool(same)}
        except Exception as e:   # noqa: BLE001
            out["reference"] = {"unavailable": repr(e)}
    return out


# ------------------------------------------------------------------ refinement leg
def refine_leg(ctx, args, lg):
    """Refiner::refine (refiner.cpp:108-194) on BASELINE config 2 (PF16593, 43 leaves, committed under tests/golden/pf16593):
    the deterministic reconstruction, then branch refinement moves until n - 1 in a row fail to improve the log-likelihood.
    Every move = conditional position weight matrices (hist_sumprod_excluded_posteriors), branch Viterbi (hist_branch_align),
    merge, log-likelihood (hist_sumprod_columns).  Reference: the binary's `recon -refine` minus `recon -norefine` wall time."""
    from historian_b200 import recon as R, refine as RF, simulator as S
    gdir = os.path.join(ROOT, "tests", "golden", "pf16593")
    gf, tf = os.path.join(gdir, "PF16593.guide.fa"), os.path.join(gdir, "PF16593.nhx")
    if not (os.path.exists(gf) and os.path.exists(tf)):
        return None
    tree = S.Tree.from_newick(open(tf).read())
    names, rows = [], []
    for ln in open(gf).read().splitlines():
        if ln.startswith(">"):
            names.append(ln[1:].split()[0]); rows.append("")
        elif names:
            rows[-1] += ln.strip()
    node_of = {nm: v for v, nm in enumerate(tree.names) if tree.is_leaf(v)}
    seqs = {node_of[nm]: r.replace("-", "").replace(".", "") for nm, r in zip(names, rows)}
    grows = {node_of[nm]: np.array([c not in "-." for c in r]) for nm, r in zip(names, rows)}
    nat = R.NativeReconstructor(ctx, lg, band=20, profile_samples=10, max_profile_states=1)
    bits, _, rec = nat.reconstruct(tree, seqs, grows)
    gapped = R.gapped_rows_dense(bits, tree, seqs, wildcard="x")
    start = np.array([np.frombuffer(gapped[v].encode(), np.uint8) for v in range(tree.n_nodes)])
    best = None
    launches0 = ctx.launch_count
    for _ in range(2):
        r = RF.Refiner(ctx, lg, tree)
        t0 = time.perf_counter()
        lp0 = r.log_likelihood(start)
        got, lp = r.refine(start)
        wall = time.perf_counter() - t0
        r.close()
        if best is None or wall < best[0]:
            best = (wall, r.moves, r.accepted, lp0, lp)
    out = {"metric": "branch refinement wall time", "value": best[0], "unit": "s", "higher_is_better": False, "moves": best[1], "accepted": best[2],
           "log_likelihood": [best[3], best[4]], "ms_per_move": 1e3 * best[0] / max(1, best[1]), "gpu_launches": int((ctx.launch_count - launches0) // 2),
           "config": {"workload": "PF16593 (43 leaves x 32-36 residues, 85 nodes), deterministic reconstruction, band 20"}}
    ref_bin = os.path.join(ROOT, "oracle", "_ref", "bin", "historian")
    if os.path.exists(ref_bin):
        t = {}
        for flag in ("-norefine", "-refine"):
            t0 = time.perf_counter()
            subprocess.run([ref_bin, "recon", "-guide", gf, "-tree", tf, "-output", "fasta", "-profmaxstates", "1", flag], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
            t[flag] = time.perf_counter() - t0
        out["cpu_baseline"] = {"value": t["-refine"] - t["-norefine"], "unit": "s", "cores": 1, "kind": "reference",
                               "sample": "oracle/_ref/bin/historian recon -refine minus -norefine on the same input, wall clock"}
    return out


# ------------------------------------------------------------------ distance-matrix leg
def distmat_leg(ctx, args, lg, guide_rows):
    """RateModel::distanceMatrix (model.cpp:506-655) of the bench family's guide alignment: n (n - 1) / 2 maximum-likelihood
    distances, each a golden-section search over matrix exponentials.  End to end: the alignment goes up, the matrix comes back."""
    from historian_b200 import descs as D, guide as G
    rows = np.ascontiguousarray(np.asarray(guide_rows)[: args.distmat_rows], np.uint8)
    n = len(rows)
    spm, rate = D.SumProdModel.from_rate_model(lg), G.expected_substitution_rate(lg)
    ctx.distance_matrix(spm, rate, rows[:8])
    walls = []
    for _ in range(2):
        t0 = time.perf_counter()
        dist = ctx.distance_matrix(spm, rate, rows)
        walls.append(time.perf_counter() - t0)
    wall = min(walls)
    pairs = n * (n - 1) // 2
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as O
    pick = np.linspace(0, n - 1, args.distmat_cpu_rows).astype(int)
    c0 = time.perf_counter()
    want = O.distance_matrix(spm, rate, rows[pick])
    cpu_s = time.perf_counter() - c0
    got = dist[np.ix_(pick, pick)]
    iu = np.triu_indices(len(pick), 1)
    rel = np.abs(got[iu] - want[iu]) / np.abs(want[iu])
    assert np.all(rel <= 0.0101) and np.mean(rel <= 1e-6) >= 0.9, "device distances differ from the CPU oracle"
    # the guide tree from that matrix (Tree::buildByUPGMA, the default of `historian recon`; tree.cpp:362-454)
    ctx.build_tree(dist[:16, :16])
    t0 = time.perf_counter()
    child, blen = ctx.build_tree(dist, 0)
    tree_s = time.perf_counter() - t0
    t0 = time.perf_counter()
    child_nj, blen_nj = ctx.build_tree(dist, 1)
    tree_nj_s = time.perf_counter() - t0
    sub = min(n, args.tree_cpu_rows)
    c0 = time.perf_counter()
    wc, wb = O.build_tree(dist[:sub, :sub], 0)
    tree_cpu_s = time.perf_counter() - c0
    gc, gb = ctx.build_tree(dist[:sub, :sub], 0)
    assert np.array_equal(gc, wc) and np.array_equal(gb, wb), "device UPGMA tree differs from the CPU oracle"
    tree = {"upgma_s": tree_s, "nj_s": tree_nj_s, "leaves": n,
            "cpu_baseline": {"value": tree_cpu_s, "unit": "s", "cores": 1, "kind": "port",
                             "sample": "UPGMA on the first %d rows through oracle/oracle_tree.cpp (O(n^3): the full matrix costs (%d/%d)^3 times that), %.2f s; "
                                       "children and branch lengths bit-identical to the device's" % (sub, n, sub, tree_cpu_s)}}
    ref_base = None
    ref_bin = os.path.join(ROOT, "oracle", "_ref", "bin", "ref_dump_distmat")
    if os.path.exists(ref_bin):
        # the unmodified reference's RateModel::distanceMatrix and tree builders on the first rows of the same alignment
        nref = min(n, args.distmat_ref_rows)
        with tempfile.TemporaryDirectory() as tmp:
            fa = os.path.join(tmp, "g.fa")
            with open(fa, "w") as f:
                for k in range(nref):
                    f.write(">s%d\n%s\n" % (k, bytes(rows[k]).decode()))
            out = subprocess.run([ref_bin, os.path.join(tmp, "d.bin"), "lg", fa, "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout
        try:
            f_ = out.split()
            ds_, us_ = float(f_[f_.index("dist_s") + 1]), float(f_[f_.index("upgma_s") + 1])
            ref_base = {"value": nref * (nref - 1) / 2 / ds_, "unit": "pairs/s", "cores": 1, "kind": "reference",
                        "sample": "the %d pairs among the first %d rows, unmodified reference RateModel::distanceMatrix via oracle/_ref/bin/ref_dump_distmat, "
                                  "%.1f s inside it (its UPGMA of those %d rows: %.4f s)" % (nref * (nref - 1) // 2, nref, ds_, nref, us_)}
        except Exception:
            ref_base = None
    return {"metric": "distance-matrix pairs/s", "value": pairs / wall, "guide_tree": tree, "unit": "pairs/s", "higher_is_better": True, "s": wall, "dtype": "f64",
            "config": {"workload": "lg_1024x1000_seed%d guide alignment: %d rows x %d columns, %d ML distances (100 golden-section iterations at most)" % (
                args.family_seed, n, rows.shape[1], pairs)},
            "e2e": {"value": pairs / wall, "unit": "pairs/s", "h2d_bytes_per_step": int(rows.size), "d2h_bytes_per_step": int(8 * n * n)},
            "gpu_launches": 1,
            "cpu_baseline": ref_base if ref_base is not None else
            {"value": len(iu[0]) / cpu_s, "unit": "pairs/s", "cores": 1, "kind": "port",
             "sample": "the %d pairs among %d evenly spaced rows through oracle/oracle_distmat.cpp (eigen-decomposed exponentials, 1 thread), %.1f s" % (len(iu[0]), len(pick), cpu_s)},
            "checked": "device distances of %d pairs equal the CPU port's (oracle/oracle_distmat.cpp, %.0f pairs/s on one core) to 1e-6 on >= 90 %% of them, "
                       "within the search's 1 %% on all" % (len(iu[0]), len(iu[0]) / cpu_s)}


# ------------------------------------------------------------------ GPU arm
def gpu_arm(args):
    import torch
    import torch.distributed as dist
    from historian_b200 import capi

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; historian_b200 has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    n_pairs = args.pairs
    # Weak scaling: every rank works on its own copy of THE family BASELINE's config names (seed 1), so the per-GPU work is
    # exactly fixed as N grows.  (--rank-families gives rank r the family of seed + r instead: families differ in work -- one
    # merge in 512 of seeds 3 and 8 has an anti-diagonal of 31-32 cells, two more slots than the chain kernel's warp holds,
    # and its two strips take the DAG path, 2.2 ms per step instead of 1.37)
    lg, guide = make_family(args.family_seed + (rank if args.rank_families else 0))
    hmm, inputs = cherry_inputs(lg, guide, n_pairs)
    ctx = capi.Context(local)
    stream = torch.cuda.ExternalStream(ctx.stream, device=local)

    # ---- resident inputs: profiles uploaded, DPMatrix tables prepared once
    profs = [(ctx.upload(x), ctx.upload(y)) for x, y, _ in inputs]
    pairs = ctx.prepare_batch([a for a, _ in profs], [b for _, b in profs], hmm, [e for _, _, e in inputs], backward=True)
    cells_per_step = 2 * sum(p.n_cells for p in pairs)

    def step_resident():
        return ctx.fill_batch(pairs, backward=True)

    handles = ctx.fill_batch_async(pairs, backward=True)   # the argument array of the batch, built once

    def step_enqueue():
        # the same C-ABI call with lp_end == NULL: fills enqueued on the context's stream, likelihoods read after the timed region
        ctx.fill_batch_async(pairs, backward=True, arr=handles)

    for _ in range(args.warmup):
        lp_end, lp_start = step_resident()
    assert np.all(np.isfinite(lp_end)) and np.allclose(lp_end, lp_start, rtol=1e-5, atol=1e-3), "Forward != Backward likelihood"
    ctx.set_timing(True)
    ctx.fill_time_ms()
    launches0 = ctx.launch_count
    clocks = ClockSampler(local)
    clocks.start()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        step_enqueue()
    e1.record(stream)
    barrier()
    ms = e0.elapsed_time(e1)
    launches = ctx.launch_count - launches0
    lp_chk, _ = step_resident()
    assert np.array_equal(lp_chk, lp_end), "likelihoods changed between steps"
    kernel_ms, kernel_launches = ctx.fill_time_ms()
    ctx.set_timing(False)

    # ---- end to end from host buffers through the public C-ABI calls
    seqs = cherry_sequences(guide, n_pairs)
    envs = [e for _, _, e in inputs]
    # the host buffer of the step: the residue strings back to back (as a FASTA reader holds them) + their offsets
    packed = np.frombuffer(b"".join(seqs), np.uint8).copy()
    packed_off = np.concatenate([[0], np.cumsum([len(q) for q in seqs])]).astype(np.int64)

    # the ctypes argument arrays of hist_pair_prepare_batch (pointers to the host descriptors) are built once;
    # every step passes them, with that step's fresh profile handles, to the C ABI
    plan = capi.PrepareBatch(ctx, [a for a, _ in profs], [b for _, b in profs], hmm, envs, backward=True)

    def step_e2e():
        # the calls a reference user makes per level: leaf Profile constructors (profile.cpp:23-76) from the residue
        # strings, DPMatrix constructors (forward.cpp:11-66), Forward + Backward fills, likelihoods read back
        h2d0 = ctx.h2d_bytes   # counted inside the library at every host -> device copy it issues
        leaves = ctx.upload_packed(packed, packed_off, lg.alphabet, lg.components)   # hist_profile_from_residues
        pr = plan.run_batch(leaves)                                                  # hist_pair_prepare_batch
        r = ctx.fill_batch(pr, backward=True)                                        # hist_fill_batch, likelihoods read back
        h2d = ctx.h2d_bytes - h2d0
        pr.close()                                                                   # hist_pairs_destroy
        leaves.close()                                                               # hist_profiles_destroy
        return r, h2d

    for _ in range(min(args.warmup, 2)):
        step_e2e()
    barrier()
    f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    f0.record(stream)
    e2e_steps = max(1, min(args.steps, 5))
    for _ in range(e2e_steps):
        (lp_e, lp_s), h2d_bytes = step_e2e()
    f1.record(stream)
    assert np.array_equal(lp_e, lp_end) and np.array_equal(lp_s, lp_start), "end-to-end likelihoods differ from the resident run"
    barrier()
    e2e_ms = f0.elapsed_time(f1)

    # The same steps with two in flight: a second context (its own stream and memory pool) on the same device, the host
    # prepares step k + 1 on one while the fills of step k run on the other; every step still uploads its inputs and reads
    # its likelihoods back inside the timed region.
    ctx_b = capi.Context(local)
    plan_b = capi.PrepareBatch(ctx_b, [a for a, _ in profs], [b for _, b in profs], hmm, envs, backward=True)
    lanes = [(ctx, plan), (ctx_b, plan_b)]

    def issue(k):
        c, pl = lanes[k % 2]
        leaves = c.upload_packed(packed, packed_off, lg.alphabet, lg.components)
        pr = pl.run_batch(leaves)
        c.fill_batch_async(pr, backward=True)
        return c, leaves, pr

    def finish(job):
        c, leaves, pr = job
        r = c.fill_results(pr, backward=True)
        pr.close()
        leaves.close()
        return r

    def run_pipelined(n):
        job, r = None, None
        for k in range(n):
            nxt = issue(k)
            if job is not None:
                r = finish(job)
            job = nxt
        return finish(job) if job is not None else r

    run_pipelined(4)
    barrier()
    pipe_steps = max(2, 2 * e2e_steps)
    h0 = ctx.h2d_bytes + ctx_b.h2d_bytes
    torch.cuda.synchronize()
    g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    g0.record(stream)
    lp_pe, lp_ps = run_pipelined(pipe_steps)
    torch.cuda.synchronize()
    g1.record(stream)
    torch.cuda.synchronize()
    assert np.array_equal(lp_pe, lp_end) and np.array_equal(lp_ps, lp_start), "pipelined end-to-end likelihoods differ from the resident run"
    pipe_ms = g0.elapsed_time(g1)
    pipe_h2d = (ctx.h2d_bytes + ctx_b.h2d_bytes - h0) // pipe_steps
    plan_b = None
    ctx_b.close()
    barrier()
    # the sampler has been running since the resident region started; a short run (tens of milliseconds) may have
    # seen a single nvidia-smi reading, so keep the same load on until it has three (untimed, at most 2 s)
    t_load = time.perf_counter()
    while len(clocks.rows) < 3 and time.perf_counter() - t_load < 2.0:
        step_resident()
    clk = clocks.summary()

    t = torch.tensor([ms, e2e_ms, kernel_ms, pipe_ms], dtype=torch.float64, device="cuda")
    c = torch.tensor([float(cells_per_step)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(c, op=dist.ReduceOp.SUM)
    ms, e2e_ms, kernel_ms, pipe_ms = t.tolist()
    total_cells_per_step = c.item()

    if rank == 0:
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        per_launch_s = kernel_ms / 1e3 / max(1, kernel_launches)
        achieved = cells_per_step * ALG_BYTES_PER_CELL / per_launch_s / 1e9
        cores = os.cpu_count() or 1
        base = cpu_baseline(guide, args.cpu_pairs, 1) if world == 1 else None
        line = {
            "metric": METRIC, "value": total_cells_per_step * args.steps / (ms / 1e3) / 1e9, "unit": UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32 mantissas + int32 exponent per cell (f64 tables and likelihoods)", "data": "synthetic",
            "config": workload_config(n_pairs),
            "e2e": {"value": max(total_cells_per_step * e2e_steps / (e2e_ms / 1e3), total_cells_per_step * pipe_steps / (pipe_ms / 1e3)) / 1e9, "unit": UNIT,
                    "mode": "two_steps_in_flight" if pipe_ms / pipe_steps < e2e_ms / e2e_steps else "one_step_at_a_time",
                    "h2d_bytes_per_step": int(h2d_bytes), "d2h_bytes_per_step": int(n_pairs * 24), "steps": pipe_steps if pipe_ms / pipe_steps < e2e_ms / e2e_steps else e2e_steps,
                    "host_memory": "one buffer of residue strings and the guide-row descriptors in pageable host memory, staged through the context's pinned buffers; profile structure, band keys and the anti-diagonal layout are built on the device",
                    "one_step_at_a_time": {"value": total_cells_per_step * e2e_steps / (e2e_ms / 1e3) / 1e9, "unit": UNIT, "steps": e2e_steps, "ms_per_step": e2e_ms / e2e_steps},
                    "two_steps_in_flight": {"value": total_cells_per_step * pipe_steps / (pipe_ms / 1e3) / 1e9, "unit": UNIT, "steps": pipe_steps, "ms_per_step": pipe_ms / pipe_steps,
                                            "h2d_bytes_per_step": int(pipe_h2d),
                                            "how": "two contexts on the device: the host prepares step k + 1 on one while the fills of step k run on the other (hist_fill_batch with lp_end == NULL, hist_fill_batch_results); same uploads and read-backs per step"}},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": FILL_TRAFFIC_BYTES if n_pairs == 512 and args.family_seed == 1 else None,
                         "kernel": "hb::fill_chain2_kernel<5> (one launch = the Forward and Backward fills of all merges of the step; absorb dot products inside)",
                         "launches_timed": int(kernel_launches),
                         "fp32": fp32_roofline(cells_per_step, per_launch_s),
                         "avg_launch_ms": 1e3 * per_launch_s, "algorithmic_bytes_per_cell": ALG_BYTES_PER_CELL,
                         "cells_per_launch": int(cells_per_step),
                         "peak_source": "MEASURED_PEAKS.json hbm_gbs (measured)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"},
            "cpu_baseline": base, "clocks": clk, "cpu_cores": cores,
        }
    em_line = None
    if not args.no_em:
        em_line = em_leg(ctx, args, rank, world, local, barrier)
    recon_line = None
    if not args.no_recon:
        recon_line = recon_leg(ctx, args, rank, world, local, barrier)
    codon_line = None
    if not args.no_codon and world == 1:
        codon_line = codon_leg(ctx, args)
    guide_line = None
    if not args.no_guide and world == 1:
        guide_line = guide_leg(ctx, args, lg, guide)
    distmat_line = None
    if not args.no_guide and world == 1:
        distmat_line = distmat_leg(ctx, args, lg, guide)
    refine_line = None
    if not args.no_guide and world == 1:
        refine_line = refine_leg(ctx, args, lg)
    if rank == 0:
        if refine_line is not None:
            line["refine"] = refine_line
        if guide_line is not None:
            line["guide_align"] = guide_line
        if distmat_line is not None:
            line["distance_matrix"] = distmat_line
        if em_line is not None:
            line["em_counts"] = em_line
        if recon_line is not None:
            line["recon"] = recon_line
        if codon_line is not None:
            line["codon"] = codon_line
        print(json.dumps(line))
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    # ONE JSON line on stdout: anything libraries print to fd 1 (e.g. NCCL's version banner) goes to stderr instead
    sys.stdout.flush()
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    sys.stdout = os.fdopen(real_stdout, "w", buffering=1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--pairs", type=int, default=512, help="sibling merges per step per GPU")
    ap.add_argument("--cpu-pairs", type=int, default=96, help="merges in the cpu_baseline sample (about 15 s on one core)")
    ap.add_argument("--cpu-cols", type=int, default=64, help="columns in the EM leg's cpu_baseline sample")
    ap.add_argument("--family-seed", type=int, default=1, help="simulation seed of the family (every rank fills its own copy)")
    ap.add_argument("--rank-families", action="store_true", help="rank r takes the family of seed + r instead of a copy of the same one")
    ap.add_argument("--no-em", action="store_true", help="skip the EM count leg (profiling runs)")
    ap.add_argument("--no-recon", action="store_true", help="skip the whole-tree reconstruction leg")
    ap.add_argument("--no-codon", action="store_true", help="skip the codon (ECMrest) leg")
    ap.add_argument("--no-guide", action="store_true", help="skip the guide-alignment (QuickAlign) and distance-matrix legs")
    ap.add_argument("--guide-stage-seqs", type=int, default=1024, help="sequences of the whole-stage guide alignment (hist_guide_align)")
    ap.add_argument("--guide-stage-ref-seqs", type=int, default=16, help="sequences the reference's AlignGraph is timed on")
    ap.add_argument("--guide-pairs", type=int, default=1024, help="sequence pairs in the guide-alignment leg")
    ap.add_argument("--guide-envelope", default="full", choices=["full", "sparse"])
    ap.add_argument("--distmat-rows", type=int, default=1024, help="rows of the guide alignment in the distance-matrix leg")
    ap.add_argument("--tree-cpu-rows", type=int, default=384, help="rows of the UPGMA cpu_baseline sample")
    ap.add_argument("--distmat-cpu-rows", type=int, default=16, help="rows in that leg's cpu_baseline sample")
    ap.add_argument("--guide-ref-seqs", type=int, default=24, help="sequences whose all-pairs alignments time the reference in the guide-alignment leg")
    ap.add_argument("--distmat-ref-rows", type=int, default=96, help="rows of the reference's distance-matrix timing sample")
    ap.add_argument("--guide-cpu-pairs", type=int, default=24, help="pairs in that leg's cpu_baseline sample")
    ap.add_argument("--codon-constructor-only", action="store_true", help=argparse.SUPPRESS)
    ap.add_argument("--recon-reps", type=int, default=3, help="timed repetitions of the whole-tree reconstruction (best is reported)")
    ap.add_argument("--recon-ref-sampled", action="store_true", help="also time the reference binary in sampled mode on the 64-leaf family (minutes)")
    args = ap.parse_args()
    if args.codon_constructor_only:
        print(codon_constructor_ms())
        return
    if args.impl == "reference":
        reference_arm(args)
    else:
        gpu_arm(args)


if __name__ == "__main__":
    main()
