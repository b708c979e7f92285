#!/usr/bin/env python3
"""Check the CPU restatement (oracle/libhist_oracle.so) against a dump made by the
real reference (oracle/_ref/bin/ref_dump): bit-level comparison of constructor
tables, envelope, Forward/Backward cells, traces, posterior cells and the parent
profile of makeProfile.  Usage: check_oracle_vs_dump.py dump.bin [node ...]"""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
from historian_b200 import descs as D
import oracle_lib as O


def same(a, b):
    a, b = np.asarray(a), np.asarray(b)
    if a.shape != b.shape:
        return False
    return bool(np.all((a == b) | (np.isnan(a) & np.isnan(b))))


def maxdiff(a, b):
    a, b = np.asarray(a, float), np.asarray(b, float)
    fin = np.isfinite(a) & np.isfinite(b)
    if not np.array_equal(np.isfinite(a), np.isfinite(b)):
        return np.inf
    return float(np.max(np.abs(a[fin] - b[fin]))) if fin.any() else 0.0


def check_node(d, node, verbose=True):
    pfx = "n%d." % node
    x, y = D.profile_from_dump(d, pfx + "x."), D.profile_from_dump(d, pfx + "y.")
    hmm, env = D.hmm_from_dump(d, pfx + "hmm."), D.env_from_dump(d, pfx + "env.")
    res = {}
    t = O.pair_tables(x, y, hmm, env)
    for k in ("subx", "suby", "insx", "insy", "rootsubx", "rootsuby"):
        res[k] = maxdiff(t[k], d[pfx + k])
    for k in ("x_near_start", "y_near_end", "in_envelope"):
        res[k] = 0.0 if same(t[k], d[pfx + k]) else np.inf
    fwd, lp_end = O.forward(x, y, hmm, env)
    res["fwd"] = maxdiff(fwd, d[pfx + "fwd"])
    res["lp_end"] = abs(lp_end - d[pfx + "lp_end"][0])
    bwd, lp_start = O.backward(x, y, hmm, env)
    res["bwd"] = maxdiff(bwd, d[pfx + "bwd"])
    res["lp_start"] = abs(lp_start - d[pfx + "lp_start"][0])
    ref_fwd = np.ascontiguousarray(d[pfx + "fwd"])
    ref_bwd = np.ascontiguousarray(d[pfx + "bwd"])
    bt = O.trace_best(x, y, hmm, env, ref_fwd, (x.n_states - 1, y.n_states - 1, D.EEE))
    res["best_trace"] = 0.0 if same(bt, d[pfx + "best_trace"]) else np.inf
    traces, st = O.trace_sample(x, y, hmm, env, ref_fwd, d[pfx + "rng_before"], 3)
    ok = all(same(traces[k], d[pfx + "sample_trace%d" % k]) for k in range(3)) and same(st, d[pfx + "rng_after"])
    res["sample_traces"] = 0.0 if ok else np.inf
    pc, pl = O.posterior_cells(x, y, hmm, env, ref_fwd, ref_bwd, float(d[pfx + "lp_end"][0]), float(d[pfx + "post_threshold"][0]))
    # the reference's priority_queue order among exact ties is heap-dependent: compare as sets + sorted lp
    a = sorted(map(tuple, pc.tolist())); b = sorted(map(tuple, d[pfx + "post_cells"].tolist()))
    res["post_cells"] = 0.0 if (a == b and maxdiff(pl, d[pfx + "post_lp"]) == 0) else np.inf
    return res


if __name__ == "__main__":
    d = D.read_dump(sys.argv[1])
    nodes = [int(a) for a in sys.argv[2:]] or [int(n) for n in d["dumped_nodes"]]
    worst = {}
    for n in nodes:
        r = check_node(d, n)
        for k, v in r.items():
            worst[k] = max(worst.get(k, 0.0), v)
        bad = {k: v for k, v in r.items() if v != 0}
        print("node %d: %s" % (n, "bit-exact" if not bad else bad))
    print("worst:", worst)
