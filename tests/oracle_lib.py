"""ctypes binding of oracle/libhist_oracle.so — the CPU checker used by the tests.
TEST INFRASTRUCTURE: nothing in historian_b200/ imports this."""
import ctypes as C
import os
import subprocess

import numpy as np

from historian_b200 import descs as D

_HERE = os.path.dirname(os.path.abspath(__file__))
ORACLE_DIR = os.path.join(os.path.dirname(_HERE), "oracle")
_lib = None


def lib():
    global _lib
    if _lib is None:
        so = os.path.join(ORACLE_DIR, "libhist_oracle.so")
        if not os.path.exists(so):
            subprocess.check_call(["make", "-s", "-C", ORACLE_DIR, "restatement"])
        _lib = C.CDLL(so)
        _lib.oracle_log_sum_exp.restype = C.c_double
        _lib.oracle_log_sum_exp.argtypes = [C.c_double, C.c_double]
        _lib.oracle_sum_path_absorb.restype = C.c_double
    return _lib


def set_exact_lse(exact):
    lib().oracle_set_exact_lse(int(bool(exact)))


def _p(a, t):
    return None if a is None else a.ctypes.data_as(C.POINTER(t))


def pair_tables(x, y, hmm, env):
    """DPMatrix constructor tables and envelope flags (forward.cpp:11-66)."""
    nx, ny, Cc, A = x.n_states, y.n_states, x.components, x.alph_size
    r = dict(subx=np.empty((nx, Cc, A)), suby=np.empty((ny, Cc, A)), insx=np.empty(nx), insy=np.empty(ny),
             rootsubx=np.empty(nx), rootsuby=np.empty(ny), x_key=np.zeros(nx, np.int32), y_key=np.zeros(ny, np.int32),
             x_near_start=np.zeros(nx, np.uint8), y_near_end=np.zeros(ny, np.uint8),
             in_envelope=np.zeros((nx - 1, ny - 1), np.uint8))
    n = C.c_int64()
    xd, yd, hd, ed = x.desc(), y.desc(), hmm.desc(), env.desc()
    rc = lib().oracle_pair_tables(C.byref(xd), C.byref(yd), C.byref(hd), C.byref(ed),
                                  _p(r["subx"], C.c_double), _p(r["suby"], C.c_double), _p(r["insx"], C.c_double),
                                  _p(r["insy"], C.c_double), _p(r["rootsubx"], C.c_double), _p(r["rootsuby"], C.c_double),
                                  _p(r["x_key"], C.c_int32), _p(r["y_key"], C.c_int32), _p(r["x_near_start"], C.c_uint8),
                                  _p(r["y_near_end"], C.c_uint8), _p(r["in_envelope"], C.c_uint8), C.byref(n))
    assert rc == 0
    r["n_cells"] = n.value
    return r


def forward(x, y, hmm, env):
    cells = np.empty((x.n_states - 1, y.n_states - 1, 5))
    lp = C.c_double()
    xd, yd, hd, ed = x.desc(), y.desc(), hmm.desc(), env.desc()
    rc = lib().oracle_forward(C.byref(xd), C.byref(yd), C.byref(hd), C.byref(ed), _p(cells, C.c_double), C.byref(lp))
    assert rc == 0
    return cells, lp.value


def backward(x, y, hmm, env):
    cells = np.empty((x.n_states - 1, y.n_states - 1, 5))
    lp = C.c_double()
    xd, yd, hd, ed = x.desc(), y.desc(), hmm.desc(), env.desc()
    rc = lib().oracle_backward(C.byref(xd), C.byref(yd), C.byref(hd), C.byref(ed), _p(cells, C.c_double), C.byref(lp))
    assert rc == 0
    return cells, lp.value


def _cells_to_np(buf, n):
    return np.array([[buf[k].xpos, buf[k].ypos, buf[k].state] for k in range(n)], np.int32).reshape(n, 3)


def trace_best(x, y, hmm, env, fwd_cells, end):
    cap = x.n_states + y.n_states + 8
    buf = (D.Cell * cap)()
    n = C.c_int64()
    xd, yd, hd, ed = x.desc(), y.desc(), hmm.desc(), env.desc()
    rc = lib().oracle_trace_best(C.byref(xd), C.byref(yd), C.byref(hd), C.byref(ed), _p(fwd_cells, C.c_double),
                                 D.Cell(*end), buf, C.c_int64(cap), C.byref(n))
    assert rc == 0, rc
    return _cells_to_np(buf, n.value)


def trace_best_forward(x, y, hmm, env, bwd_cells, start):
    cap = x.n_states + y.n_states + 8
    buf = (D.Cell * cap)()
    n = C.c_int64()
    xd, yd, hd, ed = x.desc(), y.desc(), hmm.desc(), env.desc()
    rc = lib().oracle_trace_best_forward(C.byref(xd), C.byref(yd), C.byref(hd), C.byref(ed), _p(bwd_cells, C.c_double),
                                         D.Cell(*start), buf, C.c_int64(cap), C.byref(n))
    assert rc == 0, rc
    return _cells_to_np(buf, n.value)


def trace_sample(x, y, hmm, env, fwd_cells, mt_state, n_traces):
    cap = (x.n_states + y.n_states + 8) * n_traces
    buf = (D.Cell * cap)()
    off = np.zeros(n_traces + 1, np.int64)
    st = np.ascontiguousarray(mt_state, np.uint32).copy()
    xd, yd, hd, ed = x.desc(), y.desc(), hmm.desc(), env.desc()
    rc = lib().oracle_trace_sample(C.byref(xd), C.byref(yd), C.byref(hd), C.byref(ed), _p(fwd_cells, C.c_double),
                                   _p(st, C.c_uint32), C.c_int32(n_traces), buf, C.c_int64(cap), _p(off, C.c_int64))
    assert rc == 0, rc
    allc = _cells_to_np(buf, int(off[-1]))
    return [allc[off[k]:off[k + 1]] for k in range(n_traces)], st


def posterior_cells(x, y, hmm, env, fwd_cells, bwd_cells, lp_end, min_post):
    cap = fwd_cells.size
    buf = (D.Cell * cap)()
    lpp = np.empty(cap)
    n = C.c_int64()
    xd, yd, hd, ed = x.desc(), y.desc(), hmm.desc(), env.desc()
    rc = lib().oracle_posterior_cells(C.byref(xd), C.byref(yd), C.byref(hd), C.byref(ed), _p(fwd_cells, C.c_double),
                                      _p(bwd_cells, C.c_double), C.c_double(lp_end), C.c_double(min_post), buf,
                                      _p(lpp, C.c_double), C.c_int64(cap), C.byref(n))
    assert rc == 0, rc
    return _cells_to_np(buf, n.value), lpp[:n.value].copy()


class Elim:
    pass


def make_elim_buffers(Cc, A, state_cap, trans_cap, via_cap):
    e = Elim()
    e.state_cell = (D.Cell * state_cap)()
    e.state_lp_absorb = np.empty((state_cap, Cc, A))
    e.state_fwd_lp = np.empty(state_cap)
    e.trans_src = np.empty(trans_cap, np.int32)
    e.trans_dest = np.empty(trans_cap, np.int32)
    e.trans_lp = np.empty(trans_cap)
    e.trans_best_lp = np.empty(trans_cap)
    e.via_off = np.empty(trans_cap + 1, np.int64)
    e.via = (D.Cell * via_cap)()
    r = D.ElimResult()
    r.state_cell = e.state_cell
    r.state_lp_absorb = _p(e.state_lp_absorb, C.c_double)
    r.state_fwd_lp = _p(e.state_fwd_lp, C.c_double)
    r.trans_src, r.trans_dest = _p(e.trans_src, C.c_int32), _p(e.trans_dest, C.c_int32)
    r.trans_lp, r.trans_best_lp = _p(e.trans_lp, C.c_double), _p(e.trans_best_lp, C.c_double)
    r.trans_best_via_off = _p(e.via_off, C.c_int64)
    r.trans_best_via = e.via
    r.state_capacity, r.trans_capacity, r.via_capacity = state_cap, trans_cap, via_cap
    e.r = r
    return e


def elim_to_dict(e):
    r = e.r
    ns, nt = r.n_states, r.n_trans
    return dict(state_cell=_cells_to_np(e.state_cell, ns), state_lp_absorb=e.state_lp_absorb[:ns].copy(),
                state_fwd_lp=e.state_fwd_lp[:ns].copy(), trans_src=e.trans_src[:nt].copy(),
                trans_dest=e.trans_dest[:nt].copy(), trans_lp=e.trans_lp[:nt].copy(),
                trans_best_lp=e.trans_best_lp[:nt].copy(), via_off=e.via_off[:nt + 1].copy(),
                via=_cells_to_np(e.via, int(r.n_via)))


def profile_eliminate(x, y, hmm, env, fwd_cells, lp_end, cells, strategy):
    cells = np.ascontiguousarray(cells, np.int32).reshape(-1, 3)
    n = len(cells)
    e = make_elim_buffers(x.components, x.alph_size, n + 4, 8 * n + 64, 8 * n * 8 + 64)
    xd, yd, hd, ed = x.desc(), y.desc(), hmm.desc(), env.desc()
    rc = lib().oracle_profile_eliminate(C.byref(xd), C.byref(yd), C.byref(hd), C.byref(ed), _p(fwd_cells, C.c_double),
                                        C.c_double(lp_end), cells.ctypes.data_as(C.POINTER(D.Cell)), C.c_int64(n),
                                        C.c_int32(strategy), C.byref(e.r))
    assert rc == 0, rc
    return elim_to_dict(e)


def sum_path_absorb(p, log_cpt_weight, log_ins_prob):
    cum = np.empty(p.n_states)
    w = np.ascontiguousarray(log_cpt_weight, np.float64)
    ins = np.ascontiguousarray(log_ins_prob, np.float64)
    pd = p.desc()
    tot = lib().oracle_sum_path_absorb(C.byref(pd), _p(w, C.c_double), _p(ins, C.c_double), _p(cum, C.c_double))
    return tot, cum


# ------------------------------------------------------------------ sum-product (oracle/oracle_sumprod.cpp)
SP_COUNTS, SP_MAP_STATES, SP_NODE_POST = 1, 2, 4


def sumprod_branch_tables(model, tree):
    Cc, A, N = model.components, model.alph_size, tree.n_nodes
    sub = np.zeros((Cc, N, A, A))
    esc = np.zeros((Cc, N, A, A, 2))
    md, td = model.desc(), tree.desc()
    rc = lib().oracle_sumprod_branch_tables(C.byref(md), C.byref(td), _p(sub, C.c_double), _p(esc, C.c_double))
    assert rc == 0
    return sub, esc[..., 0] + 1j * esc[..., 1]


def sumprod_columns(model, tree, columns, weights=None, flags=SP_COUNTS | SP_MAP_STATES | SP_NODE_POST, debug=False):
    """columns: uint8 [n_cols, N].  Returns dict(col_loglike, counts, map_state, node_post[, E, F, G, logE, logF, logG])."""
    columns = np.ascontiguousarray(columns, np.uint8)
    L, N = columns.shape
    Cc, A = model.components, model.alph_size
    r = dict(col_loglike=np.zeros(L), counts=np.zeros(D.counts_size(Cc, A)), map_state=np.zeros((L, N), np.uint8),
             node_post=np.zeros((L, N, A)))
    dbg = [None] * 6
    if debug:
        r.update(E=np.zeros((L, Cc, N, A)), F=np.zeros((L, Cc, N, A)), G=np.zeros((L, Cc, N, A)), logE=np.zeros((L, Cc, N)),
                 logF=np.zeros((L, Cc, N)), logG=np.zeros((L, Cc, N)))
        dbg = [r[k] for k in ("E", "F", "G", "logE", "logF", "logG")]
    w = None if weights is None else np.ascontiguousarray(weights, np.float64)
    md, td = model.desc(), tree.desc()
    rc = lib().oracle_sumprod_columns(C.byref(md), C.byref(td), columns.ctypes.data_as(C.c_char_p), _p(w, C.c_double), C.c_int64(L),
                                      C.c_int32(flags), _p(r["col_loglike"], C.c_double), _p(r["counts"], C.c_double),
                                      r["map_state"].ctypes.data_as(C.c_char_p), _p(r["node_post"], C.c_double),
                                      *[_p(a, C.c_double) for a in dbg])
    assert rc == 0, rc
    return r


def sumprod_transform_counts(model, counts):
    Cc, A = model.components, model.alph_size
    root, sub = np.zeros((Cc, A)), np.zeros((Cc, A, A))
    md = model.desc()
    counts = np.ascontiguousarray(counts, np.float64)
    rc = lib().oracle_sumprod_transform_counts(C.byref(md), _p(counts, C.c_double), _p(root, C.c_double), _p(sub, C.c_double))
    assert rc == 0
    return root, sub


def split_counts(counts, Cc, A):
    """flat EigenCounts -> (indel[7], rootCount[C][A], eigenCount[C][A][A] complex)."""
    counts = np.asarray(counts)
    root = counts[7:7 + Cc * A].reshape(Cc, A)
    ec = counts[7 + Cc * A:].reshape(Cc, A, A, 2)
    return counts[:7], root, ec[..., 0] + 1j * ec[..., 1]


# ------------------------------------------------------------------ guide-alignment stage (oracle/oracle_quickalign.cpp)
def quickalign(x_tok, y_tok, diagonals, submat, scores):
    """QuickAlignMatrix over the envelope `diagonals`: (result, (xEnd, yEnd), path uint8[] or None)."""
    xt, yt = np.ascontiguousarray(x_tok, np.int8), np.ascontiguousarray(y_tok, np.int8)
    dg = np.ascontiguousarray(diagonals, np.int32)
    sub = np.ascontiguousarray(submat, np.float64)
    sc = D.QuickAlignScores(int(sub.shape[0]), _p(sub, C.c_double), *[float(v) for v in scores])
    pr = D.QuickAlignPair(len(xt), len(yt), _p(xt, C.c_int8), _p(yt, C.c_int8), len(dg), _p(dg, C.c_int32))
    res, xe, ye, n = C.c_double(), C.c_int32(), C.c_int32(), C.c_int32()
    path = np.zeros(len(xt) + len(yt), np.uint8)
    rc = lib().oracle_quickalign(C.byref(pr), C.byref(sc), C.byref(res), C.byref(xe), C.byref(ye), _p(path, C.c_uint8), C.byref(n))
    assert rc == 0, "oracle traceback inconsistent"
    return res.value, (xe.value, ye.value), (path[:n.value].copy() if np.isfinite(res.value) else None)


# ------------------------------------------------------------------ distance-matrix stage (oracle/oracle_distmat.cpp)
def distance_matrix(sp_model, expected_sub_rate, rows, max_iterations=100):
    rows = np.ascontiguousarray(rows, np.uint8)
    n, cols = rows.shape
    dist = np.zeros((n, n))
    md = sp_model.desc()
    rc = lib().oracle_distance_matrix(C.byref(md), C.c_double(float(expected_sub_rate)), n, cols, rows.ctypes.data_as(C.c_char_p),
                                      int(max_iterations), _p(dist, C.c_double))
    assert rc == 0
    return dist


def build_tree(dist, method=0, min_branch_length=1e-9):
    """oracle/oracle_tree.cpp: (child [2n-1][2], branch_length [2n-1])."""
    dist = np.ascontiguousarray(dist, np.float64)
    n = len(dist)
    child, blen = np.zeros((2 * n - 1, 2), np.int32), np.zeros(2 * n - 1)
    rc = lib().oracle_build_tree(n, _p(dist, C.c_double), int(method), C.c_double(float(min_branch_length)), _p(child, C.c_int32), _p(blen, C.c_double))
    assert rc == 0
    return child, blen


def fb_indel_counts(x, y, hmm, env, x_trans_indel, y_trans_indel, waits):
    """The indel half of BackwardMatrix::getCounts (forward.cpp:1183-1214) with counts on the children's transitions
    ([n_trans][6] or None).  Returns [ins, del, insExt, delExt, insTime, delTime, lp]."""
    xd, yd, hd, ed = x.desc(), y.desc(), hmm.desc(), env.desc()
    xt = None if x_trans_indel is None else np.ascontiguousarray(x_trans_indel, np.float64)
    yt = None if y_trans_indel is None else np.ascontiguousarray(y_trans_indel, np.float64)
    w = np.ascontiguousarray(waits, np.float64)
    out = np.zeros(7)
    rc = lib().oracle_fb_indel_counts(C.byref(xd), C.byref(yd), C.byref(hd), C.byref(ed), _p(xt, C.c_double), _p(yt, C.c_double),
                                      _p(w, C.c_double), _p(out, C.c_double))
    assert rc == 0
    return out


def eliminate_indel_counts(x, y, hmm, env, cells, strategy, x_trans_indel, y_trans_indel, waits):
    """Indel counts of makeProfile's effective transitions (forward.cpp:753-790), [n_trans][6] in oracle_profile_eliminate's order."""
    xd, yd, hd, ed = x.desc(), y.desc(), hmm.desc(), env.desc()
    cells = np.ascontiguousarray(cells, np.int32).reshape(-1, 3)
    xt = None if x_trans_indel is None else np.ascontiguousarray(x_trans_indel, np.float64)
    yt = None if y_trans_indel is None else np.ascontiguousarray(y_trans_indel, np.float64)
    w = np.ascontiguousarray(waits, np.float64)
    cap = 4 * len(cells) + 16
    out, n = np.zeros((cap, 6)), C.c_int64()
    rc = lib().oracle_eliminate_indel_counts(C.byref(xd), C.byref(yd), C.byref(hd), C.byref(ed), cells.ctypes.data_as(C.c_void_p), C.c_int64(len(cells)),
                                             C.c_int32(int(strategy)), _p(xt, C.c_double), _p(yt, C.c_double), _p(w, C.c_double), _p(out, C.c_double),
                                             C.c_int64(cap), C.byref(n))
    assert rc == 0, rc
    return out[:n.value]
