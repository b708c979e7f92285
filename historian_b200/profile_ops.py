"""Host-side Profile bookkeeping on the flattened (T1) layout.

Mirrors the parts of the reference's profile.cpp that are pure bookkeeping and
stay on the host: addReadyStates (profile.cpp:268-319) and the consistency
checks (profile.cpp:321-354)."""
import numpy as np

from .descs import Profile


def out_lists(p):
    """(in, null_out, absorb_out) transition-id lists per state, ascending ids."""
    n = p.n_states
    ins = [[] for _ in range(n)]
    nulls = [[] for _ in range(n)]
    absorbs = [[] for _ in range(n)]
    for t in range(p.n_trans):
        s, d = int(p.trans_src[t]), int(p.trans_dest[t])
        ins[d].append(t)
        (absorbs if p.is_emit[d] else nulls)[s].append(t)
    return ins, nulls, absorbs


def add_ready_states(p, extra=None):
    """Profile::addReadyStates (profile.cpp:268-319): every state with both null and
    absorbing out-transitions is split into a wait state (keeps `in` and nullOut, gains a
    0-weight null transition) immediately followed by a ready state (takes absorbOut).

    `extra`: optional dict of per-state arrays to carry along (ready state copies the
    value of its wait state).  Returns (new_profile, old2new, extra_new) where old2new maps
    the old state index to the index of its (wait) state."""
    n = p.n_states
    _, nulls, absorbs = out_lists(p)
    mixed = [bool(nulls[s]) and bool(absorbs[s]) for s in range(n)]
    old2new = np.zeros(n, np.int64)
    ready_of = {}
    k = 0
    for s in range(n):
        old2new[s] = k
        k += 1
        if mixed[s]:
            ready_of[s] = k
            k += 1
    n_new = k
    is_emit = np.zeros(n_new, np.uint8)
    lp_absorb = np.full((n_new,) + p.lp_absorb.shape[1:], -np.inf)
    leaf = None if p.closest_leaf_pos is None else np.zeros(n_new, np.int32)
    src = p.trans_src.astype(np.int64).copy()
    dest = p.trans_dest.astype(np.int64).copy()
    lp = list(p.trans_lp)
    new_src, new_dest = [], []
    for s in range(n):
        is_emit[old2new[s]] = p.is_emit[s]
        lp_absorb[old2new[s]] = p.lp_absorb[s]
        if leaf is not None:
            leaf[old2new[s]] = p.closest_leaf_pos[s]
        if mixed[s]:
            r = ready_of[s]
            is_emit[r] = 0  # the ready state is a null state (ProfileState() has empty lpAbsorb)
            if leaf is not None:
                leaf[r] = p.closest_leaf_pos[s]
    for t in range(p.n_trans):
        s, d = int(src[t]), int(dest[t])
        # absorbing out-transitions of a mixed state move to its ready state (profile.cpp:289-291)
        ns = ready_of[s] if (mixed[s] and p.is_emit[d]) else old2new[s]
        new_src.append(ns)
        new_dest.append(old2new[d])
    for s in range(n):  # ready transitions are appended in state order (profile.cpp:283, 299)
        if mixed[s]:
            new_src.append(old2new[s])
            new_dest.append(ready_of[s])
            lp.append(0.0)
    q = Profile(is_emit, np.array(new_src, np.int32), np.array(new_dest, np.int32), np.array(lp), lp_absorb, leaf, name=p.name)
    extra_new = None
    if extra is not None:
        extra_new = {}
        for key, arr in extra.items():
            arr = np.asarray(arr)
            out = np.zeros((n_new,) + arr.shape[1:], arr.dtype)
            for s in range(n):
                out[old2new[s]] = arr[s]
                if mixed[s]:
                    out[ready_of[s]] = arr[s]
            extra_new[key] = out
    return q, old2new, extra_new


def check_profile(p):
    """assertTransitionsConsistent / assertAllStatesWaitOrReady / assertPathToEndExists
    (profile.cpp:262-354).  Raises ValueError on violation."""
    n = p.n_states
    if p.is_emit[0] or p.is_emit[n - 1]:
        raise ValueError("START/END must be null")
    if np.any(p.trans_src >= p.trans_dest):
        raise ValueError("transition violates toposort")
    _, nulls, absorbs = out_lists(p)
    for s in range(n):
        if nulls[s] and absorbs[s]:
            raise ValueError("state %d is neither wait nor ready" % s)
    reach = np.zeros(n, bool)
    reach[0] = True
    order = np.argsort(p.trans_src, kind="stable")
    for t in order:
        if reach[p.trans_src[t]]:
            reach[p.trans_dest[t]] = True
    if not reach[n - 1]:
        raise ValueError("no path from start to end")
    return True
