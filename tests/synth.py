"""Seeded synthetic inputs for parity tests: random state-DAG profiles in wait/ready
form, random PairHMMs, random guide envelopes.  TEST INFRASTRUCTURE."""
import numpy as np

from historian_b200 import descs as D
from historian_b200.profile_ops import add_ready_states, check_profile


def random_profile(rng, n_core, A, C=1, p_null=0.12, p_skip=0.25, one_hot=False, jitter_leaf=False):
    """A toposorted DAG profile: a backbone chain plus skip transitions, some null states,
    converted to wait/ready form by addReadyStates."""
    n = n_core + 2
    is_emit = (rng.random(n) >= p_null).astype(np.uint8)
    is_emit[0] = is_emit[-1] = 0
    src, dest = [], []
    for i in range(n - 1):
        src.append(i); dest.append(i + 1)
        for g in (2, 3, 4):
            if i + g < n and rng.random() < p_skip / g:
                src.append(i); dest.append(i + g)
    order = np.lexsort((dest, src))
    src, dest = np.array(src)[order], np.array(dest)[order]
    lp = np.log(rng.uniform(0.05, 1.0, len(src)))
    lp_absorb = np.full((n, C, A), -np.inf)
    for i in range(n):
        if is_emit[i]:
            if one_hot and rng.random() < 0.7:
                lp_absorb[i, :, rng.integers(A)] = 0.0
            else:
                lp_absorb[i] = np.log(rng.uniform(1e-4, 1.0, (C, A))) - rng.uniform(0, 30)
    leaf = np.concatenate([[0], np.cumsum(is_emit[1:])]).astype(np.int32)
    if jitter_leaf:  # non-monotone closest-leaf coordinates, as in internal profiles (SURVEY hard part 3)
        for i in range(2, n - 1):
            if rng.random() < 0.15:
                leaf[i] = max(0, leaf[i] - rng.integers(0, 3))
    leaf[-1] = leaf.max()
    p = D.Profile(is_emit, src, dest, lp, lp_absorb, leaf)
    q, _, _ = add_ready_states(p)
    check_profile(q)
    return q


def random_sub_matrix(rng, A, t):
    """Row-stochastic P(t) = expm(R t) of a random reversible-ish rate matrix."""
    from scipy.linalg import expm
    pi = rng.dirichlet(np.ones(A) * 5)
    S = rng.uniform(0.1, 2.0, (A, A)); S = (S + S.T) / 2
    R = S * pi[None, :]
    np.fill_diagonal(R, 0)
    np.fill_diagonal(R, -R.sum(1))
    R /= -(pi * np.diag(R)).sum()
    return expm(R * t), pi


def random_hmm(rng, A, C=1, tl=None, tr=None):
    tl = rng.uniform(0.02, 1.0) if tl is None else tl
    tr = rng.uniform(0.02, 1.0) if tr is None else tr
    cpt = rng.dirichlet(np.ones(C) * 3)
    lsub, rsub, root = [], [], []
    for c in range(C):
        mult = rng.uniform(0.3, 2.0)
        seed = int(rng.integers(1 << 30))  # same rate matrix on both branches of a component
        Pl, pi = random_sub_matrix(np.random.default_rng(seed), A, tl * mult)
        Pr, _ = random_sub_matrix(np.random.default_rng(seed), A, tr * mult)
        lsub.append(Pl); rsub.append(Pr); root.append(pi)
    root = np.array(root)
    ir, dr = rng.uniform(0.005, 0.1), rng.uniform(0.005, 0.1)
    def side(t, sub):
        return dict(ins=1 - np.exp(-ir * t), **{"del": 1 - np.exp(-dr * t)}, insExt=rng.uniform(0.3, 0.8), delExt=rng.uniform(0.3, 0.8),
                    sub=np.array(sub), insProb=root, t=t, insWait=t / 2, delWait=t / 2)
    l, r = side(tl, lsub), side(tr, rsub)
    r["insExt"], r["delExt"] = l["insExt"], l["delExt"]
    return D.PairHmm.from_probs(l, r, root, cpt)


def random_envelope(rng, x, y, max_distance):
    """A guide alignment of the two closest leaves (lengths = final closest-leaf positions)."""
    lx, ly = int(x.closest_leaf_pos[-1]), int(y.closest_leaf_pos[-1])
    row1, row2 = [], []
    a = b = 0
    while a < lx or b < ly:
        u = rng.random()
        if a < lx and b < ly and u < 0.8:
            row1.append(1); row2.append(1); a += 1; b += 1
        elif a < lx and (u < 0.9 or b >= ly):
            row1.append(1); row2.append(0); a += 1
        else:
            row1.append(0); row2.append(1); b += 1
    return D.Envelope.from_guide_rows(row1, row2, max_distance)


def leaf_pair(rng, alphabet, lx, ly, identity=0.7):
    """Two related random sequences as leaf profiles, with their true pairwise alignment as guide."""
    A = len(alphabet)
    anc = rng.integers(A, size=max(lx, ly) + 10)
    def mutate(n):
        s, rowmask = [], []
        k = 0
        while len(s) < n:
            u = rng.random()
            if u < 0.04:                # insertion
                s.append(int(rng.integers(A))); rowmask.append(-1)
            elif u < 0.08 and k < len(anc) - 1:
                k += 1                  # deletion
            else:
                s.append(int(anc[k % len(anc)]) if rng.random() < identity else int(rng.integers(A))); rowmask.append(k); k += 1
        return s, rowmask
    sx, mx = mutate(lx)
    sy, my = mutate(ly)
    # merge by ancestral coordinate -> guide rows
    row1, row2 = [], []
    i = j = 0
    while i < lx or j < ly:
        ax = mx[i] if i < lx else 1 << 30
        ay = my[j] if j < ly else 1 << 30
        if i < lx and mx[i] == -1:
            row1.append(1); row2.append(0); i += 1
        elif j < ly and my[j] == -1:
            row1.append(0); row2.append(1); j += 1
        elif ax == ay:
            row1.append(1); row2.append(1); i += 1; j += 1
        elif ax < ay:
            row1.append(1); row2.append(0); i += 1
        else:
            row1.append(0); row2.append(1); j += 1
    X = D.Profile.from_sequence("".join(alphabet[t] for t in sx), alphabet)
    Y = D.Profile.from_sequence("".join(alphabet[t] for t in sy), alphabet)
    return X, Y, (np.array(row1, bool), np.array(row2, bool))
