"""ORACLE (test infrastructure, not product code): Refiner::BranchMatrix and its traceback restated in plain Python.

Follows reference src/refiner.cpp:10-106 (fill, lpEnd, best()) over TreeAlignFuncs::BranchMatrixBase
(src/sampler.cpp:452-476 preMultiply / calcInsProbs, 1005-1031, 1162-1184; sampler.h:146-149 inEnvelope, 203-205 logMatch)
with exact log-sum-exp (the reference compiled with its own -DLOG_SUM_EXP_SLOW).  Pure Python loops: small cases only.
Pinned through tests/test_refine_gpu.py: the refinement built on it must reproduce `historian recon -refine`."""
import math

NEG = -math.inf


def lse(a, b):
    if a == NEG:
        return b
    if b == NEG:
        return a
    m = max(a, b)
    return m + math.log1p(math.exp(-abs(a - b)))


def branch_align(log_sub, log_ins, log_cw, lpt, x_pwm, y_pwm, max_distance=-1, x_key=None, y_key=None):
    """Returns (lpEnd, path codes: 3 both, 1 parent only, 2 child only)."""
    C, A = len(log_ins), len(log_ins[0])
    xS, yS = len(x_pwm) + 1, len(y_pwm) + 1
    y_sub = [[[NEG] * A for _ in range(C)] for _ in y_pwm]
    y_emit = []
    for n, lpp in enumerate(y_pwm):
        for c in range(C):
            for i in range(A):
                for j in range(A):
                    y_sub[n][c][i] = lse(y_sub[n][c][i], log_sub[c][i][j] + lpp[c][j])
        lp = NEG
        for c in range(C):
            for i in range(A):
                lp = lse(lp, log_cw[c] + log_ins[c][i] + lpp[c][i])
        y_emit.append(lp)

    def log_match(x, y):
        lip = NEG
        for c in range(C):
            v = NEG
            for i in range(A):
                v = lse(v, x_pwm[x - 1][c][i] + y_sub[y - 1][c][i])
            lip = lse(lip, v)
        return lip

    def in_env(x, y):
        return x == 0 or y == 0 or x == xS - 1 or y == yS - 1 or max_distance < 0 or abs(int(x_key[x]) - int(y_key[y])) <= max_distance

    cell = {}

    def get(x, y):
        return cell.get((x, y), (NEG, NEG, NEG))
    M, I, Dl = 0, 1, 2
    for x in range(xS):
        for y in range(yS):
            if not in_env(x, y):
                continue
            d = [0.0 if (x == 0 and y == 0) else NEG, NEG, NEG]
            if x > 0 and in_env(x - 1, y):
                s = get(x - 1, y)
                d[Dl] = max(max(s[M] + lpt[M][Dl], s[I] + lpt[I][Dl]), s[Dl] + lpt[Dl][Dl])
            if y > 0 and in_env(x, y - 1):
                s = get(x, y - 1)
                d[I] = y_emit[y - 1] + max(s[M] + lpt[M][I], s[I] + lpt[I][I])
            if x > 0 and y > 0 and in_env(x - 1, y - 1):
                s = get(x - 1, y - 1)
                d[M] = log_match(x, y) + max(max(s[M] + lpt[M][M], s[I] + lpt[I][M]), s[Dl] + lpt[Dl][M])
            cell[(x, y)] = tuple(d)
    e = get(xS - 1, yS - 1)
    lp_end = max(max(e[M] + lpt[M][3], e[I] + lpt[I][3]), e[Dl] + lpt[Dl][3])
    x, y, state, path = xS - 1, yS - 1, 3, []
    while x > 0 or y > 0:
        bx = (state == M and x > 0 and y > 0) or state == Dl
        by = (state == M and x > 0 and y > 0) or state == I
        if bx or by:
            path.append((1 if bx else 0) | (2 if by else 0))
        sx, sy = x - bx, y - by
        em = 0.0
        if state == M:
            em = log_match(x, y) if x > 0 and y > 0 else NEG
        elif state == I:
            em = y_emit[y - 1] if y > 0 else NEG
        s = get(sx, sy)
        best, bs = NEG, None
        for q in range(3):
            v = s[q] + lpt[q][state] + em
            if v > best:
                best, bs = v, q
        assert bs is not None, "could not find traceback state"
        x, y, state = sx, sy, bs
    return lp_end, path[::-1]
