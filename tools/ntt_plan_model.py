#!/usr/bin/env python3
"""Pure-Python model of the multi-pass index/twiddle plan in zkevm_prover_b200/csrc/ntt.cu (ntt_pass_kernel):
same pass decomposition, same pre-twiddle exponent, same R1 x R2 split inside a pass, same digit reversal on the
last pass.  Checked against a direct O(N^2) DFT.  Design aid / regression check for the host-side plan logic;
runs on the CPU in milliseconds (tests/test_plan_model.py)."""
import random

P = 0xFFFFFFFF00000001
W32 = 7277203076849721926


def root(logn):
    return pow(W32, 1 << (32 - logn), P)


def dft_small(a, w):
    n = len(a)
    return [sum(a[j] * pow(w, j * k, P) for j in range(n)) % P for k in range(n)]


def split(l):
    return (l + 1) // 2, l // 2  # LR1, LR2 as dispatch_pass does


def transform(x, L, C, lgs, inverse=False, pre=None, post=None):
    """x: list of N rows of C words. lgs: pass sizes. pre/post: per-row multipliers (lists) or None."""
    N = 1 << L
    wN = root(L)
    if inverse:
        wN = pow(wN, P - 2, P)
    cur = [row[:] for row in x]
    consumed = 0
    for p, l in enumerate(lgs):
        first, last = p == 0, p == len(lgs) - 1
        lgKp = consumed
        lgM = L - consumed - l
        lgNprev = lgs[p - 1] if p else 0
        lgKprev = consumed - lgNprev if p else 0
        Np, M = 1 << l, 1 << lgM
        dst = [[None] * C for _ in range(N)] if last else cur
        LR1, LR2 = split(l)
        R1, R2 = 1 << LR1, 1 << LR2
        wR = pow(wN, N >> l, P) if l else 1
        totalQ = (N >> l) * C
        F = C << lgM
        for Q in range(totalQ):
            o, q = divmod(Q, F)
            rp, c = divmod(q, C)
            tile_row0 = (o << (l + lgM)) + rp
            kprev = o & ((1 << lgNprev) - 1)
            # step 1: thread g=j2 holds a[j1], n_p = j1*R2 + j2
            B = [[None] * R2 for _ in range(R1)]
            for g in range(R2):
                a = []
                for j1 in range(R1):
                    np_ = j1 * R2 + g
                    row = tile_row0 + (np_ << lgM)
                    v = cur[row][c]
                    if p:
                        rprev = (np_ << lgM) + rp
                        xx = (rprev * kprev) << lgKprev
                        assert xx < N
                        v = v * pow(wN, xx, P) % P
                    elif pre is not None:
                        v = v * pre[row] % P
                    a.append(v)
                A = dft_small(a, pow(wR, R2, P))
                for k1 in range(R1):
                    B[k1][g] = A[k1] * pow(wR, g * k1, P) % P
            rev = 0
            if last:
                oo, sh = o, lgKp
                for qd in range(p - 1, -1, -1):
                    sh -= lgs[qd]
                    rev |= (oo & ((1 << lgs[qd]) - 1)) << sh
                    oo >>= lgs[qd]
            for g in range(R1):
                b = dft_small(B[g], pow(wR, R1, P))
                for k2 in range(R2):
                    kp = g + R1 * k2
                    v = b[k2]
                    if last:
                        drow = rev + (kp << lgKp)
                        if post is not None:
                            v = v * post[drow] % P
                    else:
                        drow = (o << (l + lgM)) + (kp << lgM) + rp
                    dst[drow][c] = v
        cur = dst
        consumed += l
    return cur


def plan(L):
    np_ = 1 if L == 0 else (L + 7) // 8
    return [L // np_ + (1 if p < L % np_ else 0) for p in range(np_)]


def direct(x, L, C, inverse=False):
    N = 1 << L
    w = root(L)
    if inverse:
        w = pow(w, P - 2, P)
    out = [[0] * C for _ in range(N)]
    for c in range(C):
        col = [x[i][c] for i in range(N)]
        r = dft_small(col, w)
        for i in range(N):
            out[i][c] = r[i]
    return out


def selftest():
    rnd = random.Random(1)
    for L, lgs in [(0, [0]), (1, [1]), (3, [3]), (5, [2, 2, 1]), (5, [1, 3, 1]), (6, [3, 3]), (7, [2, 1, 2, 2]), (6, [1, 2, 3]),
                   (4, [4]), (7, [4, 3]), (9, plan(9)), (8, [8])]:
        for C in (1, 3):
            N = 1 << L
            x = [[rnd.randrange(P) for _ in range(C)] for _ in range(N)]
            for inv in (False, True):
                assert transform(x, L, C, lgs, inv) == direct(x, L, C, inv), (L, lgs, C, inv)
    return True


if __name__ == "__main__":
    print("ok" if selftest() else "FAIL")
