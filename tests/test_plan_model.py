"""The pure-Python model of the multi-pass NTT plan (tools/ntt_plan_model.py: same pass sizes, pre-twiddle exponents,
in-tile R1 x R2 split and last-pass digit reversal as csrc/ntt.cu) against a direct DFT and the oracle."""
import sys, pathlib
import numpy as np
sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent / "tools"))
import ntt_plan_model as M
from oracle_lib import Oracle


def test_model_selftest():
    assert M.selftest()


def test_model_matches_oracle_with_scales():
    O = Oracle()
    L, C = 6, 2
    N = 1 << L
    a = O.fill(N * C).reshape(N, C)
    x = [[int(v) for v in row] for row in a]
    got = M.transform(x, L, C, [2, 3, 1])
    assert (np.array(got, dtype=np.uint64) == O.ntt(a)).all()
    ninv = pow(N, M.P - 2, M.P)
    post = [pow(7, i, M.P) * ninv % M.P for i in range(N)]
    got = M.transform(x, L, C, [3, 3], inverse=True, post=post)
    assert (np.array(got, dtype=np.uint64) == O.intt(a, extend=True)).all()
