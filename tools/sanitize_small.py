#!/usr/bin/env python3
"""Small end-to-end pass over every entry point, meant to run under `compute-sanitizer --tool memcheck` (GPU box)."""
import sys, pathlib
ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
import numpy as np
import torch
import zkevm_prover_b200 as zk
from oracle_lib import Oracle, SEED

O = Oracle()
zk.init(0)
ok = True


def check(name, got, want):
    global ok
    same = np.array_equal(np.asarray(got).reshape(-1), np.asarray(want).reshape(-1))
    ok &= same
    print(("ok   " if same else "FAIL ") + name, flush=True)


for n, c, b in [(1 << 8, 21, 2), (1 << 9, 665, 2), (1 << 10, 12, 16), (1 << 6, 3, 2), (1 << 11, 40, 2)]:
    a = O.fill(n * c, SEED, n + c).reshape(n, c)
    ext = np.zeros((n * b, c), dtype=np.uint64)
    tree = np.zeros((2 * n * b - 1) * 4, dtype=np.uint64)
    zk.extend_and_merkelize(zk.NTT_Goldilocks(n), ext, tree, a, n * b, n, c)
    want = O.extend_pol(a, n * b)
    check("commit host %dx%d b%d ext" % (n, c, b), ext, want)
    check("commit host %dx%d b%d tree" % (n, c, b), tree, O.merkletree(want, c, n * b))
    d_in = torch.from_numpy(a.view(np.int64).reshape(-1).copy()).cuda()
    d_ext = torch.empty(n * b * c, dtype=torch.int64, device="cuda")
    d_tree = torch.empty((2 * n * b - 1) * 4, dtype=torch.int64, device="cuda")
    zk.extend_and_merkelize(zk.NTT_Goldilocks(n), d_ext, d_tree, d_in, n * b, n, c)
    zk.sync()
    check("commit dev  %dx%d b%d ext" % (n, c, b), d_ext.cpu().numpy().view(np.uint64), want)
    mt = zk.MerkleTreeGL(n * b, c, d_ext, d_tree)
    check("group proof row", mt.getGroupProof(3)[:c], want[3])
for lg in (0, 1, 4, 7, 9, 13, 17):
    n, c = 1 << lg, 5 if lg < 17 else 17
    a = O.fill(n * c, SEED, lg).reshape(n, c)
    o = np.zeros_like(a)
    zk.NTT_Goldilocks(n).NTT(o, a, n, c)
    check("ntt 2^%d x %d" % (lg, c), o, O.ntt(a))
    zk.NTT_Goldilocks(n).INTT(o, a, n, c, None, 3, 1, True)
    check("intt-extend 2^%d x %d" % (lg, c), o, O.intt(a, extend=True))
a = O.fill(64 * 50, SEED, 9).reshape(64, 50)
t = np.zeros((2 * 64 - 1) * 4, dtype=np.uint64)
zk.PoseidonGoldilocks.merkletree_batch(t, a, 50, 64, 16)
check("merkletree_batch", t, O.merkletree(a, 50, 64, 1, 16))
src = O.fill(64 * 7, SEED, 11).reshape(64, 7)
dst = np.zeros((64, 3), dtype=np.uint64)
zk.NTT_Goldilocks(64).reversePermutation(dst, src, 64, 2, 3, 7)
check("reversePermutation", dst, src[[int(format(i, "06b")[::-1], 2) for i in range(64)], 2:5])
img = np.zeros(2 + 5 * 256 + (2 * 256 - 1) * 4, dtype=np.uint64)
cp = O.fill(64 * 5, SEED, 12).reshape(64, 5)
zk.build_const_tree(img, cp, 6, 8, 5)
check("const tree", img[2:2 + 5 * 256], O.extend_pol(cp, 256))
print("ALL OK" if ok else "FAILURES")
sys.exit(0 if ok else 1)
