#!/usr/bin/env python3
"""Development probe: time of the row-sponge kernel against strip width / row stride / start column (GPU box)."""
import sys, pathlib
sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent))
import torch
import zkevm_prover_b200 as zk

zk.init(0)
zk.set_stream(torch.cuda.current_stream())
L = zk.lib()
lg = int(sys.argv[1]) if len(sys.argv) > 1 else 22
rows, C = 1 << lg, 665
d = torch.empty(rows * C, dtype=torch.int64, device="cuda")
zk.fill(d, rows * C)
leaves = torch.empty(rows * 4, dtype=torch.int64, device="cuda")


def t(fn, reps=3):
    fn(); torch.cuda.synchronize()
    best = 1e9
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best


for w, c0, ld in [(8, 0, C), (16, 0, C), (40, 0, C), (41, 0, C), (41, 624, C), (48, 0, C), (104, 0, C), (208, 0, C), (208, 208, C), (416, 0, C),
                  (665, 0, C), (40, 0, 40), (41, 0, 41), (208, 0, 208), (664, 0, 664)]:
    nperm = -(-w // 8)
    ms = t(lambda: zk._check(L.zkgpu_merkle_leaves(leaves.data_ptr(), d.data_ptr() + 8 * c0, w, ld, rows if ld == C else rows * C // ld // 2)))
    r = rows if ld == C else rows * C // ld // 2
    print("w=%4d c0=%4d ld=%4d rows=%9d: %8.2f ms  %.3f ms per permutation sweep of 2^22 rows  (%.0f Mperm/s)" % (w, c0, ld, r, ms, ms / nperm * (1 << 22) / r, r * nperm / ms / 1e3))
