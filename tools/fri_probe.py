#!/usr/bin/env python3
"""Throughput of the FRI layer commit on device-resident data (development aid): folding step + getTransposed +
merkelize of the next layer, for the step sizes a 2^24-point FRI polynomial goes through (4 bits per step)."""
import sys, numpy as np, torch
sys.path.insert(0, ".")
import zkevm_prover_b200 as zk

zk.set_stream(torch.cuda.current_stream().cuda_stream)
x = np.array([3, 5, 7], dtype=np.uint64)
for pb, nb in ((24, 20), (20, 16), (16, 12), (12, 8)):
    a = torch.empty((1 << pb) * 3, dtype=torch.int64, device="cuda")
    zk.fill(a, a.numel(), 7, 0)
    o = torch.empty((1 << nb) * 3, dtype=torch.int64, device="cuda")
    t = torch.empty_like(o)
    tb = max(nb - 4, 0)
    n_groups, group = 1 << tb, (1 << nb) >> tb
    tree = torch.empty((2 * n_groups - 1) * 4, dtype=torch.int64, device="cuda")
    def step():
        zk.FRIProve.fold(o, a, pb, nb, 12345, x)
        zk.FRIProve.getTransposed(t, o, 1 << nb, tb)
        zk._check(zk.lib().zkgpu_merkletree(tree.data_ptr(), t.data_ptr(), group * 3, n_groups, 1))
    step(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        step()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    print("fold 2^%d -> 2^%d + transpose + tree(%d x %d): %.3f ms  (%.1f M ext elements/s in)" % (pb, nb, n_groups, group * 3, ms, (1 << pb) / ms / 1e3), flush=True)
