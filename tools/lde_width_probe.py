#!/usr/bin/env python3
"""LDE (extendPol, blowup 2) time per column against the width of the column strip (development aid).
The strip pipeline and the multi-GPU piece plan choose their widths from this table."""
import sys, torch
sys.path.insert(0, ".")
import zkevm_prover_b200 as zk
from zkevm_prover_b200.dist_commit import GpuOps

log_n = int(sys.argv[1]) if len(sys.argv) > 1 else 23
n = 1 << log_n
ops = GpuOps()
zk.set_stream(torch.cuda.current_stream().cuda_stream)
wmax = 136
d_in = torch.empty(n * wmax, dtype=torch.int64, device="cuda")
zk.fill(d_in, n * wmax, 1, 0)
d_out = torch.empty(2 * n * wmax, dtype=torch.int64, device="cuda")
for w in (8, 16, 24, 32, 40, 48, 56, 64, 72, 80, 88, 96, 112, 128, 136):
    ops.lde(d_out, d_in, 2 * n, n, w)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3):
        ops.lde(d_out, d_in, 2 * n, n, w)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 3
    print("w=%3d  %.2f ms  %.3f ms/col  (pitch %d B)" % (w, ms, ms / w, w * 8), flush=True)
