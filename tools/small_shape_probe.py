#!/usr/bin/env python3
"""Small-shape latency of the fused commit call (BASELINE.json configs #1 and #4): device time per call, host enqueue
time per call, and the stage split (development aid).
Round-1 note: a cooperative form of the tree levels (16 lanes per node, S-box layer 4 multiplies deep instead of 48) was
tried for the levels that cannot fill the GPU: 23 us per level against 27 us with one thread per node (dependent
IMAD.WIDE chains cost ~10 clocks per instruction whatever the lane count), so it was not kept."""
import sys, time, torch
sys.path.insert(0, ".")
import zkevm_prover_b200 as zk

zk.set_stream(torch.cuda.current_stream().cuda_stream)
for (log_n, c, log_b) in ((16, 8, 1), (17, 12, 4)):
    n, n_ext = 1 << log_n, 1 << (log_n + log_b)
    d_in = torch.empty(n * c, dtype=torch.int64, device="cuda")
    zk.fill(d_in, n * c, 1, 0)
    d_ext = torch.empty(n_ext * c, dtype=torch.int64, device="cuda")
    d_tree = torch.empty((2 * n_ext - 1) * 4, dtype=torch.int64, device="cuda")
    ntt = zk.NTT_Goldilocks(n)
    for rep in (0, 1):
        for _ in range(5):
            zk.extend_and_merkelize(ntt, d_ext, d_tree, d_in, n_ext, n, c)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record()
        for _ in range(50):
            zk.extend_and_merkelize(ntt, d_ext, d_tree, d_in, n_ext, n, c)
        e1.record()
        t1 = time.perf_counter()
        torch.cuda.synchronize()
        print("2^%d x %d b%d run %d: device %.3f ms/call, host enqueue %.3f ms/call, stages %s" %
              (log_n, c, 1 << log_b, rep, e0.elapsed_time(e1) / 50, (t1 - t0) / 50 * 1e3, zk.stage_ms()), flush=True)
