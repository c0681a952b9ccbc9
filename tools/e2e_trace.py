#!/usr/bin/env python3
"""Timeline of one end-to-end commit call with a pinned host trace (ZKGPU_PIPE_TRACE=1; development aid)."""
import os, sys, time, torch
os.environ["ZKGPU_PIPE_TRACE"] = "1"
sys.path.insert(0, ".")
import zkevm_prover_b200 as zk

log_n, c = 23, 665
n, n_ext = 1 << log_n, 2 << log_n
zk.set_stream(torch.cuda.current_stream().cuda_stream)
h_in = torch.empty(n * c, dtype=torch.int64, pin_memory=True)
d = torch.empty(n * 64, dtype=torch.int64, device="cuda")
zk.fill(d, n * 64, 1, 0)
torch.cuda.synchronize()
for k in range(0, n * c, n * 64):
    m = min(n * 64, n * c - k)
    h_in[k:k + m].copy_(d[:m])
del d
h_tree = torch.empty((2 * n_ext - 1) * 4, dtype=torch.int64, pin_memory=True)
d_ext = torch.empty(n_ext * c, dtype=torch.int64, device="cuda")
ntt = zk.NTT_Goldilocks(n)
for rep in range(2):
    t = time.perf_counter()
    zk.extend_and_merkelize(ntt, d_ext, h_tree.data_ptr(), h_in.data_ptr(), n_ext, n, c)
    torch.cuda.synchronize()
    print("call %d: %.1f ms wall" % (rep, (time.perf_counter() - t) * 1e3), file=sys.stderr, flush=True)
