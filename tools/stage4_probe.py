#!/usr/bin/env python3
"""Stage 4 of Starks::genProof (starks.cpp:261-293) at the headline size, device-resident: INTT(2^24 x 3), regroup,
NTT(2^24 x 6), merkelize — zkgpu_stage4_commit; also the same through host pointers.  python tools/stage4_probe.py"""
import sys, time, pathlib
sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent))
import numpy as np, torch
import zkevm_prover_b200 as zk

zk.init(0)
zk.set_stream(torch.cuda.current_stream())
n, ne, qd, qg = 1 << 23, 1 << 24, 3, 2
d_q = torch.empty(ne * qd, dtype=torch.int64, device="cuda")
zk.fill(d_q, ne * qd, 7, 0)
d_cm4 = torch.empty(ne * qd * qg, dtype=torch.int64, device="cuda")
d_tree = torch.empty((2 * ne - 1) * 4, dtype=torch.int64, device="cuda")
for rep in range(3):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    zk.stage4_commit(None, d_cm4, d_tree, d_q, ne, n, qd, qg)
    e1.record()
    torch.cuda.synchronize()
    print("stage 4 device-resident, 2^23 -> 2^24, qDim 3, qDeg 2: %.2f ms  stages %s" % (e0.elapsed_time(e1), zk.stage_ms()))
h_q = d_q.cpu().numpy().view(np.uint64)
h_cm4 = np.zeros(ne * qd * qg, dtype=np.uint64)
h_tree = np.zeros((2 * ne - 1) * 4, dtype=np.uint64)
for rep in range(2):
    t = time.perf_counter()
    zk.stage4_commit(None, h_cm4, h_tree, h_q, ne, n, qd, qg)
    print("stage 4 with pageable host buffers (0.4 GB in, 0.8 + 1.07 GB out): %.1f ms wall" % ((time.perf_counter() - t) * 1e3))
print("root", " ".join("%016x" % int(x) for x in h_tree[-4:]), "== device", bool((d_tree[-4:].cpu().numpy().view(np.uint64) == h_tree[-4:]).all()))
