#!/usr/bin/env python3
"""BASELINE.json config #5: standalone NTT / INTT / extendPol(blowup 2) sweep over sizes and widths on one B200, device
resident, CUDA events; prints achieved algorithmic GB/s (16 B per element for a transform, 24 B per input cell for the
LDE) and the fraction of the measured HBM peak.  Run on the GPU box: python tools/ntt_sweep.py > gpurun_out/ntt_sweep.txt"""
import json, pathlib, sys
ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import torch
import zkevm_prover_b200 as zk

zk.init(0)
zk.set_stream(torch.cuda.current_stream())
peak = 6537.6
pp = ROOT / "MEASURED_PEAKS.json"
if pp.exists():
    peak = json.loads(pp.read_text())["hbm_gbs"]


def timeit(fn, reps=5):
    fn(); torch.cuda.synchronize()
    best = 1e9
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best


print("# log2(rows) cols | NTT ms GB/s frac | INTT ms GB/s frac | extendPol(b=2) ms GB/s frac   (peak %.0f GB/s measured copy)" % peak)
import os
COLS = tuple(int(x) for x in os.environ.get("NTT_SWEEP_COLS", "1,3,8,32,100,128,665,1024").split(","))
LOGS = tuple(int(x) for x in os.environ.get("NTT_SWEEP_LOGS", "12,14,16,18,20,22,24").split(","))
for lg in LOGS:
    for c in COLS:
        n = 1 << lg
        if n * c * 8 * 4 > 60e9 or n * c < (1 << 16):
            continue
        a = torch.empty(n * c, dtype=torch.int64, device="cuda")
        b = torch.empty(n * c, dtype=torch.int64, device="cuda")
        zk.fill(a, n * c)
        ntt = zk.NTT_Goldilocks(n)
        t_f = timeit(lambda: ntt.NTT(b, a, n, c))
        t_i = timeit(lambda: ntt.INTT(b, a, n, c))
        del b
        e = torch.empty(2 * n * c, dtype=torch.int64, device="cuda")
        t_e = timeit(lambda: ntt.extendPol(e, a, 2 * n, n, c))
        g = lambda bytes_, ms: bytes_ / ms / 1e6
        print("%2d %5d | %8.3f %7.0f %.3f | %8.3f %7.0f %.3f | %8.3f %7.0f %.3f" % (
            lg, c, t_f, g(16 * n * c, t_f), g(16 * n * c, t_f) / peak, t_i, g(16 * n * c, t_i), g(16 * n * c, t_i) / peak,
            t_e, g(24 * n * c, t_e), g(24 * n * c, t_e) / peak), flush=True)
        del a, e
        torch.cuda.empty_cache()
