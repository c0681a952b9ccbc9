#!/usr/bin/env python3
"""Quick device-side timing probe (development aid, not the benchmark): stage times of extend_and_merkelize on
device-resident synthetic traces of a few shapes, plus box facts (host RAM, cores, GPU memory)."""
import os
import sys
import pathlib
import time

sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent))
import torch
import zkevm_prover_b200 as zk


def main():
    shapes = [(int(a), int(b), int(c)) for a, b, c in (s.split("x") for s in sys.argv[1:])] or [(20, 64, 1), (22, 128, 1)]
    zk.init(0)
    free, total = torch.cuda.mem_get_info()
    print("gpu", torch.cuda.get_device_name(0), "free %.1f GB total %.1f GB" % (free / 1e9, total / 1e9))
    print("host cores", os.cpu_count(), "mem", open("/proc/meminfo").readline().strip())
    zk.set_stream(torch.cuda.current_stream())
    for lg, c, lb in shapes:
        n, b = 1 << lg, 1 << lb
        d_in = torch.empty(n * c, dtype=torch.int64, device="cuda")
        d_ext = torch.empty(n * b * c, dtype=torch.int64, device="cuda")
        d_tree = torch.empty((2 * n * b - 1) * 4, dtype=torch.int64, device="cuda")
        zk.fill(d_in, n * c)
        ntt = zk.NTT_Goldilocks(n)
        for it in range(3):
            t0 = time.time()
            zk.extend_and_merkelize(ntt, d_ext, d_tree, d_in, n * b, n, c)
            ms = zk.stage_ms()
            wall = (time.time() - t0) * 1e3
            tot = ms["lde"] + ms["leaves"] + ms["nodes"]
            print("2^%d x %d b%d it%d: lde %.2f ms leaves %.2f ms nodes %.2f ms total %.2f ms wall %.1f | %.1f Mcells/s | lde alg %.0f GB/s"
                  % (lg, c, b, it, ms["lde"], ms["leaves"], ms["nodes"], tot, wall, n * c / tot / 1e3,
                     8 * c * n * (1 + b) / ms["lde"] / 1e6))
        root = d_tree[-4:].cpu().numpy().view("uint64")
        print("   root", " ".join("%016x" % int(x) for x in root))
        del d_in, d_ext, d_tree
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
