#!/usr/bin/env python3
"""Development probe (GPU box): the all-host commit (pinned trace in, pinned extended trace + tree out) at 2^23 x 665 against the
width of the column strips (ZKGPU_HOST_STRIP).  tools/pcie_duplex_probe.py says wider strided segments move faster in both
directions at once (112 columns: 33.7 GB/s each way, 176: 36.3, 332: 40.1); wider strips also expose a longer first copy and last
copy-out.  python tools/host_strip_sweep.py [widths...] > gpurun_out/host_strip_sweep.txt"""
import ctypes, os, sys, time, torch
sys.path.insert(0, ".")
import zkevm_prover_b200 as zk

log_n, c = int(os.environ.get("LOGN", "23")), 665
n, n_ext = 1 << log_n, 2 << log_n
widths = [int(x) for x in sys.argv[1:]] or [112, 144, 176, 192]
zk.init(0)
zk.set_stream(torch.cuda.current_stream().cuda_stream)


def pinned(words):
    p = ctypes.c_void_p()
    zk._check(zk.lib().zkgpu_malloc_host(ctypes.byref(p), words * 8))
    return p.value


h_in, h_ext, h_tree = pinned(n * c), pinned(n_ext * c), pinned((2 * n_ext - 1) * 4)
d = torch.empty(n * 64, dtype=torch.int64, device="cuda")
for k in range(0, n * c, n * 64):
    m = min(n * 64, n * c - k)
    zk.fill(d, m, 0x9E3779B97F4A7C15, k)
    zk.sync()
    zk._check(zk.lib().zkgpu_memcpy(h_in + 8 * k, d.data_ptr(), 8 * m))
del d
torch.cuda.empty_cache()
ntt = zk.NTT_Goldilocks(n)
root = (ctypes.c_uint64 * 4).from_address(h_tree + 8 * ((2 * n_ext - 1) * 4 - 4))
for w in widths:
    os.environ["ZKGPU_HOST_STRIP"] = str(w)
    ts = []
    for rep in range(3):
        t = time.perf_counter()
        zk.extend_and_merkelize(ntt, h_ext, h_tree, h_in, n_ext, n, c)
        ts.append((time.perf_counter() - t) * 1e3)
    free, total = torch.cuda.mem_get_info()
    print("strip %4d columns: %s ms wall (first call includes growing the staging area); device memory in use %.1f GB; root %s"
          % (w, " ".join("%.1f" % x for x in ts), (total - free) / 1e9, " ".join("%016x" % x for x in root)), flush=True)
