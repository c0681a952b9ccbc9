#!/usr/bin/env python3
"""Feasibility probe: does the sponge kernel (FMA-pipe bound) overlap with the LDE kernels (ALU-pipe bound) when both are
resident on the same SMs?  Sponge on stream B over a 2^22 x 665 matrix, LDE strips of 192 columns on stream A."""
import sys, pathlib
sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent))
import torch
import zkevm_prover_b200 as zk

zk.init(0)
L = zk.lib()
lg, C, w = 21, 665, 192
n, n_ext = 1 << lg, 2 << lg
A, B = torch.cuda.Stream(), torch.cuda.Stream()
ext = torch.empty(n_ext * C, dtype=torch.int64, device="cuda")
zk.set_stream(A)
zk.fill(ext, n_ext * C)
src = torch.empty(n * w, dtype=torch.int64, device="cuda")
zk.fill(src, n * w)
dst = torch.empty(n_ext * w, dtype=torch.int64, device="cuda")
leaves = torch.empty(n_ext * 4, dtype=torch.int64, device="cuda")
torch.cuda.synchronize()


def sponge(threads):
    L.zkgpu_tune(0, threads)
    zk.set_stream(B)
    zk._check(L.zkgpu_merkle_absorb(leaves.data_ptr(), ext.data_ptr(), 664, C, n_ext, 1))


def lde(cap, reps):
    L.zkgpu_tune(1, cap)
    zk.set_stream(A)
    for _ in range(reps):
        zk._check(L.zkgpu_lde_columns(dst.data_ptr(), w, src.data_ptr(), w, n_ext, n, w))


def timed(fn):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record(torch.cuda.default_stream()); A.wait_event(e0); B.wait_event(e0)
    fn()
    ea, eb = torch.cuda.Event(), torch.cuda.Event()
    ea.record(A); eb.record(B)
    torch.cuda.default_stream().wait_event(ea); torch.cuda.default_stream().wait_event(eb)
    e1.record(torch.cuda.default_stream()); torch.cuda.synchronize()
    return e0.elapsed_time(e1)


reps = 3  # LDE of 3 x 192 columns ~ the sponge's 664 columns
print("sponge alone 128 thr: %.1f ms" % timed(lambda: sponge(128)))
print("sponge alone 512 thr: %.1f ms" % timed(lambda: sponge(512)))
print("lde x%d alone, full SM: %.1f ms" % (reps, timed(lambda: lde(0, reps))))
for cap in (256, 512):
    print("lde x%d alone, cap %d thr/SM: %.1f ms" % (reps, cap, timed(lambda: lde(cap, reps))))
for thr, cap in ((512, 256), (512, 512), (128, 256), (128, 0), (512, 0)):
    print("sponge %d thr || lde x%d cap %d: %.1f ms" % (thr, reps, cap, timed(lambda: (sponge(thr), lde(cap, reps)))))
    print("lde first, then sponge %d thr || cap %d: %.1f ms" % (thr, cap, timed(lambda: (lde(cap, reps), sponge(thr)))))
