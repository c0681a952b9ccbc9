#!/usr/bin/env python3
"""Host->device bandwidth by NUMA node of the pinned buffer (development aid): the e2e leg is bound by this link."""
import os, glob, subprocess, time, torch
print(subprocess.run("nvidia-smi topo -m | head -20; lscpu | grep -i 'numa\\|model name\\|^CPU(s)'", shell=True, capture_output=True, text=True).stdout)
nodes = sorted(int(p.rsplit("node", 1)[1]) for p in glob.glob("/sys/devices/system/node/node[0-9]*"))
def cpus(n):
    out = []
    for part in open(f"/sys/devices/system/node/node{n}/cpulist").read().strip().split(","):
        a, _, b = part.partition("-"); out += list(range(int(a), int(b or a) + 1))
    return out
print("allowed cpus:", len(os.sched_getaffinity(0)), "numa nodes:", nodes)
try:
    print("gpu0 numa:", open(glob.glob("/sys/bus/pci/devices/*/numa_node")[0]).read().strip())
    for p in glob.glob("/sys/bus/pci/devices/*/numa_node"):
        cls = open(p.replace("numa_node", "class")).read().strip()
        if cls.startswith("0x0302") or cls.startswith("0x0300"):
            print(p, open(p).read().strip())
except Exception as e:
    print("numa probe:", e)
torch.cuda.init()
d = torch.empty(1 << 29, dtype=torch.int64, device="cuda")  # 4 GiB
allowed = os.sched_getaffinity(0)
for n in nodes:
    c = [x for x in cpus(n) if x in allowed]
    if not c:
        print("node", n, "no allowed cpus"); continue
    os.sched_setaffinity(0, c)
    h = torch.empty(1 << 29, dtype=torch.int64, pin_memory=True)
    h.zero_()
    for rep in range(2):
        torch.cuda.synchronize(); t = time.perf_counter()
        d.copy_(h, non_blocking=True); torch.cuda.synchronize()
        dt = time.perf_counter() - t
    print(f"node {n}: H2D {h.numel()*8/dt/1e9:.1f} GB/s", end="  ")
    torch.cuda.synchronize(); t = time.perf_counter()
    h.copy_(d, non_blocking=True); torch.cuda.synchronize()
    print(f"D2H {h.numel()*8/(time.perf_counter()-t)/1e9:.1f} GB/s")
    del h
os.sched_setaffinity(0, allowed)
