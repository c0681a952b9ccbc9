#!/usr/bin/env python3
"""bench.py — the STARK commit hot path on B200: LDE (extendPol) + Poseidon Merkle tree, Mcells/s.

  python bench.py --gpus N --steps K --warmup W            our arm (CUDA, libzkgpu.so)
  python bench.py --impl reference ...                      the reference's own CPU code (oracle/_ref) on host cores

One "step" = one extendAndMerkelize of a synthetic trace (in[i] = splitmix64(seed+i) mod p):
N=1: 2^23 rows x 665 columns, blowup 2 (BASELINE.json configs[1]); N>1: the same trace column-sharded over N GPUs
(configs[2]) — total work fixed, so "scaling": "strong".  A cell is one element of the un-extended trace.
Prints ONE JSON line (rank 0).  See DESIGN.md §Measurement for the roofline arithmetic.
"""
import argparse
import ctypes
import json
import os
import pathlib
import statistics
import subprocess
import sys
import tempfile
import time

ROOT = pathlib.Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))
SEED = 0x9E3779B97F4A7C15
METRIC = "lde_merkle_commit_throughput"
UNIT = "Mcells/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--log-rows", type=int, default=23)
    ap.add_argument("--cols", type=int, default=665)
    ap.add_argument("--log-blowup", type=int, default=1)
    ap.add_argument("--e2e-steps", type=int, default=5)
    ap.add_argument("--no-full-host", action="store_true", help="skip e2e_full_host (all three buffers pinned host memory)")
    ap.add_argument("--no-small", action="store_true", help="skip the small configurations (BASELINE.json configs[0], configs[3])")
    ap.add_argument("--cpu-lde-cols", type=int, default=16, help="columns of the full-height LDE strip of the CPU sample")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--cpu-log-rows", type=int, default=16, help="rows of the bounded CPU-baseline sample")
    return ap.parse_args()


def workload_name(a):
    return "2^%d rows x %d cols, blowup %d, extendPol + Poseidon merkletree" % (a.log_rows, a.cols, 1 << a.log_blowup)


# ------------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(gpu_index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                       "-lms", "200"], stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            pass

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.p.terminate()
        try:
            self.p.wait(5)
        except subprocess.TimeoutExpired:
            self.p.kill()
        self.f.flush()
        rows = [l.strip().split(", ") for l in open(self.f.name) if l.strip()]
        os.unlink(self.f.name)
        sm, mx, reasons = [], None, set()
        for r in rows:
            if len(r) < 9:
                continue
            try:
                sm.append(float(r[1]))
                mx = float(r[2])
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                if v.strip().lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------------ CPU reference leg
def reference_lib():
    """oracle/_ref/libzkref.so = the UNMODIFIED reference sources compiled by oracle/Makefile ("kind": "reference");
    falls back to the plain-C restatement oracle/liboracle.so ("kind": "port")."""
    p = ROOT / "oracle" / "_ref" / "libzkref.so"
    if p.exists():
        return ctypes.CDLL(str(p)), "reference"
    p = ROOT / "oracle" / "liboracle.so"
    if not p.exists():
        subprocess.check_call(["make", "-s", "-f", str(ROOT / "oracle" / "Makefile"), str(p)])
    return ctypes.CDLL(str(p)), "port"


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def cpu_sample(args, steps, warmup):
    """Times the reference's CPU path (oracle/_ref = the unmodified reference sources) on a bounded sample of the workload,
    with every host thread this process may use, and composes a per-cell cost the way BASELINE.md section 2 says:
      * LDE (NTT_Goldilocks::extendPol) at the FULL height 2^log_rows on a strip of --cpu-lde-cols columns — the transform is
        column-independent (ntt_goldilocks.cpp:103-110), so its cost is linear in the number of columns;
      * Merkle tree (PoseidonGoldilocks::merkletree) at the FULL width on 2^(cpu_log_rows + log_blowup) extended rows — leaf
        hashing is row-independent (poseidon_goldilocks.cpp:277-281), so its cost is linear in the number of rows.
    value = 1 / (t_lde / cells_lde + t_merkle / cells_merkle), labelled "extrapolated"."""
    import numpy as np
    # torchrun exports OMP_NUM_THREADS=1; the reference arm must use the host's cores whatever launched it
    os.environ.pop("OMP_NUM_THREADS", None)
    L, kind = reference_lib()
    cores = host_threads()
    if kind == "reference":
        L.zkref_set_threads(ctypes.c_int(cores))
        L.zkref_threads.restype = ctypes.c_int
        cores = L.zkref_threads()
    else:
        L.oracle_set_threads(ctypes.c_int(cores))
    u64, vp = ctypes.c_uint64, lambda x: x.ctypes.data_as(ctypes.c_void_p)
    olib = ctypes.CDLL(str(ROOT / "oracle" / "liboracle.so"))
    b, c = 1 << args.log_blowup, args.cols
    n_l, c_l = 1 << args.log_rows, min(args.cpu_lde_cols, c)
    n_m = 1 << min(args.cpu_log_rows, args.log_rows)
    a = np.empty(n_l * c_l, dtype=np.uint64)
    olib.oracle_fill_block(vp(a), u64(n_l), u64(0), u64(c_l), u64(c), u64(SEED))
    out = np.zeros(n_l * b * c_l, dtype=np.uint64)
    buf = np.zeros(n_l * b * c_l, dtype=np.uint64)
    rows = np.empty(n_m * b * c, dtype=np.uint64)
    olib.oracle_fill(vp(rows), u64(rows.size), u64(SEED), u64(1))
    tree = np.zeros((2 * n_m * b - 1) * 4, dtype=np.uint64)

    def once():
        t0 = time.perf_counter()
        if kind == "reference":
            L.zkref_extend_pol(vp(out), vp(a), u64(n_l * b), u64(n_l), u64(c_l), vp(buf))
        else:
            L.oracle_extend_pol(vp(out), vp(a), u64(n_l * b), u64(n_l), u64(c_l))
        t1 = time.perf_counter()
        if kind == "reference":
            L.zkref_merkletree(vp(tree), vp(rows), u64(c), u64(n_m * b), u64(1))
        else:
            L.oracle_merkletree(vp(tree), vp(rows), u64(c), u64(n_m * b), u64(1))
        return t1 - t0, time.perf_counter() - t1

    for _ in range(warmup):
        once()
    lde, mt = [], []
    for _ in range(steps):
        t_l, t_m = once()
        lde.append(t_l)
        mt.append(t_m)
    t_l, t_m = sum(lde) / steps, sum(mt) / steps
    per_cell = t_l / (n_l * c_l) + t_m / (n_m * c)
    cells = (1 << args.log_rows) * c
    return {"value": 1.0 / per_cell / 1e6, "unit": UNIT, "cores": cores, "kind": kind,
            "sample": "extrapolated: extendPol at full height 2^%d on a %d-column strip (linear in columns) + merkletree at full width "
                      "%d on 2^%d extended rows (linear in rows), blowup %d, %d step(s)" % (args.log_rows, c_l, c, args.cpu_log_rows + args.log_blowup, b, steps),
            "lde_ms": 1e3 * t_l / (n_l * c_l) * cells, "merkletree_ms": 1e3 * t_m / (n_m * c) * cells, "ms_per_step": 1e3 * per_cell * cells,
            "measured_s_per_sample": t_l + t_m, "measured_lde_s": t_l, "measured_merkletree_s": t_m,
            "extrapolated": True}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    t0 = time.time()
    r = cpu_sample(args, max(1, args.steps), min(1, max(0, args.warmup)))
    line = {"impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "u64", "data": "synthetic",
            "config": {"workload": workload_name(args), "sample": r["sample"]},
            "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "stages_ms": {"STARK_STEP_1_LDE": r["lde_ms"], "STARK_STEP_1_MERKLETREE": r["merkletree_ms"]},
            "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "measured_s_per_sample": r["measured_s_per_sample"], "measured_lde_s": r["measured_lde_s"],
            "measured_merkletree_s": r["measured_merkletree_s"], "extrapolated": True,
            "gpu_launches": 0, "wall_s": time.time() - t0}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------ our arm
def known_roots():
    """Roots computed with the COMPILED REFERENCE (tools/pin_headline_root.py -> tests/golden/headline_pin.json), keyed by
    (log_rows, cols, log_blowup).  root_ok compares the GPU's root with these, never with a GPU-made constant."""
    out = {}
    p = ROOT / "tests" / "golden" / "headline_pin.json"
    if p.exists():
        d = json.loads(p.read_text())
        out[(d["log_rows"], d["cols"], d["log_blowup"])] = " ".join(d["root"])
    return out


def pin_for(log_rows, cols, log_blowup):
    p = ROOT / "tests" / "golden" / "headline_pin.json"
    if not p.exists():
        return None
    d = json.loads(p.read_text())
    return d if (d["log_rows"], d["cols"], d["log_blowup"]) == (log_rows, cols, log_blowup) else None


def check_pin(pin, zk, torch, dist, world, rank, n_ext, c, d_tree, rows_ptr, row0, nrows):
    """Compare the committed result with the samples the COMPILED REFERENCE produced (tests/golden/headline_pin.json):
    on every rank the sampled rows of the extended trace that fall in its row shard [row0, row0 + nrows) at device address
    rows_ptr, on rank 0 the sampled leaves (level 0) and the sampled nodes of every level of the gathered tree.
    Covers what root_ok cannot see: the level slices stitched into rank 0's tree and the row shards."""
    import numpy as np
    hexw = lambda ws: np.array([int(w, 16) for w in ws], dtype=np.uint64)
    rows_ok, nrows_checked = True, 0
    buf = np.zeros(c, dtype=np.uint64)
    for idx, words in zip(pin["rows"]["index"], pin["rows"]["words"]):
        if row0 <= idx < row0 + nrows:
            zk._check(zk.lib().zkgpu_memcpy(buf.ctypes.data, rows_ptr + 8 * (idx - row0) * c, 8 * c))
            rows_ok = rows_ok and bool(np.array_equal(buf, hexw(words)))
            nrows_checked += 1
    leaves_ok = levels_ok = True
    nleaves = nnodes = 0
    if rank == 0:
        off = 0
        for lv in pin["levels"]:
            idx = torch.tensor(lv["index"], device="cuda")
            got = d_tree[off: off + lv["nodes"] * 4].view(-1, 4)[idx].cpu().numpy().view(np.uint64).reshape(-1)
            ok = bool(np.array_equal(got, hexw([w for dg in lv["digests"] for w in dg])))
            if lv["level"] == 0:
                leaves_ok, nleaves = ok, len(lv["index"])
            else:
                levels_ok, nnodes = levels_ok and ok, nnodes + len(lv["index"])
            off += lv["nodes"] * 4
    flags = torch.tensor([int(rows_ok), nrows_checked], dtype=torch.int64, device="cuda")
    if world > 1:
        allf = [torch.zeros_like(flags) for _ in range(world)]
        dist.all_gather(allf, flags)
        rows_ok = all(int(f[0]) == 1 for f in allf)
        nrows_checked = sum(int(f[1]) for f in allf)
    return {"rows_ok": rows_ok, "leaves_ok": leaves_ok, "levels_ok": levels_ok,
            "checked": {"rows": nrows_checked, "leaves": nleaves, "nodes_above_leaves": nnodes, "levels": len(pin["levels"])},
            "source": "compiled reference, tests/golden/headline_pin.json"}


def run_ours(args):
    os.environ.setdefault("PYTORCH_CUDA_ALLOC_CONF", "expandable_segments:True")
    # NCCL / torchrun may print to stdout; keep fd 1 clean for the single JSON line
    real_stdout = os.dup(1)
    os.dup2(2, 1)
    import torch
    import torch.distributed as dist
    import zkevm_prover_b200 as zk
    from zkevm_prover_b200 import dist_commit as dc

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    zk.init(local)
    stream = torch.cuda.current_stream()
    zk.set_stream(stream)

    n, b, c = 1 << args.log_rows, 1 << args.log_blowup, args.cols
    n_ext = n * b
    cells = n * c
    peaks_path = ROOT / "MEASURED_PEAKS.json"
    if peaks_path.exists():
        peak_gbs, peak_src = json.loads(peaks_path.read_text())["hbm_gbs"], "measured (MEASURED_PEAKS.json hbm_gbs)"
    else:
        peak_gbs, peak_src = 6650.0, "fallback (B200_PROFILING.md)"

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    stage_acc = {}
    ntt = zk.NTT_Goldilocks(n)
    if world == 1:
        d_in = torch.empty(cells, dtype=torch.int64, device="cuda")
        d_ext = torch.empty(n_ext * c, dtype=torch.int64, device="cuda")
        d_tree = torch.empty((2 * n_ext - 1) * 4, dtype=torch.int64, device="cuda")
        zk.fill(d_in, cells, SEED, 0)

        def step(timed):
            zk.extend_and_merkelize(ntt, d_ext, d_tree, d_in, n_ext, n, c)
            if timed:
                for k, v in zk.stage_ms().items():
                    stage_acc[k] = stage_acc.get(k, 0.0) + v
    else:
        # one process per GPU, everything through the C ABI (csrc/dist.cu): zkgpu_dist_init + zkgpu_dist_commit
        cd = dc.CDist(dist, local)
        zk.set_stream(stream)
        rounds, plan = cd.plan(c, host_input=False)

        def make_pieces(the_plan, pinned):
            """rank g holds, for every round k, columns the_plan[k][g] of the same global trace: element (i, j) = gen(i*C + j)"""
            out = [torch.empty(max(n * the_plan[k][rank][1], 1), dtype=torch.int64, device="cuda") for k in range(len(the_plan))]
            tmp = torch.empty(min(n, 1 << 16) * c, dtype=torch.int64, device="cuda")
            rows_per = tmp.numel() // c
            for r0 in range(0, n, rows_per):
                zk.fill(tmp, rows_per * c, SEED, r0 * c)
                for k in range(len(the_plan)):
                    c0, w = the_plan[k][rank]
                    if w:
                        zk._check(zk.lib().zkgpu_copy_block(out[k].data_ptr() + 8 * r0 * w, w, tmp.data_ptr() + 8 * c0, c, rows_per, w))
            zk.sync()
            del tmp
            if pinned:
                host = []
                for t in out:
                    hp = torch.empty(t.numel(), dtype=torch.int64, pin_memory=True)
                    hp.copy_(t)
                    host.append(hp)
                out = host
            return out

        pieces = make_pieces(plan, False)
        d_tree = torch.empty((2 * n_ext - 1) * 4, dtype=torch.int64, device="cuda") if rank == 0 else None

        def step(timed):
            cd.commit(pieces, n, n_ext, c, d_tree, host_input=False, gather=True)
            if timed:
                for k, v in zk.stage_ms().items():
                    stage_acc[k] = stage_acc.get(k, 0.0) + v

    for _ in range(args.warmup):
        step(False)
    barrier()
    launches0 = zk.lib().zkgpu_launch_count()
    sampler = ClockSampler(local) if rank == 0 else None
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        step(True)
    e1.record(stream)
    barrier()
    ms_total = max_over_ranks(e0.elapsed_time(e1))
    clocks = sampler.stop() if sampler else None
    launches = zk.lib().zkgpu_launch_count() - launches0
    ms_per_step = ms_total / args.steps
    value = cells / ms_per_step / 1e3
    stages = {k: v / args.steps for k, v in stage_acc.items()}
    root = None
    if rank == 0:
        root = " ".join("%016x" % (int(x) & (2**64 - 1)) for x in d_tree[-4:].cpu().tolist())
    parity_pin = None
    pin = pin_for(args.log_rows, c, args.log_blowup)
    if pin is not None:
        if world == 1:
            parity_pin = check_pin(pin, zk, torch, dist, 1, 0, n_ext, c, d_tree, d_ext.data_ptr(), 0, n_ext)
        else:
            zk.sync()
            p_rows, R, cc, _ = cd.rows()
            assert cc == c and R * world == n_ext
            parity_pin = check_pin(pin, zk, torch, dist, world, rank, n_ext, c, d_tree, p_rows, rank * R, R)
        del pin

    # ---- end to end through the C ABI with HOST buffers (pinned): H2D of the trace and D2H of the tree every step,
    # the extended trace stays resident on the device (what getGroupProof / stage 4 read later).
    e2e = None
    if not args.no_e2e:
        if world == 1:
            h_in = torch.empty(cells, dtype=torch.int64, pin_memory=True)
            h_in.copy_(d_in)
            del d_in
            torch.cuda.empty_cache()
            h_tree = torch.empty((2 * n_ext - 1) * 4, dtype=torch.int64, pin_memory=True)
            a_in, a_tree = h_in.data_ptr(), h_tree.data_ptr()

            def e2e_step():
                zk.extend_and_merkelize(ntt, d_ext, a_tree, a_in, n_ext, n, c)
            h2d, d2h = cells * 8, h_tree.numel() * 8
        else:
            h_rounds, h_plan = cd.plan(c, host_input=True)
            del pieces
            torch.cuda.empty_cache()
            h_pieces = make_pieces(h_plan, True)
            torch.cuda.empty_cache()
            h_tree = torch.empty((2 * n_ext - 1) * 4, dtype=torch.int64, pin_memory=True) if rank == 0 else None

            def e2e_step():
                # pinned host pieces in (staged round by round under the compute), the whole tree back to rank 0's host memory
                cd.commit(h_pieces, n, n_ext, c, h_tree, host_input=True, gather=True)
            h2d, d2h = cells * 8, ((2 * n_ext - 1) * 4) * 8
        e2e_step()
        barrier()
        x0, x1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        x0.record(stream)
        for _ in range(args.e2e_steps):
            e2e_step()
        x1.record(stream)
        barrier()
        e2e_ms = max_over_ranks(x0.elapsed_time(x1)) / args.e2e_steps
        e2e = {"value": cells / e2e_ms / 1e3, "unit": UNIT, "ms_per_step": e2e_ms, "h2d_bytes_per_step": h2d,
               "d2h_bytes_per_step": d2h, "steps": args.e2e_steps,
               "label": "resident" if world == 1 else "resident, pre-sharded input",
               "note": "host pinned trace -> device every step, tree -> host every step; extended trace stays device-resident (see e2e_full_host "
                       "for the all-host call)" + ("" if world == 1 else "; every rank's column pieces were packed into contiguous pinned buffers "
                       "BEFORE the timed region (1-D copies), a caller holding one row-major trace pays strided copies instead")}
        if rank == 0:
            got = " ".join("%016x" % (int(x) & (2**64 - 1)) for x in h_tree[-4:].tolist())
            assert got == root, "e2e root differs from the device-resident run"

    # ---- the same call with ALL THREE buffers in pinned host memory: what an unchanged caller of the drop-in classes pays
    # (starks.cpp:53 writes p_cm1_2ns in host memory, :255-257 reads it on the CPU): the 89.3 GB extended trace goes home too.
    e2e_full = None
    if e2e is not None and world == 1 and not args.no_full_host:
        ext_bytes = n_ext * c * 8
        avail = 0
        try:
            for l in open("/proc/meminfo"):
                if l.startswith("MemAvailable:"):
                    avail = int(l.split()[1]) * 1024
        except OSError:
            pass
        if avail < ext_bytes + (24 << 30):
            e2e_full = {"skipped": "host has %.0f GB available, the pinned extended trace needs %.0f GB" % (avail / 1e9, ext_bytes / 1e9)}
        else:
            del d_ext
            torch.cuda.empty_cache()
            p_ext = ctypes.c_void_p()
            zk._check(zk.lib().zkgpu_malloc_host(ctypes.byref(p_ext), ext_bytes))
            try:
                steps_f = max(2, min(3, args.e2e_steps))
                zk.extend_and_merkelize(ntt, p_ext.value, a_tree, a_in, n_ext, n, c)
                torch.cuda.synchronize()
                f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                f0.record(stream)
                for _ in range(steps_f):
                    zk.extend_and_merkelize(ntt, p_ext.value, a_tree, a_in, n_ext, n, c)
                f1.record(stream)
                torch.cuda.synchronize()
                f_ms = f0.elapsed_time(f1) / steps_f
                got = " ".join("%016x" % (int(x) & (2**64 - 1)) for x in h_tree[-4:].tolist())
                e2e_full = {"value": cells / f_ms / 1e3, "unit": UNIT, "ms_per_step": f_ms, "h2d_bytes_per_step": cells * 8,
                            "d2h_bytes_per_step": ext_bytes + h_tree.numel() * 8, "steps": steps_f,
                            "root_ok": got == known_roots().get((args.log_rows, c, args.log_blowup), root),
                            "note": "pinned host trace in; extended trace AND tree back to pinned host memory every step"}
            finally:
                zk.lib().zkgpu_free_host(p_ext)

    # ---- the small configurations of BASELINE.json (configs[0] 2^16 x 8 blowup 2, configs[3] 2^17 x 12 blowup 16), device-resident
    small = None
    if world == 1 and not args.no_small:
        small = {}
        for name, (ln, cc, lb) in (("cfg1_2p16x8_b2", (16, 8, 1)), ("cfg4_2p17x12_b16", (17, 12, 4))):
            nn, ne = 1 << ln, 1 << (ln + lb)
            s_in = torch.empty(nn * cc, dtype=torch.int64, device="cuda")
            s_ext = torch.empty(ne * cc, dtype=torch.int64, device="cuda")
            s_tree = torch.empty((2 * ne - 1) * 4, dtype=torch.int64, device="cuda")
            zk.fill(s_in, nn * cc, SEED, 0)
            s_ntt = zk.NTT_Goldilocks(nn)
            for _ in range(5):
                zk.extend_and_merkelize(s_ntt, s_ext, s_tree, s_in, ne, nn, cc)
            torch.cuda.synchronize()
            reps = 50
            # host time to enqueue ONE commit on an idle stream (a loop of many would fill the launch queue and measure the GPU)
            t_enq = 1e9
            for _ in range(5):
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                zk.extend_and_merkelize(s_ntt, s_ext, s_tree, s_in, ne, nn, cc)
                t_enq = min(t_enq, time.perf_counter() - t0)
            torch.cuda.synchronize()
            g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            l0 = zk.lib().zkgpu_launch_count()
            g0.record(stream)
            for _ in range(reps):
                zk.extend_and_merkelize(s_ntt, s_ext, s_tree, s_in, ne, nn, cc)
            g1.record(stream)
            l1 = zk.lib().zkgpu_launch_count()
            torch.cuda.synchronize()
            ms = g0.elapsed_time(g1) / reps
            small[name] = {"ms_per_commit": ms, "Mcells_per_s": nn * cc / ms / 1e3, "host_enqueue_ms": 1e3 * t_enq, "launches_per_commit": (l1 - l0) / reps,
                           "root": " ".join("%016x" % (int(x) & (2**64 - 1)) for x in s_tree[-4:].cpu().tolist())}
            del s_in, s_ext, s_tree

    exchange_mode = cd.exchange() if world > 1 else None
    if world > 1:
        cd.close()
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel (leaf absorption: one launch per column strip on 1 GPU — the figures below are
    # per step, i.e. summed over the strips; one launch per step on each rank in the multi-GPU path)
    leaf_ms = stages.get("leaves")
    leaf_bytes = (8 * c * n_ext + 32 * n_ext) / world  # reads its rows of the extended trace once, writes 4-word digests
    traffic, traffic_note = None, None
    tp = ROOT / "profiles" / "traffic.json"
    if tp.exists():
        tj = json.loads(tp.read_text())
        traffic = tj.get("absorb_kernel@%dx%d/%d" % (args.log_rows, c, world))
        traffic_note = "ncu capture (profiles/), dram__bytes_read.sum + dram__bytes_write.sum summed over the step's launches"
        if world > 1 and traffic is not None:
            traffic_note = tj.get("_note_multi", traffic_note)
        if traffic is None and tj.get("absorb_kernel@%dx%d/1" % (args.log_rows, c)) is not None:
            # the same kernel over 1/world of the rows per rank: the 1-GPU capture scaled, not a capture of this run
            traffic = tj["absorb_kernel@%dx%d/1" % (args.log_rows, c)] / world
            traffic_note = "1-GPU ncu capture / %d (each rank runs the same kernel over 1/%d of the rows)" % (world, world)
    achieved = leaf_bytes / (leaf_ms * 1e-3) / 1e9 if leaf_ms else None
    roofline = {"kernel": "zk::absorb_kernel (Poseidon row sponge, per step = all strips)" if world == 1 else "zk::leaves_kernel (Poseidon row sponge)", "bound": "hbm", "achieved": achieved, "peak": peak_gbs,
                "unit": "GB/s", "frac": achieved / peak_gbs if achieved else None, "traffic": traffic, "traffic_source": traffic_note, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": leaf_bytes,
                "permutations_per_s": (n_ext * (-(-c // 8)) / world) / (leaf_ms * 1e-3) if leaf_ms else None,
                "note": "arithmetic-bound kernel (about 1.4e9 Poseidon permutations per step: 472 64-bit modular multiplies on the "
                        "integer-multiply pipe + 30 MDS products on the FP64 pipe each; the two pipes share a dispatch port that ncu "
                        "shows 95 % busy: profiles/r1_ncu_quick_poseidon_fp64.txt); HBM fraction is low by nature"}
    # arithmetic roof of the sponge kernel: the integer-multiply and FP64 pipes share one dispatch port per SM sub-partition
    # (profiles/r1_int_bench_fp64.txt: IMAD.WIDE/IMAD.HI occupy it 4 clocks per warp instruction, DFMA/DADD 2).  One
    # permutation needs at least 472 products x (4 IMAD.WIDE + 1 IMAD.HI) and 8 x 156 + 11 x 256 FP64 operations on it
    #  = 17568 clocks per warp of 32 permutations; the peak below is that bound at the measured SM clock.
    # pipe rates of THIS GPU in THIS run (zkgpu_pipe_probe, csrc/probe.cu): thread-instructions per second, whole GPU
    pipe = None
    try:
        pr = (ctypes.c_double * 8)()
        zk._check(zk.lib().zkgpu_pipe_probe(pr))
        zk.sync()
        pipe = {"imad_wide": pr[0], "iadd3": pr[1], "dfma": pr[2], "dfma_plus_imad_wide_1to1": pr[3], "gl_mul_nc": pr[4], "butterfly": pr[5],
                "unit": "thread-instructions (operations for gl_mul_nc, butterfly) per second per GPU", "sms": int(pr[6]),
                "source": "measured in this run, register-only loops, 32 warps/SM (csrc/probe.cu)"}
    except Exception as e:  # the probe is evidence, not the product: say why it is missing
        pipe = {"error": str(e)}
    port_clocks = 472 * 5 * 4 + (8 * 156 + 11 * 256) * 2
    if roofline["permutations_per_s"]:
        sm_clock = 1.965e9
        try:
            sm_clock = float(clocks["sm_mhz"]) * 1e6 if clocks and clocks.get("sm_mhz") else sm_clock
        except Exception:
            pass
        peak_perm = 148 * 4 * sm_clock / port_clocks * 32
        roofline["arithmetic"] = {"bound": "IMAD+FP64 dispatch port", "achieved": roofline["permutations_per_s"], "peak": peak_perm,
                                  "unit": "permutations/s per GPU", "frac": roofline["permutations_per_s"] / peak_perm,
                                  "port_clocks_per_warp_permutation": port_clocks,
                                  "peak_source": "model: 4 port clocks per IMAD.WIDE/IMAD.HI warp instruction, 2 per DFMA/DADD, at the sampled SM clock"}
        if pipe and "imad_wide" in pipe:
            # the same operation counts (472 products x 5 multiplier instructions, 8 x 156 + 11 x 256 FP64 operations per permutation,
            # DESIGN.md section 3) priced with the MEASURED rates.  The two pipes share a dispatch port when the 1:1 mix runs no
            # faster than the sum of the separate times; if it does run faster they overlap and only the longer one counts.
            n_mul, n_fp = 472 * 5, 8 * 156 + 11 * 256
            t_mul, t_fp = n_mul / pipe["imad_wide"], n_fp / pipe["dfma"]
            t_mix_pair = 2.0 / pipe["dfma_plus_imad_wide_1to1"]                      # one DFMA + one IMAD.WIDE, interleaved
            t_sep_pair = 1.0 / pipe["imad_wide"] + 1.0 / pipe["dfma"]
            # overlap factor of the two pipes as measured by the 1:1 mix: 1.0 = fully serialised on a shared port, down to
            # max/(sum) = fully overlapped.  The peak prices the permutation's two instruction streams with that factor.
            overlap = min(1.0, max(t_mix_pair / t_sep_pair, max(t_mul, t_fp) / (t_mul + t_fp)))
            peak_meas = 1.0 / ((t_mul + t_fp) * overlap)
            roofline["arithmetic"]["measured"] = {"peak": peak_meas, "frac": roofline["permutations_per_s"] / peak_meas,
                                                   "mix_over_separate": t_mix_pair / t_sep_pair, "overlap_factor_used": overlap,
                                                   "multiplier_instructions_per_permutation": n_mul, "fp64_operations_per_permutation": n_fp,
                                                   "note": "peak = 1 / ((multiplier instructions / measured IMAD.WIDE rate + FP64 operations / measured DFMA "
                                                           "rate) x overlap), overlap = time of the interleaved 1:1 mix over the sum of the separate times "
                                                           "(1.0 = one shared dispatch port, fully serialised)"}
    lde_ms = stages.get("lde")
    lde_bytes = 8 * c * (n + n_ext) / world
    roofline_lde = {"stage": "extendPol (INTT + coset NTTs, all passes)", "bound": "hbm",
                    "achieved": lde_bytes / (lde_ms * 1e-3) / 1e9 if lde_ms else None, "peak": peak_gbs, "unit": "GB/s",
                    "frac": lde_bytes / (lde_ms * 1e-3) / 1e9 / peak_gbs if lde_ms else None, "algorithmic_bytes": lde_bytes}
    # the transform against the pipe that binds it: ncu (profiles/r1_ncu_full_2p20x665_v3.csv) has 154
    # thread-instructions per element and 8-stage pass, ALU pipe 66-71 % of its peak at 63-65 % issue-slot utilisation, i.e. 53 %
    # of the instructions are ALU-pipe instructions = 82 per element-pass; one cell costs 3 + 2*3 = 9 element-passes at blowup 2.
    # ALU pipe peak: 16 lanes per SM sub-partition = 64 thread-instructions per clock per SM (tools/int_bench.cu, IADD3 class).
    if lde_ms and args.log_blowup == 1:
        alu_per_cell = 82.0 * 9 * (-(-args.log_rows // 8)) / 3.0
        sm_clock = float(clocks["sm_mhz"]) * 1e6 if clocks and clocks.get("sm_mhz") else 1.965e9
        alu_peak = 148 * 64 * sm_clock
        alu_rate = cells / world * alu_per_cell / (lde_ms * 1e-3)
        roofline_lde["arithmetic"] = {"bound": "ALU pipe (IADD3/LOP3/SHF/SEL)", "alu_instructions_per_cell": alu_per_cell, "achieved": alu_rate,
                                      "peak": alu_peak, "unit": "ALU thread-instructions/s per GPU", "frac": alu_rate / alu_peak,
                                      "source": "instruction counts from profiles/r1_ncu_full_2p20x665_v3.csv (154 per element-pass, 53 % ALU-pipe); "
                                                "peak = 64 ALU lanes per SM at the sampled clock (model)"}
        if pipe and "iadd3" in pipe:
            # measured ALU-pipe rate instead of the 64-lane model, and the synthetic ceiling of the butterflies themselves: the
            # transform needs (1 INTT + 2 coset NTTs of size N) = 3 * N/2 * log2 N butterflies per column at blowup 2; the probe's
            # butterfly loop (general twiddle, registers only, no loads, no index arithmetic) bounds how fast they can possibly go
            bf = 3.0 * (n / 2) * args.log_rows * c / world
            roofline_lde["arithmetic"]["measured"] = {"peak": pipe["iadd3"], "frac": alu_rate / pipe["iadd3"]}
            roofline_lde["synthetic_butterfly_ceiling"] = {"butterflies": bf, "butterflies_per_s": pipe["butterfly"],
                                                           "floor_ms": 1e3 * bf / pipe["butterfly"], "frac": (1e3 * bf / pipe["butterfly"]) / lde_ms,
                                                           "note": "floor_ms = butterflies / measured register-only butterfly rate; frac = floor / measured LDE time"}

    cpu = None
    if not args.no_cpu and world == 1:
        cpu = cpu_sample(args, 1, 1)
        cpu = {k: cpu[k] for k in ("value", "unit", "cores", "kind", "sample", "lde_ms", "merkletree_ms")}

    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "u64",
            "data": "synthetic",
            "config": {"workload": workload_name(args), "rows_log2": args.log_rows, "cols": c, "blowup": b,
                       "parallelism": "1 gpu" if world == 1 else "C ABI zkgpu_dist_commit: column pieces LDE x%d, exchange = %s, row-sharded sponge, NCCL barriers + tree stitch" % (world, {"stores": "peer stores fused into the LDE's last pass (NVLink)", "dma": "copy engines, peer cudaMemcpy2DAsync (NVLink)", "nccl": "ncclSend/ncclRecv"}[exchange_mode]),
                       "l2": "inputs (%.1f GB) larger than L2, no flush" % (cells * 8 / 1e9)},
            "stages_ms": {"STARK_STEP_1_LDE": stages.get("lde"), "EXCHANGE": stages.get("exchange"),  # N>1: exposed wait only
                          "MERKLE_LEAVES": stages.get("leaves"), "MERKLE_NODES": stages.get("nodes"),
                          "STARK_STEP_1_MERKLETREE": (stages.get("leaves") or 0) + (stages.get("nodes") or 0)},
            "roofline": roofline, "roofline_lde": roofline_lde, "cpu_baseline": cpu, "e2e": e2e, "e2e_full_host": e2e_full,
            "small_configs": small, "gpu_launches": int(launches),
            "clocks": clocks, "root": root, "parity_pin": parity_pin, "pipe_rates": pipe}
    want = known_roots().get((args.log_rows, c, args.log_blowup))
    if want is not None:
        line["root_ok"] = (root == want)
        line["root_source"] = "compiled reference, tests/golden/headline_pin.json"
    os.write(real_stdout, (json.dumps(line) + "\n").encode())
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
