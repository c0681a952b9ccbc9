"""Worker processes of tests/test_gpu_dist.py (multi-GPU parity through the C ABI).  Not collected by pytest.

  python gpu_dist_worker.py spawn WORLD LOG_N COLS LOG_B   process-per-GPU form (torch.distributed NCCL + zkgpu_dist_*)
  python gpu_dist_worker.py multi WORLD LOG_N COLS LOG_B   single-process form (zkgpu_init_multi + zkgpu_extend_and_merkelize)
Each compares every word of the tree and of every row shard with the 1-GPU result of the same library (which the
other GPU tests pin to the oracle / compiled reference), and for shapes up to 2^14 x 665 that 1-GPU result is first
compared with the CPU oracle here, so that the multi-GPU words are checked against the oracle and not only against the
library itself.  The single-process form also runs zkgpu_build_const_tree with host buffers after zkgpu_init_multi
(it must hand the context over to the worker threads) and a commit in which one rank fails to allocate (every rank
must come back with the error, and the next commit must work)."""
import ctypes
import os
import pathlib
import sys

import numpy as np

ROOT = pathlib.Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))
SEED = 0x9E3779B97F4A7C15


def single_gpu_result(zk, a, n, n_ext, c):
    ext = np.zeros((n_ext, c), dtype=np.uint64)
    tree = np.zeros((2 * n_ext - 1) * 4, dtype=np.uint64)
    zk.extend_and_merkelize(zk.NTT_Goldilocks(n), ext, tree, a, n_ext, n, c)
    if n * c <= (1 << 14) * 665:
        from oracle_lib import Oracle
        O = Oracle()
        want_ext = O.extend_pol(a, n_ext)
        assert np.array_equal(ext, want_ext), "1-GPU extended trace differs from the oracle"
        assert np.array_equal(tree, O.merkletree(want_ext, c, n_ext)), "1-GPU tree differs from the oracle"
        print("oracle checked", flush=True)
    return ext, tree


def spawn_rank(rank, world, log_n, c, log_b, port, host_input, exchange):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    if exchange:
        os.environ["ZKGPU_EXCHANGE"] = exchange
    import torch
    import torch.distributed as dist
    import zkevm_prover_b200 as zk
    from zkevm_prover_b200 import dist_commit as dc
    from oracle_lib import Oracle
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", device_id=torch.device("cuda", rank))
    try:
        n, n_ext = 1 << log_n, 1 << (log_n + log_b)
        a = Oracle().fill(n * c, SEED, 0).reshape(n, c)
        zk.init(rank)
        want_ext, want_tree = single_gpu_result(zk, a, n, n_ext, c)
        cd = dc.CDist(dist, rank)
        K, plan = cd.plan(c, host_input)
        pieces = []
        for k in range(K):
            c0, w = plan[k][rank]
            m = torch.from_numpy(np.ascontiguousarray(a[:, c0:c0 + w]).view(np.int64).reshape(-1).copy()) if w else torch.zeros(1, dtype=torch.int64)
            pieces.append(m.pin_memory() if host_input else m.cuda())
        R = n_ext // world
        for rep in range(2):  # twice: the second call reuses the row shards (start-of-call barrier)
            tree = torch.zeros((2 * n_ext - 1) * 4, dtype=torch.int64, device="cuda") if rank == 0 else None
            cd.commit(pieces, n, n_ext, c, tree, host_input=host_input, gather=True)
            zk.sync()
            p, r, cc, sub = cd.rows()
            assert (r, cc) == (R, c)
            rows = np.zeros((R, c), dtype=np.uint64)
            zk._check(zk.lib().zkgpu_memcpy(rows.ctypes.data, p, rows.nbytes))
            assert np.array_equal(rows, want_ext[rank * R:(rank + 1) * R]), "rank %d: row shard differs (rep %d)" % (rank, rep)
            if rank == 0:
                got = tree.cpu().numpy().view(np.uint64)
                bad = np.nonzero(got != want_tree)[0]
                assert bad.size == 0, "tree differs at %d words, first %d (rep %d)" % (bad.size, bad[0], rep)
        print("rank %d ok: exchange=%s rounds=%d" % (rank, cd.exchange(), K), flush=True)
        cd.close()
    finally:
        dist.destroy_process_group()


def main():
    mode, world, log_n, c, log_b = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5])
    extra = sys.argv[6] if len(sys.argv) > 6 else ""
    if mode == "spawn":
        import socket
        import torch.multiprocessing as mp
        s = socket.socket()
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
        s.close()
        mp.spawn(spawn_rank, args=(world, log_n, c, log_b, port, "host" in extra, "nccl" if "nccl" in extra else "stores" if "stores" in extra else ""),
                 nprocs=world, join=True)
        print("dist spawn: ok")
    elif mode == "multi":
        if "nccl" in extra:
            os.environ["ZKGPU_EXCHANGE"] = "nccl"
        if "stores" in extra:
            os.environ["ZKGPU_EXCHANGE"] = "stores"
        import torch  # before the library maps an NCCL: torch's own libnccl.so.2 must be the one this process holds
        import zkevm_prover_b200 as zk
        from oracle_lib import Oracle
        n, n_ext = 1 << log_n, 1 << (log_n + log_b)
        a = Oracle().fill(n * c, SEED, 0).reshape(n, c)
        zk.init(0)
        want_ext, want_tree = single_gpu_result(zk, a, n, n_ext, c)
        zk._check(zk.lib().zkgpu_init_multi(world))
        for rep in range(2):
            ext = np.zeros((n_ext, c), dtype=np.uint64)
            tree = np.zeros((2 * n_ext - 1) * 4, dtype=np.uint64)
            zk.extend_and_merkelize(zk.NTT_Goldilocks(n), ext, tree, a, n_ext, n, c)
            assert np.array_equal(ext, want_ext), "extended trace differs (rep %d)" % rep
            bad = np.nonzero(tree != want_tree)[0]
            assert bad.size == 0, "tree differs at %d words, first %d (rep %d)" % (bad.size, bad[0] if bad.size else -1, rep)
        # extended trace left on the devices, tree into memory of device 0
        dtree = torch.zeros((2 * n_ext - 1) * 4, dtype=torch.int64, device="cuda:0")
        zk.extend_and_merkelize(zk.NTT_Goldilocks(n), None, dtree, a, n_ext, n, c)
        zk.sync()
        assert np.array_equal(dtree.cpu().numpy().view(np.uint64), want_tree), "device-0 tree differs"
        # buildConstTree with host buffers while the multi-GPU job exists: the commit inside it goes to the worker threads,
        # which need the context the calling thread entered with
        nb, nbe, npols = min(log_n, 10), min(log_n, 10) + 1, 5
        cp = Oracle().fill((1 << nb) * npols, SEED, 7).reshape(1 << nb, npols)
        img = np.zeros(2 + npols * (1 << nbe) + (2 * (1 << nbe) - 1) * 4, dtype=np.uint64)
        zk.build_const_tree(img, cp, nb, nbe, npols)
        cext = Oracle().extend_pol(cp, 1 << nbe)
        assert img[0] == npols and img[1] == (1 << nbe)
        assert np.array_equal(img[2:2 + npols * (1 << nbe)], cext.reshape(-1)), "const tree: extended polynomials differ"
        assert np.array_equal(img[2 + npols * (1 << nbe):], Oracle().merkletree(cext, npols, 1 << nbe)), "const tree: nodes differ"
        # one rank fails to allocate: all ranks return, the caller sees the error, the job stays usable
        os.environ["ZKGPU_TEST_FAIL_RANK"] = str(world - 1)
        try:
            zk.extend_and_merkelize(zk.NTT_Goldilocks(n), ext, tree, a, n_ext, n, c)
            raise AssertionError("the failing rank was not reported")
        except zk.ZkGpuError as e:
            assert "ZKGPU_TEST_FAIL_RANK" in str(e) or "could not allocate" in str(e), str(e)
        del os.environ["ZKGPU_TEST_FAIL_RANK"]
        tree[:] = 0
        zk.extend_and_merkelize(zk.NTT_Goldilocks(n), ext, tree, a, n_ext, n, c)
        assert np.array_equal(tree, want_tree), "tree differs after a failed commit"
        r, w, ps = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
        zk._check(zk.lib().zkgpu_dist_info(ctypes.byref(r), ctypes.byref(w), ctypes.byref(ps)))
        zk._check(zk.lib().zkgpu_dist_finalize())
        print("dist multi: ok world=%d exchange=%d" % (w.value, ps.value))
    else:
        raise SystemExit("unknown mode")


if __name__ == "__main__":
    main()
