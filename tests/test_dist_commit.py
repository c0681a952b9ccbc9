"""Host logic of the multi-GPU commit (zkevm_prover_b200/dist_commit.py) on CPU: world_size 2 and 4 over gloo, with the
oracle standing in for the three device steps, checked against the single-process oracle result (extended-trace row
shards, every tree node, root)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from zkevm_prover_b200 import dist_commit as dc
from oracle_lib import Oracle, SEED


class OracleOps:
    """Stand-in for GpuOps on CPU tensors (test only)."""

    def __init__(self):
        self.O = Oracle()

    @staticmethod
    def _np(t):
        return t.numpy().view(np.uint64)

    def empty(self, nwords):
        return torch.zeros(nwords, dtype=torch.int64)

    def lde(self, out, inp, n_ext, n, ncols):
        a = self._np(inp)[:n * ncols].reshape(n, ncols)
        self._np(out)[:n_ext * ncols] = self.O.extend_pol(a, n_ext).reshape(-1)

    def copy_block(self, dst, dst_off, dst_ld, src, src_off, src_ld, nrows, ncols):
        d, s = self._np(dst), self._np(src)
        for r in range(nrows):
            d[dst_off + r * dst_ld: dst_off + r * dst_ld + ncols] = s[src_off + r * src_ld: src_off + r * src_ld + ncols]

    def leaves(self, tree, rows, ncols, nrows):
        t = self.O.merkletree(self._np(rows)[:nrows * ncols].copy(), ncols, nrows)
        self._np(tree)[:nrows * 4] = t[:nrows * 4]

    def absorb(self, leaves, block, block_off, width, ld, nrows, first, copy_to=None, copy_off=0, copy_ld=0):
        # running digest after a whole number of 8-word blocks == linear_hash of the row prefix absorbed so far
        blk = self._np(block)[block_off + np.arange(nrows)[:, None] * ld + np.arange(width)[None, :]]
        if copy_to is not None:
            self.copy_block(copy_to, copy_off, copy_ld, block, block_off, ld, nrows, width)
        self.prefix = blk.copy() if first else np.concatenate([self.prefix, blk], axis=1)
        w = self.prefix.shape[1]
        t = self.O.merkletree(np.ascontiguousarray(self.prefix).reshape(-1), w, nrows)
        self._np(leaves)[:nrows * 4] = t[:nrows * 4]

    def stage(self, src, slot):
        return (src, None, slot)

    def staged_wait(self, st):
        return st[0]

    def staged_release(self, st):
        pass

    def levels(self, tree, nleaves, word_off=0):
        t = self._np(tree)
        lv = dc.subtree_level_offsets(nleaves)
        for (o0, cnt), (o1, _) in zip(lv[:-1], lv[1:]):
            for i in range(cnt // 2):
                st = np.concatenate([t[word_off + o0 + 8 * i: word_off + o0 + 8 * i + 8], np.zeros(4, dtype=np.uint64)])
                t[word_off + o1 + 4 * i: word_off + o1 + 4 * i + 4] = self.O.hash_full_result(st)[:4]


def _worker(rank, world, port, n, b, c, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        O = Oracle()
        full = O.fill(n * c, SEED, 0).reshape(n, c)
        cuts = dc.column_cuts(c, world)
        mine = np.ascontiguousarray(full[:, cuts[rank]:cuts[rank + 1]])
        local = torch.from_numpy(mine.view(np.int64).reshape(-1).copy()) if mine.size else torch.zeros(1, dtype=torch.int64)
        rows, tree = dc.commit_sharded(OracleOps(), dist, local, n, n * b, c)
        ext = O.extend_pol(full, n * b)
        R = n * b // world
        ok_rows = np.array_equal(rows.numpy().view(np.uint64).reshape(R, c), ext[rank * R:(rank + 1) * R])
        ok_tree = True
        if rank == 0:
            ok_tree = np.array_equal(tree.numpy().view(np.uint64), O.merkletree(ext, c, n * b))
        q.put((rank, ok_rows, ok_tree))
    finally:
        dist.destroy_process_group()


def _worker_pipe(rank, world, port, n, b, c, rounds, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        O = Oracle()
        full = O.fill(n * c, SEED, 0).reshape(n, c)
        K, plan = dc.piece_plan(c, world, rounds)
        pieces = []
        for k in range(K):
            c0, w = plan[k][rank]
            m = np.ascontiguousarray(full[:, c0:c0 + w])
            pieces.append(torch.from_numpy(m.view(np.int64).reshape(-1).copy()) if w else torch.zeros(1, dtype=torch.int64))
        rows, tree = dc.commit_sharded_pipelined(OracleOps(), dist, pieces, n, n * b, c, rounds)
        ext = O.extend_pol(full, n * b)
        R = n * b // world
        ok_rows = np.array_equal(rows.numpy().view(np.uint64).reshape(R, c), ext[rank * R:(rank + 1) * R])
        ok_tree = True
        if rank == 0:
            ok_tree = np.array_equal(tree.numpy().view(np.uint64), O.merkletree(ext, c, n * b))
        q.put((rank, ok_rows, ok_tree))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n,b,c,rounds", [(2, 64, 2, 45, 2), (4, 32, 2, 100, 3), (2, 32, 2, 21, None), (4, 16, 4, 9, None),
                                                  (2, 32, 2, 203, None)])  # 203 columns on 2 ranks: ramped rounds 16, 32, then the rest
def test_commit_sharded_pipelined_gloo(world, n, b, c, rounds):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker_pipe, args=(r, world, port, n, b, c, rounds, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in range(world)]
    for p in procs:
        p.join(60)
    for rank, ok_rows, ok_tree in res:
        assert ok_rows, "row shard of rank %d differs" % rank
        assert ok_tree, "stitched tree differs"


def test_piece_plan():
    """zkgpu_dist_plan (csrc/dist.cu make_plan; callable without a device): consecutive pieces covering the row, every
    round but the last a whole number of sponge blocks, columns dealt out evenly; host input: uniform rounds of at least
    16 columns per rank, at most 14 of them."""
    for world in (1, 2, 4, 8):
        for host in (True, False):
            K, plan = dc.piece_plan(665, world, None, host)
            flat = [p for r in plan for p in r]
            assert flat[0][0] == 0 and flat[-1][0] + flat[-1][1] == 665
            assert all(a[0] + a[1] == b[0] for a, b in zip(flat[:-1], flat[1:]))
            assert all(sum(w for _, w in r) % 8 == 0 for r in plan[:-1])
            per_rank = [sum(plan[k][g][1] for k in range(K)) for g in range(world)]
            assert max(per_rank) - min(per_rank) <= 1
            if host:
                assert K <= 14 and all(w >= 16 for r in plan[:-1] for _, w in r)
            else:
                assert K <= 2
    K, plan = dc.piece_plan(665, 8)
    assert K == 5 and [sum(w for _, w in r) for r in plan] == [136, 136, 136, 136, 121]
    assert dc.piece_plan(665, 8, 3)[1][0][0] == (0, 28)  # explicit rounds (Python model only): equal rounds
    assert dc.piece_plan(9, 4) == (1, [[(0, 3), (3, 2), (5, 2), (7, 2)]])
    assert dc.piece_plan(0, 2) == (1, [[(0, 0), (0, 0)]])


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("world,n,b,c", [(2, 64, 2, 21), (4, 32, 4, 9), (2, 16, 2, 3)])
def test_commit_sharded_gloo(world, n, b, c):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n, b, c, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in range(world)]
    for p in procs:
        p.join(60)
    for rank, ok_rows, ok_tree in res:
        assert ok_rows, "row shard of rank %d differs" % rank
        assert ok_tree, "stitched tree differs"


def test_column_cuts_and_levels():
    assert dc.column_cuts(665, 8) == [0, 88, 176, 264, 352, 440, 528, 616, 665]
    assert dc.column_cuts(665, 4) == [0, 168, 336, 504, 665]
    assert dc.column_cuts(665, 2) == [0, 336, 665]
    assert dc.column_cuts(665, 1) == [0, 665]
    assert dc.column_cuts(8, 4) == [0, 8, 8, 8, 8]
    lv = dc.subtree_level_offsets(8)
    assert lv == [(0, 8), (32, 4), (48, 2), (56, 1)]
