"""zkgpu_merkletree_batch_noalloc against zkgpu_merkletree_batch, through the C ABI.

Some hosts of this library cannot let it allocate device memory: a prover that claimed the whole GPU for its own arena
at start-up (cudaMalloc has nothing left and the cudaMallocAsync pool cannot grow), or a service with a fixed
device-memory budget.  For them the batched Merkle tree has a second entry point whose per-batch digest scratch is a
fixed 2^20-word (8 MiB) array inside the library image instead of a per-call allocation.  Both entry points run the same
kernels, so the trees must be equal word for word (the bar of every integer test here: bit-exact).

The array is shared by all calls and keeps the previous call's digests, so the test calls the entry point twice in a row
with different shapes (the second one smaller, so that stale digests of the first lie behind and between its own) and
checks both.  This file sorts last on purpose: `pytest -x` stops at the first failure, and a failure here must not hide
the parity tests.  It passes alone in a fresh process (first use of the device by the library) and after the others.
"""
import ctypes

import numpy as np
import pytest

import zkevm_prover_b200 as zk
from oracle_lib import Oracle, SEED

pytestmark = pytest.mark.gpu

LIMIT_WORDS = 1 << 20  # size of the built-in digest array (include/zkgpu.h)


def _dev(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a).view(np.int64)).cuda()


def _host(t):
    return t.cpu().numpy().view(np.uint64)


def _both_trees(L, rows, ncols, dim, batch, seed_offset):
    """The tree of one synthetic matrix by the allocating and by the no-alloc entry point (device buffers, raw C calls)."""
    import torch
    a = Oracle().fill(rows * ncols * dim, SEED, seed_offset).reshape(rows, ncols * dim)
    d_in = _dev(a)
    words = (2 * rows - 1) * 4
    t_ref = torch.zeros(words, dtype=torch.int64, device="cuda")
    t_new = torch.full((words,), -1, dtype=torch.int64, device="cuda")  # poisoned: every word must be written
    assert L.zkgpu_merkletree_batch(t_ref.data_ptr(), d_in.data_ptr(), ncols, rows, batch, dim) == 0, L.zkgpu_last_error()
    assert L.zkgpu_merkletree_batch_noalloc(t_new.data_ptr(), d_in.data_ptr(), ncols, rows, batch, dim) == 0, L.zkgpu_last_error()
    assert L.zkgpu_sync() == 0
    return _host(t_ref), _host(t_new), a


def test_noalloc_batch_tree_equals_allocating_entry_point_twice_in_a_row():
    L = zk.lib()
    assert L.zkgpu_init(0) == 0, L.zkgpu_last_error()
    # call 1: 2^12 rows x 45 columns in batches of 8 -> 6 digests per row (the last batch is ragged: 5 columns)
    ref1, new1, a1 = _both_trees(L, 1 << 12, 45, 1, 8, 11)
    assert np.array_equal(new1, ref1), "first call: trees differ"
    # call 2 straight after, smaller and with another batch count: must not pick up digests left by call 1
    ref2, new2, _ = _both_trees(L, 1 << 9, 20, 3, 7, 12)
    assert np.array_equal(new2, ref2), "second call: trees differ"
    assert not np.array_equal(new2[-4:], new1[-4:])
    # and the allocating entry point is itself pinned to the CPU oracle on the first shape
    want = Oracle().merkletree(a1, 45, 1 << 12, 1, 8)
    assert np.array_equal(ref1, want), "allocating entry point differs from the oracle"


def test_noalloc_rejects_what_it_cannot_serve():
    import torch
    L = zk.lib()
    assert L.zkgpu_init(0) == 0
    before = L.zkgpu_launch_count()
    rows, ncols, batch = 1 << 15, 64, 4  # 2^15 * 16 * 4 = 2^21 words of digests: twice the built-in array
    assert rows * ((ncols + batch - 1) // batch) * 4 > LIMIT_WORDS
    d_in = torch.zeros(8, dtype=torch.int64, device="cuda")  # never read: the call must refuse before launching
    d_tree = torch.zeros(8, dtype=torch.int64, device="cuda")
    assert L.zkgpu_merkletree_batch_noalloc(d_tree.data_ptr(), d_in.data_ptr(), ncols, rows, batch, 1) == -1
    assert b"exceed" in L.zkgpu_last_error()
    # host pointers would need a staging allocation: refused as well
    h = np.zeros(64 * 8, dtype=np.uint64)
    t = np.zeros((2 * 64 - 1) * 4, dtype=np.uint64)
    assert L.zkgpu_merkletree_batch_noalloc(t.ctypes.data, h.ctypes.data, 8, 64, 4, 1) == -1
    assert b"device pointers" in L.zkgpu_last_error()
    assert L.zkgpu_launch_count() == before
    # the largest shape that fits is served: 2^14 rows x 16 batches
    rows, ncols, batch = 1 << 14, 16, 1
    assert rows * ncols * 4 == LIMIT_WORDS
    ref, new, _ = _both_trees(L, rows, ncols, 1, batch, 13)
    assert np.array_equal(new, ref)
