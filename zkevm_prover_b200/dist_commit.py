"""Multi-GPU commit stage, host side.

The product path is in the library (csrc/dist.cu behind include/zkgpu.h): `CDist` below is the thin driver a
process-per-GPU launcher (torchrun) needs — it hands NCCL's unique id to the other ranks through torch.distributed and
then calls zkgpu_dist_init / zkgpu_dist_commit; single-process hosts call zkgpu_init_multi and the ordinary
zkgpu_extend_and_merkelize instead.

`commit_sharded` / `commit_sharded_pipelined` are the same pipeline spelled in Python over injected compute steps
(`ops`): the host-logic model the CPU test-suite runs over gloo with the oracle standing in for the device steps, and a
second implementation (torch.distributed all-to-all instead of peer stores) the GPU tests compare the library against.

  1. the trace is column-sharded (NTT columns are independent, ntt_goldilocks.cpp:103-110): rank g owns columns
     [cuts[g], cuts[g+1]) as an N x C_g row-major matrix and extends it to NExt x C_g (zkgpu_lde_columns);
  2. ONE all-to-all regroups the extended trace into row shards: rank g sends rows [h*R, (h+1)*R) of its shard
     (a contiguous block, no packing) to rank h, R = NExt / G;
  3. rank h interleaves the G received blocks into R x C row-major rows (the layout leaf hashing and later
     getGroupProof need), hashes its R leaves (poseidon_goldilocks.cpp:277-281) and builds its subtree;
  4. the G subtrees are gathered on rank 0 and stitched into the reference's level-major node array
     (poseidon_goldilocks.cpp:283-300); rank 0 hashes the top log2(G) levels.

The compute steps are injected (`ops`) so that the same host logic runs over libzkgpu.so on GPUs (GpuOps) and, in the
CPU test-suite, over a stand-in with the gloo backend."""
import numpy as np


def column_cuts(ncols, world):
    """Shard boundaries: multiples of 8 columns (one sponge block) where possible, remainder on the last rank.
    665 columns -> 336+329 (G=2), 168*3+161 (G=4), 88*7+49 (G=8)."""
    per = -(-ncols // world)
    per = -(-per // 8) * 8
    cuts = [min(g * per, ncols) for g in range(world)] + [ncols]
    return cuts


def subtree_level_offsets(nleaves):
    """Word offsets of every level of a level-major tree over nleaves leaves: [(offset, nodes)], leaves first."""
    out, off, n = [], 0, nleaves
    while n >= 1:
        out.append((off, n))
        off += n * 4
        if n == 1:
            break
        n //= 2
    return out


class GpuOps:
    """Compute steps on the local GPU through the C ABI; buffers are torch int64 CUDA tensors."""

    def __init__(self):
        import torch
        import zkevm_prover_b200 as zk
        self.torch, self.zk = torch, zk
        # the library must enqueue on the stream torch orders its collectives, copies and allocator frees against
        zk.set_stream(torch.cuda.current_stream())

    def empty(self, nwords):
        return self.torch.empty(nwords, dtype=self.torch.int64, device="cuda")

    def lde(self, out, inp, n_ext, n, ncols):
        self.zk._check(self.zk.lib().zkgpu_lde_columns(out.data_ptr(), ncols, inp.data_ptr(), ncols, n_ext, n, ncols))

    def copy_block(self, dst, dst_off, dst_ld, src, src_off, src_ld, nrows, ncols):
        self.zk._check(self.zk.lib().zkgpu_copy_block(dst.data_ptr() + 8 * dst_off, dst_ld, src.data_ptr() + 8 * src_off,
                                                      src_ld, nrows, ncols))

    def leaves(self, tree, rows, ncols, nrows):
        self.zk._check(self.zk.lib().zkgpu_merkle_leaves(tree.data_ptr(), rows.data_ptr(), ncols, ncols, nrows))

    def levels(self, tree, nleaves, word_off=0):
        self.zk._check(self.zk.lib().zkgpu_merkle_levels(tree.data_ptr() + 8 * word_off, nleaves))

    def absorb(self, leaves, block, block_off, width, ld, nrows, first, copy_to=None, copy_off=0, copy_ld=0):
        """Row sponge over one column piece; with copy_to the piece is also scattered into the row-major matrix."""
        if copy_to is None:
            self.zk._check(self.zk.lib().zkgpu_merkle_absorb(leaves.data_ptr(), block.data_ptr() + 8 * block_off, width, ld, nrows, int(first)))
        else:
            self.zk._check(self.zk.lib().zkgpu_merkle_absorb_copy(leaves.data_ptr(), block.data_ptr() + 8 * block_off, width, ld, nrows,
                                                                   int(first), copy_to.data_ptr() + 8 * copy_off, copy_ld))

    # ---- host -> device staging of input pieces on a copy stream (two slots), used by commit_sharded_pipelined
    def stage(self, src, slot):
        """Start bringing piece `src` (a pinned host tensor, or a device tensor which is used as is) to the device."""
        if src.is_cuda:
            return (src, None, slot)
        torch = self.torch
        if not hasattr(self, "_copy_stream"):
            self._copy_stream, self._bufs, self._free = torch.cuda.Stream(), [None, None], [None, None]
        if self._bufs[slot] is None or self._bufs[slot].numel() < src.numel():
            self._bufs[slot] = torch.empty(src.numel(), dtype=torch.int64, device="cuda")
        dst = self._bufs[slot][:src.numel()]
        with torch.cuda.stream(self._copy_stream):
            if self._free[slot] is not None:
                self._copy_stream.wait_event(self._free[slot])  # the LDE that last read this slot
            dst.copy_(src, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(self._copy_stream)
        return (dst, ev, slot)

    def staged_wait(self, st):
        if st[1] is not None:
            self.torch.cuda.current_stream().wait_event(st[1])
        return st[0]

    def staged_release(self, st):
        if st[1] is not None:
            ev = self.torch.cuda.Event()
            ev.record(self.torch.cuda.current_stream())
            self._free[st[2]] = ev


def exchange(dist, send, recv, send_splits, recv_splits, group=None):
    """Column shards -> row shards.  NCCL: one all_to_all_single; other backends: pairwise isend/irecv."""
    if dist.get_backend(group) == "nccl":
        dist.all_to_all_single(recv, send, output_split_sizes=recv_splits, input_split_sizes=send_splits, group=group)
        return
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    so = np.concatenate([[0], np.cumsum(send_splits)])
    ro = np.concatenate([[0], np.cumsum(recv_splits)])
    reqs = []
    for h in range(world):
        if h == rank:
            recv[ro[h]:ro[h + 1]].copy_(send[so[h]:so[h + 1]])
        else:
            reqs.append(dist.isend(send[so[h]:so[h + 1]].contiguous(), h, group=group))
            reqs.append(dist.irecv(recv[ro[h]:ro[h + 1]], h, group=group))
    for r in reqs:
        r.wait()


class _Done:
    def wait(self):
        return True


def exchange_async(dist, send, recv, send_splits, recv_splits, group=None):
    """exchange() that returns a handle: with NCCL the all-to-all runs on the communicator's stream and handle.wait()
    makes the current stream wait for it (the host does not block); other backends complete before returning."""
    if dist.get_backend(group) == "nccl":
        return dist.all_to_all_single(recv, send, output_split_sizes=recv_splits, input_split_sizes=send_splits, group=group,
                                      async_op=True)
    exchange(dist, send, recv, send_splits, recv_splits, group)
    return _Done()


def piece_plan(ncols, world, rounds=None, host_input=True):
    """Column ownership of the pipelined commit (zkgpu_dist_plan, csrc/dist.cu make_plan): the row is cut into
    rounds x world consecutive pieces, piece (k, g) belonging to rank g in round k.  Every round but the last covers a
    whole number of 8-column sponge blocks; inside a round the columns are dealt out evenly (the extra columns rotate
    over the ranks from round to round).  Returns (rounds, plan) with plan[k][g] = (first column, width).

    host_input=True: uniform small rounds — two sponge blocks (16 columns) per rank and round, wider only when that would take
    more than 14 rounds — so that little is left to compute after the last host copy lands; host_input=False: one round
    (peer stores) or two (staged exchange).  The library decides (csrc/dist.cu make_plan); this only asks it.
    An explicit `rounds` gives equal rounds (test hook, Python model only)."""
    if rounds is None:
        import ctypes
        import zkevm_prover_b200 as zk
        r = ctypes.c_int(0)
        c0 = (ctypes.c_uint64 * (16 * world))()
        w = (ctypes.c_uint64 * (16 * world))()
        zk._check(zk.lib().zkgpu_dist_plan(ncols, world, int(host_input), ctypes.byref(r), c0, w))
        return r.value, [[(int(c0[k * world + g]), int(w[k * world + g])) for g in range(world)] for k in range(r.value)]
    blocks = -(-ncols // 8)
    per_round = [blocks // rounds + (1 if k < blocks % rounds else 0) for k in range(rounds)]
    plan, c, rot = [], 0, 0
    for nb in per_round:
        cols = min(nb * 8, ncols - c)
        base, extra = divmod(cols, world)
        row = []
        for g in range(world):
            wg = base + (1 if (g - rot) % world < extra else 0)
            row.append((c, wg))
            c += wg
        plan.append(row)
        rot = (rot + extra) % world
    return rounds, plan


def commit_sharded_pipelined(ops, dist, pieces_in, n, n_ext, ncols, rounds=None, group=None, timers=None):
    """The commit stage over all ranks of `group` as a pipeline over column pieces (piece_plan):

        [host->device of piece k+1]  ||  LDE of piece k  ->  all-to-all of piece k  ||  sponge absorption of round k-1

    Round k's all-to-all hands every rank the columns plan[k][0..G-1] of its rows, which are consecutive in the row, so
    the row sponge can absorb them as soon as they land while the next round is being extended.
    pieces_in[k]: this rank's N x width(k, rank) piece of the trace (device tensor, or pinned host tensor which is staged
    on a copy stream).  Returns (rows, tree) like commit_sharded."""
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    rounds, plan = piece_plan(ncols, world, rounds)
    R = n_ext // world
    assert R * world == n_ext and R >= 1 and (R & (R - 1)) == 0, "NExt must be a power of two >= world size"
    assert ncols > 4, "rows of <= 4 words are not hashed: use commit_sharded"

    def mark(name):
        if timers is not None:
            timers(name)

    mark("start")
    sub_words = (2 * R - 1) * 4
    if rank == 0:
        tree = ops.empty((2 * n_ext - 1) * 4)
        sub = ops.empty(sub_words)
    else:
        tree = sub = ops.empty(sub_words)
    rows = ops.empty(R * ncols)
    wmax_mine = max(plan[k][rank][1] for k in range(rounds))
    wmax_round = max(sum(w for _, w in plan[k]) for k in range(rounds))
    ext_buf = [ops.empty(max(n_ext * wmax_mine, 1)) for _ in range(min(2, rounds))]
    recv_buf = [ops.empty(max(R * wmax_round, 1)) for _ in range(min(2, rounds))]

    def finish(k, recv):
        # unpack the G received blocks into the row-major row shard, then absorb the round's columns from there
        off = 0
        for g in range(world):
            c0, w = plan[k][g]
            if w:
                ops.copy_block(rows, c0, ncols, recv, off, w, R, w)
            off += R * w
        c0, wr = plan[k][0][0], sum(w for _, w in plan[k])
        if wr:
            ops.absorb(sub, rows, c0, wr, ncols, R, k == 0)
        mark("leaves")

    pending = None
    staged = ops.stage(pieces_in[0], 0)
    for k in range(rounds):
        w = plan[k][rank][1]
        nxt = ops.stage(pieces_in[k + 1], (k + 1) & 1) if k + 1 < rounds else None
        d_in = ops.staged_wait(staged)
        mark("h2d_wait")
        ext, recv = ext_buf[k & 1], recv_buf[k & 1]
        if w:
            ops.lde(ext, d_in, n_ext, n, w)
        ops.staged_release(staged)
        mark("lde")
        work = exchange_async(dist, ext[:n_ext * w], recv[:R * sum(x for _, x in plan[k])], [R * w] * world,
                              [R * x for _, x in plan[k]], group)
        if pending is not None:
            pending[0].wait()
            mark("exchange")
            finish(pending[1], pending[2])
        pending = (work, k, recv)
        staged = nxt
    pending[0].wait()
    mark("exchange")
    finish(pending[1], pending[2])
    ops.levels(sub, R)
    tree = _gather_and_stitch(ops, dist, sub, tree, R, n_ext, group)
    mark("nodes")
    return rows, tree


def _gather_and_stitch(ops, dist, sub, tree, R, n_ext, group=None):
    """Subtrees -> rank 0, stitched level by level into the reference's level-major array; rank 0 hashes the top."""
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    sub_words = (2 * R - 1) * 4
    if world == 1:
        tree[:sub_words].copy_(sub)
        return tree
    if rank == 0:
        gathered = ops.empty(sub_words * world)
        parts = [gathered[g * sub_words:(g + 1) * sub_words] for g in range(world)]
        dist.gather(sub, parts, dst=0, group=group)
        glob = subtree_level_offsets(n_ext)
        for lvl, (soff, cnt) in enumerate(subtree_level_offsets(R)):
            goff = glob[lvl][0]
            for g in range(world):
                tree[goff + g * cnt * 4: goff + (g + 1) * cnt * 4].copy_(parts[g][soff: soff + cnt * 4])
        top_off = glob[len(subtree_level_offsets(R)) - 1][0]
        ops.levels(tree, world, word_off=top_off)
    else:
        dist.gather(sub, None, dst=0, group=group)
    return tree


def commit_sharded(ops, dist, local_in, n, n_ext, ncols, group=None, timers=None):
    """Run the commit stage over all ranks of `group`.  local_in: this rank's N x C_g column shard.
    Returns (rows, tree): rows = this rank's R x C row shard of the extended trace; tree = the full node array
    ((2*NExt-1)*4 words) on rank 0, this rank's subtree elsewhere."""
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    cuts = column_cuts(ncols, world)
    cg = cuts[rank + 1] - cuts[rank]
    R = n_ext // world
    assert R * world == n_ext and R >= 1 and (R & (R - 1)) == 0, "NExt must be a power of two >= world size"

    def mark(name):
        if timers is not None:
            timers(name)

    mark("start")
    ext_local = ops.empty(max(n_ext * cg, 1))
    if cg:
        ops.lde(ext_local, local_in, n_ext, n, cg)
    mark("lde")
    recv = ops.empty(R * ncols)
    send_splits = [R * cg] * world
    recv_splits = [R * (cuts[g + 1] - cuts[g]) for g in range(world)]
    exchange(dist, ext_local[:n_ext * cg], recv, send_splits, recv_splits, group)
    mark("exchange")
    del ext_local
    rows = ops.empty(R * ncols)
    off = 0
    for g in range(world):
        w = cuts[g + 1] - cuts[g]
        if w:
            ops.copy_block(rows, cuts[g], ncols, recv, off, w, R, w)
        off += R * w
    del recv
    sub_words = (2 * R - 1) * 4
    if rank == 0:
        tree = ops.empty((2 * n_ext - 1) * 4)
        sub = ops.empty(sub_words)
    else:
        tree = sub = ops.empty(sub_words)
    ops.leaves(sub, rows, ncols, R)
    mark("leaves")
    ops.levels(sub, R)
    tree = _gather_and_stitch(ops, dist, sub, tree, R, n_ext, group)
    mark("nodes")
    return rows, tree


class CDist:
    """Driver of the library's process-per-GPU form (include/zkgpu.h: zkgpu_dist_init / zkgpu_dist_commit).

    torch.distributed is used for one thing: rank 0's NCCL unique id travels to the other ranks.  Everything else —
    the column plan, the LDE with its peer stores, the barriers, the sponge, the stitch — happens inside libzkgpu.so."""

    def __init__(self, dist, device, group=None):
        import ctypes
        import torch
        import zkevm_prover_b200 as zk
        self.zk, self.torch, self.ct = zk, torch, ctypes
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        zk.init(device)
        uid = ctypes.create_string_buffer(128)
        if self.rank == 0:
            zk._check(zk.lib().zkgpu_nccl_unique_id(uid))
        box = [bytes(uid.raw)]
        dist.broadcast_object_list(box, src=0, group=group)
        uid = ctypes.create_string_buffer(box[0], 128)
        zk._check(zk.lib().zkgpu_dist_init(uid, self.rank, self.world))

    def plan(self, ncols, host_input=False):
        return piece_plan(ncols, self.world, None, host_input)

    def commit(self, pieces, n, n_ext, ncols, tree=None, host_input=False, gather=True, ld=None):
        """pieces[k]: this rank's N x w(k, rank) piece of round k (torch tensors: CUDA, or pinned host with host_input).
        tree: rank 0's destination tensor / array when gather is set."""
        ct, zk = self.ct, self.zk
        K = len(pieces)
        ptrs = (ct.c_void_p * K)(*[zk._ptr(p) for p in pieces])
        rounds, plan = self.plan(ncols, host_input)
        assert rounds == K, "pieces do not match zkgpu_dist_plan"
        lds = (ct.c_uint64 * K)(*[(ld[k] if ld else plan[k][self.rank][1]) for k in range(K)])
        zk._check(zk.lib().zkgpu_dist_commit(None, zk._ptr(tree), ptrs, lds, int(host_input), int(gather), n_ext, n, ncols))

    def rows(self):
        """(address of this rank's R x ncols row shard, R, ncols, address of its subtree) — library-owned device memory."""
        ct = self.ct
        p, sub, r, c = ct.c_void_p(), ct.c_void_p(), ct.c_uint64(), ct.c_uint64()
        self.zk._check(self.zk.lib().zkgpu_dist_rows(ct.byref(p), ct.byref(r), ct.byref(c), ct.byref(sub)))
        return p.value, r.value, c.value, sub.value

    def close(self):
        """Leave the job (zkgpu_dist_finalize); call it before torch.distributed is torn down."""
        self.zk._check(self.zk.lib().zkgpu_dist_finalize())

    def exchange(self):
        """'stores' (peer stores fused into the LDE's last pass), 'dma' (copy engines) or 'nccl' (ncclSend/ncclRecv)."""
        ct = self.ct
        a, b, c = ct.c_int(), ct.c_int(), ct.c_int()
        self.zk._check(self.zk.lib().zkgpu_dist_info(ct.byref(a), ct.byref(b), ct.byref(c)))
        return {1: "stores", 2: "dma"}.get(c.value, "nccl")

    def peer_stores(self):
        return self.exchange() == "stores"
