"""Multi-GPU commit stage: one process per GPU, torch.distributed for the plumbing.

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
    if world == 1:
        tree[:sub_words].copy_(sub)
        mark("nodes")
        return rows, tree
    # gather the subtrees on rank 0 and stitch level by level
    if rank == 0:
        gathered = ops.empty(sub_words * world)
        parts = [gathered[g * sub_words:(g + 1) * sub_words] for g in range(world)]
        dist.gather(sub, parts, dst=0, group=group)
        glob = subtree_level_offsets(n_ext)
        for lvl, (soff, cnt) in enumerate(subtree_level_offsets(R)):
            goff = glob[lvl][0]
            for g in range(world):
                tree[goff + g * cnt * 4: goff + (g + 1) * cnt * 4].copy_(parts[g][soff: soff + cnt * 4])
        # sub-roots now sit contiguously at the level with `world` nodes: finish the top on this GPU
        top_off = glob[len(subtree_level_offsets(R)) - 1][0]
        ops.levels(tree, world, word_off=top_off)
    else:
        dist.gather(sub, None, dst=0, group=group)
    mark("nodes")
    return rows, tree
