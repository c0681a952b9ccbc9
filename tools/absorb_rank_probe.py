#!/usr/bin/env python3
"""The sponge launches ONE rank of an N-GPU commit runs (csrc/dist.cu: NExt/N rows of the row shard, all 665 columns, one launch
per round of the plan), reproduced on one GPU so that ncu can capture their DRAM traffic (ncu on a multi-rank command is not
allowed on this pool).  python tools/absorb_rank_probe.py N"""
import sys, pathlib
sys.path.insert(0, str(pathlib.Path(__file__).resolve().parent.parent))
import torch
import zkevm_prover_b200 as zk
from zkevm_prover_b200 import dist_commit as dc

world = int(sys.argv[1])
zk.init(0)
zk.set_stream(torch.cuda.current_stream())
n_ext, c = 1 << 24, 665
R = n_ext // world
rows = torch.empty(R * c, dtype=torch.int64, device="cuda")
zk.fill(rows, R * c, 3, 0)
leaves = torch.empty(R * 4, dtype=torch.int64, device="cuda")
K, plan = dc.piece_plan(c, world, None, False)
for k in range(K):
    c0, w = plan[k][0][0], sum(x[1] for x in plan[k])
    zk._check(zk.lib().zkgpu_merkle_absorb(leaves.data_ptr(), rows.data_ptr() + 8 * c0, w, c, R, int(k == 0)))
zk.sync()
print("rank shard %d rows x %d cols, %d sponge launches" % (R, c, K))
