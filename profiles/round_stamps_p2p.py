"""Per-phase timestamps of the row-sharded fused kernel (run under torchrun, one rank per GPU):
for a few rounds, when CTA 0 of each rank finished A1, its own rows, passed the rows barrier, the settle barrier."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["CBP_STAMPS_X"] = "1"
from cbc_b200 import capi, multigpu as MG  # noqa: E402
import bench  # noqa: E402

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dist.init_process_group("cpu:gloo,cuda:nccl", rank=rank, world_size=world)
w = bench.make_workload(sys.argv[1] if len(sys.argv) > 1 else "c3", 0, 1.0)
inst = w["inst"]
ctx = capi.Context(rank)
sh = MG.PeerShardedMatrix(ctx, inst)
for _ in range(4):
    r, l, u = sh.propagate(inst.lb, inst.ub)
st = sh.mat.round_stamps(64).astype(np.int64)
for rk in range(world):
    dist.barrier()
    if rk != rank:
        continue
    print(f"rank {rank}: rounds {r['rounds']}", flush=True)
    for i in [0, 1, 2, 10, 11, 20, 21]:
        if i >= r["rounds"]:
            break
        rowsDone, a1, settleDone, own = st[i]
        start = st[i - 1][2] if i > 0 else 0
        print(f"  round {i}: A1 done +{(a1-start)/1e3:6.1f}  CTA0 rows done +{(own-start)/1e3:6.1f}  rows barrier passed "
              f"+{(rowsDone-start)/1e3:6.1f}  settle+barrier +{(settleDone-rowsDone)/1e3:5.1f}", flush=True)
dist.barrier()
dist.destroy_process_group()
