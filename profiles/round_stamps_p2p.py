"""Per-phase timestamps of the row-sharded fused kernel (run under torchrun, one rank per GPU), rank 0's view:
per round, when CTA 0 finished A1, its own rows, passed the rows barrier (all ranks' proposals in), the settle barrier."""
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
best = 1e9
for _ in range(6):
    r, l, u = sh.propagate(inst.lb, inst.ub)
    best = min(best, ctx.stats()["lastRoundKernelMs"])
st = sh.mat.round_stamps(64).astype(np.int64)
if rank == 0:
    print(f"world {world}: kernel {best:.3f} ms, rounds {r['rounds']}", flush=True)
    tot = np.zeros(4)
    for i in range(min(r["rounds"], 64)):
        rowsDone, a1, settleDone, own = st[i]
        start = st[i - 1][2] if i > 0 else 0
        seg = np.array([a1 - start, own - a1, rowsDone - own, settleDone - rowsDone]) / 1e3
        tot += seg
        print(f"  round {i:2d}: A1 {seg[0]:6.1f}  CTA0 rows {seg[1]:6.1f}  wait+push+barrier {seg[2]:6.1f}  settle+sync+decide {seg[3]:5.1f}"
              f"  changed {r['roundChanged'][i] if i < len(r['roundChanged']) else -1}", flush=True)
    print(f"  totals (us): A1 {tot[0]:.0f}  rows {tot[1]:.0f}  wait+push+barrier {tot[2]:.0f}  settle {tot[3]:.0f}", flush=True)
dist.barrier()
dist.destroy_process_group()
