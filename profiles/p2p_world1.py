"""One GPU: the ordinary single-vector kernel against the row-sharded variant with a group of one rank
(same rows, same work; the difference is the price of the exchange phases alone)."""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ["CBP_STAMPS_X"] = "1"
from cbc_b200 import capi  # noqa: E402
import bench  # noqa: E402

w = bench.make_workload(sys.argv[1] if len(sys.argv) > 1 else "c3", 0, 1.0)
inst = w["inst"]
ctx = capi.Context(0)
mat = capi.Matrix(ctx, inst)
mat.p2p_attach(0, 1, [mat.p2p_export(1)])
lb0 = torch.tensor(inst.lb, device="cuda")
ub0 = torch.tensor(inst.ub, device="cuda")
lb, ub = lb0.clone(), ub0.clone()


REPS = int(os.environ.get("P2P_REPS", "20"))


def run(fn, reps=REPS):
    for _ in range(min(3, reps)):
        lb.copy_(lb0); ub.copy_(ub0); fn()
    torch.cuda.synchronize()
    ms = 0.0
    for _ in range(reps):
        lb.copy_(lb0); ub.copy_(ub0)
        r = fn()
        ms += ctx.stats()["lastRoundKernelMs"]
    return ms / reps, r["rounds"]


for name, fn in (("plain", lambda: mat.propagate_device(lb.data_ptr(), ub.data_ptr())),
                 ("group of one", lambda: mat.p2p_propagate(lb.data_ptr(), ub.data_ptr(), 0, device_ptrs=True))):
    ms, rounds = run(fn)
    st = mat.round_stamps(64).astype(np.int64)
    print(f"{name}: kernel {ms:.3f} ms, {rounds} rounds")
    for i in [0, 1, 2, 3, 10, 20, 21]:
        if i >= rounds:
            break
        rowsDone, a1, settleDone, own = st[i]
        start = st[i - 1][2] if i > 0 else 0
        print(f"  round {i}: A1 done +{(a1-start)/1e3:6.1f}  CTA0 rows done +{(own-start)/1e3:6.1f}  rows barrier passed "
              f"+{(rowsDone-start)/1e3:6.1f}  settle+barrier +{(settleDone-rowsDone)/1e3:5.1f}")
