"""Per-round device timestamps of the single-vector kernel on one workload (default c3): rows / long rows / settle /
decision per round, the changed-column count beside them, and the kernel time (best of 5).
  CBP_LIB=<variant .so> CBP_EVENT_DIV=<d> python profiles/sv_rounds.py [workload] [brief]"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from cbc_b200 import capi  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "c3"
brief = len(sys.argv) > 2
w = bench.make_workload(name)
inst = w["inst"]
ctx = capi.Context(0)
mat = capi.Matrix(ctx, inst)
best = 1e9
for _ in range(6):
    r, l, u = mat.propagate(inst.lb, inst.ub)
    best = min(best, ctx.stats()["lastRoundKernelMs"])
st = mat.round_stamps(64).astype(np.int64)
print(f"{os.environ.get('CBP_LIB', 'default')} evdiv={os.environ.get('CBP_EVENT_DIV', '-')} {name}: kernel {best:.3f} ms, "
      f"rounds {r['rounds']}, changed {r['nChanged']}")
if not brief:
    prev = 0
    for i in range(min(r["rounds"], 64)):
        a, b, c, d = st[i]
        ch = r["roundChanged"][i] if i < len(r["roundChanged"]) else -1
        print(f"  round {i:2d}: rows {(a - prev) / 1e3:7.1f} us  long {(b - a) / 1e3:6.1f}  settle {(c - b) / 1e3:6.1f}  "
              f"decide {(d - c) / 1e3:5.1f}  changed {ch}")
        prev = d
