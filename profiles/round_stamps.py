import sys, numpy as np
import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cbc_b200 import capi, instances as I
import bench
w = bench.make_workload(sys.argv[1] if len(sys.argv) > 1 else "c3")
inst = w["inst"]
ctx = capi.Context(0); mat = capi.Matrix(ctx, inst)
for _ in range(3):
    r, l, u = mat.propagate(inst.lb, inst.ub)
st = mat.round_stamps(64).astype(np.int64)
print(r)
prev = 0
for i in range(min(r["rounds"], 12)):
    a, b, c, d = st[i]
    print(f"round {i}: rows {(a-prev)/1e3:7.1f} us  long {(b-a)/1e3:6.1f}  settle {(c-b)/1e3:6.1f}  decide {(d-c)/1e3:5.1f}  changed {r['roundChanged'][i] if i < len(r['roundChanged']) else '-'}")
    prev = d
print("last stamp us", st[r["rounds"]-1][3]/1e3 if r["rounds"]<=64 else st[63][3]/1e3, ctx.stats())
