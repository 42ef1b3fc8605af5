import sys, os, numpy as np
sys.path.insert(0, '/root/repo')
from cbc_b200 import capi
import bench
w = bench.make_workload(sys.argv[1] if len(sys.argv) > 1 else "c5"); inst = w["inst"]
ctx = capi.Context(0); mat = capi.Matrix(ctx, inst)
for _ in range(3): r, l, u = mat.propagate(inst.lb, inst.ub)
st = mat.round_stamps(64).astype(np.int64)
print("rounds", r["rounds"], "changed", r["roundChanged"])
prev = 0
for i in range(min(r["rounds"], 6)):
    a, b, c, d = st[i]
    print(f"round {i}: rows phase {(a-prev)/1e3:6.1f}  long rows 2nd pass {(b-a)/1e3:6.1f}  settle {(c-b)/1e3:6.1f}  decision {(d-c)/1e3:5.1f}")
    prev = d
