import sys, os, numpy as np
sys.path.insert(0, '/root/repo')
os.environ["CBP_STAMPS_X"]="1"; os.environ["CBP_STAMPS_TOP"]="1"
from cbc_b200 import capi
import bench
w = bench.make_workload("c3"); inst = w["inst"]
ctx = capi.Context(0); mat = capi.Matrix(ctx, inst)
for _ in range(3): r, l, u = mat.propagate(inst.lb, inst.ub)
st = mat.round_stamps(64).astype(np.int64)
for i in [1,2,10,11,20,21,25]:
    top, a1, settleDone, own = st[i]
    prev = st[i-1][2]
    print(f"round {i}: prev settle barrier->loop top {(top-prev)/1e3:5.1f}  top->A1 done {(a1-top)/1e3:5.1f}  A1->CTA0 rows done {(own-a1)/1e3:6.1f}  ->settle barrier {(settleDone-own)/1e3:5.1f}")
