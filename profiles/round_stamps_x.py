import sys, os, numpy as np
sys.path.insert(0, '/root/repo')
os.environ["CBP_STAMPS_X"]="1"
from cbc_b200 import capi
import bench
w = bench.make_workload("c3"); inst = w["inst"]
ctx = capi.Context(0); mat = capi.Matrix(ctx, inst)
for _ in range(3): r, l, u = mat.propagate(inst.lb, inst.ub)
st = mat.round_stamps(64).astype(np.int64)
prev=0
for i in [0,1,2,10,11,20,21]:
    rowsDone, a1, settleDone, own = st[i]
    start = st[i-1][2] if i>0 else 0   # previous settle done ~ round start (minus decide)
    print(f"round {i}: A1 done +{(a1-start)/1e3:6.1f}  CTA0 rows done +{(own-start)/1e3:6.1f}  all rows done(after barrier) +{(rowsDone-start)/1e3:6.1f}  settle +{(settleDone-rowsDone)/1e3:5.1f}")
