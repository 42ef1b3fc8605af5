import sys, time, numpy as np
sys.path.insert(0, '/root/repo')
import bench
from cbc_b200 import capi
w = bench.make_workload("c2", 0, 1.0); inst = w["inst"]; s, v, b = w["deltas"]
ctx = capi.Context(0); mat = capi.Matrix(ctx, inst)
for _ in range(5): mat.propagate_batch_delta(w["lb"], w["ub"], s, v, b)
N = 50
tl = tr = tf = 0.0
for _ in range(N):
    t0 = time.perf_counter(); mat.batch_load_delta(w["lb"], w["ub"], s, v, b)
    t1 = time.perf_counter(); mat.batch_run()
    t2 = time.perf_counter(); r = mat.batch_fetch()
    t3 = time.perf_counter()
    tl += t1 - t0; tr += t2 - t1; tf += t3 - t2
print("load %.3f ms  run %.3f ms  fetch %.3f ms" % (tl / N * 1e3, tr / N * 1e3, tf / N * 1e3))
st = ctx.stats(); print(st)
