"""Host-side phases of one end-to-end batch call (load / run / fetch), averaged: python profiles/e2e_phases.py [c2|c4|c3b]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from cbc_b200 import capi  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "c2"
ctx = capi.Context(0)


def root(inst):
    mat = capi.Matrix(ctx, inst)
    r, lb, ub = mat.propagate(inst.lb, inst.ub)
    mat.close()
    return lb, ub


w = bench.make_workload(name, 0, float(sys.argv[2]) if len(sys.argv) > 2 else 1.0, root)
inst = w["inst"]
s, v, b = w["deltas"]
mat = capi.Matrix(ctx, inst)
for _ in range(5):
    mat.propagate_batch_delta(w["lb"], w["ub"], s, v, b)
N = 30
tl = tr = tf = 0.0
for _ in range(N):
    t0 = time.perf_counter()
    mat.batch_load_delta(w["lb"], w["ub"], s, v, b)
    t1 = time.perf_counter()
    mat.batch_run()
    t2 = time.perf_counter()
    r = mat.batch_fetch()
    t3 = time.perf_counter()
    tl += t1 - t0
    tr += t2 - t1
    tf += t3 - t2
print("%s B=%d %s: load %.3f ms  run %.3f ms  fetch %.3f ms  (updates %d)" % (name, len(s) - 1, "full-walk" if os.environ.get("CBP_NO_WORKLIST") else "worklist", tl / N * 1e3, tr / N * 1e3, tf / N * 1e3, r[2].size))
