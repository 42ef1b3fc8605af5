"""First batch call on a freshly uploaded matrix against the calls after it: python profiles/first_batch.py [scale]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from cbc_b200 import capi  # noqa: E402

ctx = capi.Context(0)


def root(inst):
    mat = capi.Matrix(ctx, inst)
    r, lb, ub = mat.propagate(inst.lb, inst.ub)
    mat.close()
    return lb, ub


for scale in (1.0, 0.03125):
    w = bench.make_workload("c2", 0, scale, root)
    s, v, b = w["deltas"]
    for rep in range(2):
        t0 = time.perf_counter()
        mat = capi.Matrix(ctx, w["inst"])
        t1 = time.perf_counter()
        ts = []
        for _ in range(4):
            ta = time.perf_counter()
            mat.propagate_batch_delta(w["lb"], w["ub"], s, v, b)
            ts.append(time.perf_counter() - ta)
        mat.close()
        print(f"c2 B={len(s) - 1}: upload {(t1 - t0) * 1e3:.2f} ms, batch calls " + " ".join(f"{x * 1e3:.2f}" for x in ts) + " ms")
