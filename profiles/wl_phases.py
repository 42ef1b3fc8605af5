"""Phase boundaries of one worklist batch run (device timestamps, CBP_WL_STAMPS=1): python profiles/wl_phases.py c4"""
import os
import sys

os.environ["CBP_WL_STAMPS"] = "1"
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
mat = capi.Matrix(ctx, w["inst"])
s, v, b = w["deltas"]
mat.batch_load_delta(w["lb"], w["ub"], s, v, b)
for _ in range(3):
    mat.batch_run()
    print(name, ctx.stats(), file=sys.stderr)
