"""Cost of cbp_matrix_upload (host re-tiling + allocations + copies): python profiles/upload_time.py"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from cbc_b200 import capi, instances as I  # noqa: E402

ctx = capi.Context(0)
for name, inst in (("100k x 100k, 1M nnz", I.mixed_bigm(m=100_000, n=100_000)),
                   ("1M x 1M, 10M nnz", I.mixed_bigm(m=1_000_000, n=1_000_000)),
                   ("507 x 63,009 set cover", I.fast0507_shaped())):
    ts = []
    for _ in range(4):
        t0 = time.perf_counter()
        mat = capi.Matrix(ctx, inst)
        ctx.synchronize()
        ts.append(time.perf_counter() - t0)
        mat.close()
    print(f"{name}: upload {min(ts) * 1e3:.1f} ms (first {ts[0] * 1e3:.1f})")
