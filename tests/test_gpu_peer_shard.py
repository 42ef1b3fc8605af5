"""GPU: rows sharded over ranks with the exchange inside the kernel (cbp_p2p_*).  World size 1 runs in
this process; world size 2 needs two GPUs and is launched with torchrun (skipped on a one-GPU box)."""
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

from cbc_b200 import capi
from cbc_b200 import instances as I
from helpers import assert_same_bounds

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("maker", [
    lambda: I.mixed_bigm(m=30000, n=25000, seed=81),
    lambda: I.power_law(n=20000, nnz_target=150000, n_dense=3, dense_len=6000, max_len=2500, seed=83, infeasible=True),
    lambda: I.micro("knapsack"),
])
def test_world_of_one_equals_oracle(oracle, gpu_ctx, maker):
    inst = maker()
    mat = capi.Matrix(gpu_ctx, inst)
    mat.p2p_attach(0, 1, [mat.p2p_export(1)])
    ro, lo, uo = oracle.propagate(oracle.Problem(inst), inst.lb, inst.ub)
    for _ in range(2):
        lb, ub = np.array(inst.lb, np.float64), np.array(inst.ub, np.float64)
        r = mat.p2p_propagate(lb, ub, 0)
        for k in ("feasible", "rounds", "nChanged", "infeasibleRow", "infeasibleCol", "stopReason", "roundChanged"):
            assert r[k] == ro[k], (k, r[k], ro[k])
        assert_same_bounds(inst.colType, lb, ub, lo, uo, "world of one")
    # the ordinary entry still works on the same matrix afterwards
    r1, lb1, ub1 = mat.propagate(inst.lb, inst.ub)
    assert r1["rounds"] == ro["rounds"]
    mat.close()


def test_unattached_matrix_is_refused(gpu_ctx):
    inst = I.micro("knapsack")
    mat = capi.Matrix(gpu_ctx, inst)
    lb, ub = np.array(inst.lb, np.float64), np.array(inst.ub, np.float64)
    with pytest.raises(capi.CbpError):
        mat.p2p_propagate(lb, ub, 0)
    mat.close()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_two_ranks_equal_oracle():
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29671", os.path.join(ROOT, "tests", "mp_peer_shard.py")]
    out = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0 and "PEER_SHARD_OK 2" in out.stdout, out.stdout[-3000:] + out.stderr[-3000:]
