"""GPU: the raw-round entry (CBP_FLAG_RAW_ROUND) and the row-sharded driver.  Two row blocks are run one
after the other on the one GPU and combined exactly as the all-reduce would combine two ranks."""
import numpy as np
import pytest
import torch

from cbc_b200 import capi
from cbc_b200 import instances as I
from cbc_b200 import multigpu as MG
from helpers import assert_same_bounds

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("maker", [
    lambda: I.mixed_bigm(m=30000, n=25000, seed=81),
    lambda: I.power_law(n=20000, nnz_target=150000, n_dense=3, dense_len=6000, max_len=2500, seed=82),
    lambda: I.power_law(n=20000, nnz_target=150000, n_dense=3, dense_len=6000, max_len=2500, seed=83, infeasible=True),
])
def test_two_row_blocks_equal_whole(oracle, gpu_ctx, maker):
    inst = maker()
    blocks = MG.row_blocks(inst.rowStart, 2)
    mats = [capi.Matrix(gpu_ctx, MG.sub_instance(inst, r0, r1)) for r0, r1 in blocks]
    rounds = [MG.cuda_local_round(m) for m in mats]

    def local_round(lb, ub):  # what max-allreduce over [lb; -ub; -row] does for two ranks
        a = rounds[0](lb, ub)
        b = rounds[1](lb, ub)
        bad = [x for x in (a[2] + blocks[0][0] if a[2] >= 0 else -1, b[2] + blocks[1][0] if b[2] >= 0 else -1) if x >= 0]
        return torch.maximum(a[0], b[0]), torch.minimum(a[1], b[1]), (min(bad) if bad else -1)

    lb = torch.tensor(inst.lb, device="cuda")
    ub = torch.tensor(inst.ub, device="cuda")
    r, lb, ub = MG.propagate_row_sharded(local_round, lb, ub, torch.tensor(inst.colType != 0, device="cuda"),
                                         inf=inst.inf)
    ro, lo, uo = oracle.propagate(oracle.Problem(inst), inst.lb, inst.ub)
    for k in ("feasible", "rounds", "nChanged", "infeasibleRow", "infeasibleCol", "stopReason", "roundChanged"):
        assert r[k] == ro[k], (k, r[k], ro[k])
    assert_same_bounds(inst.colType, lb.cpu().numpy(), ub.cpu().numpy(), lo, uo, "row blocks")
    for m in mats:
        m.close()
