"""CPU, world_size 2 over gloo: the host logic of both multi-GPU decompositions.
The per-rank engine here is the CPU oracle; on the GPU box the same drivers run the CUDA engine."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from cbc_b200 import instances as I
from cbc_b200 import multigpu as MG


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, fn, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        out[rank] = fn(rank, world)
    finally:
        dist.destroy_process_group()


def _spawn(fn, world=2):
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), fn, out), nprocs=world, join=True)
    return [out[r] for r in range(world)]


def _row_sharded(rank, world):
    from oracle import oracle_py as O

    res = {}
    for name, inst in (("mixed", I.mixed_bigm(m=3000, n=2500, seed=71)),
                       ("bad", I.power_law(n=4000, nnz_target=30000, n_dense=2, dense_len=1500, max_len=400,
                                           seed=72, infeasible=True))):
        r0, r1 = MG.row_blocks(inst.rowStart, world)[rank]
        prob = O.Problem(MG.sub_instance(inst, r0, r1))

        def local_round(lb, ub):
            nlb, nub, bad = O.round_rows(prob, lb.numpy(), ub.numpy())
            return torch.from_numpy(nlb), torch.from_numpy(nub), bad

        r, lb, ub = MG.propagate_row_sharded(local_round, torch.from_numpy(inst.lb.copy()),
                                             torch.from_numpy(inst.ub.copy()),
                                             torch.from_numpy(inst.colType != 0), row_offset=r0, inf=inst.inf)
        res[name] = (r, lb.numpy().copy(), ub.numpy().copy())
    return res


def test_row_sharded_equals_single(oracle):
    outs = _spawn(_row_sharded)
    for name, inst in (("mixed", I.mixed_bigm(m=3000, n=2500, seed=71)),
                       ("bad", I.power_law(n=4000, nnz_target=30000, n_dense=2, dense_len=1500, max_len=400,
                                           seed=72, infeasible=True))):
        ro, lo, uo = oracle.propagate(oracle.Problem(inst), inst.lb, inst.ub)
        for rank in range(2):
            r, lb, ub = outs[rank][name]
            for k in ("feasible", "rounds", "nChanged", "infeasibleRow", "infeasibleCol", "stopReason", "roundChanged"):
                assert r[k] == ro[k], (name, rank, k, r[k], ro[k])
            assert np.array_equal(lb, lo) and np.array_equal(ub, uo)


def _batch_sharded(rank, world):
    from oracle import oracle_py as O

    inst = I.knapsack_assignment(m=500, n=1200, seed=73)
    s, v, b = I.dive_deltas(inst, inst.lb, inst.ub, 37, max_depth=25, seed=74)
    ls, lv, lb_ = MG.shard_deltas(s, v, b, rank, world)
    res, _, _ = O.propagate_batch_delta(O.Problem(inst), inst.lb, inst.ub, ls, lv, lb_, want_bounds=False)
    allres = MG.gather_results(res, 37)
    return allres


def test_batch_sharded_equals_single(oracle):
    outs = _spawn(_batch_sharded)
    inst = I.knapsack_assignment(m=500, n=1200, seed=73)
    s, v, b = I.dive_deltas(inst, inst.lb, inst.ub, 37, max_depth=25, seed=74)
    ref, _, _ = oracle.propagate_batch_delta(oracle.Problem(inst), inst.lb, inst.ub, s, v, b, want_bounds=False)
    for rank in range(2):
        for k in ("feasible", "rounds", "nChanged", "infeasibleRow", "infeasibleCol", "stopReason"):
            assert np.array_equal(outs[rank][k], ref[k]), k


def test_shard_helpers():
    assert [MG.shard_range(10, r, 3) for r in range(3)] == [(0, 4), (4, 7), (7, 10)]
    rs = np.array([0, 5, 5, 9, 20, 21, 30])
    blocks = MG.row_blocks(rs, 3)
    assert blocks[0][0] == 0 and blocks[-1][1] == 6
    assert all(blocks[i][1] == blocks[i + 1][0] for i in range(2))
