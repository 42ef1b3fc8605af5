"""Worker of tests/test_gpu_peer_shard.py: one process per GPU (torchrun), rows sharded over the ranks,
the fused peer-memory kernel against the oracle over the whole matrix.  Prints PEER_SHARD_OK on rank 0."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from cbc_b200 import capi  # noqa: E402
from cbc_b200 import instances as I  # noqa: E402
from cbc_b200 import multigpu as MG  # noqa: E402
from helpers import assert_same_bounds  # noqa: E402
from oracle import oracle_py as O  # noqa: E402


def cases():
    yield I.mixed_bigm(m=30000, n=25000, seed=81)
    yield I.power_law(n=20000, nnz_target=150000, n_dense=3, dense_len=6000, max_len=2500, seed=82)
    yield I.power_law(n=20000, nnz_target=150000, n_dense=3, dense_len=6000, max_len=2500, seed=83, infeasible=True)
    yield I.set_cover(m=4000, n=9000, fix_frac=0.995, seed=84)
    yield I.knapsack_assignment(m=5000, n=12000, seed=85)
    yield I.micro("knapsack")


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dist.init_process_group("cpu:gloo,cuda:nccl", rank=rank, world_size=world)
    ctx = capi.Context(torch.cuda.current_device())
    for inst in cases():
        ro, lo, uo = O.propagate(O.Problem(inst), inst.lb, inst.ub)
        sh = MG.PeerShardedMatrix(ctx, inst)
        for rep in range(3):  # the state block is reused: run it more than once
            r, lb, ub = sh.propagate(inst.lb, inst.ub)
            for k in ("feasible", "rounds", "nChanged", "infeasibleRow", "infeasibleCol", "stopReason", "roundChanged"):
                assert r[k] == ro[k], (inst.name, rank, rep, k, r[k], ro[k])
            assert_same_bounds(inst.colType, lb, ub, lo, uo, f"{inst.name} rank {rank}")
        # round limit: the last round is applied
        r2, lb2, ub2 = sh.propagate(inst.lb, inst.ub, max_rounds=2)
        o2, l2, u2 = O.propagate(O.Problem(inst), inst.lb, inst.ub, max_rounds=2)
        assert r2["rounds"] == o2["rounds"] and r2["stopReason"] == o2["stopReason"], (inst.name, r2, o2)
        assert_same_bounds(inst.colType, lb2, ub2, l2, u2, f"{inst.name} 2 rounds rank {rank}")
        sh.close()
        dist.barrier()
    if rank == 0:
        print("PEER_SHARD_OK", world, flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
