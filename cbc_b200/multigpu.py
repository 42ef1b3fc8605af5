"""Multi-GPU decompositions of the propagation path, one process per GPU,
torch.distributed for the plumbing.

* batch mode: the node batch is split contiguously across ranks, the matrix is replicated, no
  data-path collective (`shard_deltas`, `gather_results`);
* single vector over a huge matrix: rows are split into blocks balanced by nonzeros, the bound
  vector is replicated, and each Jacobi round ends with ONE max-allreduce over [lb ; -ub ; flags]
  (`propagate_row_sharded`).  The per-rank round is whatever `local_round` does — the CUDA engine on
  a GPU, or any other implementation of one round in the CPU tests of this logic.
  `PeerShardedMatrix` is the fused form of the same decomposition for the GPUs of one box: the
  exchange happens inside the tightening kernel over NVLink peer memory (cbp_p2p_* in
  include/cbp_capi.h), torch.distributed only carries the IPC handles at set-up.
"""
import numpy as np

FEAS_MULT = 100.0
BIG = float(2 ** 40)


# ----------------------------------------------------------------------------- batch sharding
def shard_range(B, rank, world):
    """Contiguous, balanced [lo, hi) of B items for `rank`."""
    base, rem = divmod(B, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def shard_deltas(start, var, bnd, rank, world):
    """This rank's slice of a delta batch (the CbcPartialNodeInfo wire format)."""
    B = len(start) - 1
    lo, hi = shard_range(B, rank, world)
    s = np.asarray(start[lo:hi + 1], np.int64)
    return s - s[0], var[s[0]:s[-1]], bnd[s[0]:s[-1]]


def gather_results(local, B, group=None):
    """All ranks' per-vector result records, concatenated in rank order (host-side gather)."""
    import torch.distributed as dist

    world = dist.get_world_size(group)
    parts = [None] * world
    dist.all_gather_object(parts, local, group=group)
    out = np.concatenate(parts)
    assert len(out) == B
    return out


# ----------------------------------------------------------------------------- row sharding
def row_blocks(rowStart, world):
    """Row ranges [r0, r1) per rank with (nearly) equal nonzero counts."""
    rowStart = np.asarray(rowStart, np.int64)
    m = len(rowStart) - 1
    nnz = int(rowStart[-1])
    cuts = [0]
    for r in range(1, world):
        cuts.append(int(np.searchsorted(rowStart, nnz * r // world, side="left")))
    cuts.append(m)
    cuts = [min(max(c, 0), m) for c in cuts]
    for i in range(1, len(cuts)):
        cuts[i] = max(cuts[i], cuts[i - 1])
    return [(cuts[i], cuts[i + 1]) for i in range(world)]


def sub_instance(inst, r0, r1):
    """The rows [r0, r1) of `inst` as an instance of its own (all columns kept)."""
    from .instances import Instance

    s, e = int(inst.rowStart[r0]), int(inst.rowStart[r1])
    return Instance(f"{inst.name}[{r0}:{r1}]", r1 - r0, inst.n, inst.rowStart[r0:r1 + 1] - s, inst.col[s:e],
                    inst.val[s:e], inst.sense[r0:r1], inst.rhs[r0:r1], inst.range[r0:r1], inst.colType,
                    inst.lb, inst.ub, inst.inf, inst.xstar)


def prepass(lb, ub, col_is_int, tol, inf):
    """Rule R11 on torch tensors: integer columns get ceil(lb - tol) / floor(ub + tol); returns
    (lb, ub, touched mask, smallest crossing column or -1).  A crossing column keeps its input bounds."""
    import torch

    feas = FEAS_MULT * tol
    l = torch.where(col_is_int & (lb.abs() < inf), torch.ceil(lb - tol) + 0.0, lb)
    u = torch.where(col_is_int & (ub.abs() < inf), torch.floor(ub + tol) + 0.0, ub)
    cross = l > u
    hard = cross & (l > u + feas)
    bad = int(torch.nonzero(hard)[0]) if bool(hard.any()) else -1
    ub_moved = cross & (u < ub)
    x = torch.minimum(l, ub)
    u2 = torch.where(ub_moved, x, u)
    l2 = torch.where(ub_moved, x, torch.where(cross, u, l))
    l2 = torch.where(hard, lb, l2)
    u2 = torch.where(hard, ub, u2)
    return l2, u2, (l2 != lb) | (u2 != ub), bad


def propagate_row_sharded(local_round, lb, ub, col_is_int, tol=1e-7, max_rounds=0, row_offset=0, group=None,
                          inf=1e30):
    """Round loop of src/CbcBoundPropagation.cpp:107-195 over row-sharded data.

    local_round(lb, ub) -> (lb_prop, ub_prop, first violated local row or -1): the row phase of ONE Jacobi
    round (rules R3-R8) over this rank's rows from the replicated round-start bounds: per column the max of
    the accepted lower and the min of the accepted upper candidates (unchanged where nothing was accepted).
    Everything is a torch tensor on the rank's device.  R11 and R9 are applied here, once, on the combined
    bounds, so the result is what a single device computes over all rows.  Every rank returns the same values.
    """
    import torch
    import torch.distributed as dist

    n = lb.numel()
    feas = FEAS_MULT * tol
    rounds = total = 0
    per_round = []
    stop, bad_row, bad_col = 0, -1, -1
    limit = max_rounds if max_rounds > 0 else 1 << 62
    buf = torch.empty(2 * n + 1, dtype=lb.dtype, device=lb.device)
    lb, ub, pre_touched, bad0 = prepass(lb, ub, col_is_int, tol, inf)
    if bad0 >= 0:
        return ({"feasible": False, "rounds": 1, "nChanged": 0, "infeasibleRow": -1, "infeasibleCol": bad0,
                 "stopReason": 4, "roundChanged": []}, lb, ub)
    while True:
        nlb, nub, bad = local_round(lb, ub)
        rounds += 1
        # one collective per round: max over [lb ; -ub ; -(first violated row, BIG if none)]
        buf[:n] = nlb
        buf[n:2 * n] = -nub
        buf[2 * n] = -(float(bad + row_offset) if bad >= 0 else BIG)
        if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
            dist.all_reduce(buf, op=dist.ReduceOp.MAX, group=group)
        nlb, nub = buf[:n].clone(), -buf[n:2 * n]
        first_bad = -float(buf[2 * n])
        if first_bad < BIG:
            stop, bad_row = 4, int(first_bad)
            break
        # R9 on the combined proposals
        moved = (nlb != lb) | (nub != ub)
        if rounds == 1:
            moved = moved | pre_touched
        cross = nlb > nub
        if bool(cross.any()):
            hard = cross & (nlb > nub + feas)
            if bool(hard.any()):
                stop, bad_col = 4, int(torch.nonzero(hard)[0])
                break
            ub_moved = cross & (nub < ub)
            x = torch.minimum(nlb, ub)
            nub2 = torch.where(ub_moved, x, nub)
            nlb = torch.where(ub_moved, x, torch.where(cross, nub, nlb))
            nub = nub2
        changed = int(moved.sum())
        per_round.append(changed)
        if changed == 0:
            stop = 1
            break
        lb, ub = nlb, nub
        total += changed
        if rounds >= limit:
            stop = 2
            break
    res = {"feasible": stop != 4, "rounds": rounds, "nChanged": total, "infeasibleRow": bad_row,
           "infeasibleCol": bad_col, "stopReason": stop, "roundChanged": per_round}
    return res, lb, ub


def cuda_local_round(mat, tol=1e-7):
    """local_round for propagate_row_sharded backed by the CUDA engine (CBP_FLAG_RAW_ROUND)."""
    def fn(lb, ub):
        nlb, nub = lb.clone(), ub.clone()
        r = mat.propagate_device(nlb.data_ptr(), nub.data_ptr(), tol=tol, max_rounds=1, raw_round=True)
        return nlb, nub, (r["infeasibleRow"] if not r["feasible"] else -1)
    return fn


class PeerShardedMatrix:
    """One rank's block of rows of `inst`, wired to the other ranks' state blocks over peer memory.

    Collective constructor (every rank of `group` calls it with the same instance).  `propagate` is
    collective as well and returns the same result and bounds on every rank."""

    def __init__(self, ctx, inst, group=None):
        import torch.distributed as dist

        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self.r0, self.r1 = row_blocks(inst.rowStart, self.world)[self.rank]
        self.n = inst.n
        sub = sub_instance(inst, self.r0, self.r1)
        from .capi import Matrix

        self.mat = Matrix(ctx, sub)
        handles = [None] * self.world
        dist.all_gather_object(handles, self.mat.p2p_export(self.world), group=group)
        self.mat.p2p_attach(self.rank, self.world, handles)
        dist.barrier(group=group)

    def propagate(self, lb, ub, tol=1e-7, max_rounds=0, time_limit=1e100):
        lb = np.array(lb, np.float64)
        ub = np.array(ub, np.float64)
        res = self.mat.p2p_propagate(lb, ub, self.r0, tol, max_rounds, time_limit)
        return res, lb, ub

    def propagate_device(self, lb_ptr, ub_ptr, tol=1e-7, max_rounds=0, time_limit=1e100):
        """Same with the bounds already in this rank's HBM (n doubles each, updated in place)."""
        return self.mat.p2p_propagate(lb_ptr, ub_ptr, self.r0, tol, max_rounds, time_limit, device_ptrs=True)

    def close(self):
        self.mat.close()
