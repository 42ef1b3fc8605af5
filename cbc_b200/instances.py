"""Seeded synthetic MILP instances and bound-vector batches for the five
BASELINE.json configs, plus a free-format MPS writer/reader.

Every instance carries a planted feasible point `xstar`; bound propagation
must never exclude it (soundness property used by the tests).  Problem data is
laid out exactly as CbcBoundPropagation::run extracts it from the solver
(/root/reference/src/CbcBoundPropagation.cpp:90-105): CSR by row, row sense /
rhs / range, column types 0/1/2, column bounds, solver infinity.
"""
from dataclasses import dataclass, field

import numpy as np

SEED0 = 20260611
COIN_INF = 1.0e30  # stand-in for OsiSolverInterface::getInfinity()


@dataclass
class Instance:
    name: str
    m: int
    n: int
    rowStart: np.ndarray  # int64 [m+1]
    col: np.ndarray  # int32 [nnz]
    val: np.ndarray  # float64 [nnz]
    sense: np.ndarray  # uint8 [m], ASCII 'L','G','E','R','N'
    rhs: np.ndarray  # float64 [m]
    range: np.ndarray  # float64 [m]
    colType: np.ndarray  # int8 [n]: 0 continuous, 1 binary, 2 general integer
    lb: np.ndarray  # float64 [n]
    ub: np.ndarray  # float64 [n]
    inf: float = COIN_INF
    xstar: np.ndarray = None
    meta: dict = field(default_factory=dict)

    @property
    def nnz(self):
        return int(self.rowStart[-1])

    def row_bounds(self):
        """(rowLower, rowUpper) following the sense convention used by Cbc
        (/root/reference/src/Cbc_C_Interface.cpp:2219-2235)."""
        rl = np.full(self.m, -self.inf)
        ru = np.full(self.m, self.inf)
        s = self.sense
        L, G, E, R = (s == ord("L")), (s == ord("G")), (s == ord("E")), (s == ord("R"))
        ru[L | E | R] = self.rhs[L | E | R]
        rl[G | E] = self.rhs[G | E]
        rl[R] = self.rhs[R] - np.abs(self.range[R])
        return rl, ru

    def activity(self, x):
        rows = np.repeat(np.arange(self.m), np.diff(self.rowStart))
        return np.bincount(rows, weights=self.val * x[self.col], minlength=self.m)

    def contains(self, lb, ub, x=None, tol=1e-6):
        x = self.xstar if x is None else x
        return bool(np.all(x >= lb - tol) and np.all(x <= ub + tol))

    def is_feasible_point(self, x=None, tol=1e-6):
        x = self.xstar if x is None else x
        act = self.activity(x)
        rl, ru = self.row_bounds()
        keep = self.sense != ord("N")
        return bool(np.all(act[keep] <= ru[keep] + tol) and np.all(act[keep] >= rl[keep] - tol))


def _rng(seed):
    return np.random.Generator(np.random.Philox(seed))


def _csr_from_pairs(m, n, rows, cols, vals):
    """Sort by (row, col), drop duplicate (row, col) pairs, build CSR."""
    key = rows.astype(np.int64) * n + cols.astype(np.int64)
    key, idx = np.unique(key, return_index=True)
    rows = (key // n).astype(np.int64)
    cols = (key % n).astype(np.int32)
    vals = vals[idx]
    rowStart = np.zeros(m + 1, np.int64)
    np.cumsum(np.bincount(rows, minlength=m), out=rowStart[1:])
    return rowStart, cols, np.ascontiguousarray(vals, np.float64), rows


def _sense_from_bounds(rl, ru, inf):
    m = len(rl)
    sense = np.full(m, ord("N"), np.uint8)
    rhs = np.zeros(m)
    rng_ = np.zeros(m)
    lf, uf = rl > -inf, ru < inf
    L = ~lf & uf
    G = lf & ~uf
    E = lf & uf & (rl == ru)
    R = lf & uf & (rl != ru)
    sense[L], rhs[L] = ord("L"), ru[L]
    sense[G], rhs[G] = ord("G"), rl[G]
    sense[E], rhs[E] = ord("E"), ru[E]
    sense[R], rhs[R], rng_[R] = ord("R"), ru[R], ru[R] - rl[R]
    return sense, rhs, rng_


# --------------------------------------------------------------------------
# C1 / C2: set cover
# --------------------------------------------------------------------------
def set_cover(m=10_000, n=20_000, density=0.002, fix_frac=0.0, seed=SEED0 + 1, name="c1_setcover"):
    """Rows sum(x) >= 1 over binaries; every row and column non-empty.
    `fix_frac` of the columns outside the planted cover start fixed at 0."""
    g = _rng(seed)
    k = np.maximum(1, g.binomial(n, density, size=m))
    rows = np.repeat(np.arange(m, dtype=np.int64), k)
    cols = g.integers(0, n, size=rows.size)
    miss = np.setdiff1d(np.arange(n), cols)
    rows = np.concatenate([rows, g.integers(0, m, size=miss.size)])
    cols = np.concatenate([cols, miss])
    rowStart, col, val, _ = _csr_from_pairs(m, n, rows, cols, np.ones(rows.size))
    return _finish_cover(name, m, n, rowStart, col, val, fix_frac, g, seed)


def fast0507_shaped(m=507, n=63_009, col_degree=6.5, fix_frac=0.0, seed=SEED0 + 2, name="c2_fast0507"):
    """fast0507 look-alike (507 x 63,009, ~409k nonzeros; shape from
    /root/reference/src/unitTestClp.cpp:317): column degree 1+Poisson, skewed
    row lengths."""
    g = _rng(seed)
    deg = 1 + g.poisson(col_degree - 1.0, size=n)
    deg = np.minimum(deg, m)
    cols = np.repeat(np.arange(n, dtype=np.int64), deg)
    w = g.lognormal(0.0, 0.6, size=m)
    rows = g.choice(m, size=cols.size, p=w / w.sum())
    miss = np.setdiff1d(np.arange(m), rows)
    rows = np.concatenate([rows, miss])
    cols = np.concatenate([cols, g.integers(0, n, size=miss.size)])
    rowStart, col, val, _ = _csr_from_pairs(m, n, rows, cols, np.ones(rows.size))
    return _finish_cover(name, m, n, rowStart, col, val, fix_frac, g, seed)


def _finish_cover(name, m, n, rowStart, col, val, fix_frac, g, seed):
    # planted cover, kept small: visit rows in random order, add a column only
    # for rows that the cover so far misses
    xstar = np.zeros(n)
    r01 = g.random(m)
    for i in g.permutation(m):
        s, e = rowStart[i], rowStart[i + 1]
        if not xstar[col[s:e]].any():
            xstar[col[s + int(r01[i] * (e - s))]] = 1.0
    lb, ub = np.zeros(n), np.ones(n)
    if fix_frac > 0:
        cand = np.flatnonzero(xstar == 0)
        fixed = cand[g.random(cand.size) < fix_frac]
        ub[fixed] = 0.0
    return Instance(
        name, m, n, rowStart, col, val, np.full(m, ord("G"), np.uint8), np.ones(m), np.zeros(m),
        np.ones(n, np.int8), lb, ub, COIN_INF, xstar, {"seed": seed, "fix_frac": fix_frac},
    )


# --------------------------------------------------------------------------
# C3: mixed-integer with big-M rows
# --------------------------------------------------------------------------
def mixed_bigm(m=1_000_000, n=1_000_000, avg_len=10.0, fix_frac=0.05, tight_frac=0.35,
               seed=SEED0 + 3, name="c3_mixed", locality=0.0, band=2048):
    """50% binary / 20% general integer in [0,U] / 30% continuous in [0,1e3]
    (2% of those with an infinite upper bound).  Rows: 40% knapsack 'L' with
    integer weights, 30% big-M rows x - M*y + extras <= rhs, 20% equalities,
    10% ranged.  All data is integral or dyadic so row sums are exact.
    Columns of a row are uniform over all n by default (no locality at all: the worst case for
    the bound gathers); with locality=p a fraction p of them falls within +-band of the row's
    diagonal position, the staircase / block structure most real MILPs have."""
    g = _rng(seed)
    u = g.random(n)
    colType = np.where(u < 0.5, 1, np.where(u < 0.7, 2, 0)).astype(np.int8)
    lb = np.zeros(n)
    ub = np.ones(n)
    gi = colType == 2
    ub[gi] = g.integers(5, 101, size=int(gi.sum())).astype(np.float64)
    ct = colType == 0
    ub[ct] = 1000.0
    xstar = np.zeros(n)
    xstar[colType == 1] = (g.random(int((colType == 1).sum())) < 0.3).astype(np.float64)
    xstar[gi] = np.floor(g.random(int(gi.sum())) * (ub[gi] + 1.0))
    xstar[ct] = np.floor(g.random(int(ct.sum())) * 8000.0) / 8.0
    infub = ct & (g.random(n) < 0.02)
    ub[infub] = COIN_INF

    ln = np.maximum(2, g.poisson(avg_len, size=m))
    rows = np.repeat(np.arange(m, dtype=np.int64), ln)
    cols = g.integers(0, n, size=rows.size)
    if locality > 0.0:
        near = g.random(rows.size) < locality
        centre = (rows * n) // m
        local = np.clip(centre + g.integers(-band, band + 1, size=rows.size), 0, n - 1)
        cols = np.where(near, local, cols)
    rowStart, col, _, rows = _csr_from_pairs(m, n, rows, cols, np.zeros(rows.size))
    nnz = col.size
    rt = g.random(m)
    rtype = np.where(rt < 0.4, 0, np.where(rt < 0.7, 1, np.where(rt < 0.9, 2, 3)))
    val = g.integers(1, 21, size=nnz).astype(np.float64)
    first = rowStart[:-1]
    bm = np.flatnonzero(rtype == 1)
    # big-M rows: first entry +1, second entry -M, remaining entries small positive
    inrow = np.arange(nnz) - np.repeat(first, np.diff(rowStart))
    isbm = np.repeat(rtype == 1, np.diff(rowStart))
    val[isbm] = g.integers(1, 4, size=int(isbm.sum())).astype(np.float64)
    val[first[bm]] = 1.0
    M = np.round(10.0 ** g.uniform(3.0, 6.0, size=bm.size))
    val[first[bm] + 1] = -M
    # equalities / ranged: signed small integer weights
    iseq = np.repeat(rtype >= 2, np.diff(rowStart))
    sgn = np.where(g.random(nnz) < 0.25, -1.0, 1.0)
    val[iseq] = val[iseq] % 5 + 1
    val[iseq] *= sgn[iseq]
    del inrow

    act = np.bincount(rows, weights=val * xstar[col], minlength=m)
    tight = g.random(m) < tight_frac
    slack = np.where(tight, 0.0, np.floor(g.random(m) * 40.0) + 1.0)
    rl = np.full(m, -COIN_INF)
    ru = np.full(m, COIN_INF)
    k0, k1, k2, k3 = rtype == 0, rtype == 1, rtype == 2, rtype == 3
    ru[k0] = act[k0] + slack[k0]
    ru[k1] = np.maximum(act[k1], 0.0) + np.where(tight[k1], 0.0, slack[k1])
    ru[k2] = act[k2]
    rl[k2] = act[k2]
    ru[k3] = act[k3] + slack[k3]
    rl[k3] = act[k3] - (np.floor(g.random(int(k3.sum())) * 20.0) + 1.0)
    sense, rhs, rge = _sense_from_bounds(rl, ru, COIN_INF)
    if fix_frac > 0:
        fx = g.random(n) < fix_frac
        lb[fx] = xstar[fx]
        ub[fx] = xstar[fx]
    return Instance(name, m, n, rowStart, col, val, sense, rhs, rge, colType, lb, ub, COIN_INF, xstar,
                    {"seed": seed, "fix_frac": fix_frac})


# --------------------------------------------------------------------------
# C4: knapsack + assignment
# --------------------------------------------------------------------------
def knapsack_assignment(m=200_000, n=500_000, seed=SEED0 + 4, name="c4_knapassign", knap_per_col=5,
                        slack_max=6, root_fixpoint=False):
    """Every binary column sits in 1 assignment row (sum x = 1) and
    `knap_per_col` knapsack rows (integer weights 1..20).  root_fixpoint=True gives every knapsack
    at least the capacity of its heaviest item, so the root bounds are already a propagation
    fixpoint (a node as the solver would hand it to a dive) and only the dive's own fixings propagate."""
    g = _rng(seed)
    mE = max(1, m // 4)
    mK = m - mE
    arow = g.integers(0, mE, size=n)
    arow[:mE] = np.arange(mE)  # no empty assignment row
    rows = [arow.astype(np.int64)]
    cols = [np.arange(n, dtype=np.int64)]
    vals = [np.ones(n)]
    for _ in range(knap_per_col):
        rows.append(mE + g.integers(0, mK, size=n))
        cols.append(np.arange(n, dtype=np.int64))
        vals.append(g.integers(1, 21, size=n).astype(np.float64))
    rows, cols, vals = np.concatenate(rows), np.concatenate(cols), np.concatenate(vals)
    rowStart, col, val, rws = _csr_from_pairs(m, n, rows, cols, vals)
    # planted assignment: one column per assignment row
    xstar = np.zeros(n)
    order = g.permutation(n)
    first_of_row = np.full(mE, -1, np.int64)
    first_of_row[arow[order][::-1]] = order[::-1]
    xstar[first_of_row[first_of_row >= 0]] = 1.0
    act = np.bincount(rws, weights=val * xstar[col], minlength=m)
    rl = np.full(m, -COIN_INF)
    ru = np.full(m, COIN_INF)
    rl[:mE] = 1.0
    ru[:mE] = 1.0
    ru[mE:] = act[mE:] + g.integers(0, slack_max + 1, size=mK)
    if root_fixpoint:
        heaviest = np.zeros(m)
        np.maximum.at(heaviest, rws, val)
        ru[mE:] = np.maximum(ru[mE:], heaviest[mE:])
    sense, rhs, rge = _sense_from_bounds(rl, ru, COIN_INF)
    return Instance(name, m, n, rowStart, col, val, sense, rhs, rge, np.ones(n, np.int8),
                    np.zeros(n), np.ones(n), COIN_INF, xstar, {"seed": seed, "mE": mE})


# --------------------------------------------------------------------------
# C5: power-law row lengths with a few very long rows
# --------------------------------------------------------------------------
def power_law(n=1_000_000, nnz_target=5_000_000, n_dense=8, dense_len=100_000, alpha=1.6,
              max_len=10_000, infeasible=False, seed=SEED0 + 5, name="c5_powerlaw"):
    """Zipf(alpha) row lengths (truncated) plus `n_dense` rows of `dense_len`
    nonzeros; mixed column types; knapsack-style 'L' rows and some 'G' rows.
    infeasible=True lowers one dense row's rhs below its minimum activity."""
    g = _rng(seed)
    u = g.random(n)
    colType = np.where(u < 0.5, 1, np.where(u < 0.7, 2, 0)).astype(np.int8)
    lb = np.zeros(n)
    ub = np.ones(n)
    gi = colType == 2
    ub[gi] = g.integers(5, 101, size=int(gi.sum())).astype(np.float64)
    ct = colType == 0
    ub[ct] = 100.0
    xstar = np.where(colType == 1, (g.random(n) < 0.3).astype(np.float64),
                     np.where(gi, np.floor(g.random(n) * (ub + 1.0)), np.floor(g.random(n) * 800.0) / 8.0))
    dense_len = min(dense_len, n)
    budget = max(0, nnz_target - n_dense * dense_len)
    lens = []
    tot = 0
    while tot < budget:
        z = np.minimum(g.zipf(alpha, size=65536), max_len)
        lens.append(z)
        tot += int(z.sum())
    lens = np.concatenate(lens) if lens else np.zeros(0, np.int64)
    cut = int(np.searchsorted(np.cumsum(lens), budget)) + 1
    lens = lens[:cut]
    mS = lens.size
    m = mS + n_dense
    # dense rows are spread through the row order
    pos = np.sort(g.choice(m, size=n_dense, replace=False)) if n_dense else np.zeros(0, np.int64)
    rl_len = np.zeros(m, np.int64)
    mask = np.ones(m, bool)
    mask[pos] = False
    rl_len[mask] = lens
    rl_len[pos] = dense_len
    rows = np.repeat(np.arange(m, dtype=np.int64), rl_len)
    cols = g.integers(0, n, size=rows.size)
    for p_ in pos:  # dense rows: distinct columns
        s = int(np.sum(rl_len[:p_]))
        cols[s:s + dense_len] = g.choice(n, size=dense_len, replace=False)
    vals = g.integers(1, 10, size=rows.size).astype(np.float64)
    vals *= np.where(g.random(rows.size) < 0.15, -1.0, 1.0)
    rowStart, col, val, rws = _csr_from_pairs(m, n, rows, cols, vals)
    act = np.bincount(rws, weights=val * xstar[col], minlength=m)
    isG = g.random(m) < 0.2
    slack = np.floor(g.random(m) * 10.0)
    rl = np.where(isG, act - slack, -COIN_INF)
    ru = np.where(isG, COIN_INF, act + slack)
    bad_row = -1
    if infeasible and n_dense:
        bad_row = int(pos[n_dense // 2])
        s, e = rowStart[bad_row], rowStart[bad_row + 1]
        a = val[s:e]
        j = col[s:e]
        min_act = np.sum(np.where(a > 0, a * lb[j], a * ub[j]))
        rl[bad_row], ru[bad_row] = -COIN_INF, min_act - 5.0
    sense, rhs, rge = _sense_from_bounds(rl, ru, COIN_INF)
    return Instance(name, m, n, rowStart, col, val, sense, rhs, rge, colType, lb, ub, COIN_INF,
                    None if infeasible else xstar,
                    {"seed": seed, "dense_rows": [int(x) for x in pos], "bad_row": bad_row})


# --------------------------------------------------------------------------
# bound-vector batches in the node bound-diff wire format of
# /root/reference/src/CbcNode.cpp:265-273 (column | 0x80000000 = upper bound)
# --------------------------------------------------------------------------
UPPER_FLAG = np.uint32(0x80000000)
ODD_WAY_FLAG = np.uint32(0x40000000)  # set by CbcPartialNodeInfo::applyBounds (src/CbcPartialNodeInfo.cpp:208,220); not part of the column
COL_MASK = np.uint32(0x3FFFFFFF)      # the column is `variable & 0x3fffffff` (src/CbcPartialNodeInfo.cpp:139,196)


def strong_branch_deltas(inst, lb, ub, n_candidates, seed=SEED0 + 11):
    """2*n_candidates vectors: candidate column j gives a down arm (ub=floor)
    and an up arm (lb=ceil) as CbcIntegerBranchingObject::branch would set them
    (/root/reference/src/CbcSimpleInteger.cpp:450-454,499-501)."""
    g = _rng(seed)
    free = np.flatnonzero((inst.colType != 0) & (ub - lb >= 1.0) & (ub < inst.inf) & (lb > -inst.inf))
    if free.size == 0:
        raise ValueError("no free integer column to branch on")
    cand = g.choice(free, size=n_candidates, replace=free.size < n_candidates)
    mid = lb[cand] + np.floor(g.random(n_candidates) * (ub[cand] - lb[cand])) + 0.5
    var = np.empty(2 * n_candidates, np.uint32)
    bnd = np.empty(2 * n_candidates)
    var[0::2] = cand.astype(np.uint32) | UPPER_FLAG
    bnd[0::2] = np.floor(mid)
    var[1::2] = cand.astype(np.uint32)
    bnd[1::2] = np.ceil(mid)
    start = np.arange(2 * n_candidates + 1, dtype=np.int64)
    return start, var, bnd


def dive_deltas(inst, lb, ub, B, max_depth=64, seed=SEED0 + 12, follow_xstar=0.9):
    """B dive states: vector b = base + d_b cumulative fixings (d_b geometric,
    1..max_depth), mostly towards the planted point so many stay feasible."""
    g = _rng(seed)
    free = np.flatnonzero((inst.colType != 0) & (ub - lb >= 1.0) & (ub < inst.inf) & (lb > -inst.inf))
    depth = np.minimum(max_depth, g.geometric(0.12, size=B))
    start = np.zeros(B + 1, np.int64)
    np.cumsum(depth, out=start[1:])
    tot = int(start[-1])
    cols = g.choice(free, size=tot)
    if inst.xstar is not None:
        target = inst.xstar[cols]
        flip = g.random(tot) > follow_xstar
        target = np.where(flip, np.clip(target + np.where(g.random(tot) < 0.5, 1.0, -1.0), lb[cols], ub[cols]), target)
    else:
        target = lb[cols] + np.floor(g.random(tot) * (ub[cols] - lb[cols] + 1.0))
    up = g.random(tot) < 0.5
    # fixing = one bound moved onto the target value (the other side follows in later fixings)
    var = cols.astype(np.uint32)
    var = np.where(up, var, var | UPPER_FLAG).astype(np.uint32)
    # "up" moves the lower bound to target, "down" moves the upper bound to target
    return start, var, target.astype(np.float64)


def probe_deltas(inst, lb, ub, n_probe, seed=SEED0 + 13):
    """Probing vectors in the style of fixVubs
    (/root/reference/src/CbcSolverHeuristics.cpp:270-276): for a binary column
    fake ub=0, and the complementary lb=1 probe."""
    g = _rng(seed)
    free = np.flatnonzero((inst.colType == 1) & (lb == 0.0) & (ub == 1.0))
    cand = g.choice(free, size=n_probe, replace=free.size < n_probe)
    var = np.empty(2 * n_probe, np.uint32)
    bnd = np.empty(2 * n_probe)
    var[0::2] = cand.astype(np.uint32) | UPPER_FLAG
    bnd[0::2] = 0.0
    var[1::2] = cand.astype(np.uint32)
    bnd[1::2] = 1.0
    return np.arange(2 * n_probe + 1, dtype=np.int64), var, bnd


def expand_deltas(lb, ub, start, var, bnd):
    """Dense (B, n) bound arrays from base + deltas."""
    B = len(start) - 1
    L = np.repeat(lb[None, :], B, axis=0)
    U = np.repeat(ub[None, :], B, axis=0)
    b = np.repeat(np.arange(B), np.diff(start))
    j = (var & COL_MASK).astype(np.int64)
    isU = (var & UPPER_FLAG) != 0
    # later deltas of the same vector win: apply in order
    for k in np.flatnonzero(isU):
        U[b[k], j[k]] = bnd[k]
    for k in np.flatnonzero(~isU):
        L[b[k], j[k]] = bnd[k]
    return L, U


# --------------------------------------------------------------------------
# free-format MPS (config 1 is "written as MPS")
# --------------------------------------------------------------------------
def write_mps(inst, path):
    m, n = inst.m, inst.n
    csc_order = np.argsort(inst.col, kind="stable")
    rows = np.repeat(np.arange(m), np.diff(inst.rowStart))[csc_order]
    cols = inst.col[csc_order]
    vals = inst.val[csc_order]
    colStart = np.zeros(n + 1, np.int64)
    np.cumsum(np.bincount(cols, minlength=n), out=colStart[1:])
    with open(path, "w") as f:
        f.write(f"NAME {inst.name}\nROWS\n N OBJ\n")
        for i in range(m):
            s = chr(inst.sense[i])
            f.write(f" {'L' if s == 'R' else s} R{i}\n")
        f.write("COLUMNS\n")
        in_int = False
        for j in range(n):
            isint = inst.colType[j] != 0
            if isint != in_int:
                f.write(f" M{j} 'MARKER' '{'INTORG' if isint else 'INTEND'}'\n")
                in_int = isint
            f.write(f" C{j} OBJ 1\n")
            for k in range(colStart[j], colStart[j + 1]):
                f.write(f" C{j} R{rows[k]} {float(vals[k])!r}\n")
        if in_int:
            f.write(" MEND 'MARKER' 'INTEND'\n")
        f.write("RHS\n")
        for i in range(m):
            if inst.rhs[i] != 0.0:
                f.write(f" RHS R{i} {float(inst.rhs[i])!r}\n")
        f.write("RANGES\n")
        for i in np.flatnonzero(inst.sense == ord("R")):
            f.write(f" RNG R{i} {float(abs(inst.range[i]))!r}\n")
        f.write("BOUNDS\n")
        for j in range(n):
            l, u = float(inst.lb[j]), float(inst.ub[j])
            if l == u:
                f.write(f" FX BND C{j} {l!r}\n")
                continue
            if l <= -inst.inf:
                f.write(f" MI BND C{j}\n")
            elif l != 0.0:
                f.write(f" LO BND C{j} {l!r}\n")
            if u >= inst.inf:
                if inst.colType[j] != 0:
                    f.write(f" PL BND C{j}\n")
            else:
                f.write(f" UP BND C{j} {u!r}\n")
        f.write("ENDATA\n")


def read_mps(path, inf=COIN_INF):
    """Minimal free-format reader for files produced by write_mps."""
    name = "mps"
    section = None
    row_idx, row_sense = {}, []
    col_idx, col_int = {}, []
    ents = []
    rhs, rngs, bnds = {}, {}, []
    in_int = False
    with open(path) as f:
        for line in f:
            t = line.split()
            if not t or t[0].startswith("*"):
                continue
            if not line[0].isspace():
                section = t[0]
                if section == "NAME" and len(t) > 1:
                    name = t[1]
                continue
            if section == "ROWS":
                if t[0] != "N":
                    row_idx[t[1]] = len(row_sense)
                    row_sense.append(t[0])
            elif section == "COLUMNS":
                if len(t) >= 3 and t[1] == "'MARKER'":
                    in_int = "INTORG" in t[2]
                    continue
                c = t[0]
                if c not in col_idx:
                    col_idx[c] = len(col_int)
                    col_int.append(in_int)
                for r, v in zip(t[1::2], t[2::2]):
                    if r in row_idx:
                        ents.append((row_idx[r], col_idx[c], float(v)))
            elif section == "RHS":
                for r, v in zip(t[1::2], t[2::2]):
                    if r in row_idx:
                        rhs[row_idx[r]] = float(v)
            elif section == "RANGES":
                for r, v in zip(t[1::2], t[2::2]):
                    rngs[row_idx[r]] = float(v)
            elif section == "BOUNDS":
                bnds.append(t)
    m, n = len(row_sense), len(col_int)
    colType = np.array([2 if x else 0 for x in col_int], np.int8)
    lb = np.zeros(n)
    ub = np.full(n, inf)
    ub[colType != 0] = 1.0  # MPS default for marker integers without bounds
    for t in bnds:
        j = col_idx[t[2]]
        kind = t[0]
        if kind == "FX":
            lb[j] = ub[j] = float(t[3])
        elif kind == "LO":
            lb[j] = float(t[3])
        elif kind == "UP":
            ub[j] = float(t[3])
        elif kind == "MI":
            lb[j] = -inf
        elif kind == "PL":
            ub[j] = inf
        elif kind == "BV":
            lb[j], ub[j] = 0.0, 1.0
    colType[(colType != 0) & (lb == 0.0) & (ub == 1.0)] = 1
    e = np.array(ents, dtype=[("r", "i8"), ("c", "i8"), ("v", "f8")]) if ents else np.zeros(0, [("r", "i8"), ("c", "i8"), ("v", "f8")])
    rowStart, col, val, _ = _csr_from_pairs(m, n, e["r"], e["c"], e["v"])
    sense = np.array([ord(s) for s in row_sense], np.uint8)
    rhs_a = np.zeros(m)
    for i, v in rhs.items():
        rhs_a[i] = v
    rng_a = np.zeros(m)
    for i, v in rngs.items():
        rng_a[i] = abs(v)
        if sense[i] == ord("L"):
            sense[i] = ord("R")
        elif sense[i] == ord("G"):  # G with range r: [rhs, rhs+r]
            sense[i] = ord("R")
            rhs_a[i] = rhs_a[i] + abs(v)
    return Instance(name, m, n, rowStart, col, val, sense, rhs_a, rng_a, colType, lb, ub, inf)


# hand-checkable micro-instances lifted from the reference's C-API tests
def micro(name):
    inf = COIN_INF
    if name == "knapsack":  # /root/reference/test/CInterfaceTest.c:32-59
        return Instance("knapsack", 1, 5, np.array([0, 5], np.int64), np.arange(5, dtype=np.int32),
                        np.array([2.0, 8, 4, 2, 5]), np.array([ord("L")], np.uint8), np.array([10.0]),
                        np.zeros(1), np.ones(5, np.int8), np.zeros(5), np.ones(5), inf,
                        np.array([1.0, 1, 0, 0, 0]))
    if name == "integer_infeasible":  # test/CInterfaceTest.c:328-358
        return Instance("integer_infeasible", 1, 1, np.array([0, 1], np.int64), np.zeros(1, np.int32),
                        np.ones(1), np.array([ord("L")], np.uint8), np.array([-10.0]), np.zeros(1),
                        np.ones(1, np.int8), np.zeros(1), np.ones(1), inf)
    if name == "integer_bounds":  # test/CInterfaceTest.c:401-436
        return Instance("integer_bounds", 1, 2, np.array([0, 2], np.int64), np.array([0, 1], np.int32),
                        np.ones(2), np.array([ord("L")], np.uint8), np.array([3.0]), np.zeros(1),
                        np.array([0, 2], np.int8), np.zeros(2), np.array([3.0, 1.5]), inf,
                        np.array([2.0, 1.0]))
    if name == "integer_unbounded":  # test/CInterfaceTest.c:362-399
        return Instance("integer_unbounded", 2, 2, np.array([0, 2, 4], np.int64),
                        np.array([0, 1, 0, 1], np.int32), np.array([1.0, 1, 1, -1]),
                        np.array([ord("L"), ord("E")], np.uint8), np.array([3.0, 0.0]), np.zeros(2),
                        np.array([2, 0], np.int8), np.full(2, -inf), np.full(2, inf), inf,
                        np.array([1.0, 1.0]))
    if name == "infeasible_preprocess":  # test/CInterfaceTest.c:261-326 (8 binaries a..h, 14 rows, infeasible only by search)
        rows = [(1, 2, 5, 6), (0, 3, 4, 7), (0, 1), (2, 3), (4, 5), (6, 7), (0, 2), (1, 3), (2, 4), (3, 5), (4, 6),
                (5, 7), (0, 6), (1, 7)]
        start = np.zeros(15, np.int64)
        np.cumsum([len(r) for r in rows], out=start[1:])
        col = np.array([j for r in rows for j in r], np.int32)
        sense = np.frombuffer(b"LLEEEELLLLLLLL", np.uint8).copy()
        rhs = np.array([3.0, 3, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1])
        return Instance("infeasible_preprocess", 14, 8, start, col, np.ones(col.size), sense, rhs, np.zeros(14),
                        np.ones(8, np.int8), np.zeros(8), np.ones(8), inf)
    raise KeyError(name)
