"""ctypes binding of libcbp.so (include/cbp_capi.h) — the only way Python code
in this repo reaches the CUDA engine.  There is no CPU fallback: if the
library is missing, or the machine has no CUDA device, this raises.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("CBP_LIB", os.path.join(_HERE, "lib", "libcbp.so"))  # CBP_LIB: A/B builds while tuning

FLAG_DEVICE_PTRS = 1
FLAG_RAW_ROUND = 2
ERR_CAPACITY = -4
STOP_NAMES = {0: "NotRun", 1: "ReachedFixpoint", 2: "HitMaxRounds", 3: "HitTimeLimit", 4: "InfeasibleDetected"}

RESULT_DTYPE = np.dtype(
    [("feasible", "i4"), ("rounds", "i4"), ("nChanged", "i4"), ("infeasibleRow", "i4"),
     ("infeasibleCol", "i4"), ("stopReason", "i4"), ("seconds", "f8")],
    align=True,
)


class Result(C.Structure):
    _fields_ = [("feasible", C.c_int), ("rounds", C.c_int), ("nChanged", C.c_int),
                ("infeasibleRow", C.c_int), ("infeasibleCol", C.c_int), ("stopReason", C.c_int),
                ("seconds", C.c_double)]


class Stats(C.Structure):
    _fields_ = [("kernelLaunches", C.c_int64), ("lastRunMs", C.c_double), ("lastRoundKernelMs", C.c_double),
                ("lastRoundKernelLaunches", C.c_int), ("lastRoundsTotal", C.c_int64),
                ("lastFromBase", C.c_int), ("lastGroupsBuilt", C.c_int), ("lastMode", C.c_int),
                ("lastHostSyncs", C.c_int), ("lastPairsWalked", C.c_int64), ("lastItems", C.c_int64)]


# every symbol include/cbp_capi.h declares (tests check that the library exports them all)
SYMBOLS = [
    "cbp_create", "cbp_destroy", "cbp_set_stream", "cbp_last_error", "cbp_get_stats", "cbp_synchronize",
    "cbp_matrix_upload", "cbp_matrix_free", "cbp_singletons", "cbp_propagate", "cbp_propagate_batch",
    "cbp_propagate_batch_delta", "cbp_batch_load_delta", "cbp_batch_run",
    "cbp_batch_fetch", "cbp_batch_fetch_dense", "cbp_row_activity", "cbp_sv_round_stamps", "cbp_singletons_noalloc",
    "cbp_p2p_export", "cbp_p2p_attach", "cbp_p2p_propagate",
]
P2P_HANDLE_BYTES = 64
ROUND_CAP = 1024

_lib = None


class CbpError(RuntimeError):
    pass


def load():
    """dlopen libcbp.so; raises if it has not been built (python __graft_entry__.py)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise CbpError(f"{LIB_PATH} is missing: build it with cbc_b200/csrc/build.sh (no CPU fallback exists)")
    lib = C.CDLL(LIB_PATH)
    vp, i32, i64, f64 = C.c_void_p, C.c_int, C.c_int64, C.c_double
    lib.cbp_create.restype = vp
    lib.cbp_create.argtypes = [i32]
    lib.cbp_destroy.argtypes = [vp]
    lib.cbp_destroy.restype = None
    lib.cbp_set_stream.argtypes = [vp, vp]
    lib.cbp_last_error.restype = C.c_char_p
    lib.cbp_last_error.argtypes = [vp]
    lib.cbp_get_stats.argtypes = [vp, C.POINTER(Stats)]
    lib.cbp_synchronize.argtypes = [vp]
    lib.cbp_matrix_upload.restype = vp
    lib.cbp_matrix_upload.argtypes = [vp, i32, i32, vp, vp, vp, vp, vp, vp, vp, vp, f64]
    lib.cbp_matrix_free.argtypes = [vp]
    lib.cbp_matrix_free.restype = None
    lib.cbp_singletons.argtypes = [vp, vp, vp, f64, C.POINTER(i32), C.POINTER(i32), i32]
    lib.cbp_propagate.argtypes = [vp, vp, vp, f64, i32, f64, C.POINTER(Result), vp, i32, i32]
    lib.cbp_propagate_batch.argtypes = [vp, i32, vp, vp, f64, i32, f64, vp, i32]
    lib.cbp_propagate_batch_delta.argtypes = [vp, vp, vp, i32, vp, vp, vp, f64, i32, f64, vp, i64, vp, vp, vp, vp, i32]
    lib.cbp_batch_load_delta.argtypes = [vp, vp, vp, i32, vp, vp, vp, i32]
    lib.cbp_batch_run.argtypes = [vp, f64, i32, f64]
    lib.cbp_batch_fetch.argtypes = [vp, vp, i64, vp, vp, vp, vp, i32]
    lib.cbp_batch_fetch_dense.argtypes = [vp, vp, vp, i32]
    lib.cbp_row_activity.argtypes = [vp, vp, vp, vp, vp, vp, vp, i32]
    lib.cbp_sv_round_stamps.argtypes = [vp, vp, i32]
    lib.cbp_p2p_export.argtypes = [vp, i32, vp]
    lib.cbp_p2p_attach.argtypes = [vp, i32, i32, vp]
    lib.cbp_p2p_propagate.argtypes = [vp, vp, vp, f64, i32, f64, i32, C.POINTER(Result), vp, i32, i32]
    lib.cbp_singletons_noalloc.argtypes = [i32, i32, vp, vp, vp, vp, vp, vp, vp, vp, f64, vp, vp, f64,
                                           C.POINTER(i32), C.POINTER(i32), vp]
    _lib = lib
    return lib


def _ptr(a):
    if a is None:
        return None
    if isinstance(a, int):
        return C.c_void_p(a)
    return C.c_void_p(a.ctypes.data)


NOALLOC_MAX_COLS = 262144


def singletons_noalloc(m, n, row_start, col, val, sense, rhs, rng, col_type, inf, lb, ub, tol=1e-7, stream=0):
    """cbp_singletons_noalloc: every array argument is a DEVICE pointer (int) on the current device; lb/ub are
    updated in place.  Returns (rc, nTightened, nFixed); rc is the library's return code (0 = ok)."""
    lib = load()
    nt, nf = C.c_int(0), C.c_int(0)
    rc = lib.cbp_singletons_noalloc(m, n, C.c_void_p(row_start), None, C.c_void_p(col), C.c_void_p(val),
                                    C.c_void_p(sense), C.c_void_p(rhs), C.c_void_p(rng), C.c_void_p(col_type),
                                    float(inf), C.c_void_p(lb), C.c_void_p(ub), float(tol), C.byref(nt), C.byref(nf),
                                    C.c_void_p(stream))
    return rc, nt.value, nf.value


class Context:
    def __init__(self, device=-1, stream=None):
        self.lib = load()
        self.h = self.lib.cbp_create(device)
        if not self.h:
            raise CbpError("cbp_create failed: " + (self.lib.cbp_last_error(None) or b"").decode())
        if stream is not None:
            self.lib.cbp_set_stream(self.h, C.c_void_p(stream))

    def check(self, rc, what):
        if rc != 0:
            raise CbpError(f"{what} failed ({rc}): " + (self.lib.cbp_last_error(self.h) or b"").decode())

    def stats(self):
        s = Stats()
        self.lib.cbp_get_stats(self.h, C.byref(s))
        return {"kernelLaunches": s.kernelLaunches, "lastRunMs": s.lastRunMs,
                "lastRoundKernelMs": s.lastRoundKernelMs, "lastRoundKernelLaunches": s.lastRoundKernelLaunches,
                "lastRoundsTotal": s.lastRoundsTotal, "lastFromBase": s.lastFromBase,
                "lastGroupsBuilt": s.lastGroupsBuilt, "lastMode": s.lastMode, "lastHostSyncs": s.lastHostSyncs,
                "lastPairsWalked": s.lastPairsWalked, "lastItems": s.lastItems,
                "lastModeName": ("classic", "overlay-full-walk", "worklist")[s.lastMode] if 0 <= s.lastMode < 3 else "?"}

    def synchronize(self):
        self.check(self.lib.cbp_synchronize(self.h), "cbp_synchronize")

    def close(self):
        if self.h:
            self.lib.cbp_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def _res_dict(r):
    return {"feasible": bool(r.feasible), "rounds": r.rounds, "nChanged": r.nChanged,
            "infeasibleRow": r.infeasibleRow, "infeasibleCol": r.infeasibleCol,
            "stopReason": r.stopReason, "seconds": r.seconds}


class Matrix:
    """Device-resident copy of one instance's constraint data."""

    def __init__(self, ctx, inst):
        self.ctx = ctx
        self.m, self.n = int(inst.m), int(inst.n)
        rs = np.ascontiguousarray(inst.rowStart, np.int64)
        col = np.ascontiguousarray(inst.col, np.int32)
        val = np.ascontiguousarray(inst.val, np.float64)
        sense = np.ascontiguousarray(inst.sense, np.uint8)
        rhs = np.ascontiguousarray(inst.rhs, np.float64)
        rng = np.ascontiguousarray(inst.range, np.float64)
        ct = np.ascontiguousarray(inst.colType, np.int8)
        self.nnz = int(rs[-1])
        self.h = ctx.lib.cbp_matrix_upload(ctx.h, self.m, self.n, _ptr(rs), None, _ptr(col), _ptr(val),
                                           _ptr(sense), _ptr(rhs), _ptr(rng), _ptr(ct), float(inst.inf))
        if not self.h:
            raise CbpError("cbp_matrix_upload failed: " + (ctx.lib.cbp_last_error(ctx.h) or b"").decode())

    def close(self):
        if self.h:
            self.ctx.lib.cbp_matrix_free(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- single vector -----------------------------------------------------
    def singletons(self, lb, ub, tol=1e-7):
        lb = np.array(lb, np.float64)
        ub = np.array(ub, np.float64)
        nt, nf = C.c_int(0), C.c_int(0)
        self.ctx.check(self.ctx.lib.cbp_singletons(self.h, _ptr(lb), _ptr(ub), tol, C.byref(nt), C.byref(nf), 0),
                       "cbp_singletons")
        return nt.value, nf.value, lb, ub

    def propagate(self, lb, ub, tol=1e-7, max_rounds=0, time_limit=1e100, round_cap=1024):
        lb = np.array(lb, np.float64)
        ub = np.array(ub, np.float64)
        r = Result()
        rc_arr = np.full(round_cap, -1, np.int32)
        self.ctx.check(self.ctx.lib.cbp_propagate(self.h, _ptr(lb), _ptr(ub), tol, max_rounds, time_limit,
                                                  C.byref(r), _ptr(rc_arr), round_cap, 0), "cbp_propagate")
        out = _res_dict(r)
        done = r.rounds - (0 if r.feasible else 1)  # the failing round reports no count
        out["roundChanged"] = [int(x) for x in rc_arr[: min(done, round_cap)] if x >= 0]
        return out, lb, ub

    def propagate_ptrs(self, lb_ptr, ub_ptr, tol=1e-7, max_rounds=0, time_limit=1e100):
        """lb/ub are HOST pointers (ints), e.g. into pinned memory; updated in place."""
        r = Result()
        self.ctx.check(self.ctx.lib.cbp_propagate(self.h, C.c_void_p(lb_ptr), C.c_void_p(ub_ptr), tol, max_rounds,
                                                  time_limit, C.byref(r), None, 0, 0), "cbp_propagate")
        return _res_dict(r)

    def propagate_device(self, lb_ptr, ub_ptr, tol=1e-7, max_rounds=0, time_limit=1e100, raw_round=False):
        """lb/ub are device pointers (ints) to n doubles each, updated in place.  raw_round: see
        CBP_FLAG_RAW_ROUND (one round's unsettled proposals, for row-sharded runs)."""
        r = Result()
        flags = FLAG_DEVICE_PTRS | (FLAG_RAW_ROUND if raw_round else 0)
        self.ctx.check(self.ctx.lib.cbp_propagate(self.h, C.c_void_p(lb_ptr), C.c_void_p(ub_ptr), tol, max_rounds,
                                                  time_limit, C.byref(r), None, 0, flags),
                       "cbp_propagate")
        return _res_dict(r)

    # -- rows sharded over the GPUs of one box (peer memory) ----------------
    def p2p_export(self, world):
        """This rank's state block (for a group of `world` ranks) as a CUDA IPC handle (bytes)."""
        h = (C.c_ubyte * P2P_HANDLE_BYTES)()
        self.ctx.check(self.ctx.lib.cbp_p2p_export(self.h, world, h), "cbp_p2p_export")
        return bytes(h)

    def p2p_attach(self, rank, world, handles):
        """handles: list of `world` handle byte strings in rank order."""
        blob = b"".join(handles)
        assert len(blob) == world * P2P_HANDLE_BYTES
        self.ctx.check(self.ctx.lib.cbp_p2p_attach(self.h, rank, world, blob), "cbp_p2p_attach")

    def p2p_propagate(self, lb, ub, row_offset, tol=1e-7, max_rounds=0, time_limit=1e100, device_ptrs=False):
        """Collective over the attached group.  lb/ub: writable contiguous float64 numpy arrays (or device
        pointers) holding the input bounds; they receive the final bounds."""
        r = Result()
        per_round = np.zeros(ROUND_CAP, np.int32)
        if device_ptrs:
            a, b, flags = C.c_void_p(lb), C.c_void_p(ub), FLAG_DEVICE_PTRS
        else:
            a, b, flags = _ptr(lb), _ptr(ub), 0
        self.ctx.check(self.ctx.lib.cbp_p2p_propagate(self.h, a, b, tol, max_rounds, time_limit, row_offset,
                                                      C.byref(r), _ptr(per_round), ROUND_CAP, flags),
                       "cbp_p2p_propagate")
        d = _res_dict(r)
        done = d["rounds"] - (0 if d["feasible"] else 1)
        d["roundChanged"] = per_round[:max(0, min(done, ROUND_CAP))].tolist()
        return d

    def round_stamps(self, max_rounds=64):
        """(rounds, 4) ns since loop start: after rows, after long rows, after settle, after decision."""
        out = np.zeros((max_rounds, 4), np.uint32)
        k = self.ctx.lib.cbp_sv_round_stamps(self.h, _ptr(out), max_rounds)
        if k < 0:
            self.ctx.check(k, "cbp_sv_round_stamps")
        return out[:k]

    def row_activity(self, lb, ub):
        lb = np.ascontiguousarray(lb, np.float64)
        ub = np.ascontiguousarray(ub, np.float64)
        m = self.m
        mn, mx = np.zeros(m), np.zeros(m)
        im, ix = np.zeros(m, np.int32), np.zeros(m, np.int32)
        self.ctx.check(self.ctx.lib.cbp_row_activity(self.h, _ptr(lb), _ptr(ub), _ptr(mn), _ptr(mx), _ptr(im),
                                                     _ptr(ix), 0), "cbp_row_activity")
        return mn, mx, im, ix

    # -- batch -------------------------------------------------------------
    def propagate_batch(self, lb, ub, tol=1e-7, max_rounds=0, time_limit=1e100):
        lb = np.array(lb, np.float64, order="C")
        ub = np.array(ub, np.float64, order="C")
        B = lb.shape[0]
        res = np.zeros(B, RESULT_DTYPE)
        self.ctx.check(self.ctx.lib.cbp_propagate_batch(self.h, B, _ptr(lb), _ptr(ub), tol, max_rounds, time_limit,
                                                        _ptr(res), 0), "cbp_propagate_batch")
        return res, lb, ub

    def propagate_batch_delta(self, base_lb, base_ub, start, var, bnd, tol=1e-7, max_rounds=0, time_limit=1e100,
                              upd_cap=None):
        """Returns (results, updStart, updCol, updLb, updUb)."""
        self.batch_load_delta(base_lb, base_ub, start, var, bnd)
        self.batch_run(tol, max_rounds, time_limit)
        return self.batch_fetch(upd_cap)

    def batch_load_delta(self, base_lb, base_ub, start, var, bnd):
        base_lb = np.ascontiguousarray(base_lb, np.float64)
        base_ub = np.ascontiguousarray(base_ub, np.float64)
        start = np.ascontiguousarray(start, np.int64)
        var = np.ascontiguousarray(var, np.uint32)
        bnd = np.ascontiguousarray(bnd, np.float64)
        self._B = len(start) - 1
        self.ctx.check(self.ctx.lib.cbp_batch_load_delta(self.h, _ptr(base_lb), _ptr(base_ub), self._B, _ptr(start),
                                                         _ptr(var), _ptr(bnd), 0), "cbp_batch_load_delta")

    def batch_run(self, tol=1e-7, max_rounds=0, time_limit=1e100):
        self.ctx.check(self.ctx.lib.cbp_batch_run(self.h, tol, max_rounds, time_limit), "cbp_batch_run")

    def batch_fetch(self, upd_cap=None):
        B = self._B
        res = np.zeros(B, RESULT_DTYPE)
        start = np.zeros(B + 1, np.int64)
        # (the size the last fetch needed is remembered: a batch that comes back does not ask twice)
        cap = int(upd_cap) if upd_cap is not None else max(1024, 4 * B, getattr(self, "_upd_cap_hint", 0))
        while True:
            col = np.empty(cap, np.int32)  # (only the first start[B] entries are written and returned)
            lb = np.empty(cap)
            ub = np.empty(cap)
            rc = self.ctx.lib.cbp_batch_fetch(self.h, _ptr(res), cap, _ptr(start), _ptr(col), _ptr(lb), _ptr(ub), 0)
            if rc == ERR_CAPACITY and upd_cap is None:
                cap = int(start[B]) + int(start[B]) // 8 + 64
                self._upd_cap_hint = cap
                continue
            self.ctx.check(rc, "cbp_batch_fetch")
            k = int(start[B])
            return res, start, col[:k], lb[:k], ub[:k]

    def batch_fetch_results(self):
        res = np.zeros(self._B, RESULT_DTYPE)
        self.ctx.check(self.ctx.lib.cbp_batch_fetch(self.h, _ptr(res), 0, None, None, None, None, 0),
                       "cbp_batch_fetch")
        return res

    def batch_fetch_dense(self):
        lb = np.empty((self._B, self.n))
        ub = np.empty((self._B, self.n))
        self.ctx.check(self.ctx.lib.cbp_batch_fetch_dense(self.h, _ptr(lb), _ptr(ub), 0), "cbp_batch_fetch_dense")
        return lb, ub
