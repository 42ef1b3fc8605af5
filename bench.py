#!/usr/bin/env python
"""bench.py — bound-propagation throughput on B200 (metric of BASELINE.json:
nnz·rounds·nodes processed per second, beside the HBM roofline and the CPU baseline).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--impl reference]

A step is one pass of the hot path over one batch of synthetic input that is already
resident in HBM.  Without --workload the HEADLINE is `c2` — BASELINE.json's configs[1]: the
fast0507-shaped set cover at a deep node (99.8 % of the columns outside the planted cover fixed
at 0, node bounds = the propagated root), 4096 strong-branching bound vectors per GPU, a share of
which tighten and go through a second round — and the same invocation also times, as
sub-results under "workloads":
    c2q  the same shape at a shallow node where no arm tightens (round 1's headline; kept as a label)
    c3   1M x 1M, 10M nnz mixed-integer, single-vector fixpoint
    c3b  the north-star target shape: a batch of dive vectors on that 1M x 1M matrix
    c4   knapsack/assignment 200k x 500k, 8,192 dive vectors per GPU (the config's share of 65,536 over 8 GPUs)
    c5   power-law rows with 100k-nnz dense rows, single-vector fixpoint
Under torchrun (N > 1) the node batches are sharded (one process per GPU, no data-path collective,
weak scaling) and the sub-results are c4 (sharded) and c3_rowshard: the c3 fixpoint with the ROWS
sharded over the N GPUs and the per-round exchange inside the persistent kernel (strong scaling).

Every workload is measured twice: `sustained` = steps repeated back to back for >= 1 s, and the
contract's EXACTLY K steps bracketed by CUDA events (`value`, `ms_per_step`); `e2e` is the same
call through the C ABI with HOST buffers (copies inside the timed region).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from cbc_b200 import instances as I  # noqa: E402

METRIC = "nnz*rounds*nodes processed/sec"
UNIT = "nnz*rounds*nodes/s"
SUBS_1GPU = ["c2q", "c3", "c3b", "c4", "c5"]
SUBS_NGPU = ["c4", "c3_rowshard"]
KEYS = ("feasible", "rounds", "nChanged", "infeasibleRow", "infeasibleCol", "stopReason")


# ------------------------------------------------------------------ workloads
def make_workload(name, rank=0, scale=1.0, root_propagate=None):
    """dict(inst, kind, lb, ub, deltas or None, B, desc).  `root_propagate(inst) -> (lb, ub)` gives the node
    bounds for the workloads whose base is a propagated root (the GPU arm uses the CUDA path for it, the
    reference arm the CPU oracle)."""
    if name in ("c2", "c2q"):
        deep = name == "c2"
        inst = I.fast0507_shaped(fix_frac=0.998 if deep else 0.3)
        lb, ub = (root_propagate(inst) if deep else (inst.lb, inst.ub))
        B = max(64, int(4096 * scale))
        s, v, b = I.strong_branch_deltas(inst, lb, ub, B // 2, seed=I.SEED0 + 100 + rank)
        what = ("deep node (99.8% of the non-cover columns fixed at 0, node bounds = propagated root)" if deep
                else "shallow node (30% of the columns fixed at 0: no arm tightens)")
        return dict(name=name, inst=inst, kind="batch", lb=lb, ub=ub, deltas=(s, v, b), B=B,
                    desc=f"{name}: fast0507-shaped set cover {inst.m}x{inst.n}, {inst.nnz} nnz, {what}, "
                         f"{B} strong-branching bound vectors per GPU (base + one bound change each)")
    if name == "c4":
        inst = I.knapsack_assignment(root_fixpoint=True)
        lb, ub = root_propagate(inst)  # (a handful of one-column assignment rows still fix their column at the root)
        B = max(64, int(8192 * scale))
        s, v, b = I.dive_deltas(inst, lb, ub, B, max_depth=64, seed=I.SEED0 + 200 + rank)
        return dict(name=name, inst=inst, kind="batch", lb=lb, ub=ub, deltas=(s, v, b), B=B,
                    desc=f"c4: knapsack/assignment {inst.m}x{inst.n}, {inst.nnz} nnz, node bounds = propagated root, "
                         f"{B} dive vectors per GPU (1..64 cumulative fixings)")
    if name == "c3b":
        sz = max(1000, int(1_000_000 * scale))
        inst = I.mixed_bigm(m=sz, n=sz)
        lb, ub = root_propagate(inst)
        B = max(64, int(1024 * scale))
        s, v, b = I.dive_deltas(inst, lb, ub, B, max_depth=16, seed=I.SEED0 + 300 + rank)
        return dict(name=name, inst=inst, kind="batch", lb=lb, ub=ub, deltas=(s, v, b), B=B,
                    desc=f"c3b: mixed-integer {inst.m}x{inst.n}, {inst.nnz} nnz, node bounds = propagated root, "
                         f"{B} dive vectors per GPU (1..16 cumulative fixings)")
    if name in ("c3", "c3_rowshard"):
        sz = max(1000, int(1_000_000 * scale))
        inst = I.mixed_bigm(m=sz, n=sz)
        return dict(name=name, inst=inst, kind="single", lb=inst.lb, ub=inst.ub, deltas=None, B=1,
                    desc=f"c3: mixed-integer {inst.m}x{inst.n}, {inst.nnz} nnz, single-vector fixpoint")
    if name == "c3s":  # same shape as c3, but 90% of a row's columns lie near its diagonal position
        sz = max(1000, int(1_000_000 * scale))
        inst = I.mixed_bigm(m=sz, n=sz, locality=0.9, band=2048)
        return dict(name=name, inst=inst, kind="single", lb=inst.lb, ub=inst.ub, deltas=None, B=1,
                    desc=f"c3s: mixed-integer {inst.m}x{inst.n}, {inst.nnz} nnz, staircase structure, single-vector fixpoint")
    if name == "c5":
        inst = I.power_law(n=max(1000, int(1_000_000 * scale)), nnz_target=max(10000, int(5_000_000 * scale)),
                           dense_len=max(100, int(100_000 * scale)))
        return dict(name=name, inst=inst, kind="single", lb=inst.lb, ub=inst.ub, deltas=None, B=1,
                    desc=f"c5: power-law rows {inst.m}x{inst.n}, {inst.nnz} nnz, 8 rows of "
                         f"{max(100, int(100_000 * scale))} nnz, single-vector fixpoint")
    if name == "c1":
        inst = I.set_cover(fix_frac=0.93)
        return dict(name=name, inst=inst, kind="single", lb=inst.lb, ub=inst.ub, deltas=None, B=1,
                    desc=f"c1: set cover {inst.m}x{inst.n}, {inst.nnz} nnz, root bound vector")
    raise SystemExit(f"unknown workload {name}")


def config_for(work, world=1):
    """The `config` object of the JSON line: the same for the GPU arm and the reference arm."""
    batch = work["kind"] == "batch"
    cfg = {"workload": work["desc"], "vectors_per_gpu": work["B"]}
    if work["name"] == "c3_rowshard":
        cfg["workload"] += f", rows sharded over {world} GPUs"
    cfg["l2"] = ("matrix, shared node bounds and per-vector diff cells are L2-resident by design (per-vector bound arrays "
                 "are never materialised), so there is no larger-than-L2 input to stream; the state every step "
                 "rebuilds is re-read each round" if batch else
                 ("inputs larger than L2 (matrix + bounds re-streamed every round)"
                  if 12 * work["inst"].nnz + 33 * work["inst"].n > 126e6 else
                  "working set fits L2; matrix re-read every round by design"))
    return cfg


def algorithmic_bytes(inst, rounds_launched, vector_rounds, batch):
    """per round 12*nnz + 20*m + 33*n (single vector) or
    12*nnz + 20*m + n + 32*n*B (batch; B = vectors active in that round, so the sum over
    rounds of B is the vector-rounds total)."""
    if batch:
        return rounds_launched * (12 * inst.nnz + 20 * inst.m + inst.n) + 32 * inst.n * vector_rounds
    return rounds_launched * (12 * inst.nnz + 20 * inst.m + 33 * inst.n)


# ------------------------------------------------------------------ clocks
class ClockSampler:
    """SM clock and throttle reasons sampled DURING the timed region (NVML, ~1 ms per sample;
    falls back to nvidia-smi if NVML is not importable)."""

    def __init__(self, index):
        self.index, self.rows, self.stop = index, [], False
        self.t = threading.Thread(target=self._run, daemon=True)
        self.nvml = None
        try:
            import pynvml

            pynvml.nvmlInit()
            self.nvml = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
        except Exception:
            self.nvml = None

    def _sample(self):
        if self.nvml is not None:
            n = self.nvml
            sm = n.nvmlDeviceGetClockInfo(self.h, n.NVML_CLOCK_SM)
            mx = n.nvmlDeviceGetMaxClockInfo(self.h, n.NVML_CLOCK_SM)
            r = n.nvmlDeviceGetCurrentClocksEventReasons(self.h) if hasattr(n, "nvmlDeviceGetCurrentClocksEventReasons") \
                else n.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
            names = []
            for bit, name in ((0x8, "hw_slowdown"), (0x40, "hw_thermal_slowdown"), (0x20, "sw_thermal_slowdown"),
                              (0x4, "sw_power_cap")):
                if r & bit:
                    names.append(name)
            return sm, mx, names
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        out = subprocess.run(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-i",
                              str(self.index)], capture_output=True, text=True, timeout=5).stdout
        f = [x.strip() for x in out.strip().split(",")]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        return int(f[0]), int(f[1]), [names[i] for i in range(4) if f[2 + i] == "Active"]

    def _run(self):
        while not self.stop:
            try:
                self.rows.append(self._sample())
            except Exception:
                pass
            time.sleep(0.002 if self.nvml is not None else 0.1)

    def __enter__(self):
        self.t.start()
        return self

    def __exit__(self, *a):
        self.stop = True
        self.t.join(timeout=6)

    def summary(self):
        sm = [r[0] for r in self.rows]
        mx = [r[1] for r in self.rows]
        reasons = sorted({x for r in self.rows for x in r[2]})
        return {"sm_mhz": int(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


# ------------------------------------------------------------------ CPU arms (oracle/ — the only place it is used here)
def _cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def cpu_run(work, threads, target_seconds, want_bounds=False, max_vectors=None):
    """Time the CPU restatement (oracle/) on a bounded sample of the workload: the first k vectors of a batch, or
    the first R rounds of a single-vector fixpoint (R chosen so that the sample takes about target_seconds)."""
    from oracle import oracle_py as O

    inst = work["inst"]
    prob = O.Problem(inst)
    if work["kind"] == "single":
        t0 = time.perf_counter()
        r1, _, _ = O.propagate(prob, work["lb"], work["ub"], max_rounds=1)
        t1 = max(time.perf_counter() - t0, 1e-6)
        cap = max(1, int(target_seconds / t1))
        t0 = time.perf_counter()
        r, lb, ub = O.propagate(prob, work["lb"], work["ub"], max_rounds=cap)
        dt = time.perf_counter() - t0
        whole = r["stopReason"] != 2
        return dict(units=inst.nnz * r["rounds"], seconds=dt, threads=1, results=r, lb=lb, ub=ub, max_rounds=cap,
                    sample=(f"whole instance, {r['rounds']} rounds" if whole else
                            f"first {cap} rounds of the same fixpoint run (maxRounds={cap})"))
    s, v, b = work["deltas"]
    B = len(s) - 1
    probe = min(B, max(threads, 4))
    t0 = time.perf_counter()
    O.propagate_batch_delta(prob, work["lb"], work["ub"], s[: probe + 1], v[: s[probe]], b[: s[probe]],
                            threads=threads, want_bounds=False)
    dt = max(time.perf_counter() - t0, 1e-6)
    k = int(min(B, max(probe, probe * target_seconds / dt)))
    if max_vectors:
        k = min(k, max_vectors)
    if want_bounds:  # the dense bounds of the sample must fit comfortably in host memory
        k = max(1, min(k, int(2e9 // (16 * inst.n))))
    t0 = time.perf_counter()
    res, lbo, ubo = O.propagate_batch_delta(prob, work["lb"], work["ub"], s[: k + 1], v[: s[k]], b[: s[k]],
                                            threads=threads, want_bounds=want_bounds)
    dt = time.perf_counter() - t0
    return dict(units=inst.nnz * int(res["rounds"].sum()), seconds=dt, threads=threads, results=res, lb=lbo, ub=ubo,
                sample=f"first {k} of {B} vectors of the same batch", k=k)


def _close(a, b, isint):
    """integer columns exact, continuous within 1e-9 relative / 1e-12 absolute (north_star)."""
    a, b = np.asarray(a), np.asarray(b)
    ok = (a == b) | (~isint & (np.abs(a - b) <= 1e-12 + 1e-9 * np.maximum(np.abs(a), np.abs(b))))
    return int(np.count_nonzero(~ok))


def compare_with_cpu(work, c, gpu):
    """Bounds-level agreement of the CUDA path with the CPU leg on the CPU leg's sample.
    gpu: for a batch (results, updStart, updCol, updLb, updUb) of the whole batch; for a single vector
    (result, lb, ub) of a run with the same maxRounds as the CPU sample."""
    inst = work["inst"]
    isint = inst.colType != 0
    if work["kind"] == "single":
        r, lb, ub = gpu
        o = c["results"]
        return {"vectors": 1, "mismatched_verdicts": sum(int(r[k] != o[k]) for k in KEYS),
                "mismatched_bounds": _close(lb, c["lb"], isint) + _close(ub, c["ub"], isint),
                "compared": "verdict fields + all column bounds (integer exact, continuous 1e-9 rel / 1e-12 abs)"}
    res, us, uc, ul, uu = gpu
    o, k = c["results"], c["k"]
    bad_v = sum(int(np.count_nonzero(np.asarray(res[key][:k]) != np.asarray(o[key]))) for key in KEYS)
    bad_b = 0
    s, v, b = work["deltas"]
    for i in range(k):
        lb, ub = np.array(work["lb"]), np.array(work["ub"])
        for q in range(s[i], s[i + 1]):
            j = int(v[q] & I.COL_MASK)
            if v[q] & I.UPPER_FLAG:
                ub[j] = b[q]
            else:
                lb[j] = b[q]
        if res["feasible"][i]:  # (a failing vector lists no updates; the oracle leaves its bounds at the last good round)
            cols = uc[us[i]:us[i + 1]]
            lb[cols] = ul[us[i]:us[i + 1]]
            ub[cols] = uu[us[i]:us[i + 1]]
            bad_b += _close(lb, c["lb"][i], isint) + _close(ub, c["ub"][i], isint)
    return {"vectors": int(k), "mismatched_verdicts": int(bad_v), "mismatched_bounds": int(bad_b),
            "compared": "verdict fields of every sampled vector + the update lists (cols, lb, ub) applied to base+diffs "
                        "against the oracle's dense bounds (integer exact, continuous 1e-9 rel / 1e-12 abs)"}


def run_reference(args, rank):
    """--impl reference: the reference's CPU path.  The reference itself cannot be built here (its
    arithmetic lives in CoinUtils/Osi, absent), so this times the CPU restatement in oracle/ with all
    the host threads it can use, on the GPU arm's headline config."""
    if rank != 0:
        return
    from oracle import oracle_py as O

    def root(inst):
        r, lb, ub = O.propagate(O.Problem(inst), inst.lb, inst.ub)
        return lb, ub

    name = args.workload or "c2"
    work = make_workload(name, 0, args.scale, root)
    # all the cores this process may run on — not omp_get_max_threads(): torchrun exports OMP_NUM_THREADS=1
    threads = max(_cores(), O.max_threads()) if work["kind"] == "batch" else 1
    per_step = max(2.0, min(20.0, 120.0 / max(1, args.steps + args.warmup)))
    for _ in range(args.warmup):
        cpu_run(work, threads, per_step / 4)
    units, secs, sample = 0, 0.0, ""
    for _ in range(args.steps):
        r = cpu_run(work, threads, per_step)
        units += r["units"]
        secs += r["seconds"]
        sample = r["sample"]
    val = units / secs
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * secs / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": config_for(work, args.gpus),
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


# ------------------------------------------------------------------ GPU arm
class Bench:
    def __init__(self, args):
        import torch
        import torch.distributed as dist

        from cbc_b200 import capi

        self.torch, self.dist, self.capi = torch, dist, capi
        self.args = args
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py needs a CUDA device; there is no CPU fallback (use --impl reference for the CPU arm)")
        torch.cuda.set_device(self.local)
        if self.world > 1:
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local))
        self.ctx = capi.Context(self.local, stream=torch.cuda.current_stream().cuda_stream)
        self.peaks = {}
        pk = os.path.join(ROOT, "MEASURED_PEAKS.json")
        if os.path.exists(pk):
            self.peaks = json.load(open(pk))
        self.props = torch.cuda.get_device_properties(self.local)

    def barrier(self):
        self.torch.cuda.synchronize()
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def root_propagate(self, inst):
        """node bounds = the propagated root, computed by the CUDA path itself"""
        mat = self.capi.Matrix(self.ctx, inst)
        r, lb, ub = mat.propagate(inst.lb, inst.ub)
        mat.close()
        if not r["feasible"]:
            raise SystemExit("root propagation of a bench instance reported infeasible")
        return lb, ub

    # -------------------------------------------------------------- one workload
    def run(self, name, steps, warmup, sustain_s, with_cpu):
        torch, capi, ctx = self.torch, self.capi, self.ctx
        rank, world = self.rank, self.world
        t_setup = time.perf_counter()
        work = make_workload(name, rank, self.args.scale, self.root_propagate)
        inst = work["inst"]
        batch = work["kind"] == "batch"
        rowshard = name == "c3_rowshard" and world > 1
        sh = None
        if rowshard:
            from cbc_b200 import multigpu as MG

            sh = MG.PeerShardedMatrix(ctx, inst)
            mat = sh.mat
        else:
            mat = capi.Matrix(ctx, inst)
        info = {}

        def pinned(a, dt):
            return torch.from_numpy(np.ascontiguousarray(a, dt)).pin_memory()

        if batch:
            s, v, b = work["deltas"]
            mat.batch_load_delta(work["lb"], work["ub"], s, v, b)
            ctx.synchronize()

            def step():
                mat.batch_run()
                st = ctx.stats()
                info["stats"] = st
                return st["lastRoundsTotal"], st["lastRoundKernelMs"], st["lastRoundKernelLaunches"]

            lb_h, ub_h = pinned(work["lb"], np.float64).numpy(), pinned(work["ub"], np.float64).numpy()
            s_h, v_h, b_h = pinned(s, np.int64).numpy(), pinned(v, np.uint32).numpy(), pinned(b, np.float64).numpy()

            def e2e_step():
                out = mat.propagate_batch_delta(lb_h, ub_h, s_h, v_h, b_h)
                info["last"] = out
                return int(out[0]["rounds"].sum()), out[2].size

            h2d = 16 * inst.n + 8 * len(s) + 12 * len(v)
        else:
            dlb0 = torch.tensor(work["lb"], device="cuda")
            dub0 = torch.tensor(work["ub"], device="cuda")
            dlb, dub = dlb0.clone(), dub0.clone()
            lb_pin0, ub_pin0 = pinned(work["lb"], np.float64), pinned(work["ub"], np.float64)
            lb_pin, ub_pin = lb_pin0.clone().pin_memory(), ub_pin0.clone().pin_memory()

            def step():
                dlb.copy_(dlb0)
                dub.copy_(dub0)
                if rowshard:
                    r = sh.propagate_device(dlb.data_ptr(), dub.data_ptr())
                else:
                    r = mat.propagate_device(dlb.data_ptr(), dub.data_ptr())
                st = ctx.stats()
                info["stats"] = st
                info["last_r"] = r
                # row-sharded: every rank does all rounds; count them once so the SUM over ranks is the job's
                return (r["rounds"] if (not rowshard or rank == 0) else 0), st["lastRoundKernelMs"], r["rounds"]

            def e2e_step():
                lb_pin.copy_(lb_pin0)
                ub_pin.copy_(ub_pin0)
                if rowshard:
                    r = mat.p2p_propagate(lb_pin.numpy(), ub_pin.numpy(), sh.r0)
                else:
                    r = mat.propagate_ptrs(lb_pin.data_ptr(), ub_pin.data_ptr())
                info["last"] = (r, lb_pin.numpy(), ub_pin.numpy())
                return (r["rounds"] if (not rowshard or rank == 0) else 0), 0

            h2d = 16 * inst.n
        setup_s = time.perf_counter() - t_setup

        for _ in range(max(3, warmup)):
            step()
        self.barrier()
        # sustained: back-to-back steps for >= sustain_s (every rank the same count: rank 0's clock decides)
        n_sus, sus_units = 0, 0.0
        with ClockSampler(self.local) as clk:
            t0 = time.perf_counter()
            first = time.perf_counter()
            vr, _, _ = step()
            torch.cuda.synchronize()
            one = max(time.perf_counter() - first, 1e-5)
            n_more = int(min(20000, max(0, sustain_s / one - 1)))
            if world > 1:
                tn = torch.tensor([n_more], device="cuda")
                self.dist.broadcast(tn, 0)
                n_more = int(tn.item())
            sus_units += vr
            for _ in range(n_more):
                sus_units += step()[0]
            torch.cuda.synchronize()
            sus_dt = time.perf_counter() - t0
            n_sus = 1 + n_more
            self.barrier()
            # the contract's region: exactly `steps` steps between two events
            l0 = ctx.stats()["kernelLaunches"]
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            vec_rounds = kernel_ms = 0.0
            round_launches = 0
            ev0.record()
            for _ in range(steps):
                vr, kms, kl = step()
                vec_rounds += vr
                kernel_ms += kms
                round_launches += kl
            ev1.record()
            self.barrier()
        ms = ev0.elapsed_time(ev1)
        launches = ctx.stats()["kernelLaunches"] - l0
        units = float(inst.nnz) * vec_rounds
        t = torch.tensor([ms, units, float(launches), sus_dt, float(inst.nnz) * sus_units], dtype=torch.float64,
                         device="cuda")
        if world > 1:
            tmax = t.clone()
            self.dist.all_reduce(tmax, op=self.dist.ReduceOp.MAX)
            self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM)
            ms, units, launches, sus_dt, sus_u = float(tmax[0]), float(t[1]), float(t[2]), float(tmax[3]), float(t[4])
        else:
            sus_u = float(t[4])
        value = units / (ms * 1e-3)

        # ---- end to end through the C ABI with HOST (pinned) buffers, copies inside the timed region
        e2e_steps = max(2, min(steps, 5))
        for _ in range(3):  # warm-up: the first calls size the list buffers and switch the runs to writing the lists
            e2e_step()
        self.barrier()
        t0 = time.perf_counter()
        e_units, n_upd = 0, 0
        for _ in range(e2e_steps):
            rr, nu = e2e_step()
            e_units += inst.nnz * rr
            n_upd = nu
        torch.cuda.synchronize()
        e_dt = time.perf_counter() - t0
        d2h = (32 * work["B"] + 8 * (work["B"] + 1) + 20 * n_upd) if batch else (16 * inst.n + 64)
        e = torch.tensor([e_dt, float(e_units)], dtype=torch.float64, device="cuda")
        if world > 1:
            emax = e.clone()
            self.dist.all_reduce(emax, op=self.dist.ReduceOp.MAX)
            self.dist.all_reduce(e, op=self.dist.ReduceOp.SUM)
            e_dt, e_units = float(emax[0]), float(e[1])
        e2e_val = e_units / e_dt

        out = None
        if rank == 0:
            st = info.get("stats", {})
            out = {
                "value": value, "unit": UNIT, "steps": steps, "ms_per_step": ms / steps,
                "scaling": "weak" if batch else "strong",
                "sustained": {"value": sus_u / sus_dt, "steps": n_sus, "seconds": sus_dt},
                "config": config_for(work, world),
                "e2e": {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                        "steps": e2e_steps},
                "gpu_launches": int(launches), "clocks": clk.summary(), "setup_s": round(setup_s, 2),
            }
            out["work"] = self.work_stats(work, info, mat, batch)
            out["roofline"] = self.roofline(work, batch, kernel_ms, round_launches, vec_rounds, st, clk, rowshard)
            if with_cpu and world == 1:
                out["cpu_baseline"] = self.cpu_leg(work, mat, info, headline=(name == (self.args.workload or "c2")))
        mat.close() if sh is None else sh.mat.close()
        return out

    def work_stats(self, work, info, mat, batch):
        """what the timed step did: rounds / changed columns / infeasible share of its vectors"""
        if batch:
            res = mat.batch_fetch_results()
            B = len(res)
            st = info.get("stats", {})
            w = {"rounds_per_vector": {"mean": float(res["rounds"].mean()), "max": int(res["rounds"].max())},
                 "changed_per_vector": {"mean": float(res["nChanged"].mean()), "max": int(res["nChanged"].max())},
                 "vectors_with_more_than_one_round": float((res["rounds"] > 1).mean()),
                 "vectors_with_updates": float((res["nChanged"] > 0).mean()),
                 "infeasible_frac": float((res["feasible"] == 0).mean()), "vectors": B,
                 "step": "one cbp_batch_run on the resident batch: all rounds on the device, every vector's verdict and "
                         "counts on the host when it returns; the vectors' update lists are written and copied in the "
                         "e2e leg (cbp_propagate_batch_delta), not here"}
            for k in ("lastMode", "lastPairsWalked", "lastHostSyncs"):
                if k in st:
                    w[k] = st[k]
            if st.get("lastPairsWalked"):
                w["pairs_walked_over_nnz_x_vector_rounds"] = st["lastPairsWalked"] / max(
                    1.0, float(work["inst"].nnz) * float(res["rounds"].sum()))
            return w
        r = info.get("last_r", {})
        return {"rounds": r.get("rounds"), "changed_total": r.get("nChanged"), "feasible": r.get("feasible")}

    def roofline(self, work, batch, kernel_ms, round_launches, vec_rounds, st, clk, rowshard):
        inst = work["inst"]
        peak = float(self.peaks.get("hbm_gbs", 6650.0))
        if rowshard:
            peak *= self.world  # the whole matrix is streamed once per round by all ranks together
        alg = algorithmic_bytes(inst, round_launches, vec_rounds, batch)
        achieved = alg / (kernel_ms * 1e-3) / 1e9 if kernel_ms > 0 else 0.0
        mode = st.get("lastModeName", "")
        rf = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
              "traffic": None,
              "peak_source": "MEASURED_PEAKS.json hbm_gbs (measured copy)" if self.peaks else "fallback 6650 GB/s",
              "kernel": (st.get("lastKernelName") or ("batch round kernels" if batch else "sv_fixpoint_kernel (per round)")),
              "kernel_ms_per_launch": kernel_ms / max(1, round_launches),
              "algorithmic_bytes_per_launch": alg / max(1, round_launches),
              "model": "SURVEY 8d: 12 nnz + 20 m + n + 32 n B per batch round" if batch else
                       "SURVEY 8d: 12 nnz + 20 m + 33 n per round"}
        if not batch:
            # The second ruler of the single-vector kernel: every nonzero is one 16-byte gather at an unrelated address,
            # i.e. one L1/TEX wavefront; a warp-wide gather of 32 different lines is replayed at ~2.07 cycles per
            # wavefront (B300_MICROARCH.md, "L1tex wavefront queue"; 1.0 across instructions).  Rounds that are
            # event-driven gather far fewer than nnz, so this too is an upper estimate of the work per round.
            sm_mhz = clk.summary().get("sm_mhz") or 1965
            wf_peak = self.props.multi_processor_count * sm_mhz * 1e6 / 2.07
            wf = float(inst.nnz) * round_launches
            if kernel_ms > 0:
                rf["l1tex_wavefront"] = {"achieved": wf / (kernel_ms * 1e-3) / 1e9, "peak": wf_peak / 1e9,
                                         "unit": "G gathers/s", "frac": wf / (kernel_ms * 1e-3) / wf_peak,
                                         "model": "one wavefront per nonzero and round (pass 1 only); peak = SMs x SM clock / 2.07"}
        if batch:
            sm_mhz = clk.summary().get("sm_mhz") or 1965
            fp64_peak = self.props.multi_processor_count * 64 * 2 * sm_mhz * 1e6 / 1e12  # DFMA: 64 per clock and SM
            pairs = st.get("lastPairsWalked")
            if not pairs and st.get("lastMode", 1) == 1 and st.get("lastFromBase"):
                # full-walk overlay rounds: every (nonzero, vector) pair of every launched round, vectors padded to the
                # 128 a warp carries
                pairs = float(inst.nnz) * ((work["B"] + 127) // 128 * 128) * round_launches
            if pairs and kernel_ms > 0:
                fl = 4.0 * float(pairs)  # two DFMA per (nonzero, vector) pair actually walked
                rf["fp64"] = {"achieved": fl / (kernel_ms * 1e-3) / 1e12, "peak": fp64_peak, "unit": "TFLOP/s",
                              "model": "2 DFMA per (nonzero, vector) pair walked; peak = SMs x 64 DFMA/clk x 2 x SM clock"}
                rf["fp64"]["frac"] = rf["fp64"]["achieved"] / fp64_peak
            if rf["frac"] > 1.2 or mode:
                rf["void"] = ("the 8d byte model assumes every vector's bounds are read and written every round; this "
                              "path keeps per-vector bounds as sparse cells over shared node bounds and evaluates only the "
                              "(row, vector) pairs a changed column touches, so the model is not a lower bound here: "
                              "see work.pairs_walked_over_nnz_x_vector_rounds, roofline.traffic (measured DRAM bytes) and "
                              "roofline.fp64")
        prof = os.path.join(ROOT, "profiles", f"traffic_{work['name']}.json")
        if os.path.exists(prof):
            try:
                rf["traffic"] = json.load(open(prof)).get("dram_bytes_per_launch")
            except Exception:
                pass
        return rf

    def cpu_leg(self, work, mat, info, headline):
        """the CPU restatement on a bounded sample, and its bounds against what the CUDA path returned"""
        batch = work["kind"] == "batch"
        threads = 1 if (headline or not batch) else _cores()
        c = cpu_run(work, threads, 12.0 if headline else 5.0, want_bounds=True, max_vectors=64 if batch else None)
        leg = {"value": c["units"] / c["seconds"], "unit": UNIT, "cores": threads, "kind": "port", "sample": c["sample"]}
        if batch:
            gpu = info.get("last")
        else:
            lb, ub = np.array(work["lb"]), np.array(work["ub"])
            r, lb, ub = mat.propagate(lb, ub, max_rounds=c["max_rounds"] if c["results"]["stopReason"] == 2 else 0)
            gpu = (r, lb, ub)
        if gpu is not None:
            leg["agrees_with_gpu"] = compare_with_cpu(work, c, gpu)
        return leg


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="cbc_b200")
    ap.add_argument("--workload", default=None, help="time only this workload (default: headline c2 + sub-results)")
    ap.add_argument("--scale", type=float, default=1.0)
    ap.add_argument("--sustain", type=float, default=1.0, help="seconds of back-to-back steps per workload")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline legs")
    ap.add_argument("--no-subs", action="store_true", help="headline only")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))

    if args.impl == "reference":
        run_reference(args, rank)
        return

    bn = Bench(args)
    W = max(3, args.warmup)
    head_name = args.workload or "c2"
    head = bn.run(head_name, args.steps, W, args.sustain, not args.no_cpu)
    subs = {}
    if args.workload is None and not args.no_subs:
        for name in (SUBS_1GPU if bn.world == 1 else SUBS_NGPU):
            try:
                sub_steps = args.steps if name != "c3b" else max(3, min(args.steps, 10))
                subs[name] = bn.run(name, sub_steps, 3, min(args.sustain, 1.0), not args.no_cpu)
            except Exception as e:  # a sub-result must not take the headline down with it
                subs[name] = {"error": f"{type(e).__name__}: {e}"}
                if bn.world > 1:
                    raise
    if bn.rank == 0:
        line = {
            "metric": METRIC, "value": head["value"], "unit": UNIT, "n_gpus": bn.world, "steps": args.steps, "warmup": W,
            "ms_per_step": head["ms_per_step"], "higher_is_better": True, "scaling": head["scaling"],
            "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": head["config"],
            "sustained": head["sustained"], "work": head["work"], "roofline": head["roofline"], "e2e": head["e2e"],
            "gpu_launches": head["gpu_launches"], "clocks": head["clocks"],
        }
        if "cpu_baseline" in head:
            line["cpu_baseline"] = head["cpu_baseline"]
        if subs:
            line["workloads"] = subs
        print(json.dumps(line))
    if bn.world > 1:
        bn.dist.destroy_process_group()


if __name__ == "__main__":
    main()
