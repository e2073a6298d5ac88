#!/usr/bin/env python
"""Benchmark of the LM refinement hot path (BASELINE.json metric: LM iterations/s and Jacobian
rows/s).  One "step" = one complete run of the LM loop (Calibrator.refineCalibrationParameters,
maxIters=100, until the reference's own termination rule fires) on the named workload.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload cfg3|cfg5|cfg4|itest|cfg1|cfg2]
                  [--impl reference]

value   device-resident: detections already in HBM, parameters uploaded before the timed region,
        timed with CUDA events on the library's stream (max over ranks).  L2 is flushed (256 MB
        write) before every step; inside a step the detections stay L2-resident across the LM
        iterations, as in production use.
ms_per_refinement   the robust figure beside `value`: the iteration COUNT of a noisy problem after
        the cost plateau is rounding noise in every implementation.
e2e     through the C-ABI with HOST buffers (page-locked CSR arrays): calib_create (H2D of the
        detections) + calib_lm_run (H2D of P, the loop, D2H of P) + calib_destroy, every step.
        e2e.python_api is the same through Calibrator.refineCalibrationParameters with the
        reference's list-of-arrays argument (adds the host-side flattening of the list);
        e2e.single_process (N > 1) drives all N GPUs from rank 0's process alone through the
        multi-device handle (calib_create_multi), the way the reference's one synchronous call would.
parity_vs_oracle (N = 1) / parity_vs_single (N > 1): the checker's verdict on the run that was timed.
secondary  cfg5 (the workload that shards: 10 000 views x 64x64, fisheye) at every N and, at N = 1,
        cfg4 (8192 independent 20-view calibrations): value, ms per refinement, kernel times, parity.
--impl reference  times the CPU restatement of the reference's path (oracle/calib_oracle.c,
        OpenMP over all host threads) on the same workload: the reference itself is Python/sympy,
        cannot travel to the GPU box and cannot hold the dense J of this workload (154 GB).  That arm
        imports nothing of the product (no camera_calibration_b200, no libcalib_b200.so).
"""
import argparse
import ctypes
import importlib
import json
import os
import subprocess
import sys
import threading
import time
import types

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

MAX_ITERS = 100
# Algorithmic FP64 work per point of the fused evaluation pass (DESIGN.md
# "Roofline"): F_model + 2 (L+6)(L+7) + 4 (L+6) + 4
FLOPS_PER_POINT = {0: 400 + 2 * 16 * 17 + 4 * 16 + 4, 1: 450 + 2 * 15 * 16 + 4 * 15 + 4}
COST_FLOPS_PER_POINT = 100    # residual-only sweep
BYTES_PER_POINT = 40          # X, Y, Z, u, v as float64, read once per sweep
# FP64 peak: not in MEASURED_PEAKS.json (the driver measures HBM and bf16 only), so the denominator is
# the library's own DFMA micro-benchmark taken in the same run; nominal for context.
FP64_NOMINAL_TFLOPS = 148 * 64 * 2 * 1.965e9 / 1e12


# ---- host-only pieces shared by both arms (no product library behind any of them) ------------------
def host_only():
    """synth.py and the two numpy modules it needs, loaded as a bare namespace WITHOUT executing the
    package's __init__: nothing that could load libcalib_b200.so gets imported (the reference arm must
    not touch the product)."""
    name = "calib_hostonly"
    if name not in sys.modules:
        pkg = types.ModuleType(name)
        pkg.__path__ = [os.path.join(ROOT, "camera_calibration_b200")]
        sys.modules[name] = pkg
    return importlib.import_module(name + ".synth"), importlib.import_module(name + ".mathutils")


def compose_P(mu, A, W, k):
    """src/calibrate.py:199-229: [alpha beta gamma uc vc k..., (rho_x rho_y rho_z[deg] tx ty tz) per view]"""
    A = np.asarray(A, dtype=np.float64)
    k = np.asarray(k, dtype=np.float64).ravel()
    W = np.asarray(W, dtype=np.float64)
    P = np.empty(5 + len(k) + 6 * len(W))
    P[:5] = A[0, 0], A[1, 1], A[0, 1], A[0, 2], A[1, 2]
    P[5:5 + len(k)] = k
    ext = P[5 + len(k):].reshape(-1, 6)
    ext[:, :3] = mu.rotationMatrixToEulerBatch(W[:, :3, :3])
    ext[:, 3:] = W[:, :3, 3]
    return P


def flatten_np(dets):
    counts = np.array([len(s) for s, m in dets], dtype=np.int64)
    offs = np.concatenate([[0], np.cumsum(counts)]).astype(np.int64)
    mp = np.ascontiguousarray(np.concatenate([m for s, m in dets], axis=0), dtype=np.float64).reshape(-1, 3)
    sp = np.ascontiguousarray(np.concatenate([s for s, m in dets], axis=0), dtype=np.float64).reshape(-1, 2)
    return offs, mp, sp


def load_workload(name, scale):
    synth, mu = host_only()
    kw = {"procs": min(host_threads(), 16)} if name == "cfg5" else {}
    c = synth.config(name, scale, **kw) if name in ("cfg3", "cfg5") else synth.config(name)
    if c["init"] is None:
        P0 = np.load(os.path.join(ROOT, "tests", "golden", name + ".npz"))["P0"].copy()
    else:
        P0 = compose_P(mu, *c["init"])
    return c, P0


def shard_bounds(counts, world):
    """contiguous view ranges balanced by point count (the rule of sharding.view_bounds / the library)"""
    cum = np.concatenate([[0], np.cumsum(counts)])
    total = cum[-1]
    b = [int(np.searchsorted(cum, total * r / world, side="left")) for r in range(world)] + [len(counts)]
    b[0] = 0
    for r in range(1, world + 1):
        b[r] = max(b[r], b[r - 1])
    return b


def shard_views(dets, P0, L, world, rank):
    """every rank keeps the L shared values and the extrinsics of its own views"""
    counts = np.array([len(s) for s, m in dets], dtype=np.int64)
    b = shard_bounds(counts, world)
    lo, hi = b[rank], b[rank + 1]
    Ploc = np.concatenate([P0[:L], P0[L + 6 * lo:L + 6 * hi]])
    return dets[lo:hi], Ploc, (lo, hi)


class ClockSampler:
    QUERY = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.stop = index, [], False
        self.t = threading.Thread(target=self._run, daemon=True)

    def _run(self):
        while not self.stop:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.QUERY,
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True,
                                     timeout=5).stdout.strip()
                if out:
                    self.rows.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            time.sleep(0.05)

    def __enter__(self):
        self.t.start()
        return self

    def __exit__(self, *a):
        self.stop = True
        self.t.join(timeout=6)

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        sm = sorted(float(r[0]) for r in self.rows)
        reasons = []
        for i, name in enumerate(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]):
            if any(r[3 + i].lower().startswith("active") for r in self.rows):
                reasons.append(name)
        pw = [float(r[2]) for r in self.rows if r[2].replace(".", "", 1).isdigit()]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(self.rows[0][1]), "reasons": reasons,
                "samples": len(self.rows), "power_w_max": max(pw) if pw else None}


def host_threads():
    """all the host threads this process may use (torchrun sets OMP_NUM_THREADS=1 for its children:
    the CPU arms ask for the cores explicitly)"""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def oracle_full_run(model, offs, mp, sp, P0):
    import oracle_c
    t0 = time.perf_counter()
    sse, P, it, tr = oracle_c.lm(model, P0, offs, mp, sp, MAX_ITERS, nthreads=host_threads())
    dt = time.perf_counter() - t0
    evals = 1 + int(tr[:-1, 3].sum())
    return it, evals, dt, (sse, P, tr)


def plateau_stop(tr):
    """first iteration whose trial cost is within 1e-10 relative of the current cost: from there on every
    accept/reject decision is the 13th digit of a sum"""
    idx = np.where(np.abs(tr[:, 1] - tr[:, 0]) <= 1e-10 * np.abs(tr[:, 0]))[0]
    return int(idx[0]) if len(idx) else len(tr)


def parity_block(P, sse, it, tr, Pr, sser, itr, trr, other):
    """the check of tests/mp/sharded_worker.py / tests/test_gpu_parity.py on the timed run: identical
    (accepted, lambda) sequence and costs up to the cost plateau of the checker's trace, converged
    parameters, final cost"""
    n = min(len(tr), len(trr))
    stop = min(n, plateau_stop(trr))
    same = bool(np.array_equal(tr[:stop, 3], trr[:stop, 3]) and np.array_equal(tr[:stop, 2], trr[:stop, 2]))
    cost_rel = float(np.max(np.abs(tr[:stop, 0] - trr[:stop, 0]) / np.abs(trr[:stop, 0]))) if stop else 0.0
    d = np.abs(np.asarray(P) - np.asarray(Pr))
    d = np.minimum(d, np.abs(d - 360.0))
    scale = np.maximum(np.abs(Pr), 1e-3)
    scale[2] = max(scale[2], abs(Pr[0]))            # the skew is measured at the scale of alpha (tests/conftest.py)
    rel = float((d / scale).max())
    sse_rel = float(abs(sse - sser) / abs(sser)) if sser else float(abs(sse - sser))
    # parameters: 1e-5 -- the two runs stop at different points of the flat valley past the cost plateau (their
    # post-plateau iteration counts differ, see iters / iters_*); up to the plateau they are the same run
    ok = bool(same and cost_rel <= 1e-9 and rel <= 1e-5 and sse_rel <= 1e-9)
    return {"ok": ok, "against": other, "pre_plateau_trace_equal": same, "pre_plateau_iterations": int(stop),
            "pre_plateau_cost_max_rel_diff": cost_rel, "max_rel_param_diff": rel, "sse_rel_diff": sse_rel,
            "iters": int(it), "iters_" + ("oracle" if "oracle" in other else "single"): int(itr)}


def workload_config(name, scale, c, offs, gpus):
    return {"workload": name + ("" if scale == 1.0 else f"@scale{scale}"),
            "model": "radtan" if c["model"] == 0 else "fisheye", "views": int(len(offs) - 1),
            "points": int(offs[-1]), "params": int((10 if c["model"] == 0 else 9) + 6 * (len(offs) - 1)),
            "max_iters": MAX_ITERS,
            "init": "perturbed ground truth, RandomState(12345)" if c["init"] is not None
                    else "the reference's linear stage (tests/golden)",
            "step": "one complete LM refinement (until the reference's termination rule fires)",
            "parallelism": f"views sharded over {gpus} GPU(s)" if gpus > 1 else "1 GPU",
            "l2": "flushed (256 MB write) before every timed step; within a step the detections stay "
                  "L2-resident across LM iterations, as in production use"}


def reference_python_seconds():
    """What the UNMODIFIED Python/sympy reference itself took on its own CPU-runnable cases: measured while
    the golden fixtures were generated (oracle/make_golden.py, build container, one core as the reference
    runs) and committed inside tests/golden/*.npz as ref_seconds = [initial guess, first Jacobian,
    refinement loop, Jacobian share of the loop].  The reference cannot travel to the GPU box, so these are
    quoted, not re-measured; BASELINE.md's 0.2-5 iterations/s sits beside them."""
    out = {}
    for case in ("cfg1", "cfg1n", "cfg2", "itest"):
        f = os.path.join(ROOT, "tests", "golden", case + ".npz")
        if not os.path.exists(f):
            continue
        z = np.load(f)
        if "ref_seconds" not in z.files:
            continue
        loop_s, its = float(z["ref_seconds"][2]), int(z["iters"])
        out[case] = {"loop_s": round(loop_s, 3), "iterations": its, "iterations_per_s": round(its / loop_s, 3),
                     "points": int(z["view_offsets"][-1]), "views": int(len(z["view_offsets"]) - 1)}
    return {"source": "tests/golden/*.npz ref_seconds (src/calibrate.py:117-171 run unmodified by "
                      "oracle/make_golden.py in the build container; not re-measured here)", "cases": out}


def time_reference_python(case="cfg1", timeout=240):
    """Runs the UNMODIFIED Python reference (vendored, git-ignored copy under baseline/_ref, see
    oracle/vendor_reference.py) on one of its own CPU-runnable cases in a subprocess on this box's host
    cores and returns its JSON (oracle/time_reference.py), or None when the copy or sympy is absent."""
    ref_root = os.path.join(ROOT, "baseline", "_ref")
    if not os.path.isdir(os.path.join(ref_root, "src")):
        return None
    env = dict(os.environ, CALIB_REF_ROOT=ref_root, CUDA_VISIBLE_DEVICES="")
    try:
        res = subprocess.run([sys.executable, os.path.join(ROOT, "oracle", "time_reference.py"), case],
                             capture_output=True, text=True, timeout=timeout, env=env)
        if res.returncode != 0:
            return {"case": case, "error": (res.stderr or "").strip().splitlines()[-1:]}
        return json.loads(res.stdout.strip().splitlines()[-1])
    except Exception as e:      # a missing sympy, a time-out: the quoted figures remain
        return {"case": case, "error": repr(e)}


def reference_python_block(product=None, case="cfg1"):
    """cpu_baseline.reference_python: the real reference timed HERE on `case` (kind "reference"), with the
    product's own run of the same case beside it when `product` (a callable returning sse, P, iterations,
    ms) is given."""
    r = time_reference_python(case)
    if r is None:
        return {"unavailable": "baseline/_ref not present (oracle/vendor_reference.py was not run where "
                               "/root/reference exists)", "quoted": reference_python_seconds()}
    if "error" in r:
        return {"unavailable": r["error"], "quoted": reference_python_seconds()}
    P_ref = np.asarray(r.pop("P"))
    out = {"kind": "reference", "cores": 1, "what": "Calibrator.refineCalibrationParameters of the unmodified "
           "Python/sympy reference (baseline/_ref), its own case " + case + ", this box's host cores", **r}
    if product is not None:
        sse, P, iters, ms = product()
        out["product_same_case"] = {
            "ms_per_refinement": ms, "iterations": int(iters), "sse": float(sse),
            "iterations_equal": bool(int(iters) == int(r["iterations"])),
            # Euler angles (degrees) compared modulo 360
            "max_abs_param_diff": float(np.max(np.minimum(np.abs(np.asarray(P) - P_ref),
                                                          np.abs(np.abs(np.asarray(P) - P_ref) - 360.0)))),
            "sse_abs_diff": abs(float(sse) - float(r["sse"])),
            "speedup_by_time": 1e3 * r["refine_call_s"] / ms if ms else None}
    return out


# ---- reference arm: the CPU restatement, no product import --------------------------------------------
def run_reference_arm(args):
    if int(os.environ.get("RANK", "0")) != 0:
        return
    import oracle_c
    c, P0 = load_workload(args.workload, args.scale)
    offs, mp, sp = flatten_np(c["dets"])
    for _ in range(min(args.warmup, 1)):
        oracle_c.lm(c["model"], P0, offs, mp, sp, 2)
    tot_it = evals = 0
    t0 = time.perf_counter()
    for _ in range(args.steps):
        it, ev, _, _ = oracle_full_run(c["model"], offs, mp, sp, P0)
        tot_it += it
        evals += ev
    dt = time.perf_counter() - t0
    val = tot_it / dt
    line = {
        "impl": "reference", "metric": "lm_iterations_per_sec", "value": val, "unit": "LM iterations/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
        "ms_per_refinement": 1e3 * dt / args.steps,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "jacobian_rows_per_sec": 2.0 * offs[-1] * evals / dt, "lm_iterations_per_step": tot_it / args.steps,
        "config": workload_config(args.workload, args.scale, c, offs, args.gpus),
        "cpu_baseline": {"value": val, "unit": "LM iterations/s", "cores": host_threads(), "kind": "port",
                         "sample": "each step = one complete LM run of the workload; C restatement of the "
                                   "reference's path (oracle/calib_oracle.c, OpenMP over all host threads); the "
                                   "Python/sympy reference cannot travel to this box nor hold this workload's dense J",
                         "reference_python_s": reference_python_seconds(),
                         "reference_python": None if args.no_python_reference else reference_python_block()},
        "e2e": {"value": val, "unit": "LM iterations/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "product_imported": bool("camera_calibration_b200" in sys.modules),
    }
    print(json.dumps(line), flush=True)


def pinned_copy(L, arr):
    """numpy view of a page-locked copy of arr (calib_host_alloc); returns (view, ptr)."""
    ptr = L.calib_host_alloc(max(arr.nbytes, 8))
    if not ptr:
        raise RuntimeError("calib_host_alloc failed")
    buf = (ctypes.c_char * max(arr.nbytes, 8)).from_address(ptr)
    view = np.frombuffer(buf, dtype=arr.dtype, count=arr.size).reshape(arr.shape)
    view[...] = arr
    return view, ptr


def fp64_peak(cc, device):
    t = cc.calibrate.fp64_peak_tflops(device)
    return {"tflops": t, "of_nominal": t / FP64_NOMINAL_TFLOPS, "nominal_tflops": FP64_NOMINAL_TFLOPS,
            "source": "DFMA micro-benchmark in this run (calib_measure_fp64_peak), builder-measured: FP64 is not in "
                      "MEASURED_PEAKS.json; nominal = 148 SMs x 64 lanes x 2 x 1.965 GHz"}


# ---- config 4: batch of independent calibrations, split over GPUs with no communication ----------
def load_batch(n_problems):
    synth, mu = host_only()
    probs = synth.batch_config(n_problems, procs=min(host_threads(), 32))
    P0 = [compose_P(mu, *p["init"]) for p in probs]
    return probs, P0


def batch_cpu_baseline(probs, P0, sample):
    """oracle port on the host cores: one problem per thread (ctypes releases the GIL)."""
    import oracle_c
    from concurrent.futures import ThreadPoolExecutor
    threads = host_threads()
    idx = list(range(min(sample, len(probs))))
    flat = [flatten_np(probs[j]["dets"]) for j in idx]

    def one(k):
        offs, mp, sp = flat[k]
        return oracle_c.lm(0, P0[idx[k]], offs, mp, sp, MAX_ITERS, nthreads=1)[2]
    one(0)
    t0 = time.perf_counter()
    with ThreadPoolExecutor(threads) as ex:
        its = list(ex.map(one, range(len(idx))))
    dt = time.perf_counter() - t0
    return sum(its) / dt, threads, len(idx), dt


def batch_config_block(n_all, probs, world):
    return {"workload": f"cfg4: {n_all} independent 20-view calibrations, 9x6 board, radtan, sigma 0.1 px",
            "model": "radtan", "problems": n_all, "views_per_problem": 20,
            "points": int(sum(len(s) for p in probs for s, m in p["dets"])), "max_iters": MAX_ITERS,
            "init": "perturbed ground truth, RandomState(problem index)",
            "step": "all problems refined to termination in one launch",
            "parallelism": f"problems split over {world} GPU(s), no communication",
            "l2": "flushed (256 MB write) before every timed step"}


def bench_batch(args, env, steps, light=False):
    """cfg4 on this process group.  Returns the bench line (rank 0) or None."""
    import torch
    import camera_calibration_b200 as cc
    from camera_calibration_b200 import _lib
    rank, world, local, dist = env["rank"], env["world"], env["local"], env["dist"]
    n_all = args.problems
    probs, P0 = load_batch(n_all)
    cfg = batch_config_block(n_all, probs, world)
    lo, hi = n_all * rank // world, n_all * (rank + 1) // world
    mine, P0m = probs[lo:hi], np.concatenate(P0[lo:hi])
    views = np.array([len(p["dets"]) for p in mine], dtype=np.int64)
    pv = np.concatenate([[0], np.cumsum(views)]).astype(np.int64)
    voffs, mp, sp = flatten_np([det for p in mine for det in p["dets"]])
    lib = _lib.lib()
    mp_pin, mp_ptr = pinned_copy(lib, mp)
    sp_pin, sp_ptr = pinned_copy(lib, sp)
    bp = cc.BatchProblem(0, None, device=local, csr=(pv, voffs, mp_pin, sp_pin))
    flush = torch.empty(64 << 20, dtype=torch.float32, device="cuda")
    barrier = env["barrier"]

    for _ in range(3):
        bp.lm_run(P0m, MAX_ITERS)
    barrier()
    tot_ms = tot_it = 0
    with ClockSampler(local) as clk:
        for _ in range(steps):
            flush.fill_(1.0)
            torch.cuda.synchronize()
            sse, P, iters, _, ms = bp.lm_run(P0m, MAX_ITERS)
            tot_ms += ms
            tot_it += int(iters.sum())
        barrier()
    if world > 1:
        t = torch.tensor([tot_ms, float(tot_it)], dtype=torch.float64, device="cuda")
        tm = t.clone()
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        tot_ms, tot_it = float(tm[0].item()), int(t[1].item())
    value = tot_it / (tot_ms * 1e-3)

    def e2e_step():
        b2 = cc.BatchProblem(0, None, device=local, csr=(pv, voffs, mp_pin, sp_pin))
        _, _, it, _, _ = b2.lm_run(P0m, MAX_ITERS)
        b2.close()
        return int(it.sum())
    e2e_step()
    barrier()
    n_e2e = max(3, min(steps, 10))
    t0 = time.perf_counter()
    it_e2e = sum(e2e_step() for _ in range(n_e2e))
    barrier()
    dt = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([dt, float(it_e2e)], dtype=torch.float64, device="cuda")
        tm = t.clone()
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        dt, it_e2e = float(tm[0].item()), int(t[1].item())
    e2e = {"value": it_e2e / dt, "unit": "LM iterations/s",
           "h2d_bytes_per_step": int(40 * voffs[-1] + 8 * len(P0m) + 8 * (len(voffs) + len(pv))),
           "d2h_bytes_per_step": int(8 * len(P0m) + 12 * len(mine)), "ms_per_step": 1e3 * dt / n_e2e,
           "api": "C-ABI per step and rank: calib_batch_create(page-locked CSR host buffers) + "
                  "calib_batch_lm_run(host P) + calib_batch_destroy"}
    # roofline: FP64 work actually required by the traces of one run / kernel time
    sse, P, iters, tr, ms = bp.lm_run(P0m, MAX_ITERS, want_trace=True)
    npts = np.array([sum(len(s) for s, m in p["dets"]) for p in mine], dtype=np.float64)
    acc = tr[:, :, 3].sum(axis=1)
    last_acc = np.array([tr[j, iters[j] - 1, 3] for j in range(len(mine))])
    evals = 1 + acc - last_acc
    flops = float((npts * (evals * FLOPS_PER_POINT[0] + iters * COST_FLOPS_PER_POINT)).sum())
    line = None
    if rank == 0:
        pk = fp64_peak(cc, local)
        sse2, _, iters2, _, ms2 = bp.lm_run(P0m, MAX_ITERS)
        roof = {"bound": "fp64", "kernel": "k_batch_lm", "achieved": flops / (ms2 * 1e-3) / 1e12, "peak": pk["tflops"],
                "unit": "TFLOP/s", "frac": flops / (ms2 * 1e-3) / 1e12 / pk["tflops"], "traffic": None,
                "peak_source": pk["source"], "peak_of_nominal": pk["of_nominal"],
                "flops": "sum over problems of points x (Jacobian passes x 1012 + iterations x 100)",
                "launch_ms": ms2, "problems_per_launch": len(mine),
                "mean_iterations": float(iters.mean()), "mean_jacobian_passes": float(evals.mean())}
        # parity of the timed launch against the oracle on a bounded sample of its problems
        import oracle_c
        ns = min(16 if light else 64, len(mine))
        sl = bp.param_slices()
        worst = {"rel": 0.0, "trace_equal": True}
        for j in range(ns):
            offs_j, mp_j, sp_j = flatten_np(mine[j]["dets"])
            so, Po, io, to = oracle_c.lm(0, P0[lo + j], offs_j, mp_j, sp_j, MAX_ITERS, nthreads=1)
            pb = parity_block(P[sl[j]], float(sse[j]), int(iters[j]), tr[j, :iters[j]], Po, so, io, to, "oracle")
            worst["rel"] = max(worst["rel"], pb["max_rel_param_diff"])
            worst["trace_equal"] = worst["trace_equal"] and pb["pre_plateau_trace_equal"]
        parity = {"ok": bool(worst["trace_equal"] and worst["rel"] <= 1e-6), "against": "oracle/calib_oracle.c",
                  "problems_checked": ns, "pre_plateau_trace_equal": worst["trace_equal"],
                  "max_rel_param_diff": worst["rel"]}
        cpu = None
        if world == 1 and not args.no_cpu_baseline and not light:
            v, threads, nsb, dtc = batch_cpu_baseline(probs, P0, args.cpu_sample)
            cpu = {"value": v, "unit": "LM iterations/s", "cores": threads, "kind": "port",
                   "sample": f"the first {nsb} problems, one per host thread ({dtc:.2f} s); C restatement of the "
                             "reference's path (oracle/calib_oracle.c)"}
        line = {"metric": "lm_iterations_per_sec", "value": value, "unit": "LM iterations/s", "n_gpus": world,
                "steps": steps, "warmup": 3, "ms_per_step": tot_ms / steps, "ms_per_refinement": tot_ms / steps / n_all,
                "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
                "data": "synthetic", "config": cfg, "calibrations_per_sec": n_all * steps / (tot_ms * 1e-3),
                "e2e": e2e, "gpu_launches": int(steps), "roofline": roof, "parity_vs_oracle": parity,
                "cpu_baseline": cpu, "clocks": clk.summary()}
    bp.close()
    lib.calib_host_free(mp_ptr)
    lib.calib_host_free(sp_ptr)
    return line


def run_batch_reference(args):
    if int(os.environ.get("RANK", "0")) != 0:
        return
    import oracle_c  # noqa: F401
    probs, P0 = load_batch(args.problems)
    val, threads, ns, dt = batch_cpu_baseline(probs, P0, args.cpu_sample)
    line = {"impl": "reference", "metric": "lm_iterations_per_sec", "value": val, "unit": "LM iterations/s",
            "n_gpus": args.gpus, "steps": 1, "warmup": 0, "ms_per_step": 1e3 * dt, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": batch_config_block(args.problems, probs, 1),
            "cpu_baseline": {"value": val, "unit": "LM iterations/s", "cores": threads, "kind": "port",
                             "sample": f"the first {ns} problems, one per host thread ({dt:.2f} s); C restatement "
                                       "of the reference's path (oracle/calib_oracle.c)"},
            "e2e": {"value": val, "unit": "LM iterations/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "product_imported": bool("camera_calibration_b200" in sys.modules)}
    print(json.dumps(line), flush=True)


# ---- one large calibration (cfg3 / cfg5 / the small golden cases), views sharded over the ranks ----------
def bench_calibration(args, env, name, scale, steps, warmup, light=False):
    """Returns the bench line (rank 0) or None.  light: the secondary block (no e2e, no CPU baseline)."""
    import torch
    import camera_calibration_b200 as cc
    from camera_calibration_b200 import _lib
    rank, world, local, dist, barrier = env["rank"], env["world"], env["local"], env["dist"], env["barrier"]
    c, P0 = load_workload(name, scale)
    model = c["model"]
    L = 10 if model == 0 else 9
    offs_all = np.concatenate([[0], np.cumsum([len(s) for s, m in c["dets"]])]).astype(np.int64)
    dets, Ploc, (lo, hi) = shard_views(c["dets"], P0, L, world, rank)
    offs, mp, sp = flatten_np(dets)
    lib = _lib.lib()
    mp_pin, mp_ptr = pinned_copy(lib, mp)
    sp_pin, sp_ptr = pinned_copy(lib, sp)
    prob = cc.Problem(model, offs, mp_pin, sp_pin, device=local)
    if args.segment:
        prob.set_segment(args.segment)
    if args.loop_mode:
        prob.set_loop_mode(args.loop_mode)
    prob.set_ladder(args.ladder)
    transport = "none"
    if world > 1:
        from camera_calibration_b200 import sharding
        transport = sharding.connect(prob, dist, args.transport)
        prob.comm_set_default(local)

    def step_resident():
        prob.upload(Ploc)
        prob.flush_l2()
        sse, iters = prob.lm_run_resident(MAX_ITERS)
        ms, launches, evals = prob.last_run_stats()
        return sse, iters, ms, launches, evals

    for _ in range(warmup):
        step_resident()
    barrier()
    tot_ms = tot_it = tot_launch = tot_eval = 0
    t_wall = time.perf_counter()
    with ClockSampler(local) as clk:
        for _ in range(steps):
            sse, iters, ms, launches, evals = step_resident()
            tot_ms += ms; tot_it += iters; tot_launch += launches; tot_eval += evals
        barrier()
    t_wall = time.perf_counter() - t_wall
    P_run = prob.download()
    tr_run = prob.download_trace(iters)
    if world > 1:
        t = torch.tensor([tot_ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        tot_ms = float(t.item())
    value = tot_it / (tot_ms * 1e-3)
    N_all = int(offs_all[-1])
    rows_per_s = 2.0 * N_all * tot_eval / (tot_ms * 1e-3)

    # ---- e2e: host buffers in, host buffers out, every step
    e2e = None
    if not light:
        def e2e_step():
            p2 = cc.Problem(model, offs, mp_pin, sp_pin, device=local)     # borrows prob's communicator (set_default)
            p2.set_ladder(args.ladder)
            if args.loop_mode:
                p2.set_loop_mode(args.loop_mode)
            _, _, it, _ = p2.lm_run(Ploc, MAX_ITERS, want_trace=False)
            p2.close()
            return it

        n_e2e = max(3, min(steps, 50))
        for _ in range(3):
            e2e_step()
        barrier()
        t0 = time.perf_counter()
        it_e2e = 0
        for _ in range(n_e2e):
            it_e2e += e2e_step()
        barrier()
        dt = time.perf_counter() - t0
        if world > 1:
            t = torch.tensor([dt], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt = float(t.item())
        e2e = {"value": it_e2e / dt, "unit": "LM iterations/s",
               "h2d_bytes_per_step": int(40 * offs[-1] + 8 * len(Ploc) + 8 * len(offs)),
               "d2h_bytes_per_step": int(8 * len(Ploc) + 96), "ms_per_step": 1e3 * dt / n_e2e,
               "ms_per_refinement": 1e3 * dt / n_e2e, "steps": n_e2e,
               "api": "C-ABI per step" + (" and rank" if world > 1 else "") +
                      ": calib_create(page-locked CSR host buffers) + calib_lm_run(host P) + calib_destroy"}
        if world > 1:
            # the drop-in's own way to several GPUs: ONE process (rank 0's), the library shards the views and
            # drives all N devices (calib_multi_*); the other ranks' processes sit at the barrier meanwhile
            sp_block = None
            env["cpu_barrier"]()
            if rank == 0:
                offs_a, mp_a, sp_a = flatten_np(c["dets"])
                mpa, mpa_ptr = pinned_copy(lib, mp_a)
                spa, spa_ptr = pinned_copy(lib, sp_a)

                def sp_step():
                    mprob = cc.MultiProblem(model, offs_a, mpa, spa, list(range(world)))
                    mprob.set_ladder(args.ladder)
                    _, _, it, _ = mprob.lm_run(P0, MAX_ITERS, want_trace=False)
                    ms = mprob.last_device_ms
                    mprob.close()
                    return it, ms
                for _ in range(2):
                    sp_step()
                n_sp = max(3, min(steps, 20))
                t0 = time.perf_counter()
                it_sp = ms_sp = 0
                for _ in range(n_sp):
                    it, ms = sp_step()
                    it_sp += it
                    ms_sp += ms
                dts = time.perf_counter() - t0
                lib.calib_host_free(mpa_ptr)
                lib.calib_host_free(spa_ptr)
                sp_block = {"value": it_sp / dts, "unit": "LM iterations/s", "ms_per_refinement": 1e3 * dts / n_sp,
                            "device_ms_per_refinement": ms_sp / n_sp, "steps": n_sp,
                            "h2d_bytes_per_step": int(40 * offs_a[-1] + 8 * len(P0)), "d2h_bytes_per_step": int(8 * len(P0)),
                            "api": "ONE process: calib_multi_create(page-locked CSR host buffers, all N devices) + "
                                   "calib_multi_lm_run(full-length host P) + calib_multi_destroy"}
            env["cpu_barrier"]()
            e2e["single_process"] = sp_block
        if world == 1:
            cal = cc.Calibrator(model, device=local)
            A0, W0, k0 = cal._decomposeParameterVector(P0)
            for _ in range(2):
                cal.refineCalibrationParameters(A0, W0, k0, c["dets"], MAX_ITERS)
            n_py = max(3, min(steps, 10))
            t0 = time.perf_counter()
            it_py = 0
            for _ in range(n_py):
                cal.refineCalibrationParameters(A0, W0, k0, c["dets"], MAX_ITERS)
                it_py += cal.lastIters
            dtp = time.perf_counter() - t0
            e2e["python_api"] = {"value": it_py / dtp, "unit": "LM iterations/s", "ms_per_step": 1e3 * dtp / n_py,
                                 "ms_per_refinement": 1e3 * dtp / n_py,
                                 "api": "Calibrator.refineCalibrationParameters(A, W, k, allDetections, maxIters): "
                                        "list of per-view arrays in, (sse, A, W, k) out"}

    # ---- per-kernel times of the pass (the kernels inside the WHILE graph's body, launched one by one): CUDA events around back-to-back launches on the library's stream, L2-warm as between the
    # iterations of a run, and after an L2 flush before every launch
    kt = {}
    for i, nm in enumerate(["k_eval", "k_schur", "k_backsub", "k_cost"]):
        kt[nm] = {"us_l2_warm": prob.time_kernel(Ploc, i, 30), "us_after_l2_flush": prob.time_kernel(Ploc, i, 10, flush=True)}
    roof = None
    if rank == 0:
        pk = fp64_peak(cc, local)
        peak64 = pk["tflops"]
        n_loc = int(offs[-1])
        avg_s = 1e-6 * kt["k_eval"]["us_l2_warm"]
        cold_s = 1e-6 * kt["k_eval"]["us_after_l2_flush"]
        fl = FLOPS_PER_POINT[model] * n_loc
        by = 2 * BYTES_PER_POINT * n_loc          # two sweeps over the detections
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(
            os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
        hbm_peak = peaks.get("hbm_gbs", 6650.0)
        mix = ncu_mix(model)
        issue = None
        if mix:
            issue_peak = peak64 * 1e12 / 2.0       # thread-level FP64 instructions per second
            ipp = mix["instructions_per_point"]
            issue = {"instructions_per_point": ipp, "achieved_ginst_per_s": ipp * n_loc / avg_s / 1e9,
                     "peak_ginst_per_s": issue_peak / 1e9, "frac": ipp * n_loc / avg_s / issue_peak,
                     "from_capture": mix.get("capture")}
        roof = {"bound": "fp64", "kernel": "k_eval",
                "achieved": fl / avg_s / 1e12, "peak": peak64, "unit": "TFLOP/s",
                # headline fraction: executed FP64 instructions against the measured issue rate; the
                # algorithmic figure (1012 FLOP / point) is kept beside it
                "frac": issue["frac"] if issue else fl / avg_s / 1e12 / peak64,
                "frac_algorithmic": fl / avg_s / 1e12 / peak64,
                "traffic": mix.get("dram_bytes_per_launch") if mix and mix.get("points") == n_loc else None,
                "traffic_source": (mix.get("capture") if mix and mix.get("points") == n_loc else None),
                "peak_source": pk["source"], "peak_of_nominal": pk["of_nominal"],
                "fp64_pipe_pct_of_peak_elapsed_ncu": mix.get("pipe_fp64_pct_elapsed") if mix else None,
                "flops_per_point": FLOPS_PER_POINT[model], "points_per_launch": n_loc,
                "avg_launch_us": avg_s * 1e6, "avg_launch_us_after_l2_flush": cold_s * 1e6,
                "frac_algorithmic_after_l2_flush": fl / cold_s / 1e12 / peak64,
                "fp64_issue": issue,
                # what the DFMA-only peak leaves out (micro-benchmark, not measured in this run)
                "issue_port_note": "an FP64 instruction holds a scheduler's issue port for 2 cycles and other instructions "
                                   "hide behind it only a little (profiles/r4n_fp64_issue_mix.txt): a 32-point step of a warp "
                                   "pair (349 FP64 + 187 other warp instructions, radtan) needs >= ~866 cycles and takes "
                                   "~1100 -- 0.79 of the issue bound at frac 0.60 of the DFMA-only rate (DESIGN.md 4.1)",
                # the whole refinement as one launch: algorithmic work of all its sweeps / its duration
                "whole_refinement": {
                    "kernel": "all kernels of the loop (k_eval, k_schur, k_backsub, k_cost inside the graph)",
                    "tflops_algorithmic": (tot_eval * FLOPS_PER_POINT[model] + tot_it * COST_FLOPS_PER_POINT) * n_loc
                                          / (tot_ms * 1e-3) / 1e12,
                    "frac_algorithmic": (tot_eval * FLOPS_PER_POINT[model] + tot_it * COST_FLOPS_PER_POINT) * n_loc
                                        / (tot_ms * 1e-3) / 1e12 / peak64},
                "hbm": {"achieved": by / cold_s / 1e9, "peak": hbm_peak, "unit": "GB/s",
                        "frac": by / cold_s / 1e9 / hbm_peak, "bytes_per_point_per_sweep": BYTES_PER_POINT,
                        "peak_source": "MEASURED_PEAKS.json (of measured)" if peaks else "fallback 6650 (of fallback)"}}

    # ---- parity of the timed run
    parity_key, parity = None, None
    cpu = None
    if world == 1:
        if rank == 0 and not args.no_cpu_baseline:
            offs_a, mp_a, sp_a = flatten_np(c["dets"])
            it_o, ev_o, dt_o, (sse_o, P_o, tr_o) = oracle_full_run(model, offs_a, mp_a, sp_a, P0)
            cpu = dict(value=it_o / dt_o, unit="LM iterations/s", cores=host_threads(), kind="port",
                       ms_per_refinement=1e3 * dt_o,
                       sample=f"one complete LM run of the workload: {it_o} iterations, {ev_o} Jacobian passes, "
                              f"{dt_o:.2f} s; C restatement of the reference's path (oracle/calib_oracle.c, OpenMP)",
                       jacobian_rows_per_s=2.0 * offs_a[-1] * ev_o / dt_o,
                       reference_python_s=reference_python_seconds())
            parity_key = "parity_vs_oracle"
            parity = parity_block(P_run, sse, iters, tr_run, P_o, sse_o, it_o, tr_o, "oracle/calib_oracle.c (CPU)")
    else:
        parts = [None] * world
        dist.all_gather_object(parts, (lo, hi, P_run.tolist(), float(sse), int(iters), tr_run.tolist()))
        parity_key = "parity_vs_single"
        if rank == 0:       # merged parameters; every rank must hold the same trace and shared parameters
            Pfull = np.empty(L + 6 * (len(offs_all) - 1))
            Pfull[:L] = P_run[:L]
            agree = True
            for a, b, p, s, it, t in parts:
                Pfull[L + 6 * a:L + 6 * b] = np.asarray(p)[L:]
                agree = agree and it == iters and s == sse and np.array_equal(np.asarray(t), tr_run) \
                    and np.array_equal(np.asarray(p)[:L], P_run[:L])
            sharded = {"P": Pfull, "sse": float(sse), "iters": int(iters), "trace": tr_run, "agree": bool(agree)}
    line = None
    if rank == 0:
        line = {
            "metric": "lm_iterations_per_sec", "value": value, "unit": "LM iterations/s", "n_gpus": world,
            "steps": steps, "warmup": warmup, "ms_per_step": tot_ms / steps, "ms_per_refinement": tot_ms / steps,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": dict(workload_config(name, scale, c, offs_all, world),
                                                damping_ladder=args.ladder, exchange={
                "none": "none (one GPU)",
                "peer": "reduced system (161 doubles) + trial cost summed inside the kernel through NVLink peer memory",
                "nccl": "two ncclAllReduce per iteration (161 doubles, 1 double)"}[transport]),
            "jacobian_rows_per_sec": rows_per_s, "lm_iterations_per_step": tot_it / steps,
            "jacobian_passes_per_step": tot_eval / steps, "final_sse": sse,
            "wall_ms_per_step_incl_flush": 1e3 * t_wall / steps,
            "e2e": e2e, "gpu_launches": int(tot_launch), "roofline": roof, "kernel_times": kt,
            "cpu_baseline": cpu, "clocks": clk.summary(),
        }
        if world > 1:
            line["_sharded"] = sharded
    env["_after"] = (prob, lib, mp_ptr, sp_ptr, c, P0, parity_key, parity, line)
    return line


def ncu_mix(model):
    """numbers that only a profiler gives, from the committed capture named in the entry itself
    (profiles/ncu_traffic.json): executed FP64 instructions per point of the evaluation pass, its DRAM
    bytes per launch at `points`, the FP64 pipe's share of the elapsed cycles"""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
        return t["k_eval"][str(int(model))]
    except Exception:
        return None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default="cfg3")
    ap.add_argument("--scale", type=float, default=1.0)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--segment", type=int, default=0)
    ap.add_argument("--loop-mode", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-secondary", action="store_true")
    ap.add_argument("--no-python-reference", action="store_true",
                    help="skip timing the vendored Python reference on cfg1 (about 25 s of host time)")
    ap.add_argument("--problems", type=int, default=8192)
    ap.add_argument("--cpu-sample", type=int, default=512)
    ap.add_argument("--ladder", type=int, default=4, choices=[1, 2, 4],
                    help="damping ladder: trial steps taken ahead of the accept/reject decisions (1 = off)")
    ap.add_argument("--transport", default="auto", choices=["auto", "peer", "nccl"],
                    help="sharded views: in-kernel all-reduce over NVLink peer memory, or NCCL all-reduces")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_batch_reference(args) if args.workload == "cfg4" else run_reference_arm(args)
    args.warmup = max(args.warmup, 3)

    import torch
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist = None
    cpu_group = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        cpu_group = dist.new_group(backend="gloo")      # host-side waits that leave the GPUs idle (see cpu_barrier)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def cpu_barrier():
        """A barrier that does NOT run on the GPUs: while rank 0's process drives every device by itself (single-
        process handle, one-GPU parity run) the other ranks must not sit in an NCCL barrier -- its kernel would
        spin on their GPU and time-slice with rank 0's kernels there."""
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier(group=cpu_group)

    env = {"rank": rank, "world": world, "local": local, "dist": dist, "barrier": barrier, "cpu_barrier": cpu_barrier}
    if args.workload == "cfg4":
        line = bench_batch(args, env, 10 if args.steps == 100 else args.steps)
        if rank == 0:
            print(json.dumps(line), flush=True)
    else:
        line = run_calibration_with_parity(args, env, args.workload, args.scale, args.steps, args.warmup, light=False)
        if not args.no_secondary and args.workload == "cfg3" and args.scale == 1.0:
            sec = {}
            s5 = run_calibration_with_parity(args, env, "cfg5", 1.0, 5, 3, light=True)
            if rank == 0:
                sec["cfg5"] = {k: s5[k] for k in ("value", "unit", "ms_per_refinement", "lm_iterations_per_step",
                                                  "jacobian_passes_per_step", "jacobian_rows_per_sec", "final_sse",
                                                  "gpu_launches", "kernel_times", "config") if k in s5}
                for k in ("parity_vs_oracle", "parity_vs_single"):
                    if k in s5:
                        sec["cfg5"][k] = s5[k]
            if world == 1:
                s4 = bench_batch(args, env, 3, light=True)
                if rank == 0:
                    sec["cfg4"] = {k: s4[k] for k in ("value", "unit", "ms_per_step", "ms_per_refinement",
                                                      "calibrations_per_sec", "roofline", "parity_vs_oracle",
                                                      "e2e", "config")}
            if rank == 0:
                line["secondary"] = sec
        if (rank == 0 and world == 1 and not args.no_cpu_baseline and not args.no_python_reference
                and isinstance(line.get("cpu_baseline"), dict)):
            line["cpu_baseline"]["reference_python"] = reference_python_block(lambda: product_on_golden("cfg1", local))
        if rank == 0:
            print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def product_on_golden(case, device):
    """the product on one of the reference's own cases (tests/golden/<case>.npz): sse, P, iterations, device ms"""
    import camera_calibration_b200 as cc
    z = np.load(os.path.join(ROOT, "tests", "golden", case + ".npz"))
    prob = cc.Problem(int(z["model"]), z["view_offsets"], z["model_pts"], z["sensor_pts"], device=device)
    for _ in range(3):
        sse, P, iters, _ = prob.lm_run(z["P0"].copy(), int(z["maxIters"]))
    ms = prob.last_run_stats()[0]
    prob.close()
    return sse, P, iters, ms


def run_calibration_with_parity(args, env, name, scale, steps, warmup, light):
    """bench_calibration + the parity block of the timed run + teardown (collective: every rank calls it)."""
    import camera_calibration_b200 as cc
    line = bench_calibration(args, env, name, scale, steps, warmup, light)
    prob, lib, mp_ptr, sp_ptr, c, P0, parity_key, parity, _ = env.pop("_after")
    rank, world, local, dist = env["rank"], env["world"], env["local"], env["dist"]
    if world > 1:
        # rank 0 alone runs the whole workload on its one GPU (the communicator is not lent to this handle)
        prob.comm_set_default(local, on=False)
        env["cpu_barrier"]()
        if rank == 0:
            model = c["model"]
            L = 10 if model == 0 else 9
            sh = line.pop("_sharded")
            offs_a, mp_a, sp_a = flatten_np(c["dets"])
            one = cc.Problem(model, offs_a, mp_a, sp_a, device=local)
            one.set_ladder(args.ladder)
            s1, P1, i1, t1 = one.lm_run(P0, MAX_ITERS)
            ms1 = one.last_run_stats()[0]
            one.close()
            parity = parity_block(sh["P"], sh["sse"], sh["iters"], sh["trace"], P1, s1, i1, t1,
                                  "the same workload on one GPU (rank 0, outside the timed region)")
            parity["ranks_agree"] = sh["agree"]
            parity["ok"] = bool(parity["ok"] and sh["agree"])
            parity["single_gpu_ms_per_refinement"] = ms1
        env["cpu_barrier"]()
    if rank == 0 and parity_key and parity is not None:
        line[parity_key] = parity
    prob.close()
    lib.calib_host_free(mp_ptr)
    lib.calib_host_free(sp_ptr)
    return line


if __name__ == "__main__":
    main()
