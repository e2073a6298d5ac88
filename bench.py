#!/usr/bin/env python
"""Benchmark of the LM refinement hot path (BASELINE.json metric: LM iterations/s and Jacobian
rows/s).  One "step" = one complete run of the LM loop (Calibrator.refineCalibrationParameters,
maxIters=100, until the reference's own termination rule fires) on the named workload.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload cfg3|cfg5|itest|cfg1|cfg2]
                  [--impl reference]

value   device-resident: detections already in HBM, parameters uploaded before the timed region,
        timed with CUDA events on the library's stream (max over ranks).  L2 is flushed (256 MB
        write) before every step; inside a step the detections stay L2-resident across the LM
        iterations, as in production use.
e2e     through the C-ABI with HOST buffers (page-locked CSR arrays): calib_create (H2D of the
        detections) + calib_lm_run (H2D of P, the loop, D2H of P) + calib_destroy, every step.
        e2e.python_api is the same through Calibrator.refineCalibrationParameters with the
        reference's list-of-arrays argument (adds the host-side flattening of the list).
--impl reference  times the CPU restatement of the reference's path (oracle/calib_oracle.c,
        OpenMP over all host threads) on the same workload: the reference itself is Python/sympy,
        cannot travel to the GPU box and cannot hold the dense J of this workload (154 GB).
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

MAX_ITERS = 100
# Algorithmic FP64 work per point of the fused evaluation pass (DESIGN.md
# "Roofline"): F_model + 2 (L+6)(L+7) + 4 (L+6) + 4
FLOPS_PER_POINT = {0: 400 + 2 * 16 * 17 + 4 * 16 + 4, 1: 450 + 2 * 15 * 16 + 4 * 15 + 4}
BYTES_PER_POINT = 40          # X, Y, Z, u, v as float64, read once per sweep


def load_workload(name, scale):
    from camera_calibration_b200 import synth
    import camera_calibration_b200 as cc
    c = synth.config(name, scale) if name in ("cfg3", "cfg5") else synth.config(name)
    if c["init"] is None:
        g = np.load(os.path.join(ROOT, "tests", "golden", name + ".npz"))
        P0 = g["P0"].copy()
    else:
        cal = cc.Calibrator(c["model"])
        P0 = cal._composeParameterVector(*c["init"]).ravel()
    return c, P0


def shard_bounds(counts, world):
    """contiguous view ranges balanced by point count"""
    cum = np.concatenate([[0], np.cumsum(counts)])
    total = cum[-1]
    b = [int(np.searchsorted(cum, total * r / world, side="left")) for r in range(world)] + [len(counts)]
    b[0] = 0
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
    _, _, it, tr = oracle_c.lm(model, P0, offs, mp, sp, MAX_ITERS, nthreads=host_threads())
    dt = time.perf_counter() - t0
    evals = 1 + int(tr[:-1, 3].sum())
    return it, evals, dt


def cpu_baseline(model, offs, mp, sp, P0):
    """The oracle port timed on the host cores: one complete LM run of the workload (bounded:
    a fraction of a second to a few seconds for every workload benched here)."""
    import oracle_c
    it, evals, dt = oracle_full_run(model, offs, mp, sp, P0)
    return dict(value=it / dt, unit="LM iterations/s", cores=host_threads(), kind="port",
                sample=f"one complete LM run of the workload: {it} iterations, {evals} Jacobian passes, "
                       f"{dt:.2f} s; C restatement of the reference's path (oracle/calib_oracle.c, OpenMP)",
                jacobian_rows_per_s=2.0 * offs[-1] * evals / dt)


def workload_config(args, c, offs):
    return {"workload": args.workload + ("" if args.scale == 1.0 else f"@scale{args.scale}"),
            "model": "radtan" if c["model"] == 0 else "fisheye", "views": int(len(offs) - 1),
            "points": int(offs[-1]), "params": int((10 if c["model"] == 0 else 9) + 6 * (len(offs) - 1)),
            "max_iters": MAX_ITERS,
            "init": "perturbed ground truth, RandomState(12345)" if c["init"] is not None
                    else "the reference's linear stage (tests/golden)",
            "step": "one complete LM refinement (until the reference's termination rule fires)",
            "parallelism": f"views sharded over {args.gpus} GPU(s)" if args.gpus > 1 else "1 GPU",
            "l2": "flushed (256 MB write) before every timed step; within a step the detections stay "
                  "L2-resident across LM iterations, as in production use"}


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import camera_calibration_b200 as cc
    import oracle_c
    c, P0 = load_workload(args.workload, args.scale)
    offs, mp, sp = cc.flatten_detections(c["dets"])
    for _ in range(min(args.warmup, 1)):
        oracle_c.lm(c["model"], P0, offs, mp, sp, 2)
    tot_it = evals = 0
    t0 = time.perf_counter()
    for _ in range(args.steps):
        it, ev, _ = oracle_full_run(c["model"], offs, mp, sp, P0)
        tot_it += it
        evals += ev
    dt = time.perf_counter() - t0
    val = tot_it / dt
    line = {
        "impl": "reference", "metric": "lm_iterations_per_sec", "value": val, "unit": "LM iterations/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "jacobian_rows_per_sec": 2.0 * offs[-1] * evals / dt, "lm_iterations_per_step": tot_it / args.steps,
        "config": workload_config(args, c, offs),
        "cpu_baseline": {"value": val, "unit": "LM iterations/s", "cores": host_threads(), "kind": "port",
                         "sample": "each step = one complete LM run of the workload; C restatement of the "
                                   "reference's path (oracle/calib_oracle.c, OpenMP over all host threads); the "
                                   "Python/sympy reference cannot travel to this box nor hold this workload's dense J"},
        "e2e": {"value": val, "unit": "LM iterations/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def ncu_traffic(workload, n_points):
    """DRAM bytes per launch of k_eval from the committed ncu capture (profiles/ncu_traffic.json), or
    None when there is no capture of this workload at this size."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))[workload]
        return int(t["k_eval"]) if int(t["points"]) == int(n_points) else None
    except Exception:
        return None


def fp64_issue(model, n_points, seconds, peak_tflops):
    """Executed FP64 instructions (ncu count per point, profiles/r1w_fp64_mix.md) against the FP64 issue
    rate the DFMA micro-benchmark reached: the kernel exploits more structure than the algorithmic FLOP
    figure assumes, so this is the fraction of the pipe it really keeps busy."""
    try:
        ipp = float(json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
                    ["fp64_instructions_per_point_k_eval"][str(int(model))])
    except Exception:
        return None
    issue_peak = peak_tflops * 1e12 / 2.0          # thread-level FP64 instructions per second
    return {"instructions_per_point": ipp, "achieved_ginst_per_s": ipp * n_points / seconds / 1e9,
            "peak_ginst_per_s": issue_peak / 1e9, "frac": ipp * n_points / seconds / issue_peak}


def pinned_copy(L, arr):
    """numpy view of a page-locked copy of arr (calib_host_alloc); returns (view, ptr)."""
    ptr = L.calib_host_alloc(max(arr.nbytes, 8))
    if not ptr:
        raise RuntimeError("calib_host_alloc failed")
    buf = (ctypes.c_char * max(arr.nbytes, 8)).from_address(ptr)
    view = np.frombuffer(buf, dtype=arr.dtype, count=arr.size).reshape(arr.shape)
    view[...] = arr
    return view, ptr


# ---- config 4: batch of independent calibrations, split over GPUs with no communication ----------
COST_FLOPS_PER_POINT = 100     # residual-only sweep


def load_batch(n_problems):
    from camera_calibration_b200 import synth
    import camera_calibration_b200 as cc
    probs = synth.batch_config(n_problems, procs=min(os.cpu_count() or 1, 32))
    cal = cc.Calibrator(0)
    P0 = [cal._composeParameterVector(*p["init"]).ravel() for p in probs]
    return probs, P0


def batch_cpu_baseline(probs, P0, sample):
    """oracle port on the host cores: one problem per thread (ctypes releases the GIL)."""
    import oracle_c
    import camera_calibration_b200 as cc
    from concurrent.futures import ThreadPoolExecutor
    threads = os.cpu_count() or 1
    idx = list(range(min(sample, len(probs))))
    flat = [cc.flatten_detections(probs[j]["dets"]) for j in idx]

    def one(k):
        offs, mp, sp = flat[k]
        return oracle_c.lm(0, P0[idx[k]], offs, mp, sp, MAX_ITERS, nthreads=1)[2]
    one(0)
    t0 = time.perf_counter()
    with ThreadPoolExecutor(threads) as ex:
        its = list(ex.map(one, range(len(idx))))
    dt = time.perf_counter() - t0
    return sum(its) / dt, threads, len(idx), dt


def run_batch(args):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    n_all = args.problems
    probs, P0 = load_batch(n_all)
    cfg = {"workload": f"cfg4: {n_all} independent 20-view calibrations, 9x6 board, radtan, sigma 0.1 px",
           "model": "radtan", "problems": n_all, "views_per_problem": 20,
           "points": int(sum(len(s) for p in probs for s, m in p["dets"])), "max_iters": MAX_ITERS,
           "init": "perturbed ground truth, RandomState(problem index)",
           "step": "all problems refined to termination in one launch",
           "parallelism": f"problems split over {world} GPU(s), no communication",
           "l2": "flushed (256 MB write) before every timed step"}
    if args.impl == "reference":
        if rank != 0:
            return
        import oracle_c  # noqa: F401
        val, threads, ns, dt = batch_cpu_baseline(probs, P0, args.cpu_sample)
        line = {"impl": "reference", "metric": "lm_iterations_per_sec", "value": val, "unit": "LM iterations/s",
                "n_gpus": args.gpus, "steps": 1, "warmup": 0, "ms_per_step": 1e3 * dt, "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": cfg,
                "cpu_baseline": {"value": val, "unit": "LM iterations/s", "cores": threads, "kind": "port",
                                 "sample": f"the first {ns} problems, one per host thread ({dt:.2f} s); C restatement "
                                           "of the reference's path (oracle/calib_oracle.c)"},
                "e2e": {"value": val, "unit": "LM iterations/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "gpu_launches": 0}
        print(json.dumps(line), flush=True)
        return
    import torch
    import camera_calibration_b200 as cc
    from camera_calibration_b200 import _lib
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    lo, hi = n_all * rank // world, n_all * (rank + 1) // world
    mine, P0m = probs[lo:hi], np.concatenate(P0[lo:hi])
    views = np.array([len(p["dets"]) for p in mine], dtype=np.int64)
    pv = np.concatenate([[0], np.cumsum(views)]).astype(np.int64)
    voffs, mp, sp = cc.flatten_detections([det for p in mine for det in p["dets"]])
    lib = _lib.lib()
    mp_pin, mp_ptr = pinned_copy(lib, mp)
    sp_pin, sp_ptr = pinned_copy(lib, sp)
    bp = cc.BatchProblem(0, None, device=local, csr=(pv, voffs, mp_pin, sp_pin))
    flush = torch.empty(64 << 20, dtype=torch.float32, device="cuda")

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        bp.lm_run(P0m, MAX_ITERS)
    barrier()
    tot_ms = tot_it = 0
    with ClockSampler(local) as clk:
        for _ in range(args.steps):
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
    n_e2e = max(3, min(args.steps, 10))
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
    roof = None
    cpu = None
    if rank == 0:
        peak64 = cc.calibrate.fp64_peak_tflops(local)
        sse2, _, iters2, _, ms2 = bp.lm_run(P0m, MAX_ITERS)
        roof = {"bound": "fp64", "kernel": "k_batch_lm", "achieved": flops / (ms2 * 1e-3) / 1e12, "peak": peak64,
                "unit": "TFLOP/s", "frac": flops / (ms2 * 1e-3) / 1e12 / peak64, "traffic": None,
                "peak_source": "DFMA micro-benchmark in this run (calib_measure_fp64_peak)",
                "flops": "sum over problems of points x (Jacobian passes x 1012 + iterations x 100)",
                "launch_ms": ms2, "problems_per_launch": len(mine),
                "mean_iterations": float(iters.mean()), "mean_jacobian_passes": float(evals.mean())}
        if world == 1 and not args.no_cpu_baseline:
            v, threads, ns, dtc = batch_cpu_baseline(probs, P0, args.cpu_sample)
            cpu = {"value": v, "unit": "LM iterations/s", "cores": threads, "kind": "port",
                   "sample": f"the first {ns} problems, one per host thread ({dtc:.2f} s); C restatement of the "
                             "reference's path (oracle/calib_oracle.c)"}
        line = {"metric": "lm_iterations_per_sec", "value": value, "unit": "LM iterations/s", "n_gpus": world,
                "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": tot_ms / args.steps,
                "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
                "data": "synthetic", "config": cfg, "calibrations_per_sec": n_all * args.steps / (tot_ms * 1e-3),
                "e2e": e2e, "gpu_launches": int(args.steps), "roofline": roof, "cpu_baseline": cpu,
                "clocks": clk.summary()}
        print(json.dumps(line), flush=True)
    bp.close()
    lib.calib_host_free(mp_ptr)
    lib.calib_host_free(sp_ptr)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


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
    ap.add_argument("--problems", type=int, default=8192)
    ap.add_argument("--cpu-sample", type=int, default=512)
    ap.add_argument("--ladder", type=int, default=4, choices=[1, 2, 4],
                    help="damping ladder: trial steps taken ahead of the accept/reject decisions (1 = off)")
    ap.add_argument("--pdl", type=int, default=0, choices=[0, 1], help="programmatic dependent launch inside the graphs")
    ap.add_argument("--transport", default="auto", choices=["auto", "peer", "nccl"],
                    help="sharded views: in-kernel all-reduce over NVLink peer memory, or NCCL all-reduces")
    args = ap.parse_args()
    if args.workload == "cfg4":
        if args.steps == 100:
            args.steps = 10
        return run_batch(args)
    if args.impl == "reference":
        return run_reference_arm(args)
    args.warmup = max(args.warmup, 3)

    import torch
    import camera_calibration_b200 as cc
    from camera_calibration_b200 import _lib
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    c, P0 = load_workload(args.workload, args.scale)
    model = c["model"]
    L = 10 if model == 0 else 9
    offs_all, mp_all, sp_all = cc.flatten_detections(c["dets"])
    dets, Ploc, (lo, hi) = shard_views(c["dets"], P0, L, world, rank)
    offs, mp, sp = cc.flatten_detections(dets)
    lib = _lib.lib()
    mp_pin, mp_ptr = pinned_copy(lib, mp)
    sp_pin, sp_ptr = pinned_copy(lib, sp)
    prob = cc.Problem(model, offs, mp_pin, sp_pin, device=local)
    if args.segment:
        prob.set_segment(args.segment)
    if args.loop_mode:
        prob.set_loop_mode(args.loop_mode)
    prob.set_ladder(args.ladder)
    prob.set_pdl(args.pdl)
    transport = "none"
    if world > 1:
        from camera_calibration_b200 import sharding
        transport = sharding.connect(prob, dist, args.transport)
        prob.comm_set_default(local)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def step_resident():
        prob.upload(Ploc)
        prob.flush_l2()
        sse, iters = prob.lm_run_resident(MAX_ITERS)
        ms, launches, evals = prob.last_run_stats()
        return sse, iters, ms, launches, evals

    for _ in range(args.warmup):
        step_resident()
    barrier()
    tot_ms = tot_it = tot_launch = tot_eval = 0
    t_wall = time.perf_counter()
    with ClockSampler(local) as clk:
        for _ in range(args.steps):
            sse, iters, ms, launches, evals = step_resident()
            tot_ms += ms; tot_it += iters; tot_launch += launches; tot_eval += evals
        barrier()
    t_wall = time.perf_counter() - t_wall
    if world > 1:
        t = torch.tensor([tot_ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        tot_ms = float(t.item())
    value = tot_it / (tot_ms * 1e-3)
    N_all = int(offs_all[-1])
    rows_per_s = 2.0 * N_all * tot_eval / (tot_ms * 1e-3)

    # ---- e2e: host buffers in, host buffers out, every step
    def e2e_step():
        p2 = cc.Problem(model, offs, mp_pin, sp_pin, device=local)     # borrows prob's communicator (set_default)
        p2.set_ladder(args.ladder)
        p2.set_pdl(args.pdl)
        _, _, it, _ = p2.lm_run(Ploc, MAX_ITERS, want_trace=False)
        p2.close()
        return it

    n_e2e = max(3, min(args.steps, 50))
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
           "d2h_bytes_per_step": int(8 * len(Ploc) + 96), "ms_per_step": 1e3 * dt / n_e2e, "steps": n_e2e,
           "api": "C-ABI per step" + (" and rank" if world > 1 else "") +
                  ": calib_create(page-locked CSR host buffers) + calib_lm_run(host P) + calib_destroy"}
    if world == 1:
        cal = cc.Calibrator(model, device=local)
        A0, W0, k0 = cal._decomposeParameterVector(P0)
        for _ in range(2):
            cal.refineCalibrationParameters(A0, W0, k0, c["dets"], MAX_ITERS)
        n_py = max(3, min(args.steps, 10))
        t0 = time.perf_counter()
        it_py = 0
        for _ in range(n_py):
            cal.refineCalibrationParameters(A0, W0, k0, c["dets"], MAX_ITERS)
            it_py += cal.lastIters
        dtp = time.perf_counter() - t0
        e2e["python_api"] = {"value": it_py / dtp, "unit": "LM iterations/s", "ms_per_step": 1e3 * dtp / n_py,
                             "api": "Calibrator.refineCalibrationParameters(A, W, k, allDetections, maxIters): "
                                    "list of per-view arrays in, (sse, A, W, k) out"}

    # ---- roofline of the dominant kernel (k_eval): CUDA events around back-to-back launches on the
    # library's stream (calib_time_kernel), L2-warm as between the iterations of a run, and after
    # an L2 flush before every launch
    roof = None
    kt = {}
    for i, nm in enumerate(["k_eval", "k_schur", "k_backsub", "k_cost"]):
        kt[nm] = {"us_l2_warm": prob.time_kernel(Ploc, i, 30), "us_after_l2_flush": prob.time_kernel(Ploc, i, 10, flush=True)}
    if rank == 0:
        peak64 = cc.calibrate.fp64_peak_tflops(local)
        n_loc = int(offs[-1])
        avg_s = 1e-6 * kt["k_eval"]["us_l2_warm"]
        cold_s = 1e-6 * kt["k_eval"]["us_after_l2_flush"]
        fl = FLOPS_PER_POINT[model] * n_loc
        by = 2 * BYTES_PER_POINT * n_loc          # two sweeps over the detections
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(
            os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
        hbm_peak = peaks.get("hbm_gbs", 6650.0)
        roof = {"bound": "fp64", "kernel": "k_eval", "achieved": fl / avg_s / 1e12, "peak": peak64,
                "unit": "TFLOP/s", "frac": fl / avg_s / 1e12 / peak64, "traffic": ncu_traffic(args.workload, n_loc),
                "peak_source": "DFMA micro-benchmark in this run (calib_measure_fp64_peak): FP64 is not in "
                               "MEASURED_PEAKS.json and the kernel is FP64-pipe bound, not HBM bound",
                "flops_per_point": FLOPS_PER_POINT[model], "points_per_launch": n_loc,
                "avg_launch_us": avg_s * 1e6, "avg_launch_us_after_l2_flush": cold_s * 1e6,
                "frac_after_l2_flush": fl / cold_s / 1e12 / peak64,
                "fp64_pipe_active_pct_ncu": "see profiles/README.md (ncu sm__inst_executed_pipe_fp64)",
                "fp64_issue": fp64_issue(model, n_loc, avg_s, peak64),
                "hbm": {"achieved": by / cold_s / 1e9, "peak": hbm_peak, "unit": "GB/s",
                        "frac": by / cold_s / 1e9 / hbm_peak, "bytes_per_point_per_sweep": BYTES_PER_POINT,
                        "peak_source": "MEASURED_PEAKS.json (of measured)" if peaks else "fallback 6650 (of fallback)"}}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu = cpu_baseline(model, offs_all, mp_all, sp_all, P0)

    if rank == 0:
        line = {
            "metric": "lm_iterations_per_sec", "value": value, "unit": "LM iterations/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": tot_ms / args.steps,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": dict(workload_config(args, c, offs_all), damping_ladder=args.ladder, exchange={
                "none": "none (one GPU)",
                "peer": "reduced system (161 doubles) + trial cost summed inside k_schur / k_cost through NVLink peer memory",
                "nccl": "two ncclAllReduce per iteration (161 doubles, 1 double)"}[transport]),
            "jacobian_rows_per_sec": rows_per_s, "lm_iterations_per_step": tot_it / args.steps,
            "jacobian_passes_per_step": tot_eval / args.steps, "final_sse": sse,
            "wall_ms_per_step_incl_flush": 1e3 * t_wall / args.steps,
            "e2e": e2e, "gpu_launches": int(tot_launch), "roofline": roof, "kernel_times": kt,
            "cpu_baseline": cpu, "clocks": clk.summary(),
        }
        print(json.dumps(line), flush=True)
    prob.close()
    lib.calib_host_free(mp_ptr)
    lib.calib_host_free(sp_ptr)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
