#!/usr/bin/env python
"""Benchmark of the LM refinement hot path (BASELINE.json metric: LM iterations/s and Jacobian
rows/s).  One "step" = one complete run of the LM loop (Calibrator.refineCalibrationParameters,
maxIters=100, until the reference's own termination rule fires) on the named workload.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload cfg3|cfg5|itest] [--impl reference]

value   device-resident: detections already in HBM, parameters uploaded before the timed region,
        timed with CUDA events on the library's stream (max over ranks).
e2e     through the public API with host numpy buffers: flattening, H2D of the detections,
        the loop, D2H of the result -- all inside the timed region.
--impl reference  times the CPU oracle port (oracle/calib_oracle.c, OpenMP, all host threads) on
        a bounded sample of the same workload (the reference itself is Python and cannot travel).
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
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

MAX_ITERS = 100
# algorithmic FP64 work per point of the passes (DESIGN.md "Kernels"): dense-block formulation of
# F_model + 2(L+6)(L+7) + 4(L+6) + 4, split over the two fused passes
FLOPS = {0: dict(model=400, A=2 * 2 * (55 + 10) + 4, BD=2 * 2 * (60 + 21 + 6)),
         1: dict(model=450, A=2 * 2 * (45 + 9) + 4, BD=2 * 2 * (54 + 21 + 6))}
BYTES_PER_POINT = 40


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


def shard_views(dets, P0, L, world, rank):
    """contiguous view ranges balanced by point count; every rank keeps the L shared values"""
    counts = np.array([len(s) for s, m in dets], dtype=np.int64)
    cum = np.concatenate([[0], np.cumsum(counts)])
    total = cum[-1]
    bounds = [int(np.searchsorted(cum, total * r / world, side="left")) for r in range(world)] + [len(dets)]
    bounds[0] = 0
    lo, hi = bounds[rank], bounds[rank + 1]
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
            time.sleep(0.1)

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
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(self.rows[0][1]), "reasons": reasons,
                "samples": len(self.rows)}


def cpu_baseline(model, offs, mp, sp, P0, budget_s=12.0):
    """The oracle port timed on the host cores (bounded sample: a few LM iterations)."""
    import oracle_c
    threads = oracle_c.num_threads()
    t0 = time.perf_counter()
    _, _, it, _ = oracle_c.lm(model, P0, offs, mp, sp, 2)
    dt = time.perf_counter() - t0
    n = int(max(2, min(MAX_ITERS, budget_s / max(dt / it, 1e-6))))
    t0 = time.perf_counter()
    _, _, it, tr = oracle_c.lm(model, P0, offs, mp, sp, n)
    dt = time.perf_counter() - t0
    evals = 1 + int(tr[:-1, 3].sum())
    return dict(value=it / dt, unit="LM iterations/s", cores=threads, kind="port",
                sample=f"first {it} LM iterations of the workload ({dt:.2f} s), C oracle with OpenMP",
                jacobian_rows_per_s=2.0 * offs[-1] * evals / dt), it, dt


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import camera_calibration_b200 as cc
    c, P0 = load_workload(args.workload, args.scale)
    offs, mp, sp = cc.flatten_detections(c["dets"])
    import oracle_c
    per_step = max(2, int(args.ref_iters))
    for _ in range(args.warmup):
        oracle_c.lm(c["model"], P0, offs, mp, sp, 1)
    tot_it, evals = 0, 0
    t0 = time.perf_counter()
    for _ in range(args.steps):
        _, _, it, tr = oracle_c.lm(c["model"], P0, offs, mp, sp, per_step)
        tot_it += it
        evals += 1 + int(tr[:-1, 3].sum())
    dt = time.perf_counter() - t0
    val = tot_it / dt
    line = {
        "impl": "reference", "metric": "lm_iterations_per_sec", "value": val, "unit": "LM iterations/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "jacobian_rows_per_sec": 2.0 * offs[-1] * evals / dt,
        "config": workload_config(args, c, offs),
        "cpu_baseline": {"value": val, "unit": "LM iterations/s", "cores": oracle_c.num_threads(), "kind": "port",
                         "sample": f"each step = first {per_step} LM iterations of the workload; C oracle "
                                   f"(oracle/calib_oracle.c, OpenMP); the Python reference cannot travel to this box"},
        "e2e": {"value": val, "unit": "LM iterations/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def workload_config(args, c, offs):
    return {"workload": args.workload + ("" if args.scale == 1.0 else f"@scale{args.scale}"),
            "model": "radtan" if c["model"] == 0 else "fisheye", "views": int(len(offs) - 1),
            "points": int(offs[-1]), "params": int((10 if c["model"] == 0 else 9) + 6 * (len(offs) - 1)),
            "max_iters": MAX_ITERS, "init": "perturbed ground truth, RandomState(12345)",
            "parallelism": f"views sharded over {args.gpus} GPU(s)" if args.gpus > 1 else "1 GPU",
            "l2": "flushed (256 MB write) before every timed step; within a step the detections stay "
                  "L2-resident across LM iterations, as in production use"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default="cfg3")
    ap.add_argument("--scale", type=float, default=1.0)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--ref-iters", type=int, default=3)
    ap.add_argument("--segment", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl != "reference" else args.warmup
    if args.impl == "reference":
        return run_reference_arm(args)

    import torch
    import camera_calibration_b200 as cc
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    torch.cuda.set_device(local)

    c, P0 = load_workload(args.workload, args.scale)
    model = c["model"]
    L = 10 if model == 0 else 9
    offs_all, mp_all, sp_all = cc.flatten_detections(c["dets"])
    dets, Ploc, (lo, hi) = shard_views(c["dets"], P0, L, world, rank)
    offs, mp, sp = cc.flatten_detections(dets)
    prob = cc.Problem(model, offs, mp, sp, device=local)
    if args.segment:
        prob.set_segment(args.segment)
    if world > 1:
        idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            idt = torch.frombuffer(bytearray(cc.calibrate.comm_unique_id()), dtype=torch.uint8).cuda()
        dist.broadcast(idt, 0)
        prob.comm_init(world, rank, bytes(idt.cpu().numpy().tobytes()))

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
    Pfinal = prob.download()

    # ---- e2e: the public API with host buffers (flatten + H2D + loop + D2H every step)
    e2e = None
    if world == 1:
        cal = cc.Calibrator(model, device=local)
        A0, W0, k0 = cal._decomposeParameterVector(P0)
        for _ in range(2):
            cal.refineCalibrationParameters(A0, W0, k0, c["dets"], MAX_ITERS)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        it_e2e = 0
        n_e2e = max(3, min(args.steps, 10))
        for _ in range(n_e2e):
            cal.refineCalibrationParameters(A0, W0, k0, c["dets"], MAX_ITERS)
            it_e2e += cal.lastIters
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        e2e = {"value": it_e2e / dt, "unit": "LM iterations/s", "h2d_bytes_per_step": int(40 * N_all + 8 * len(P0)),
               "d2h_bytes_per_step": int(8 * len(P0) + 32 * MAX_ITERS), "ms_per_step": 1e3 * dt / n_e2e,
               "api": "Calibrator.refineCalibrationParameters(A, W, k, allDetections, maxIters)"}
    else:
        # sharded: every step creates the rank's problem from host buffers (H2D), borrows the
        # long-lived communicator, runs the loop and reads the result back; max over ranks
        def e2e_step():
            p2 = cc.Problem(model, offs, mp, sp, device=local)
            p2.comm_share(prob)
            _, _, it, _ = p2.lm_run(Ploc, MAX_ITERS, want_trace=False)
            p2.close()
            return it
        e2e_step()
        barrier()
        t0 = time.perf_counter()
        it_e2e = 0
        n_e2e = max(3, min(args.steps, 10))
        for _ in range(n_e2e):
            it_e2e += e2e_step()
        barrier()
        t = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt = float(t.item())
        e2e = {"value": it_e2e / dt, "unit": "LM iterations/s",
               "h2d_bytes_per_step": int(40 * offs[-1] + 8 * len(Ploc)), "d2h_bytes_per_step": int(8 * len(Ploc)),
               "ms_per_step": 1e3 * dt / n_e2e,
               "api": "per rank: Problem(host detections) + comm_share + lm_run(host P) + close"}

    # ---- roofline of the dominant kernel (pass BD), events around every launch, separate run
    roof = roof_hbm = None
    kt = {}
    if rank == 0 or world > 1:
        prob.set_profiling(True)
        prob.upload(Ploc)
        prob.flush_l2()
        prob.lm_run_resident(MAX_ITERS)
        for i, nm in enumerate(["pass_A", "pass_BD", "cost", "schur", "solve_backsub"]):
            ms_k, n_k = prob.kernel_time(i)
            kt[nm] = {"launches": n_k, "avg_us": 1e3 * ms_k / max(n_k, 1)}
        prob.set_profiling(False)
    if rank == 0:
        peak64 = cc.calibrate.fp64_peak_tflops(local)
        n_loc = int(offs[-1])
        # only launches that did work: evals of the profiled run
        _, _, ev_prof = prob.last_run_stats()
        ms_bd, n_bd = prob.kernel_time(1)
        # skipped launches (rejected iterations reuse the normal equations) return immediately;
        # charge the whole measured time to the launches that evaluated
        avg_bd_s = 1e-3 * ms_bd / max(ev_prof, 1)
        fl = (FLOPS[model]["model"] + FLOPS[model]["BD"]) * n_loc
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(
            os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
        hbm_peak = peaks.get("hbm_gbs", 6650.0)
        roof = {"bound": "fp64", "kernel": "k_eval_BD", "achieved": fl / avg_bd_s / 1e12, "peak": peak64,
                "unit": "TFLOP/s", "frac": fl / avg_bd_s / 1e12 / peak64, "traffic": None,
                "peak_source": "DFMA micro-benchmark in this run (calib_measure_fp64_peak); FP64 is not in "
                               "MEASURED_PEAKS.json",
                "flops_per_point": FLOPS[model]["model"] + FLOPS[model]["BD"], "avg_launch_us": avg_bd_s * 1e6,
                "hbm": {"achieved": BYTES_PER_POINT * n_loc / avg_bd_s / 1e9, "peak": hbm_peak, "unit": "GB/s",
                        "frac": BYTES_PER_POINT * n_loc / avg_bd_s / 1e9 / hbm_peak,
                        "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback"}}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu, _, _ = cpu_baseline(model, offs_all, mp_all, sp_all, P0)

    if rank == 0:
        line = {
            "metric": "lm_iterations_per_sec", "value": value, "unit": "LM iterations/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": tot_ms / args.steps,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": workload_config(args, c, offs_all),
            "jacobian_rows_per_sec": rows_per_s, "lm_iterations_per_step": tot_it / args.steps,
            "jacobian_passes_per_step": tot_eval / args.steps, "final_sse": sse,
            "wall_ms_per_step_incl_flush": 1e3 * t_wall / args.steps,
            "e2e": e2e, "gpu_launches": int(tot_launch), "roofline": roof, "kernel_times": kt,
            "cpu_baseline": cpu, "clocks": clk.summary(),
        }
        print(json.dumps(line), flush=True)
    prob.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
