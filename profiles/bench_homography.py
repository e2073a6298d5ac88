"""Per-view homography refinement (calib_refine_homographies): device path through
the C-ABI (host H in, host H out, detections resident) against the numpy oracle on a sample of views.
usage: python profiles/bench_homography.py [cfg3] [scale] [cpu_views]"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import camera_calibration_b200 as cc  # noqa: E402
from camera_calibration_b200 import synth  # noqa: E402
import calib_oracle as onp  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "cfg3"
scale = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
cpu_views = int(sys.argv[3]) if len(sys.argv) > 3 else 20
c = synth.config(name, scale)
dets = c["dets"]
offs, mp, sp = cc.flatten_detections(dets)
M = len(dets)
# initial homographies: A [r1 r2 t] of the perturbed initial pose (what the DLT would be close to)
A0, W0, k0 = c["init"]
H0 = np.array([(A0 @ np.asarray(W)[:3, [0, 1, 3]]).ravel() for W in W0])
H0 /= H0[:, 8:9]
prob = cc.Problem(c["model"], offs, mp, sp)
for _ in range(3):
    Href, iters, _ = prob.refine_homographies(H0, 20)
t0 = time.perf_counter()
reps = 20
for _ in range(reps):
    Href, iters, _ = prob.refine_homographies(H0, 20)
gpu_s = (time.perf_counter() - t0) / reps
prob.close()
t0 = time.perf_counter()
worst = 0.0
for v in range(min(cpu_views, M)):
    Ho, ito, _ = onp.homography_lm(H0[v], dets[v][0], dets[v][1])
    worst = max(worst, float((np.abs(Href[v] - Ho) / np.maximum(np.abs(Ho), 1e-3)).max()))
cpu_s_per_view = (time.perf_counter() - t0) / min(cpu_views, M)
print(json.dumps({"workload": name, "scale": scale, "views": M, "points": int(offs[-1]),
                  "gpu_ms_all_views_e2e": 1e3 * gpu_s, "gpu_views_per_s": M / gpu_s,
                  "lm_iterations_total": int(iters.sum()),
                  "cpu_oracle_ms_per_view": 1e3 * cpu_s_per_view, "cpu_views_per_s": 1.0 / cpu_s_per_view,
                  "cpu_sample_views": min(cpu_views, M), "max_rel_diff_vs_oracle_on_sample": worst}))
