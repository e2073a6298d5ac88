"""Small driver for ncu / timing: one batch refinement (cfg4 shape: 20-view, 9x6-board problems).
usage: python profiles/run_batch.py [n_problems] [runs]"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import camera_calibration_b200 as cc  # noqa: E402
from camera_calibration_b200 import synth  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 512
runs = int(sys.argv[2]) if len(sys.argv) > 2 else 2
t0 = time.time()
probs = synth.batch_config(n, procs=min(os.cpu_count() or 1, 16))
cal = cc.Calibrator(0)
P0 = np.concatenate([cal._composeParameterVector(*p["init"]).ravel() for p in probs])
bp = cc.BatchProblem(0, [p["dets"] for p in probs])
print(f"cfg4 x {n} problems (data {time.time() - t0:.1f}s)", flush=True)
for i in range(runs):
    sse, P, iters, tr, ms = bp.lm_run(P0, 100)
    print(f"run {i}: device_ms={ms:.3f} mean iters {iters.mean():.2f}  {n / ms * 1e3:.0f} calibrations/s", flush=True)
bp.close()
