"""Small driver for ncu: one (or a few) LM runs of a named workload through the C-ABI.
usage: python profiles/run_lm.py [cfg3|cfg5|itest|cfg1|cfg2] [scale] [runs] [loop_mode]
(loop_mode 1 = host-driven iteration graphs: ncu cannot see kernels inside a WHILE node's body)"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import camera_calibration_b200 as cc  # noqa: E402
from camera_calibration_b200 import synth  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "cfg3"
scale = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
runs = int(sys.argv[3]) if len(sys.argv) > 3 else 2
loop_mode = int(sys.argv[4]) if len(sys.argv) > 4 else 0
t0 = time.time()
c = synth.config(name, scale) if name in ("cfg3", "cfg5") else synth.config(name)
if c["init"] is None:
    P0 = np.load(os.path.join(ROOT, "tests", "golden", name + ".npz"))["P0"]
else:
    P0 = cc.Calibrator(c["model"])._composeParameterVector(*c["init"]).ravel()
offs, mp, sp = cc.flatten_detections(c["dets"])
print(f"{name}: M={len(offs) - 1} N={offs[-1]} (data {time.time() - t0:.1f}s)", flush=True)
prob = cc.Problem(c["model"], offs, mp, sp)
prob.set_loop_mode(loop_mode)
for i in range(runs):
    sse, P, iters, tr = prob.lm_run(P0, 100)
    ms, launches, evals = prob.last_run_stats()
    print(f"run {i}: iters={iters} sse={sse:.6g} device_ms={ms:.3f} launches={launches} evals={evals}", flush=True)
prob.close()
