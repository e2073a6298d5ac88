"""Device time of each kernel of the iteration (CUDA events around back-to-back launches).
usage: python profiles/time_kernels.py [cfg3|cfg5|itest|...] [scale] [reps]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import camera_calibration_b200 as cc  # noqa: E402
from camera_calibration_b200 import synth  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "cfg3"
scale = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 50
c = synth.config(name, scale) if name in ("cfg3", "cfg5") else synth.config(name)
if c["init"] is None:
    P0 = np.load(os.path.join(ROOT, "tests", "golden", name + ".npz"))["P0"]
else:
    P0 = cc.Calibrator(c["model"])._composeParameterVector(*c["init"]).ravel()
offs, mp, sp = cc.flatten_detections(c["dets"])
prob = cc.Problem(c["model"], offs, mp, sp)
print(f"{name}: M={len(offs) - 1} N={offs[-1]}")
for which, nm in enumerate(["k_eval", "k_schur", "k_backsub", "k_cost"]):
    warm = prob.time_kernel(P0, which, reps)
    cold = prob.time_kernel(P0, which, max(reps // 5, 3), flush=True)
    print(f"{nm:10s} L2-warm {warm:8.2f} us   after L2 flush {cold:8.2f} us")
sse, P, iters, tr = prob.lm_run(P0, 100)
ms, launches, evals = prob.last_run_stats()
print(f"lm_run: iters={iters} evals={evals} device_ms={ms:.3f}")
prob.close()
