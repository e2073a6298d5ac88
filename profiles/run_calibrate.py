"""Small driver for ncu: the whole Calibrator.calibrate() (DLT, homography LM, intrinsics / extrinsics,
distortion estimate, LM refinement) once on a named workload, so that every kernel of the path shows up.
usage: CALIB_LOOP_MODE=1 python profiles/run_calibrate.py [cfg3|cfg5] [scale]
(loop mode 1 = host-driven iteration graphs: ncu cannot see kernels inside a WHILE node's body)"""
import contextlib
import io
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import camera_calibration_b200 as cc  # noqa: E402
from camera_calibration_b200 import synth  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "cfg3"
scale = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
c = synth.config(name, scale)
cal = cc.Calibrator(c["model"])
for rep in range(2):
    t0 = time.perf_counter()
    with contextlib.redirect_stdout(io.StringIO()):
        sse, A, W, k = cal.calibrate(c["dets"], 100)
    print(f"calibrate() {name} x{scale}: {len(c['dets'])} views, sse {sse:.6g}, {cal.lastIters} LM iterations, "
          f"{1e3 * (time.perf_counter() - t0):.2f} ms", flush=True)
