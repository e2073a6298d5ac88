"""Latency study: globaltimer stamps inside the four kernels of one pass (build with -DCALIB_STAMPS,
same nvcc line as camera_calibration_b200/_lib.py NVCC_FLAGS plus -DCALIB_STAMPS, output e.g.
camera_calibration_b200/libcalib_b200_stamps.so; the product library carries no stamps).
usage: CALIB_B200_SO=$PWD/camera_calibration_b200/libcalib_b200_stamps.so python profiles/stamps.py [cfg3] [scale]"""
import ctypes
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import camera_calibration_b200 as cc  # noqa: E402
from camera_calibration_b200 import _lib, synth  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "cfg3"
scale = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
c = synth.config(name, scale) if name in ("cfg3", "cfg5") else synth.config(name)
if c["init"] is None:
    P0 = np.load(os.path.join(ROOT, "tests", "golden", name + ".npz"))["P0"]
else:
    P0 = cc.Calibrator(c["model"])._composeParameterVector(*c["init"]).ravel()
offs, mp, sp = cc.flatten_detections(c["dets"])
prob = cc.Problem(c["model"], offs, mp, sp)
L = _lib.lib()
KB = 1280
out = np.zeros((4, KB, 8), dtype=np.uint64)
fn = L.calib_debug_stamps
fn.restype = ctypes.c_int
fn.argtypes = [ctypes.c_void_p, ctypes.POINTER(ctypes.c_double), ctypes.c_double, ctypes.c_void_p]
P = np.ascontiguousarray(P0, dtype=np.float64)


def report(title):
    t = out.astype(np.float64)
    t[t == 0] = np.nan
    t0 = np.nanmin(t[1:]) if np.isnan(t[0]).all() else np.nanmin(t)
    t = (t - t0) * 1e-3   # us since the first stamp of the pass
    names = ["k_eval", "k_schur", "k_backsub", "k_cost"]
    slots = {0: ["entry", "sweep1 done", "sweep2 done", "after block barrier", "end"],
             1: ["entry", "view rows loaded (last view)", "view phase done", "ticket taken", "last: A gathered",
                 "last: red written", "last: solved"],
             2: ["entry", "end"],
             3: ["entry", "sweep done", "ticket taken", "last: end"]}
    print(f"{title}: M={len(offs) - 1} N={offs[-1]}   (us since the first stamp; min / median / max over blocks)")
    for k in range(4):
        print(names[k])
        for s_, label in enumerate(slots[k]):
            col = t[k, :, s_]
            col = col[~np.isnan(col)]
            if len(col) == 0:
                continue
            print(f"   {label:34s} n={len(col):5d}  {col.min():8.2f} {np.median(col):8.2f} {col.max():8.2f}")


rc = fn(prob._h, P.ctypes.data_as(ctypes.POINTER(ctypes.c_double)), 1e-3, out.ctypes.data_as(ctypes.c_void_p))
assert rc == 0
report(f"{name} x{scale}, one pass, plain stream launches, J = 1")
# the last pass of a whole refinement inside the WHILE graph (rejected steps: no evaluation, J = 4):
# stamps of earlier passes are overwritten, k_eval's are from the last evaluation (ignore them)
for mode, label in ((2, "WHILE graph"), (1, "host-driven iteration graphs")):
    prob.set_loop_mode(mode)
    sse, Pr, iters, tr = prob.lm_run(P0, 100)
    rc = fn(prob._h, None, 0.0, out.ctypes.data_as(ctypes.c_void_p))
    assert rc == 0
    out[0] = 0
    report(f"{name} x{scale}, last pass of a {iters}-iteration run, {label}")
prob.close()
