"""Where an end-to-end refinement spends its time (host side included).
usage: CALIB_DEBUG_TIMING=1 python profiles/e2e_probe.py [cfg3] [scale]"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import camera_calibration_b200 as cc  # noqa: E402
from camera_calibration_b200 import synth  # noqa: E402
from camera_calibration_b200.calibrate import PinnedBuffers  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "cfg3"
scale = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
c = synth.config(name, scale)
cal = cc.Calibrator(c["model"])
A0, W0, k0 = c["init"]


def t(label, fn, n=3):
    best = 1e9
    for _ in range(n):
        t0 = time.perf_counter()
        out = fn()
        best = min(best, time.perf_counter() - t0)
    print(f"{label:40s} {best * 1e3:9.3f} ms", flush=True)
    return out


P0 = t("compose", lambda: cal._composeParameterVector(A0, W0, k0))
pin = PinnedBuffers()
n_points = sum(len(m) for s, m in c["dets"])
offs, mp, sp = t("flatten -> pinned", lambda: cc.flatten_detections(c["dets"], out=pin.reserve(n_points)))
offs2, mp2, sp2 = t("flatten -> pageable", lambda: cc.flatten_detections(c["dets"]))
for label, (m_, s_) in (("pinned", (mp, sp)), ("pageable", (mp2, sp2))):
    prob = t(f"Problem create ({label})", lambda: cc.Problem(c["model"], offs, m_, s_), n=1)
    prob.close()
    prob = t(f"Problem create ({label}, again)", lambda: cc.Problem(c["model"], offs, m_, s_), n=1)
    res = t("lm_run (first: builds the graph)", lambda: prob.lm_run(P0, 100), n=1)
    res = t("lm_run (again)", lambda: prob.lm_run(P0, 100), n=3)
    print("   iters", res[2], "device ms", prob.last_run_stats()[0])
    t("close", prob.close, n=1)
t("decompose", lambda: cal._decomposeParameterVector(res[1]))
t("Calibrator.refineCalibrationParameters", lambda: cal.refineCalibrationParameters(A0, W0, k0, c["dets"], 100), n=4)


def cabi():
    p = cc.Problem(c["model"], offs, mp, sp)
    r = p.lm_run(P0, 100, want_trace=False)
    p.close()
    return r


t("C-ABI create+lm_run+destroy (pinned CSR)", cabi, n=5)


def views():
    p = cc.Problem.from_detections(c["model"], c["dets"])
    r = p.lm_run(P0, 100, want_trace=False)
    p.close()
    return r


from camera_calibration_b200 import calibrate as _cal  # noqa: E402
t("_view_pointers (per-view addresses)", lambda: _cal._view_pointers(c["dets"]), n=5)
t("from_detections+lm_run+destroy (per-view arrays)", views, n=5)
t("Calibrator.refineCalibrationParameters (again)", lambda: cal.refineCalibrationParameters(A0, W0, k0, c["dets"], 100), n=5)
