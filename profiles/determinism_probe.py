"""Run-to-run and staging-mode determinism of a refinement (bit level).  usage: determinism_probe.py cfg3 0.03"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import camera_calibration_b200 as cc
from camera_calibration_b200 import synth
name = sys.argv[1]; scale = float(sys.argv[2])
c = synth.config(name, scale)
P0 = cc.Calibrator(c["model"])._composeParameterVector(*c["init"]).ravel()
offs, mp, sp = cc.flatten_detections(c["dets"])
prob = cc.Problem(c["model"], offs, mp, sp)
ref = None
for mode in (0, 0, 1, 1, 0):
    prob.set_cost_staging(mode)
    sse, P, it, tr = prob.lm_run(P0, 100)
    A, B, D, g, E = prob.normal_equations(P0)
    if ref is None:
        ref = (sse, P, it, tr, A, B, D, g)
    same_tr = np.array_equal(tr, ref[3]) if len(tr) == len(ref[3]) else False
    first = None
    if not same_tr:
        n = min(len(tr), len(ref[3]))
        d = np.where((tr[:n] != ref[3][:n]).any(axis=1))[0]
        first = (int(d[0]), tr[d[0]].tolist(), ref[3][d[0]].tolist()) if len(d) else ("len", len(tr), len(ref[3]))
    print(f"staging {mode}: sse {sse!r} iters {it} trace_equal {same_tr} first_diff {first} "
          f"normal_eq_equal {np.array_equal(A, ref[4]) and np.array_equal(B, ref[5]) and np.array_equal(D, ref[6]) and np.array_equal(g, ref[7])}")
prob.close()
