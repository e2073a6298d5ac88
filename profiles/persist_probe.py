import sys, time, os, numpy as np
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/oracle')
import camera_calibration_b200 as cc
from camera_calibration_b200 import synth
def run(name, scale, modes=(1,2,3)):
    c = synth.config(name, scale) if name in ("cfg3","cfg5") else synth.config(name)
    if c["init"] is None:
        g = np.load(f"/root/repo/tests/golden/{name}.npz"); P0 = g["P0"].copy()
    else:
        P0 = cc.Calibrator(c["model"])._composeParameterVector(*c["init"]).ravel()
    offs, mp, sp = cc.flatten_detections(c["dets"])
    prob = cc.Problem(c["model"], offs, mp, sp)
    ref = None
    for mode in modes:
        prob.set_loop_mode(mode)
        best = 1e9
        for rep in range(6):
            s, P, it, tr = prob.lm_run(P0, 100)
            ms, nl, ne = prob.last_run_stats()
            best = min(best, ms)
        same = None
        if ref is None: ref = (s, P, it, tr)
        else: same = (ref[0] == s and ref[2] == it and np.array_equal(ref[1], P) and np.array_equal(ref[3], tr))
        print(f"{name}@{scale} mode {mode}: iters {it} evals {ne} launches {nl} best {best*1e3:.1f} us  ({best*1e3/it:.2f} us/iter) identical_to_mode1={same}", flush=True)
    prob.close()
for a in sys.argv[1:]:
    n, s = a.split("@"); run(n, float(s))
