"""Device time of a refinement against the damping ladder's cap and its start after an accepted step.
usage: python profiles/ladder_probe.py cfg3@1 cfg1@1 ..."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import camera_calibration_b200 as cc
from camera_calibration_b200 import synth
for a in sys.argv[1:]:
    name, scale = a.split("@"); scale = float(scale)
    c = synth.config(name, scale) if name in ("cfg3", "cfg5") else synth.config(name)
    if c["init"] is None:
        P0 = np.load(os.path.join(ROOT, "tests", "golden", name + ".npz"))["P0"]
    else:
        P0 = cc.Calibrator(c["model"])._composeParameterVector(*c["init"]).ravel()
    offs, mp, sp = cc.flatten_detections(c["dets"])
    prob = cc.Problem(c["model"], offs, mp, sp)
    for cap, start in ((1, 1), (4, 1), (4, 2), (4, 0)):
        prob.set_ladder(cap); prob.set_ladder_start(start)
        best = 1e9
        for rep in range(8):
            sse, P, it, tr = prob.lm_run(P0, 100)
            ms, nl, ne = prob.last_run_stats()
            best = min(best, ms)
        acc = tr[:, 3]
        bounce = int(sum(1 for i in range(len(acc) - 1) if acc[i] == 1 and acc[i + 1] == 0))
        print(f"{name}@{scale} ladder cap {cap} start {start}: {it} iterations ({int(acc.sum())} accepted, {bounce} of them followed by a "
              f"rejected one), {ne} evaluations, {(nl - 1) // 4} passes, best {best * 1e3:.1f} us", flush=True)
    prob.close()
