"""k_eval variants: kernel time and bit-identity of a run (hash of the final parameters and the trace).
Captures r4a-r4e were taken with builds that still had every variant behind CALIB_EVAL_ROLES (2 = warp pairs,
3 = warp triples, 4 = three-stage pipeline); the product now has the triples only -- compare other builds of the
library with CALIB_B200_SO=<path> python profiles/eval_roles.py child <workload> <scale>.
usage: python profiles/eval_roles.py [variants]     (spawns itself once per variant and workload)"""
import hashlib
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

if len(sys.argv) > 1 and sys.argv[1] == "child":
    import numpy as np
    import camera_calibration_b200 as cc
    from camera_calibration_b200 import synth
    name, scale = sys.argv[2], float(sys.argv[3])
    c = synth.config(name, scale) if name in ("cfg3", "cfg5") else synth.config(name)
    if c["init"] is None:
        P0 = np.load(os.path.join(ROOT, "tests", "golden", name + ".npz"))["P0"]
    else:
        P0 = cc.Calibrator(c["model"])._composeParameterVector(*c["init"]).ravel()
    offs, mp, sp = cc.flatten_detections(c["dets"])
    prob = cc.Problem(c["model"], offs, mp, sp)
    warm = prob.time_kernel(P0, 0, 50)
    cold = prob.time_kernel(P0, 0, 10, flush=True)
    best = 1e9
    for _ in range(5):
        sse, P, iters, tr = prob.lm_run(P0, 100)
        best = min(best, prob.last_run_stats()[0])
    h = hashlib.sha1(P.tobytes() + tr.tobytes()).hexdigest()[:16]
    print(f"{name}@{scale} roles={os.environ.get('CALIB_EVAL_ROLES', '3')}: N={offs[-1]} k_eval warm {warm:.2f} us cold {cold:.2f} us; "
          f"lm_run {iters} iterations best {best * 1e3:.1f} us sse={sse!r} hash={h}", flush=True)
    prob.close()
else:
    variants = sys.argv[1].split(",") if len(sys.argv) > 1 else ["2", "3", "4"]
    for name, scale in (("cfg3", 1.0), ("cfg5", 0.2), ("cfg1", 1.0), ("cfg3", 0.125)):
        for roles in variants:
            env = dict(os.environ, CALIB_EVAL_ROLES=roles)
            subprocess.run([sys.executable, os.path.abspath(__file__), "child", name, str(scale)], env=env, check=False)
