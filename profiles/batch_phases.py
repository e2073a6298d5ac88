"""Where k_batch_lm spends its cycles (profiling build: -DCALIB_STAMPS, see profiles/stamps.py).
usage: CALIB_B200_SO=$PWD/camera_calibration_b200/libcalib_b200_stamps.so python profiles/batch_phases.py [n_problems]"""
import ctypes
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import camera_calibration_b200 as cc  # noqa: E402
from camera_calibration_b200 import synth, _lib  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
probs = synth.batch_config(n, procs=min(os.cpu_count() or 1, 16))
cal = cc.Calibrator(0)
P0 = np.concatenate([cal._composeParameterVector(*p["init"]).ravel() for p in probs])
bp = cc.BatchProblem(0, [p["dets"] for p in probs])
L = _lib.lib()
fn = L.calib_batch_debug_phases
fn.restype = ctypes.c_int
fn.argtypes = [ctypes.c_void_p, ctypes.POINTER(ctypes.c_uint64)]
out = (ctypes.c_uint64 * 12)()
for i in range(2):
    sse, P, iters, tr, ms = bp.lm_run(P0, 100)
assert fn(bp._h, out) == 0          # allocates the counters
sse, P, iters, tr, ms = bp.lm_run(P0, 100)
assert fn(bp._h, out) == 0
v = np.array(list(out), dtype=np.float64)
names = ["evaluation sweeps (2)", "sum A + 6x6 factors + forward substitutions", "-", "W", "-",
         "dense form + reduced solve (1 warp)", "back-substitution + view constants", "trial cost sweep", "accept / reject"]
nprob, its, total = v[11], v[10], v[9]
print(f"cfg4 x {n}: device {ms:.3f} ms, mean iterations {its / nprob:.2f}; cycles per problem {total / nprob:.0f} "
      f"(block-resident time; 2 blocks share an SM)")
print(f"{'phase':45s} {'cycles/problem':>15s} {'share':>7s} {'cycles/iteration':>17s}")
for i, nm in enumerate(names):
    if nm != "-":
        print(f"{nm:45s} {v[i] / nprob:15.0f} {100 * v[i] / total:6.1f}% {v[i] / its:17.0f}")
print(f"{'sum of phases':45s} {v[:9].sum() / nprob:15.0f} {100 * v[:9].sum() / total:6.1f}%")
bp.close()
