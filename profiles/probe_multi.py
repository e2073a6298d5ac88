"""Host-side breakdown of the single-process multi-device path (calib_multi_*).
usage: [CALIB_DEBUG_TIMING=1] python profiles/probe_multi.py [n_gpus] [cfg3] [scale]"""
import ctypes
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import camera_calibration_b200 as cc  # noqa: E402
from camera_calibration_b200 import _lib, synth  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 2
name = sys.argv[2] if len(sys.argv) > 2 else "cfg3"
scale = float(sys.argv[3]) if len(sys.argv) > 3 else 1.0
c = synth.config(name, scale)
P0 = cc.Calibrator(c["model"])._composeParameterVector(*c["init"]).ravel()
offs, mp, sp = cc.flatten_detections(c["dets"])
L = _lib.lib()


def pinned(a):
    ptr = L.calib_host_alloc(a.nbytes)
    v = np.frombuffer((ctypes.c_char * a.nbytes).from_address(ptr), dtype=a.dtype).reshape(a.shape)
    v[...] = a
    return v


mpp, spp = pinned(mp), pinned(sp)
devs = list(range(n))
for rep in range(6):
    quiet = rep < 4
    if not quiet:
        print(f"--- rep {rep}", flush=True)
    t0 = time.perf_counter()
    m = cc.MultiProblem(c["model"], offs, mpp, spp, devs)
    t1 = time.perf_counter()
    sse, P, it, tr = m.lm_run(P0, 100, want_trace=False)
    t2 = time.perf_counter()
    sse, P, it, tr = m.lm_run(P0, 100, want_trace=False)
    t3 = time.perf_counter()
    dms = m.last_device_ms
    m.close()
    t4 = time.perf_counter()
    if not quiet:
        print(f"create {1e3 * (t1 - t0):.3f} ms, lm_run {1e3 * (t2 - t1):.3f} ms, lm_run again {1e3 * (t3 - t2):.3f} ms "
              f"(device {dms:.3f} ms, {it} iterations), destroy {1e3 * (t4 - t3):.3f} ms", flush=True)
one = cc.Problem(c["model"], offs, mpp, spp, device=0)
t0 = time.perf_counter()
one.lm_run(P0, 100, want_trace=False)
print(f"one GPU lm_run {1e3 * (time.perf_counter() - t0):.3f} ms (device {one.last_run_stats()[0]:.3f} ms)")
one.close()
