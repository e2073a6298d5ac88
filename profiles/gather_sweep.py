"""How the per-view upload (calib_create_from_views: threaded gather into page-locked staging, pipelined with
the PCIe copies) depends on the helper threads, the staging slots per thread and the chunk size.
usage: python profiles/gather_sweep.py           (spawns one process per setting: the library reads the
environment once)"""
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

if len(sys.argv) > 1 and sys.argv[1] == "one":
    sys.path.insert(0, ROOT)
    import numpy as np
    import camera_calibration_b200 as cc
    from camera_calibration_b200 import synth
    c = synth.config("cfg3", 1.0)
    ts = []
    for i in range(23):
        t0 = time.perf_counter()
        p = cc.Problem.from_detections(c["model"], c["dets"])
        ts.append(time.perf_counter() - t0)
        p.close()
    ts = np.array(ts[3:]) * 1e3
    print(f"{np.mean(ts):7.3f} {np.median(ts):7.3f} {np.min(ts):7.3f}")
    sys.exit(0)

print("threads slots chunk_KB | create from views, ms: mean median min (20 calls, cfg3: 62 MB in 2000 arrays)")
for T in (2, 4, 6, 8):
    for S in (2, 3, 4):
        for C in (256, 512, 1024, 2048):
            env = dict(os.environ, CALIB_GATHER_THREADS=str(T), CALIB_GATHER_SLOTS=str(S), CALIB_GATHER_CHUNK_KB=str(C))
            out = subprocess.run([sys.executable, __file__, "one"], env=env, capture_output=True, text=True).stdout.strip()
            print(f"{T:7d} {S:5d} {C:8d} | {out}", flush=True)
