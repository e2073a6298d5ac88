import os, sys, time
sys.path.insert(0, os.getcwd())
import camera_calibration_b200 as cc
from camera_calibration_b200 import synth
c = synth.config("cfg3", 1.0)
cal = cc.Calibrator(c["model"]); P0 = cal._composeParameterVector(*c["init"])
def go():
    p = cc.Problem.from_detections(c["model"], c["dets"]); p.close()
for _ in range(3): go()
best = 1e9
for _ in range(10):
    t0 = time.perf_counter(); go(); best = min(best, time.perf_counter() - t0)
print("threads", os.environ.get("CALIB_GATHER_THREADS"), "create_from_views+destroy %.3f ms" % (best * 1e3))
