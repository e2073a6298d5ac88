import os, sys, time
sys.path.insert(0, os.getcwd())
import numpy as np
import camera_calibration_b200 as cc
from camera_calibration_b200 import synth
from camera_calibration_b200.calibrate import PinnedBuffers
c = synth.config("cfg3", 1.0)
cal = cc.Calibrator(c["model"]); P0 = cal._composeParameterVector(*c["init"]).ravel()
pin = PinnedBuffers()
n = sum(len(m) for s, m in c["dets"])
offs, mp, sp = cc.flatten_detections(c["dets"], out=pin.reserve(n))
def go():
    p = cc.Problem(c["model"], offs, mp, sp); r = p.lm_run(P0, 100, want_trace=False); p.close()
for _ in range(5): go()
os.environ["X"]="1"
ts=[]
for _ in range(20):
    t0 = time.perf_counter(); go(); ts.append(time.perf_counter() - t0)
print("C-ABI create+run+destroy: best %.3f ms median %.3f ms" % (min(ts)*1e3, sorted(ts)[10]*1e3))
