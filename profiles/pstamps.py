"""Latency study of the persistent kernel: globaltimer stamps at every phase boundary of every pass
(first and last block).  Needs the -DCALIB_STAMPS build of the library:
  nvcc <NVCC_FLAGS of camera_calibration_b200/_lib.py> -DCALIB_STAMPS -o camera_calibration_b200/libcalib_b200_stamps.so ...
usage: CALIB_B200_SO=$PWD/camera_calibration_b200/libcalib_b200_stamps.so python profiles/pstamps.py cfg3 1.0"""
import ctypes
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import camera_calibration_b200 as cc  # noqa: E402
from camera_calibration_b200 import _lib, synth  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "cfg3"
scale = float(sys.argv[2]) if len(sys.argv) > 2 else 1.0
c = synth.config(name, scale) if name in ("cfg3", "cfg5") else synth.config(name)
if c["init"] is None:
    P0 = np.load(os.path.join(ROOT, "tests", "golden", name + ".npz"))["P0"]
else:
    P0 = cc.Calibrator(c["model"])._composeParameterVector(*c["init"]).ravel()
offs, mp, sp = cc.flatten_detections(c["dets"])
prob = cc.Problem(c["model"], offs, mp, sp)
prob.set_loop_mode(3)
L = _lib.lib()
NP, NS = 128, 16
fn = L.calib_debug_pstamps
fn.restype = ctypes.c_int
fn.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
out = np.zeros((2, NP, NS), dtype=np.uint64)
for rep in range(3):
    sse, P, iters, tr = prob.lm_run(P0, 100)
    assert fn(prob._h, out.ctypes.data_as(ctypes.c_void_p)) == 0
ms, launches, evals = prob.last_run_stats()
t = out.astype(np.float64)
t[t == 0] = np.nan
npass = int(np.sum(~np.isnan(t[0, :, 0])))
print(f"{name} x{scale}: M={len(offs) - 1} N={offs[-1]}  {iters} iterations, {evals} evaluations, {npass} passes, {ms * 1e3:.1f} us")
labels = ["eval", "bar", "views", "bar", "tail(b0)", "bar", "backsub", "bar", "cost", "bar", "accept->next"]
edges = [(0, 1), (1, 2), (2, 3), (3, 4), (4, 5), (5, 6), (6, 7), (7, 8), (8, 9), (9, 10), (10, 16)]
for blk in (0, 1):
    print(f"block {'0' if blk == 0 else 'last'}: per pass, us  [" + " | ".join(labels) + "]")
    tot = np.zeros(len(edges))
    for p in range(npass):
        row = t[blk, p]
        nxt = t[blk, p + 1, 0] if p + 1 < npass else np.nan
        vals = []
        for k, (a, b) in enumerate(edges):
            ta = row[a]
            if a == 2 and np.isnan(row[2]):
                ta = row[0]             # no evaluation in this pass
            tb = nxt if b == 16 else row[b]
            v = (tb - ta) * 1e-3
            vals.append(v)
            if not np.isnan(v):
                tot[k] += v
        J = "eval" if not np.isnan(row[1]) else "    "
        if p < 12 or p >= npass - 3:
            print(f"  pass {p:3d} {J} " + " ".join(f"{v:7.2f}" if not np.isnan(v) else "      -" for v in vals))
    print("  total     " + " ".join(f"{v:7.1f}" for v in tot) + f"   sum {tot.sum():.1f}")
prob.close()

# in-phase stamps (STAMP in the phase functions) of the LAST pass of a run, both loop modes
KB = 1280
fn2 = L.calib_debug_stamps
fn2.restype = ctypes.c_int
fn2.argtypes = [ctypes.c_void_p, ctypes.POINTER(ctypes.c_double), ctypes.c_double, ctypes.c_void_p]
st = np.zeros((4, KB, 8), dtype=np.uint64)
prob2 = cc.Problem(c["model"], offs, mp, sp)
for mode, label in ((3, "persistent"), (2, "WHILE graph")):
    prob2.set_loop_mode(mode)
    for rep in range(2):
        sse, P, iters, tr = prob2.lm_run(P0, 100)
    if mode == 3:
        assert fn(prob2._h, out.ctypes.data_as(ctypes.c_void_p)) == 0
    assert fn2(prob2._h, None, 0.0, st.ctypes.data_as(ctypes.c_void_p)) == 0
    t = st.astype(np.float64)
    t[t == 0] = np.nan
    print(f"--- last pass, {label}: in-phase stamps (us)")
    # tail: block that ran it = the one with slot 6 set
    tail_blk = np.where(~np.isnan(t[1, :, 6]))[0]
    if len(tail_blk):
        b = tail_blk[0]
        r = t[1, b]
        base = r[4] if not np.isnan(r[4]) else r[3]
        print(f"  schur tail (block {b}): A/redA gathered -> red written {1e-3 * (r[5] - r[4]):.2f}, red written -> solved {1e-3 * (r[6] - r[5]):.2f}"
              + (f", ticket -> A gathered {1e-3 * (r[4] - r[3]):.2f}" if not np.isnan(r[3]) else ""))
        if mode == 3:
            po = out.astype(np.float64)
            lastp = int(np.sum(po[0, :, 0] > 0)) - 1
            print(f"     persistent: phase start -> A gathered {1e-3 * (r[4] - po[0, lastp, 4]):.2f}, solved -> phase end {1e-3 * (po[0, lastp, 5] - r[6]):.2f}")
    v = t[1, :, 2] - t[1, :, 1]
    v = v[~np.isnan(v)]
    if len(v):
        print(f"  views stage: rows loaded -> view phase done: median {1e-3 * np.median(v):.2f} max {1e-3 * v.max():.2f} (n={len(v)})")
    cs = t[3, :, 1] - t[3, :, 0]
    cs = cs[~np.isnan(cs)]
    if len(cs):
        print(f"  cost sweep (entry -> sweep done): min {1e-3 * cs.min():.2f} median {1e-3 * np.median(cs):.2f} max {1e-3 * cs.max():.2f} (n={len(cs)})")
    bs = t[2, :, 1] - t[2, :, 0]
    bs = bs[~np.isnan(bs)]
    if len(bs):
        print(f"  backsub (entry -> end): median {1e-3 * np.median(bs):.2f} max {1e-3 * bs.max():.2f}")
prob2.close()
