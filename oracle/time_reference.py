"""TEST / BASELINE INFRASTRUCTURE ONLY -- times the UNMODIFIED Python reference on one of its own
CPU-runnable cases and prints one JSON line.  Run as a subprocess by bench.py (cpu_baseline leg at N = 1 and
--impl reference) with CALIB_REF_ROOT pointing at the vendored copy (baseline/_ref, see
oracle/vendor_reference.py) or at /root/reference.

    CALIB_REF_ROOT=baseline/_ref python oracle/time_reference.py cfg1

The case's detections and the initial parameter vector come from tests/golden/<case>.npz (the P0 the
reference's own linear stage produced when the fixture was made).  What is timed is exactly the call the
product replaces: Calibrator.refineCalibrationParameters(A0, W0, k0, allDetections, maxIters)
(src/calibrate.py:117-171), once with the symbolic Jacobian set-up it performs on first use
(src/calibrate.py:173-176, jacobian.py:10-36) reported separately.
"""
import contextlib
import io
import json
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)


def main(case):
    import ref_harness
    ref_harness.load()
    from src import calibrate, distortion
    z = np.load(os.path.join(HERE, "..", "tests", "golden", case + ".npz"))
    model = distortion.RadialTangentialModel() if int(z["model"]) == 0 else distortion.FisheyeModel()
    offs = z["view_offsets"]
    dets = [(z["sensor_pts"][a:b].copy(), z["model_pts"][a:b].copy()) for a, b in zip(offs[:-1], offs[1:])]
    cal = calibrate.Calibrator(model)
    A0, W0, k0 = cal._decomposeParameterVector(z["P0"].copy())
    t0 = time.perf_counter()
    cal._initializeJacobian()
    t_setup = time.perf_counter() - t0
    out = io.StringIO()
    t0 = time.perf_counter()
    with contextlib.redirect_stdout(out):
        sse, A, W, k = cal.refineCalibrationParameters(A0, W0, k0, dets, int(z["maxIters"]), shouldPrint=True)
    t_call = time.perf_counter() - t0
    iters = sum(1 for ln in out.getvalue().splitlines() if ln.startswith("iter "))
    P = cal._composeParameterVector(A, W, k).ravel()
    print(json.dumps({
        "case": case, "points": int(offs[-1]), "views": int(len(offs) - 1), "model": int(z["model"]),
        "jacobian_setup_s": t_setup, "refine_call_s": t_call, "iterations": iters,
        "iterations_per_s": iters / t_call, "sse": float(sse), "P": P.tolist(),
        "ref_root": os.path.basename(os.path.normpath(ref_harness.REF_ROOT)),
    }))


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else "cfg1")
