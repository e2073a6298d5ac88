"""TEST INFRASTRUCTURE ONLY -- captures what the UNMODIFIED reference prints when
Calibrator.refineCalibrationParameters runs with shouldPrint=True (src/calibrate.py:158-159,269-274:
per iteration `iter N: (T s), error=E, lambda=L`, then `A:` and `k:` of the iterate BEFORE the update)
on the cfg1 golden case, from the reference's own P0.  Output: tests/golden/print_cfg1.txt.

    python oracle/make_golden_print.py
"""
import contextlib
import io
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import ref_harness  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")


def main():
    src = ref_harness.load()
    from src import calibrate, distortion
    g = np.load(os.path.join(OUT, "cfg1.npz"))
    offs = g["view_offsets"]
    dets = [(g["sensor_pts"][offs[i]:offs[i + 1]], g["model_pts"][offs[i]:offs[i + 1]]) for i in range(len(offs) - 1)]
    cal = calibrate.Calibrator(distortion.RadialTangentialModel())
    A0, W0, k0 = cal._decomposeParameterVector(g["P0"].reshape(-1, 1))
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf):
        sse, A, W, k = cal.refineCalibrationParameters(A0, W0, k0, dets, 100, shouldPrint=True)
    text = buf.getvalue()
    with open(os.path.join(OUT, "print_cfg1.txt"), "w") as f:
        f.write(text)
    print("wrote print_cfg1.txt:", len(text.splitlines()), "lines; sse", sse)


if __name__ == "__main__":
    main()
