"""TEST INFRASTRUCTURE ONLY -- writes tests/golden/detections_cfg1.json with the UNMODIFIED reference's
own Dataset.exportDetections (/root/reference/src/dataset.py:97-109) for the cfg1 dataset (9x6 board, 10
views, radial-tangential, no noise: the detections of tests/golden/cfg1.npz), as the fixture that pins the
detections wire format {"views": [{"sensorPoints": [[u, v], ...], "modelPoints": [[X, Y, Z], ...]}, ...]}.
usage: python oracle/make_golden_json.py"""
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import ref_harness  # noqa: E402
from make_golden import A_SYN, K_RADTAN  # noqa: E402

src = ref_harness.load()
from src import checkerboard, dataset, distortion, virtualcamera  # noqa: E402

cam = virtualcamera.VirtualCamera(A_SYN, K_RADTAN, distortion.RadialTangentialModel(), 1440, 1080, None)
ds = dataset.Dataset(checkerboard.Checkerboard(9, 6, 0.050), cam, 10)
out = os.path.join(HERE, "..", "tests", "golden", "detections_cfg1.json")
ds.exportDetections(out)
back = dataset.createDetectionsFromPath(out)
print("views", len(back), "bytes", os.path.getsize(out))
