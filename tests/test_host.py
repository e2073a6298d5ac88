"""CPU: host-side mirror (parameter layout, angle convention KATs), synthetic inputs bit-identical
to the reference's, and the C-ABI library exporting every symbol of include/calib_b200.h."""
import ctypes
import os
import subprocess

import numpy as np
import pytest

import camera_calibration_b200 as cc
from camera_calibration_b200 import _lib, distortion, synth
from camera_calibration_b200 import mathutils as mu


def test_library_builds_and_exports_every_declared_symbol():
    so = _lib.build()
    assert os.path.exists(so)
    names = _lib.exported_symbols()
    assert "calib_lm_run" in names and "calib_create" in names and len(names) >= 25
    L = ctypes.CDLL(so)
    for n in names:
        assert hasattr(L, n), n
    out = subprocess.run(["nm", "-D", "--defined-only", so], capture_output=True, text=True).stdout
    for n in names:
        assert f" T {n}" in out, n
    # every declared symbol has a ctypes signature in the binding
    assert set(names) == set(_lib.lib()._signatures)


def test_built_kernels_contain_the_instructions_design_md_names():
    """What DESIGN.md says about the kernels, read off the SASS of the library that ships (no GPU needed):
    built for sm_100a; k_eval hands values from its producer warps to its consumer warps through mbarriers
    (SYNCS), spills nothing and has no call left in it -- the reciprocal and the reciprocal square root of
    point_eval are branch-free, which is what lets ptxas schedule a warp's independent sums into the stalls of its
    chain; the bulk-copy variants of k_cost use the 1-D TMA (UBLKCP)."""
    so = _lib.build()
    cuobjdump = "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(cuobjdump):
        pytest.skip("no cuobjdump")
    elf = subprocess.run([cuobjdump, "-lelf", so], capture_output=True, text=True).stdout
    assert "sm_100a" in elf, elf
    sass = subprocess.run([cuobjdump, "-sass", so], capture_output=True, text=True).stdout
    funcs = {}
    name = None
    for line in sass.splitlines():
        if "Function :" in line:
            name = line.split("Function :")[1].strip()
            funcs[name] = []
        elif name is not None:
            funcs[name].append(line)
    evals = {n: "\n".join(b) for n, b in funcs.items() if "6k_evalILi" in n}
    assert len(evals) == 2, list(evals)
    for n, body in evals.items():
        assert "SYNCS.ARRIVE" in body and "SYNCS.PHASECHK" in body
        assert "STL" not in body and "LDL" not in body, n            # no spill
        assert "CALL" not in body, n                                 # no slow path of a reciprocal / rsqrt
        assert "MUFU.RCP64H" in body
        assert body.count("DFMA") > 200                              # ~250 DFMA + ~100 DMUL / DADD per point
    assert any("MUFU.RSQ64H" in b for b in evals.values())           # fisheye
    costs = {n: "\n".join(b) for n, b in funcs.items() if "6k_costILi" in n}
    assert any("UBLKCP" in b for b in costs.values()) and len(costs) == 8


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    c = synth.config("cfg1")
    with pytest.raises(cc.CalibError):
        cc.Problem.from_detections(0, c["dets"])


def test_euler_known_answers():
    # tests/test_mathutils.py:112-152 of the reference
    assert np.allclose(mu.eulerToRotationMatrix((0, 0, 90)), [[0, -1, 0], [1, 0, 0], [0, 0, 1]], atol=1e-12)
    assert np.allclose(mu.eulerToRotationMatrix((0, 180, 0)), [[-1, 0, 0], [0, 1, 0], [0, 0, -1]], atol=1e-12)
    assert np.allclose(mu.eulerToRotationMatrix((0, 90, 90)), [[0, -1, 0], [0, 0, 1], [-1, 0, 0]], atol=1e-12)
    rs = np.random.RandomState(0)
    for _ in range(20):
        e = rs.uniform(-89, 89, 3) * np.array([2, 1, 2])
        R = mu.eulerToRotationMatrix(e)
        assert np.allclose(R @ R.T, np.eye(3), atol=1e-12)
        assert np.allclose(mu.eulerToRotationMatrix(mu.rotationMatrixToEuler(R)), R, atol=1e-9)
    # gimbal cases, tests/test_mathutils.py:275-319
    for e in ((10, 90, 0), (10, -90, 0)):
        R = mu.eulerToRotationMatrix(e)
        assert np.allclose(mu.eulerToRotationMatrix(mu.rotationMatrixToEuler(R)), R, atol=1e-9)


@pytest.mark.parametrize("model,k", [(0, (-0.25, 0.2, 0.07, -0.03, 0.05)), (1, (-0.155, -0.02, 0.0, -0.03))])
def test_parameter_vector_round_trip(model, k):
    # tests/test_calibrate.py:63-78 of the reference
    cal = cc.Calibrator(model)
    A = np.array([[400, 0.3, 320], [0, 410, 240], [0, 0, 1.0]])
    W = [mu.poseFromRT(mu.eulerToRotationMatrix(e), t) for e, t in
         (((170, 5, -30), (0.1, 0.2, 1.0)), ((-160, -12, 100), (-0.2, 0.1, 0.7)))]
    P = cal._composeParameterVector(A, W, k)
    assert P.shape == (6 * 2 + 5 + len(k), 1)
    A2, W2, k2 = cal._decomposeParameterVector(P)
    assert np.allclose(A2, A, atol=1e-9) and np.allclose(k2, k, atol=1e-9)
    for a, b in zip(W, W2):
        assert np.allclose(a, b, atol=1e-9)
    with pytest.raises(ValueError):
        cal._composeParameterVector(A, W, k[:-1])


def test_compose_matches_reference_layout(golden):
    g = golden("cfg1")
    cal = cc.Calibrator(distortion.RadialTangentialModel())
    A, W, k = cal._decomposeParameterVector(g["P0"])
    P = cal._composeParameterVector(A, W, k).ravel()
    d = np.abs(P - g["P0"])
    d = np.minimum(d, np.abs(d - 360))
    assert d.max() < 1e-9


def test_model_selection():
    assert distortion.model_id_of("radtan") == 0 and distortion.model_id_of("fisheye") == 1
    assert distortion.model_id_of(distortion.FisheyeModel()) == 1
    with pytest.raises(ValueError):   # src/main.py:33
        cc.calibrateCamera([], "pinhole", 10)

    class RefLike:   # the reference's own model objects are duck-typed
        def getDistortionSymbols(self):
            return ("k1", "k2", "p1", "p2", "k3")
    assert distortion.model_id_of(RefLike()) == 0


@pytest.mark.parametrize("case", ["cfg1", "cfg1n", "cfg2", "itest"])
def test_synthetic_inputs_bit_identical_to_reference(golden, case):
    g = golden(case)
    c = synth.config(case)
    offs, mp, sp = cc.flatten_detections(c["dets"])
    assert np.array_equal(offs, g["view_offsets"])
    assert np.array_equal(mp, g["model_pts"])
    assert np.array_equal(sp, g["sensor_pts"])
    assert np.array_equal(np.array(c["poses"]), g["W_true"])
    assert np.array_equal(cc.getSensorPoints(c["dets"]), g["sensor_pts"])


def test_large_config_shapes():
    c = synth.config("cfg3", scale=0.01)
    assert len(c["dets"]) == 10
    n = [len(s) for s, m in c["dets"]]
    assert max(n) <= 1600 and min(n) > 100     # ragged views, most corners in view
    A0, W0, k0 = c["init"]
    assert len(W0) == 10 and np.all(k0 == 0)
    b = synth.batch_config(2)
    assert len(b) == 2 and len(b[0]["dets"]) == 20


def test_detections_json_wire_format(tmp_path):
    """The detections JSON of the reference (src/dataset.py:97-109,133-141): the fixture was written by
    the reference's own Dataset.exportDetections (oracle/make_golden_json.py); our reader returns the
    cfg1 detections bit for bit and our writer reproduces the file byte for byte."""
    import os
    import numpy as np
    from camera_calibration_b200 import dataset
    here = os.path.dirname(__file__)
    fixture = os.path.join(here, "golden", "detections_cfg1.json")
    dets = dataset.createDetectionsFromPath(fixture)
    g = np.load(os.path.join(here, "golden", "cfg1.npz"))
    offs = g["view_offsets"]
    assert len(dets) == len(offs) - 1
    for v, (s, m) in enumerate(dets):
        assert s.shape[1] == 2 and m.shape[1] == 3
        assert np.array_equal(s, g["sensor_pts"][offs[v]:offs[v + 1]])
        assert np.array_equal(m, g["model_pts"][offs[v]:offs[v + 1]])
    out = tmp_path / "d.json"
    dataset.exportDetections(dets, str(out))
    assert open(out, "rb").read() == open(fixture, "rb").read()


def test_flatten_detections_threaded_gather():
    """flatten_detections on a large list goes through calib_host_gather (threads in the library, no
    device involved) and must equal numpy's concatenate; odd inputs (float32, non-contiguous, lists)
    take the numpy path and give the same CSR arrays."""
    import numpy as np
    import camera_calibration_b200 as cc
    rs = np.random.RandomState(3)
    dets = []
    for v in range(300):
        n = int(rs.randint(1, 900))
        dets.append((rs.rand(n, 2), rs.rand(n, 3)))
    dets[7] = (dets[7][0][:0], dets[7][1][:0])                      # an empty view
    offs, m, s = cc.flatten_detections(dets)
    assert offs[-1] >= 100000
    assert np.array_equal(m, np.concatenate([d[1] for d in dets])) and np.array_equal(s, np.concatenate([d[0] for d in dets]))
    assert np.array_equal(np.diff(offs), [len(d[1]) for d in dets])
    odd = list(dets)
    odd[3] = (np.asfortranarray(dets[3][0]), dets[3][1][::1].astype(np.float32))
    odd[5] = (dets[5][0].tolist(), dets[5][1].tolist())
    o2, m2, s2 = cc.flatten_detections(odd)
    assert np.array_equal(o2, offs) and np.array_equal(s2, s)
    assert np.allclose(m2, m, rtol=1e-6)


def test_view_pointers_of_detection_lists():
    """calibrate._view_pointers: addresses and counts of the per-view arrays handed to
    calib_create_from_views; anything that is not a C-contiguous float64 (n,2)/(n,3) pair -> None
    (the numpy concatenate path takes over)."""
    import ctypes
    import numpy as np
    from camera_calibration_b200 import calibrate
    rs = np.random.RandomState(5)
    dets = [(rs.rand(n, 2), rs.rand(n, 3)) for n in (4, 0, 7, 1)]
    dets[2][1].flags.writeable = False                                # read-only: the slower pointer route
    counts, pm, ps = calibrate._view_pointers(dets)
    assert counts.tolist() == [4, 0, 7, 1] and counts.dtype == np.int64
    for i, (s, m) in enumerate(dets):
        if len(m):
            assert pm[i] == m.ctypes.data and ps[i] == s.ctypes.data
            assert np.array_equal(np.ctypeslib.as_array(ctypes.cast(int(pm[i]), ctypes.POINTER(ctypes.c_double)), (len(m) * 3,)), m.ravel())
    assert calibrate._view_pointers([]) is None
    assert calibrate._view_pointers([(dets[0][0].tolist(), dets[0][1])]) is None
    assert calibrate._view_pointers([(dets[0][0], dets[0][1].astype(np.float32))]) is None
    assert calibrate._view_pointers([(dets[0][0], np.asfortranarray(dets[0][1]))]) is None
    assert calibrate._view_pointers([(dets[0][0][:3], dets[0][1])]) is None
    wide = rs.rand(4, 6)
    assert calibrate._view_pointers([(dets[0][0], wide[:, :3])]) is None


def test_view_range_refuses_empty_shards_on_every_rank():
    """more ranks than views (or all points in few views): every rank must raise the same error before any
    handle is created -- a rank that alone failed in calib_create would leave its peers in a collective"""
    from camera_calibration_b200 import sharding
    dets = [(np.zeros((n, 2)), np.zeros((n, 3))) for n in (6, 6, 6)]
    for rank in range(3):
        lo, hi = sharding.view_range(dets, 3, rank)
        assert hi == lo + 1
    for rank in range(4):
        with pytest.raises(ValueError, match="without views"):
            sharding.view_range(dets, 4, rank)
    lopsided = [(np.zeros((n, 2)), np.zeros((n, 3))) for n in (1000, 1, 1)]
    for rank in range(3):
        with pytest.raises(ValueError, match="without views"):
            sharding.view_range(lopsided, 3, rank)
