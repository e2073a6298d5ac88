"""GPU, needs >= 2 devices (skipped on a one-GPU box): the view-sharded path over NCCL against
the single-GPU path, launched with torchrun on 127.0.0.1."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ngpu():
    import torch
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


@pytest.mark.parametrize("case,scale,loop_mode,transport", [
    ("cfg3", 0.04, 1, "nccl"), ("cfg5", 0.003, 1, "nccl"),
    ("cfg3", 0.04, 0, "peer"), ("cfg5", 0.003, 0, "peer"), ("cfg3", 0.04, 1, "peer")])
def test_sharded_equals_single(case, scale, loop_mode, transport):
    n = _ngpu()
    if n < 2:
        pytest.skip("needs 2 GPUs")
    world = 2 if n < 4 else 4
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", "29541",
           os.path.join(ROOT, "tests", "mp", "sharded_worker.py"), case, str(scale), str(loop_mode), transport]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    lines = [ln for ln in res.stdout.splitlines() if ln.startswith("{")]
    assert res.returncode == 0 and lines, res.stdout[-2000:] + res.stderr[-6000:]
    out = json.loads(lines[-1])
    assert out["ok"], out


@pytest.mark.parametrize("mode", ["soak", "timeout"])
def test_peer_exchange_soak_and_timeout(mode):
    """tests/mp/soak_worker.py: the flag-in-data exchange under deliberate rank skew (200 runs, bit-identical),
    and a rank that never arrives (CALIB_E_COMM after the exchange's timeout, no hang)."""
    n = _ngpu()
    if n < 2:
        pytest.skip("needs 2 GPUs")
    world = 2 if n < 4 else 4
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", "29543",
           os.path.join(ROOT, "tests", "mp", "soak_worker.py"), mode]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=400)
    lines = [ln for ln in res.stdout.splitlines() if ln.startswith("{")]
    assert res.returncode == 0 and lines, res.stdout[-2000:] + res.stderr[-6000:]
    assert json.loads(lines[-1])["ok"], lines[-1]


def test_single_process_multi_device_handle():
    """calib_multi_*: ONE Python process, no torch, several GPUs behind the reference's method.  Results must
    repeat run to run, agree with the one-GPU path up to the cost plateau, and come out the same through
    Calibrator(devices=[...]).refineCalibrationParameters and through the per-view-array entry point."""
    n = _ngpu()
    if n < 2:
        pytest.skip("needs 2 GPUs")
    import numpy as np
    sys.path.insert(0, ROOT)
    import camera_calibration_b200 as cc
    from camera_calibration_b200 import synth
    devices = list(range(min(n, 4)))
    for name, scale in (("cfg3", 0.04), ("cfg5", 0.003)):
        c = synth.config(name, scale)
        model = c["model"]
        cal = cc.Calibrator(model, devices=devices)
        # start from a vector that survives the decompose / compose round trip of the drop-in bit for bit
        P0 = cal._composeParameterVector(*c["init"]).ravel()
        for _ in range(8):
            A0, W0, k0 = cal._decomposeParameterVector(P0)
            P1 = cal._composeParameterVector(A0, W0, k0).ravel()
            if np.array_equal(P1, P0):
                break
            P0 = P1
        assert np.array_equal(cal._composeParameterVector(A0, W0, k0).ravel(), P0)
        offs, mp, sp = cc.flatten_detections(c["dets"])
        mprob = cc.MultiProblem(model, offs, mp, sp, devices)
        assert mprob.n_gpus == len(devices)
        b = mprob.view_bounds()
        assert b[0] == 0 and b[-1] == len(c["dets"]) and np.all(np.diff(b) > 0)
        ref = None
        for mode in (0, 2):
            mprob.set_loop_mode(mode)
            for rep in range(2):
                got = mprob.lm_run(P0, 100)
                if ref is None:
                    ref = got
                assert got[0] == ref[0] and got[2] == ref[2], (name, mode)
                assert np.array_equal(got[1], ref[1]) and np.array_equal(got[3], ref[3]), (name, mode)
        mprob.close()
        s, P, it, tr = ref
        one = cc.Problem(model, offs, mp, sp, device=devices[0])
        s1, P1, i1, t1 = one.lm_run(P0, 100)
        one.close()
        m = min(len(tr), len(t1))
        plateau = np.where(np.abs(t1[:, 1] - t1[:, 0]) <= 1e-10 * t1[:, 0])[0]
        stop = min(m, plateau[0]) if len(plateau) else m
        assert np.array_equal(tr[:stop, 3], t1[:stop, 3]) and np.array_equal(tr[:stop, 2], t1[:stop, 2])
        assert np.allclose(tr[:stop, 0], t1[:stop, 0], rtol=1e-9)
        assert (np.abs(P - P1) / np.maximum(np.abs(P1), 1e-3)).max() < 1e-6 and abs(s - s1) <= 1e-9 * s1
        # the drop-in: same call as the reference's, devices given once at construction
        sse, A, W, k = cal.refineCalibrationParameters(A0, W0, k0, c["dets"], 100)
        Pc = cal._composeParameterVector(A, W, k).ravel()
        assert sse == s and cal.lastIters == it and np.array_equal(cal.lastTrace, tr)
        d = np.abs(Pc - P)
        d = np.minimum(d, np.abs(d - 360))
        assert (d / np.maximum(np.abs(P), 1e-3)).max() < 1e-9      # Euler angles go through R and back
    # more devices than views: the library uses as many as there are views
    c = synth.config("cfg3", 0.002)
    offs, mp, sp = cc.flatten_detections(c["dets"])
    if len(c["dets"]) < len(devices):
        mprob = cc.MultiProblem(0, offs, mp, sp, devices)
        assert mprob.n_gpus == len(c["dets"])
        mprob.close()
    with pytest.raises(cc.CalibError):
        cc.MultiProblem(0, offs, mp, sp, [devices[0], devices[0]])
