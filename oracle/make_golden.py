"""TEST INFRASTRUCTURE ONLY -- generates tests/golden/*.npz by running the UNMODIFIED
reference (through oracle/ref_harness.py) in this container.  The reference's own tests
pin no Jacobian values, traces or iteration counts, so these
fixtures are what pins the oracle and, through it, the CUDA path.

    python oracle/make_golden.py [case ...]      # cases: kat cfg1 cfg1n cfg2 itest

Every case stores the inputs in the CSR layout the C-ABI takes (view_offsets, model_pts,
sensor_pts), the initial parameter vector P0 the reference's own linear stage produced
(Calibrator.estimateCalibrationParameters, calibrate.py:41-58), the reference's
residual / Jacobian blocks / first step at P0, the per-iteration trace of the loop body
of Calibrator.refineCalibrationParameters (calibrate.py:143-168, re-run line by line so
E, E', lambda and the decision can be recorded), and its outputs.
"""
import contextlib
import io
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import ref_harness  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")

A_SYN = np.array([[803.1, 0, 700.5], [0, 803.1, 529.2], [0, 0, 1]], dtype=np.float64)
K_RADTAN = (-0.25, 0.2, 0.07, -0.03, 0.05)
K_FISHEYE = (-0.155, -0.02, 0.0, -0.03)


def _flatten(allDetections):
    counts = [len(s) for s, m in allDetections]
    offs = np.zeros(len(counts) + 1, dtype=np.int64)
    offs[1:] = np.cumsum(counts)
    sensor = np.vstack([s for s, m in allDetections]).astype(np.float64)
    model = np.vstack([m for s, m in allDetections]).astype(np.float64)
    return offs, model, sensor


def _blocks(src, jac, P, allModelPoints, L):
    """Per-view compact blocks from the reference's ProjectionJacobian (jacobian.py:38-46)."""
    P = np.asarray(P).ravel()
    M = len(allModelPoints)
    intr = P[:L]
    Jc, Jv = [], []
    for i in range(M):
        ext = P[L + 6 * i: L + 6 * i + 6]
        Jc.append(jac._createIntrinsicsJacobianBlock(intr, ext, allModelPoints[i]))
        Jv.append(jac._createExtrinsicsJacobianBlock(intr, ext, allModelPoints[i]))
    return np.vstack(Jc), np.vstack(Jv)


def run_case(name, src, model_name, ds, maxIters, check_entry_point=True):
    from src import calibrate, distortion
    model = (distortion.RadialTangentialModel() if model_name == "radtan"
             else distortion.FisheyeModel())
    L = 5 + len(model.getDistortionSymbols())
    allDetections = ds.getCornerDetectionsInSensorCoordinates()
    offs, model_pts, sensor_pts = _flatten(allDetections)
    cal = calibrate.Calibrator(model)
    t0 = time.time()
    with contextlib.redirect_stdout(io.StringIO()):
        A0, W0, k0 = cal.estimateCalibrationParameters(allDetections)
    t_init = time.time() - t0
    t0 = time.time()
    cal._initializeJacobian()
    t_jacinit = time.time() - t0
    P0 = cal._composeParameterVector(A0, W0, k0).ravel().copy()
    allModelPoints = [m for s, m in allDetections]
    ydot = calibrate.getSensorPoints(allDetections)

    Jc0, Jv0 = _blocks(src, cal._jac, P0, allModelPoints, L)
    y0 = cal.projectAllPoints(P0, allModelPoints)
    r0 = (ydot.reshape(-1, 1) - y0.reshape(-1, 1)).ravel()

    # loop body of calibrate.py:143-168, verbatim calls, with a trace
    Pt = P0.reshape(-1, 1).copy()
    lam = cal._λinitial
    trace = []
    delta0 = None
    t0 = time.time()
    t_jac = 0.0
    for it in range(maxIters):
        tj = time.time()
        J = cal._jac.compute(Pt, allModelPoints)
        t_jac += time.time() - tj
        JTJ = J.T @ J
        diagJTJ = np.diag(np.diagonal(JTJ))
        y = cal.projectAllPoints(Pt, allModelPoints)
        r = ydot.reshape(-1, 1) - y.reshape(-1, 1)
        delta = np.linalg.inv(JTJ + lam * diagJTJ) @ J.T @ r
        if it == 0:
            delta0 = delta.ravel().copy()
            JTJ0 = JTJ.copy()
            JTr0 = (J.T @ r).ravel().copy()
        E = cal._computeReprojectionError(Pt, allDetections)
        E1 = cal._computeReprojectionError(Pt + delta, allDetections)
        acc = bool(E1 < E)
        trace.append((E, E1, lam, 1.0 if acc else 0.0))
        if acc:
            Pt += delta
            lam /= 10
        else:
            lam *= 10
        if not (cal._λmin < lam < cal._λmax) or E < cal._Pt_error_min:
            break
    t_loop = time.time() - t0
    iters = len(trace)
    sse = trace[-1][0]
    Pfinal = Pt.ravel().copy()
    out = dict(
        model=np.array(0 if model_name == "radtan" else 1),
        view_offsets=offs, model_pts=model_pts, sensor_pts=sensor_pts,
        P0=P0, r0=r0, Jc0=Jc0, Jv0=Jv0, y0=y0, delta0=delta0,
        JTJ0=JTJ0 if JTJ0.shape[0] <= 200 else np.zeros(0), JTr0=JTr0,
        trace=np.array(trace), iters=np.array(iters), sse=np.array(sse), Pfinal=Pfinal,
        A_true=np.asarray(ds.getIntrinsicMatrix(), dtype=np.float64),
        k_true=np.asarray(ds.getDistortionVector(), dtype=np.float64),
        W_true=np.array(ds.getAllBoardPosesInCamera()),
        maxIters=np.array(maxIters),
        ref_seconds=np.array([t_init, t_jacinit, t_loop, t_jac]),
    )
    if check_entry_point:
        # the untouched entry point must give what the traced loop gave
        cal2 = calibrate.Calibrator(model)
        cal2._jac = cal._jac
        with contextlib.redirect_stdout(io.StringIO()):
            sse2, A2, W2, k2 = cal2.refineCalibrationParameters(
                A0, W0, k0, allDetections, maxIters, shouldPrint=False)
        assert sse2 == sse, (sse2, sse)
        A3, W3, k3 = cal._decomposeParameterVector(Pfinal)
        assert np.array_equal(A2, A3) and np.array_equal(k2, k3)
        assert all(np.array_equal(a, b) for a, b in zip(W2, W3))
        out["entry_point_checked"] = np.array(1)
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **out)
    print(f"{name}: M={len(offs)-1} N={offs[-1]} K={len(P0)} iters={iters} sse={sse:.6g} "
          f"init={t_init:.1f}s jacinit={t_jacinit:.1f}s loop={t_loop:.1f}s (jac {t_jac:.1f}s)",
          flush=True)


def make_kat(src):
    """Known-answer vectors for the per-point maps: the reference test's own vector
    (tests/test_jacobian.py:19-24) and a general one (Z != 0, skew != 0, negative angles)."""
    from src import calibrate, distortion, jacobian
    out = {}
    rs = np.random.RandomState(7)
    for mname, model, kk in (("radtan", distortion.RadialTangentialModel(), [-0.5, 0.2, 0, 0, 0]),
                             ("fisheye", distortion.FisheyeModel(), [-0.1, 0.02, 0.01, -0.005])):
        jac = jacobian.ProjectionJacobian(model)
        cal = calibrate.Calibrator(model)
        L = 5 + len(kk)
        # (a) reference test vector
        intr = [400, 400, 0, 320, 240] + kk
        ext = [180, 0, 0, 0.1, 0.2, 1.0]
        pts = np.array([[0.1, 0.1, 0], [0.1, 0.2, 0]], dtype=np.float64)
        out[f"{mname}_a_P"] = np.array(intr + ext, dtype=np.float64)
        out[f"{mname}_a_pts"] = pts
        out[f"{mname}_a_Jc"] = jac._createIntrinsicsJacobianBlock(intr, ext, pts)
        out[f"{mname}_a_Jv"] = jac._createExtrinsicsJacobianBlock(intr, ext, pts)
        out[f"{mname}_a_y"] = cal.projectAllPoints(np.array(intr + ext, dtype=np.float64), [pts])
        # (b) general: 3 views, 40 points each, Z != 0, gamma != 0
        if mname == "radtan":
            intr = [812.3, 795.1, 1.7, 651.2, 489.9, -0.21, 0.11, 0.003, -0.002, 0.03]
        else:
            intr = [812.3, 795.1, 1.7, 651.2, 489.9, -0.12, 0.015, 0.004, -0.002]
        exts = [[-158.0, 12.0, -33.0, -0.21, 0.13, 0.92],
                [171.0, -24.0, 77.0, 0.05, -0.18, 0.71],
                [-12.0, 9.0, -110.0, 0.11, 0.07, 0.65]]
        allpts = []
        for e in exts:
            p = rs.uniform(-0.3, 0.3, (40, 3))
            p[:, 2] *= 0.2
            if e[0] < 0 and abs(e[0]) < 90:
                pass
            allpts.append(p)
        P = np.array(intr + sum(exts, []), dtype=np.float64)
        Jc = np.vstack([jac._createIntrinsicsJacobianBlock(intr, e, p) for e, p in zip(exts, allpts)])
        Jv = np.vstack([jac._createExtrinsicsJacobianBlock(intr, e, p) for e, p in zip(exts, allpts)])
        out[f"{mname}_b_P"] = P
        out[f"{mname}_b_pts"] = np.vstack(allpts)
        out[f"{mname}_b_offs"] = np.array([0, 40, 80, 120], dtype=np.int64)
        out[f"{mname}_b_Jc"] = Jc
        out[f"{mname}_b_Jv"] = Jv
        out[f"{mname}_b_y"] = cal.projectAllPoints(P, allpts)
        J = jac.compute(P, allpts)
        assert J.shape == (240, L + 18)
        assert np.array_equal(J[:, :L], Jc) and np.array_equal(J[80:160, L + 6:L + 12], Jv[80:160])
        assert J[0, L + 6] == 0
    np.savez_compressed(os.path.join(OUT, "kat.npz"), **out)
    print("kat: done", flush=True)


def main(argv):
    cases = argv or ["kat", "cfg1", "cfg1n", "cfg2", "itest"]
    os.makedirs(OUT, exist_ok=True)
    src = ref_harness.load()
    from src import checkerboard, dataset, distortion, noise, virtualcamera
    for c in cases:
        if c == "kat":
            make_kat(src)
        elif c in ("cfg1", "cfg1n"):
            cam = virtualcamera.VirtualCamera(A_SYN, K_RADTAN, distortion.RadialTangentialModel(),
                                              1440, 1080, noise.NoiseModel(0.1) if c == "cfg1n" else None)
            ds = dataset.Dataset(checkerboard.Checkerboard(9, 6, 0.050), cam, 10)
            run_case(c, src, "radtan", ds, 100)
        elif c == "cfg2":
            cam = virtualcamera.VirtualCamera(A_SYN, K_FISHEYE, distortion.FisheyeModel(),
                                              1440, 1080, noise.NoiseModel(0.1))
            ds = dataset.Dataset(checkerboard.Checkerboard(9, 6, 0.050), cam, 25)
            run_case(c, src, "fisheye", ds, 100)
        elif c == "itest":
            ds = dataset.createRealisticRadTanDataset()
            run_case(c, src, "radtan", ds, 100, check_entry_point=False)
        else:
            raise SystemExit("unknown case " + c)


if __name__ == "__main__":
    main(sys.argv[1:])
