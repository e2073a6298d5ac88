"""GPU (-m gpu): the CUDA path, called through the C-ABI, against the golden vectors generated
from the unmodified reference and against the CPU oracle on the same seeded inputs.
Tolerances are north_star's: residuals / Jacobian entries 1e-10 relative, converged parameters
1e-8 relative, same iteration count (exact where the reference's own count is well defined)."""
import os

import numpy as np
import pytest

import calib_oracle as onp
import oracle_c as oc
import camera_calibration_b200 as cc
from camera_calibration_b200 import synth
from conftest import jac_relerr, param_relerr

pytestmark = pytest.mark.gpu

TOL_J = 1e-10
TOL_P = 1e-8
CASES = ["cfg1", "cfg1n", "cfg2", "itest"]


def _problem(g):
    return cc.Problem(int(g["model"]), g["view_offsets"], g["model_pts"], g["sensor_pts"])


@pytest.mark.parametrize("model,name", [(0, "radtan"), (1, "fisheye")])
def test_kat_vectors(golden, model, name):
    g = golden("kat")
    for tag in ("a", "b"):
        P, pts = g[f"{name}_{tag}_P"], g[f"{name}_{tag}_pts"]
        offs = np.array([0, len(pts)]) if tag == "a" else g[f"{name}_b_offs"]
        prob = cc.Problem(model, offs, pts, np.zeros((len(pts), 2)))
        r, Jc, Jv = prob.residual_jacobian(P)
        assert not np.isnan(Jc).any() and not np.isnan(Jv).any()
        assert jac_relerr(Jc, g[f"{name}_{tag}_Jc"]) < TOL_J
        assert jac_relerr(Jv, g[f"{name}_{tag}_Jv"]) < TOL_J
        y = prob.project(P)
        assert np.allclose(y, g[f"{name}_{tag}_y"], rtol=1e-12, atol=1e-9)
        assert np.allclose(r, -y.reshape(-1), rtol=0, atol=1e-9)
        prob.close()


@pytest.mark.parametrize("case", CASES)
def test_residual_jacobian_vs_reference(golden, case):
    g = golden(case)
    prob = _problem(g)
    r, Jc, Jv = prob.residual_jacobian(g["P0"])
    assert jac_relerr(Jc, g["Jc0"]) < TOL_J
    assert jac_relerr(Jv, g["Jv0"]) < TOL_J
    assert np.abs(r - g["r0"]).max() <= 1e-10 * max(1.0, np.abs(g["r0"]).max())
    y = prob.project(g["P0"])
    assert np.abs(y - g["y0"]).max() <= 1e-10 * np.abs(g["y0"]).max()
    E = prob.cost(g["P0"])
    assert abs(E - g["trace"][0, 0]) <= 1e-12 * g["trace"][0, 0]
    prob.close()


def test_dense_jacobian_layout(golden):
    # tests/test_jacobian.py:56-70 of the reference: shape and block sparsity
    g = golden("cfg1")
    offs = g["view_offsets"]
    allModelPoints = [g["model_pts"][offs[i]:offs[i + 1]] for i in range(len(offs) - 1)]
    J = cc.ProjectionJacobian("radtan").compute(g["P0"], allModelPoints)
    MN = int(offs[-1])
    assert J.shape == (2 * MN, 6 * 10 + 10)
    assert J[0, 0] != 0 and J[-1, -1] != 0 and J[0, 16] == 0
    Jo = onp.dense_jacobian(0, g["P0"], offs, g["model_pts"])
    assert jac_relerr(J, Jo) < TOL_J


@pytest.mark.parametrize("case", ["cfg1", "cfg1n", "cfg2"])
def test_normal_equations_vs_reference(golden, case):
    g = golden(case)
    prob = _problem(g)
    L = prob.L
    A, B, D, gg, E = prob.normal_equations(g["P0"])
    JTJ, JTr = g["JTJ0"], g["JTr0"]
    sc = np.abs(JTJ).max()
    assert np.abs(A - JTJ[:L, :L]).max() < 1e-12 * sc
    assert np.allclose(A, A.T, rtol=0, atol=0)
    for v in range(prob.M):
        c = L + 6 * v
        assert np.abs(B[v] - JTJ[:L, c:c + 6]).max() < 1e-12 * sc
        assert np.abs(D[v] - JTJ[c:c + 6, c:c + 6]).max() < 1e-12 * sc
    assert np.abs(gg - JTr).max() < 1e-12 * np.abs(JTr).max()
    assert abs(E - g["trace"][0, 0]) <= 1e-12 * g["trace"][0, 0]
    # the damped step: against the oracle's Schur solve tightly, against the reference's
    # explicit inverse loosely (cond ~5e9)
    d = prob.lm_step(g["P0"], 1e-3)
    Ao, Bo, Do, go, _ = oc.normal_equations(int(g["model"]), g["P0"], g["view_offsets"], g["model_pts"], g["sensor_pts"])
    do = oc.lm_step(int(g["model"]), Ao, Bo, Do, go, 1e-3)
    assert np.abs(d - do).max() < 1e-8 * np.abs(do).max()
    assert np.abs(d - g["delta0"]).max() < 1e-6 * np.abs(g["delta0"]).max()
    prob.close()


def _check_trace(tr, gt):
    n = min(len(tr), len(gt))
    plateau = np.where(np.abs(gt[:, 1] - gt[:, 0]) <= 1e-10 * gt[:, 0])[0]
    stop = min(n, plateau[0]) if len(plateau) else n
    assert np.array_equal(tr[:stop, 3], gt[:stop, 3])
    assert np.array_equal(tr[:stop, 2], gt[:stop, 2])   # bit-identical lambda recurrence
    big = gt[:stop, 0] > 1e-9
    assert np.allclose(tr[:stop, 0][big], gt[:stop, 0][big], rtol=1e-6)
    return stop


@pytest.mark.parametrize("case", CASES)
def test_lm_run_vs_reference(golden, case):
    g = golden(case)
    prob = _problem(g)
    sse, P, iters, tr = prob.lm_run(g["P0"], int(g["maxIters"]))
    gt = g["trace"]
    stop = _check_trace(tr, gt)
    if case in ("cfg1", "itest"):      # noise-free: the count is well defined -> exact
        assert param_relerr(P, g["Pfinal"]) < TOL_P
        assert iters == int(g["iters"])
        assert sse < 1e-12
    else:
        # noisy: past the plateau (|E'-E| <= 1e-10 E) accept/reject is rounding noise in the
        # reference too.  Same iterates up to there (against the oracle,
        # itself pinned on this golden trace), same minimum after it.
        _, Ps, _, _ = prob.lm_run(g["P0"], stop)
        _, Pso, _, _ = oc.lm(int(g["model"]), g["P0"], g["view_offsets"], g["model_pts"], g["sensor_pts"], stop)
        assert param_relerr(Ps, Pso) < TOL_P
        assert param_relerr(P, g["Pfinal"]) < 1e-6
        assert abs(sse - float(g["sse"])) < 1e-9 * float(g["sse"])
    # internal consistency the reference has by construction: the cost after an accepted step
    # is the cost the next iteration starts from (same function, same rounding)
    for i in range(iters - 1):
        if tr[i, 3] == 1.0:
            assert tr[i + 1, 0] == tr[i, 1]
        else:
            assert tr[i + 1, 0] == tr[i, 0]
    assert sse == tr[-1, 0]             # stale sse, src/calibrate.py:155,171
    prob.close()


@pytest.mark.parametrize("case", ["cfg1", "cfg2"])
def test_lm_run_vs_oracle_c(golden, case):
    g = golden(case)
    prob = _problem(g)
    sse, P, iters, tr = prob.lm_run(g["P0"], 100)
    so, Po, io, to = oc.lm(int(g["model"]), g["P0"], g["view_offsets"], g["model_pts"], g["sensor_pts"], 100)
    _check_trace(tr, to)
    assert param_relerr(P, Po) < TOL_P
    prob.close()


def test_calibrator_drop_in(golden, capsys):
    """The reference's own integration assertions (tests/itest_main.py:12-29) on the realistic
    radtan dataset, through the mirrored entry point."""
    g = golden("itest")
    c = synth.config("itest")
    cal = cc.Calibrator("radtan")
    A0, W0, k0 = cal._decomposeParameterVector(g["P0"])
    sse, A, W, k = cal.refineCalibrationParameters(A0, W0, k0, c["dets"], 100, shouldPrint=True)
    assert "iter 0:" in capsys.readouterr().out
    assert abs(sse) < 1e-7
    assert np.allclose(A, c["A"], atol=1e-9)
    assert np.allclose(k, c["k"], atol=1e-9)
    for we, wc in zip(c["poses"], W):
        assert np.allclose(we, wc, atol=1)
    assert cal.lastIters == int(g["iters"])
    with pytest.raises(UnboundLocalError):
        cal.refineCalibrationParameters(A0, W0, k0, c["dets"], 0)
    # zero residual at the truth, tests/test_calibrate.py:123-133
    Ptruth = cal._composeParameterVector(c["A"], c["poses"], c["k"])
    assert cal._computeReprojectionError(Ptruth, c["dets"]) < 1e-12
    y = cal.projectAllPoints(Ptruth, [m for s, m in c["dets"]])
    assert y.shape == cc.getSensorPoints(c["dets"]).shape


def test_should_print_matches_the_reference(golden, capsys):
    """shouldPrint=True: what the unmodified reference printed on cfg1 (tests/golden/print_cfg1.txt, captured by
    oracle/make_golden_print.py; src/calibrate.py:158-159, 269-274) against what the drop-in prints: the
    same lines in the same order -- `iter N: (T s), error=E, lambda=L`, `A:`, the 3 x 3 matrix, `k:`, the
    distortion vector of the iterate BEFORE the update -- with every number but the seconds equal to 1e-6."""
    import re
    g = golden("cfg1")
    offs = g["view_offsets"]
    dets = [(g["sensor_pts"][offs[i]:offs[i + 1]], g["model_pts"][offs[i]:offs[i + 1]]) for i in range(len(offs) - 1)]
    cal = cc.Calibrator("radtan")
    A0, W0, k0 = cal._decomposeParameterVector(g["P0"])
    capsys.readouterr()
    cal.refineCalibrationParameters(A0, W0, k0, dets, 100, shouldPrint=True)
    got = capsys.readouterr().out
    want = open(os.path.join(os.path.dirname(__file__), "golden", "print_cfg1.txt")).read()
    num = re.compile(r"[-+]?(?:\d+\.\d*|\.\d+|\d+)(?:[eE][-+]?\d+)?")

    def split(text):
        text = re.sub(r"\(\d+\.\d+s\)", "(Ts)", text)
        vals = [float(x) for x in num.findall(text)]
        skeleton = re.sub(r"[ ]+", " ", num.sub("#", text))
        return skeleton, np.array(vals)
    sk_g, v_g = split(got)
    sk_w, v_w = split(want)
    assert sk_g == sk_w
    assert len(v_g) == len(v_w)
    # printed numbers: errors to 3 decimals (the last iterations print 0.000), parameters to 8 digits
    assert np.allclose(v_g, v_w, rtol=1e-6, atol=2e-3)


def test_resident_detections_reentry(golden):
    """The reference's re-entry pattern (src/animate.py:40-42: maxIters=1, 50 times over one allDetections)
    with keepResident=True: the detections go to the device once; every call gives what a fresh upload gives."""
    c = synth.config("cfg3", 0.25)
    cal = cc.Calibrator(0, keepResident=True)
    fresh = cc.Calibrator(0)
    A, W, k = c["init"]
    for i in range(50):
        sse, A, W, k = cal.refineCalibrationParameters(A, W, k, c["dets"], 1)
        assert cal.lastIters == 1 and cal.lastTrace[0, 2] == 1e-3
        if i in (0, 1, 49):
            Ai, Wi, ki = c["init"] if i == 0 else prev
            s2, A2, W2, k2 = fresh.refineCalibrationParameters(Ai, Wi, ki, c["dets"], 1)
            assert s2 == sse and np.array_equal(A2, A) and np.array_equal(k2, k)
            assert all(np.array_equal(a, b) for a, b in zip(W2, W))
        prev = (A, W, k)
    assert cal.uploadCount == 1 and fresh.uploadCount == 3
    # another list object (even with the same arrays) is a new problem
    cal.refineCalibrationParameters(A, W, k, list(c["dets"]), 1)
    assert cal.uploadCount == 2
    cal.release()


def test_one_iteration_reentry(golden):
    """animate.py:40-42 re-enters the path with maxIters=1 (lambda resets each call)."""
    g = golden("cfg1")
    prob = _problem(g)
    P = g["P0"].copy()
    for i in range(3):
        sse, P, iters, tr = prob.lm_run(P, 1)
        assert iters == 1 and tr[0, 2] == 1e-3
    prob.close()


def test_error_paths(golden):
    g = golden("cfg1")
    with pytest.raises(ValueError):
        cc.Problem(0, g["view_offsets"], g["model_pts"][:-1], g["sensor_pts"])
    with pytest.raises(cc.CalibError):
        cc.Problem(0, np.array([0, 5, 3]), g["model_pts"][:3], g["sensor_pts"][:3])
    prob = _problem(g)
    with pytest.raises(ValueError):
        prob.cost(g["P0"][:-1])
    with pytest.raises(cc.CalibError):
        prob.lm_run(g["P0"], 0)
    prob.close()


def test_segment_size_independence(golden):
    """Task size is a tuning knob: results may differ only in summation order."""
    g = golden("itest")
    prob = _problem(g)
    ref = prob.normal_equations(g["P0"])
    for seg in (32, 96, 512):
        prob.set_segment(seg)
        out = prob.normal_equations(g["P0"])
        for a, b in zip(ref[:4], out[:4]):
            assert np.abs(a - b).max() <= 1e-12 * np.abs(a).max()
        sse, P, iters, tr = prob.lm_run(g["P0"], 100)
        assert iters == int(g["iters"]) and param_relerr(P, g["Pfinal"]) < TOL_P
    prob.close()


def test_ragged_and_degenerate_views(golden):
    g = golden("cfg1")
    offs = g["view_offsets"].copy()
    # an empty view in the middle: its 6x6 block is singular -> every step rejected, P unchanged
    offs2 = np.concatenate([offs[:3], [offs[2]], offs[3:]])
    P0 = np.concatenate([g["P0"][:10 + 12], g["P0"][10 + 6:10 + 12], g["P0"][10 + 12:]])
    prob = cc.Problem(0, offs2, g["model_pts"], g["sensor_pts"])
    sse, P, iters, tr = prob.lm_run(P0, 100)
    assert np.all(tr[:, 3] == 0.0) and np.array_equal(P, P0)
    assert iters == 13 and tr[-1, 2] == 1e9     # lambda walks 1e-3 -> 1e10, IEEE *10 chain
    prob.close()
    # single-point and 1-view problems do not crash
    prob = cc.Problem(0, np.array([0, 1]), g["model_pts"][:1], g["sensor_pts"][:1])
    prob.lm_run(g["P0"][:16], 5)
    prob.close()


def test_mid_size_vs_oracle():
    """Down-scaled cfg3 (30 views x 40x40 board, noise, perturbed-truth start): the dense reference
    cannot run here in reasonable time; the oracle (validated on the goldens) is the checker."""
    c = synth.config("cfg3", scale=0.03)
    cal = cc.Calibrator("radtan")
    P0 = cal._composeParameterVector(*c["init"]).ravel()
    offs, mp, sp = cc.flatten_detections(c["dets"])
    prob = cc.Problem(0, offs, mp, sp)
    r, Jc, Jv = prob.residual_jacobian(P0)
    Jco, Jvo = oc.jacobian_blocks(0, P0, offs, mp)
    assert jac_relerr(Jc, Jco) < TOL_J and jac_relerr(Jv, Jvo) < TOL_J
    sse, P, iters, tr = prob.lm_run(P0, 100)
    so, Po, io, to = oc.lm(0, P0, offs, mp, sp, 100)
    stop = _check_trace(tr, to)
    assert stop >= 5
    # same iterate sequence: after the last pre-plateau iteration both are at the same point
    _, Ps, _, _ = prob.lm_run(P0, stop)
    _, Pso, _, _ = oc.lm(0, P0, offs, mp, sp, stop)
    assert param_relerr(Ps, Pso) < TOL_P
    # past the plateau (|E'-E| < 1e-10 E) accept/reject is rounding noise in the reference too; a
    # cost that is flat to 1e-13 pins the parameters only to its square root
    assert param_relerr(P, Po) < 1e-6
    assert abs(sse - so) < 1e-9 * so
    A, W, k = cal._decomposeParameterVector(P)
    assert np.allclose(A, c["A"], atol=2.0) and np.allclose(k, c["k"], atol=0.05)
    prob.close()


def test_loop_modes_identical(golden):
    """The device-driven loop (one graph with a WHILE node) and the host-driven loop run the same
    kernels in the same order: bit-identical traces and parameters."""
    for case in ("cfg1", "itest", "cfg2"):
        g = golden(case)
        prob = _problem(g)
        prob.set_loop_mode(1)
        s1, P1, i1, t1 = prob.lm_run(g["P0"], 100)
        for mode in (2, 0):
            prob.set_loop_mode(mode)
            s2, P2, i2, t2 = prob.lm_run(g["P0"], 100)
            assert i1 == i2 and s1 == s2, (case, mode, i1, i2)
            assert np.array_equal(t1, t2) and np.array_equal(P1, P2), (case, mode)
            # and a run that is cut by maxIters stops at exactly that many iterations
            s3, P3, i3, t3 = prob.lm_run(g["P0"], 7)
            assert i3 == 7 and np.array_equal(t3, t1[:7]), (case, mode)
        prob.close()


def test_damping_ladder_is_invisible(golden):
    """Trial steps taken ahead of the accept/reject decisions (calib_set_ladder) change the number of
    launches and nothing else: traces, iterates and iteration counts are bit-identical to the
    one-candidate-per-pass loop, in both loop modes and when maxIters cuts a pass short."""
    for case in ("cfg1", "cfg1n", "itest", "cfg2"):
        g = golden(case)
        prob = _problem(g)
        prob.set_ladder(1)
        prob.set_loop_mode(2)
        s1, P1, i1, t1 = prob.lm_run(g["P0"], 100)
        n1 = prob.last_run_stats()[1]
        for cap in (2, 4):
            for start in (1, 2, 0):
                for mode in (1, 2):
                    prob.set_ladder(cap)
                    prob.set_ladder_start(start)
                    prob.set_loop_mode(mode)
                    s2, P2, i2, t2 = prob.lm_run(g["P0"], 100)
                    assert i1 == i2 and s1 == s2, (case, cap, start, mode, i1, i2)
                    assert np.array_equal(t1, t2) and np.array_equal(P1, P2), (case, cap, start, mode)
                assert prob.last_run_stats()[1] <= n1
        rejected = np.where(t1[:, 3] == 0)[0]
        cuts = {3, i1 - 1} | ({int(rejected[0]) + 2, int(rejected[0]) + 3} if len(rejected) else set())
        for cut in sorted(c for c in cuts if 0 < c < i1):
            prob.set_ladder(1)
            sa, Pa, ia, ta = prob.lm_run(g["P0"], cut)
            prob.set_ladder(4)
            sb, Pb, ib, tb = prob.lm_run(g["P0"], cut)
            assert ia == ib == cut and sa == sb and np.array_equal(ta, tb) and np.array_equal(Pa, Pb), (case, cut)
        prob.close()


# ---- batch of independent calibrations (BASELINE config 4) -------------------------------------
def _batch_inputs(n, views=20):
    probs = synth.batch_config(n, views)
    cal = cc.Calibrator("radtan")
    P0 = [cal._composeParameterVector(*p["init"]).ravel() for p in probs]
    return probs, P0


def test_batch_vs_oracle_and_single():
    """Every problem of a batch must follow the same iterates as the oracle and as the
    single-problem path (same device math, different mapping onto the GPU)."""
    probs, P0 = _batch_inputs(6)
    bp = cc.BatchProblem(0, [p["dets"] for p in probs])
    sse, P, iters, tr, ms = bp.lm_run(np.concatenate(P0), 100, want_trace=True)
    for j, sl in enumerate(bp.param_slices()):
        offs, mp, sp = cc.flatten_detections(probs[j]["dets"])
        so, Po, io, to = oc.lm(0, P0[j], offs, mp, sp, 100)
        t = tr[j, :iters[j]]
        stop = _check_trace(t, to)
        assert stop >= 5
        assert abs(sse[j] - so) < 1e-9 * so
        assert param_relerr(P[sl], Po) < 1e-6         # past the plateau: see test_mid_size_vs_oracle
        assert sse[j] == t[-1, 0]
        # same iterates at the last pre-plateau iteration, against the single-problem CUDA path
        prob = cc.Problem(0, offs, mp, sp)
        _, Ps, _, _ = prob.lm_run(P0[j], stop)
        prob.close()
        bp1 = cc.BatchProblem(0, [probs[j]["dets"]])
        _, Pb, itb, _, _ = bp1.lm_run(P0[j], stop)
        bp1.close()
        assert itb[0] == stop
        assert param_relerr(Pb, Ps) < TOL_P
        _, Pso, _, _ = oc.lm(0, P0[j], offs, mp, sp, stop)
        assert param_relerr(Pb, Pso) < TOL_P
    bp.close()


def test_batch_golden_cases(golden):
    """The reference's own cases as one-problem batches: radtan and fisheye, noise-free counts exact."""
    for case in ("cfg1", "cfg2", "itest"):
        g = golden(case)
        offs = g["view_offsets"]
        dets = [(g["sensor_pts"][offs[i]:offs[i + 1]], g["model_pts"][offs[i]:offs[i + 1]]) for i in range(len(offs) - 1)]
        bp = cc.BatchProblem(int(g["model"]), [dets])
        sse, P, iters, tr, ms = bp.lm_run(g["P0"], int(g["maxIters"]), want_trace=True)
        _check_trace(tr[0, :iters[0]], g["trace"])
        assert param_relerr(P, g["Pfinal"]) < TOL_P
        if case in ("cfg1", "itest"):
            assert iters[0] == int(g["iters"]) and sse[0] < 1e-12
        bp.close()


def test_batch_ragged_and_errors():
    probs, P0 = _batch_inputs(3)
    # problems with different numbers of views in one batch
    dets = [probs[0]["dets"][:7], probs[1]["dets"], probs[2]["dets"][:13]]
    Pcat = np.concatenate([P0[0][:10 + 42], P0[1], P0[2][:10 + 78]])
    bp = cc.BatchProblem(0, dets)
    sse, P, iters, tr, ms = bp.lm_run(Pcat, 100, want_trace=True)
    for j, sl in enumerate(bp.param_slices()):
        offs, mp, sp = cc.flatten_detections(dets[j])
        so, Po, io, to = oc.lm(0, Pcat[sl], offs, mp, sp, 100)
        _check_trace(tr[j, :iters[j]], to)
        assert abs(sse[j] - so) < 1e-9 * so
    with pytest.raises(ValueError):
        bp.lm_run(Pcat[:-1], 10)
    with pytest.raises(cc.CalibError):
        bp.lm_run(Pcat, 0)
    bp.close()
    with pytest.raises(cc.CalibError):
        cc.BatchProblem(0, [probs[0]["dets"] * 8])      # 160 views: beyond the one-block path


def test_cost_staging_modes_identical(golden):
    """The trial-cost sweep through the bulk-copy (TMA) ring and through plain loads evaluates the same
    points with the same code; a lane sums the points at the same offset from the start of the warp's
    chunk in one and from the start of the view's piece in the other, so the costs agree to rounding
    (1e-13), the accept / lambda sequence is the same up to the cost plateau and the refined
    parameters agree; each mode repeats itself bit for bit."""
    c = synth.config("cfg3", scale=0.03)
    cal = cc.Calibrator("radtan")
    P0 = cal._composeParameterVector(*c["init"]).ravel()
    offs, mp, sp = cc.flatten_detections(c["dets"])
    prob = cc.Problem(0, offs, mp, sp)
    out = []
    for mode in (0, 1):
        prob.set_cost_staging(mode)
        E = prob.cost(P0)
        sse, P, iters, tr = prob.lm_run(P0, 100)
        sse2, P2, iters2, tr2 = prob.lm_run(P0, 100)
        assert sse2 == sse and iters2 == iters and np.array_equal(tr2, tr) and np.array_equal(P2, P)
        out.append((E, sse, iters, tr, P))
    assert abs(out[0][0] - out[1][0]) <= 1e-13 * out[0][0] and abs(out[0][1] - out[1][1]) <= 1e-12 * out[0][1]
    t0, t1 = out[0][3], out[1][3]
    n = min(len(t0), len(t1))
    plateau = np.where(np.abs(t0[:, 1] - t0[:, 0]) <= 1e-10 * t0[:, 0])[0]
    stop = min(n, plateau[0]) if len(plateau) else n
    assert stop >= 5 and np.array_equal(t0[:stop, 2:], t1[:stop, 2:])
    assert np.allclose(t0[:stop, :2], t1[:stop, :2], rtol=1e-12, atol=0)
    assert param_relerr(out[1][4], out[0][4]) < 1e-7
    prob.close()
    for case in ("cfg1", "cfg2"):            # ragged tiny views, odd point counts, both models
        g = golden(case)
        prob = _problem(g)
        prob.set_cost_staging(1)
        sse, P, iters, tr = prob.lm_run(g["P0"], int(g["maxIters"]))
        _check_trace(tr, g["trace"])
        assert param_relerr(P, g["Pfinal"]) < TOL_P
        prob.close()


# ---- per-view homography refinement --------------------------------
@pytest.mark.parametrize("case", ["cfg1", "cfg1n", "cfg2", "itest"])
def test_homography_refinement_vs_reference(golden, case):
    """calib_refine_homographies against the reference's own refined homographies and traces
    (tests/golden/homography.npz) and against the numpy oracle: identical accept/reject and lambda
    up to the cost plateau, costs <= 1e-9 relative there, refined H <= 1e-6 relative (after the
    plateau every implementation takes rounding-level steps along a flat valley; the oracle itself is
    within 3e-7 of the reference)."""
    import os
    hg = np.load(os.path.join(os.path.dirname(__file__), "golden", "homography.npz"))
    g = golden(case)
    prob = _problem(g)
    Href, iters, tr = prob.refine_homographies(hg[case + "_H0"], 20, want_trace=True)
    prob.close()
    offs, mp, sp = g["view_offsets"], g["model_pts"], g["sensor_pts"]
    for v in range(len(offs) - 1):
        gt = hg[case + "_trace"][v][:int(hg[case + "_iters"][v])]
        pl = np.where(np.abs(gt[:, 1] - gt[:, 0]) <= 1e-10 * gt[:, 0])[0]
        stop = min(len(gt), int(iters[v]), int(pl[0]) if len(pl) else 99)
        assert stop >= 2
        assert np.array_equal(gt[:stop, 3], tr[v, :stop, 3]) and np.array_equal(gt[:stop, 2], tr[v, :stop, 2])
        assert np.allclose(gt[:stop, :2], tr[v, :stop, :2], rtol=1e-9)
        ref = hg[case + "_Href"][v]
        assert (np.abs(Href[v].ravel() - ref) / np.maximum(np.abs(ref), 1e-3)).max() < 1e-6, (case, v)
        assert Href[v][2, 2] == 1.0
        if not len(pl):      # no plateau inside the run: the count itself must match
            assert int(iters[v]) == len(gt)
        Ho, ito, tro = onp.homography_lm(hg[case + "_H0"][v], sp[offs[v]:offs[v + 1]], mp[offs[v]:offs[v + 1]])
        assert (np.abs(Href[v] - Ho) / np.maximum(np.abs(Ho), 1e-3)).max() < 1e-6


def test_homography_mirror_and_edge_cases(golden):
    """Calibrator._refineHomographies returns the reference's list of (3,3) arrays; a view with too
    few points (singular system) keeps a finite or rejected result without disturbing its neighbours."""
    import os
    hg = np.load(os.path.join(os.path.dirname(__file__), "golden", "homography.npz"))
    g = golden("cfg1")
    offs, mp, sp = g["view_offsets"], g["model_pts"], g["sensor_pts"]
    dets = [(sp[offs[v]:offs[v + 1]], mp[offs[v]:offs[v + 1]]) for v in range(len(offs) - 1)]
    cal = cc.Calibrator("radtan")
    Hs = [hg["cfg1_H0"][v].reshape(3, 3) for v in range(len(dets))]
    out = cal._refineHomographies(Hs, dets)
    assert isinstance(out, list) and out[0].shape == (3, 3)
    for v, H in enumerate(out):
        ref = hg["cfg1_Href"][v]
        assert (np.abs(H.ravel() - ref) / np.maximum(np.abs(ref), 1e-3)).max() < 1e-6
    one = cal._refineHomography(Hs[3], *dets[3])
    assert np.array_equal(one, out[3])
    # ragged: view 1 cut to 3 points (under-determined), view 2 empty
    dets2 = [dets[0], (dets[1][0][:3], dets[1][1][:3]), (dets[2][0][:0], dets[2][1][:0]), dets[3]]
    out2 = cal._refineHomographies(Hs[:4], dets2)
    assert np.array_equal(out2[0], out[0]) and np.array_equal(out2[3], out[3])


# ---- closed-form initial estimate ----------------------------------
@pytest.mark.parametrize("case", ["cfg1", "cfg1n", "cfg2", "itest"])
def test_initial_estimate_vs_reference(golden, case):
    """Calibrator.estimateCalibrationParameters (DLT Gram matrices, homography LM and distortion rows
    on the device, O(M) algebra on the host) against the reference's own P0 and against the oracle.
    Tolerances: DLT homographies <= 1e-7 relative to the reference's (eigenvector of M^T M instead of
    the singular vector of M), P0 <= 5e-6 relative (dominated by the rounding-level plateau steps of the
    homography refinement on both sides; angles compared modulo 360 degrees)."""
    import os
    import linear_oracle as lo
    from camera_calibration_b200 import linearcalibrate
    hg = np.load(os.path.join(os.path.dirname(__file__), "golden", "homography.npz"))
    g = golden(case)
    offs, mp, sp, model = g["view_offsets"], g["model_pts"], g["sensor_pts"], int(g["model"])
    dets = [(sp[offs[v]:offs[v + 1]], mp[offs[v]:offs[v + 1]]) for v in range(len(offs) - 1)]
    prob = _problem(g)
    H0 = linearcalibrate.estimateHomographies(prob).reshape(-1, 9)
    ref = hg[case + "_H0"]
    assert (np.abs(H0 - ref) / np.maximum(np.abs(ref), 1e-3)).max() < 1e-7
    # the device's Jacobi eigenvectors against LAPACK's on the same Gram sums
    H1 = linearcalibrate.estimateHomographiesHostEig(prob).reshape(-1, 9)
    assert (np.abs(H0 - H1) / np.maximum(np.abs(H1), 1e-3)).max() < 1e-8
    # distortion rows against the oracle at the oracle's own (A, W)
    A, Ws, k, Hs = lo.estimate_calibration_parameters(model, dets)
    kd = linearcalibrate.estimateDistortion(prob, A, Ws)
    assert np.allclose(kd, k, rtol=1e-6, atol=1e-9)
    prob.close()
    cal = cc.Calibrator("radtan" if model == 0 else "fisheye")
    A0, W0, k0 = cal.estimateCalibrationParameters(dets)
    P = cal._composeParameterVector(A0, W0, k0).ravel()
    d = np.abs(P - g["P0"])
    d = np.minimum(d, np.abs(d - 360))
    assert (d / np.maximum(np.abs(g["P0"]), 1e-3)).max() < 5e-6


def test_calibrate_end_to_end_like_itest(golden):
    """The reference's own acceptance test (tests/itest_main.py:12-29) on our Calibrator.calibrate:
    noise-free synthetic data -> intrinsics and distortion within 1e-9 of the ground truth, sse ~ 0."""
    c = synth.config("cfg1")
    cal = cc.Calibrator("radtan")
    import contextlib
    import io
    with contextlib.redirect_stdout(io.StringIO()):
        sse, A, W, k = cal.calibrate(c["dets"], 100)
    assert sse < 1e-10
    assert np.allclose(A, c["A"], atol=1e-9) and np.allclose(k, c["k"], atol=1e-9)


# ---- BASELINE full sizes: size-independent properties + the block-sparse oracle --------------------
@pytest.mark.parametrize("name,model_name", [("cfg3", "radtan"), ("cfg5", "fisheye")])
def test_full_size_properties_and_oracle(name, model_name):
    """cfg3 (1000 views x 40x40 grid, 1.55 M points, K = 6010) and cfg5 (10 000 views x 64x64 grid,
    fisheye, 40.9 M points, K = 60 009; the trial-cost sweep goes through the TMA ring) at their full
    sizes: the dense reference cannot hold J (154 GB / 28.8 GB for JTJ alone).  Checked here: the C oracle's trace up to the plateau and iterate there
    (<= 1e-8), and properties that hold at any size -- E(i+1) = E'(i) after every accepted step,
    accepted costs strictly decrease, calib_cost(P_final) reproduces the last accepted cost bit for bit,
    the gradient J^T r at the solution is ~0 relative to its start, the sum of the per-view block costs
    (each view refined alone is NOT needed: one calib_cost per half of the views) equals the whole."""
    c = synth.config(name, scale=1.0)
    cal = cc.Calibrator(model_name)
    mid = c["model"]
    L = 10 if mid == 0 else 9
    P0 = cal._composeParameterVector(*c["init"]).ravel()
    offs, mp, sp = cc.flatten_detections(c["dets"])
    M = len(offs) - 1
    prob = cc.Problem(mid, offs, mp, sp)
    sse, P, iters, tr = prob.lm_run(P0, 100)
    so, Po, io, to = oc.lm(mid, P0, offs, mp, sp, 100)
    stop = _check_trace(tr, to)
    assert stop >= 5
    _, Ps, _, _ = prob.lm_run(P0, stop)
    _, Pso, _, _ = oc.lm(mid, P0, offs, mp, sp, stop)
    assert param_relerr(Ps, Pso) < TOL_P
    assert abs(sse - so) < 1e-9 * so
    acc = tr[:, 3] == 1
    for i in range(iters - 1):
        assert tr[i + 1, 0] == (tr[i, 1] if acc[i] else tr[i, 0])
    assert np.all(tr[acc, 1] < tr[acc, 0])
    last = tr[acc][-1, 1] if acc.any() else tr[0, 0]
    assert prob.cost(P) == last
    A0, B0, D0, g0, E0 = prob.normal_equations(P0)
    A1, B1, D1, g1, E1 = prob.normal_equations(P)
    assert np.linalg.norm(g1) < 1e-6 * np.linalg.norm(g0)
    prob.close()
    # additivity over views: cost(all) = cost(first half) + cost(second half) to rounding
    h = M // 2
    parts = 0.0
    for lo, hi in ((0, h), (h, M)):
        o2 = offs[lo:hi + 1] - offs[lo]
        p2 = cc.Problem(mid, o2, mp[offs[lo]:offs[hi]], sp[offs[lo]:offs[hi]])
        parts += p2.cost(np.concatenate([P[:L], P[L + 6 * lo:L + 6 * hi]]))
        p2.close()
    assert abs(parts - last) < 1e-11 * last


def test_calibrate_camera_from_json():
    """main.calibrateCamera (src/main.py:11-36) fed from the reference's detections JSON: the user-facing
    path end to end (file -> initial estimate -> refinement), ground truth recovered as in itest_main."""
    import contextlib
    import io
    import os
    from camera_calibration_b200 import dataset
    dets = dataset.createDetectionsFromPath(os.path.join(os.path.dirname(__file__), "golden", "detections_cfg1.json"))
    with contextlib.redirect_stdout(io.StringIO()):
        sse, A, W, k = cc.calibrateCamera(dets, "radtan", 100)
    c = synth.config("cfg1")
    assert sse < 1e-10 and np.allclose(A, c["A"], atol=1e-9) and np.allclose(k, c["k"], atol=1e-9)
    assert len(W) == len(dets) and W[0].shape == (4, 4)
    with pytest.raises(ValueError):
        cc.calibrateCamera(dets, "pinhole", 10)


@pytest.mark.parametrize("model,k,sigma", [("radtan", synth.K_RADTAN, None), ("fisheye", synth.K_FISHEYE, 0.1)])
def test_dataset_generation_on_device(model, k, sigma):
    """Dataset._computeDetections / VirtualCamera._measurePoints (src/dataset.py:59-82,
    src/virtualcamera.py:42-55) with the projection on the device: same poses, same in-image masks, same
    noise as the host generator (which is bit-identical to the reference's), coordinates <= 1e-9 px."""
    corners = synth.checkerboard_corners(9, 6, 0.050)
    host, ph = synth.make_dataset(corners, synth.A_SYNTHETIC, k, model, 1440, 1080, 12, sigma)
    dev, pd = synth.make_dataset_on_device(corners, synth.A_SYNTHETIC, k, model, 1440, 1080, 12, sigma)
    assert len(host) == len(dev) == 12
    for (sh, mh), (sd, md), a, b in zip(host, dev, ph, pd):
        assert np.array_equal(a, b) and np.array_equal(mh, md)
        assert sh.shape == sd.shape and np.abs(sh - sd).max() < 1e-9


def test_full_size_batch_cfg4():
    """BASELINE config 4 at its full size: 8192 independent 20-view calibrations in one launch.  A
    problem's result must not depend on the batch around it (bit-identical to the same problem run in a
    batch of one -- checked on a sample), the oracle must agree on the sample, and over ALL problems:
    iterations > 0, a finite cost that is below the cost at the initial guess, sse equal to the last
    trace row's E."""
    probs, P0 = _batch_inputs(8192)
    bp = cc.BatchProblem(0, [p["dets"] for p in probs])
    sse, P, iters, tr, ms = bp.lm_run(np.concatenate(P0), 100, want_trace=True)
    slices = bp.param_slices()
    bp.close()
    assert np.all(iters > 0) and np.all(np.isfinite(sse))
    last = tr[np.arange(len(iters)), iters - 1]
    assert np.array_equal(sse, last[:, 0])
    assert np.all(sse <= tr[:, 0, 0])
    for j in (0, 1, 4095, 8191):
        offs, mp, sp = cc.flatten_detections(probs[j]["dets"])
        b1 = cc.BatchProblem(0, [probs[j]["dets"]])
        s1, P1, i1, t1, _ = b1.lm_run(P0[j], 100, want_trace=True)
        b1.close()
        assert i1[0] == iters[j] and s1[0] == sse[j] and np.array_equal(P1, P[slices[j]])
        assert np.array_equal(t1[0, :i1[0]], tr[j, :iters[j]])
        so, Po, io, to = oc.lm(0, P0[j], offs, mp, sp, 100)
        assert _check_trace(tr[j, :iters[j]], to) >= 5
        assert abs(sse[j] - so) < 1e-9 * max(so, 1e-12) or (sse[j] < 1e-10 and so < 1e-10)


@pytest.mark.gpu
def test_create_from_views_matches_flat_upload(golden):
    """calib_create_from_views (the reference's list of per-view arrays, gathered and uploaded by the
    library in pipelined chunks) must leave exactly the detections calib_create gets from the
    concatenated arrays: projections, costs and a whole LM run are bit-identical.  Sizes on both sides
    of the 2 MB chunk, ragged and empty views, read-only arrays, and the numpy fallback for odd inputs."""
    g = golden("cfg1")
    offs = g["view_offsets"]
    dets = [(g["sensor_pts"][offs[i]:offs[i + 1]].copy(), g["model_pts"][offs[i]:offs[i + 1]].copy())
            for i in range(len(offs) - 1)]
    flat = cc.Problem(0, offs, g["model_pts"], g["sensor_pts"])
    views = cc.Problem.from_detections(0, dets)
    assert views.M == flat.M and views.N == flat.N and np.array_equal(views.view_offsets, offs)
    assert np.array_equal(views.project(g["P0"]), flat.project(g["P0"]))
    ra, rb = flat.lm_run(g["P0"], 100), views.lm_run(g["P0"], 100)
    assert ra[0] == rb[0] and np.array_equal(ra[1], rb[1]) and np.array_equal(ra[3], rb[3])
    flat.close(), views.close()
    # read-only arrays and the list fallback
    ro = [(s.copy(), m.copy()) for s, m in dets]
    for s, m in ro:
        s.flags.writeable = False
        m.flags.writeable = False
    odd = [(s.tolist(), m.tolist()) for s, m in dets]
    for d in (ro, odd):
        p = cc.Problem.from_detections(0, d)
        r = p.lm_run(g["P0"], 100)
        assert r[0] == ra[0] and np.array_equal(r[1], ra[1])
        p.close()
    # several chunks per array, views that straddle chunk boundaries, an empty view
    c = synth.config("cfg3", 0.25)
    big = list(c["dets"])
    big[5] = (big[5][0][:0], big[5][1][:0])
    offs2, mp, sp = cc.flatten_detections(big)
    assert mp.nbytes > 3 * (2 << 20)
    P0 = cc.Calibrator(c["model"])._composeParameterVector(*c["init"])
    a, b = cc.Problem(c["model"], offs2, mp, sp), cc.Problem.from_detections(c["model"], big)
    assert np.array_equal(a.project(P0), b.project(P0))
    assert a.cost(P0) == b.cost(P0)
    a.close()
    b2 = cc.Problem.from_detections(c["model"], big)       # staging slots reused straight away
    assert np.array_equal(b.project(P0), b2.project(P0))
    b.close(), b2.close()
    with pytest.raises(cc.CalibError):
        L = cc._lib.lib()
        import ctypes
        h = ctypes.c_void_p()
        cnt = np.array([3], dtype=np.int64)
        nul = np.zeros(1, dtype=np.uint64)
        rc = L.calib_create_from_views(ctypes.byref(h), 0, 1, cnt.ctypes.data_as(ctypes.POINTER(ctypes.c_int64)),
                                       nul.ctypes.data, nul.ctypes.data, 0)
        assert rc == -1 and not h
        raise cc.CalibError(rc, L.calib_last_error(None).decode())


def test_fisheye_wide_angle_vs_oracle():
    """Fisheye points far off the axis (r = |(x,y)|/z up to ~8, where the device atan takes its
    pi/2 - atan(1/r) branch) and close to it: projection, residuals and Jacobian against the numpy
    oracle at the parity tolerance; a point ON the axis gives NaN in both (src/distortion.py:215)."""
    rs = np.random.RandomState(21)
    M, n = 6, 400
    offs = np.arange(M + 1) * n
    pts = np.zeros((M * n, 3))
    ang = rs.uniform(0, 2 * np.pi, M * n)
    rad = np.concatenate([rs.uniform(1e-4, 0.3, M * n // 2), rs.uniform(0.3, 8.0, M * n - M * n // 2)])
    pts[:, 0], pts[:, 1] = rad * np.cos(ang), rad * np.sin(ang)
    intr = np.array([420.0, 415.0, 0.3, 640.0, 512.0, -0.03, 0.004, -0.0007, 0.0001])
    ext = np.column_stack([rs.uniform(-3, 3, M), rs.uniform(-3, 3, M), rs.uniform(-20, 20, M),
                           rs.uniform(-0.02, 0.02, M), rs.uniform(-0.02, 0.02, M), rs.uniform(0.9, 1.1, M)])
    P = np.concatenate([intr, ext.ravel()])
    prob = cc.Problem(1, offs, pts, np.zeros((M * n, 2)))
    y = prob.project(P)
    y_ref = onp.project(1, P, offs, pts)
    assert np.abs(y - y_ref).max() <= 1e-11 * np.abs(y_ref).max()
    r, Jc, Jv = prob.residual_jacobian(P)
    Jc_ref, Jv_ref = onp.jacobian_blocks(1, P, offs, pts)
    assert jac_relerr(Jc, np.asarray(Jc_ref).reshape(Jc.shape)) < TOL_J
    assert jac_relerr(Jv, np.asarray(Jv_ref).reshape(Jv.shape)) < TOL_J
    prob.close()
    on_axis = np.zeros((1, 3))
    P1 = np.concatenate([intr, [0, 0, 0, 0, 0, 1.0]])
    prob = cc.Problem(1, np.array([0, 1]), on_axis, np.zeros((1, 2)))
    with np.errstate(all="ignore"):
        assert np.isnan(prob.project(P1)).all() and np.isnan(onp.project(1, P1, np.array([0, 1]), on_axis)).all()
    prob.close()


def test_many_handles_created_and_destroyed(golden):
    """Handles come and go (the e2e path creates one per refinement): destroyed handles park their
    loop graphs for the next calib_create to tear down, streams and events are recycled.  Forty
    create / run / destroy rounds over three problems, some destroyed out of order, give the same
    answers as the first round, and more than the parked-graph bound are never kept."""
    gs = [golden(n) for n in ("cfg1", "cfg2", "itest")]
    first = {}
    held = []
    for k in range(40):
        g = gs[k % 3]
        prob = _problem(g)
        sse, P, iters, tr = prob.lm_run(g["P0"], 30)
        key = k % 3
        if key in first:
            assert sse == first[key][0] and iters == first[key][2] and np.array_equal(P, first[key][1])
        else:
            first[key] = (sse, P, iters)
        if k % 5 == 0:
            held.append(prob)          # destroyed later, out of order
        else:
            prob.close()
    for prob in reversed(held):
        prob.close()
    prob = _problem(gs[0])
    assert prob.lm_run(gs[0]["P0"], 30)[0] == first[0][0]
    prob.close()


def test_trim_returns_the_pool_and_the_library_goes_on(golden):
    """The library's own memory pool (one per device, never the device's default pool) keeps what destroyed
    handles free; calib_trim hands it back to the driver, with a live handle untouched and the next create
    working as before.  The default pool's release threshold is what the process set it to (0 here)."""
    import ctypes
    from camera_calibration_b200 import _lib
    L = _lib.lib()
    g = golden("cfg2")
    live = _problem(g)
    ref = live.lm_run(g["P0"], 30)
    for _ in range(3):
        prob = _problem(g)
        prob.lm_run(g["P0"], 5)
        prob.close()
    assert L.calib_trim(0) == 0
    assert L.calib_trim(-1) != 0 and L.calib_trim(63) == 0        # bad index refused; unused device: nothing to do
    again = live.lm_run(g["P0"], 30)                                # the live handle's memory was not trimmed
    assert again[0] == ref[0] and np.array_equal(again[1], ref[1])
    live.close()
    prob = _problem(g)
    assert prob.lm_run(g["P0"], 30)[0] == ref[0]
    prob.close()
    rt = ctypes.CDLL("libcudart.so.12")
    pool = ctypes.c_void_p()
    assert rt.cudaDeviceGetDefaultMemPool(ctypes.byref(pool), 0) == 0
    thr = ctypes.c_uint64(123)
    assert rt.cudaMemPoolGetAttribute(pool, 4, ctypes.byref(thr)) == 0     # cudaMemPoolAttrReleaseThreshold = 4
    assert thr.value == 0


def test_soak_repeats_bit_for_bit(golden):
    """Stand-in for the race checker that is closed on this pool (compute-sanitizer): the kernels' cross-block
    and cross-warp hand-offs -- the ticketed last-block reductions of k_schur / k_cost, the producer / consumer
    rings of k_eval -- are timing dependent if they are wrong.  300 refinements back to back of a 48 000-point
    problem (40 views, several pieces per view and per warp) and of the small fisheye case, through fresh
    handles and through one long-lived handle, both loop modes: every run must repeat run 0 bit for bit."""
    c = synth.config("cfg3", 0.04)
    P3 = cc.Calibrator(0)._composeParameterVector(*c["init"]).ravel()
    offs, mp, sp = cc.flatten_detections(c["dets"])
    g2 = golden("cfg2")
    cases = [(0, offs, mp, sp, P3), (int(g2["model"]), g2["view_offsets"], g2["model_pts"], g2["sensor_pts"], g2["P0"])]
    first = {}
    held = [cc.Problem(m, o, a, b) for m, o, a, b, _ in cases]
    for k in range(300):
        ci = k % 2
        m, o, a, b, P0 = cases[ci]
        if k % 3 == 0:
            prob, own = held[ci], False
        else:
            prob, own = cc.Problem(m, o, a, b), True
        prob.set_loop_mode(1 if k % 7 == 0 else 0)
        sse, P, iters, tr = prob.lm_run(P0, 100)
        if ci in first:
            assert sse == first[ci][0] and iters == first[ci][2], (k, sse, first[ci][0])
            assert np.array_equal(P, first[ci][1]) and np.array_equal(tr, first[ci][3]), k
        else:
            first[ci] = (sse, P, iters, tr)
        if own:
            prob.close()
    for prob in held:
        prob.close()


def test_reference_itest_cases_through_calibrate_camera():
    """The three cases of the reference's integration test (tests/itest_main.py), through the same
    entry point (main.calibrateCamera) and with the same assertions:
    realistic radtan, 25x18 board, 15 views, no noise, maxIters 100 -> sse ~ 0 (7 places), A and k to
    1e-9 of the ground truth, every board pose to atol 1 (:12-29); the 9x6 board with 0.1 px noise -> A
    within 2.0, k within 0.05 (:31-52); fisheye, maxIters 10 -> runs and returns finite values (:54-79)."""
    import contextlib
    import io
    out = io.StringIO()
    c = synth.config("itest")
    with contextlib.redirect_stdout(out):
        sse, A, W, k = cc.calibrateCamera(c["dets"], "radtan", 100)
    assert round(sse, 7) == 0.0
    assert np.allclose(A, c["A"], atol=1e-9) and np.allclose(k, c["k"], atol=1e-9)
    assert len(W) == len(c["poses"])
    for Wc, Wt in zip(W, c["poses"]):
        assert np.allclose(Wc, Wt, atol=1.0)
    import re                                          # Calibrator.calibrate prints (shouldPrint=True, :36-38, 269-274)
    assert re.search(r"\niter 0: \(\d+\.\d{3}s\), error=\d+\.\d{3}, λ=1\.000000e-03\nA:\n\[\[", out.getvalue())
    c = synth.config("cfg1n")
    with contextlib.redirect_stdout(io.StringIO()):
        sse, A, W, k = cc.calibrateCamera(c["dets"], "radtan", 100)
    assert np.allclose(A, c["A"], atol=2.0) and np.allclose(k, c["k"], atol=0.05)
    c = synth.config("cfg2")
    with contextlib.redirect_stdout(io.StringIO()):
        sse, A, W, k = cc.calibrateCamera(c["dets"], "fisheye", 10)
    assert np.isfinite(sse) and np.isfinite(A).all() and np.isfinite(np.asarray(k)).all() and len(k) == 4
