"""CPU: the oracle (numpy and C restatements) against the golden vectors generated from the
unmodified reference (oracle/make_golden.py).  This is what pins the oracle."""
import numpy as np
import pytest

import calib_oracle as onp
import oracle_c as oc
from conftest import jac_relerr, param_relerr

CASES = ["cfg1", "cfg1n", "cfg2", "itest"]
TOL_J = 1e-10   # north_star: residuals and Jacobian entries to 1e-10 relative
TOL_P = 1e-8    # north_star: converged parameters to 1e-8 relative


@pytest.mark.parametrize("model,name", [(0, "radtan"), (1, "fisheye")])
def test_kat_vectors(golden, model, name):
    g = golden("kat")
    L = onp.num_intrinsics(model)
    # (a) the reference test's own vector, tests/test_jacobian.py:19-24
    P, pts = g[name + "_a_P"], g[name + "_a_pts"]
    offs = np.array([0, len(pts)])
    for impl in (onp, oc):
        Jc, Jv = impl.jacobian_blocks(model, P, offs, pts)
        assert Jc.shape == (2 * len(pts), L) and Jv.shape == (2 * len(pts), 6)
        assert jac_relerr(Jc, g[name + "_a_Jc"]) < TOL_J
        assert jac_relerr(Jv, g[name + "_a_Jv"]) < TOL_J
        assert np.allclose(impl.project(model, P, offs, pts), g[name + "_a_y"], rtol=1e-12, atol=1e-9)
    # (b) general vector: Z != 0, skew != 0, negative angles
    P, pts, offs = g[name + "_b_P"], g[name + "_b_pts"], g[name + "_b_offs"]
    for impl in (onp, oc):
        Jc, Jv = impl.jacobian_blocks(model, P, offs, pts)
        assert jac_relerr(Jc, g[name + "_b_Jc"]) < TOL_J
        assert jac_relerr(Jv, g[name + "_b_Jv"]) < TOL_J
        assert np.allclose(impl.project(model, P, offs, pts), g[name + "_b_y"], rtol=1e-12, atol=1e-9)


@pytest.mark.parametrize("case", CASES)
def test_residual_and_jacobian_at_initial_guess(golden, case):
    g = golden(case)
    m, offs, mp, sp, P0 = int(g["model"]), g["view_offsets"], g["model_pts"], g["sensor_pts"], g["P0"]
    for impl in (onp, oc):
        Jc, Jv = impl.jacobian_blocks(m, P0, offs, mp)
        assert jac_relerr(Jc, g["Jc0"]) < TOL_J
        assert jac_relerr(Jv, g["Jv0"]) < TOL_J
        y = impl.project(m, P0, offs, mp)
        r = (sp - y).reshape(-1)
        assert np.abs(r - g["r0"]).max() <= 1e-10 * max(1.0, np.abs(g["r0"]).max())
        E = impl.cost(m, P0, offs, mp, sp)
        assert abs(E - g["trace"][0, 0]) <= 1e-12 * g["trace"][0, 0]


@pytest.mark.parametrize("case", ["cfg1", "cfg1n", "cfg2"])
def test_normal_equations_and_first_step(golden, case):
    g = golden(case)
    m, offs, mp, sp, P0 = int(g["model"]), g["view_offsets"], g["model_pts"], g["sensor_pts"], g["P0"]
    L = onp.num_intrinsics(m)
    JTJ, JTr = g["JTJ0"], g["JTr0"]
    for impl in (onp, oc):
        A, B, D, gg, E = impl.normal_equations(m, P0, offs, mp, sp)
        sc = np.abs(JTJ).max()
        assert np.abs(A - JTJ[:L, :L]).max() < 1e-12 * sc
        for v in range(len(offs) - 1):
            c = L + 6 * v
            assert np.abs(B[v] - JTJ[:L, c:c + 6]).max() < 1e-12 * sc
            assert np.abs(D[v] - JTJ[c:c + 6, c:c + 6]).max() < 1e-12 * sc
        assert np.abs(gg - JTr).max() < 1e-12 * np.abs(JTr).max()
    # first step: the reference's explicit inverse at cond ~5e9 limits agreement
    d = oc.lm_step(m, A, B, D, gg, 1e-3)
    assert np.abs(d - g["delta0"]).max() < 1e-6 * np.abs(g["delta0"]).max()
    d2 = onp.schur_step(A, B, D, gg, 1e-3)
    assert np.abs(d - d2).max() < 1e-9 * np.abs(d2).max()


def _check_trace(tr, gt):
    """identical (accepted, lambda) sequence up to the first plateau iteration
    (|E'-E| <= 1e-10 E), where the reference's own decisions are LAPACK rounding noise."""
    n = min(len(tr), len(gt))
    plateau = np.where(np.abs(gt[:, 1] - gt[:, 0]) <= 1e-10 * gt[:, 0])[0]
    stop = min(n, plateau[0]) if len(plateau) else n
    assert np.array_equal(tr[:stop, 3], gt[:stop, 3])
    assert np.array_equal(tr[:stop, 2], gt[:stop, 2])
    big = gt[:stop, 0] > 1e-9   # costs comparable while they are above rounding level
    assert np.allclose(tr[:stop, 0][big], gt[:stop, 0][big], rtol=1e-6)
    return stop


@pytest.mark.parametrize("case,impl_name", [("cfg1", "dense"), ("cfg1", "c"), ("cfg1n", "dense"),
                                            ("cfg1n", "c"), ("cfg2", "c"), ("itest", "c")])
def test_lm_loop_against_reference(golden, case, impl_name):
    g = golden(case)
    m, offs, mp, sp, P0 = int(g["model"]), g["view_offsets"], g["model_pts"], g["sensor_pts"], g["P0"]
    if impl_name == "dense":
        sse, P, iters, tr = onp.lm_dense(m, P0, offs, mp, sp, int(g["maxIters"]))
    else:
        sse, P, iters, tr = oc.lm(m, P0, offs, mp, sp, int(g["maxIters"]))
    gt = g["trace"]
    _check_trace(tr, gt)
    # the explicit inverse at cond ~5e9 makes the plateau steps of the dense variant (and of the
    # reference itself) noisy at the 1e-8 level in the weakly determined skew; the Cholesky/Schur
    # solve lands within 1e-10 of the reference's fixed point
    assert param_relerr(P, g["Pfinal"]) < (1e-6 if (impl_name == "dense" and case == "cfg1n") else TOL_P)
    if case in ("cfg1", "itest"):   # noise-free: the count is solver independent
        assert iters == int(g["iters"])
        assert sse < 1e-12
    else:
        assert abs(sse - float(g["sse"])) < 1e-9 * float(g["sse"])
    if case == "cfg1n" and impl_name == "dense":
        assert iters == int(g["iters"])   # same explicit-inverse solve -> same plateau walk here


def test_sparse_numpy_equals_c(golden):
    g = golden("cfg1")
    m, offs, mp, sp, P0 = int(g["model"]), g["view_offsets"], g["model_pts"], g["sensor_pts"], g["P0"]
    s1, P1, i1, t1 = onp.lm_sparse(m, P0, offs, mp, sp, 100)
    s2, P2, i2, t2 = oc.lm(m, P0, offs, mp, sp, 100)
    assert i1 == i2 and np.array_equal(t1[:, 3], t2[:, 3])
    assert param_relerr(P1, P2) < 1e-9


def test_reference_quirks():
    g_offs = np.array([0, 2])
    pts = np.array([[0.1, 0.1, 0], [0.1, 0.2, 0]])
    P = np.array([400, 400, 0, 320, 240, -0.5, 0.2, 0, 0, 0, 180, 0, 0, 0.1, 0.2, 1.0])
    with pytest.raises(UnboundLocalError):   # maxIters=0, src/calibrate.py:171
        onp.lm_dense(0, P, g_offs, pts, np.zeros((2, 2)), 0)
    with pytest.raises(UnboundLocalError):
        oc.lm(0, P, g_offs, pts, np.zeros((2, 2)), 0)
    # block sparsity of the dense layout, tests/test_jacobian.py:56-70
    P2 = np.concatenate([P, [170, 5, 3, 0.1, 0.2, 1.1]])
    J = onp.dense_jacobian(0, P2, np.array([0, 2, 4]), np.vstack([pts, pts]))
    assert J.shape == (8, 22) and J[0, 0] != 0 and J[-1, -1] != 0 and J[0, 16] == 0


# ---- per-view homography refinement --------------------------------
def _hom_cases():
    import os
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "homography.npz"))
    return g


@pytest.mark.parametrize("case", ["cfg1", "cfg1n", "cfg2", "itest"])
def test_homography_oracle_vs_reference(golden, case):
    """The numpy restatement of Calibrator._refineHomography / HomographyJacobian against the
    reference's own outputs (oracle/make_golden_homography.py): Jacobian entries <= 1e-12 relative,
    identical accept/reject and costs up to the cost plateau, refined homography <= 1e-6 relative
    (after the plateau both sides take rounding-level steps along a flat valley)."""
    import calib_oracle as onp
    hg = _hom_cases()
    g = golden(case)
    offs, mp, sp = g["view_offsets"], g["model_pts"], g["sensor_pts"]
    J = onp.homography_jacobian(hg[case + "_H0"][0], mp[offs[0]:offs[1]])
    assert np.abs(J - hg[case + "_J0"]).max() <= 1e-12 * np.abs(hg[case + "_J0"]).max()
    for v in range(len(offs) - 1):
        H, it, tr = onp.homography_lm(hg[case + "_H0"][v], sp[offs[v]:offs[v + 1]], mp[offs[v]:offs[v + 1]])
        gt = hg[case + "_trace"][v][:int(hg[case + "_iters"][v])]
        pl = np.where(np.abs(gt[:, 1] - gt[:, 0]) <= 1e-10 * gt[:, 0])[0]
        stop = min(len(gt), len(tr), int(pl[0]) if len(pl) else 99)
        assert stop >= 2
        assert np.array_equal(gt[:stop, 3], tr[:stop, 3]) and np.array_equal(gt[:stop, 2], tr[:stop, 2])
        assert np.allclose(gt[:stop, :2], tr[:stop, :2], rtol=1e-9)
        ref = hg[case + "_Href"][v]
        assert (np.abs(H.ravel() - ref) / np.maximum(np.abs(ref), 1e-3)).max() < 1e-6


# ---- closed-form initial estimate ----------------------------------
@pytest.mark.parametrize("case", ["cfg1", "cfg1n", "cfg2", "itest"])
def test_linear_oracle_vs_reference(golden, case):
    """oracle/linear_oracle.py against the reference: its DLT homographies are the reference's bit for
    bit (same normalisation, same SVD), and the whole estimateCalibrationParameters chain reproduces the
    reference's P0 (what its own linear stage handed to the LM) to <= 2e-6 relative -- the limit is the
    homography refinement's rounding-level plateau steps (see test_homography_oracle_vs_reference)."""
    import linear_oracle as lo
    import camera_calibration_b200 as cc
    hg = _hom_cases()
    g = golden(case)
    offs, mp, sp, model = g["view_offsets"], g["model_pts"], g["sensor_pts"], int(g["model"])
    dets = [(sp[offs[v]:offs[v + 1]], mp[offs[v]:offs[v + 1]]) for v in range(len(offs) - 1)]
    H0 = np.array([lo.estimate_homography(s[:, :2], m[:, :2]).ravel() for s, m in dets])
    assert np.abs(H0 - hg[case + "_H0"]).max() <= 1e-12 * np.abs(hg[case + "_H0"]).max()
    A, Ws, k, Hs = lo.estimate_calibration_parameters(model, dets)
    P = cc.Calibrator("radtan" if model == 0 else "fisheye")._composeParameterVector(A, Ws, k).ravel()
    d = np.abs(P - g["P0"])
    d = np.minimum(d, np.abs(d - 360))
    assert (d / np.maximum(np.abs(g["P0"]), 1e-3)).max() < 2e-6


def test_device_atan_polynomial():
    """The fisheye kernels use their own atan (calib_device.cuh: atan_pos, Estrin scheme instead of
    libdevice's Horner chain).  The coefficients are read from the source and the same evaluation
    order is replayed in float64 here against 40-digit mpmath: <= 3e-16 relative on [0, 1] and, through
    pi/2 - atan(1/r), on (1, 1e6].  (The device evaluates with fused multiply-adds, one rounding fewer
    per step; the GPU check is tests/test_gpu_parity.py::test_fisheye_wide_angle_vs_oracle.)"""
    import os
    import re
    import mpmath as mp
    import numpy as np
    src = open(os.path.join(os.path.dirname(__file__), "..", "camera_calibration_b200", "csrc",
                            "calib_device.cuh")).read()
    body = src[src.index("double atan_pos("):]
    body = body[:body.index("return big")]
    pairs = re.findall(r"a(\d) = fma\(([-0-9.e]+), q, ([-0-9.e]+)\)", body)
    assert [int(p[0]) for p in pairs] == list(range(10))
    c = []
    for n, hi, lo in pairs:
        c += [float(lo), float(hi)]
    tail = src[src.index("return big", src.index("double atan_pos(")):]
    assert "1.5707963267948966" in tail[:120] and "6.123233995736766e-17" in tail[:120]

    def atan_small(z):
        q = z * z
        q2 = q * q
        q4 = q2 * q2
        q8 = q4 * q4
        q16 = q8 * q8
        a = [c[2 * i] + c[2 * i + 1] * q for i in range(10)]
        b = [a[2 * j] + a[2 * j + 1] * q2 for j in range(5)]
        d0, d1 = b[0] + b[1] * q4, b[2] + b[3] * q4
        return z + (z * q) * ((d0 + d1 * q8) + b[4] * q16)

    def atan_pos(r):
        if r > 1.0:
            return (1.5707963267948966 - atan_small(1.0 / r)) + 6.123233995736766e-17
        return atan_small(r)

    mp.mp.dps = 40
    rs = np.random.RandomState(11)
    xs = np.concatenate([rs.rand(3000), np.linspace(0.95, 1.05, 400), 1.0 + 10.0 ** rs.uniform(-3, 6, 1500),
                         [0.0, 1.0, 1e-300, 1e-9, 0.5, 2.0, 1e6]])
    worst = 0.0
    for x in xs:
        ref = mp.atan(mp.mpf(float(x)))
        got = atan_pos(float(x))
        if x == 0.0:
            assert got == 0.0
            continue
        worst = max(worst, float(abs((mp.mpf(got) - ref) / ref)))
    assert worst <= 3e-16, worst


@pytest.mark.parametrize("case", ["cfg1", "cfg1n", "cfg2", "itest"])
def test_jacobi_eigenvector_scheme_vs_lapack(golden, case):
    """The device takes the DLT homography as the eigenvector of the smallest eigenvalue of M^T M by
    cyclic Jacobi rotations (calib_linear.cuh: warp_jacobi9).  Its CPU restatement
    (linear_oracle.jacobi_smallest_eigenvector: same pivot order, skip rule and rotation formulas) on
    every view of the golden cases -- from noise-free ones without distortion residual (smallest eigenvalue
    at rounding level) to the strongly distorted itest board -- converges in well under the 30-sweep cap and gives the reference's homography
    (last right-singular vector of M) to 1e-7 relative, LAPACK's eigh on the same matrix to 1e-8."""
    import linear_oracle as lo
    hg = _hom_cases()
    g = golden(case)
    offs, mp, sp = g["view_offsets"], g["model_pts"], g["sensor_pts"]
    ref = hg[case + "_H0"].reshape(-1, 9)
    worst_sweeps = 0
    for v in range(len(offs) - 1):
        s, m = sp[offs[v]:offs[v + 1], :2], mp[offs[v]:offs[v + 1], :2]
        H, lam, sweeps = lo.estimate_homography_jacobi(s, m)
        worst_sweeps = max(worst_sweeps, sweeps)
        assert (np.abs(H.ravel() - ref[v]) / np.maximum(np.abs(ref[v]), 1e-3)).max() < 1e-7
        Na, Nb = lo.normalization_matrix(s), lo.normalization_matrix(m)
        M = lo.dlt_rows(s * Na[0, 0] + Na[:2, 2], m * Nb[0, 0] + Nb[:2, 2])
        w, vec = np.linalg.eigh(M.T @ M)
        Hl = np.linalg.inv(Na) @ vec[:, 0].reshape(3, 3) @ Nb
        Hl = (Hl / Hl[2, 2]).ravel()
        assert (np.abs(H.ravel() - Hl) / np.maximum(np.abs(Hl), 1e-3)).max() < 1e-8
        assert abs(lam - w[0]) <= 1e-12 * w[-1]
    assert worst_sweeps <= 15
