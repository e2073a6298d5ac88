"""TEST INFRASTRUCTURE ONLY -- numpy restatement of the reference's closed-form initial estimate
(`Calibrator.estimateCalibrationParameters`, /root/reference/src/calibrate.py:41-58), vectorised over
points so that it also runs at the large configurations.  Only tests/ may import this module.

  estimate_homography      linearcalibrate.estimateHomography        src/linearcalibrate.py:24-58
  normalization_matrix     linearcalibrate.computeNormalizationMatrix  :61-91
  intrinsic_matrix         linearcalibrate.computeIntrinsicMatrix      :94-155, vecHomography :158-185,
                           computeIntrinsicMatrixFrombCholesky         :256-292
  extrinsics               linearcalibrate.computeExtrinsics           :295-325, approximateRotationMatrix :328-372
  estimate_distortion      RadialTangentialModel.estimateDistortion   src/distortion.py:110-191
                           FisheyeModel.estimateDistortion             src/distortion.py:222-271
Pinned on the P0 of tests/golden/{cfg1,cfg1n,cfg2,itest}.npz, which the reference's own
estimateCalibrationParameters produced (oracle/make_golden.py).
"""
import numpy as np

import calib_oracle as onp


def normalization_matrix(X):
    mean = X.mean(axis=0)
    scale = np.sqrt(2) / np.mean(np.linalg.norm(X - mean, axis=1))
    return np.array([[scale, 0, -scale * mean[0]], [0, scale, -scale * mean[1]], [0, 0, 1.0]])


def dlt_rows(xa, xb):
    """The 2N x 9 matrix M of src/linearcalibrate.py:46-52 for normalised sensor (xa) / model (xb) points."""
    u, v, X, Y = xa[:, 0], xa[:, 1], xb[:, 0], xb[:, 1]
    N = len(u)
    M = np.zeros((2 * N, 9))
    M[0::2, 0], M[0::2, 1], M[0::2, 2] = -X, -Y, -1
    M[0::2, 6], M[0::2, 7], M[0::2, 8] = u * X, u * Y, u
    M[1::2, 3], M[1::2, 4], M[1::2, 5] = -X, -Y, -1
    M[1::2, 6], M[1::2, 7], M[1::2, 8] = v * X, v * Y, v
    return M


def estimate_homography(Xa, Xb):
    Na, Nb = normalization_matrix(Xa), normalization_matrix(Xb)
    xa = Xa * Na[0, 0] + Na[:2, 2]
    xb = Xb * Nb[0, 0] + Nb[:2, 2]
    _, _, Vt = np.linalg.svd(dlt_rows(xa, xb))
    H = np.linalg.inv(Na) @ Vt[-1].reshape(3, 3) @ Nb
    return H / H[2, 2]


def jacobi_smallest_eigenvector(G, max_sweeps=30, eps=1e-17):
    """CPU restatement of the device's 9 x 9 symmetric eigen-solver (warp_jacobi9 in
    camera_calibration_b200/csrc/calib_linear.cuh): cyclic Jacobi rotations in the same pivot order with
    the same skip rule |a_pq| <= eps sqrt(|a_pp a_qq|) and the same rotation formulas; returns
    (eigenvector of the smallest eigenvalue, its eigenvalue, sweeps used).  Test infrastructure: it pins
    the convergence and accuracy of that scheme against LAPACK on the CPU."""
    A = np.array(G, dtype=np.float64)
    n = A.shape[0]
    V = np.eye(n)
    sweeps = 0
    for sweeps in range(1, max_sweeps + 1):
        rotated = False
        for p in range(n - 1):
            for q in range(p + 1, n):
                apq, app, aqq = A[p, q], A[p, p], A[q, q]
                if not abs(apq) > eps * np.sqrt(abs(app * aqq)):
                    continue
                rotated = True
                theta = (aqq - app) / (2.0 * apq)
                t = np.copysign(1.0, theta) / (abs(theta) + np.sqrt(theta * theta + 1.0))
                c = 1.0 / np.sqrt(t * t + 1.0)
                sn = t * c
                vp, vq = V[:, p].copy(), V[:, q].copy()
                V[:, p], V[:, q] = c * vp - sn * vq, sn * vp + c * vq
                akp, akq = A[:, p].copy(), A[:, q].copy()
                newp, newq = c * akp - sn * akq, sn * akp + c * akq
                for k in range(n):
                    if k != p and k != q:
                        A[k, p] = A[p, k] = newp[k]
                        A[k, q] = A[q, k] = newq[k]
                A[p, p], A[q, q] = app - t * apq, aqq + t * apq
                A[p, q] = A[q, p] = 0.0
        if not rotated:
            break
    j = int(np.argmin(np.diag(A)))
    return V[:, j], A[j, j], sweeps


def estimate_homography_jacobi(Xa, Xb):
    """estimate_homography with the eigenvector of M^T M taken by jacobi_smallest_eigenvector (what the
    device does) instead of the last right-singular vector of M (what the reference does)."""
    Na, Nb = normalization_matrix(Xa), normalization_matrix(Xb)
    xa = Xa * Na[0, 0] + Na[:2, 2]
    xb = Xb * Nb[0, 0] + Nb[:2, 2]
    M = dlt_rows(xa, xb)
    h, lam, sweeps = jacobi_smallest_eigenvector(M.T @ M)
    H = np.linalg.inv(Na) @ h.reshape(3, 3) @ Nb
    return H / H[2, 2], lam, sweeps


def _vec(H, p, q):
    return np.stack([H[:, 0, p] * H[:, 0, q],
                     H[:, 0, p] * H[:, 1, q] + H[:, 1, p] * H[:, 0, q],
                     H[:, 1, p] * H[:, 1, q],
                     H[:, 2, p] * H[:, 0, q] + H[:, 0, p] * H[:, 2, q],
                     H[:, 2, p] * H[:, 1, q] + H[:, 1, p] * H[:, 2, q],
                     H[:, 2, p] * H[:, 2, q]], axis=1)


def intrinsic_matrix(Hs):
    H = np.asarray(Hs, dtype=np.float64).reshape(-1, 3, 3)
    V = np.empty((2 * len(H), 6))
    V[0::2] = _vec(H, 0, 1)
    V[1::2] = _vec(H, 0, 0) - _vec(H, 1, 1)
    _, _, Vt = np.linalg.svd(V)
    B0, B1, B2, B3, B4, B5 = Vt[-1]
    sign = -1.0 if (B0 < 0 or B2 < 0 or B5 < 0) else 1.0
    B = sign * np.array([[B0, B1, B3], [B1, B2, B4], [B3, B4, B5]])
    A = np.linalg.inv(np.linalg.cholesky(B).T)
    return A / A[2, 2]


def extrinsics(Hs, A):
    H = np.asarray(Hs, dtype=np.float64).reshape(-1, 3, 3)
    Ainv = np.linalg.inv(A)
    G = Ainv @ H                                     # columns: lam r0, lam r1, lam t
    lam = np.linalg.norm(G[:, :, 0], axis=1)
    r0, r1, t = G[:, :, 0] / lam[:, None], G[:, :, 1] / lam[:, None], G[:, :, 2] / lam[:, None]
    Q = np.stack([r0, r1, np.cross(r0, r1)], axis=2)
    U, _, Vt = np.linalg.svd(Q)
    R = U @ Vt
    W = np.tile(np.eye(4), (len(H), 1, 1))
    W[:, :3, :3], W[:, :3, 3] = R, t
    return [W[i] for i in range(len(W))]


def distortion_rows(model, A, W, sensor, model_pts):
    """Rows (D (2N, nd), Ddot (2N,)) of one view, src/distortion.py:149-188 | 231-268."""
    fx, fy, uc, vc = A[0, 0], A[1, 1], A[0, 2], A[1, 2]
    Xc = model_pts @ W[:3, :3].T + W[:3, 3]
    xn, yn = Xc[:, 0] / Xc[:, 2], Xc[:, 1] / Xc[:, 2]
    r = np.sqrt(xn * xn + yn * yn)
    u = A[0, 0] * xn + A[0, 1] * yn + uc
    v = fy * yn + vc
    N = len(xn)
    if model == 0:
        D = np.empty((2 * N, 5))
        D[0::2] = np.stack([(u - uc) * r**2, (u - uc) * r**4, fx * (2 * xn * yn), fx * (r**2 + 2 * xn**2),
                            (u - uc) * r**6], axis=1)
        D[1::2] = np.stack([(v - vc) * r**2, (v - vc) * r**4, fy * (r**2 + 2 * yn**2), fy * (2 * xn * yn),
                            (v - vc) * r**6], axis=1)
    else:
        th = np.arctan(r)
        with np.errstate(invalid="ignore", divide="ignore"):
            q = th / r
        D = np.empty((2 * N, 4))
        for c in range(4):
            D[0::2, c] = fx * (u - uc) * q * th**(2 * c + 2)
            D[1::2, c] = fy * (v - vc) * q * th**(2 * c + 2)
    Ddot = np.empty(2 * N)
    Ddot[0::2], Ddot[1::2] = sensor[:, 0] - u, sensor[:, 1] - v
    return D, Ddot


def estimate_distortion(model, A, allDetections, Ws):
    rows = [distortion_rows(model, A, W, s, m) for (s, m), W in zip(allDetections, Ws)]
    D = np.vstack([r[0] for r in rows])
    Ddot = np.concatenate([r[1] for r in rows])
    return tuple((np.linalg.pinv(D) @ Ddot).ravel())


def estimate_calibration_parameters(model, allDetections, refine=True):
    """src/calibrate.py:41-58: DLT -> homography LM -> intrinsics -> extrinsics -> distortion."""
    Hs = [estimate_homography(s[:, :2], m[:, :2]) for s, m in allDetections]
    if refine:
        Hs = [onp.homography_lm(H, s, m)[0] for H, (s, m) in zip(Hs, allDetections)]
    A = intrinsic_matrix(Hs)
    Ws = extrinsics(Hs, A)
    k = estimate_distortion(model, A, allDetections, Ws)
    return A, Ws, k, Hs
