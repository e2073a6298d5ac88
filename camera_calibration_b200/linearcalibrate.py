"""Closed-form initial estimate of Zhang's method on the GPU handle:
host-side mirror of the reference's src/linearcalibrate.py and DistortionModel.estimateDistortion.

Everything that touches the points runs on the device, on the detections a `Problem` already holds
(calib_dlt_gram, calib_refine_homographies, calib_distortion_rows); this module is the O(M) small
dense algebra in between (numpy / LAPACK, as in the reference) -- there are no per-point Python loops
and no CPU fallback for the per-point parts.
"""
import ctypes

import numpy as np

from . import _lib
from ._lib import dptr


def _sym3(p):
    """(M, 6) packed (00, 01, 02, 11, 12, 22) -> (M, 3, 3) symmetric"""
    S = np.empty((len(p), 3, 3))
    S[:, 0, 0], S[:, 0, 1], S[:, 0, 2] = p[:, 0], p[:, 1], p[:, 2]
    S[:, 1, 0], S[:, 1, 1], S[:, 1, 2] = p[:, 1], p[:, 3], p[:, 4]
    S[:, 2, 0], S[:, 2, 1], S[:, 2, 2] = p[:, 2], p[:, 4], p[:, 5]
    return S


def estimateHomographies(prob):
    """linearcalibrate.estimateHomographies / estimateHomography (src/linearcalibrate.py:7-58) for all
    views of a Problem, entirely on the device (calib_dlt_homographies): point normalisations, the Gram
    matrix of the DLT system, its eigenvector of the smallest eigenvalue by Jacobi rotations (the
    reference: last right-singular vector of the 2N x 9 matrix; same vector up to sign, removed by
    H /= H[2,2]) and the de-normalisation.  Returns (M, 3, 3)."""
    H = np.zeros((prob.M, 9))
    prob._check(prob._L.calib_dlt_homographies(prob._h, dptr(H)))
    return H.reshape(prob.M, 3, 3)


def estimateHomographiesHostEig(prob):
    """The same with the 9 x 9 eigenvectors taken by LAPACK on the host from the device's Gram sums
    (calib_dlt_gram): the cross-check of the device's Jacobi solver in the tests."""
    M = prob.M
    gram = np.zeros((M, 24))
    norm = np.zeros((M, 6))
    prob._check(prob._L.calib_dlt_gram(prob._h, dptr(gram), dptr(norm)))
    P1, Pu, Pv, Pq = (_sym3(gram[:, 6 * k:6 * k + 6]) for k in range(4))
    G = np.zeros((M, 9, 9))
    G[:, 0:3, 0:3] = P1
    G[:, 3:6, 3:6] = P1
    G[:, 0:3, 6:9] = -Pu
    G[:, 6:9, 0:3] = -Pu
    G[:, 3:6, 6:9] = -Pv
    G[:, 6:9, 3:6] = -Pv
    G[:, 6:9, 6:9] = Pq
    _, vecs = np.linalg.eigh(G)                      # ascending eigenvalues
    Hp = vecs[:, :, 0].reshape(M, 3, 3)
    mu, mv, sa, mx, my, sb = (norm[:, k] for k in range(6))
    Na_inv = np.tile(np.eye(3), (M, 1, 1))           # inverse of [[s, 0, -s m0], [0, s, -s m1], [0, 0, 1]]
    Na_inv[:, 0, 0] = Na_inv[:, 1, 1] = 1.0 / sa
    Na_inv[:, 0, 2], Na_inv[:, 1, 2] = mu, mv
    Nb = np.tile(np.eye(3), (M, 1, 1))
    Nb[:, 0, 0] = Nb[:, 1, 1] = sb
    Nb[:, 0, 2], Nb[:, 1, 2] = -sb * mx, -sb * my
    H = Na_inv @ Hp @ Nb
    return H / H[:, 2:3, 2:3]


def _vec(H, p, q):
    """vecHomography, src/linearcalibrate.py:158-185, for all homographies at once: (M, 6)"""
    return np.stack([H[:, 0, p] * H[:, 0, q],
                     H[:, 0, p] * H[:, 1, q] + H[:, 1, p] * H[:, 0, q],
                     H[:, 1, p] * H[:, 1, q],
                     H[:, 2, p] * H[:, 0, q] + H[:, 0, p] * H[:, 2, q],
                     H[:, 2, p] * H[:, 1, q] + H[:, 1, p] * H[:, 2, q],
                     H[:, 2, p] * H[:, 2, q]], axis=1)


def computeIntrinsicMatrix(Hs):
    """src/linearcalibrate.py:94-155 with computeIntrinsicMatrixFrombCholesky (:256-292)."""
    H = np.asarray(Hs, dtype=np.float64).reshape(-1, 3, 3)
    V = np.empty((2 * len(H), 6))
    V[0::2] = _vec(H, 0, 1)
    V[1::2] = _vec(H, 0, 0) - _vec(H, 1, 1)
    b = np.linalg.svd(V, full_matrices=False)[2][-1]      # 2M >= 6: the same V^T as the full SVD
    sign = -1.0 if (b[0] < 0 or b[2] < 0 or b[5] < 0) else 1.0
    B = sign * np.array([[b[0], b[1], b[3]], [b[1], b[2], b[4]], [b[3], b[4], b[5]]])
    A = np.linalg.inv(np.linalg.cholesky(B).T)
    A /= A[2, 2]
    if np.isnan(A).any():
        raise ValueError(f"Computed intrinsic matrix contains NaN: \n{A}")
    return A


def computeExtrinsics(Hs, A):
    """src/linearcalibrate.py:295-325 with approximateRotationMatrix (:328-372), all views at once.
    Returns the reference's list of (4,4) world-to-camera transforms."""
    H = np.asarray(Hs, dtype=np.float64).reshape(-1, 3, 3)
    G = np.linalg.inv(A) @ H
    lam = np.linalg.norm(G[:, :, 0], axis=1)[:, None]
    r0, r1, t = G[:, :, 0] / lam, G[:, :, 1] / lam, G[:, :, 2] / lam
    Q = np.stack([r0, r1, np.cross(r0, r1)], axis=2)
    U, _, Vt = np.linalg.svd(Q)
    W = np.tile(np.eye(4), (len(H), 1, 1))
    W[:, :3, :3] = U @ Vt
    W[:, :3, 3] = t
    return [W[i] for i in range(len(W))]


def estimateDistortion(prob, A, allBoardPosesInCamera):
    """DistortionModel.estimateDistortion (src/distortion.py:110-191 | 222-271): the device forms the
    rows of D k = Ddot and accumulates D^T D and D^T Ddot per view; k solves the summed nd x nd system
    (the reference: pinv(D) @ Ddot, the same least-squares solution)."""
    M = prob.M
    nd = 5 if prob.model == 0 else 4
    W = np.asarray(allBoardPosesInCamera, dtype=np.float64).reshape(M, 4, 4)
    W12 = np.ascontiguousarray(np.concatenate([W[:, :3, :3].reshape(M, 9), W[:, :3, 3]], axis=1))
    A9 = np.ascontiguousarray(np.asarray(A, dtype=np.float64).reshape(9))
    rows = np.zeros((M, 20))
    prob._check(prob._L.calib_distortion_rows(prob._h, dptr(A9), dptr(W12), dptr(rows)))
    tot = rows.sum(axis=0)
    nt = nd * (nd + 1) // 2
    DtD = np.zeros((nd, nd))
    DtD[np.triu_indices(nd)] = tot[:nt]
    DtD = DtD + np.triu(DtD, 1).T
    return tuple(np.linalg.solve(DtD, tot[nt:nt + nd]))
