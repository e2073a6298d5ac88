"""Host-side helpers with the reference's names and conventions (src/mathutils.py), only the
ones the parameter-vector pack/unpack of the hot path needs.  Euler angles are in DEGREES and
R = Rz @ Ry @ Rx (src/mathutils.py:36-51)."""
import numpy as np


def col(v):
    """src/mathutils.py:54-56"""
    return np.array(v).reshape(-1, 1)


def _axisRotation(angleRadians, axis):
    """Rodrigues' formula for a rotation about a unit coordinate axis, evaluated the way the
    reference's numeric branch does (src/mathutils.py:59-81: I + K/|w| sin|w| + K.K/|w|^2 (1-cos|w|)
    with K = angle * skew(axis)), so that poses are bit-identical to the reference's."""
    K = np.zeros((3, 3))
    i, j = [(1, 2), (2, 0), (0, 1)][axis]
    K[j, i] = angleRadians
    K[i, j] = -angleRadians
    wNorm = np.linalg.norm(np.array([K[2, 1], K[0, 2], K[1, 0]]))
    if np.isclose(wNorm, 0):
        return np.eye(3)
    return np.eye(3) + (K / wNorm) * np.sin(wNorm) + ((K @ K) / (wNorm**2)) * (1 - np.cos(wNorm))


def eulerToRotationMatrix(rXrYrZDegrees):
    """src/mathutils.py:36-51 (numeric branch): R = Rz @ Ry @ Rx, angles in degrees."""
    rx, ry, rz = rXrYrZDegrees
    Rx = _axisRotation(np.radians(rx), 0)
    Ry = _axisRotation(np.radians(ry), 1)
    Rz = _axisRotation(np.radians(rz), 2)
    return Rz @ Ry @ Rx


def rotationMatrixToEuler(R):
    """src/mathutils.py:13-33 (Slabaugh), returns degrees (psi, theta, phi) = (rx, ry, rz)."""
    R = np.asarray(R, dtype=np.float64)
    R11, R12, R13, R21, R22, R23, R31, R32, R33 = R.ravel()
    if not (np.isclose(R31, +1) or np.isclose(R31, -1)):
        th = -np.arcsin(R31)
        psi = np.arctan2(R32 / np.cos(th), R33 / np.cos(th))
        phi = np.arctan2(R21 / np.cos(th), R11 / np.cos(th))
    else:
        phi = 0
        if np.isclose(R31, -1):
            th = np.pi / 2
            psi = phi + np.arctan2(R12, R13)
        else:
            th = -np.pi / 2
            psi = -phi + np.arctan2(-R12, -R13)
    return tuple(np.degrees((psi, th, phi)))


def poseFromRT(R, T):
    """src/mathutils.py:144-150"""
    M = np.eye(4)
    M[:3, :3] = R
    if isinstance(T, np.ndarray):
        T = T.ravel()
    M[:3, 3] = T
    return M


def hom(v):
    """src/mathutils.py:120-127 for (N,d) arrays"""
    v = np.asarray(v)
    if v.ndim == 1:
        return np.append(v, 1)
    return np.hstack((v, np.ones((v.shape[0], 1))))


def unhom(vHom):
    """src/mathutils.py:130-137"""
    if vHom.ndim == 1:
        return vHom[:-1] / vHom[-1]
    return vHom[:, :-1] / col(vHom[:, -1])


def transform(b_M_a, aX):
    """src/mathutils.py:195-208"""
    return unhom((b_M_a @ hom(aX).T).T)


# ---- batched forms (same arithmetic, one numpy call for all views) --------------------------
def eulerToRotationMatrixBatch(eulerDegrees):
    """(M,3) Euler angles in degrees -> (M,3,3), R = Rz @ Ry @ Rx per view
    (src/mathutils.py:36-51).  Closed form of the three axis rotations; agrees with the scalar
    Rodrigues evaluation above to rounding.  (Angles transposed to three contiguous rows first: two trig
    calls over 3M contiguous values instead of six over strided columns.)"""
    e = np.radians(np.ascontiguousarray(np.asarray(eulerDegrees, dtype=np.float64).reshape(-1, 3).T))
    s, c = np.sin(e), np.cos(e)
    sa, sb, sg = s
    ca, cb, cg = c
    R = np.empty((9, e.shape[1]))
    cgsb, sgsb = cg * sb, sg * sb
    np.multiply(cg, cb, out=R[0])
    R[1] = cgsb * sa - sg * ca
    R[2] = cgsb * ca + sg * sa
    np.multiply(sg, cb, out=R[3])
    R[4] = sgsb * sa + cg * ca
    R[5] = sgsb * ca - cg * sa
    np.negative(sb, out=R[6])
    np.multiply(cb, sa, out=R[7])
    np.multiply(cb, ca, out=R[8])
    return R.T.reshape(-1, 3, 3)


def rotationMatrixToEulerBatch(R):
    """(M,3,3) -> (M,3) degrees, the branches of src/mathutils.py:13-33 applied per view."""
    R = np.asarray(R, dtype=np.float64).reshape(-1, 3, 3)
    R31 = R[:, 2, 0]
    # np.isclose(R31, +-1) with its default tolerances: |R31 -+ 1| <= 1e-8 + 1e-5 * 1
    reg = np.abs(np.abs(R31) - 1.0) > (1e-8 + 1e-5)
    out = np.empty((len(R), 3))
    with np.errstate(invalid="ignore", divide="ignore"):
        th = -np.arcsin(np.clip(R31, -1, 1))
        c = np.cos(th)
        psi = np.arctan2(R[:, 2, 1] / c, R[:, 2, 2] / c)
        phi = np.arctan2(R[:, 1, 0] / c, R[:, 0, 0] / c)
    out[:, 0], out[:, 1], out[:, 2] = psi, th, phi
    for i in np.nonzero(~reg)[0]:          # gimbal lock: rare, scalar path
        out[i] = np.radians(rotationMatrixToEuler(R[i]))
    return np.degrees(out)
