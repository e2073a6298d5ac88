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
    Rodrigues evaluation above to rounding."""
    e = np.radians(np.asarray(eulerDegrees, dtype=np.float64).reshape(-1, 3))
    sa, ca = np.sin(e[:, 0]), np.cos(e[:, 0])
    sb, cb = np.sin(e[:, 1]), np.cos(e[:, 1])
    sg, cg = np.sin(e[:, 2]), np.cos(e[:, 2])
    R = np.empty((len(e), 3, 3))
    R[:, 0, 0] = cg * cb
    R[:, 0, 1] = cg * sb * sa - sg * ca
    R[:, 0, 2] = cg * sb * ca + sg * sa
    R[:, 1, 0] = sg * cb
    R[:, 1, 1] = sg * sb * sa + cg * ca
    R[:, 1, 2] = sg * sb * ca - cg * sa
    R[:, 2, 0] = -sb
    R[:, 2, 1] = cb * sa
    R[:, 2, 2] = cb * ca
    return R


def rotationMatrixToEulerBatch(R):
    """(M,3,3) -> (M,3) degrees, the branches of src/mathutils.py:13-33 applied per view."""
    R = np.asarray(R, dtype=np.float64).reshape(-1, 3, 3)
    R31 = R[:, 2, 0]
    reg = ~(np.isclose(R31, +1) | np.isclose(R31, -1))
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
