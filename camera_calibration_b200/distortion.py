"""Distortion-model tags with the reference's class names (src/distortion.py:74-76,194-196).
On the GPU path the model is a template parameter of the kernels; these classes only carry
the identity and the coefficient names, plus a numpy forward projection used by the synthetic
data generator (src/distortion.py:42-59,78-108,198-220)."""
import numpy as np

RADTAN, FISHEYE = 0, 1


class DistortionModel:
    model_id = None
    _names = ()

    def getIntrinsicSymbols(self):
        return ("α", "β", "γ", "uc", "vc")

    def getDistortionSymbols(self):
        return self._names

    def distortPoints(self, x, k):
        raise NotImplementedError()

    def projectWithDistortion(self, A, X, k):
        """src/distortion.py:42-59.  X: (N,3) camera coordinates -> (N,2) pixels."""
        X = np.asarray(X, dtype=np.float64)
        x = X[:, :2] / X[:, 2:3]
        xd = self.distortPoints(x, k)
        Ap = np.asarray(A, dtype=np.float64)[:2, :3]
        return (Ap @ np.hstack((xd, np.ones((len(xd), 1)))).T).T


class RadialTangentialModel(DistortionModel):
    model_id = RADTAN
    _names = ("k1", "k2", "p1", "p2", "k3")

    def distortPoints(self, x, k):
        k1, k2, p1, p2, k3 = k
        xn, yn = x[:, 0], x[:, 1]
        r = np.linalg.norm(x, axis=1)
        radial = (1 + k1 * r**2 + k2 * r**4 + k3 * r**6)
        tx = (2 * p1 * xn * yn + p2 * (r**2 + 2 * xn**2))
        ty = (p1 * (r**2 + 2 * yn**2) + 2 * p2 * xn * yn)
        return np.stack((radial * xn + tx, radial * yn + ty), axis=1)


class FisheyeModel(DistortionModel):
    model_id = FISHEYE
    _names = ("k1", "k2", "k3", "k4")

    def distortPoints(self, x, k):
        k1, k2, k3, k4 = k
        xn, yn = x[:, 0], x[:, 1]
        r = np.linalg.norm(x, axis=1)
        th = np.arctan(r)
        with np.errstate(invalid="ignore", divide="ignore"):
            radial = (th / r) * (1 + k1 * th**2 + k2 * th**4 + k3 * th**6 + k4 * th**8)
        return np.stack((radial * xn, radial * yn), axis=1)


def model_id_of(distortionModel):
    """Accepts our tags, the reference's own model objects (duck-typed on the number of
    distortion symbols, src/distortion.py:76,196), or the strings main.calibrateCamera takes."""
    if isinstance(distortionModel, str):
        if distortionModel == "radtan":
            return RADTAN
        if distortionModel == "fisheye":
            return FISHEYE
        raise ValueError(f"Distortion type: {distortionModel} unknown")   # src/main.py:33
    if isinstance(distortionModel, int):
        if distortionModel in (RADTAN, FISHEYE):
            return distortionModel
        raise ValueError(f"Distortion type: {distortionModel} unknown")
    mid = getattr(distortionModel, "model_id", None)
    if mid in (RADTAN, FISHEYE):
        return mid
    n = len(distortionModel.getDistortionSymbols())
    if n == 5:
        return RADTAN
    if n == 4:
        return FISHEYE
    raise ValueError("unsupported distortion model with %d coefficients" % n)
