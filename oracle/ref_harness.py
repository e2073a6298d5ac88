"""TEST INFRASTRUCTURE ONLY -- harness that imports the UNMODIFIED reference
(pvphan/camera-calibration, mounted read-only at /root/reference) in THIS container so
that golden vectors can be generated from it (oracle/make_golden.py).

Nothing in the product path, bench.py or the -m gpu tests may import this module: the
reference does not exist on the GPU box.  No reference file is edited; two harness-side
shims are needed because the reference pins numpy 1.21 / imageio (Dockerfile:8-12):

  * a stub ``imageio`` module (src/visualize.py:5 imports it at module scope);
  * ``src.jacobian.createLambdaFunction`` (jacobian.py:195-199) is replaced by a version
    whose lambdified 2xT matrix is assembled as a dtype=object array entry by entry,
    which is what numpy<=1.23 did with the ragged nested list that lambdify emits
    (numpy>=1.24 raises "inhomogeneous shape").  The arithmetic is unchanged.
"""
import os
import sys
import types
import warnings

import numpy as np

REF_ROOT = os.environ.get("CALIB_REF_ROOT", "/root/reference")


def available():
    return os.path.isdir(os.path.join(REF_ROOT, "src"))


def _obj_array(rows):
    out = np.empty((len(rows), len(rows[0])), dtype=object)
    for i, r in enumerate(rows):
        for j, v in enumerate(r):
            out[i, j] = v
    return out


_loaded = None


def load():
    """Returns the reference's ``src`` namespace package with the shims applied."""
    global _loaded
    if _loaded is not None:
        return _loaded
    if not available():
        raise RuntimeError("reference tree not present at %s" % REF_ROOT)
    sys.dont_write_bytecode = True
    if "imageio" not in sys.modules:
        stub = types.ModuleType("imageio")
        stub.imwrite = lambda *a, **k: None
        stub.mimsave = lambda *a, **k: None

        def _imread(*a, **k):
            raise RuntimeError("imageio stub")
        stub.imread = _imread
        sys.modules["imageio"] = stub
    tests_dir = os.path.join(REF_ROOT, "tests")
    if tests_dir not in sys.path:
        sys.path.insert(0, tests_dir)
    import sympy
    from __context__ import src  # noqa: F401  (tests/__context__.py adds REF_ROOT)
    from src import jacobian

    def createLambdaFunction(expression, orderedSymbols):
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            return sympy.lambdify(orderedSymbols, expression,
                                  [{"array": _obj_array}, "numpy"])
    jacobian.createLambdaFunction = createLambdaFunction
    from src import calibrate, checkerboard, dataset, distortion, mathutils  # noqa
    from src import noise, virtualcamera, main  # noqa
    _loaded = src
    return src
