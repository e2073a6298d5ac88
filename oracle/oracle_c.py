"""TEST INFRASTRUCTURE ONLY -- ctypes binding of oracle/calib_oracle.c (the C restatement
of the reference's LM hot path; see that file's header for the reference file:line it
follows and its parity status).  Used by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs only."""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liboracle.so")
_SRC = os.path.join(_HERE, "calib_oracle.c")
_lib = None


def build(force=False):
    if not force and os.path.exists(_SO) and os.path.getmtime(_SO) >= os.path.getmtime(_SRC):
        return _SO
    last = None
    for cc in ("/usr/bin/gcc", "gcc", "cc"):
        for omp in (["-fopenmp"], []):
            cmd = [cc, "-O2", "-ffp-contract=off", *omp, "-shared", "-fPIC", "-o", _SO, _SRC, "-lm"]
            try:
                subprocess.run(cmd, check=True, capture_output=True)
                return _SO
            except (OSError, subprocess.CalledProcessError) as e:  # try the next recipe
                last = e
    raise RuntimeError("could not build the C oracle: %r" % (last,))


_d = ctypes.POINTER(ctypes.c_double)
_i64 = ctypes.POINTER(ctypes.c_int64)


def _p(a):
    return a.ctypes.data_as(_d)


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(_SO)
        for name in ("oracle_project", "oracle_jacobian_blocks", "oracle_cost",
                     "oracle_normal_equations", "oracle_lm", "oracle_lm_step", "oracle_num_threads"):
            getattr(_lib, name).restype = ctypes.c_int
    return _lib


def _prep(view_offsets, model_pts, sensor_pts=None):
    offs = np.ascontiguousarray(view_offsets, dtype=np.int64)
    mp = np.ascontiguousarray(model_pts, dtype=np.float64)
    sp = None if sensor_pts is None else np.ascontiguousarray(sensor_pts, dtype=np.float64)
    return offs, mp, sp


def nintr(model):
    return 10 if model == 0 else 9


def num_threads():
    return lib().oracle_num_threads()


def project(model, P, view_offsets, model_pts):
    offs, mp, _ = _prep(view_offsets, model_pts)
    P = np.ascontiguousarray(P, dtype=np.float64).ravel()
    y = np.empty((offs[-1], 2))
    lib().oracle_project(ctypes.c_int(model), ctypes.c_int64(len(offs) - 1),
                         offs.ctypes.data_as(_i64), _p(mp), _p(P), _p(y))
    return y


def jacobian_blocks(model, P, view_offsets, model_pts):
    offs, mp, _ = _prep(view_offsets, model_pts)
    P = np.ascontiguousarray(P, dtype=np.float64).ravel()
    L = nintr(model)
    N = int(offs[-1])
    Jc = np.empty((2 * N, L))
    Jv = np.empty((2 * N, 6))
    lib().oracle_jacobian_blocks(ctypes.c_int(model), ctypes.c_int64(len(offs) - 1),
                                 offs.ctypes.data_as(_i64), _p(mp), _p(P), _p(Jc), _p(Jv))
    return Jc, Jv


def cost(model, P, view_offsets, model_pts, sensor_pts):
    offs, mp, sp = _prep(view_offsets, model_pts, sensor_pts)
    P = np.ascontiguousarray(P, dtype=np.float64).ravel()
    E = ctypes.c_double()
    lib().oracle_cost(ctypes.c_int(model), ctypes.c_int64(len(offs) - 1), offs.ctypes.data_as(_i64),
                      _p(mp), _p(sp), _p(P), ctypes.byref(E))
    return E.value


def normal_equations(model, P, view_offsets, model_pts, sensor_pts):
    offs, mp, sp = _prep(view_offsets, model_pts, sensor_pts)
    P = np.ascontiguousarray(P, dtype=np.float64).ravel()
    L = nintr(model)
    M = len(offs) - 1
    A = np.empty((L, L)); B = np.empty((M, L, 6)); D = np.empty((M, 6, 6)); g = np.empty(L + 6 * M)
    E = ctypes.c_double()
    lib().oracle_normal_equations(ctypes.c_int(model), ctypes.c_int64(M), offs.ctypes.data_as(_i64),
                                  _p(mp), _p(sp), _p(P), _p(A), _p(B), _p(D), _p(g), ctypes.byref(E))
    return A, B, D, g, E.value


def lm_step(model, A, B, D, g, lam):
    M = B.shape[0]
    delta = np.empty(nintr(model) + 6 * M)
    lib().oracle_lm_step(ctypes.c_int(model), ctypes.c_int64(M), _p(np.ascontiguousarray(A)),
                         _p(np.ascontiguousarray(B)), _p(np.ascontiguousarray(D)),
                         _p(np.ascontiguousarray(g)), ctypes.c_double(lam), _p(delta))
    return delta


def lm(model, P0, view_offsets, model_pts, sensor_pts, maxIters, lam0=1e-3, lam_min=1e-10,
       lam_max=1e10, err_min=1e-12, nthreads=0):
    """Returns (sse, P, iters, trace[iters,4])."""
    offs, mp, sp = _prep(view_offsets, model_pts, sensor_pts)
    P = np.array(P0, dtype=np.float64).ravel().copy()
    trace = np.zeros((max(maxIters, 1), 4))
    sse = ctypes.c_double()
    iters = ctypes.c_int()
    rc = lib().oracle_lm(ctypes.c_int(model), ctypes.c_int64(len(offs) - 1), offs.ctypes.data_as(_i64),
                         _p(mp), _p(sp), _p(P), ctypes.c_int(maxIters), ctypes.c_double(lam0),
                         ctypes.c_double(lam_min), ctypes.c_double(lam_max), ctypes.c_double(err_min),
                         ctypes.byref(sse), ctypes.byref(iters), _p(trace), ctypes.c_int(nthreads))
    if rc != 0:
        raise UnboundLocalError("maxIters=0: the reference has no Pt_error to return")
    return sse.value, P, iters.value, trace[:iters.value].copy()
