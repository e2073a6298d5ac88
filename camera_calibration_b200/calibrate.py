"""Host-side mirror of the reference's interface for the LM refinement path.

  Calibrator.refineCalibrationParameters   src/calibrate.py:117-171   (same signature / returns)
  Calibrator.projectAllPoints              src/calibrate.py:190-197
  Calibrator._compose/_decomposeParameterVector   src/calibrate.py:199-267
  ProjectionJacobian.compute               src/jacobian.py:48-85      (parity only: dense J)
  getSensorPoints                          src/calibrate.py:277-282

Everything numeric runs in libcalib_b200.so on the GPU; there is no CPU fallback here.
`Problem` is the thin object-oriented face of one C-ABI handle."""
import ctypes
import time

import numpy as np

from . import _lib
from . import distortion
from . import mathutils as mu
from ._lib import CalibError, as_f64, dptr


def flatten_detections(allDetections, out=None):
    """list of (sensorPoints (Ni,2), modelPoints (Ni,3)) -> CSR (view_offsets, model, sensor).
    out = (model (>=N,3), sensor (>=N,2)) float64 buffers to fill (e.g. page-locked ones)."""
    counts = np.fromiter((len(m) for s, m in allDetections), dtype=np.int64, count=len(allDetections))
    offs = np.zeros(len(counts) + 1, dtype=np.int64)
    np.cumsum(counts, out=offs[1:])
    N = int(offs[-1])
    if out is None:
        model = np.empty((N, 3), dtype=np.float64)
        sensor = np.empty((N, 2), dtype=np.float64)
    else:
        model, sensor = out[0][:N], out[1][:N]
    if N:
        if N >= 100000 and _gather_fast(allDetections, counts, model, sensor):
            return offs, model, sensor
        try:
            np.concatenate([m for s, m in allDetections], axis=0, out=model)
            np.concatenate([s for s, m in allDetections], axis=0, out=sensor)
        except (ValueError, TypeError):      # odd shapes / dtypes / lists: the careful way
            for i, (s, m) in enumerate(allDetections):
                model[offs[i]:offs[i + 1]] = np.asarray(m, dtype=np.float64).reshape(-1, 3)
                sensor[offs[i]:offs[i + 1]] = np.asarray(s, dtype=np.float64).reshape(-1, 2)
    return offs, model, sensor


def _gather_fast(allDetections, counts, model, sensor):
    """Large inputs: threaded gather in the library (calib_host_gather) when every per-view array is a
    C-contiguous float64 array of the expected shape; False = let numpy do it."""
    f8 = np.dtype(np.float64)
    M = len(allDetections)
    pm = np.empty(M, dtype=np.uint64)
    ps = np.empty(M, dtype=np.uint64)
    for i, (s, m) in enumerate(allDetections):
        if not (type(m) is np.ndarray and type(s) is np.ndarray and m.dtype == f8 and s.dtype == f8
                and m.ndim == 2 and s.ndim == 2 and m.shape[1] == 3 and s.shape[1] == 2 and len(s) == len(m)
                and m.flags.c_contiguous and s.flags.c_contiguous):
            return False
        pm[i] = m.__array_interface__["data"][0]
        ps[i] = s.__array_interface__["data"][0]
    try:
        L = _lib.lib()
    except CalibError:
        return False
    i64 = ctypes.POINTER(ctypes.c_int64)
    bm, bs = counts * 24, counts * 16
    ok = L.calib_host_gather(model.ctypes.data, pm.ctypes.data, bm.ctypes.data_as(i64), M, 0) == 0
    return ok and L.calib_host_gather(sensor.ctypes.data, ps.ctypes.data, bs.ctypes.data_as(i64), M, 0) == 0


_ptr_cache = {"key": None, "views": None, "out": None}
_helper_pool = None


def _helper():
    """one helper thread per process for work that overlaps a library call (created on first use)"""
    global _helper_pool
    if _helper_pool is None:
        from concurrent.futures import ThreadPoolExecutor
        _helper_pool = ThreadPoolExecutor(max_workers=1, thread_name_prefix="calib-b200")
    return _helper_pool


def _view_pointers(allDetections):
    """(counts, model pointers, sensor pointers) of a list of per-view (sensor (n,2), model (n,3))
    C-contiguous float64 arrays, or None when any view is something else (or there are no views).
    Reading 2 M array addresses in Python costs ~0.5 us each (1 ms at 1000 views); the table of the last list
    is kept together with references to its arrays and reused when the same list holds the same array
    objects (an identity check per array: ~8x cheaper)."""
    M = len(allDetections)
    if M == 0:
        return None
    c = _ptr_cache
    if c["key"] == (id(allDetections), M):
        same = True
        for det, (cs, cm) in zip(allDetections, c["views"]):
            if det[0] is not cs or det[1] is not cm:
                same = False
                break
        if same:
            return c["out"]
    out = _view_pointers_walk(allDetections, M)
    if out is not None:
        c["key"], c["views"], c["out"] = (id(allDetections), M), [(s, m) for s, m in allDetections], out
    return out


def _view_pointers_walk(allDetections, M):
    f8 = np.dtype(np.float64)
    addressof, byte = ctypes.addressof, ctypes.c_char
    counts, pm, ps = [0] * M, [0] * M, [0] * M
    for i, (s, m) in enumerate(allDetections):
        if not (type(m) is np.ndarray and type(s) is np.ndarray and m.dtype == f8 and s.dtype == f8
                and m.ndim == 2 and s.ndim == 2 and m.shape[1] == 3 and s.shape[1] == 2 and len(s) == len(m)
                and m.flags.c_contiguous and s.flags.c_contiguous):
            return None
        n = len(m)
        counts[i] = n
        if n:
            try:                                   # ~4x cheaper than __array_interface__ / .ctypes.data
                pm[i] = addressof(byte.from_buffer(m))
                ps[i] = addressof(byte.from_buffer(s))
            except (TypeError, ValueError):        # read-only arrays do not export a writable buffer
                pm[i] = m.__array_interface__["data"][0]
                ps[i] = s.__array_interface__["data"][0]
    return (np.array(counts, dtype=np.int64), np.array(pm, dtype=np.uint64), np.array(ps, dtype=np.uint64))


class PinnedBuffers:
    """Grow-only page-locked staging for the detections of one Calibrator (calib_host_alloc)."""

    def __init__(self):
        self._ptr = None
        self._cap = 0
        self.model = self.sensor = None

    def reserve(self, n_points):
        if n_points <= self._cap:
            return self.model, self.sensor
        self.release()
        L = _lib.lib()
        cap = max(int(n_points * 1.25), 1024)
        ptr = L.calib_host_alloc(cap * 5 * 8)
        if not ptr:
            raise CalibError(-2, "calib_host_alloc failed")
        buf = (ctypes.c_double * (cap * 5)).from_address(ptr)
        flat = np.frombuffer(buf, dtype=np.float64)
        self._ptr, self._cap = ptr, cap
        self.model = flat[:3 * cap].reshape(cap, 3)
        self.sensor = flat[3 * cap:].reshape(cap, 2)
        return self.model, self.sensor

    def release(self):
        if self._ptr:
            self.model = self.sensor = None
            _lib.lib().calib_host_free(self._ptr)
            self._ptr, self._cap = None, 0

    def __del__(self):
        try:
            self.release()
        except Exception:
            pass


def getSensorPoints(allDetections):
    """src/calibrate.py:277-282"""
    return flatten_detections(allDetections)[2]


class Problem:
    """One calibration problem resident on one GPU (a calib_t handle)."""

    def __init__(self, model, view_offsets, model_pts, sensor_pts, device=0):
        self.model = distortion.model_id_of(model)
        self.L = 10 if self.model == distortion.RADTAN else 9
        self.view_offsets = np.ascontiguousarray(view_offsets, dtype=np.int64)
        self.M = len(self.view_offsets) - 1
        self.N = int(self.view_offsets[-1]) if self.M >= 0 else 0
        self.K = self.L + 6 * self.M
        mp = as_f64(model_pts).reshape(-1, 3)
        sp = as_f64(sensor_pts).reshape(-1, 2)
        if len(mp) != self.N or len(sp) != self.N:
            raise ValueError("view_offsets[-1] must equal the number of points")
        self._L = _lib.lib()
        self._h = ctypes.c_void_p()
        rc = self._L.calib_create(ctypes.byref(self._h), self.model, self.M,
                                  self.view_offsets.ctypes.data_as(ctypes.POINTER(ctypes.c_int64)),
                                  dptr(mp), dptr(sp), int(device))
        if rc != 0:
            msg = self._L.calib_last_error(None)
            self._h = None
            raise CalibError(rc, msg.decode() if msg else "calib_create failed")

    @classmethod
    def from_detections(cls, model, allDetections, device=0, pinned=None):
        """From the reference's own container (a list of per-view (sensorPoints, modelPoints) arrays).
        When every array is a C-contiguous float64 array the library gathers and uploads the views itself
        (calib_create_from_views: threaded gather into its page-locked staging, pipelined with the copies);
        anything else (lists, other dtypes, strided views) is concatenated by numpy first, into the
        caller's PinnedBuffers when given."""
        vp = _view_pointers(allDetections)
        if vp is None:
            out = None
            if pinned is not None:
                out = pinned.reserve(sum(len(m) for s, m in allDetections))
            offs, mp, sp = flatten_detections(allDetections, out=out)
            return cls(model, offs, mp, sp, device)
        counts, pm, ps = vp
        self = cls.__new__(cls)
        self.model = distortion.model_id_of(model)
        self.L = 10 if self.model == distortion.RADTAN else 9
        self.view_offsets = np.zeros(len(counts) + 1, dtype=np.int64)
        np.cumsum(counts, out=self.view_offsets[1:])
        self.M = len(counts)
        self.N = int(self.view_offsets[-1])
        self.K = self.L + 6 * self.M
        self._L = _lib.lib()
        self._h = ctypes.c_void_p()
        rc = self._L.calib_create_from_views(ctypes.byref(self._h), self.model, self.M,
                                             counts.ctypes.data_as(ctypes.POINTER(ctypes.c_int64)),
                                             pm.ctypes.data, ps.ctypes.data, int(device))
        if rc != 0:
            msg = self._L.calib_last_error(None)
            self._h = None
            raise CalibError(rc, msg.decode() if msg else "calib_create_from_views failed")
        return self

    def close(self):
        if getattr(self, "_h", None):
            self._L.calib_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc != 0:
            msg = self._L.calib_last_error(self._h)
            raise CalibError(rc, msg.decode() if msg else "")

    def _P(self, P):
        P = as_f64(P).ravel()
        if P.size != self.K:
            raise ValueError(f"Expected a parameter vector of {self.K} values, got {P.size}")
        return P

    def set_segment(self, n):
        self._check(self._L.calib_set_segment(self._h, int(n)))

    def set_ladder(self, max_candidates):
        """damping ladder: trial steps taken ahead of the accept/reject decisions (1 = off, 2, 4)"""
        self._check(self._L.calib_set_ladder(self._h, int(max_candidates)))

    def set_ladder_start(self, candidates):
        """candidates of the pass that follows an accepted step (1, 2, or 0 = adaptive, the default)"""
        self._check(self._L.calib_set_ladder_start(self._h, int(candidates)))

    def set_loop_mode(self, mode):
        """0 auto, 1 host-driven iteration graphs, 2 device-driven WHILE graph."""
        self._check(self._L.calib_set_loop_mode(self._h, int(mode)))

    def project(self, P):
        P = self._P(P)
        y = np.empty((self.N, 2))
        self._check(self._L.calib_project(self._h, dptr(P), dptr(y)))
        return y

    def cost(self, P):
        P = self._P(P)
        E = ctypes.c_double()
        self._check(self._L.calib_cost(self._h, dptr(P), ctypes.cast(ctypes.byref(E), _lib._d)))
        return E.value

    def residual_jacobian(self, P):
        P = self._P(P)
        r = np.empty(2 * self.N)
        Jc = np.empty((2 * self.N, self.L))
        Jv = np.empty((2 * self.N, 6))
        self._check(self._L.calib_eval_residual_jacobian(self._h, dptr(P), dptr(r), dptr(Jc), dptr(Jv)))
        return r, Jc, Jv

    def normal_equations(self, P):
        P = self._P(P)
        A = np.empty((self.L, self.L))
        B = np.empty((self.M, self.L, 6))
        D = np.empty((self.M, 6, 6))
        g = np.empty(self.K)
        E = ctypes.c_double()
        self._check(self._L.calib_eval_normal_equations(
            self._h, dptr(P), dptr(A), dptr(B), dptr(D), dptr(g), ctypes.cast(ctypes.byref(E), _lib._d)))
        return A, B, D, g, E.value

    def lm_step(self, P, lam):
        P = self._P(P)
        delta = np.empty(self.K)
        self._check(self._L.calib_lm_step(self._h, dptr(P), float(lam), dptr(delta)))
        return delta

    def lm_run(self, P0, maxIters, lam0=1e-3, lam_min=1e-10, lam_max=1e10, err_min=1e-12,
               want_trace=True):
        """Returns (sse, P, iters, trace[iters,4])."""
        P = self._P(P0).copy()
        sse = ctypes.c_double()
        iters = ctypes.c_int()
        trace = np.zeros((max(int(maxIters), 1), 4)) if want_trace else None
        self._check(self._L.calib_lm_run(
            self._h, dptr(P), int(maxIters), lam0, lam_min, lam_max, err_min,
            ctypes.cast(ctypes.byref(sse), _lib._d), ctypes.byref(iters),
            dptr(trace) if want_trace else None))
        return sse.value, P, iters.value, (trace[:iters.value].copy() if want_trace else None)

    # resident variants (bench: device part alone)
    def upload(self, P):
        self._check(self._L.calib_upload_params(self._h, dptr(self._P(P))))

    def lm_run_resident(self, maxIters, lam0=1e-3, lam_min=1e-10, lam_max=1e10, err_min=1e-12):
        sse = ctypes.c_double()
        iters = ctypes.c_int()
        self._check(self._L.calib_lm_run_resident(
            self._h, int(maxIters), lam0, lam_min, lam_max, err_min,
            ctypes.cast(ctypes.byref(sse), _lib._d), ctypes.byref(iters)))
        return sse.value, iters.value

    def download(self):
        P = np.empty(self.K)
        self._check(self._L.calib_download_params(self._h, dptr(P)))
        return P

    def download_trace(self, iters):
        t = np.empty((iters, 4))
        self._check(self._L.calib_download_trace(self._h, int(iters), dptr(t)))
        return t

    def set_param_trace(self, on):
        """record A and k of the iterate beside every trace row (what the reference prints per iteration)"""
        self._check(self._L.calib_set_param_trace(self._h, 1 if on else 0))

    def param_trace(self, iters):
        """(iters, L): shared parameters of the iterate each iteration started from (after set_param_trace)"""
        out = np.empty((int(iters), self.L))
        self._check(self._L.calib_download_param_trace(self._h, int(iters), dptr(out)))
        return out

    def last_run_stats(self):
        ms = ctypes.c_double()
        nl = ctypes.c_int64()
        ne = ctypes.c_int64()
        self._check(self._L.calib_last_run_stats(self._h, ctypes.cast(ctypes.byref(ms), _lib._d),
                                                 ctypes.byref(nl), ctypes.byref(ne)))
        return ms.value, nl.value, ne.value

    def set_profiling(self, on):
        self._check(self._L.calib_set_profiling(self._h, 1 if on else 0))

    def kernel_time(self, which):
        ms = ctypes.c_double()
        n = ctypes.c_int64()
        self._check(self._L.calib_kernel_time(self._h, int(which), ctypes.cast(ctypes.byref(ms), _lib._d),
                                              ctypes.byref(n)))
        return ms.value, n.value

    def set_cost_staging(self, mode):
        """-1 auto, 0 plain loads, 1 bulk-copy (TMA) ring for the trial-cost sweep"""
        self._check(self._L.calib_set_cost_staging(self._h, int(mode)))

    def time_kernel(self, P, which, reps=20, lam=1e-3, flush=False):
        """average device microseconds of one kernel (0 eval, 1 schur, 2 backsub, 3 cost)"""
        us = ctypes.c_double()
        self._check(self._L.calib_time_kernel(self._h, dptr(self._P(P)), float(lam), int(which), int(reps),
                                              1 if flush else 0, ctypes.cast(ctypes.byref(us), _lib._d)))
        return us.value

    def flush_l2(self):
        self._check(self._L.calib_flush_l2(self._h))

    def comm_init(self, nranks, rank, id128: bytes):
        buf = ctypes.create_string_buffer(id128, 128)
        self._check(self._L.calib_comm_init(self._h, int(nranks), int(rank), ctypes.cast(buf, ctypes.c_void_p)))

    def comm_share(self, owner):
        self._check(self._L.calib_comm_share(self._h, owner._h))

    def refine_homographies(self, H, maxIters=20, lam0=1e-3, lam_min=1e-10, lam_max=1e10, err_min=1e-12,
                            want_trace=False):
        """Calibrator._refineHomographies on the handle's detections (src/calibrate.py:60-115).
        H: (M, 9) | (M, 3, 3) initial homographies -> (Href (M, 3, 3), iters (M,), trace | None)"""
        Hb = np.array(H, dtype=np.float64).reshape(-1, 9).copy()
        M = self.M
        if Hb.shape[0] != M:
            raise ValueError("expected %d homographies, got %d" % (M, Hb.shape[0]))
        iters = np.zeros(M, dtype=np.int32)
        trace = np.zeros((M, int(maxIters), 4)) if want_trace else None
        self._check(self._L.calib_refine_homographies(
            self._h, dptr(Hb), int(maxIters), float(lam0), float(lam_min), float(lam_max), float(err_min),
            iters.ctypes.data_as(ctypes.POINTER(ctypes.c_int)), dptr(trace) if want_trace else None))
        return Hb.reshape(M, 3, 3), iters, trace

    def comm_set_default(self, device, on=True):
        """handles created on `device` from now on borrow this handle's communicators"""
        self._check(self._L.calib_comm_set_default(self._h if on else None, int(device)))

    def peer_export(self) -> bytes:
        """64-byte CUDA IPC handle of this rank's inbox for the in-kernel all-reduce"""
        buf = ctypes.create_string_buffer(64)
        self._check(self._L.calib_peer_export(self._h, ctypes.cast(buf, ctypes.c_void_p)))
        return buf.raw

    def peer_attach(self, nranks, rank, handles: bytes):
        assert len(handles) == 64 * nranks
        buf = ctypes.create_string_buffer(handles, len(handles))
        self._check(self._L.calib_peer_attach(self._h, int(nranks), int(rank), ctypes.cast(buf, ctypes.c_void_p)))

    def set_transport(self, mode):
        """0 automatic, 1 NCCL all-reduces, 2 in-kernel exchange over NVLink peer memory"""
        self._check(self._L.calib_set_transport(self._h, int(mode)))

    def transport(self):
        return int(self._L.calib_transport(self._h))


class MultiProblem:
    """One calibration problem whose views are sharded over several GPUs of THIS process (a calib_multi_t
    handle): the library cuts the views into contiguous ranges balanced by point count, uploads every range
    to its device and sums the reduced system over the devices inside the kernels (NVLink peer memory).
    Takes and returns the full-length parameter vector, like Problem."""

    def __init__(self, model, view_offsets, model_pts, sensor_pts, devices):
        self._init_common(model, np.ascontiguousarray(view_offsets, dtype=np.int64))
        mp = as_f64(model_pts).reshape(-1, 3)
        sp = as_f64(sensor_pts).reshape(-1, 2)
        if len(mp) != self.N or len(sp) != self.N:
            raise ValueError("view_offsets[-1] must equal the number of points")
        dev = (ctypes.c_int * len(devices))(*[int(x) for x in devices])
        rc = self._L.calib_multi_create(ctypes.byref(self._h), self.model, self.M,
                                        self.view_offsets.ctypes.data_as(ctypes.POINTER(ctypes.c_int64)),
                                        dptr(mp), dptr(sp), len(devices), dev)
        self._created(rc)

    def _init_common(self, model, view_offsets):
        self.model = distortion.model_id_of(model)
        self.L = 10 if self.model == distortion.RADTAN else 9
        self.view_offsets = view_offsets
        self.M = len(view_offsets) - 1
        self.N = int(view_offsets[-1])
        self.K = self.L + 6 * self.M
        self._L = _lib.lib()
        self._h = ctypes.c_void_p()
        self.last_device_ms = 0.0

    def _created(self, rc):
        if rc != 0:
            msg = self._L.calib_multi_last_error(None)
            self._h = None
            raise CalibError(rc, msg.decode() if msg else "calib_multi_create failed")
        self.n_gpus = int(self._L.calib_multi_num_gpus(self._h))

    @classmethod
    def from_detections(cls, model, allDetections, devices, pinned=None):
        """From the reference's own container; see Problem.from_detections."""
        vp = _view_pointers(allDetections)
        if vp is None:
            out = None
            if pinned is not None:
                out = pinned.reserve(sum(len(m) for s, m in allDetections))
            offs, mp, sp = flatten_detections(allDetections, out=out)
            return cls(model, offs, mp, sp, devices)
        counts, pm, ps = vp
        self = cls.__new__(cls)
        offs = np.zeros(len(counts) + 1, dtype=np.int64)
        np.cumsum(counts, out=offs[1:])
        self._init_common(model, offs)
        dev = (ctypes.c_int * len(devices))(*[int(x) for x in devices])
        rc = self._L.calib_multi_create_from_views(ctypes.byref(self._h), self.model, self.M,
                                                   counts.ctypes.data_as(ctypes.POINTER(ctypes.c_int64)),
                                                   pm.ctypes.data, ps.ctypes.data, len(devices), dev)
        self._created(rc)
        return self

    def _check(self, rc):
        if rc != 0:
            msg = self._L.calib_multi_last_error(self._h)
            raise CalibError(rc, msg.decode() if msg else "")

    def view_bounds(self):
        b = np.zeros(self.n_gpus + 1, dtype=np.int64)
        self._check(self._L.calib_multi_view_bounds(self._h, b.ctypes.data_as(ctypes.POINTER(ctypes.c_int64))))
        return b

    def set_loop_mode(self, mode):
        self._check(self._L.calib_multi_set_loop_mode(self._h, int(mode)))

    def set_ladder(self, max_candidates):
        self._check(self._L.calib_multi_set_ladder(self._h, int(max_candidates)))

    def lm_run(self, P0, maxIters, lam0=1e-3, lam_min=1e-10, lam_max=1e10, err_min=1e-12, want_trace=True):
        """Returns (sse, P, iters, trace[iters,4])."""
        P = as_f64(P0).ravel().copy()
        if P.size != self.K:
            raise ValueError(f"Expected a parameter vector of {self.K} values, got {P.size}")
        sse = ctypes.c_double()
        iters = ctypes.c_int()
        ms = ctypes.c_double()
        trace = np.zeros((max(int(maxIters), 1), 4)) if want_trace else None
        self._check(self._L.calib_multi_lm_run(
            self._h, dptr(P), int(maxIters), lam0, lam_min, lam_max, err_min,
            ctypes.cast(ctypes.byref(sse), _lib._d), ctypes.byref(iters),
            dptr(trace) if want_trace else None, ctypes.cast(ctypes.byref(ms), _lib._d)))
        self.last_device_ms = ms.value
        return sse.value, P, iters.value, (trace[:iters.value].copy() if want_trace else None)

    def flush_l2(self):
        self._check(self._L.calib_multi_flush_l2(self._h))

    def set_param_trace(self, on):
        self._check(self._L.calib_multi_set_param_trace(self._h, 1 if on else 0))

    def param_trace(self, iters):
        out = np.empty((int(iters), self.L))
        self._check(self._L.calib_multi_download_param_trace(self._h, int(iters), dptr(out)))
        return out

    def last_run_stats(self):
        return self.last_device_ms, self.n_gpus, 0

    def close(self):
        if getattr(self, "_h", None):
            self._L.calib_multi_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class BatchProblem:
    """A batch of independent calibrations resident on one GPU (calib_batch_t): one thread block
    per problem, the whole LM loop in one launch.  problems: list of allDetections lists."""

    def __init__(self, model, problems, device=0, csr=None):
        """problems: list of allDetections lists; or csr = (problem_view_offsets, view_offsets,
        model_pts, sensor_pts) already flattened (e.g. in page-locked memory)."""
        self.model = distortion.model_id_of(model)
        self.L = 10 if self.model == distortion.RADTAN else 9
        if csr is None:
            views = [len(p) for p in problems]
            self.n = len(problems)
            self.prob_view = np.zeros(self.n + 1, dtype=np.int64)
            np.cumsum(views, out=self.prob_view[1:])
            flat = [det for p in problems for det in p]
            self.view_offsets, mp, sp = flatten_detections(flat)
        else:
            self.prob_view = np.ascontiguousarray(csr[0], dtype=np.int64)
            self.view_offsets = np.ascontiguousarray(csr[1], dtype=np.int64)
            self.n = len(self.prob_view) - 1
            mp, sp = as_f64(csr[2]).reshape(-1, 3), as_f64(csr[3]).reshape(-1, 2)
        self.total_params = self.L * self.n + 6 * int(self.prob_view[-1])
        self._L = _lib.lib()
        self._h = ctypes.c_void_p()
        i64 = ctypes.POINTER(ctypes.c_int64)
        rc = self._L.calib_batch_create(ctypes.byref(self._h), self.model, self.n,
                                        self.prob_view.ctypes.data_as(i64),
                                        self.view_offsets.ctypes.data_as(i64), dptr(mp), dptr(sp), int(device))
        if rc != 0:
            msg = self._L.calib_batch_last_error(None)
            self._h = None
            raise CalibError(rc, msg.decode() if msg else "calib_batch_create failed")

    def param_slices(self):
        """slice of problem p inside the concatenated parameter vector"""
        starts = self.L * np.arange(self.n + 1) + 6 * self.prob_view
        return [slice(int(starts[p]), int(starts[p + 1])) for p in range(self.n)]

    def lm_run(self, P0, maxIters, lam0=1e-3, lam_min=1e-10, lam_max=1e10, err_min=1e-12, want_trace=False):
        """P0: concatenated parameter vectors.  Returns (sse[n], P, iters[n], trace[n,maxIters,4] | None, device_ms)."""
        P = as_f64(P0).ravel().copy()
        if P.size != self.total_params:
            raise ValueError(f"Expected {self.total_params} parameters, got {P.size}")
        sse = np.zeros(self.n)
        iters = np.zeros(self.n, dtype=np.int32)
        trace = np.zeros((self.n, max(int(maxIters), 1), 4)) if want_trace else None
        ms = ctypes.c_double()
        rc = self._L.calib_batch_lm_run(self._h, dptr(P), int(maxIters), lam0, lam_min, lam_max, err_min,
                                        dptr(sse), iters.ctypes.data_as(ctypes.POINTER(ctypes.c_int)),
                                        dptr(trace) if want_trace else None,
                                        ctypes.cast(ctypes.byref(ms), _lib._d))
        if rc != 0:
            msg = self._L.calib_batch_last_error(self._h)
            raise CalibError(rc, msg.decode() if msg else "")
        return sse, P, iters, trace, ms.value

    def close(self):
        if getattr(self, "_h", None):
            self._L.calib_batch_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def comm_unique_id():
    buf = ctypes.create_string_buffer(128)
    rc = _lib.lib().calib_comm_unique_id(ctypes.cast(buf, ctypes.c_void_p))
    if rc != 0:
        raise CalibError(rc, (_lib.lib().calib_last_error(None) or b"").decode())
    return buf.raw


def fp64_peak_tflops(device=0):
    t = ctypes.c_double()
    rc = _lib.lib().calib_measure_fp64_peak(int(device), ctypes.cast(ctypes.byref(t), _lib._d))
    if rc != 0:
        raise CalibError(rc, (_lib.lib().calib_last_error(None) or b"").decode())
    return t.value


class ProjectionJacobian:
    """Parity face of src/jacobian.py:12-85: compute(P, allModelPoints) -> dense J (2N, L+6M).
    The product path never builds this matrix; it exists so the reference's Jacobian tests can
    be run against the CUDA kernels."""
    _numExtrinsicParamsPerView = 6

    def __init__(self, distortionModel, device=0):
        self._model = distortion.model_id_of(distortionModel)
        self._device = device

    def blocks(self, P, allModelPoints):
        dets = [(np.zeros((len(m), 2)), m) for m in allModelPoints]
        prob = Problem.from_detections(self._model, dets, self._device)
        try:
            _, Jc, Jv = prob.residual_jacobian(P)
        finally:
            prob.close()
        return Jc, Jv

    def compute(self, P, allModelPoints):
        Jc, Jv = self.blocks(P, allModelPoints)
        L = Jc.shape[1]
        M = len(allModelPoints)
        J = np.zeros((Jc.shape[0], L + 6 * M))
        J[:, :L] = Jc
        row = 0
        for i, m in enumerate(allModelPoints):
            n = 2 * len(m)
            J[row:row + n, L + 6 * i:L + 6 * i + 6] = Jv[row:row + n]
            row += n
        return J


class Calibrator:
    """Drop-in for the reference's Calibrator on the refinement path (src/calibrate.py:12-283)."""
    _λinitial = 1e-3
    _λmin = 1e-10
    _λmax = 1e+10
    _Pt_error_min = 1e-12

    def __init__(self, distortionModel, device=0, devices=None, keepResident=False):
        """devices=[0, 1, ...]: refineCalibrationParameters shards the views over these GPUs of this process
        (MultiProblem); everything else runs on devices[0].
        keepResident=True: the detections of the last refineCalibrationParameters call stay on the device(s);
        a following call that passes THE SAME list object (same length, same first / last arrays) runs on
        them without gathering and uploading again -- the reference's re-entry pattern, src/animate.py:40-42
        (maxIters=1, 50 times over one allDetections).  Opt-in because it trusts the caller not to rewrite
        the arrays in place between calls; release() drops the device copy."""
        self._distortionModel = distortionModel
        self._model = distortion.model_id_of(distortionModel)
        self._devices = [int(x) for x in devices] if devices is not None and len(devices) > 1 else None
        self._device = int(devices[0]) if devices is not None and len(devices) >= 1 else device
        self._jac = None
        self.lastTrace = None     # (iters,4): E, E', lambda, accepted
        self.lastIters = 0
        self.lastDeviceMs = 0.0
        self._pinned = PinnedBuffers()
        self._keep = bool(keepResident)
        self._resident = None      # (key, problem)
        self.uploadCount = 0       # detections sent to the device(s) by refineCalibrationParameters so far

    def release(self):
        """drops the device copy kept by keepResident"""
        if self._resident is not None:
            self._resident[1].close()
            self._resident = None

    def __del__(self):
        try:
            self.release()
        except Exception:
            pass

    @staticmethod
    def _detections_key(allDetections):
        n = len(allDetections)
        if n == 0:
            return (id(allDetections), 0)
        s0, m0 = allDetections[0]
        s1, m1 = allDetections[-1]
        return (id(allDetections), n, id(s0), id(m0), id(s1), id(m1), np.shape(s0), np.shape(s1))

    def _problem_for(self, allDetections):
        """(problem, owned): the device-side problem of these detections; owned = close it after the call"""
        key = self._detections_key(allDetections) if self._keep else None
        if self._keep and self._resident is not None:
            if self._resident[0] == key:
                return self._resident[1], False
            self.release()
        if self._devices is not None:
            prob = MultiProblem.from_detections(self._model, allDetections, self._devices, pinned=self._pinned)
        else:
            prob = Problem.from_detections(self._model, allDetections, self._device, pinned=self._pinned)
        self.uploadCount += 1
        if self._keep:
            self._resident = (key, prob)
            return prob, False
        return prob, True

    # -- entry points -----------------------------------------------------------------------
    def calibrate(self, allDetections, maxIters, initial=None):
        """src/calibrate.py:21-39.  The detections are uploaded once; the initial estimate
        (estimateCalibrationParameters) and the refinement run on the same device handle.
        initial=(A, W, k) skips the closed-form estimate."""
        prob = Problem.from_detections(self._model, allDetections, self._device, pinned=self._pinned)
        try:
            if initial is None:
                initial = self._estimate_on(prob)
            Ainitial, Winitial, kInitial = initial
            return self._refine_on(prob, Ainitial, Winitial, kInitial, maxIters, True)
        finally:
            prob.close()

    def estimateCalibrationParameters(self, allDetections):
        """src/calibrate.py:41-58: DLT homographies -> per-view LM refinement -> intrinsics ->
        extrinsics -> linear distortion estimate.  Returns (Ainitial, Winitial, kInitial)."""
        prob = Problem.from_detections(self._model, allDetections, self._device)
        try:
            return self._estimate_on(prob)
        finally:
            prob.close()

    def _estimate_on(self, prob):
        from . import linearcalibrate
        Hs = linearcalibrate.estimateHomographies(prob)
        Hsref, iters, _ = prob.refine_homographies(Hs, 20, 1e-3, self._λmin, self._λmax, self._Pt_error_min)
        self.lastHomographyIters = iters
        Ainitial = linearcalibrate.computeIntrinsicMatrix(Hsref)
        Winitial = linearcalibrate.computeExtrinsics(Hsref, Ainitial)
        kInitial = linearcalibrate.estimateDistortion(prob, Ainitial, Winitial)
        return Ainitial, Winitial, kInitial

    def _refineHomographies(self, Hs, allDetections):
        """src/calibrate.py:60-68: per-view LM refinement of the DLT homographies, all views in one
        launch (calib_refine_homographies).  Returns a list of (3,3) arrays like the reference."""
        prob = Problem.from_detections(self._model, allDetections, self._device)
        try:
            Href, iters, _ = prob.refine_homographies(np.array([np.asarray(H).ravel() for H in Hs]), 20,
                                                      1e-3, self._λmin, self._λmax, self._Pt_error_min)
        finally:
            prob.close()
        self.lastHomographyIters = iters
        return [Href[i] for i in range(len(Href))]

    def _refineHomography(self, H, sensorPoints, modelPoints, jac=None):
        """src/calibrate.py:70-110 for one view."""
        return self._refineHomographies([H], [(sensorPoints, modelPoints)])[0]

    def refineCalibrationParameters(self, Ainitial, Winitial, kInitial, allDetections,
                                    maxIters, shouldPrint=False):
        """src/calibrate.py:117-171.  Returns (sse, Arefined, Wrefined, kRefined)."""
        if int(maxIters) <= 0:
            # the reference returns an unbound Pt_error (src/calibrate.py:171)
            raise UnboundLocalError("cannot access local variable 'Pt_error' where it is not "
                                    "associated with a value")
        # The upload of the detections (gather + PCIe, inside the library, GIL released) runs on a helper
        # thread while this one composes the parameter vector (0.3-0.4 ms at 1000 views) -- for many views.
        if len(allDetections) < 256 or (self._keep and self._resident is not None):
            # few views: composing costs less than the hand-over to the helper thread
            prob, owned = self._problem_for(allDetections)
            try:
                return self._refine_on(prob, Ainitial, Winitial, kInitial, maxIters, shouldPrint)
            finally:
                if owned:
                    prob.close()
        fut = _helper().submit(self._problem_for, allDetections)
        try:
            Pt = self._composeParameterVector(Ainitial, Winitial, kInitial)
        except BaseException:
            try:
                prob, owned = fut.result()
                if owned:
                    prob.close()
            except Exception:
                pass
            raise
        prob, owned = fut.result()
        try:
            return self._refine_on(prob, Ainitial, Winitial, kInitial, maxIters, shouldPrint, Pt=Pt)
        finally:
            if owned:
                prob.close()

    def _refine_on(self, prob, Ainitial, Winitial, kInitial, maxIters, shouldPrint, Pt=None):
        if int(maxIters) <= 0:
            raise UnboundLocalError("cannot access local variable 'Pt_error' where it is not "
                                    "associated with a value")
        if Pt is None:
            Pt = self._composeParameterVector(Ainitial, Winitial, kInitial)
        prob.set_param_trace(bool(shouldPrint))
        ts = time.time()
        sse, P, iters, trace = prob.lm_run(Pt, int(maxIters), self._λinitial, self._λmin,
                                           self._λmax, self._Pt_error_min)
        elapsed = time.time() - ts
        self.lastDeviceMs = prob.last_run_stats()[0]
        self.lastTrace, self.lastIters = trace, iters
        if shouldPrint:
            # src/calibrate.py:158-159, 269-274: per iteration the time since the loop started, min(E', E),
            # lambda, then A and k of the iterate BEFORE the update.  The loop ran on the device in one go:
            # the per-iteration times are its wall time spread evenly over the iterations.
            nd = self._numDistortion()
            for it, ((E, E1, lam, acc), row) in enumerate(zip(trace, prob.param_trace(iters))):
                α, β, γ, uc, vc = row[:5]
                At = np.array([[α, γ, uc], [0, β, vc], [0, 0, 1]])
                print(f"\niter {it}: ({elapsed * (it + 1) / iters:0.3f}s), error={min(E1, E):0.3f}, "
                      f"λ={lam:e}")
                print(f"A:\n{At}")
                print(f"k:\n{row[5:5 + nd]}")
        Arefined, Wrefined, kRefined = self._decomposeParameterVector(P)
        return sse, Arefined, Wrefined, kRefined

    def _initializeJacobian(self):
        """No symbolic build on this path (src/calibrate.py:173-176 takes ~7 s there)."""
        if self._jac is None:
            self._jac = ProjectionJacobian(self._model, self._device)

    def _computeReprojectionError(self, P, allDetections):
        """src/calibrate.py:178-183"""
        prob = Problem.from_detections(self._model, allDetections, self._device)
        try:
            return prob.cost(P)
        finally:
            prob.close()

    def projectAllPoints(self, P, allModelPoints):
        """src/calibrate.py:190-197"""
        dets = [(np.zeros((len(m), 2)), m) for m in allModelPoints]
        prob = Problem.from_detections(self._model, dets, self._device)
        try:
            return prob.project(P)
        finally:
            prob.close()

    # -- parameter vector layout (host) -----------------------------------------------------
    def _numDistortion(self):
        return 5 if self._model == distortion.RADTAN else 4

    def _composeParameterVector(self, A, W, k):
        """src/calibrate.py:199-229: (K,1) column [α β γ uc vc k..., (ρx ρy ρz tx ty tz) per view]."""
        A = np.asarray(A, dtype=np.float64)
        k = list(np.asarray(k, dtype=np.float64).ravel())
        if len(k) != self._numDistortion():
            raise ValueError(f"Expected {self._numDistortion()} distortion coefficients, got {len(k)}")
        W = np.asarray(W, dtype=np.float64)
        if W.ndim != 3 or W.shape[1:] != (4, 4):
            raise ValueError(f"Expected shape (4, 4), got {W.shape[1:] if W.ndim == 3 else W.shape}")
        P = np.empty(5 + len(k) + 6 * len(W))
        P[:5] = A[0, 0], A[1, 1], A[0, 1], A[0, 2], A[1, 2]
        P[5:5 + len(k)] = k
        ext = P[5 + len(k):].reshape(-1, 6)
        ext[:, :3] = mu.rotationMatrixToEulerBatch(W[:, :3, :3])
        ext[:, 3:] = W[:, :3, 3]
        return P.reshape(-1, 1)

    def _decomposeParameterVector(self, P):
        """src/calibrate.py:231-267"""
        P = np.asarray(P, dtype=np.float64).ravel()
        nd = self._numDistortion()
        start = 5 + nd
        α, β, γ, uc, vc = P[:5]
        A = np.array([[α, γ, uc], [0, β, vc], [0, 0, 1]])
        ext = P[start:].reshape(-1, 6)
        Wall = np.zeros((len(ext), 4, 4))
        Wall[:, :3, :3] = mu.eulerToRotationMatrixBatch(ext[:, :3])
        Wall[:, :3, 3] = ext[:, 3:]
        Wall[:, 3, 3] = 1.0
        return A, list(Wall), P[5:start]


def calibrateCamera(allDetections, distortionType, maxIters, initial=None, device=0, devices=None):
    """src/main.py:11-36 on the GPU path: closed-form initial estimate and LM refinement on one device
    handle.  `initial` = (A, W, k) skips the estimate; ValueError on an unknown distortion type as in
    the reference.  devices=[...]: the refinement shards the views over these GPUs (the initial estimate
    runs on the first).  Returns (sse, Afinal, Wfinal, kFinal)."""
    model = distortion.model_id_of(distortionType)
    calibrator = Calibrator(model, device, devices=devices)
    return calibrator.calibrate(allDetections, maxIters, initial=initial)
