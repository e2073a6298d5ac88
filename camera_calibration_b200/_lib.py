"""ctypes binding of libcalib_b200.so (include/calib_b200.h).  There is no CPU fallback:
if the library is missing or no CUDA device is visible, the calls fail loudly."""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
SO_PATH = os.environ.get("CALIB_B200_SO") or os.path.join(_HERE, "libcalib_b200.so")   # override: kernel-variant experiments
HEADER = os.path.join(_HERE, "..", "include", "calib_b200.h")

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-std=c++17", "-O3", "-lineinfo",
              "-shared", "-Xcompiler", "-fPIC"]


class CalibError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("libcalib_b200 error %d: %s" % (code, msg))
        self.code = code


def _nccl_include():
    incs = ["/usr/include"]
    try:
        import importlib.util
        spec = importlib.util.find_spec("nvidia.nccl")
        if spec and spec.submodule_search_locations:
            incs.append(os.path.join(list(spec.submodule_search_locations)[0], "include"))
    except Exception:
        pass
    return [i for i in incs if os.path.exists(os.path.join(i, "nccl.h"))]


def sources():
    return [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC)) if f.endswith((".cu", ".cuh", ".inc"))] + [HEADER]


def build(force=False, verbose=False):
    """Compiles csrc/calib_api.cu for sm_100a into the in-tree shared library."""
    srcs = sources()
    if (not force and os.path.exists(SO_PATH)
            and os.path.getmtime(SO_PATH) >= max(os.path.getmtime(s) for s in srcs)):
        return SO_PATH
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    cmd = [nvcc, *NVCC_FLAGS]
    for inc in _nccl_include():
        cmd += ["-I", inc]
    if verbose:
        cmd += ["-Xptxas", "-v"]
    cmd += ["-o", SO_PATH, os.path.join(CSRC, "calib_api.cu"), "-ldl"]
    env = dict(os.environ)
    env.pop("CC", None)
    env.pop("CXX", None)
    res = subprocess.run(cmd, capture_output=True, text=True, env=env)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + res.stdout + res.stderr)
    if verbose:
        print(res.stderr)
    return SO_PATH


_lib = None
_d = ctypes.POINTER(ctypes.c_double)
_i64 = ctypes.POINTER(ctypes.c_int64)
_vp = ctypes.c_void_p


def lib():
    """Loads the library (never builds implicitly on a GPU box: build() is explicit)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(SO_PATH):
        raise CalibError(-2, "libcalib_b200.so is not built (run __graft_entry__.build()); "
                             "there is no CPU fallback")
    # make sure the NCCL that torch ships is the one dlopen("libnccl.so.2") finds
    L = ctypes.CDLL(SO_PATH, mode=ctypes.RTLD_GLOBAL)
    c_int, c_dbl, c_i64 = ctypes.c_int, ctypes.c_double, ctypes.c_int64
    sig = {
        "calib_last_error": (ctypes.c_char_p, [_vp]),
        "calib_create": (c_int, [ctypes.POINTER(_vp), c_int, c_i64, _i64, _d, _d, c_int]),
        "calib_create_from_views": (c_int, [ctypes.POINTER(_vp), c_int, c_i64, _i64, _vp, _vp, c_int]),
        "calib_destroy": (None, [_vp]),
        "calib_trim": (c_int, [c_int]),
        "calib_num_params": (c_i64, [_vp]),
        "calib_num_points": (c_i64, [_vp]),
        "calib_set_segment": (c_int, [_vp, c_int]),
        "calib_set_loop_mode": (c_int, [_vp, c_int]),
        "calib_set_cost_staging": (c_int, [_vp, c_int]),
        "calib_set_ladder": (c_int, [_vp, c_int]),
        "calib_set_ladder_start": (c_int, [_vp, c_int]),
        "calib_host_alloc": (_vp, [ctypes.c_size_t]),
        "calib_host_free": (None, [_vp]),
        "calib_host_gather": (c_int, [_vp, _vp, _i64, c_i64, c_int]),
        "calib_project": (c_int, [_vp, _d, _d]),
        "calib_cost": (c_int, [_vp, _d, _d]),
        "calib_eval_residual_jacobian": (c_int, [_vp, _d, _d, _d, _d]),
        "calib_eval_normal_equations": (c_int, [_vp, _d, _d, _d, _d, _d, _d]),
        "calib_lm_step": (c_int, [_vp, _d, c_dbl, _d]),
        "calib_lm_run": (c_int, [_vp, _d, c_int, c_dbl, c_dbl, c_dbl, c_dbl, _d,
                                 ctypes.POINTER(c_int), _d]),
        "calib_upload_params": (c_int, [_vp, _d]),
        "calib_lm_run_resident": (c_int, [_vp, c_int, c_dbl, c_dbl, c_dbl, c_dbl, _d,
                                          ctypes.POINTER(c_int)]),
        "calib_download_params": (c_int, [_vp, _d]),
        "calib_download_trace": (c_int, [_vp, c_int, _d]),
        "calib_set_param_trace": (c_int, [_vp, c_int]),
        "calib_download_param_trace": (c_int, [_vp, c_int, _d]),
        "calib_multi_flush_l2": (c_int, [_vp]),
        "calib_multi_set_param_trace": (c_int, [_vp, c_int]),
        "calib_multi_download_param_trace": (c_int, [_vp, c_int, _d]),
        "calib_last_run_stats": (c_int, [_vp, _d, _i64, _i64]),
        "calib_set_profiling": (c_int, [_vp, c_int]),
        "calib_kernel_time": (c_int, [_vp, c_int, _d, _i64]),
        "calib_time_kernel": (c_int, [_vp, _d, c_dbl, c_int, c_int, c_int, _d]),
        "calib_measure_fp64_peak": (c_int, [c_int, _d]),
        "calib_flush_l2": (c_int, [_vp]),
        "calib_refine_homographies": (c_int, [_vp, _d, c_int, c_dbl, c_dbl, c_dbl, c_dbl,
                                              ctypes.POINTER(c_int), _d]),
        "calib_dlt_gram": (c_int, [_vp, _d, _d]),
        "calib_dlt_homographies": (c_int, [_vp, _d]),
        "calib_distortion_rows": (c_int, [_vp, _d, _d, _d]),
        "calib_comm_unique_id": (c_int, [_vp]),
        "calib_comm_init": (c_int, [_vp, c_int, c_int, _vp]),
        "calib_comm_share": (c_int, [_vp, _vp]),
        "calib_comm_set_default": (c_int, [_vp, c_int]),
        "calib_peer_export": (c_int, [_vp, _vp]),
        "calib_peer_attach": (c_int, [_vp, c_int, c_int, _vp]),
        "calib_set_transport": (c_int, [_vp, c_int]),
        "calib_transport": (c_int, [_vp]),
        "calib_multi_create": (c_int, [ctypes.POINTER(_vp), c_int, c_i64, _i64, _d, _d, c_int, ctypes.POINTER(c_int)]),
        "calib_multi_create_from_views": (c_int, [ctypes.POINTER(_vp), c_int, c_i64, _i64, _vp, _vp, c_int,
                                                  ctypes.POINTER(c_int)]),
        "calib_multi_destroy": (None, [_vp]),
        "calib_multi_last_error": (ctypes.c_char_p, [_vp]),
        "calib_multi_num_gpus": (c_int, [_vp]),
        "calib_multi_num_params": (c_i64, [_vp]),
        "calib_multi_view_bounds": (c_int, [_vp, _i64]),
        "calib_multi_set_loop_mode": (c_int, [_vp, c_int]),
        "calib_multi_set_ladder": (c_int, [_vp, c_int]),
        "calib_multi_lm_run": (c_int, [_vp, _d, c_int, c_dbl, c_dbl, c_dbl, c_dbl, _d, ctypes.POINTER(c_int), _d, _d]),
        "calib_batch_create": (c_int, [ctypes.POINTER(_vp), c_int, c_i64, _i64, _i64, _d, _d, c_int]),
        "calib_batch_destroy": (None, [_vp]),
        "calib_batch_last_error": (ctypes.c_char_p, [_vp]),
        "calib_batch_lm_run": (c_int, [_vp, _d, c_int, c_dbl, c_dbl, c_dbl, c_dbl, _d,
                                       ctypes.POINTER(c_int), _d, _d]),
        # device pointers as integers: (model, M, view_offsets_dev, N, model_pts_dev, sensor_pts_dev, P_dev, E_dev,
        # device_id, cuda_stream)
        "calib_cost_noalloc": (c_int, [c_int, c_i64, _vp, c_i64, _vp, _vp, _vp, _vp, c_int, _vp]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)
        fn.restype = res
        fn.argtypes = args
    L._signatures = sig
    _lib = L
    return L


def exported_symbols():
    """Names declared in include/calib_b200.h (used by the CPU-side ABI test)."""
    import re
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(calib_[a-z0-9_]+)\s*\(", text)))


def dptr(a):
    return a.ctypes.data_as(_d)


def as_f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)
