"""GPU (-m gpu), collected last: calib_cost_noalloc -- the allocation-free variant of the cost sweep whose
scratch is a static device array (include/calib_b200.h; no counterpart in the reference)
-- called through the C-ABI with DEVICE pointers and checked against the existing entry
point calib_cost on the same inputs, two (and more) calls in a row.  torch is used only to hold the caller's
device buffers.  Tolerance: 1e-12 relative (same per-point arithmetic as calib_cost, another summation order)."""
import ctypes

import numpy as np
import pytest

import camera_calibration_b200 as cc
from camera_calibration_b200 import _lib

pytestmark = pytest.mark.gpu

TOL_E = 1e-12


class DeviceCase:
    """the caller's side: detections, parameters and the result slot in device memory (torch tensors)"""

    def __init__(self, model, offs, model_pts, sensor_pts, device=0):
        import torch
        self.torch = torch
        self.dev = torch.device("cuda", device)
        self.model, self.device = int(model), device
        self.M, self.N = len(offs) - 1, int(offs[-1])
        self.offs = torch.as_tensor(np.asarray(offs, dtype=np.int64)).to(self.dev)
        self.mp = torch.as_tensor(np.ascontiguousarray(model_pts, dtype=np.float64).reshape(-1, 3)).to(self.dev)
        self.sp = torch.as_tensor(np.ascontiguousarray(sensor_pts, dtype=np.float64).reshape(-1, 2)).to(self.dev)
        self.E = torch.full((1,), -1.0, dtype=torch.float64, device=self.dev)
        torch.cuda.synchronize(self.dev)

    def cost(self, P, M=None):
        torch = self.torch
        Pd = torch.as_tensor(np.asarray(P, dtype=np.float64).ravel()).to(self.dev)
        self.E.fill_(-1.0)
        stream = torch.cuda.current_stream(self.dev)
        rc = _lib.lib().calib_cost_noalloc(self.model, self.M if M is None else M, self.offs.data_ptr(), self.N,
                                           self.mp.data_ptr(), self.sp.data_ptr(), Pd.data_ptr(), self.E.data_ptr(),
                                           self.device, ctypes.c_void_p(stream.cuda_stream))
        if rc != 0:
            raise cc.CalibError(rc, (_lib.lib().calib_last_error(None) or b"").decode())
        stream.synchronize()
        return float(self.E.item())


def _rel(a, b):
    return abs(a - b) / max(abs(b), 1e-300)


@pytest.mark.parametrize("name", ["cfg2", "itest"])
def test_noalloc_cost_matches_calib_cost_two_calls_in_a_row(golden, name):
    g = golden(name)
    prob = cc.Problem(int(g["model"]), g["view_offsets"], g["model_pts"], g["sensor_pts"])
    E0, E1 = prob.cost(g["P0"]), prob.cost(g["Pfinal"])
    prob.close()
    case = DeviceCase(g["model"], g["view_offsets"], g["model_pts"], g["sensor_pts"])
    a0 = case.cost(g["P0"])            # first call: sets up the static scratch's address for this device
    a1 = case.cost(g["Pfinal"])        # second call in a row, other parameters: nothing of call 1 may survive
    assert _rel(a0, E0) < TOL_E, (a0, E0)
    # at the optimum of a noise-free case the cost is ~1e-15 of rounding: compare at the scale of the initial cost
    assert abs(a1 - E1) < TOL_E * max(E0, E1), (a1, E1)
    # and again: the same inputs give the same bits (fixed summation order, scratch rewritten every call)
    assert case.cost(g["P0"]) == a0 and case.cost(g["Pfinal"]) == a1


def test_noalloc_cost_fisheye_ragged_views(golden):
    g = golden("kat")
    P, pts, offs = g["fisheye_b_P"], g["fisheye_b_pts"], np.asarray(g["fisheye_b_offs"], dtype=np.int64)
    rng = np.random.RandomState(7)
    sensor = rng.uniform(0, 1000, size=(len(pts), 2))
    prob = cc.Problem(1, offs, pts, sensor)
    E = prob.cost(P)
    prob.close()
    case = DeviceCase(1, offs, pts, sensor)
    assert _rel(case.cost(P), E) < TOL_E
    assert _rel(case.cost(P), E) < TOL_E
    # an empty view in the middle (its six parameters are there, it has no points): same total
    v = len(offs) // 2
    offs2 = np.concatenate([offs[:v + 1], offs[v:]])
    L = 9
    P2 = np.concatenate([P[:L + 6 * v], [10.0, -20.0, 30.0, 0.1, 0.2, 1.5], P[L + 6 * v:]])
    case2 = DeviceCase(1, offs2, pts, sensor)
    assert _rel(case2.cost(P2), E) < TOL_E


def test_noalloc_cost_larger_than_one_block_and_no_allocation():
    from camera_calibration_b200 import synth
    c = synth.config("cfg3", 0.05)      # 50 views, ~78 k points: many blocks, the last-block reduction
    cal = cc.Calibrator(c["model"])
    P0 = cal._composeParameterVector(*c["init"]).ravel()
    offs, mp, sp = cc.flatten_detections(c["dets"])
    prob = cc.Problem(c["model"], offs, mp, sp)
    E = prob.cost(P0)
    prob.close()
    case = DeviceCase(c["model"], offs, mp, sp)
    first = case.cost(P0)
    assert _rel(first, E) < TOL_E
    import torch
    Pd_keep = [torch.as_tensor(P0).to(case.dev) for _ in range(4)]      # torch's pool is warm before the measurement
    del Pd_keep
    free0 = torch.cuda.mem_get_info(case.dev)[0]
    for _ in range(5):
        assert case.cost(P0) == first
    assert torch.cuda.mem_get_info(case.dev)[0] == free0          # the calls took no device memory


def test_noalloc_cost_rejects_what_the_scratch_cannot_hold(golden):
    g = golden("cfg1")
    case = DeviceCase(g["model"], g["view_offsets"], g["model_pts"], g["sensor_pts"])
    with pytest.raises(cc.CalibError) as ei:
        case.cost(g["P0"], M=1025)
    assert "1024" in str(ei.value) and "views" in str(ei.value)
    with pytest.raises(cc.CalibError):
        case.cost(g["P0"], M=0)
    # still usable after a refused call
    prob = cc.Problem(int(g["model"]), g["view_offsets"], g["model_pts"], g["sensor_pts"])
    assert _rel(case.cost(g["P0"]), prob.cost(g["P0"])) < TOL_E
    prob.close()
