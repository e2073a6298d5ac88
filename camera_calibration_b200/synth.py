"""Synthetic checkerboard inputs of the named benchmark shapes.  The reference's generators
(src/dataset.py:59-95, src/virtualcamera.py:33-55, src/checkerboard.py:9-17, src/noise.py:8-18)
cannot travel to the GPU box, so their procedure is restated here with the same seeded numpy
RNG calls in the same order: the detections are bit-identical to the reference's (pinned by
tests/test_synth.py against the detections stored in tests/golden/*.npz)."""
import numpy as np

from . import distortion
from . import mathutils as mu

A_SYNTHETIC = np.array([[803.1, 0, 700.5], [0, 803.1, 529.2], [0, 0, 1]], dtype=np.float64)
K_RADTAN = (-0.25, 0.2, 0.07, -0.03, 0.05)
K_FISHEYE = (-0.155, -0.02, 0.0, -0.03)
A_REALISTIC = np.array([[1432.1, 0, 719.2], [0, 1432.1, 564.3], [0, 0, 1]], dtype=np.float64)
K_REALISTIC = (-0.2674, 0.1716, 1.4287e-05, 0.000177, -0.052701)


def checkerboard_corners(numCornersWidth, numCornersHeight, spacing):
    """src/checkerboard.py:9-17: row-major planar grid, Z = 0."""
    pts = [(i * spacing, j * spacing, 0) for j in range(numCornersHeight) for i in range(numCornersWidth)]
    return np.array(pts).reshape(-1, 3)


def _camera_pose_in_board(aim, euler, distance):
    """src/dataset.py:84-95"""
    perturb = mu.poseFromRT(mu.eulerToRotationMatrix(euler), (0, 0, 0))
    facing = mu.poseFromRT(mu.eulerToRotationMatrix((180, 0, 0)), aim)
    back = mu.poseFromRT(np.eye(3), (0, 0, -distance))
    return facing @ perturb @ back


def make_dataset(corners, A, k, model, width, height, numViews, noise_sigma=None,
                 roll_pitch=(-30, 30), yaw=(-180, 180), dist=(0.5, 1.0), view_seed_base=0):
    """src/dataset.py:59-82 + src/virtualcamera.py:42-55.  Returns (allDetections, poses) with
    allDetections = [(sensorPoints (Ni,2), modelPoints (Ni,3))] and poses = board-in-camera 4x4.
    Uses (and reseeds) numpy's global RNG exactly like the reference: seed = view index."""
    mdl = (distortion.RadialTangentialModel() if distortion.model_id_of(model) == distortion.RADTAN
           else distortion.FisheyeModel())
    n = corners.shape[0]
    dets, poses = [], []
    for viewIndex in range(numViews):
        np.random.seed(view_seed_base + viewIndex)
        aimIndex = np.random.choice(n)
        rx = np.random.uniform(*roll_pitch)
        ry = np.random.uniform(*roll_pitch)
        rz = np.random.uniform(*yaw)
        d = np.random.uniform(*dist)
        cMb = np.linalg.inv(_camera_pose_in_board(corners[aimIndex], (rx, ry, rz), d))
        poses.append(cMb)
        cP = mu.transform(cMb, corners)
        u = mdl.projectWithDistortion(A, cP, k)
        if noise_sigma is not None:
            u = u + np.random.normal(0.0, noise_sigma, u.shape)
        keep = ((u[:, 0] > 0) & (u[:, 0] < width) & (u[:, 1] > 0) & (u[:, 1] < height) & (cP[:, 2] > 0))
        dets.append((u[keep], corners[keep]))
    return dets, poses


def make_dataset_on_device(corners, A, k, model, width, height, numViews, noise_sigma=None,
                           roll_pitch=(-30, 30), yaw=(-180, 180), dist=(0.5, 1.0), view_seed_base=0, device=0):
    """The same generator with the projection of ALL views done by one
    device pass (calib_project = Calibrator.projectAllPoints, the kernels' own point_eval) instead of
    one numpy projectWithDistortion per view.  The seeded RNG stays on the host and is consumed in the
    reference's order (pose draws, then the view's noise), so poses, noise and masks are the reference's;
    the projected coordinates agree with the host path to rounding (<= 1e-9 px, tests), not bit for bit
    -- the parity fixtures therefore keep using make_dataset."""
    from . import calibrate
    mid = distortion.model_id_of(model)
    n = corners.shape[0]
    poses = []
    for viewIndex in range(numViews):
        np.random.seed(view_seed_base + viewIndex)
        aimIndex = np.random.choice(n)
        rx = np.random.uniform(*roll_pitch)
        ry = np.random.uniform(*roll_pitch)
        rz = np.random.uniform(*yaw)
        d = np.random.uniform(*dist)
        poses.append(np.linalg.inv(_camera_pose_in_board(corners[aimIndex], (rx, ry, rz), d)))
    W = np.array(poses)
    offs = np.arange(numViews + 1, dtype=np.int64) * n
    model_pts = np.tile(np.asarray(corners, dtype=np.float64), (numViews, 1))
    prob = calibrate.Problem(mid, offs, model_pts, np.zeros((numViews * n, 2)), device)
    try:
        P = calibrate.Calibrator(mid, device)._composeParameterVector(A, poses, k).ravel()
        u_all = prob.project(P).reshape(numViews, n, 2)
    finally:
        prob.close()
    z_all = np.einsum("vj,nj->vn", W[:, 2, :3], corners) + W[:, 2, 3:4]      # camera-frame depth
    dets = []
    for viewIndex in range(numViews):
        u = u_all[viewIndex]
        if noise_sigma is not None:
            np.random.seed(view_seed_base + viewIndex)      # replay the pose draws, then the noise
            np.random.choice(n)
            for rng in (roll_pitch, roll_pitch, yaw, dist):
                np.random.uniform(*rng)
            u = u + np.random.normal(0.0, noise_sigma, u.shape)
        keep = ((u[:, 0] > 0) & (u[:, 0] < width) & (u[:, 1] > 0) & (u[:, 1] < height) & (z_all[viewIndex] > 0))
        dets.append((u[keep], corners[keep]))
    return dets, poses


def perturbed_initial(A, k, poses, rs, rel_A=0.01, euler_deg=0.5, trans=0.005, zero_k=True):
    """Initial guess for sizes where the reference's O(N^2) linear stage cannot run
    (SURVEY.md section 8d): ground truth perturbed with a fixed RandomState."""
    A0 = np.array(A, dtype=np.float64)
    A0[0, 0] *= 1 + rel_A * rs.uniform(-1, 1)
    A0[1, 1] *= 1 + rel_A * rs.uniform(-1, 1)
    A0[0, 2] *= 1 + rel_A * rs.uniform(-1, 1)
    A0[1, 2] *= 1 + rel_A * rs.uniform(-1, 1)
    k0 = np.zeros(len(k)) if zero_k else np.array(k, dtype=np.float64)
    W0 = []
    for cMb in poses:
        e = np.array(mu.rotationMatrixToEuler(cMb[:3, :3])) + rs.uniform(-euler_deg, euler_deg, 3)
        t = cMb[:3, 3] + rs.uniform(-trans, trans, 3)
        W0.append(mu.poseFromRT(mu.eulerToRotationMatrix(e), t))
    return A0, W0, k0


# ---- the named configurations (BASELINE.json "configs") -------------------------------------
def _dataset_range(args):
    corners, A, k, model, W, H, lo, hi, sigma = args
    return make_dataset(corners, A, k, model, W, H, hi - lo, sigma, view_seed_base=lo)


def make_dataset_parallel(corners, A, k, model, width, height, numViews, noise_sigma, procs):
    """make_dataset with the views generated by `procs` worker processes: every view is seeded by its own
    index (src/dataset.py:64), so the result does not depend on how the views are split."""
    if procs <= 1 or numViews < 4 * procs:
        return make_dataset(corners, A, k, model, width, height, numViews, noise_sigma)
    import multiprocessing as mp
    cuts = [numViews * i // (4 * procs) for i in range(4 * procs + 1)]
    jobs = [(corners, A, k, model, width, height, cuts[i], cuts[i + 1], noise_sigma) for i in range(4 * procs)]
    with mp.get_context("fork").Pool(procs) as pool:
        parts = pool.map(_dataset_range, jobs)
    dets = [d for p in parts for d in p[0]]
    poses = [w for p in parts for w in p[1]]
    return dets, poses


def config(name, scale=1.0, procs=1):
    """Returns dict(model, dets, poses, A, k, init=(A0,W0,k0) or None, width, height).
    cfg1/cfg1n/cfg2/itest use the reference's linear stage for the initial guess (stored in
    tests/golden); cfg3/cfg4/cfg5 use perturbed ground truth.  `scale` shrinks the number of
    views of the large configurations for tests; procs > 1 generates the views of the large
    configurations in worker processes (same result)."""
    W, H = 1440, 1080
    if name in ("cfg1", "cfg1n"):
        c = checkerboard_corners(9, 6, 0.050)
        dets, poses = make_dataset(c, A_SYNTHETIC, K_RADTAN, "radtan", W, H, 10,
                                   0.1 if name == "cfg1n" else None)
        return dict(model=0, dets=dets, poses=poses, A=A_SYNTHETIC, k=K_RADTAN, init=None)
    if name == "cfg2":
        c = checkerboard_corners(9, 6, 0.050)
        dets, poses = make_dataset(c, A_SYNTHETIC, K_FISHEYE, "fisheye", W, H, 25, 0.1)
        return dict(model=1, dets=dets, poses=poses, A=A_SYNTHETIC, k=K_FISHEYE, init=None)
    if name == "itest":
        c = checkerboard_corners(25, 18, 0.030)
        dets, poses = make_dataset(c, A_REALISTIC, K_REALISTIC, "radtan", W, H, 15, None)
        return dict(model=0, dets=dets, poses=poses, A=A_REALISTIC, k=K_REALISTIC, init=None)
    if name == "cfg3":
        views = max(2, int(round(1000 * scale)))
        c = checkerboard_corners(40, 40, 0.30 / 39)   # ~98 % of the corners in view, ragged
        dets, poses = make_dataset_parallel(c, A_REALISTIC, K_REALISTIC, "radtan", W, H, views, 0.1, procs)
        init = perturbed_initial(A_REALISTIC, K_REALISTIC, poses, np.random.RandomState(12345))
        return dict(model=0, dets=dets, poses=poses, A=A_REALISTIC, k=K_REALISTIC, init=init)
    if name == "cfg5":
        views = max(2, int(round(10000 * scale)))
        c = checkerboard_corners(64, 64, 0.40 / 63)
        dets, poses = make_dataset_parallel(c, A_SYNTHETIC, K_FISHEYE, "fisheye", W, H, views, 0.1, procs)
        init = perturbed_initial(A_SYNTHETIC, K_FISHEYE, poses, np.random.RandomState(12345))
        return dict(model=1, dets=dets, poses=poses, A=A_SYNTHETIC, k=K_FISHEYE, init=init)
    raise ValueError("unknown config " + name)


def _batch_one(args):
    j, views = args
    c = checkerboard_corners(9, 6, 0.050)
    dets, poses = make_dataset(c, A_SYNTHETIC, K_RADTAN, "radtan", 1440, 1080, views, 0.1,
                               view_seed_base=views * j)
    init = perturbed_initial(A_SYNTHETIC, K_RADTAN, poses, np.random.RandomState(j))
    return dict(model=0, dets=dets, poses=poses, init=init)


def batch_config(n_problems, views=20, procs=1):
    """cfg4: independent 20-view calibrations of the 9x6 board; problem j uses view seeds
    20*j .. 20*j+19 and RandomState(j) for its perturbed initial guess.  procs > 1 generates the
    problems in worker processes (the per-view seeding makes the result independent of it)."""
    jobs = [(j, views) for j in range(n_problems)]
    if procs <= 1 or n_problems < 64:
        return [_batch_one(a) for a in jobs]
    import multiprocessing as mp
    with mp.get_context("fork").Pool(procs) as pool:
        return pool.map(_batch_one, jobs, chunksize=max(1, n_problems // (8 * procs)))
