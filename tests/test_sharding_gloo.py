"""CPU: the N>1 host logic (view sharding, parameter split/merge, all-reduce of the reduced
system) with torch.distributed gloo, world_size 2, against the single-rank oracle."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

from camera_calibration_b200 import sharding

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_view_bounds_balance_and_cover():
    rs = np.random.RandomState(3)
    for world in (1, 2, 3, 4, 8):
        counts = rs.randint(0, 2000, size=37)
        b = sharding.view_bounds(counts, world)
        assert b[0] == 0 and b[-1] == len(counts) and all(x <= y for x, y in zip(b, b[1:]))
        loads = [counts[b[r]:b[r + 1]].sum() for r in range(world)]
        assert max(loads) - min(loads) <= 2 * counts.max()
    assert sharding.view_bounds([5], 4) == [0, 0, 0, 0, 1] or sharding.view_bounds([5], 4)[-1] == 1
    P = np.arange(10 + 6 * 7, dtype=float)
    parts = [(lo, hi, sharding.local_params(P, 10, lo, hi)) for lo, hi in ((0, 3), (3, 7))]
    assert np.array_equal(sharding.merge_params(parts, 10, 7), P)


@pytest.mark.parametrize("case,iters", [("cfg1", 30), ("cfg2", 12)])
def test_sharded_lm_world2_gloo(case, iters):
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", OMP_NUM_THREADS="2")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
           "--master-addr", "127.0.0.1", "--master-port", "29533",
           os.path.join(ROOT, "tests", "mp", "gloo_worker.py"), case, str(iters)]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=300, env=env)
    lines = [ln for ln in res.stdout.splitlines() if ln.startswith("{")]
    assert res.returncode == 0 and lines, res.stdout[-2000:] + res.stderr[-3000:]
    out = json.loads(lines[-1])
    assert out["ok"], out
    if case == "cfg1":
        assert out["iters"] == out["iters_single"] == 25
