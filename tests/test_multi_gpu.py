"""GPU, needs >= 2 devices (skipped on a one-GPU box): the view-sharded path over NCCL against
the single-GPU path, launched with torchrun on 127.0.0.1."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ngpu():
    import torch
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


@pytest.mark.parametrize("case,scale,loop_mode,transport", [
    ("cfg3", 0.04, 1, "nccl"), ("cfg5", 0.003, 1, "nccl"),
    ("cfg3", 0.04, 0, "peer"), ("cfg5", 0.003, 0, "peer"), ("cfg3", 0.04, 1, "peer")])
def test_sharded_equals_single(case, scale, loop_mode, transport):
    n = _ngpu()
    if n < 2:
        pytest.skip("needs 2 GPUs")
    world = 2 if n < 4 else 4
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", "29541",
           os.path.join(ROOT, "tests", "mp", "sharded_worker.py"), case, str(scale), str(loop_mode), transport]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    lines = [ln for ln in res.stdout.splitlines() if ln.startswith("{")]
    assert res.returncode == 0 and lines, res.stdout[-2000:] + res.stderr[-6000:]
    out = json.loads(lines[-1])
    assert out["ok"], out
