"""Worker of tests/test_multi_gpu.py::test_peer_exchange_soak_and_timeout: the in-kernel exchange over NVLink
peer memory (calib_peer.cuh) under rank skew and with a rank that never arrives.
  soak     200 sharded refinements on one communicator; before every 5th run one rank sleeps a few
           milliseconds (the others then spin in peer_allreduce: the path a race checker would exercise);
           every run must repeat run 0 bit for bit on every rank.
  timeout  rank 1 never calls lm_run: the other ranks must come back with CALIB_E_COMM (-4) after the
           exchange's own timeout (20 s) instead of hanging the GPU; the process then exits cleanly."""
import json
import os
import sys
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import camera_calibration_b200 as cc  # noqa: E402
from camera_calibration_b200 import sharding, synth  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    mode = sys.argv[1]
    c = synth.config("cfg3", 0.04)
    P0 = cc.Calibrator(0)._composeParameterVector(*c["init"]).ravel()
    lo, hi = sharding.view_range(c["dets"], world, rank)
    Ploc = sharding.local_params(P0, 10, lo, hi)
    offs, mp, sp = cc.flatten_detections(c["dets"][lo:hi])
    prob = cc.Problem(0, offs, mp, sp, device=local)
    assert sharding.connect(prob, dist, "peer") == "peer"
    ok, msg = True, ""
    if mode == "soak":
        first = None
        for k in range(200):
            if k % 5 == 0 and rank == (k // 5) % world:
                time.sleep(0.003)
            got = prob.lm_run(Ploc, 100)
            if first is None:
                first = got
            elif not (got[0] == first[0] and got[2] == first[2] and np.array_equal(got[1], first[1])
                      and np.array_equal(got[3], first[3])):
                ok, msg = False, f"run {k} differs from run 0 on rank {rank}"
                break
        flags = [None] * world
        dist.all_gather_object(flags, (ok, msg))
        ok = all(f[0] for f in flags)
        msg = "; ".join(f[1] for f in flags if f[1])
    else:
        code = None
        if rank != 1:
            t0 = time.time()
            try:
                prob.lm_run(Ploc, 100)
                ok, msg = False, "lm_run returned although a rank never arrived"
            except cc.CalibError as e:
                code = e.code
                ok = code == -4 and 5.0 < time.time() - t0 < 60.0
                msg = "" if ok else f"code {code} after {time.time() - t0:.1f} s"
        flags = [None] * world
        dist.all_gather_object(flags, (ok, msg))
        ok = all(f[0] for f in flags)
        msg = "; ".join(f[1] for f in flags if f[1])
    if rank == 0:
        print(json.dumps({"ok": ok, "msg": msg, "mode": mode, "world": world}), flush=True)
    dist.barrier()
    prob.close()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
