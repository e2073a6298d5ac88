"""Worker of tests/test_multi_gpu.py: one rank of a view-sharded refinement (torchrun, NCCL).
Rank 0 also runs the same problem on one GPU and compares."""
import json
import os
import sys

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
    cpu = dist.new_group(backend="gloo")     # waits that leave the GPUs idle while rank 0 drives all of them alone
    name, scale, loop_mode = sys.argv[1], float(sys.argv[2]), int(sys.argv[3])
    c = synth.config(name, scale)
    model = c["model"]
    L = 10 if model == 0 else 9
    P0 = cc.Calibrator(model)._composeParameterVector(*c["init"]).ravel()
    lo, hi = sharding.view_range(c["dets"], world, rank)
    Ploc = sharding.local_params(P0, L, lo, hi)
    offs, mp, sp = cc.flatten_detections(c["dets"][lo:hi])
    prob = cc.Problem(model, offs, mp, sp, device=local)
    prob.set_loop_mode(loop_mode)
    transport = sys.argv[4] if len(sys.argv) > 4 else "nccl"
    used = sharding.connect(prob, dist, transport)
    assert used == transport, (used, transport)
    sse, P, iters, tr = prob.lm_run(Ploc, 100)
    launches = prob.last_run_stats()[1]
    # a second run on the same communicator (sequence numbers carry on) must repeat the first bit for bit
    sse2, P2, iters2, tr2 = prob.lm_run(Ploc, 100)
    assert iters2 == iters and sse2 == sse and np.array_equal(P2, P) and np.array_equal(tr2, tr)
    # a handle created afterwards borrows the communicator of the device's default owner
    prob.comm_set_default(local)
    p3 = cc.Problem(model, offs, mp, sp, device=local)
    assert p3.transport() == prob.transport() != 0
    sse3, P3, iters3, tr3 = p3.lm_run(Ploc, 100)
    assert iters3 == iters and sse3 == sse and np.array_equal(P3, P) and np.array_equal(tr3, tr)
    p3.close()
    prob.comm_set_default(local, on=False)
    # gather the extrinsics of every rank on rank 0
    parts = [None] * world
    dist.all_gather_object(parts, (lo, hi, P.tolist(), sse, iters, tr.tolist()), group=cpu)
    torch.cuda.synchronize()
    ok, msg = True, ""
    if rank == 0:
        Pfull = sharding.merge_params([(a, b, np.array(p)) for a, b, p, *_ in parts], L, len(c["dets"]))
        # every rank must have taken the same decisions and hold the same shared parameters
        for a, b, p, s, it, t in parts:
            if it != iters or s != sse or not np.array_equal(np.array(t), tr) or not np.array_equal(np.array(p)[:L], P[:L]):
                ok, msg = False, "ranks disagree on the trace or the shared parameters"
        offs_a, mp_a, sp_a = cc.flatten_detections(c["dets"])
        one = cc.Problem(model, offs_a, mp_a, sp_a, device=local)
        s1, P1, i1, t1 = one.lm_run(P0, 100)
        one.close()
        n = min(len(tr), len(t1))
        plateau = np.where(np.abs(t1[:, 1] - t1[:, 0]) <= 1e-10 * t1[:, 0])[0]
        stop = min(n, plateau[0]) if len(plateau) else n
        if not (np.array_equal(tr[:stop, 3], t1[:stop, 3]) and np.array_equal(tr[:stop, 2], t1[:stop, 2])):
            ok, msg = False, "sharded and single-GPU traces differ before the plateau"
        if not np.allclose(tr[:stop, 0], t1[:stop, 0], rtol=1e-9):
            ok, msg = False, "costs differ"
        d = np.abs(Pfull - P1)
        rel = float((d / np.maximum(np.abs(P1), 1e-3)).max())
        if rel > 1e-6 or abs(sse - s1) > 1e-9 * s1:
            ok, msg = False, f"parameters differ: {rel}"
        # the single-process handle over the same devices (calib_multi_*: library-side sharding, peer access
        # instead of IPC handles) must repeat the one-process-per-GPU result bit for bit
        if transport == "peer":
            mprob = cc.MultiProblem(model, offs_a, mp_a, sp_a, devices=list(range(world)))
            mprob.set_loop_mode(2 if loop_mode == 2 else 0)
            sm, Pm, im, tm = mprob.lm_run(P0, 100)
            mprob.close()
            if not (im == iters and sm == sse and np.array_equal(tm, tr) and np.array_equal(Pm, Pfull)):
                ok, msg = False, "single-process multi-device handle differs from the one-process-per-GPU path"
        print(json.dumps({"ok": ok, "msg": msg, "world": world, "iters_sharded": iters, "iters_single": i1,
                          "pre_plateau_iterations": int(stop), "max_rel_param_diff": rel,
                          "transport": used, "launches": int(launches)}), flush=True)
    dist.barrier(group=cpu)      # nobody unmaps an inbox a peer may still write to (host-side barrier: rank 0 may
                                 # still be running on every GPU through the single-process handle)
    prob.close()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
