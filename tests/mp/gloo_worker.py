"""Worker of tests/test_sharding_gloo.py (CPU, gloo, world_size 2): the view-sharded LM iteration
stated on the host -- local block normal equations (oracle), local Schur complement, all-reduce of
the reduced system, redundant L x L solve, local back-substitution, all-reduce of the trial cost,
identical accept decision on every rank -- must reproduce the single-rank oracle loop."""
import json
import os
import sys

import numpy as np
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    sys.path.insert(0, p)
import oracle_c as oc  # noqa: E402
from camera_calibration_b200 import sharding  # noqa: E402
from camera_calibration_b200.calibrate import flatten_detections  # noqa: E402


def main():
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    g = dict(np.load(os.path.join(ROOT, "tests", "golden", sys.argv[1] + ".npz")))
    model, offs_all = int(g["model"]), g["view_offsets"]
    L = 10 if model == 0 else 9
    M = len(offs_all) - 1
    dets = [(g["sensor_pts"][offs_all[i]:offs_all[i + 1]], g["model_pts"][offs_all[i]:offs_all[i + 1]]) for i in range(M)]
    lo, hi = sharding.view_range(dets, world, rank)
    offs, mp, sp = flatten_detections(dets[lo:hi])
    P = sharding.local_params(g["P0"], L, lo, hi)
    lam, trace = 1e-3, []
    for it in range(int(sys.argv[2])):
        A, B, D, gg, E = oc.normal_equations(model, P, offs, mp, sp)
        # local share of the reduced system: A' and g_c are sums over points, the Schur terms over views
        S = A + lam * np.diag(np.diag(A))
        s = gg[:L].copy()
        Dd = D + lam * np.stack([np.diag(np.diag(d)) for d in D]) if len(D) else D
        for v in range(len(D)):
            X = np.linalg.solve(Dd[v], np.column_stack([B[v].T, gg[L + 6 * v:L + 6 * v + 6]]))
            S -= B[v] @ X[:, :L]
            s -= B[v] @ X[:, L]
        S, s, E = sharding.allreduce_reduced_system(S, s, E, dist)
        dc = np.linalg.solve(S, s)                       # every rank, redundantly
        P1 = P.copy()
        P1[:L] += dc
        for v in range(len(D)):
            P1[L + 6 * v:L + 6 * v + 6] += np.linalg.solve(Dd[v], gg[L + 6 * v:L + 6 * v + 6] - B[v].T @ dc)
        E1 = np.array([oc.cost(model, P1, offs, mp, sp)])
        import torch
        t = torch.from_numpy(E1)
        dist.all_reduce(t)
        E1 = float(t[0])
        acc = E1 < E
        trace.append((E, E1, lam, float(acc)))
        if acc:
            P, lam = P1, lam / 10.0
        else:
            lam = lam * 10.0
        if not (1e-10 < lam < 1e10) or E < 1e-12:
            break
    parts = [None] * world
    dist.all_gather_object(parts, (lo, hi, P.tolist(), trace))
    ok, msg = True, ""
    if rank == 0:
        for a, b, p, t in parts:
            if t != trace or p[:L] != P[:L].tolist():
                ok, msg = False, "ranks disagree"
        Pfull = sharding.merge_params([(a, b, np.array(p)) for a, b, p, t in parts], L, M)
        so, Po, io, to = oc.lm(model, g["P0"], offs_all, g["model_pts"], g["sensor_pts"], int(sys.argv[2]))
        tr = np.array(trace)
        n = min(len(tr), len(to))
        plateau = np.where(np.abs(to[:, 1] - to[:, 0]) <= 1e-10 * to[:, 0])[0]
        stop = min(n, plateau[0]) if len(plateau) else n
        if not (np.array_equal(tr[:stop, 3], to[:stop, 3]) and np.array_equal(tr[:stop, 2], to[:stop, 2])):
            ok, msg = False, "sharded trace differs from the single-rank oracle"
        big = to[:stop, 0] > 1e-9
        if not np.allclose(tr[:stop, 0][big], to[:stop, 0][big], rtol=1e-7):
            ok, msg = False, "costs differ"
        d = np.abs(Pfull - Po)
        d = np.minimum(d, np.abs(d - 360))
        scale = np.maximum(np.abs(Po), 1e-3)
        scale[2] = max(scale[2], abs(Po[0]))
        rel = float((d / scale).max())
        if rel > 1e-7:
            ok, msg = False, f"parameters differ {rel}"
        print(json.dumps({"ok": ok, "msg": msg, "iters": len(tr), "iters_single": int(io), "rel": rel}), flush=True)
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
