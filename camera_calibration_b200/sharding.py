"""Host-side sharding of one large calibration over ranks: contiguous view
ranges balanced by point count; every rank keeps the L shared parameters and the extrinsics of
its own views.  The reduced system is a sum over views, so a rank's share is the same computation
on fewer views followed by an all-reduce (the C-ABI does it with NCCL on the device;
`allreduce_reduced_system` is the host-side statement of the same step for gloo tests)."""
import numpy as np


def view_bounds(counts, world):
    """[b_0=0, b_1, ..., b_world=M]: rank r owns views b_r .. b_{r+1}-1."""
    counts = np.asarray(counts, dtype=np.int64)
    cum = np.concatenate([[0], np.cumsum(counts)])
    total = cum[-1]
    b = [int(np.searchsorted(cum, total * r / world, side="left")) for r in range(world)] + [len(counts)]
    b[0] = 0
    for r in range(1, world + 1):        # monotone even for degenerate inputs
        b[r] = max(b[r], b[r - 1])
    return b


def view_range(allDetections, world, rank):
    """Views of `rank`.  Every rank computes the same bounds, so a split that would leave ANY rank without
    views (more ranks than views, or all points in a few views) is refused on all ranks alike -- not by
    calib_create on the empty rank alone while its peers wait in collectives.  (The single-process handle,
    calib_multi_create, drops such ranks instead.)"""
    b = view_bounds([len(s) for s, m in allDetections], world)
    empty = [r for r in range(world) if b[r + 1] <= b[r]]
    if empty:
        raise ValueError(f"view sharding over {world} ranks leaves rank(s) {empty} without views "
                         f"({len(allDetections)} views): use fewer ranks")
    return b[rank], b[rank + 1]


def local_params(P, L, lo, hi):
    P = np.asarray(P, dtype=np.float64).ravel()
    return np.concatenate([P[:L], P[L + 6 * lo:L + 6 * hi]])


def merge_params(parts, L, M):
    """parts: iterable of (lo, hi, Plocal) -> the full parameter vector (shared values from the first)."""
    parts = list(parts)
    P = np.empty(L + 6 * M)
    P[:L] = parts[0][2][:L]
    for lo, hi, p in parts:
        P[L + 6 * lo:L + 6 * hi] = np.asarray(p)[L:]
    return P


def allreduce_reduced_system(local_S, local_s, local_E, dist, group=None):
    """Sum of the ranks' Schur complements, right-hand sides and costs (what ncclAllReduce does
    on the device to the 161-double `red` block)."""
    import torch
    flat = torch.from_numpy(np.concatenate([np.asarray(local_S).ravel(), np.asarray(local_s).ravel(), [local_E]]))
    dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
    out = flat.numpy()
    n = np.asarray(local_s).size
    return out[:n * n].reshape(n, n), out[n * n:n * n + n], float(out[-1])


def connect(prob, dist, transport="auto"):
    """Joins the ranks of `dist` (an initialised torch.distributed) for a view-sharded Problem.
    transport "peer": in-kernel all-reduce through NVLink peer memory (CUDA IPC handles gathered
    over `dist`); "nccl": library all-reduces; "auto": peer memory if every rank can map every
    inbox, else NCCL.  Returns the transport in use ("peer" | "nccl" | "none")."""
    import torch
    world, rank = dist.get_world_size(), dist.get_rank()
    if world == 1:
        return "none"
    from . import calibrate
    dev = torch.device("cuda", torch.cuda.current_device())
    got_peer = False
    if transport in ("auto", "peer") and world <= 8:
        mine = torch.frombuffer(bytearray(prob.peer_export()), dtype=torch.uint8).to(dev)
        allh = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(allh, mine)
        ok, err = 1, None
        try:
            prob.peer_attach(world, rank, b"".join(bytes(t.cpu().numpy().tobytes()) for t in allh))
        except Exception as e:       # every rank still takes part in the vote below: the others are entering it
            ok, err = 0, e
        flag = torch.tensor([ok], dtype=torch.int32, device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        got_peer = bool(flag.item())
        if not got_peer and transport == "peer":
            raise RuntimeError("peer-memory transport not available on every rank") from err
    if got_peer and transport != "nccl":
        prob.set_transport(2)
        return "peer"
    if transport == "peer":
        raise RuntimeError("peer-memory transport not available on every rank")
    idt = torch.zeros(128, dtype=torch.uint8, device=dev)
    if rank == 0:
        idt = torch.frombuffer(bytearray(calibrate.comm_unique_id()), dtype=torch.uint8).to(dev)
    dist.broadcast(idt, 0)
    prob.comm_init(world, rank, bytes(idt.cpu().numpy().tobytes()))
    prob.set_transport(1)
    return "nccl"
