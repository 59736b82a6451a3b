"""Multi-GPU plumbing for the localisation path: one process per GPU, torch.distributed
(NCCL on GPUs, gloo in the CPU tests).  The path has no exchange step -- frames are independent -- so the data path
uses no collective at all; only results cross NVLink:

  * frame sharding (primary): rank r localises frames r, r+W, r+2W, ... and the poses are all_gather'ed
    (7 f64 per frame);
  * hypothesis-block sharding (single-frame latency / RANSAC microbench): every rank scores its own block of the
    same counter-based sample stream and the winners are combined by ``best_of_ranks`` with Open3D's
    IsBetterRANSACThan order (fitness desc, inlier_rmse asc) and the lowest hypothesis index on ties.  That equals
    what the sequential loop would have kept ONLY with a fixed hypothesis count (``registration.ransac_early_exit =
    False``, the mode configs[4] and tests/test_scale_gpu.py run): with the adaptive stop each rank's estimated stop
    index is local to its block, ranks above 0 score hypotheses the sequential run would never have drawn, and the
    winner then depends on the world size.  Callers of ``hypothesis_block`` must switch the adaptive stop off.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.distributed as dist


def world():
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def shard_indices(n_items, rank=None, world_size=None):
    """Round-robin shard: indices of the items this rank owns."""
    if rank is None:
        rank, world_size = world()
    return list(range(rank, n_items, world_size))


def hypothesis_block(max_iteration, rank=None, world_size=None):
    """Contiguous block [h0, h1) of the hypothesis stream owned by this rank."""
    if rank is None:
        rank, world_size = world()
    per = (max_iteration + world_size - 1) // world_size
    return min(rank * per, max_iteration), min((rank + 1) * per, max_iteration)


def gather_frame_results(local_indices, local_poses, n_items, device=None):
    """all_gather of the per-frame poses -> [n_items, 7] float64 on every rank (rows this job did not compute stay
    NaN).  Traffic: 64 B per frame."""
    rank, W = world()
    device = device or ("cuda" if torch.cuda.is_available() and dist.is_initialized() and dist.get_backend() == "nccl" else "cpu")
    per = (n_items + W - 1) // W
    buf = torch.full((per, 8), float("nan"), dtype=torch.float64, device=device)
    for k, (i, p) in enumerate(zip(local_indices, local_poses)):
        buf[k, 0] = float(i)
        buf[k, 1:] = torch.as_tensor(np.asarray(p, np.float64), device=device)
    out = torch.full((n_items, 7), float("nan"), dtype=torch.float64)
    if W == 1:
        parts = [buf]
    else:
        parts = [torch.empty_like(buf) for _ in range(W)]
        dist.all_gather(parts, buf)
    for part in parts:
        part = part.cpu()
        for row in part:
            if not torch.isnan(row[0]):
                out[int(row[0])] = row[1:]
    return out.numpy()


def best_of_ranks(fitness, inlier_rmse, hyp_index, T, device=None):
    """Combine per-rank RANSAC winners: 19 f64 per rank through one all_gather, then the deterministic order
    (fitness desc, rmse asc, hypothesis index asc).  -> (fitness, rmse, hyp_index, T[4,4], winning rank)."""
    rank, W = world()
    device = device or ("cuda" if torch.cuda.is_available() and dist.is_initialized() and dist.get_backend() == "nccl" else "cpu")
    rec = torch.tensor([fitness, inlier_rmse, float(hyp_index)] + list(np.asarray(T, np.float64).reshape(16)),
                       dtype=torch.float64, device=device)
    if W == 1:
        recs = [rec]
    else:
        recs = [torch.empty_like(rec) for _ in range(W)]
        dist.all_gather(recs, rec)
    recs = [r.cpu().numpy() for r in recs]
    order = sorted(range(W), key=lambda r: (-recs[r][0], recs[r][1], recs[r][2] if recs[r][2] >= 0 else float("inf")))
    b = recs[order[0]]
    return float(b[0]), float(b[1]), int(b[2]), b[3:].reshape(4, 4).copy(), order[0]
