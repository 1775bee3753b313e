"""Monte-Carlo sweep of independent planning queries sharded over ranks (BASELINE config 5).

Queries are independent (own start / goal / obstacles / Philox query id), so the path shards with NO data-path
collective: rank g plans the contiguous block [g*Q/G, (g+1)*Q/G) and the only exchange is one all-gather of the
fixed-size result records (krrt_result, 1304 B each) at the end — NCCL over NVLink on GPUs, gloo in the CPU tests.
Results do not depend on the number of ranks because the Philox stream is keyed by the GLOBAL query id.
"""
import numpy as np

from .capi import RESULT_DTYPE


def shard_bounds(q_total, rank, world):
    """Contiguous block of rank `rank` (sizes differ by at most one query)."""
    return (q_total * rank) // world, (q_total * (rank + 1)) // world


def max_shard(q_total, world):
    return max(shard_bounds(q_total, r, world)[1] - shard_bounds(q_total, r, world)[0] for r in range(world))


def gather_results(local, q_total, rank, world, dist=None, device=None):
    """All-gather the per-rank result records into the full, globally ordered array (on every rank).

    `local`: structured numpy array (RESULT_DTYPE) of this rank's block, or a torch uint8 tensor
    [n_local * itemsize] already on `device` (no host round trip before the collective)."""
    import torch
    item = RESULT_DTYPE.itemsize
    pad = max_shard(q_total, world)
    if isinstance(local, np.ndarray):
        t = torch.from_numpy(local.view(np.uint8).reshape(-1).copy())
        if device is not None:
            t = t.to(device)
    else:
        t = local.reshape(-1)
    buf = torch.zeros(pad * item, dtype=torch.uint8, device=t.device)
    buf[: t.numel()] = t
    if world == 1 or dist is None:
        full = buf
    else:
        full = torch.empty(world * pad * item, dtype=torch.uint8, device=t.device)
        dist.all_gather_into_tensor(full, buf)
    full = full.cpu().numpy().reshape(world, pad * item)
    out = np.zeros(q_total, dtype=RESULT_DTYPE)
    for r in range(world):
        lo, hi = shard_bounds(q_total, r, world)
        out[lo:hi] = full[r, : (hi - lo) * item].view(RESULT_DTYPE)
    return out


def run_sweep(plan_fn, q_total, n_iter, m, scene_seed, rank=0, world=1, dist=None, device=None, dynamic=1):
    """plan_fn(starts, goals, obstacles, dynamic, n_iter, query_base) -> RESULT_DTYPE array for this rank's block.
    The product passes Planner.plan_batch; the CPU (gloo) tests pass the oracle to exercise the sharding logic."""
    from . import scenes
    lo, hi = shard_bounds(q_total, rank, world)
    starts, goals, obs = scenes.batch_queries(hi - lo, m, seed=scene_seed, first=lo)
    local = plan_fn(starts, goals, obs, dynamic, n_iter, lo)
    return gather_results(local, q_total, rank, world, dist=dist, device=device)
