"""world_size-2 (and 3, uneven shards) gloo test of the multi-GPU sweep's host logic: sharding by contiguous
blocks + one all-gather of result records.  The CPU oracle stands in for the per-rank planner (tests may use it)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle_binding as ob
from robosoccer_motion_planning_b200 import scenes, sweep

Q, N_ITER, M = 37, 60, 5


def oracle_plan(starts, goals, obs, dynamic, n_iter, query_base):
    return ob.plan_batch(ob.default_config(p_direct=0.1, seed=9), starts, goals, obs, dynamic, n_iter,
                         query_base=query_base, nthreads=1)


def _worker(rank, world, port, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        res = sweep.run_sweep(oracle_plan, Q, N_ITER, M, scene_seed=4, rank=rank, world=world, dist=dist)
        np.save(os.path.join(out_dir, f"r{rank}.npy"), res.view(np.uint8))
    finally:
        dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_sweep_equals_single_rank(tmp_path, world):
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    starts, goals, obs = scenes.batch_queries(Q, M, seed=4)
    ref = oracle_plan(starts, goals, obs, 1, N_ITER, 0)
    for r in range(world):
        got = np.load(os.path.join(tmp_path, f"r{r}.npy")).view(sweep.RESULT_DTYPE)
        assert np.array_equal(got.view(np.uint8), ref.view(np.uint8)), f"rank {r}"


def test_shard_bounds_cover_everything():
    for q in (0, 1, 7, 64, 65536):
        for w in (1, 2, 3, 4, 8):
            b = [sweep.shard_bounds(q, r, w) for r in range(w)]
            assert b[0][0] == 0 and b[-1][1] == q
            assert all(b[i][1] == b[i + 1][0] for i in range(w - 1))
            sizes = [hi - lo for lo, hi in b]
            assert max(sizes) - min(sizes) <= 1
