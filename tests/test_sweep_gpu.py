"""GPU tests of the sweep (configs C4/C5): the sharded plan equals the one-shot plan and the oracle byte for byte, through
libkrrt's batched planner with query_base = lo (any number of shards on one GPU) and through libkrrt_mgpu.so (1 GPU here;
2+ GPUs when the box has them: `gpurun --gpus 2 -- python -m pytest tests/test_sweep_gpu.py -m gpu`)."""
import ctypes
import hashlib

import numpy as np
import pytest

import oracle_binding as ob
from robosoccer_motion_planning_b200 import capi, scenes, sweep

pytestmark = pytest.mark.gpu

Q, N_ITER = 4096, 200


@pytest.fixture(scope="module")
def one_shot(gpu_required):
    starts, goals, obs = scenes.batch_queries(Q, 9, seed=0)
    G = capi.Planner(p_direct=0.1)
    out = G.plan_batch(starts, goals, obs, 1, N_ITER, query_base=0)
    return starts, goals, obs, out


def test_one_shot_equals_oracle(one_shot):
    """all 4096 records of C4 against the CPU checker (every host core: well under a second), not a sample of them"""
    import os
    starts, goals, obs, out = one_shot
    ref = ob.plan_batch(ob.default_config(p_direct=0.1), starts, goals, obs, 1, N_ITER, query_base=0,
                        nthreads=os.cpu_count() or 1)
    differ = np.flatnonzero((out.view(np.uint8).reshape(Q, -1) != ref.view(np.uint8).reshape(Q, -1)).any(axis=1))
    assert differ.size == 0, f"{differ.size} of {Q} records differ, first: query {differ[:5]}"
    # (and the checker plans a block the way it plans its queries one by one: query i draws with query_base + i)
    for i in (0, 1, 777, Q - 1):
        one = ob.plan_batch(ob.default_config(p_direct=0.1), starts[i:i + 1], goals[i:i + 1], obs[i:i + 1], 1, N_ITER, query_base=int(i))
        assert one.tobytes() == ref[i:i + 1].tobytes()


@pytest.mark.parametrize("shards", [1, 2, 3, 8])
def test_contiguous_shards_equal_one_shot(one_shot, shards):
    """the 4096-query sweep planned as 1, 2, 3 and 8 contiguous blocks with query_base = lo: byte equality"""
    starts, goals, obs, out = one_shot
    G = capi.Planner(p_direct=0.1)
    got = np.zeros(Q, dtype=capi.RESULT_DTYPE)
    for r in range(shards):
        lo, hi = sweep.shard_bounds(Q, r, shards)
        got[lo:hi] = G.plan_batch(starts[lo:hi], goals[lo:hi], obs[lo:hi], 1, N_ITER, query_base=lo)
    assert got.tobytes() == out.tobytes()


def test_mgpu_one_gpu(one_shot):
    """krrt_sweep_multi_gpu on one GPU: no gather, the same records; load / run / results and the counters"""
    starts, goals, obs, out = one_shot
    S = sweep.Sweep.single_process(1, p_direct=0.1)
    assert S.info()["world"] == 1 and S.info()["nccl"] == 0
    got = S.sweep(starts, goals, obs, 1, N_ITER)
    assert got.tobytes() == out.tobytes()
    S.load(starts[:1000], goals[:1000], obs[:1000])
    ms = S.run(1, N_ITER, query_base=0)
    assert ms > 0.0 and S.results().tobytes() == out[:1000].tobytes()
    G = capi.Planner(p_direct=0.1)
    G.plan_batch(starts[:1000], goals[:1000], obs[:1000], 1, N_ITER)
    assert S.stats() == G.batch_stats()
    assert S.launch_count >= 2
    # ragged and tiny sweeps
    for q in (1, 7):
        assert S.sweep(starts[:q], goals[:q], obs[:q], 1, N_ITER).tobytes() == out[:q].tobytes()
    S.close()


def test_chunk_counters(one_shot):
    """krrt_plan_batch_device_chunks: every chunk's counter ends at the chunk's size, results unchanged"""
    import torch
    starts, goals, obs, out = one_shot
    q, chunk = 1000, 96
    dev = torch.device("cuda", 0)
    d = [torch.from_numpy(np.ascontiguousarray(x[:q])).to(dev) for x in (starts, goals, obs)]
    d_out = torch.zeros(q * capi.RESULT_DTYPE.itemsize, dtype=torch.uint8, device=dev)
    n_chunks = (q + chunk - 1) // chunk
    d_done = torch.zeros(n_chunks, dtype=torch.int32, device=dev)
    G = capi.Planner(p_direct=0.1)
    st = torch.cuda.Stream(device=dev)
    torch.cuda.synchronize()
    G._ck(capi.lib().krrt_plan_batch_device_chunks(G.h, d[0].data_ptr(), d[1].data_ptr(), d[2].data_ptr(), q, 9, 1, N_ITER,
                                                   0, d_out.data_ptr(), chunk, d_done.data_ptr(), st.cuda_stream))
    st.synchronize()
    want = [min(chunk, q - c * chunk) for c in range(n_chunks)]
    assert d_done.cpu().tolist() == want
    assert d_out.cpu().numpy().tobytes() == out[:q].tobytes()


def test_mgpu_all_gpus_of_the_box(one_shot):
    """2+ GPUs in ONE process (mode A): blocks + chunked NCCL gather == the one-shot plan; sha256 of the records"""
    n = capi.lib().krrt_device_count()
    if n < 2:
        pytest.skip("one GPU on this box; run under gpurun --gpus 2")
    starts, goals, obs, out = one_shot
    for g in sorted({2, min(n, 3), n}):
        S = sweep.Sweep.single_process(g, p_direct=0.1)
        info = S.info()
        assert info["world"] == g and info["nccl"] == 1 and info["nccl_version"] >= 22700
        got = S.sweep(starts, goals, obs, 1, N_ITER)
        assert hashlib.sha256(got.tobytes()).hexdigest() == hashlib.sha256(out.tobytes()).hexdigest()
        # uneven blocks, fewer queries than one chunk per rank, fewer queries than GPUs
        for q in (4095, 1001, g - 1, 1):
            assert S.sweep(starts[:q], goals[:q], obs[:q], 1, N_ITER).tobytes() == out[:q].tobytes()
        S.load(starts, goals, obs)
        S.run(1, N_ITER)
        G = capi.Planner(p_direct=0.1)
        G.plan_batch(starts, goals, obs, 1, N_ITER)
        assert S.stats() == G.batch_stats()
        assert S.max_over_ranks(3.5) == 3.5
        S.barrier()
        S.close()


@pytest.mark.gpu
def test_single_process_over_all_gpus_equals_one_gpu(gpu_required):
    """krrt_mgpu_create: ONE process drives every GPU of the box (a host thread + context + stream per GPU, ncclCommInitAll),
    the way a C++ caller without a launcher does; records byte-equal to the sweep on one GPU.  Skipped on a one-GPU box."""
    n = capi.lib().krrt_device_count()
    if n < 2:
        pytest.skip("one GPU visible")
    q = 4096 + 37
    s, g, o = scenes.batch_queries(q, 9, seed=4)
    S1 = sweep.Sweep.single_process(1, p_direct=0.1)
    ref = S1.sweep(s, g, o, 1, 120)
    S1.close()
    Sn = sweep.Sweep.single_process(n, p_direct=0.1)
    out = Sn.sweep(s, g, o, 1, 120)
    info = Sn.info()
    Sn.close()
    assert info["world"] == n and info["nccl"] == 1
    assert np.array_equal(out.view(np.uint8), ref.view(np.uint8))
