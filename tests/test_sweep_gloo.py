"""world_size-2 (and 3, uneven shards) gloo test of the multi-GPU sweep's host logic (libkrrt_mgpu.so, include/krrt_mgpu.h):
sharding by contiguous blocks (krrt_mgpu_shard), the chunked gather into a [chunk][rank][records] buffer and its
un-permutation into global query order (krrt_mgpu_chunk_layout).  The C functions are the library's own (pure host
logic, callable without a GPU); gloo stands in for NCCL and the CPU oracle for the per-rank planner (tests may use it)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle_binding as ob
from robosoccer_motion_planning_b200 import capi, scenes, sweep

Q, N_ITER, M, CHUNK = 37, 60, 5, 4


def oracle_plan(starts, goals, obs, dynamic, n_iter, query_base):
    return ob.plan_batch(ob.default_config(p_direct=0.1, seed=9), starts, goals, obs, dynamic, n_iter,
                         query_base=query_base, nthreads=1)


def _worker(rank, world, port, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        item = capi.RESULT_DTYPE.itemsize
        lo, hi = sweep.shard_bounds(Q, rank, world)
        starts, goals, obs = scenes.batch_queries(hi - lo, M, seed=4, first=lo)
        pad = (Q + world - 1) // world
        block = np.zeros(pad, dtype=capi.RESULT_DTYPE)            # this rank's padded block (krrt_mgpu.cu: d_out)
        block[:hi - lo] = oracle_plan(starts, goals, obs, 1, N_ITER, lo)
        gathered = np.zeros(pad * world, dtype=capi.RESULT_DTYPE)  # krrt_mgpu.cu: d_all, [chunk][rank][records]
        n_chunks = (pad + CHUNK - 1) // CHUNK
        for c in range(n_chunks):                                  # one all-gather per chunk, as krrt_sweep_run issues them
            c0, cc = c * CHUNK, min(CHUNK, pad - c * CHUNK)
            send = torch.from_numpy(block[c0:c0 + cc].view(np.uint8).reshape(-1).copy())
            recv = torch.empty(world * cc * item, dtype=torch.uint8)
            dist.all_gather_into_tensor(recv, send)
            gathered[c0 * world:(c0 + cc) * world] = recv.numpy().view(capi.RESULT_DTYPE)
        res = sweep.unpermute(gathered, Q, world, CHUNK)
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
        got = np.load(os.path.join(tmp_path, f"r{r}.npy")).view(capi.RESULT_DTYPE)
        assert np.array_equal(got.view(np.uint8), ref.view(np.uint8)), f"rank {r}"


def test_shard_bounds_cover_everything():
    for q in (0, 1, 7, 64, 65536):
        for w in (1, 2, 3, 4, 8):
            b = [sweep.shard_bounds(q, r, w) for r in range(w)]
            assert b[0][0] == 0 and b[-1][1] == q
            assert all(b[i][1] == b[i + 1][0] for i in range(w - 1))
            sizes = [hi - lo for lo, hi in b]
            assert max(sizes) - min(sizes) <= 1


def test_chunk_layout_is_a_bijection():
    """every query has exactly one home in the gathered buffer, blocks of a chunk are contiguous by rank"""
    for q, w, ch in ((37, 3, 4), (65536, 8, 1024), (65536, 3, 1000), (5, 8, 2), (1, 1, 7), (4096, 2, 4096)):
        pad = (q + w - 1) // w
        n_chunks = max(1, (pad + ch - 1) // ch)
        seen = np.full(max(pad * w, 1), -1, dtype=np.int64)
        covered = np.zeros(q, dtype=bool)
        for r in range(w):
            lo, hi = sweep.shard_bounds(q, r, w)
            for c in range(n_chunks):
                off, cnt, glob = sweep.chunk_layout(q, w, ch, r, c)
                cc = min(ch, pad - c * ch)
                assert off == c * ch * w + r * cc and glob == lo + c * ch
                assert cnt == max(0, min(cc, (hi - lo) - c * ch))
                assert np.all(seen[off:off + cnt] == -1)
                seen[off:off + cnt] = np.arange(glob, glob + cnt)
                covered[glob:glob + cnt] = True
        assert covered.all()
    with pytest.raises(capi.KrrtError):
        sweep.chunk_layout(10, 2, 4, 2, 0)        # rank out of range


def test_mgpu_library_exports_every_declared_symbol():
    import re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    hdr = open(os.path.join(root, "include", "krrt_mgpu.h")).read()
    names = sorted(set(re.findall(r"\b(krrt_(?:mgpu|sweep)_\w+)\s*\(", hdr)))
    L = sweep.lib()
    for n in names:
        assert hasattr(L, n), n
    assert sorted(sweep.EXPORTS) == names
    # libkrrt.so itself still needs nothing but the CUDA runtime; the NCCL dependency lives in libkrrt_mgpu.so only
    import subprocess
    ldd = subprocess.run(["ldd", capi.LIB_PATH], capture_output=True, text=True).stdout
    assert "nccl" not in ldd and "torch" not in ldd
    ldd = subprocess.run(["ldd", sweep.MGPU_LIB_PATH], capture_output=True, text=True).stdout
    assert "libnccl" in ldd and "libkrrt.so" in ldd and "torch" not in ldd


# ---- the NCCL id hand-over of Sweep.from_env (sweep.exchange_id / release_id): one file per JOB ----------------------
def _id_reader(rank, path_dir, q):
    os.environ["KRRT_ID_DIR"] = path_dir
    os.environ["MASTER_PORT"] = "29500"                     # torchrun's defaults: the same for every launch
    os.environ.pop("TORCHELASTIC_RUN_ID", None)
    q.put(sweep.exchange_id(rank + 1, 2, timeout=20.0))


def test_nccl_id_file_belongs_to_one_job(tmp_path, monkeypatch):
    """a second launch with the same MASTER_PORT / run id must not pick up the first launch's id"""
    import multiprocessing
    import time
    monkeypatch.setenv("KRRT_ID_DIR", str(tmp_path))
    monkeypatch.setenv("MASTER_PORT", "29500")
    monkeypatch.delenv("TORCHELASTIC_RUN_ID", raising=False)
    monkeypatch.delenv("KRRT_MGPU_JOB", raising=False)
    ctx = multiprocessing.get_context("spawn")
    for job in range(2):                                     # two jobs in a row from the same launcher (this process)
        ident = bytes([job + 1]) * sweep.ID_BYTES
        q = ctx.Queue()
        p = ctx.Process(target=_id_reader, args=(0, str(tmp_path), q))
        p.start()                                            # the reader is up first and polls, as non-zero ranks do
        time.sleep(0.5)
        # rank 0 of the job lives in a child too, so that its launcher is this process as well
        w = ctx.Process(target=_id_writer, args=(str(tmp_path), ident))
        w.start()
        got = q.get(timeout=60)
        p.join()
        w.join()
        assert got == ident, f"job {job}: the reader took another job's id"
        assert not [f for f in os.listdir(tmp_path) if f.startswith("krrt_mgpu_id_")], "rank 0 leaves no id file behind"
    # a file from before this launcher existed (a crashed job that used the same name) is not accepted
    path, born = sweep._id_path()
    assert born is not None and str(os.getppid()) in os.path.basename(path)
    with open(path, "wb") as f:
        f.write(b"\xee" * sweep.ID_BYTES)
    os.utime(path, (born - 3600, born - 3600))
    with pytest.raises(capi.KrrtError):
        sweep.exchange_id(1, 2, timeout=0.3)
    os.utime(path, None)                                     # written during this launch: accepted
    assert sweep.exchange_id(1, 2, timeout=5.0) == b"\xee" * sweep.ID_BYTES
    # another launch (another nonce) never sees it
    monkeypatch.setenv("KRRT_MGPU_JOB", "job-b")
    with pytest.raises(capi.KrrtError):
        sweep.exchange_id(1, 2, timeout=0.3)


def _id_writer(path_dir, ident):
    os.environ["KRRT_ID_DIR"] = path_dir
    os.environ["MASTER_PORT"] = "29500"
    os.environ.pop("TORCHELASTIC_RUN_ID", None)
    import time
    assert sweep.exchange_id(0, 2, make_id=lambda: ident) == ident
    time.sleep(1.0)                                          # (krrt_mgpu_create_rank: the ranks meet in ncclCommInitRank)
    sweep.release_id(0, 2)
