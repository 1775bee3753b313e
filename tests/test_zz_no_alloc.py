"""krrt_plan_batch_static — the batched planner for callers that may not allocate (include/krrt.h).

Some users of libkrrt cannot afford the cudaMalloc / cudaFree a krrt_ctx does behind their back: a planner that runs inside
a robot's fixed-rate control cycle (an allocation takes a driver-wide lock and may wait for the whole GPU), or a process
whose device memory is already owned by another allocator's pool (here: torch's, which stands in for such a caller — it
hands out the input and output buffers).  For them the library carries its scratch as a fixed array inside the .so and
plans without a context.  The array is shared by all calls and keeps the last call's contents, so the thing to test is not
one call but calls in a row whose table, counters and redo list differ: each must equal what krrt_plan_batch_device writes
for the same inputs on a fresh context, byte for byte (the tolerance of every planner test here: none).

This file sorts last on purpose: pytest -x stops at the first failure, and a failure here must not hide another test.
"""
import ctypes

import numpy as np
import pytest

from robosoccer_motion_planning_b200 import capi, scenes


def _static_rc(cfg, q, n_iter, ptr=0):
    """the raw status of a call that must be turned down before any CUDA call is made (so it runs without a GPU too)"""
    return capi.lib().krrt_plan_batch_static(ctypes.byref(cfg) if cfg is not None else None, ptr, ptr, ptr, q, 9, 1,
                                             n_iter, 0, ptr, None)


def test_static_limits_are_checked_before_the_gpu_is_touched():
    """beyond the compiled-in scratch the call says so — with both limits in the text — and does nothing"""
    L = capi.lib()
    cfg = capi.default_config()
    assert _static_rc(cfg, capi.KRRT_STATIC_MAX_QUERIES + 1, 200) == capi.KRRT_ERR_INVALID
    msg = L.krrt_last_error(None).decode()
    assert str(capi.KRRT_STATIC_MAX_QUERIES) in msg and str(capi.KRRT_STATIC_MAX_ITER) in msg, msg
    assert _static_rc(cfg, 16, capi.KRRT_STATIC_MAX_ITER + 1) == capi.KRRT_ERR_INVALID
    assert "beyond the static scratch" in L.krrt_last_error(None).decode()
    assert _static_rc(cfg, -1, 200) == capi.KRRT_ERR_INVALID
    assert _static_rc(None, 16, 200) == capi.KRRT_ERR_INVALID
    assert _static_rc(capi.default_config(precision=capi.KRRT_F32), 16, 200) == capi.KRRT_ERR_UNSUPPORTED
    assert _static_rc(capi.default_config(device=16), 16, 200) == capi.KRRT_ERR_INVALID
    assert _static_rc(cfg, 0, 200) == capi.KRRT_OK          # an empty batch is fine and is no work
    assert _static_rc(cfg, 16, 200) == capi.KRRT_ERR_INVALID  # null buffers


# three batches that leave different things behind in the shared array:
#   A  the soccer-field configuration (C4's shape): 1000 queries, nothing on the redo list
#   B  a near radius that covers the whole field: near sets outgrow k_plan_batch2's registers, so the redo list and the
#      redo pass's own work counter are used; fewer queries, fewer iterations, another radius table
#   C  no obstacles, tolerance goal rule, 7 queries: smaller than anything before it
CASES = {
    "A": dict(q=1000, n_iter=200, m=9, dynamic=1, base=0, cfg=dict(p_direct=0.1)),
    "B": dict(q=300, n_iter=150, m=9, dynamic=0, base=5000, cfg=dict(p_direct=0.05, gamma=60.0, rho_max=30.0, seed=7)),
    "C": dict(q=7, n_iter=64, m=0, dynamic=0, base=11, cfg=dict(goal_mode=capi.KRRT_GOAL_TOLERANCE, goal_tol=3.0)),
}


def _inputs(case, seed):
    starts, goals, obs = scenes.batch_queries(case["q"], max(case["m"], 1), seed=seed)
    return starts, goals, np.ascontiguousarray(obs[:, :case["m"]])


@pytest.mark.gpu
def test_static_calls_in_a_row_equal_the_context_planner(gpu_required):
    import torch
    dev = torch.device("cuda", 0)
    rec = capi.RESULT_DTYPE.itemsize
    want, bufs = {}, {}
    for seed, (name, case) in enumerate(CASES.items()):
        starts, goals, obs = _inputs(case, seed)
        d = [torch.from_numpy(x).to(dev) if x.size else torch.zeros(8, dtype=torch.float64, device=dev)
             for x in (starts, goals, obs)]
        # the existing entry point, on its own fresh context
        G = capi.Planner(**case["cfg"])
        d_ref = torch.zeros(case["q"] * rec, dtype=torch.uint8, device=dev)
        torch.cuda.synchronize()   # (the context's stream does not wait for torch's)
        G.plan_batch_device(d[0].data_ptr(), d[1].data_ptr(), d[2].data_ptr(), case["q"], case["m"], case["dynamic"],
                            case["n_iter"], case["base"], d_ref.data_ptr())
        G.synchronize()
        want[name] = d_ref.cpu().numpy().tobytes()
        if name == "B":   # the case is only worth its place if the redo pass really planned something
            # (the radius covers the field, so a tree of more than 65 nodes has had a near set of more than 64)
            assert d_ref.cpu().numpy().view(capi.RESULT_DTYPE)["n_nodes"].min() > 66
        # the caller's buffers for the static call: all of them exist before the first call is made
        bufs[name] = (d, torch.full((case["q"] * rec,), 0xA5, dtype=torch.uint8, device=dev))
        del G
    st = torch.cuda.Stream(device=dev)
    torch.cuda.synchronize()

    def run(name):
        case = CASES[name]
        d, d_out = bufs[name]
        d_out.fill_(0xA5)     # (no allocation: in place)
        torch.cuda.synchronize()
        capi.plan_batch_static(capi.default_config(**case["cfg"]), d[0].data_ptr(), d[1].data_ptr(), d[2].data_ptr(),
                               case["q"], case["m"], case["dynamic"], case["n_iter"], case["base"], d_out.data_ptr(),
                               st.cuda_stream)
        st.synchronize()

    def check(name):
        got = bufs[name][1].cpu().numpy()
        assert got.tobytes() == want[name], f"batch {name}: krrt_plan_batch_static differs from krrt_plan_batch_device"
        return got.view(capi.RESULT_DTYPE)

    run("A")                                   # (first call on the device: also the symbol look-up and the module load)
    assert int(check("A")["solved"].sum()) > 0
    free0 = torch.cuda.mem_get_info(dev)[0]
    for name in ("B", "A", "C", "B", "B"):     # back to back, each on what the one before left in the array
        run(name)
        assert torch.cuda.mem_get_info(dev)[0] == free0, "the call changed the amount of free device memory"
        check(name)
    # two calls queued on one stream with no host synchronisation between them are ordered by the stream
    for name in ("A", "C"):
        bufs[name][1].fill_(0xA5)
    torch.cuda.synchronize()
    for name in ("A", "C"):
        case = CASES[name]
        d, d_out = bufs[name]
        capi.plan_batch_static(capi.default_config(**case["cfg"]), d[0].data_ptr(), d[1].data_ptr(), d[2].data_ptr(),
                               case["q"], case["m"], case["dynamic"], case["n_iter"], case["base"], d_out.data_ptr(),
                               st.cuda_stream)
    st.synchronize()
    check("A")
    check("C")


@pytest.mark.gpu
def test_static_default_stream_and_largest_batch(gpu_required):
    """cuda_stream = NULL queues on the default stream; the largest tree the scratch allows plans like the context's"""
    import torch
    dev = torch.device("cuda", 0)
    rec = capi.RESULT_DTYPE.itemsize
    q, n_iter = 40, capi.KRRT_STATIC_MAX_ITER
    starts, goals, obs = scenes.batch_queries(q, 9, seed=3)
    d = [torch.from_numpy(x).to(dev) for x in (starts, goals, obs)]
    G = capi.Planner(p_direct=0.1)
    d_ref = torch.zeros(q * rec, dtype=torch.uint8, device=dev)
    torch.cuda.synchronize()
    G.plan_batch_device(d[0].data_ptr(), d[1].data_ptr(), d[2].data_ptr(), q, 9, 1, n_iter, 0, d_ref.data_ptr())
    G.synchronize()
    d_out = torch.zeros(q * rec, dtype=torch.uint8, device=dev)
    torch.cuda.synchronize()
    for _ in range(2):
        d_out.zero_()
        torch.cuda.synchronize()
        capi.plan_batch_static(capi.default_config(p_direct=0.1), d[0].data_ptr(), d[1].data_ptr(), d[2].data_ptr(), q, 9,
                               1, n_iter, 0, d_out.data_ptr(), 0)
        torch.cuda.synchronize()
        assert d_out.cpu().numpy().tobytes() == d_ref.cpu().numpy().tobytes()
