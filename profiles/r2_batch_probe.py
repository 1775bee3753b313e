"""C4 timing probe of the batched planner: device time of krrt_plan_batch_device for 4096 x 200, per goal mode."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
from robosoccer_motion_planning_b200 import capi, scenes

q, it = 4096, 200
bs, bg, bo = scenes.batch_queries(q, 9, seed=0)
dev = torch.device("cuda", 0)
d_s, d_g, d_o = (torch.from_numpy(x).to(dev) for x in (bs, bg, bo))
item = capi.RESULT_DTYPE.itemsize
d_out = torch.zeros(q * item, dtype=torch.uint8, device=dev)
st = torch.cuda.Stream(device=dev)
res = {}
for mode in (0, 1):
    G = capi.Planner(p_direct=0.1, goal_mode=mode, goal_tol=0.5)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ms = []
    for k in range(6):
        with torch.cuda.stream(st):
            e0.record(st)
            G.plan_batch_device(d_s.data_ptr(), d_g.data_ptr(), d_o.data_ptr(), q, 9, 1, it, 0, d_out.data_ptr(), st.cuda_stream)
            e1.record(st)
        e1.synchronize()
        ms.append(e0.elapsed_time(e1))
    out = d_out.cpu().numpy().view(capi.RESULT_DTYPE)
    res[f"goal_mode_{mode}"] = {"ms": ms, "solved": float(out["solved"].mean()), "mean_cost": float(out["cost"][out["solved"] != 0].mean())}
print(json.dumps(res))
