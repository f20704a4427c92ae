#!/usr/bin/env python
"""Times K0 policy_eval on 2^30 device-resident requests (5 B/request) — for ncu captures."""
import importlib, os, sys, json
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
bgc = importlib.import_module("bacchus-gpu-controller_b200")
m = 1 << 30
eng = bgc.Engine(0); eng.use_torch_stream()
op = torch.empty(m, dtype=torch.uint8, device="cuda"); fl = torch.empty(m, dtype=torch.int16, device="cuda")
c = torch.empty(m, dtype=torch.uint8, device="cuda"); p = torch.empty(m, dtype=torch.uint8, device="cuda")
eng.synth_policy_dev(1, 0, m, op, fl)
for _ in range(3): eng.policy_eval_dev(op, fl, c, p)
torch.cuda.synchronize()
ts = []
for _ in range(5):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); eng.policy_eval_dev(op, fl, c, p); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
ms = sorted(ts)[2]
print(json.dumps({"kernel": "k0_policy_eval", "n": m, "ms": ms, "GBps": m * 5 / ms / 1e6, "frac_of_6538": m * 5 / ms / 1e6 / 6538.3}))
