#!/usr/bin/env python
"""K1/K2 over compact columns RESIDENT in HBM (bgc_usage_reduce_typed_dev / bgc_quota_eval_typed_dev) next to the
int64 entry points on the same C3 workload: u16 namespace ids, u8 gpu count, u16 cpu millicores, u32 memory MiB << 20
(9 B per pod instead of 28).  Results are compared with the int64 kernels' (which the test-suite holds to the oracle).
Usage: python tools/typed_probe.py [--pods N] [--reps K]   (also the target of the ncu captures of the typed kernels)"""
import argparse, importlib, json, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle import quota_ref as Q
from workloads import KINDS, UNITS, make_hard
bgc = importlib.import_module("bacchus-gpu-controller_b200")

ap = argparse.ArgumentParser(); ap.add_argument("--pods", type=int, default=100_000_000); ap.add_argument("--reps", type=int, default=7)
args = ap.parse_args()
n, NS, R = args.pods, 10_000, 3
eng = bgc.Engine(0); eng.use_torch_stream()
a, b = Q.scramble_params(NS)
cdf = Q.zipf_cdf(NS, 1.1); d_cdf = torch.from_numpy(cdf.view(np.int64)).cuda()
ns = torch.empty(n, dtype=torch.int32, device="cuda"); req = [torch.empty(n, dtype=torch.int64, device="cuda") for _ in range(R)]
eng.quota_table_set(np.zeros((R, NS), np.int64), np.zeros(NS, np.uint8), UNITS[R])
eng.synth_pods_dev(bgc.synth_params(0xB2000003, 0, NS, 1, 0, n, a, b, d_cdf, KINDS[R]), 0, n, ns, req)
eng.usage_reduce_dev(ns, req); torch.cuda.synchronize()
used = eng.usage_get()
hard, hm = make_hard(used, 3)
eng.quota_table_set(hard, hm, UNITS[R]); eng.usage_set(used)
v64 = torch.empty(n, dtype=torch.uint8, device="cuda"); vt = torch.empty(n, dtype=torch.uint8, device="cuda")
# compact forms (exactness checked: widening them back gives the int64 columns)
c_ns = ns.to(torch.int16)
cols = [req[0].to(torch.uint8), req[1].to(torch.int16), (req[2] >> 20).to(torch.int32)]
types, shifts = [bgc.T_U8, bgc.T_U16, bgc.T_U32], [0, 0, 20]
assert bool(((c_ns.to(torch.int32) & 0xFFFF) == ns).all()) and bool((cols[0].to(torch.int64) == req[0]).all())
assert bool(((cols[1].to(torch.int64) & 0xFFFF) == req[1]).all()) and bool((((cols[2].to(torch.int64) & 0xFFFFFFFF) << 20) == req[2]).all())


def med(fn):
    fn(); fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(args.reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    return sorted(ts)[len(ts) // 2]


k1_64 = med(lambda: eng.usage_reduce_dev(ns, req)); u64 = eng.usage_get()
k1_t = med(lambda: eng.usage_reduce_typed_dev(c_ns, cols, shifts, 0, None, types=types)); ut = eng.usage_get()
k2_64 = med(lambda: eng.quota_eval_dev(ns, req, None, v64))
k2_t = med(lambda: eng.quota_eval_typed_dev(c_ns, cols, shifts, None, vt, types=types))
same = bool(np.array_equal(u64, ut) and np.array_equal(u64, used) and bool((v64 == vt).all()))
print(json.dumps({"workload": "C3: 1e8 pods, Zipf(1.1) over 1e4 namespaces, R=3, columns resident in HBM", "pods": n,
                  "int64_columns": {"bytes_per_pod": [28, 29], "k1_ms": k1_64, "k2_ms": k2_64, "k1_GBps": n * 28 / k1_64 / 1e6, "k2_GBps": n * 29 / k2_64 / 1e6},
                  "compact_columns": {"bytes_per_pod": [9, 10], "k1_ms": k1_t, "k2_ms": k2_t, "k1_GBps": n * 9 / k1_t / 1e6, "k2_GBps": n * 10 / k2_t / 1e6},
                  "typed_equals_int64": same, "library": bgc.version()}))
eng.close()
