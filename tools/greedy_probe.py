#!/usr/bin/env python
"""Times K3 admit_greedy (row N1) on device-generated candidates next to the oracle's literal loop
(single thread: the recurrence is sequential per namespace) on a bounded sample."""
import argparse, importlib, json, os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle
from oracle import quota_ref as Q
from workloads import KINDS, UNITS, make_hard
bgc = importlib.import_module("bacchus-gpu-controller_b200")

ap = argparse.ArgumentParser(); ap.add_argument("--pods", type=int, default=100_000_000)
ap.add_argument("--verify", action="store_true", help="check EVERY verdict and the final usage table against the literal loop (about 10 s per case)")
args = ap.parse_args()
n = args.pods
eng = bgc.Engine(0); eng.use_torch_stream()
R = 3
for NS, dist, s in [(10_000, 0, 0.0), (10_000, 1, 1.1), (1_000_000, 0, 0.0), (1_000_000, 1, 1.1)]:
    a, b = Q.scramble_params(NS)
    cdf = Q.zipf_cdf(NS, s) if dist else None
    d_cdf = torch.from_numpy(cdf.view(np.int64)).cuda() if dist else None
    ns = torch.empty(n, dtype=torch.int32, device="cuda"); req = [torch.empty(n, dtype=torch.int64, device="cuda") for _ in range(R)]
    v = torch.empty(n, dtype=torch.uint8, device="cuda")
    eng.quota_table_set(np.zeros((R, NS), np.int64), np.zeros(NS, np.uint8), UNITS[R])
    eng.synth_pods_dev(bgc.synth_params(0xB2000003, 0, NS, dist, 0, n, a, b, d_cdf, KINDS[R]), 0, n, ns, req)
    eng.usage_reduce_dev(ns, req)
    torch.cuda.synchronize()
    used0 = eng.usage_get()
    hard, hm = make_hard(used0, 3)
    hard = hard + hard // 8
    eng.quota_table_set(hard, hm, UNITS[R])
    eng.synth_pods_dev(bgc.synth_params(0xB2000003, 1, NS, dist, 0, n, a, b, d_cdf, KINDS[R]), 0, n, ns, req)
    ts = []
    for it in range(4):
        eng.usage_set(used0)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); eng.admit_greedy_dev(ns, req, None, v); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ms = sorted(ts[1:])[1]
    adm = float((v[: 1 << 22] == 0).float().mean())
    # CPU literal loop on a bounded sample of the same stream
    m = min(n, 1 << 23)
    c = oracle.gen_fill(oracle.gen_params(0xB2000003, 1, NS, dist, 0, n, a, b, cdf, KINDS[R]), 0, m, threads=8)
    t0 = time.perf_counter(); wv, wu = oracle.admit_greedy(c[0], c[1], None, NS, used0, hard, hm, 6); cpu_s = time.perf_counter() - t0
    same = bool(np.array_equal(wv[: 1 << 16], v[: 1 << 16].cpu().numpy())) if m == n else None
    full = None
    if args.verify:     # the device's own columns, every pod, through the literal loop
        fv, fu = oracle.admit_greedy(ns.cpu().numpy().view(np.uint32), [c.cpu().numpy() for c in req], None, NS, used0, hard, hm, 6)
        full = bool(np.array_equal(fv, v.cpu().numpy()) and np.array_equal(fu, eng.usage_get()))
    print(json.dumps({"kernel": "k3_admit_greedy", "n": n, "NS": NS, "dist": "zipf%.1f" % s if dist else "uniform", "ms": round(ms, 3),
                      "Mpods_s": round(n / ms / 1e3, 1), "admitted_frac": round(adm, 3), "cpu_literal_loop_Mpods_s": round(m / cpu_s / 1e6, 1),
                      "prefix_matches_oracle": same, "every_verdict_and_usage_match_the_literal_loop": full,
                      "long_segments": os.environ.get("BGC_K3_LONG", "chain")}), flush=True)
    del ns, req, v
