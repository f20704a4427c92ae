#!/usr/bin/env python
"""Times every K1 variant (and K2, K0) on device-generated workloads; prints one JSON line per case.
Development aid for choosing the automatic K1 variant — numbers quoted in DESIGN.md come from here
and from bench.py.  Usage: python tools/k1_probe.py [--pods N]"""
import argparse
import importlib
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle import quota_ref as Q  # noqa: E402  (zipf table / scramble parameters only)
from workloads import KINDS, UNITS, make_hard  # noqa: E402

bgc = importlib.import_module("bacchus-gpu-controller_b200")
PEAK = 6538.3


def timeit(fn, reps=5, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    return ts[len(ts) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--pods", type=int, default=100_000_000)
    ap.add_argument("--cases", default="all")
    args = ap.parse_args()
    n = args.pods
    eng = bgc.Engine(0)
    eng.use_torch_stream()
    names = {0: "auto", 1: "global_red", 2: "smem_direct", 3: "smem_hash", 4: "cluster8", 5: "cluster16"}
    cases = [  # (NS, R, dist, s, grouped)
        (10_000, 3, 1, 1.1, 0), (10_000, 3, 0, 0, 0), (10_000, 3, 1, 1.1, 1), (10_000, 2, 0, 0, 0),
        (100, 3, 1, 1.1, 0), (1_000_000, 3, 1, 1.1, 0), (1_000_000, 3, 0, 0, 0), (10_000, 3, 1, 1.5, 0),
        (125_000, 3, 1, 1.1, 0), (500_000, 3, 1, 1.1, 0), (20_000, 3, 1, 1.1, 0), (50_000, 3, 0, 0, 0), (125_000, 3, 0, 0, 0), (200_000, 3, 1, 1.1, 0), (10_000, 4, 1, 1.1, 0), (10_000, 1, 1, 1.1, 0), (5_000, 6, 1, 1.1, 0), (10_001, 6, 1, 1.1, 0), (10_001, 8, 0, 0, 0),
    ]
    if args.cases != "all":
        cases = [c for c in cases if str(c[0]) in args.cases.split(",")]
    for NS, R, dist, s, grouped in cases:
        a, b = Q.scramble_params(NS)
        cdf = Q.zipf_cdf(NS, s) if dist else None
        d_cdf = torch.from_numpy(cdf.view(np.int64)).cuda() if dist else None
        ns = torch.empty(n, dtype=torch.int32, device="cuda")
        req = [torch.empty(n, dtype=torch.int64, device="cuda") for _ in range(R)]
        eng.synth_pods_dev(bgc.synth_params(0xB2000003, 0, NS, dist, grouped, n, a, b, d_cdf, KINDS[R]), 0, n, ns, req)
        eng.quota_table_set(np.zeros((R, NS), np.int64), np.zeros(NS, np.uint8), UNITS[R])
        ref = None
        todo = [1, 2, 3] if grouped else [1] + [v | (sh << 8) for v in (2, 3) for sh in (0, 1)] + [4, 5]
        if not grouped and (dist and NS > 100):
            todo = todo[1:]          # global reds on a Zipf head take tens of ms: measured once, skip
        for variant in todo:
            eng.set_k1_variant(variant)
            flags = 1 if grouped else 0
            try:
                ms = timeit(lambda: eng.usage_reduce_dev(ns, req, flags))
            except Exception as e:  # noqa: BLE001
                print(json.dumps({"case": [NS, R, dist, s, grouped], "variant": names.get(variant & 0xFF), "error": str(e)[:80]}), flush=True)
                continue
            used = eng.usage_get()
            if ref is None:
                ref = used
            gbs = n * (4 + 8 * R) / ms / 1e6
            print(json.dumps({"kernel": "k1", "NS": NS, "R": R, "dist": "zipf%.1f" % s if dist else "uniform", "grouped": grouped,
                              "variant": eng.last_k1_variant(), "shape": variant >> 8, "ms": round(ms, 4), "GBps": round(gbs, 1), "frac": round(gbs / PEAK, 3),
                              "Gpods_s": round(n / ms / 1e6, 2), "same_as_first": bool(np.array_equal(used, ref))}), flush=True)
        # K2 on the same columns
        hard, hm = make_hard(ref, 1)
        eng.quota_table_set(hard, hm, UNITS[R])
        eng.usage_set(ref)
        v = torch.empty(n, dtype=torch.uint8 if R <= 4 else torch.int16, device="cuda")
        ms = timeit(lambda: eng.quota_eval_dev(ns, req, None, v))
        gbs = n * (4 + 8 * R + (1 if R <= 4 else 2)) / ms / 1e6
        print(json.dumps({"kernel": "k2", "NS": NS, "R": R, "dist": "zipf%.1f" % s if dist else "uniform", "ms": round(ms, 4),
                          "GBps": round(gbs, 1), "frac": round(gbs / PEAK, 3), "deny": float((v[: 1 << 20] != 0).float().mean())}), flush=True)
        del ns, req, v
    # K0
    m = 1 << 30
    op = torch.empty(m, dtype=torch.uint8, device="cuda"); fl = torch.empty(m, dtype=torch.int16, device="cuda")
    c = torch.empty(m, dtype=torch.uint8, device="cuda"); p = torch.empty(m, dtype=torch.uint8, device="cuda")
    eng.synth_policy_dev(1, 0, m, op, fl)
    ms = timeit(lambda: eng.policy_eval_dev(op, fl, c, p))
    print(json.dumps({"kernel": "k0", "n": m, "ms": round(ms, 4), "GBps": round(m * 5 / ms / 1e6, 1), "frac": round(m * 5 / ms / 1e6 / PEAK, 3)}), flush=True)


if __name__ == "__main__":
    main()
