#!/usr/bin/env python
"""Round-2 probe: K1 scatter kernel (hash and direct sinks) over launch shapes and L2-prefetch distances, and
K2 over prefetch distances; one JSON line per case.  python tools/k1_hash_probe.py [--pods N] > out.jsonl"""
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


def timeit(fn, reps=7, warm=2):
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
    ap.add_argument("--shapes", default="0,1,3")
    ap.add_argument("--pf", default="0,1,2,4")
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--cases", default="", help="NS:dist[,NS:dist...] subset, dist 0 uniform / 1 zipf")
    args = ap.parse_args()
    n = args.pods
    shapes = [int(x) for x in args.shapes.split(",")]
    pfs = [int(x) for x in args.pf.split(",")]
    eng = bgc.Engine(0)
    eng.use_torch_stream()
    cases = [(1_000_000, 3, 1, 1.1), (20_000, 3, 1, 1.1), (125_000, 3, 1, 1.1), (500_000, 3, 1, 1.1), (10_000, 3, 1, 1.1),
             (1_000_000, 3, 0, 0.0), (125_000, 3, 0, 0.0), (10_000, 2, 0, 0.0), (10_000, 3, 0, 0.0)]
    if args.quick:
        cases = cases[:2]
    if args.cases:
        want = [tuple(int(y) for y in x.split(":")) for x in args.cases.split(",")]   # NS:dist pairs
        cases = [c for c in cases if (c[0], c[2]) in want]
    for NS, R, dist, s in cases:
        a, b = Q.scramble_params(NS)
        cdf = Q.zipf_cdf(NS, s) if dist else None
        d_cdf = torch.from_numpy(cdf.view(np.int64)).cuda() if dist else None
        ns = torch.empty(n, dtype=torch.int32, device="cuda")
        req = [torch.empty(n, dtype=torch.int64, device="cuda") for _ in range(R)]
        eng.synth_pods_dev(bgc.synth_params(0xB2000003, 0, NS, dist, 0, n, a, b, d_cdf, KINDS[R]), 0, n, ns, req)
        eng.quota_table_set(np.zeros((R, NS), np.int64), np.zeros(NS, np.uint8), UNITS[R])
        ref = None
        for shape in shapes:
            for pf in pfs:
                eng.set_k1_variant(0 | (shape << 8) | ((pf + 1) << 16))
                ms = timeit(lambda: eng.usage_reduce_dev(ns, req, 0))
                used = eng.usage_get()
                if ref is None:
                    ref = used
                gbs = n * (4 + 8 * R) / ms / 1e6
                print(json.dumps({"kernel": "k1", "NS": NS, "R": R, "dist": "zipf%.1f" % s if dist else "uniform", "variant": eng.last_k1_variant(),
                                  "shape": shape, "pf": pf, "ms": round(ms, 4), "GBps": round(gbs, 1), "frac": round(gbs / PEAK, 3),
                                  "same_as_first": bool(np.array_equal(used, ref))}), flush=True)
        hard, hm = make_hard(ref, 1)
        eng.quota_table_set(hard, hm, UNITS[R])
        eng.usage_set(ref)
        v = torch.empty(n, dtype=torch.uint8, device="cuda")
        vref = None
        for pf in pfs:
            eng.set_k1_variant((pf + 1) << 24)
            ms = timeit(lambda: eng.quota_eval_dev(ns, req, None, v))
            cs = eng.checksum_u8_dev(v)
            if vref is None:
                vref = cs
            gbs = n * (4 + 8 * R + 1) / ms / 1e6
            print(json.dumps({"kernel": "k2", "NS": NS, "R": R, "dist": "zipf%.1f" % s if dist else "uniform", "pf": pf, "ms": round(ms, 4),
                              "GBps": round(gbs, 1), "frac": round(gbs / PEAK, 3), "same_as_first": cs == vref}), flush=True)
        del ns, req, v


if __name__ == "__main__":
    main()
