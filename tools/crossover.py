#!/usr/bin/env python
"""BASELINE.json configs[3] (C4): latency mode.  For batch sizes 1..65536 (and beyond, until the GPU
wins or 2^24) time one quota evaluation end to end through the host-pointer C ABI — H2D, launch,
D2H and sync included, quota tables already resident — from pageable and from pinned host memory,
next to the CPU oracle port (1 thread and all cores) on the same batch.  Reports median / p99 and
the crossover batch size (smallest N where the GPU median beats the all-core CPU median).
Also times the policy decision (K0 vs the restated mutate()).  Writes profiles/<tag>_crossover.json/.md.
"""
import argparse
import importlib
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle  # noqa: E402  (the CPU arm being timed beside the GPU)
from oracle import quota_ref as Q  # noqa: E402
from workloads import KINDS, UNITS, make_hard  # noqa: E402

bgc = importlib.import_module("bacchus-gpu-controller_b200")


def stats(fn, reps):
    for _ in range(min(20, reps)):
        fn()
    ts = np.empty(reps)
    for i in range(reps):
        t0 = time.perf_counter()
        fn()
        ts[i] = time.perf_counter() - t0
    ts *= 1e6
    return float(np.median(ts)), float(np.percentile(ts, 99))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--tag", default="r01")
    ap.add_argument("--max-log2", type=int, default=24)
    args = ap.parse_args()
    cores = os.cpu_count() or 1
    eng = bgc.Engine(0)
    out = {"cores": cores, "rows": [], "policy_rows": [], "note": "times in microseconds per call, python ctypes call overhead (~2 us) included on both sides"}
    R = 2
    for NS in (100, 10_000):
        a, b = Q.scramble_params(NS)
        nmax = 1 << args.max_log2
        run = oracle.gen_fill(oracle.gen_params(0xB2000004, 0, NS, 0, 0, nmax, a, b, None, KINDS[R]), 0, 1 << 20, threads=cores)
        used = oracle.usage_reduce(run[0], run[1], NS, threads=cores)
        hard, hm = make_hard(used, 4)
        eng.quota_table_set(hard, hm, UNITS[R])
        eng.usage_set(used)
        cand = oracle.gen_fill(oracle.gen_params(0xB2000004, 1, NS, 0, 0, nmax, a, b, None, KINDS[R]), 0, nmax, threads=cores)
        pin_ns = torch.from_numpy(cand[0].view(np.int32)).pin_memory()
        pin_req = [torch.from_numpy(c).pin_memory() for c in cand[1]]
        pin_v = torch.empty(nmax, dtype=torch.uint8).pin_memory()
        crossed = {"pinned": None, "pageable": None}
        for lg in range(0, args.max_log2 + 1):
            n = 1 << lg
            reps = 1000 if n <= 4096 else 200 if n <= 65536 else 30 if n <= (1 << 20) else 8
            ns = cand[0][:n]; req = [c[:n] for c in cand[1]]
            want = oracle.quota_eval(ns, req, None, NS, used, hard, hm, 2)
            got = eng.quota_eval(ns, req)
            assert np.array_equal(got, want)
            eng.quota_eval_ptr(pin_ns.data_ptr(), [t.data_ptr() for t in pin_req], 0, n, pin_v.data_ptr())
            assert np.array_equal(pin_v[:n].numpy(), want)
            g_page = stats(lambda: eng.quota_eval(ns, req), reps)
            ptrs = [t.data_ptr() for t in pin_req]
            g_pin = stats(lambda: eng.quota_eval_ptr(pin_ns.data_ptr(), ptrs, 0, n, pin_v.data_ptr()), reps)
            c1 = stats(lambda: oracle.quota_eval(ns, req, None, NS, used, hard, hm, 2, threads=1), reps)
            call = stats(lambda: oracle.quota_eval(ns, req, None, NS, used, hard, hm, 2, threads=cores), reps)
            best_cpu = min(c1[0], call[0])
            row = {"NS": NS, "n": n, "gpu_pinned_med": g_pin[0], "gpu_pinned_p99": g_pin[1], "gpu_pageable_med": g_page[0],
                   "gpu_pageable_p99": g_page[1], "cpu_1t_med": c1[0], "cpu_1t_p99": c1[1], "cpu_all_med": call[0], "cpu_all_p99": call[1]}
            out["rows"].append(row)
            print(json.dumps(row), flush=True)
            if crossed["pinned"] is None and g_pin[0] < best_cpu:
                crossed["pinned"] = n
            if crossed["pageable"] is None and g_page[0] < best_cpu:
                crossed["pageable"] = n
            if lg >= 16 and crossed["pinned"] and crossed["pageable"] and lg >= 20:
                break
        out[f"crossover_NS{NS}"] = crossed
    # policy decision (K0) vs the restated mutate()
    nmax = 1 << 22
    op, fl = oracle.gen_policy_fill(0xB2000001, 0, nmax, threads=cores)
    crossed = None
    for lg in range(0, 23):
        n = 1 << lg
        reps = 1000 if n <= 4096 else 200 if n <= 65536 else 30
        o, f = op[:n], fl[:n]
        g = stats(lambda: eng.policy_eval(o, f), reps)
        c1 = stats(lambda: oracle.policy_bits(o, f, 1), reps)
        call = stats(lambda: oracle.policy_bits(o, f, cores), reps)
        row = {"n": n, "gpu_pageable_med": g[0], "gpu_pageable_p99": g[1], "cpu_1t_med": c1[0], "cpu_all_med": call[0]}
        out["policy_rows"].append(row)
        print(json.dumps(row), flush=True)
        if crossed is None and g[0] < min(c1[0], call[0]):
            crossed = n
    out["crossover_policy"] = crossed
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(out, open(os.path.join(ROOT, "gpurun_out", f"{args.tag}_crossover.json"), "w"), indent=1)
    with open(os.path.join(ROOT, "gpurun_out", f"{args.tag}_crossover.md"), "w") as f:
        f.write(f"# C4 latency sweep (quota_eval end to end, tables resident; {cores} host cores; median us / p99 us)\n\n")
        for NS in (100, 10_000):
            f.write(f"\n## NS = {NS}  — crossover (GPU median < best CPU median): {out[f'crossover_NS{NS}']}\n\n")
            f.write("| N | GPU pinned | GPU pageable | CPU 1 thread | CPU all cores |\n|---|---|---|---|---|\n")
            for r in out["rows"]:
                if r["NS"] == NS:
                    f.write(f"| {r['n']} | {r['gpu_pinned_med']:.1f} / {r['gpu_pinned_p99']:.1f} | {r['gpu_pageable_med']:.1f} / {r['gpu_pageable_p99']:.1f} | {r['cpu_1t_med']:.1f} / {r['cpu_1t_p99']:.1f} | {r['cpu_all_med']:.1f} / {r['cpu_all_p99']:.1f} |\n")
        f.write(f"\n## policy decision (K0 vs restated mutate()) — crossover: {out['crossover_policy']}\n\n| N | GPU pageable | CPU 1 thread | CPU all cores |\n|---|---|---|---|\n")
        for r in out["policy_rows"]:
            f.write(f"| {r['n']} | {r['gpu_pageable_med']:.1f} / {r['gpu_pageable_p99']:.1f} | {r['cpu_1t_med']:.1f} | {r['cpu_all_med']:.1f} |\n")
    print("crossover:", {k: v for k, v in out.items() if k.startswith("crossover")})


if __name__ == "__main__":
    main()
