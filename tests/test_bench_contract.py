"""The bench.py contract on the CPU side: `--impl reference` (the CPU arm, no GPU needed) prints ONE JSON line with
the keys the driver reads, the same `config` the GPU arm prints for the same command line, and a whole-workload
cpu_baseline; the helpers that decide the compact host columns are lossless on the generator's values."""
import json
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_prints_the_contract_line():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--pods", "300000", "--steps", "2", "--warmup", "1"],
                         capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.strip().startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    for k in ("impl", "metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline",
              "dtype", "data", "config", "cpu_baseline", "e2e"):
        assert k in d, k
    assert d["impl"] == "reference" and d["metric"] == "pods_evaluated_per_s" and d["unit"] == "pods/s" and d["higher_is_better"] is True
    assert d["vs_baseline"] is None and d["dtype"] == "int64" and d["steps"] == 2
    assert d["e2e"] == {"value": d["value"], "unit": "pods/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == d["value"] and "whole workload" in cb["sample"]
    assert d["config"]["pods_per_gpu"] == 300000 and d["config"]["namespaces_per_gpu"] == 10000 and d["config"]["resources"] == 3
    assert abs(d["value"] - 300000 / (d["ms_per_step"] * 1e-3)) / d["value"] < 1e-6


def test_both_arms_share_one_config_and_the_compact_columns_are_lossless(oracle):
    sys.path.insert(0, ROOT)
    import bench
    from oracle import quota_ref as Q
    from workloads import KINDS
    n, NS, R = 50_000, 10_000, 3
    cfg = bench.make_config("C3", n, NS, R)
    assert set(cfg) == {"workload", "pods_per_gpu", "namespaces_per_gpu", "resources", "step", "running_set", "l2", "e2e_host_columns", "sharding"}
    ns_t, plan, bpp = bench.narrow_plan(NS, R)
    assert ns_t is np.uint16 and bpp == 9
    assert bench.narrow_plan(70_000, R)[0] is np.uint32                    # more than 65,536 namespaces: u32 ids
    a, b = Q.scramble_params(NS)
    cdf = Q.zipf_cdf(NS, 1.1)
    ns, req, _ = oracle.gen_fill(oracle.gen_params(bench.SEED["c3"], 0, NS, 1, 0, n, a, b, cdf, KINDS[R]), 0, n)
    ns2, cols, shifts = bench.narrow_numpy(ns, req, NS, R)
    assert np.array_equal(ns2.astype(np.uint32), ns)
    for c, d, s in zip(req, cols, shifts):
        assert np.array_equal(d.astype(np.int64) << s, c)
    # and the CPU arm on the compact columns decides what it decides on the int64 columns
    arm = bench.CpuArm(oracle, NS, R, 2)
    arm.prepare(ns2, cols, shifts, 7)
    _, u1, v1 = arm.step_typed(ns2, cols, ns2, cols, shifts)
    _, u2, v2 = arm.step_i64(ns, req, ns, req)
    assert np.array_equal(u1, u2) and np.array_equal(v1, v2)
