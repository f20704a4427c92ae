#!/usr/bin/env python
"""C1 (BASELINE.json configs[0]) measured: an in-process replay of 10k AdmissionReviews, wire bytes in,
response JSON out, no cluster, timed per stage on both arms.

  (1a) 10k UserBootstrap reviews — the only thing the reference's admission handler can replay
       (src/admission.rs:184-204 -> mutate :241-431): parse -> flatten (row A2) -> decide -> render.
       The decide stage is K0 through bgc_policy_eval (host pointers: copies + launch + sync inside) on the
       GPU arm and policy_ref_bits_batch (oracle/policy_ref.c) on the CPU arm; parse, flatten and render are
       host C++ (csrc/bgc_host.cpp, called through ctypes) and are the SAME on both arms.
  (1b) 10k Pod-create reviews against 100 namespace quotas (not a reference path):
       parse -> flatten (row A5) -> verdicts -> render; verdicts = K2 (bgc_quota_eval, host pointers) vs
       quota_ref_quota_eval, usage over 10k running pods = K1 vs quota_ref_usage_reduce.

Prints one JSON document.  Expectation (north_star): at this size the CPU wins the decide stage by a wide
margin and the stage itself is a rounding error next to JSON parsing.
"""
import importlib
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle  # noqa: E402
from oracle import policy_ref as P  # noqa: E402
from workloads import C1_KEYS, c1_namespaces, c1_pods  # noqa: E402

bgc = importlib.import_module("bacchus-gpu-controller_b200")


def med(fn, reps=21):
    fn()
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter()
        fn()
        ts.append(time.perf_counter() - t0)
    ts.sort()
    return ts[len(ts) // 2]


def main():
    n = 10_000
    eng = bgc.Engine(0)
    cores = len(os.sched_getaffinity(0))
    out = {"config": "C1: in-process replay of 10k AdmissionReviews, no cluster, no TLS, no HTTP", "requests": n, "host_cores": cores}

    # ---------------- (1a) UserBootstrap reviews
    cfg_py, cfg = P.Config(), bgc.admission.AdmissionConfig()
    base = []
    for op in range(4):
        for fl in range(1024):
            r = P.request_from_bits(op, fl, cfg_py)
            if r is not None:
                base.append(r)
    rng = np.random.default_rng(1)
    bodies = []
    for i in range(n):
        r = dict(base[int(rng.integers(0, len(base)))])
        r["uid"] = f"uid-{i}"
        bodies.append(json.dumps({"apiVersion": "admission.k8s.io/v1", "kind": "AdmissionReview", "request": r}).encode())
    wire_bytes = sum(len(b) for b in bodies)
    t_parse = med(lambda: [bgc.review.Review(b) for b in bodies], 5)
    reviews = [bgc.review.Review(b) for b in bodies]
    reqs = [bgc.admission.AdmissionRequest.from_review(json.loads(b)["request"]) for b in bodies]
    t_flatten = med(lambda: bgc.admission.flatten_requests(cfg, reqs), 5)
    op, flags = bgc.admission.flatten_requests(cfg, reqs)
    t_gpu = med(lambda: eng.policy_eval(op, flags))
    t_cpu1 = med(lambda: oracle.policy_bits(op, flags, threads=1))
    t_cpuN = med(lambda: oracle.policy_bits(op, flags, threads=cores))
    code, pm = eng.policy_eval(op, flags)
    rc, rp = oracle.policy_bits(op, flags)
    same = bool(np.array_equal(code, rc) and np.array_equal(pm, rp))
    t_e2e_gpu = med(lambda: bgc.review.handle_reviews(eng, cfg, bodies), 9)
    t_e2e_step = med(lambda: bgc.review.handle_reviews_stepwise(eng, cfg, bodies), 5)
    one_call = {}
    bgc.review.handle_reviews(eng, cfg, bodies, stage_ms=one_call)
    out["userbootstrap_reviews"] = {
        "wire_bytes": wire_bytes, "gpu_equals_cpu": same,
        "stage_ms": {"parse (C++ parser, one ctypes call per body)": t_parse * 1e3, "flatten strings -> predicate bits (C++, one call)": t_flatten * 1e3,
                     "decide GPU: bgc_policy_eval from host memory (copies, K0 launch, sync)": t_gpu * 1e3,
                     "decide CPU: policy_ref_bits_batch, 1 thread": t_cpu1 * 1e3, f"decide CPU: policy_ref_bits_batch, {cores} threads": t_cpuN * 1e3},
        "wire_to_wire_ms_gpu_arm (review.handle_reviews = ONE C call, bgc_handle_reviews: parse + flatten + K0 + patch + render; ctypes marshalling of the bodies included)": t_e2e_gpu * 1e3,
        "one_call_stage_ms (host clock inside bgc_handle_reviews)": one_call,
        "wire_to_wire_ms_call_by_call (review.handle_reviews_stepwise: the same five C calls made from Python per body)": t_e2e_step * 1e3,
        "reviews_per_s_wire_to_wire_gpu_arm": n / t_e2e_gpu,
        "decide_stage_reviews_per_s": {"gpu": n / t_gpu, "cpu_1_thread": n / t_cpu1, "cpu_all": n / t_cpuN},
        "who_wins_the_decide_stage": "CPU" if min(t_cpu1, t_cpuN) < t_gpu else "GPU",
    }
    del reviews

    # ---------------- (1b) Pod-create reviews against 100 namespace quotas
    nss = c1_namespaces()
    m = bgc.pods.NamespaceMap()
    for ns in nss:
        m.intern(ns["name"])
    R = len(C1_KEYS)
    hard = np.zeros((R, len(nss)), np.int64)
    hm = np.zeros(len(nss), np.uint8)
    for i, ns in enumerate(nss):
        h, unit = bgc.pods.hard_from_sheet_row(ns["cpu"], ns["memory"], ns["gpu"], ns["storage"], ns["mig"], C1_KEYS)
        hard[:, i] = h
        hm[i] = (1 << R) - 1 if ns["has_quota"] else 0
    eng.quota_table_set(hard, hm, unit)
    running = c1_pods(n, nss, running=True)
    r_ns, r_req, _, _ = bgc.pods.flatten_pods(running, C1_KEYS, m, intern=False)
    pods = c1_pods(n, nss, running=False)
    pbodies = [json.dumps({"apiVersion": "admission.k8s.io/v1", "kind": "AdmissionReview",
                           "request": {"uid": f"u{i}", "operation": "CREATE", "kind": {"kind": "Pod"}, "namespace": p["metadata"]["namespace"], "object": p}}).encode()
               for i, p in enumerate(pods)]
    t_pparse = med(lambda: [bgc.review.Review(b) for b in pbodies], 5)
    previews = [bgc.review.Review(b) for b in pbodies]
    t_pflat = med(lambda: bgc.review.flatten_pod_reviews(None, C1_KEYS, m, reviews=previews), 5)
    c_ns, c_req, c_mask, _ = bgc.review.flatten_pod_reviews(None, C1_KEYS, m, reviews=previews)
    cm = sum(1 << r for r, u in enumerate(unit) if u)
    t_k1 = med(lambda: eng.usage_reduce(r_ns, r_req))
    t_c1 = med(lambda: oracle.usage_reduce(r_ns, r_req, len(nss)))
    used = eng.usage_reduce(r_ns, r_req)
    t_k2 = med(lambda: eng.quota_eval(c_ns, c_req, c_mask))
    t_c2 = med(lambda: oracle.quota_eval(c_ns, c_req, c_mask, len(nss), used, hard, hm, cm))
    t_c2N = med(lambda: oracle.quota_eval(c_ns, c_req, c_mask, len(nss), used, hard, hm, cm, threads=cores))
    same = bool(np.array_equal(used, oracle.usage_reduce(r_ns, r_req, len(nss))) and
                np.array_equal(eng.quota_eval(c_ns, c_req, c_mask), oracle.quota_eval(c_ns, c_req, c_mask, len(nss), used, hard, hm, cm)))
    t_pe2e = med(lambda: bgc.review.handle_pod_reviews(eng, pbodies, C1_KEYS, m, used, hard), 9)
    t_pstep = med(lambda: bgc.review.handle_pod_reviews_stepwise(eng, pbodies, C1_KEYS, m, used, hard), 3)
    p_one = {}
    bgc.review.handle_pod_reviews(eng, pbodies, C1_KEYS, m, used, hard, stage_ms=p_one)
    out["pod_reviews"] = {
        "wire_bytes": sum(len(b) for b in pbodies), "namespaces": len(nss), "running_pods": int(r_ns.shape[0]), "gpu_equals_cpu": same,
        "stage_ms": {"parse (C++ parser, one ctypes call per body)": t_pparse * 1e3, "flatten pod specs -> columns (C++, row A5)": t_pflat * 1e3,
                     "usage GPU: bgc_usage_reduce from host memory": t_k1 * 1e3, "usage CPU: quota_ref_usage_reduce, 1 thread": t_c1 * 1e3,
                     "verdicts GPU: bgc_quota_eval from host memory": t_k2 * 1e3, "verdicts CPU: quota_ref_quota_eval, 1 thread": t_c2 * 1e3,
                     f"verdicts CPU: quota_ref_quota_eval, {cores} threads": t_c2N * 1e3},
        "wire_to_wire_ms_gpu_arm (review.handle_pod_reviews = ONE C call, bgc_handle_pod_reviews; ctypes marshalling of the bodies included)": t_pe2e * 1e3,
        "one_call_stage_ms (host clock inside bgc_handle_pod_reviews)": p_one,
        "wire_to_wire_ms_call_by_call (review.handle_pod_reviews_stepwise, Python glue per body)": t_pstep * 1e3,
        "reviews_per_s_wire_to_wire_gpu_arm": n / t_pe2e,
        "verdict_stage_pods_per_s": {"gpu": n / t_k2, "cpu_1_thread": n / t_c2, "cpu_all": n / t_c2N},
        "who_wins_the_verdict_stage": "CPU" if min(t_c2, t_c2N) < t_k2 else "GPU",
    }
    print(json.dumps(out, indent=1))
    eng.close()


if __name__ == "__main__":
    main()
