"""BASELINE.json configs[0] (C1): in-process replay, no cluster, of
  (1a) 10k UserBootstrap AdmissionReviews through the batched mutate() (the only thing the reference
       admission handler can replay: src/admission.rs:184-431) — covered in test_gpu_policy.py; and
  (1b) 10k Pod-create records against 100 namespace quotas: pod JSON → columns by the C++ flattener
       (row A5), usage over the running pods (K1), verdicts for the candidates (K2), all compared with
       an independent pure-Python evaluation of the same pod objects.  Not a reference path: the
       reference never evaluates pods."""
import json
from fractions import Fraction

import numpy as np
import pytest

from workloads import C1_KEYS, c1_namespaces, c1_pods

pytestmark = pytest.mark.gpu

_SUFFIX = {"": Fraction(1), "m": Fraction(1, 1000), "k": Fraction(1000), "M": Fraction(10**6), "G": Fraction(10**9),
           "Ki": Fraction(1 << 10), "Mi": Fraction(1 << 20), "Gi": Fraction(1 << 30)}


def _quantity(s: str) -> Fraction:
    """Tiny exact Kubernetes-quantity reader for the strings the C1 generator emits."""
    if "e" in s:                       # decimal-exponent form (no SI suffix allowed with it)
        mant, exp = s.split("e")
        return Fraction(mant) * Fraction(10) ** int(exp)
    i = len(s)
    while i and not (s[i - 1].isdigit() or s[i - 1] == "."):
        i -= 1
    return Fraction(s[:i]) * _SUFFIX[s[i:]]


def _ceil(f: Fraction) -> int:
    return -((-f.numerator) // f.denominator)


def _pod_usage(pod, key):
    """max(sum of containers, max of init containers) + overhead, in canonical units (python ints)."""
    res = key.split(".", 1)[1] if key.startswith(("requests.", "limits.")) else key
    milli = res == "cpu"
    def val(d):
        q = (d or {}).get(res)
        return None if q is None else _ceil(_quantity(q) * (1000 if milli else 1))
    spec = pod["spec"]
    app = [val(c["resources"].get("requests")) for c in spec["containers"]]
    init = [val(c["resources"].get("requests")) for c in spec.get("initContainers", [])]
    stated = all(v is not None for v in app + init)
    total = max(sum(v or 0 for v in app), max([v or 0 for v in init], default=0)) + (val(spec.get("overhead")) or 0)
    return total, stated


def test_c1_pod_reviews_against_100_namespace_quotas(bgc, engine):
    namespaces = c1_namespaces()
    nsmap = bgc.pods.NamespaceMap()
    for ns in namespaces:
        nsmap.intern(ns["name"])
    R = len(C1_KEYS)
    hard = np.zeros((R, len(namespaces)), np.int64)
    hard_mask = np.zeros(len(namespaces), np.uint8)
    for i, ns in enumerate(namespaces):
        h, unit = bgc.pods.hard_from_sheet_row(ns["cpu"], ns["memory"], ns["gpu"], ns["storage"], ns["mig"], C1_KEYS)
        hard[:, i] = h
        hard_mask[i] = (1 << R) - 1 if ns["has_quota"] else 0
    engine.quota_table_set(hard, hard_mask, unit)

    running = c1_pods(10_000, namespaces, running=True)
    cands = c1_pods(10_000, namespaces, running=False)
    cands[17]["metadata"]["namespace"] = "not-a-namespace"

    r_ns, r_req, r_mask, _ = bgc.pods.flatten_pods(running, C1_KEYS, nsmap, intern=False)
    used = engine.usage_reduce(r_ns, r_req)
    c_ns, c_req, c_mask, _ = bgc.pods.flatten_pods(cands, C1_KEYS, nsmap, intern=False)
    verdict = engine.quota_eval(c_ns, c_req, c_mask)

    # independent evaluation in plain Python on the pod objects themselves
    idx = {ns["name"]: i for i, ns in enumerate(namespaces)}
    want_used = [[0] * len(namespaces) for _ in range(R)]
    live = 0
    for pod in running:
        if pod["status"]["phase"] in ("Succeeded", "Failed"):
            continue
        live += 1
        for r, key in enumerate(C1_KEYS):
            want_used[r][idx[pod["metadata"]["namespace"]]] += _pod_usage(pod, key)[0]
    assert live == r_ns.shape[0]
    assert used.tolist() == want_used
    compute = [False, True, True, False]
    want = []
    for pod in cands:
        i = idx.get(pod["metadata"]["namespace"])
        if i is None:
            want.append(0xFF)
            continue
        missing = exceeded = 0
        for r, key in enumerate(C1_KEYS):
            if not (hard_mask[i] >> r) & 1:
                continue
            q, stated = _pod_usage(pod, key)
            if not stated:
                if compute[r]:
                    missing |= 1 << r
                continue
            if q > 0 and want_used[r][i] + q > int(hard[r][i]):
                exceeded |= 1 << r
        want.append((missing << 4) | exceeded)
    assert verdict.tolist() == want
    deny = np.mean(np.array(want) != 0)
    assert 0.05 < deny < 0.999
    assert verdict[17] == bgc.VERDICT_BAD_NS


@pytest.mark.parametrize("wide", [False, True])
def test_pod_reviews_wire_to_wire_on_the_device(bgc, engine, oracle, wide):
    """review.handle_pod_reviews with the real Engine: AdmissionReview bytes -> parse -> flatten (row A5) -> K2 on
    the GPU -> response JSON, against the oracle's verdicts on the flattened columns; one malformed body and one
    body without an object in the middle of the batch are answered `invalid` on their own.  Seam:
    src/admission.rs:184-204; webhook registration charts/.../templates/webhook.yaml:10-27."""
    from workloads import C1_KEYS6
    keys = C1_KEYS6 if wide else C1_KEYS
    nss = c1_namespaces()
    m = bgc.pods.NamespaceMap()
    for ns in nss:
        m.intern(ns["name"])
    R = len(keys)
    hard = np.zeros((R, len(nss)), np.int64)
    hm = np.zeros(len(nss), np.uint8)
    for i, ns in enumerate(nss):
        h, unit = bgc.pods.hard_from_sheet_row(ns["cpu"], ns["memory"], ns["gpu"], ns["storage"], ns["mig"], keys)
        hard[:, i] = h
        hm[i] = (1 << R) - 1 if ns["has_quota"] else 0
    engine.quota_table_set(hard, hm, unit)
    r_ns, r_req, _, _ = bgc.pods.flatten_pods(c1_pods(10_000, nss, running=True), keys, m, intern=False)
    used = engine.usage_reduce(r_ns, r_req)
    assert np.array_equal(used, oracle.usage_reduce(r_ns, r_req, len(nss)))
    pods = c1_pods(10_000, nss, running=False)
    pods[5]["metadata"]["namespace"] = "ghost"
    bodies = [json.dumps({"apiVersion": "admission.k8s.io/v1", "kind": "AdmissionReview",
                          "request": {"uid": f"u{i}", "operation": "CREATE", "kind": {"kind": "Pod"}, "namespace": p["metadata"]["namespace"], "object": p}})
              for i, p in enumerate(pods)]
    bodies[100] = '{"apiVersion": "admission.k8s.io/v1", "kind": "AdmissionReview", "request": {"uid": '      # truncated JSON
    bodies[200] = json.dumps({"apiVersion": "admission.k8s.io/v1", "kind": "AdmissionReview", "request": {"uid": "u200", "operation": "DELETE"}})
    outs = bgc.review.handle_pod_reviews(engine, bodies, keys, m, used, hard, wide=wide)
    assert len(outs) == len(bodies)
    keep = [i for i in range(len(pods)) if i not in (100, 200)]
    c_ns, c_req, c_mask, _ = bgc.pods.flatten_pods([pods[i] for i in keep], keys, m, intern=False)
    cm = sum(1 << r for r, u in enumerate(unit) if u)
    want = (oracle.quota_eval16 if wide else oracle.quota_eval)(c_ns, c_req, c_mask, len(nss), used, hard, hm, cm)
    kinds = set()
    for k, i in enumerate(keep):
        resp = json.loads(outs[i])["response"]
        assert resp["uid"] == f"u{i}" and resp["allowed"] == (want[k] == 0), i
        if want[k] != 0 and i != 5:
            kinds.add(resp["status"]["message"].split(":")[0])
    assert kinds == {"exceeded quota", "failed quota"}
    assert "not known" in json.loads(outs[5])["response"]["status"]["message"]
    for i in (100, 200):
        resp = json.loads(outs[i])["response"]
        assert resp["allowed"] is False and resp["status"]["reason"] == "InvalidRequest" and resp["uid"] == ""
    assert 0.02 < float(np.mean(want != 0)) < 0.999
    # the one-call form (bgc_handle_pod_reviews) answers byte for byte like the call-by-call form
    stages = {}
    assert bgc.review.handle_pod_reviews(engine, bodies, keys, m, used, hard, wide=wide, stage_ms=stages) == \
        bgc.review.handle_pod_reviews_stepwise(engine, bodies, keys, m, used, hard, wide=wide)
    assert set(stages) == {"parse_flatten_release", "gather", "verdicts", "render", "copy"}
    few = bodies[:7] + bodies[98:102]                              # below the OpenMP threshold, no denial text
    assert bgc.review.handle_pod_reviews(engine, few, keys, m, wide=wide) == bgc.review.handle_pod_reviews_stepwise(engine, few, keys, m, wide=wide)
    assert bgc.review.handle_pod_reviews(engine, [], keys, m, used, hard, wide=wide) == []
    done = dict(pods[3]); done["status"] = {"phase": "Succeeded"}
    gone = json.dumps({"apiVersion": "admission.k8s.io/v1", "kind": "AdmissionReview",
                       "request": {"uid": "t", "operation": "CREATE", "kind": {"kind": "Pod"}, "namespace": done["metadata"]["namespace"], "object": done}})
    with pytest.raises(ValueError):
        bgc.review.handle_pod_reviews(engine, [gone], keys, m, used, hard, wide=wide)   # a terminal pod is not a create review


def test_c1_pods_through_compact_columns(bgc, engine, oracle):
    """The whole host-to-device route a caller with pod objects would take for the narrow-column ABI: pod JSON ->
    bgc_flatten_pods (int64 columns) -> bgc_compact_plan/column (narrowest exact representation) ->
    bgc_usage_reduce_typed / bgc_quota_eval_typed; results equal the int64 entry points and the oracle."""
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
    engine.quota_table_set(hard, hm, unit)
    r_ns, r_req, _, _ = bgc.pods.flatten_pods(c1_pods(10_000, nss, running=True), C1_KEYS, m, intern=False)
    cands = c1_pods(10_000, nss, running=False)
    cands[3]["metadata"]["namespace"] = "nowhere"
    c_ns, c_req, c_mask, _ = bgc.pods.flatten_pods(cands, C1_KEYS, m, intern=False)
    rn, rc, rs = bgc.pods.compact_columns(r_ns, r_req, len(nss))
    cn, cc, cs = bgc.pods.compact_columns(c_ns, c_req, len(nss))
    assert rn.dtype == np.uint16 and sum(c.dtype.itemsize for c in rc) < 8 * R     # narrower than int64 columns
    used = engine.usage_reduce(r_ns, r_req)
    assert np.array_equal(engine.usage_reduce_typed(rn, rc, rs), used)
    assert np.array_equal(used, oracle.usage_reduce(r_ns, r_req, len(nss)))
    want = engine.quota_eval(c_ns, c_req, c_mask)
    got = engine.quota_eval_typed(cn, cc, cs, c_mask)
    assert np.array_equal(got, want) and got[3] == bgc.VERDICT_BAD_NS
    cm = sum(1 << r for r, u in enumerate(unit) if u)
    assert np.array_equal(got, oracle.quota_eval(c_ns, c_req, c_mask, len(nss), used, hard, hm, cm))
