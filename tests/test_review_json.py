"""Row N4: AdmissionReview JSON in/out.  CPU tests cover the C++ parser against Python's json module
and the Python mirror; the GPU test runs wire bytes → K0 → response JSON against the oracle."""
import base64
import json

import numpy as np
import pytest

from oracle import policy_ref as P


def _reviews():
    cfg = P.Config()
    out = []
    for op in range(4):
        for fl in range(1024):
            r = P.request_from_bits(op, fl, cfg)
            if r is not None:
                out.append({"apiVersion": "admission.k8s.io/v1", "kind": "AdmissionReview", "request": r})
    return out


def test_parser_extracts_the_fields_mutate_reads(bgc):
    A, Rv = bgc.admission, bgc.review
    for rv in _reviews():
        for text in (json.dumps(rv), json.dumps(rv, indent=2, ensure_ascii=True)):
            got = Rv.Review(text).request_fields()
            want = A.AdmissionRequest.from_review(rv["request"])
            assert got["username"] == want.username and got["groups"] == (None if want.groups is None else list(want.groups))
            assert got["operation"] == A.OPERATIONS[want.operation]
            assert (got["has_object"], got["name"], got["parse_ok"]) == (want.has_object, want.name, want.parse_ok)
            assert (got["spec_kube_username"], got["has_quota"], got["has_rolebinding"]) == (want.spec_kube_username, want.has_quota, want.has_rolebinding)


def test_parser_json_details(bgc):
    Rv = bgc.review
    body = r'{"request":{"uid":"é😀-\"q\"","operation":"CREATE","kind":{"kind":"UserBootstrap"},"userInfo":{"username":"oidc:알리스","groups":[]},"object":{"metadata":{"name":"알리스"},"spec":{"kube_username":null,"quota":null}}},"extra":[1,2.5e3,true,null]}'
    r = Rv.Review(body)
    assert r.uid == 'é😀-"q"' and r.kind == "UserBootstrap"
    f = r.request_fields()
    assert f["username"] == "oidc:알리스" and f["groups"] == [] and f["name"] == "알리스" and f["parse_ok"] and not f["has_quota"]
    assert not Rv.Review('{"apiVersion":"admission.k8s.io/v1"}').has_request
    for bad in ["", "{", '{"a":}', '{"a":1,}', "[1,2", '{"a":"\\x"}', b'{"a":"\xff"}', b'{"a":"\xed\xa0\x80"}', b'{"a":"\xc0\xaf"}', '{"a":"\\ud800"}', "nul", '{"a":1} x', '"str"']:
        with pytest.raises(ValueError):
            Rv.Review(bad)


def test_pod_review_flattening_equals_dict_flattening(bgc):
    import sys, os
    sys.path.insert(0, os.path.dirname(__file__))
    from workloads import C1_KEYS, c1_namespaces, c1_pods
    nss = c1_namespaces()
    pods = c1_pods(500, nss, running=False)
    m = bgc.pods.NamespaceMap()
    for ns in nss:
        m.intern(ns["name"])
    bodies = [json.dumps({"apiVersion": "admission.k8s.io/v1", "kind": "AdmissionReview",
                          "request": {"uid": f"u{i}", "operation": "CREATE", "kind": {"kind": "Pod"}, "namespace": p["metadata"]["namespace"], "object": p}})
              for i, p in enumerate(pods)]
    a = bgc.review.flatten_pod_reviews(bodies, C1_KEYS, m)
    b = bgc.pods.flatten_pods(pods, C1_KEYS, m, intern=False)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[2], b[2])
    for x, y in zip(a[1], b[1]):
        assert np.array_equal(x, y)


def test_pod_review_with_thousands_of_containers(bgc):
    """The per-review stores of resource lists hand out pointers while they grow (a deque: nothing moves): a pod with
    3,000 containers and 500 init containers flattens to the same row through the review path and the dict path."""
    from workloads import C1_KEYS
    m = bgc.pods.NamespaceMap()
    m.intern("big")
    ctrs = [{"name": f"c{i}", "resources": {"requests": {"cpu": f"{1 + i % 7}m", "memory": f"{1 + i % 5}Mi", "nvidia.com/gpu": str(i % 2)},
                                            "limits": {"nvidia.com/gpu": str(i % 2)}}} for i in range(3000)]
    inits = [{"name": f"i{i}", "resources": {"requests": {"cpu": "9", "memory": "1Gi"}}} for i in range(500)]
    pod = {"metadata": {"name": "p", "namespace": "big"}, "spec": {"containers": ctrs, "initContainers": inits}}
    body = json.dumps({"apiVersion": "admission.k8s.io/v1", "kind": "AdmissionReview",
                       "request": {"uid": "u", "operation": "CREATE", "kind": {"kind": "Pod"}, "namespace": "big", "object": pod}})
    a = bgc.review.flatten_pod_reviews([body], C1_KEYS, m)
    b = bgc.pods.flatten_pods([pod], C1_KEYS, m, intern=False)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[2], b[2])
    for x, y in zip(a[1], b[1]):
        assert np.array_equal(x, y) and x.shape == (1,)
    assert int(a[1][C1_KEYS.index("requests.cpu")][0]) == max(sum(1 + i % 7 for i in range(3000)), 9000)   # max(sum of containers, largest init container)


@pytest.mark.gpu
def test_wire_bytes_to_response_json(bgc, engine):
    cfg_py, cfg = P.Config(), bgc.admission.AdmissionConfig()
    reviews = _reviews()
    bodies = [json.dumps(rv) for rv in reviews] + ['{"apiVersion":"admission.k8s.io/v1","kind":"AdmissionReview"}']
    outs = bgc.review.handle_reviews(engine, cfg, bodies)
    assert len(outs) == len(bodies)
    for rv, text in zip(reviews, outs):
        doc = json.loads(text)
        assert doc["apiVersion"] == "admission.k8s.io/v1" and doc["kind"] == "AdmissionReview"
        resp = doc["response"]
        d = P.mutate(rv["request"], cfg_py)
        assert resp["allowed"] == (d.klass in (P.ALLOW, P.ALLOW_PATCH))
        if d.klass == P.INVALID:
            assert resp["uid"] == "" and resp["status"]["code"] == 400 and resp["status"]["reason"] == "InvalidRequest"
            if d.reason != 8:
                assert resp["status"]["message"] == d.message
        else:
            assert resp["uid"] == rv["request"]["uid"]
            if d.klass == P.DENY:
                assert resp["status"]["message"] == d.message
        if d.patch_mask:
            assert resp["patchType"] == "JSONPatch" and json.loads(base64.b64decode(resp["patch"])) == d.patches
        else:
            assert "patch" not in resp
    last = json.loads(outs[-1])["response"]
    assert last["allowed"] is False and last["status"]["reason"] == "InvalidRequest"


@pytest.mark.gpu
def test_one_call_handler_equals_the_stepwise_handler(bgc, engine):
    """bgc_handle_reviews (one C call per batch) against the same handler spelled out call by call: byte-equal
    responses, bad bodies answered on their own, an `out` that is too small reports the size it needs."""
    import ctypes as C
    cfg = bgc.admission.AdmissionConfig()
    reviews = _reviews()
    rng = np.random.default_rng(7)
    bodies = [json.dumps(reviews[int(i)]) for i in rng.integers(0, len(reviews), 3000)]
    long_name = "n" * 5000                      # a patch and a response longer than the first render buffers
    big = json.loads(bodies[0]); big["request"]["object"] = {"metadata": {"name": long_name}, "spec": {"kube_username": None, "quota": None}}
    big["request"]["operation"] = "CREATE"; big["request"]["userInfo"] = {"username": "oidc:" + long_name, "groups": []}
    bodies[17] = json.dumps(big)
    bodies[5] = "{"; bodies[6] = '{"kind":"AdmissionReview"}'; bodies[7] = ""; bodies[2999] = b'{"a":"\xff"}'
    stages = {}
    got = bgc.review.handle_reviews(engine, cfg, bodies, stage_ms=stages)
    want = bgc.review.handle_reviews_stepwise(engine, cfg, bodies)
    assert got == want
    assert set(stages) == {"parse_flatten_release", "decide", "render", "copy"} and all(v >= 0 for v in stages.values())
    assert bgc.review.handle_reviews(engine, cfg, []) == []
    assert bgc.review.handle_reviews(engine, cfg, bodies[:3]) == want[:3]       # below the OpenMP threshold
    raw = [b.encode() if isinstance(b, str) else b for b in bodies[:10]]
    arr = (C.c_char_p * 10)(*raw); lens = (C.c_size_t * 10)(*[len(b) for b in raw]); off = (C.c_size_t * 11)()
    cc, keep = cfg._c()
    small = C.create_string_buffer(16)
    assert bgc.lib.bgc_handle_reviews(engine._h, C.byref(cc), arr, lens, 10, small, 16, off, None) == bgc.E_BADARG
    assert off[10] == sum(len(w.encode()) + 1 for w in want[:10])
    assert bgc.lib.bgc_handle_reviews(None, C.byref(cc), arr, lens, 10, small, 16, off, None) == bgc.E_BADARG


def test_pod_reviews_wire_to_wire_with_a_stand_in_engine(bgc, oracle):
    """`review.handle_pod_reviews` composes parse -> flatten -> K2 -> render.  K2 itself is covered by the
    GPU tests; here a stand-in engine whose quota_eval is the oracle lets the composition run on CPU."""
    import os, sys
    sys.path.insert(0, os.path.dirname(__file__))
    from workloads import C1_KEYS, C1_KEYS6, c1_namespaces, c1_pods
    nss = c1_namespaces()
    m = bgc.pods.NamespaceMap()
    for ns in nss:
        m.intern(ns["name"])
    for keys, wide in ((C1_KEYS, False), (C1_KEYS6, True)):
        R = len(keys)
        hard = np.zeros((R, len(nss)), np.int64)
        hm = np.zeros(len(nss), np.uint8)
        for i, ns in enumerate(nss):
            h, unit = bgc.pods.hard_from_sheet_row(ns["cpu"], ns["memory"], ns["gpu"], ns["storage"], ns["mig"], keys)
            hard[:, i] = h
            hm[i] = (1 << R) - 1 if ns["has_quota"] else 0
        r_ns, r_req, _, _ = bgc.pods.flatten_pods(c1_pods(3000, nss, running=True), keys, m, intern=False)
        used = oracle.usage_reduce(r_ns, r_req, len(nss))
        cm = sum(1 << r for r, u in enumerate(unit) if u)

        class StandIn:
            def quota_eval(self, ns_id, req, mask, wide=False):
                f = oracle.quota_eval16 if wide else oracle.quota_eval
                return f(ns_id, req, mask, len(nss), used, hard, hm, cm)

        pods = c1_pods(400, nss, running=False)
        pods[5]["metadata"]["namespace"] = "ghost"
        bodies = [json.dumps({"apiVersion": "admission.k8s.io/v1", "kind": "AdmissionReview",
                              "request": {"uid": f"u{i}", "operation": "CREATE", "kind": {"kind": "Pod"}, "namespace": p["metadata"]["namespace"], "object": p}})
                  for i, p in enumerate(pods)]
        bodies.append("{")        # a malformed body costs only itself an answer (per-body parse failures)
        outs = bgc.review.handle_pod_reviews(StandIn(), bodies, keys, m, used, hard, wide=wide)
        bad = json.loads(outs.pop())["response"]
        assert bad["allowed"] is False and bad["status"]["reason"] == "InvalidRequest"
        bodies.pop()
        c_ns, c_req, c_mask, _ = bgc.pods.flatten_pods(pods, keys, m, intern=False)
        verdict = StandIn().quota_eval(c_ns, c_req, c_mask, wide=wide)
        seen = set()
        for i, text in enumerate(outs):
            resp = json.loads(text)["response"]
            assert resp["uid"] == f"u{i}" and resp["allowed"] == (verdict[i] == 0)
            if verdict[i] == 0:
                assert resp["status"] == {}
            elif i == 5:
                assert "not known" in resp["status"]["message"]
            else:
                msg = resp["status"]["message"]
                ns_name = pods[i]["metadata"]["namespace"]
                assert msg.startswith(f"exceeded quota: {ns_name}, requested: ") or msg.startswith(f"failed quota: {ns_name}: must specify ")
                seen.add(msg.split(":")[0])
        assert seen == {"exceeded quota", "failed quota"}


def test_parser_is_as_strict_as_serde_on_groups_subjects_and_quota(bgc):
    """a wrong-typed userInfo.groups fails the whole review (AdmissionReview deserialisation,
    src/admission.rs:186); an object whose rolebinding subjects lack kind/name, or whose quota.hard holds
    non-quantities, does not parse as UserBootstrap (src/admission.rs:341-347 answers `invalid`)."""
    from bacchus_helpers import ub_review
    for groups in ("gpu", ["gpu", 7], [None]):
        with pytest.raises(ValueError):
            bgc.review.Review(json.dumps(ub_review(groups=groups)))
    assert bgc.review.Review(json.dumps(ub_review(groups=None))).request_fields()["groups"] is None
    good_rb = {"role_ref": {"apiGroup": "rbac.authorization.k8s.io", "kind": "ClusterRole", "name": "edit"},
               "subjects": [{"kind": "User", "name": "alice"}]}
    cases = [({"rolebinding": good_rb}, True),
             ({"rolebinding": {**good_rb, "subjects": [{"kind": "User"}]}}, False),
             ({"rolebinding": {**good_rb, "subjects": ["alice"]}}, False),
             ({"rolebinding": {**good_rb, "subjects": None}}, True),
             ({"quota": {"hard": {"requests.cpu": "8", "requests.nvidia.com/gpu": 2}}}, True),
             ({"quota": {"hard": {"requests.cpu": ["8"]}}}, False),
             ({"quota": {"hard": "lots"}}, False)]
    for spec, ok in cases:
        rv = ub_review(spec=spec)
        f = bgc.review.Review(json.dumps(rv)).request_fields()
        assert f["parse_ok"] is ok, spec
        assert P._parses_as_userbootstrap(rv["request"]["object"]) is ok, spec
        assert bgc.admission._parses_as_userbootstrap(rv["request"]["object"]) is ok, spec
