"""Committed regression vectors for rows A3/A4/N1 (tests/golden/quota_cases.json, produced by the numpy /
pure-Python restatements): the C oracle must reproduce them on CPU, the CUDA kernels on the GPU."""
import json
import os

import numpy as np
import pytest

CASES = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "quota_cases.json")))


def _arrays(c):
    ns = np.array(c["ns_id"], np.uint32)
    req = [np.array(x, np.int64) for x in c["req"]]
    mask = np.array(c["req_mask"], np.uint8)
    used = np.array(c["used"], np.int64).reshape(c["R"], c["NS"])
    hard = np.array(c["hard"], np.int64).reshape(c["R"], c["NS"])
    hm = np.array(c["hard_mask"], np.uint8)
    cm = sum(1 << r for r, u in enumerate(c["units"]) if u)
    return ns, req, mask, used, hard, hm, cm


@pytest.mark.parametrize("i", range(len(CASES)))
def test_c_oracle_reproduces_the_committed_vectors(oracle, i):
    c = CASES[i]
    ns, req, mask, used, hard, hm, cm = _arrays(c)
    assert np.array_equal(oracle.usage_reduce(ns, req, c["NS"]), used)
    assert oracle.quota_eval16(ns, req, mask, c["NS"], used, hard, hm, cm).tolist() == c["verdict16"]
    if c["R"] <= 4:
        assert oracle.quota_eval(ns, req, mask, c["NS"], used, hard, hm, cm).tolist() == c["verdict8"]
        gv, gu = oracle.admit_greedy(ns, req, mask, c["NS"], np.zeros_like(used), hard, hm, cm)
        assert gv.tolist() == c["greedy_verdict"] and gu.tolist() == c["greedy_used"]


@pytest.mark.gpu
@pytest.mark.parametrize("i", range(len(CASES)))
def test_kernels_reproduce_the_committed_vectors(engine, i):
    c = CASES[i]
    ns, req, mask, used, hard, hm, cm = _arrays(c)
    engine.quota_table_set(hard, hm, c["units"])
    for variant in (0, 1, 2, 3):
        engine.set_k1_variant(variant)
        assert np.array_equal(engine.usage_reduce(ns, req), used)
    engine.set_k1_variant(0)
    assert engine.quota_eval(ns, req, mask, wide=True).tolist() == c["verdict16"]
    if c["R"] <= 4:
        assert engine.quota_eval(ns, req, mask).tolist() == c["verdict8"]
        engine.usage_set(np.zeros_like(used))
        assert engine.admit_greedy(ns, req, mask).tolist() == c["greedy_verdict"]
        assert engine.usage_get().tolist() == c["greedy_used"]
