#!/usr/bin/env python
"""Small invocations of every kernel and variant, for compute-sanitizer runs (memcheck / racecheck /
synccheck).  Results are compared with the oracle as well, so a sanitizer-clean run is also a correct one."""
import importlib, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle
from workloads import UNITS, make_hard
bgc = importlib.import_module("bacchus-gpu-controller_b200")

eng = bgc.Engine(0)
rng = np.random.default_rng(0)
# K0
op = rng.integers(0, 5, 5003, dtype=np.uint8); fl = rng.integers(0, 1024, 5003, dtype=np.uint16)
c, m = eng.policy_eval(op, fl)
rc, rm = oracle.policy_bits(np.minimum(op, 3).astype(np.uint8), fl)
assert np.array_equal(c, rc) and np.array_equal(m, rm)
# K1 / K2 / K3
for R in (1, 3, 4, 6, 8):
    for NS, n in ((7, 3001), (3000, 20011), (40000, 20011)):
        ns = rng.integers(0, NS + 1, n, dtype=np.uint32)
        req = [rng.integers(-5, 1 << 34, n, dtype=np.int64) for _ in range(R)]
        mask = rng.integers(0, 256, n, dtype=np.uint8)
        want = oracle.usage_reduce(ns, req, NS)
        hard, hm = make_hard(want, R)
        eng.quota_table_set(hard, hm, UNITS[R])
        variants = [0, 1, 2, 3, 2 | (1 << 8), 3 | (1 << 8)] + ([4, 5] if NS <= 40000 else [])
        for v in variants:
            if R > 4 and (v >> 8):
                continue
            eng.set_k1_variant(v)
            for flags in (0, 1):
                got = eng.usage_reduce(ns, req, flags=flags)
                assert np.array_equal(got, want), (R, NS, v, flags, eng.last_k1_variant())
        eng.set_k1_variant(0)
        cm = sum(1 << r for r, u in enumerate(UNITS[R]) if u)
        w16 = oracle.quota_eval16(ns, req, mask, NS, want, hard, hm, cm)
        assert np.array_equal(eng.quota_eval(ns, req, mask, wide=True), w16)
        if R <= 4:
            assert np.array_equal(eng.quota_eval(ns, req, mask), oracle.quota_eval(ns, req, mask, NS, want, hard, hm, cm))
            gv, gu = oracle.admit_greedy(ns, req, mask, NS, want, hard * 2, hm, cm)
            eng.quota_table_set(hard * 2, hm, UNITS[R]); eng.usage_set(want)
            assert np.array_equal(eng.admit_greedy(ns, req, mask), gv) and np.array_equal(eng.usage_get(), gu)
# hot segment for the CTA-wide greedy walker (>= 32768 pods in one namespace), tame and wild tiles
R, NS, n = 3, 2, 70001
ns = np.zeros(n, np.uint32); ns[::9] = 1
req = [rng.integers(0, 1 << 20, n, dtype=np.int64) for _ in range(R)]
req[1][40000:40100] = -3
used0 = np.zeros((R, NS), np.int64)
hard = np.full((R, NS), 1 << 33, np.int64); hm = np.full(NS, 7, np.uint8)
eng.quota_table_set(hard, hm, UNITS[R])
gv, gu = oracle.admit_greedy(ns, req, None, NS, used0, hard, hm, 6)
assert np.array_equal(eng.admit_greedy(ns, req), gv) and np.array_equal(eng.usage_get(), gu)
eng.close()
print("sanitize cases ok")
