"""Regenerates tests/golden/quota_cases.json — small fixed cases for rows A3/A4/N1 with the answers of
the pure-numpy / pure-Python restatements (oracle/quota_ref.py).  The reference has no implementation
of this path, so these are regression vectors for THIS repo's spec, not reference
outputs: the C oracle (CPU test) and the CUDA kernels (GPU test) must both reproduce them.
Run:  python tests/golden/make_quota_cases.py"""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
from oracle import quota_ref as Q  # noqa: E402

rng = np.random.default_rng(20261106)
cases = []
for R, NS, n, wild in [(1, 1, 5, False), (2, 7, 64, False), (3, 16, 200, False), (4, 5, 150, True), (6, 9, 120, False), (8, 3, 90, True), (3, 40, 300, True)]:
    ns = rng.integers(0, NS + (2 if n > 50 else 0), n, dtype=np.uint32)
    if wild:
        req = [rng.integers(-2**62, 2**62, n, dtype=np.int64) for _ in range(R)]
        for c in req:
            c[::3] = rng.integers(0, 1000, c[::3].shape[0])
    else:
        req = [rng.choice(np.array([0, 1, 2, 250, 1000, 1 << 28, 1 << 33], np.int64), n) for _ in range(R)]
    mask = rng.integers(0, 256, n, dtype=np.uint8)
    units = [int(u) for u in rng.integers(0, 3, R)]
    cm = sum(1 << r for r, u in enumerate(units) if u)
    used = Q.usage_reduce(ns, req, NS)
    with np.errstate(over="ignore"):
        hard = (used // 2 + rng.integers(0, 5000, used.shape)).astype(np.int64)
    hm = rng.integers(0, 1 << R, NS, dtype=np.uint8)
    v16 = Q.quota_eval(ns, req, mask, NS, used, hard, hm, cm, wide=True)
    case = dict(R=R, NS=NS, units=units, ns_id=ns.tolist(), req=[c.tolist() for c in req], req_mask=mask.tolist(),
                used=used.tolist(), hard=hard.tolist(), hard_mask=hm.tolist(), verdict16=v16.tolist())
    if R <= 4:
        case["verdict8"] = Q.quota_eval(ns, req, mask, NS, used, hard, hm, cm).tolist()
        gv, gu = Q.admit_greedy(ns, req, mask, NS, np.zeros_like(used), hard, hm, cm)
        case["greedy_verdict"] = gv.tolist()
        case["greedy_used"] = gu.tolist()
    cases.append(case)
out = os.path.join(os.path.dirname(__file__), "quota_cases.json")
json.dump(cases, open(out, "w"))
print("wrote", out, os.path.getsize(out), "bytes")
