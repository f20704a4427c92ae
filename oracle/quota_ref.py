"""oracle/quota_ref.py — TEST INFRASTRUCTURE ONLY (second, independent numpy statement of rows A3/A4).

PARITY UNPINNED: the reference has no implementation of this path (see oracle/quota_ref.h).
Vectorised numpy, sharing no code with oracle/quota_ref.c.
"""
from __future__ import annotations

import numpy as np

VERDICT_BAD_NS = 0xFF
_M64 = (1 << 64) - 1


def usage_reduce(ns_id: np.ndarray, req: list, NS: int) -> np.ndarray:
    """used[r][ns] = Σ req[r][i] over pods of ns, wrapping int64 (numpy int64 add wraps)."""
    R = len(req)
    used = np.zeros((R, NS), dtype=np.int64)
    ok = ns_id < NS
    idx = ns_id[ok].astype(np.int64)
    for r in range(R):
        with np.errstate(over="ignore"):
            np.add.at(used[r], idx, req[r][ok])
    return used


def quota_eval(ns_id, req, req_mask, NS, used, hard, hard_mask, compute_mask, wide: bool = False) -> np.ndarray:
    R = len(req)
    n = ns_id.shape[0]
    bad = ns_id >= NS
    ns = np.where(bad, 0, ns_id).astype(np.int64)
    hm = hard_mask[ns] if NS else np.zeros(n, np.uint8)
    rm = req_mask if req_mask is not None else np.full(n, 0xFF, np.uint8)
    missing = np.zeros(n, np.uint8)
    exceeded = np.zeros(n, np.uint8)
    for r in range(R):
        constrained = ((hm >> r) & 1).astype(bool)
        present = ((rm >> r) & 1).astype(bool)
        if (compute_mask >> r) & 1:
            missing |= ((constrained & ~present).astype(np.uint8) << r)
        u = used[r][ns]
        with np.errstate(over="ignore"):
            total = u + req[r]                       # wraps; a positive overflow shows as total < u
        overflow = (req[r] > 0) & (total < u)       # true sum > INT64_MAX >= hard
        ex = constrained & present & (req[r] > 0) & (overflow | (total > hard[r][ns]))
        exceeded |= (ex.astype(np.uint8) << r)
    if wide:   # u16 form, up to 8 resources
        out = ((missing.astype(np.uint16) << 8) | exceeded).astype(np.uint16)
        out[bad] = 0xFFFF
        return out
    out = ((missing << 4) | exceeded).astype(np.uint8)
    out[bad] = VERDICT_BAD_NS
    return out


def admit_greedy(ns_id, req, req_mask, NS, used, hard, hard_mask, compute_mask):
    """N1, pure-Python statement (small cases only): arrival order, only admitted pods consume."""
    R = len(req)
    used = [[int(x) for x in row] for row in used]
    out = np.zeros(ns_id.shape[0], np.uint8)
    wrap = lambda x: ((x + (1 << 63)) & _M64) - (1 << 63)
    for j in range(ns_id.shape[0]):
        ns = int(ns_id[j])
        if ns >= NS:
            out[j] = VERDICT_BAD_NS
            continue
        rm = 0xFF if req_mask is None else int(req_mask[j])
        missing = exceeded = 0
        for r in range(R):
            if not (int(hard_mask[ns]) >> r) & 1:
                continue
            if not (rm >> r) & 1:
                if (compute_mask >> r) & 1:
                    missing |= 1 << r
                continue
            q = int(req[r][j])
            if q > 0 and used[r][ns] + q > int(hard[r][ns]):
                exceeded |= 1 << r
        out[j] = (missing << 4) | exceeded
        if out[j] == 0:
            for r in range(R):
                used[r][ns] = wrap(used[r][ns] + int(req[r][j]))
    return out, np.array(used, dtype=np.int64)


# ---- generator spec, numpy restatement (cross-checks gen_ref_* in quota_ref.c) ----
def _sm64(x: np.ndarray) -> np.ndarray:
    x = (x + np.uint64(0x9E3779B97F4A7C15))
    x = (x ^ (x >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
    x = (x ^ (x >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    return x ^ (x >> np.uint64(31))


def gen_hash(seed: int, stream: int, index: np.ndarray) -> np.ndarray:
    with np.errstate(over="ignore"):
        key = _sm64(np.array([(seed ^ (((stream + 1) * 0xD1B54A32D192ED03) & _M64)) & _M64], dtype=np.uint64))[0]
        return _sm64(key ^ (index.astype(np.uint64) * np.uint64(0x9E3779B97F4A7C15)))


def gen_value(kind: int, h: np.ndarray) -> np.ndarray:
    t = (h >> np.uint64(32)).astype(np.int64)
    if kind == 0:
        c = t % 100
        return np.select([c < 50, c < 80, c < 90, c < 97], [0, 1, 2, 4], 8).astype(np.int64)
    if kind == 1:
        return (250 << (t % 7)).astype(np.int64)
    if kind == 2:
        return (np.int64(1) << (28 + (t % 9))).astype(np.int64)
    c = t % 100
    return np.select([c < 90, c < 97], [0, 1], 2).astype(np.int64)


def gen_fill(seed, set_, NS, dist, grouped, n_total, a, b, cdf, kinds, first, count, want_mask=False):
    i = np.arange(first, first + count, dtype=np.uint64)
    sbase = set_ * 16
    with np.errstate(over="ignore"):
        if grouped:
            u = i * np.uint64(_M64 // n_total)
        else:
            u = gen_hash(seed, sbase, i)
    if dist == 1:
        rank = np.minimum(np.searchsorted(cdf, u, side="right"), NS - 1).astype(np.uint64)
    else:
        rank = ((u >> np.uint64(32)) * np.uint64(NS)) >> np.uint64(32)
    ns = ((rank * np.uint64(a) + np.uint64(b)) % np.uint64(NS)).astype(np.uint32)
    req = [gen_value(k, gen_hash(seed, sbase + 1 + r, i)) for r, k in enumerate(kinds)]
    mask = None
    if want_mask:
        h = gen_hash(seed, sbase + 8, i)
        mask = np.full(count, 0xFF, np.uint8)
        for r in range(len(kinds)):
            drop = ((h >> np.uint64(8 * r)) & np.uint64(0xFF)) < np.uint64(8)
            mask[drop] &= np.uint8(~(1 << r) & 0xFF)
    return ns, req, mask


def zipf_cdf(NS: int, s: float) -> np.ndarray:
    """Host-built u64 CDF table for Zipf(s) over ranks 1..NS (an INPUT to both CPU and GPU)."""
    w = np.arange(1, NS + 1, dtype=np.float64) ** (-s)
    c = np.cumsum(w)
    c /= c[-1]
    out = np.minimum(np.floor(c * 18446744073709551616.0), 18446744073709549568.0).astype(np.uint64)
    out[-1] = np.uint64(_M64)
    return out


def scramble_params(NS: int) -> tuple:
    """Multiplier coprime to NS near NS*0.618 (rank → id bijection), and an offset."""
    from math import gcd
    a = max(1, int(NS * 0.6180339887))
    while gcd(a, NS) != 1:
        a += 1
    return a, NS // 3
