"""GPU parity tests for K3 admit_greedy: Kubernetes-faithful sequential
admission inside a batch.  Order-dependent, so the checker is the oracle's literal loop
(oracle/quota_ref.c::quota_ref_admit_greedy); bit-exact verdicts and final usage."""
import numpy as np
import pytest

from oracle import quota_ref as Q
from workloads import make_hard, KINDS, UNITS, SEEDS

pytestmark = pytest.mark.gpu


def _cm(units):
    return sum(1 << r for r, u in enumerate(units) if u != 0)


@pytest.mark.parametrize("R", [1, 2, 3, 4])
@pytest.mark.parametrize("with_mask", [False, True])
@pytest.mark.parametrize("n,NS", [(1, 1), (31, 2), (32, 2), (33, 2), (1000, 7), (50_000, 300), (200_003, 10_000), (160_001, 3), (32768, 1), (32769, 1)])
def test_greedy_vs_literal_loop(engine, oracle, R, with_mask, n, NS):
    rng = np.random.default_rng(R * 7 + n)
    ns = rng.integers(0, NS, n, dtype=np.uint32)
    req = [rng.integers(0, 1 << 30, n, dtype=np.int64) for _ in range(R)]
    for r in range(R):
        req[r][::9] = 0
        req[r][3::41] = -rng.integers(1, 1000, req[r][3::41].shape[0])     # negative requests reduce usage when admitted
    if n > 100:
        ns[5] = NS
        ns[n // 3] = 0xFFFFFFFF
    mask = rng.integers(0, 256, n, dtype=np.uint8) | np.uint8(0b0110 if n > 40 else 0) if with_mask else None
    if with_mask and n > 40:
        mask[::17] &= np.uint8(0xF9)                                        # some pods omit cpu / memory
    used0 = oracle.usage_reduce(ns[: n // 4 + 1], [c[: n // 4 + 1] for c in req], NS)
    hard, hm = make_hard(used0 * 4 + 1000, 3)
    hm[::6] = rng.integers(0, 1 << R, hm[::6].shape[0], dtype=np.uint8)
    engine.quota_table_set(hard, hm, UNITS[R])
    engine.usage_set(used0)
    want_v, want_u = oracle.admit_greedy(ns, req, mask, NS, used0, hard, hm, _cm(UNITS[R]))
    got_v = engine.admit_greedy(ns, req, mask)
    assert np.array_equal(got_v, want_v)
    assert np.array_equal(engine.usage_get(), want_u)
    if n <= 1000:  # the pure-Python statement agrees too
        pv, pu = Q.admit_greedy(ns, req, mask, NS, used0, hard, hm, _cm(UNITS[R]))
        assert np.array_equal(pv, want_v) and np.array_equal(pu, want_u)


def test_greedy_differs_from_snapshot_and_is_order_dependent(engine, oracle):
    """Two pods that each fit alone but not together: snapshot semantics (K2) admits both,
    sequential semantics admits the first only; swapping them changes who wins."""
    hard = np.array([[10]], np.int64)
    engine.quota_table_set(hard, np.array([1], np.uint8), [0])
    ns = np.zeros(2, np.uint32)
    req = [np.array([6, 5], np.int64)]
    assert engine.quota_eval(ns, req).tolist() == [0, 0]
    assert engine.admit_greedy(ns, req).tolist() == [0, 1]
    assert engine.usage_get().tolist() == [[6]]
    engine.usage_set(np.zeros((1, 1), np.int64))
    assert engine.admit_greedy(ns, [np.array([5, 6], np.int64)]).tolist() == [0, 1]
    assert engine.usage_get().tolist() == [[5]]
    # a later, smaller pod still gets in after a rejection
    engine.usage_set(np.zeros((1, 1), np.int64))
    assert engine.admit_greedy(np.zeros(4, np.uint32), [np.array([6, 5, 4, 1], np.int64)]).tolist() == [0, 1, 0, 1]


@pytest.mark.parametrize("dist", [0, 1])
def test_greedy_generated_workload_device_pointers(bgc, engine, oracle, dist):
    import torch
    R, NS, n = 3, 10_000, 2_000_003
    a, b = Q.scramble_params(NS)
    cdf = Q.zipf_cdf(NS, 1.1)
    run = oracle.gen_fill(oracle.gen_params(SEEDS["C3"], 0, NS, dist, 0, n, a, b, cdf if dist else None, KINDS[R]), 0, n, threads=8)
    cand = oracle.gen_fill(oracle.gen_params(SEEDS["C3"], 1, NS, dist, 0, n, a, b, cdf if dist else None, KINDS[R]), 0, n, want_mask=True, threads=8)
    used0 = oracle.usage_reduce(run[0], run[1], NS, threads=8)
    hard, hm = make_hard(used0, SEEDS["C3"])
    hard = hard + (hard // 8)               # leave some room so that admissions and rejections interleave
    engine.quota_table_set(hard, hm, UNITS[R])
    engine.usage_set(used0)
    want_v, want_u = oracle.admit_greedy(cand[0], cand[1], cand[2], NS, used0, hard, hm, 6)
    d_ns = torch.from_numpy(cand[0].view(np.int32)).cuda()
    d_req = [torch.from_numpy(c).cuda() for c in cand[1]]
    d_mask = torch.from_numpy(cand[2]).cuda()
    d_v = torch.empty(n, dtype=torch.uint8, device="cuda")
    engine.use_torch_stream()
    engine.admit_greedy_dev(d_ns, d_req, d_mask, d_v)
    torch.cuda.synchronize()
    assert np.array_equal(d_v.cpu().numpy(), want_v)
    assert np.array_equal(engine.usage_get(), want_u)
    adm = np.mean(want_v == 0)
    assert 0.02 < adm < 0.98, adm
    # K2 right after sees the updated usage (headroom table is rebuilt)
    v2 = engine.quota_eval(cand[0][:1000], [c[:1000] for c in cand[1]], cand[2][:1000])
    assert np.array_equal(v2, oracle.quota_eval(cand[0][:1000], [c[:1000] for c in cand[1]], cand[2][:1000], NS, want_u, hard, hm, 6))


@pytest.mark.parametrize("R", [1, 3, 4])
@pytest.mark.parametrize("n,NS", [(300_000, 2), (70_000, 1), (100_000, 40)])
def test_greedy_monotone_fast_path_long_segments(engine, oracle, R, n, NS):
    """Non-negative requests only: every tile takes the strike-out fast path, in the warp kernel
    (short segments) and in the CTA-wide kernel (segments >= 32768 pods)."""
    rng = np.random.default_rng(R + n)
    ns = rng.integers(0, NS, n, dtype=np.uint32)
    req = [rng.choice(np.array([0, 1, 2, 4, 8, 250, 1000, 1 << 20, 1 << 32], np.int64), n) for _ in range(R)]
    mask = rng.integers(0, 256, n, dtype=np.uint8) | np.uint8(0b0110)
    mask[::29] &= np.uint8(0xFD)
    used0 = oracle.usage_reduce(ns[: n // 8], [c[: n // 8] for c in req], NS)
    hard, hm = make_hard(used0 * 3, 11)
    engine.quota_table_set(hard, hm, UNITS[R])
    engine.usage_set(used0)
    want_v, want_u = oracle.admit_greedy(ns, req, mask, NS, used0, hard, hm, _cm(UNITS[R]))
    assert np.array_equal(engine.admit_greedy(ns, req, mask), want_v)
    assert np.array_equal(engine.usage_get(), want_u)
    assert 0.01 < np.mean(want_v == 0) < 0.99


@pytest.mark.parametrize("R", [2, 5, 6, 8])
@pytest.mark.parametrize("with_mask", [False, True])
@pytest.mark.parametrize("n,NS", [(33, 2), (50_000, 300), (200_003, 70_000), (70_000, 1)])
def test_greedy_wide_u16_verdicts_up_to_eight_resources(engine, oracle, R, with_mask, n, NS):
    """bgc_admit_greedy16: the six pod-relevant keys of src/synchronizer.rs:249-281 (and up to 8 columns) in one
    pass, u16 verdicts; warp walker, CTA-team walker (one 70,000-pod namespace) and the 3-pass sort (70,000
    namespaces need 17 key bits) against the literal loop."""
    rng = np.random.default_rng(R * 13 + n)
    ns = rng.integers(0, NS, n, dtype=np.uint32)
    req = [rng.integers(0, 1 << 30, n, dtype=np.int64) for _ in range(R)]
    for r in range(R):
        req[r][::9] = 0
        if n < 60_000:
            req[r][3::41] = -rng.integers(1, 1000, req[r][3::41].shape[0])
    if n > 100:
        ns[5] = NS
        ns[n // 3] = 0xFFFFFFFF
    mask = rng.integers(0, 256, n, dtype=np.uint8) | np.uint8(0b0110) if with_mask else None
    used0 = oracle.usage_reduce(ns[: n // 4 + 1], [c[: n // 4 + 1] for c in req], NS)
    hard, hm = make_hard(used0 * 4 + 1000, 3)
    hm[::6] = rng.integers(0, 1 << R, hm[::6].shape[0], dtype=np.uint8)
    engine.quota_table_set(hard, hm, UNITS[R])
    engine.usage_set(used0)
    want_v, want_u = oracle.admit_greedy16(ns, req, mask, NS, used0, hard, hm, _cm(UNITS[R]))
    got_v = engine.admit_greedy(ns, req, mask, wide=True)
    assert np.array_equal(got_v, want_v)
    assert np.array_equal(engine.usage_get(), want_u)
    if R > 4:
        with pytest.raises(Exception):
            engine.admit_greedy(ns, req, mask)          # u8 verdicts cannot hold more than 4 resources


@pytest.mark.parametrize("R,wide", [(1, False), (3, False), (4, False), (6, True), (8, True)])
def test_greedy_long_segments_in_and_out_of_the_monotone_range(engine, oracle, R, wide):
    """Four namespaces of ~60k pods each (all above the 32,768-pod threshold of the CTA-wide walker): ns 0 stays
    inside the monotone fast path (requests in [0, 2^40], usage in [0, 2^62]); ns 1 has one negative request (that
    tile takes the exact alternating path); ns 2 starts with usage above 2^62 (exact path throughout); ns 3 has no
    quota object and its usage crosses 2^62 half-way (fast path first, exact path after).  Verdicts and final usage
    against the literal loop."""
    rng = np.random.default_rng(R)
    n, NS = 240_017, 4
    ns = rng.integers(0, NS, n, dtype=np.uint32)
    req = [rng.choice(np.array([0, 1, 2, 250, 4000, 1 << 20, 1 << 33, 1 << 40], np.int64), n) for _ in range(R)]
    mask = rng.integers(0, 256, n, dtype=np.uint8) | np.uint8(0b0110)
    neg = np.nonzero(ns == 1)[0][777]
    req[0][neg] = -5
    used0 = np.zeros((R, NS), np.int64)
    used0[:, 2] = (1 << 62) + 12345
    used0[:, 3] = (1 << 62) - (30_000 << 40)                # crossed after ~30k admitted 2^40-byte requests
    hard = np.full((R, NS), np.iinfo(np.int64).max, np.int64)
    hard[:, 0] = 7 * (1 << 40)
    hard[:, 1] = 5 * (1 << 40)
    hm = np.array([(1 << R) - 1, (1 << R) - 1, (1 << R) - 1, 0], np.uint8)   # ns 3 has no quota object: everything is admitted
    req[0][ns == 3] = 1 << 40
    engine.quota_table_set(hard, hm, UNITS[R])
    engine.usage_set(used0)
    f = oracle.admit_greedy16 if wide else oracle.admit_greedy
    want_v, want_u = f(ns, req, mask, NS, used0, hard, hm, _cm(UNITS[R]))
    got_v = engine.admit_greedy(ns, req, mask, wide=wide)
    assert np.array_equal(got_v, want_v)
    assert np.array_equal(engine.usage_get(), want_u)
    assert 0.0 < np.mean(want_v[ns == 0] == 0) < 1.0 and np.all(want_v[ns == 3] == 0)


@pytest.mark.parametrize("R,wide", [(1, False), (3, False), (4, False), (6, True)])
@pytest.mark.parametrize("scenario", ["fill_then_slip", "saturated_from_start", "never_full", "negatives", "wrap"])
def test_greedy_chained_tiles(engine, oracle, R, wide, scenario):
    """Segments of hundreds to thousands of 1024-pod tiles, walked by many CTAs at once through the tile descriptors
    (AGG / IDENT / EXACT / OPAQUE, csrc/bgc_k3.cuh): every regime of the chain against the literal loop.
      fill_then_slip        quota reached a third of the way in, small pods keep slipping into what is left
      saturated_from_start  usage starts at the quota: every tile is IDENT, verdict bits still need the exact usage
      never_full            no tile ever rejects: pure prefix sums over AGG descriptors (one namespace without a quota object)
      negatives             a negative request every ~3000 pods: OPAQUE tiles force the sequential hand-over
      wrap                  an unconstrained column crosses 2^63 and wraps (more than 2048 tiles of 2^40-byte requests)"""
    rng = np.random.default_rng(R * 31 + len(scenario))
    NS = 3
    n = 4_600_000 if scenario == "wrap" else 1_500_000
    ns = rng.choice(np.array([0, 0, 0, 1, 2], np.uint32), n) if scenario != "wrap" else np.zeros(n, np.uint32)
    sizes = np.array([0, 1, 1, 2, 3, 8, 64, 250, 1000, 4000], np.int64)
    req = [rng.choice(sizes, n) for _ in range(R)]
    mask = rng.integers(0, 256, n, dtype=np.uint8) | np.uint8(0b0110)
    mask[::53] &= np.uint8(0xFD)                     # some pods omit a required request
    used0 = np.zeros((R, NS), np.int64)
    hm = np.full(NS, (1 << R) - 1, np.uint8)
    total = np.stack([np.bincount(ns, weights=c.astype(np.float64), minlength=NS) for c in req]).astype(np.int64)
    hard = total // 3
    if scenario == "saturated_from_start":
        used0[:] = hard - 1
        used0[0, 1] = hard[0, 1] + 5                 # one namespace already over its quota
    elif scenario == "never_full":
        hard[:, 0] = total[:, 0] + 1                 # fits to the last pod
        hm[1] = 0                                    # no quota object
        hard[:, 2] = np.iinfo(np.int64).max
    elif scenario == "negatives":
        for r in range(R):
            req[r][rng.integers(0, n, n // 3000)] = -rng.integers(1, 5000, n // 3000)
    elif scenario == "wrap":
        hm[:] = 1 if R > 1 else 0                    # the last column is unconstrained (R = 1: no quota at all)
        req[R - 1][:] = 1 << 40
        used0[R - 1, 0] = (1 << 62) - 12345
        hard = np.full((R, NS), np.iinfo(np.int64).max, np.int64)
        if R > 1:
            hard[0, 0] = total[0, 0] - total[0, 0] // 40    # the constrained column fills up near the end
    engine.quota_table_set(hard, hm, UNITS[R])
    engine.usage_set(used0)
    f = oracle.admit_greedy16 if wide else oracle.admit_greedy
    want_v, want_u = f(ns, req, mask, NS, used0, hard, hm, _cm(UNITS[R]))
    got_v = engine.admit_greedy(ns, req, mask, wide=wide)
    assert np.array_equal(got_v, want_v)
    assert np.array_equal(engine.usage_get(), want_u)
    if scenario == "fill_then_slip":
        assert 0.2 < np.mean(want_v == 0) < 0.9
    if scenario == "wrap":
        assert want_u[R - 1, 0] < 0                  # it did wrap


def test_greedy_one_cta_per_long_segment_form(bgc, oracle, monkeypatch):
    """BGC_K3_LONG=team: the round-2 walker that gives a long segment to ONE CTA is kept for comparison; it must stay
    bit-exact too (a fresh context reads the knob at its first greedy call)."""
    monkeypatch.setenv("BGC_K3_LONG", "team")
    eng = bgc.Engine(0)
    try:
        rng = np.random.default_rng(77)
        R, NS, n = 3, 3, 400_000
        ns = rng.choice(np.array([0, 0, 0, 1, 2], np.uint32), n)
        req = [rng.choice(np.array([0, 1, 2, 8, 250, 4000], np.int64), n) for _ in range(R)]
        req[1][rng.integers(0, n, 50)] = -7
        mask = rng.integers(0, 256, n, dtype=np.uint8) | np.uint8(0b0110)
        total = np.stack([np.bincount(ns, weights=c.astype(np.float64), minlength=NS) for c in req]).astype(np.int64)
        hard, hm = total // 3, np.full(NS, 7, np.uint8)
        used0 = np.zeros((R, NS), np.int64)
        eng.quota_table_set(hard, hm, UNITS[R])
        eng.usage_set(used0)
        want_v, want_u = oracle.admit_greedy(ns, req, mask, NS, used0, hard, hm, _cm(UNITS[R]))
        assert np.array_equal(eng.admit_greedy(ns, req, mask), want_v)
        assert np.array_equal(eng.usage_get(), want_u)
    finally:
        eng.close()
