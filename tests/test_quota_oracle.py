"""CPU tests: the quota oracle (rows A3/A4; PARITY UNPINNED — no reference implementation) in C
against its independent numpy statement, the generator spec in C against numpy, and the
size-independent properties SURVEY.md §8(c) lists."""
import numpy as np
import pytest
from hypothesis import given, settings, strategies as st

from oracle import quota_ref as Q
from workloads import make_hard, KINDS


def _rand_case(rng, n, NS, R, wild=False):
    ns = rng.integers(0, NS, n, dtype=np.uint32)
    if wild:
        req = [rng.integers(-2**63, 2**63 - 1, n, dtype=np.int64) for _ in range(R)]
    else:
        req = [rng.integers(0, 1 << 36, n, dtype=np.int64) for _ in range(R)]
    return ns, req


@pytest.mark.parametrize("R", [1, 2, 3, 4])
@pytest.mark.parametrize("n,NS", [(0, 5), (1, 1), (1000, 7), (50_000, 1000)])
def test_usage_reduce_c_vs_numpy(oracle, R, n, NS):
    rng = np.random.default_rng(R * 1000 + n + NS)
    ns, req = _rand_case(rng, n, NS, R, wild=True)
    if n > 10:
        ns[3] = NS + 5  # out-of-range id contributes nothing
    a = oracle.usage_reduce(ns, req, NS)
    b = Q.usage_reduce(ns, req, NS)
    assert np.array_equal(a, b)
    c = oracle.usage_reduce(ns, req, NS, threads=4)
    assert np.array_equal(a, c)


@pytest.mark.parametrize("R", [1, 2, 3, 4])
@pytest.mark.parametrize("with_mask", [False, True])
def test_quota_eval_c_vs_numpy(oracle, R, with_mask):
    rng = np.random.default_rng(77 + R)
    n, NS = 40_000, 300
    ns, req = _rand_case(rng, n, NS, R)
    ns[::97] = NS + rng.integers(0, 10, ns[::97].shape[0], dtype=np.uint32)
    for r in range(R):
        req[r][::13] = 0
        req[r][5::31] = -rng.integers(1, 100, req[r][5::31].shape[0])
    used = oracle.usage_reduce(ns, req, NS)
    hard, hard_mask = make_hard(used, 5)
    hard_mask[::7] = rng.integers(0, 1 << R, hard_mask[::7].shape[0], dtype=np.uint8)
    mask = rng.integers(0, 256, n, dtype=np.uint8) if with_mask else None
    cm = 0b0110 & ((1 << R) - 1)
    cns, creq = _rand_case(rng, n, NS, R)
    cns[::101] = 0xFFFFFFFF
    a = oracle.quota_eval(cns, creq, mask, NS, used, hard, hard_mask, cm)
    b = Q.quota_eval(cns, creq, mask, NS, used, hard, hard_mask, cm)
    assert np.array_equal(a, b)
    assert np.array_equal(a, oracle.quota_eval(cns, creq, mask, NS, used, hard, hard_mask, cm, threads=4))
    assert (a[::101] == 0xFF).all()


def test_quota_eval_is_overflow_free(oracle):
    # used + req is compared with hard as integers (no 64-bit wrap): a sum past INT64_MAX exceeds
    NS = 1
    used = np.array([[2**63 - 10]], np.int64)
    hard = np.array([[2**63 - 1]], np.int64)
    hm = np.array([1], np.uint8)
    ns = np.zeros(3, np.uint32)
    req = [np.array([5, 9, 100], np.int64)]
    v = oracle.quota_eval(ns, req, None, NS, used, hard, hm, 0)
    assert v.tolist() == [0, 0, 1]          # fits, exactly hard, overflows → exceeded
    assert np.array_equal(v, Q.quota_eval(ns, req, None, NS, used, hard, hm, 0))
    # negative usage (a wrapped or corrected table) and extreme limits
    used = np.array([[-2**63, -5, 2**62, 0]], np.int64)
    hard = np.array([[2**63 - 1, -1, -2**63, 2**63 - 1]], np.int64)
    hm = np.ones(4, np.uint8)
    ns = np.array([0, 0, 1, 1, 2, 3], np.uint32)
    req = [np.array([2**63 - 1, 1, 4, 5, 1, 2**63 - 1], np.int64)]
    v = oracle.quota_eval(ns, req, None, 4, used, hard, hm, 0)
    assert v.tolist() == [0, 0, 0, 1, 1, 0]
    assert np.array_equal(v, Q.quota_eval(ns, req, None, 4, used, hard, hm, 0))


def test_semantics_examples(oracle):
    """Hand-worked cases of row A4."""
    NS, R = 3, 3
    used = np.array([[1, 0, 0], [4000, 0, 0], [8 << 30, 0, 0]], np.int64)
    hard = np.array([[2, 0, 5], [8000, 0, 5], [16 << 30, 0, 5]], np.int64)
    hard_mask = np.array([0b111, 0b000, 0b010], np.uint8)   # ns1 has no quota; ns2 constrains cpu only
    cm = 0b110
    ns = np.array([0, 0, 0, 1, 2, 2, 0, 0], np.uint32)
    req = [np.array([1, 2, 0, 99, 9, 9, 1, 1], np.int64),
           np.array([4000, 4001, 0, 10**9, 5, 6, 1, 1], np.int64),
           np.array([8 << 30, (8 << 30) + 1, 0, 1 << 50, 1 << 50, 0, 1, 1], np.int64)]
    mask = np.array([0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0xFF, 0b001, 0b110], np.uint8)
    v = oracle.quota_eval(ns, req, mask, NS, used, hard, hard_mask, cm)
    assert v.tolist() == [
        0,            # exactly fills every limit
        0b111,        # exceeds all three
        0,            # zero requests never exceed
        0,            # namespace without quota → allow
        0,            # only cpu constrained: 5 <= 5
        0b010,        # cpu 6 > 5
        0b0110 << 4,  # cpu and memory not stated but constrained → missing; gpu 1+1<=2 ok
        0,            # gpu not stated → counts as 0 (extended resource); cpu/mem fit
    ]
    assert np.array_equal(v, Q.quota_eval(ns, req, mask, NS, used, hard, hard_mask, cm))


@settings(max_examples=25, deadline=None)
@given(st.integers(0, 2**32 - 1), st.integers(1, 400), st.integers(1, 40), st.integers(1, 4))
def test_properties(seed, n, NS, R):
    import oracle
    rng = np.random.default_rng(seed)
    ns, req = _rand_case(rng, n, NS, R, wild=True)
    used = oracle.usage_reduce(ns, req, NS)
    # permutation invariance (wrapping add is associative + commutative)
    perm = rng.permutation(n)
    assert np.array_equal(used, oracle.usage_reduce(ns[perm], [c[perm] for c in req], NS))
    # shard union == unsharded (row E)
    G = int(rng.integers(1, 5))
    shard = ns % G
    total = np.zeros_like(used)
    for g in range(G):
        sel = shard == g
        with np.errstate(over="ignore"):
            total += oracle.usage_reduce(ns[sel], [c[sel] for c in req], NS)
    assert np.array_equal(total, used)
    # hard_mask == 0 → allow; monotone in hard
    ns2, req2 = _rand_case(rng, n, NS, R)
    used2 = oracle.usage_reduce(ns2, req2, NS)
    hard, _ = make_hard(used2, seed)
    zero = np.zeros(NS, np.uint8)
    assert not oracle.quota_eval(ns2, req2, None, NS, used2, hard, zero, 0).any()
    full = np.full(NS, (1 << R) - 1, np.uint8)
    v1 = oracle.quota_eval(ns2, req2, None, NS, used2, hard, full, 0)
    v2 = oracle.quota_eval(ns2, req2, None, NS, used2, hard + 1000, full, 0)
    assert ((v2 & ~v1) == 0).all()   # raising hard never adds an exceeded bit


@pytest.mark.parametrize("dist,grouped", [(0, 0), (1, 0), (0, 1), (1, 1)])
def test_generator_c_vs_numpy(oracle, dist, grouped):
    NS, n_total, R = 1000, 20_000, 4
    cdf = Q.zipf_cdf(NS, 1.1)
    a, b = Q.scramble_params(NS)
    p = oracle.gen_params(0xB2000003, 1, NS, dist, grouped, n_total, a, b, cdf if dist else None, KINDS[R])
    ns, req, mask = oracle.gen_fill(p, 5_000, 7_001, want_mask=True)
    ns2, req2, mask2 = Q.gen_fill(0xB2000003, 1, NS, dist, grouped, n_total, a, b, cdf, KINDS[R], 5_000, 7_001, True)
    assert np.array_equal(ns, ns2) and np.array_equal(mask, mask2)
    for x, y in zip(req, req2):
        assert np.array_equal(x, y)
    # threads and slicing do not change values (counter-based)
    ns3, req3, _ = oracle.gen_fill(p, 5_000, 7_001, want_mask=False, threads=4)
    assert np.array_equal(ns, ns3) and np.array_equal(req[2], req3[2])
    ns4, _, _ = oracle.gen_fill(p, 6_000, 100)
    assert np.array_equal(ns[1000:1100], ns4)
    assert ns.max() < NS
    if grouped:
        # grouped mode: each namespace forms one contiguous run over the whole batch
        full, _, _ = oracle.gen_fill(p, 0, n_total)
        change = np.count_nonzero(full[1:] != full[:-1]) + 1
        assert change == len(np.unique(full))


def test_generator_distributions(oracle):
    NS, n = 100, 200_000
    a, b = Q.scramble_params(NS)
    p = oracle.gen_params(1, 0, NS, 0, 0, n, a, b, None, KINDS[3])
    ns, req, _ = oracle.gen_fill(p, 0, n)
    assert abs(np.mean(req[0] == 0) - 0.5) < 0.01 and set(np.unique(req[0])) == {0, 1, 2, 4, 8}
    assert set(np.unique(req[1])) == {250, 500, 1000, 2000, 4000, 8000, 16000}
    assert set(np.unique(req[2])) == {1 << k for k in range(28, 37)}
    cnt = np.bincount(ns, minlength=NS)
    assert cnt.min() > 0.8 * n / NS and cnt.max() < 1.2 * n / NS
    cdf = Q.zipf_cdf(NS, 1.1)
    pz = oracle.gen_params(1, 0, NS, 1, 0, n, a, b, cdf, KINDS[3])
    nz, _, _ = oracle.gen_fill(pz, 0, n)
    top = np.bincount(nz, minlength=NS).max() / n
    assert 0.15 < top < 0.25      # Zipf(1.1) over 100 ranks: hottest ≈ 19 %


def test_checksum_is_additive_and_position_keyed(oracle):
    rng = np.random.default_rng(3)
    v = rng.integers(0, 256, 10_000, dtype=np.uint8)
    whole = oracle.checksum_u8(v, 100)
    parts = (oracle.checksum_u8(v[:4000], 100) + oracle.checksum_u8(v[4000:], 4100)) & (2**64 - 1)
    assert whole == parts
    w = v.copy(); w[[1, 2]] = w[[2, 1]]
    if v[1] != v[2]:
        assert oracle.checksum_u8(w, 100) != whole


@pytest.mark.parametrize("R", [1, 2, 3, 4])
def test_greedy_oracle_c_vs_python(oracle, R):
    rng = np.random.default_rng(50 + R)
    n, NS = 4000, 30
    ns = rng.integers(0, NS + 2, n, dtype=np.uint32)
    req = [rng.integers(-3, 1 << 20, n, dtype=np.int64) for _ in range(R)]
    mask = rng.integers(0, 256, n, dtype=np.uint8)
    used0 = oracle.usage_reduce(ns[:500], [c[:500] for c in req], NS)
    hard, hm = make_hard(used0 * 3, R)
    cm = 6 & ((1 << R) - 1)
    a = oracle.admit_greedy(ns, req, mask, NS, used0, hard, hm, cm)
    b = Q.admit_greedy(ns, req, mask, NS, used0, hard, hm, cm)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    # with unlimited quota the greedy pass admits everything and ends at used0 + usage of the batch
    big = np.full_like(hard, 2**62)
    v, u = oracle.admit_greedy(ns, req, None, NS, used0, big, hm, cm)
    ok = ns < NS
    assert (v[ok] == 0).all() and (v[~ok] == 0xFF).all()
    with np.errstate(over="ignore"):
        assert np.array_equal(u, used0 + oracle.usage_reduce(ns, req, NS))
