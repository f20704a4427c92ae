"""GPU parity tests for more than four resources (the six pod-relevant quota keys
of src/synchronizer.rs:249-281 in one pass; u16 verdicts = missing << 8 | exceeded)."""
import numpy as np
import pytest

from oracle import quota_ref as Q
from workloads import make_hard, KINDS, UNITS, SEEDS, C1_KEYS6, c1_namespaces, c1_pods

pytestmark = pytest.mark.gpu


def _cm(units):
    return sum(1 << r for r, u in enumerate(units) if u != 0)


@pytest.mark.parametrize("R", [5, 6, 7, 8])
@pytest.mark.parametrize("variant", [0, 1, 2, 3])
@pytest.mark.parametrize("n,NS", [(1, 1), (3, 2), (129, 10), (4099, 1000), (250_007, 5_000), (300_000, 70_000)])
def test_k1_wide_random_values(engine, oracle, R, variant, n, NS):
    rng = np.random.default_rng(R * 100 + variant * 10 + n)
    ns = rng.integers(0, NS, n, dtype=np.uint32)
    req = [rng.integers(-2**63, 2**63 - 1, n, dtype=np.int64) for _ in range(R)]
    if n > 50:
        ns[7] = NS
        ns[n // 2] = 0xFFFFFFFF
    want = oracle.usage_reduce(ns, req, NS, threads=4)
    engine.quota_table_set(np.zeros((R, NS), np.int64), np.zeros(NS, np.uint8), UNITS[R])
    engine.set_k1_variant(variant)
    for flags in (0, 1):
        assert np.array_equal(engine.usage_reduce(ns, req, flags=flags), want), (engine.last_k1_variant(), flags)


@pytest.mark.parametrize("R", [1, 2, 3, 4, 5, 6, 7, 8])
@pytest.mark.parametrize("with_mask", [False, True])
@pytest.mark.parametrize("n", [1, 2, 3, 5, 1023, 300_009])
def test_k2_u16_verdicts_vs_oracle(engine, oracle, R, with_mask, n):
    rng = np.random.default_rng(R + 10 * n)
    NS = 777
    ns = rng.integers(0, NS, max(n, 5000), dtype=np.uint32)
    req = [rng.integers(0, 1 << 36, ns.shape[0], dtype=np.int64) for _ in range(R)]
    used = oracle.usage_reduce(ns, req, NS, threads=4)
    hard, hm = make_hard(used, 1)
    hm[::5] = rng.integers(0, 1 << R, hm[::5].shape[0], dtype=np.uint8)
    engine.quota_table_set(hard, hm, UNITS[R])
    engine.usage_set(used)
    cns = rng.integers(0, NS, n, dtype=np.uint32)
    creq = [rng.integers(-5, 1 << 34, n, dtype=np.int64) for _ in range(R)]
    if n > 100:
        cns[::53] = NS + 1
    mask = rng.integers(0, 256, n, dtype=np.uint8) if with_mask else None
    want = oracle.quota_eval16(cns, creq, mask, NS, used, hard, hm, _cm(UNITS[R]), threads=4)
    got = engine.quota_eval(cns, creq, mask, wide=True)
    assert got.dtype == np.uint16 and np.array_equal(got, want)
    assert np.array_equal(want, Q.quota_eval(cns, creq, mask, NS, used, hard, hm, _cm(UNITS[R]), wide=True))
    if R <= 4:   # the u8 form says the same thing
        v8 = engine.quota_eval(cns, creq, mask)
        assert np.array_equal(np.where(want == 0xFFFF, 0xFF, ((want >> 8) << 4) | (want & 0xFF)).astype(np.uint8), v8)


def test_u8_entry_points_refuse_more_than_four_resources(bgc, engine):
    R, NS = 6, 10
    engine.quota_table_set(np.zeros((R, NS), np.int64), np.zeros(NS, np.uint8), UNITS[R])
    cols = [np.zeros(4, np.int64)] * R
    with pytest.raises(bgc.BgcError) as e:
        engine.quota_eval(np.zeros(4, np.uint32), cols)
    assert e.value.code == bgc.E_BADARG and "bgc_quota_eval16" in str(e.value)
    with pytest.raises(bgc.BgcError):
        engine.admit_greedy(np.zeros(4, np.uint32), cols)
    with pytest.raises(bgc.BgcError):
        bgc.Engine(0).quota_table_set(np.zeros((9, NS), np.int64), np.zeros(NS, np.uint8), [0] * 9)


def test_six_quota_keys_in_one_pass_device_pointers(bgc, engine, oracle):
    """Generated workload, R = 6, device pointers, misaligned slices, every K1 variant."""
    import torch
    R, NS, n = 6, 10_000, 2_000_003
    a, b = Q.scramble_params(NS)
    cdf = Q.zipf_cdf(NS, 1.1)
    run = oracle.gen_fill(oracle.gen_params(SEEDS["C3"], 0, NS, 1, 0, n, a, b, cdf, KINDS[R]), 0, n + 8, threads=8)
    cand = oracle.gen_fill(oracle.gen_params(SEEDS["C3"], 1, NS, 1, 0, n, a, b, cdf, KINDS[R]), 0, n + 8, want_mask=True, threads=8)
    engine.use_torch_stream()
    d_ns = torch.from_numpy(run[0].view(np.int32)).cuda()
    d_req = [torch.from_numpy(c).cuda() for c in run[1]]
    d_used = torch.empty(R * NS, dtype=torch.int64, device="cuda")
    engine.quota_table_set(np.zeros((R, NS), np.int64), np.zeros(NS, np.uint8), UNITS[R])
    for off in (0, 1, 3):
        want = oracle.usage_reduce(run[0][off:off + n], [c[off:off + n] for c in run[1]], NS, threads=8)
        for variant in (0, 1, 3):
            engine.set_k1_variant(variant)
            engine.usage_reduce_dev(d_ns[off:off + n], [c[off:off + n] for c in d_req], 0, d_used)
            torch.cuda.synchronize()
            assert np.array_equal(d_used.cpu().numpy().reshape(R, NS), want), (off, engine.last_k1_variant())
    used = oracle.usage_reduce(run[0][:n], [c[:n] for c in run[1]], NS, threads=8)
    hard, hm = make_hard(used, 5)
    engine.quota_table_set(hard, hm, UNITS[R])
    engine.usage_set(used)
    c_ns = torch.from_numpy(cand[0].view(np.int32)).cuda()
    c_req = [torch.from_numpy(c).cuda() for c in cand[1]]
    c_mask = torch.from_numpy(cand[2]).cuda()
    for off in (0, 1, 2):
        d_v = torch.zeros(n + 8, dtype=torch.int16, device="cuda")
        engine.quota_eval_dev(c_ns[off:off + n], [c[off:off + n] for c in c_req], c_mask[off:off + n], d_v[off:off + n])
        torch.cuda.synchronize()
        want = oracle.quota_eval16(cand[0][off:off + n], [c[off:off + n] for c in cand[1]], cand[2][off:off + n], NS, used, hard, hm, _cm(UNITS[R]), threads=8)
        assert np.array_equal(d_v[off:off + n].cpu().numpy().view(np.uint16), want), off
        assert d_v[:off].abs().sum().item() == 0 and d_v[off + n:].abs().sum().item() == 0


def test_c1_pods_with_all_six_keys(bgc, engine, oracle):
    """C1b with the full pod-relevant key set: flattener (R = 6) -> K1 -> K2 u16, vs the oracle on the same columns."""
    namespaces = c1_namespaces()
    nsmap = bgc.pods.NamespaceMap()
    for ns in namespaces:
        nsmap.intern(ns["name"])
    R = len(C1_KEYS6)
    hard = np.zeros((R, len(namespaces)), np.int64)
    hm = np.zeros(len(namespaces), np.uint8)
    for i, ns in enumerate(namespaces):
        h, unit = bgc.pods.hard_from_sheet_row(ns["cpu"], ns["memory"], ns["gpu"], ns["storage"], ns["mig"], C1_KEYS6)
        hard[:, i] = h
        hm[i] = (1 << R) - 1 if ns["has_quota"] else 0
    assert unit.tolist() == UNITS[6]
    engine.quota_table_set(hard, hm, unit)
    r_ns, r_req, r_mask, _ = bgc.pods.flatten_pods(c1_pods(10_000, namespaces, running=True), C1_KEYS6, nsmap, intern=False)
    c_ns, c_req, c_mask, _ = bgc.pods.flatten_pods(c1_pods(10_000, namespaces, running=False), C1_KEYS6, nsmap, intern=False)
    used = engine.usage_reduce(r_ns, r_req)
    assert np.array_equal(used, oracle.usage_reduce(r_ns, r_req, len(namespaces)))
    got = engine.quota_eval(c_ns, c_req, c_mask, wide=True)
    want = oracle.quota_eval16(c_ns, c_req, c_mask, len(namespaces), used, hard, hm, _cm(UNITS[6]))
    assert np.array_equal(got, want)
    assert (got != 0).any() and (got == 0).any()
    # limits.* columns really came from the containers' limits (init containers carry none -> mask bit clear)
    assert (c_mask & 0b110000 != 0b110000).any()
