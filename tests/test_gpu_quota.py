"""GPU parity tests for K1 usage_reduce (row A3) and K2 quota_eval (row A4), through the C ABI,
against the CPU oracle on the same seeded inputs.  Bit-exact (int64 / byte work)."""
import numpy as np
import pytest

from oracle import quota_ref as Q
from workloads import make_hard, KINDS, UNITS, SEEDS

pytestmark = pytest.mark.gpu

VARIANTS = [1, 2, 3]  # BGC_K1_GLOBAL, SMEM_DIRECT, SMEM_HASH
SHAPES = [v | (sh << 8) for v in (2, 3) for sh in (1, 2, 3)]  # scatter-kernel launch shapes (bits 8..15)


def _table(engine, oracle, run, NS, R, seed=1, units=None):
    used = oracle.usage_reduce(run[0], run[1], NS, threads=4)
    hard, hm = make_hard(used, seed)
    engine.quota_table_set(hard, hm, units or UNITS[R])
    return used, hard, hm


def _compute_mask(units):
    return sum(1 << r for r, u in enumerate(units) if u != 0)


@pytest.mark.parametrize("R", [1, 2, 3, 4])
@pytest.mark.parametrize("variant", VARIANTS + SHAPES)
@pytest.mark.parametrize("n,NS", [(1, 1), (127, 3), (128, 10), (129, 10), (4099, 1000), (250_007, 10_000), (300_000, 70_000)])
def test_k1_variants_random_values(engine, oracle, R, variant, n, NS):
    rng = np.random.default_rng(R * 100 + variant * 10 + n)
    ns = rng.integers(0, NS, n, dtype=np.uint32)
    req = [rng.integers(-2**63, 2**63 - 1, n, dtype=np.int64) for _ in range(R)]   # full-range: wraps, carries, negatives
    if n > 50:
        ns[7] = NS            # out-of-range ids are ignored
        ns[n // 2] = 0xFFFFFFFF
    want = oracle.usage_reduce(ns, req, NS, threads=4)
    engine.quota_table_set(np.zeros((R, NS), np.int64), np.zeros(NS, np.uint8), UNITS[R])
    engine.set_k1_variant(variant)
    got = engine.usage_reduce(ns, req)
    assert np.array_equal(got, want), engine.last_k1_variant()
    assert np.array_equal(engine.usage_get(), want)
    if variant < 256:
        got = engine.usage_reduce(ns, req, flags=1)      # grouped hint on shuffled input: slower, still exact
        assert np.array_equal(got, want), engine.last_k1_variant()


@pytest.mark.parametrize("variant", [0] + VARIANTS)
@pytest.mark.parametrize("dist,grouped", [(0, 0), (1, 0), (0, 1), (1, 1)])
@pytest.mark.parametrize("NS,R", [(100, 3), (10_000, 2), (10_000, 3), (200_000, 3)])
def test_k1_generated_distributions(engine, oracle, variant, dist, grouped, NS, R):
    n = 600_011
    a, b = Q.scramble_params(NS)
    cdf = Q.zipf_cdf(NS, 1.1)
    p = oracle.gen_params(SEEDS["C3"], 0, NS, dist, grouped, n, a, b, cdf if dist else None, KINDS[R])
    ns, req, _ = oracle.gen_fill(p, 0, n, threads=4)
    want = oracle.usage_reduce(ns, req, NS, threads=4)
    engine.quota_table_set(np.zeros((R, NS), np.int64), np.zeros(NS, np.uint8), UNITS[R])
    engine.set_k1_variant(variant)
    got = engine.usage_reduce(ns, req, flags=1 if grouped else 0)
    assert np.array_equal(got, want), engine.last_k1_variant()


def test_k1_long_runs_and_run_boundaries(engine, oracle):
    """Runs that straddle lane, tile (128) and warp-chunk boundaries; one namespace owning everything."""
    R, NS = 2, 50
    for n, lens in [(5000, [1, 2, 3, 4, 5, 127, 128, 129, 255, 256, 257, 1000]), (100_000, [100_000]), (70_000, [33_333, 1, 36_666])]:
        ns = np.concatenate([np.full(L, i % NS, np.uint32) for i, L in enumerate(lens * (n // sum(lens) + 1))])[:n]
        req = [np.arange(n, dtype=np.int64) * 3 + 1, np.full(n, 1 << 40, np.int64)]
        want = oracle.usage_reduce(ns, req, NS)
        engine.quota_table_set(np.zeros((R, NS), np.int64), np.zeros(NS, np.uint8), UNITS[R])
        for variant in [0] + VARIANTS:
            engine.set_k1_variant(variant)
            for flags in (0, 1):
                assert np.array_equal(engine.usage_reduce(ns, req, flags=flags), want), (n, variant, flags)


def test_k1_accumulate_flag_and_empty_batch(engine, oracle):
    R, NS, n = 3, 321, 50_000
    rng = np.random.default_rng(4)
    ns = rng.integers(0, NS, n, dtype=np.uint32)
    req = [rng.integers(0, 1 << 40, n, dtype=np.int64) for _ in range(R)]
    want = oracle.usage_reduce(ns, req, NS)
    engine.quota_table_set(np.zeros((R, NS), np.int64), np.zeros(NS, np.uint8), UNITS[R])
    h = n // 3
    engine.usage_reduce(ns[:h], [c[:h] for c in req], want_used=False)
    got = engine.usage_reduce(ns[h:], [c[h:] for c in req], flags=2)        # BGC_REDUCE_ACCUMULATE
    assert np.array_equal(got, want)
    assert np.array_equal(engine.usage_reduce(ns[:0], [c[:0] for c in req]), np.zeros((R, NS), np.int64))
    engine.usage_set(want)
    assert np.array_equal(engine.usage_get(), want)


def test_k1_device_pointers_misaligned_and_multichunk(engine, oracle):
    import torch
    R, NS, n = 3, 5000, 9_000_001          # > 2 host chunks of 4 Mi pods
    a, b = Q.scramble_params(NS)
    p = oracle.gen_params(11, 0, NS, 0, 0, n, a, b, None, KINDS[R])
    ns, req, _ = oracle.gen_fill(p, 0, n + 8, threads=4)
    engine.quota_table_set(np.zeros((R, NS), np.int64), np.zeros(NS, np.uint8), UNITS[R])
    want = oracle.usage_reduce(ns[:n], [c[:n] for c in req], NS, threads=4)
    assert np.array_equal(engine.usage_reduce(ns[:n], [c[:n] for c in req]), want)   # host path, 3 chunks
    t = engine.timings()
    assert t["h2d_ns"] > 0 and t["kernel_ns"] > 0
    d_ns = torch.from_numpy(ns.view(np.int32)).cuda()
    d_req = [torch.from_numpy(c).cuda() for c in req]
    d_used = torch.empty(R * NS, dtype=torch.int64, device="cuda")
    engine.use_torch_stream()
    for off in (0, 1, 3):
        w = oracle.usage_reduce(ns[off:off + n], [c[off:off + n] for c in req], NS, threads=4)
        engine.usage_reduce_dev(d_ns[off:off + n], [c[off:off + n] for c in d_req], 0, d_used)
        torch.cuda.synchronize()
        assert np.array_equal(d_used.cpu().numpy().reshape(R, NS), w), off


@pytest.mark.parametrize("R", [1, 2, 3, 4])
@pytest.mark.parametrize("with_mask", [False, True])
@pytest.mark.parametrize("n", [1, 3, 4, 5, 1023, 500_009])
def test_k2_vs_oracle(engine, oracle, R, with_mask, n):
    rng = np.random.default_rng(R + 10 * n)
    NS = 777
    ns = rng.integers(0, NS, max(n, 5000), dtype=np.uint32)
    req = [rng.integers(0, 1 << 36, ns.shape[0], dtype=np.int64) for _ in range(R)]
    used, hard, hm = _table(engine, oracle, (ns, req), NS, R)
    hm[::5] = rng.integers(0, 1 << R, hm[::5].shape[0], dtype=np.uint8)      # partially constrained namespaces
    engine.quota_table_set(hard, hm, UNITS[R])
    engine.usage_set(used)
    cns = rng.integers(0, NS, n, dtype=np.uint32)
    creq = [rng.integers(-5, 1 << 34, n, dtype=np.int64) for _ in range(R)]
    if n > 100:
        cns[::53] = NS + 1
        cns[1::97] = 0xFFFFFFFF
    mask = rng.integers(0, 256, n, dtype=np.uint8) if with_mask else None
    want = oracle.quota_eval(cns, creq, mask, NS, used, hard, hm, _compute_mask(UNITS[R]))
    got = engine.quota_eval(cns, creq, mask)
    assert np.array_equal(got, want)


def test_k2_overflow_free_compare_and_extremes(engine, oracle):
    NS, R = 4, 2
    used = np.array([[2**63 - 10, -5, 0, 2**62], [0, 0, -2**63, 1]], np.int64)
    hard = np.array([[2**63 - 1, 0, 0, 2**62], [0, -1, 2**63 - 1, 1]], np.int64)
    hm = np.array([3, 3, 3, 3], np.uint8)
    engine.quota_table_set(hard, hm, [0, 0])
    engine.usage_set(used)
    ns = np.array([0, 0, 0, 1, 1, 2, 2, 3, 3], np.uint32)
    req = [np.array([5, 9, 100, 5, 6, 1, 0, 1, 2**62], np.int64), np.array([0, 1, 2**63 - 1, 0, 1, 2**63 - 1, 5, 0, 0], np.int64)]
    want = oracle.quota_eval(ns, req, None, NS, used, hard, hm, 0)
    assert np.array_equal(engine.quota_eval(ns, req), want)
    # saturating headroom: every combination of extreme used / hard / req
    ext = np.array([-2**63, -2**63 + 1, -1, 0, 1, 2**62, 2**63 - 2, 2**63 - 1], np.int64)
    U, H = np.meshgrid(ext, ext, indexing="ij")
    NS2 = U.size
    engine.quota_table_set(H.reshape(1, NS2).copy(), np.ones(NS2, np.uint8), [0])
    engine.usage_set(U.reshape(1, NS2).copy())
    ns2 = np.repeat(np.arange(NS2, dtype=np.uint32), ext.size)
    q2 = [np.tile(ext, NS2)]
    want2 = oracle.quota_eval(ns2, q2, None, NS2, U.reshape(1, NS2), H.reshape(1, NS2), np.ones(NS2, np.uint8), 0)
    assert np.array_equal(engine.quota_eval(ns2, q2), want2)
    py = np.array([int(q > 0 and int(u) + int(q) > int(h)) for u, h in zip(U.ravel(), H.ravel()) for q in ext], np.uint8)
    assert np.array_equal(want2, py)


def test_k2_device_pointers_misaligned(engine, oracle):
    import torch
    R, NS, n = 3, 2000, 400_003
    rng = np.random.default_rng(8)
    ns = rng.integers(0, NS, n + 8, dtype=np.uint32)
    req = [rng.integers(0, 1 << 36, n + 8, dtype=np.int64) for _ in range(R)]
    mask = rng.integers(0, 256, n + 8, dtype=np.uint8)
    used, hard, hm = _table(engine, oracle, (ns, req), NS, R)
    engine.usage_set(used)
    d_ns = torch.from_numpy(ns.view(np.int32)).cuda()
    d_req = [torch.from_numpy(c).cuda() for c in req]
    d_mask = torch.from_numpy(mask).cuda()
    engine.use_torch_stream()
    for off in (0, 1, 2, 4):
        d_v = torch.zeros(n + 8, dtype=torch.uint8, device="cuda")
        engine.quota_eval_dev(d_ns[off:off + n], [c[off:off + n] for c in d_req], d_mask[off:off + n], d_v[off:off + n])
        torch.cuda.synchronize()
        want = oracle.quota_eval(ns[off:off + n], [c[off:off + n] for c in req], mask[off:off + n], NS, used, hard, hm, 6)
        assert np.array_equal(d_v[off:off + n].cpu().numpy(), want), off
        assert d_v[:off].sum().item() == 0 and d_v[off + n:].sum().item() == 0


def test_errors_are_reported_not_thrown(bgc, engine):
    R, NS = 2, 10
    with pytest.raises(bgc.BgcError) as e:
        engine.usage_get()
    assert e.value.code == bgc.E_STATE
    with pytest.raises(bgc.BgcError) as e:
        bgc.Engine(0).quota_table_set(np.zeros((9, NS), np.int64), np.zeros(NS, np.uint8), [0] * 9)
    assert e.value.code == bgc.E_BADARG
    engine.quota_table_set(np.zeros((R, NS), np.int64), np.zeros(NS, np.uint8), [0, 1])
    with pytest.raises(ValueError):
        engine.usage_reduce(np.zeros(4, np.uint32), [np.zeros(4, np.int64)])      # wrong column count


def test_c2_full_size_bit_exact(engine, oracle):
    """BASELINE.json configs[1]: 1M pods x 10k namespaces, uniform, R=2 (gpu, cpu), bit-exact."""
    n, NS, R = 1_048_576, 10_000, 2
    a, b = Q.scramble_params(NS)
    run = oracle.gen_fill(oracle.gen_params(SEEDS["C2"], 0, NS, 0, 0, n, a, b, None, KINDS[R]), 0, n, threads=4)
    cand = oracle.gen_fill(oracle.gen_params(SEEDS["C2"], 1, NS, 0, 0, n, a, b, None, KINDS[R]), 0, n, threads=4)
    used, hard, hm = _table(engine, oracle, run, NS, R, seed=SEEDS["C2"])
    for variant in [0] + VARIANTS:
        engine.set_k1_variant(variant)
        assert np.array_equal(engine.usage_reduce(run[0], run[1]), used)
    got = engine.quota_eval(cand[0], cand[1])
    want = oracle.quota_eval(cand[0], cand[1], None, NS, used, hard, hm, 2, threads=4)
    assert np.array_equal(got, want)
    deny = np.mean(want != 0)
    assert 0.3 < deny < 0.7, deny            # the workload is meant to split verdicts roughly in half
    # exactly 10^6 pods as well
    m = 1_000_000
    u2 = oracle.usage_reduce(run[0][:m], [c[:m] for c in run[1]], NS, threads=4)
    assert np.array_equal(engine.usage_reduce(run[0][:m], [c[:m] for c in run[1]]), u2)


def test_device_generator_matches_oracle(engine, oracle):
    import torch
    R, NS, n = 4, 10_000, 1_000_003
    a, b = Q.scramble_params(NS)
    cdf = Q.zipf_cdf(NS, 1.1)
    d_cdf = torch.from_numpy(cdf.view(np.int64)).cuda()
    engine.use_torch_stream()
    for dist, grouped in [(0, 0), (1, 0), (0, 1), (1, 1)]:
        d_ns = torch.empty(n, dtype=torch.int32, device="cuda")
        d_req = [torch.empty(n, dtype=torch.int64, device="cuda") for _ in range(R)]
        d_mask = torch.empty(n, dtype=torch.uint8, device="cuda")
        first = 777
        sp = bgc_params(engine, SEEDS["C3"], 1, NS, dist, grouped, n + 5000, a, b, d_cdf if dist else None, KINDS[R])
        engine.synth_pods_dev(sp, first, n, d_ns, d_req, d_mask)
        torch.cuda.synchronize()
        p = oracle.gen_params(SEEDS["C3"], 1, NS, dist, grouped, n + 5000, a, b, cdf if dist else None, KINDS[R])
        ns, req, mask = oracle.gen_fill(p, first, n, want_mask=True, threads=4)
        assert np.array_equal(d_ns.cpu().numpy().view(np.uint32), ns), (dist, grouped)
        assert np.array_equal(d_mask.cpu().numpy(), mask)
        for x, y in zip(d_req, req):
            assert np.array_equal(x.cpu().numpy(), y)


def bgc_params(engine, *a):
    import importlib
    return importlib.import_module("bacchus-gpu-controller_b200").synth_params(*a)


@pytest.mark.parametrize("dist", [0, 1])
def test_c3_scale_properties(engine, oracle, dist):
    """BASELINE.json configs[2] shape (R=3, Zipf, NS=10^4) at 2^26 pods generated on the device:
    full usage table == oracle streaming the same generator in chunks; verdict stream checked by
    slice checksums and dense slices; grouped and arrival-order inputs give the same usage."""
    import torch
    R, NS, n = 3, 10_000, 1 << 26
    a, b = Q.scramble_params(NS)
    cdf = Q.zipf_cdf(NS, 1.1)
    d_cdf = torch.from_numpy(cdf.view(np.int64)).cuda()
    engine.use_torch_stream()
    d_ns = torch.empty(n, dtype=torch.int32, device="cuda")
    d_req = [torch.empty(n, dtype=torch.int64, device="cuda") for _ in range(R)]
    d_used = torch.empty(R * NS, dtype=torch.int64, device="cuda")
    sp = bgc_params(engine, SEEDS["C3"], 0, NS, dist, 0, n, a, b, d_cdf if dist else None, KINDS[R])
    engine.synth_pods_dev(sp, 0, n, d_ns, d_req)
    engine.quota_table_set(np.zeros((R, NS), np.int64), np.zeros(NS, np.uint8), UNITS[R])
    # oracle: stream the generator in chunks, accumulate
    p = oracle.gen_params(SEEDS["C3"], 0, NS, dist, 0, n, a, b, cdf if dist else None, KINDS[R])
    want = np.zeros((R, NS), np.int64)
    step = 1 << 23
    for first in range(0, n, step):
        ns, req, _ = oracle.gen_fill(p, first, step, threads=8)
        want += oracle.usage_reduce(ns, req, NS, threads=8)
    results = {}
    for variant in [0] + VARIANTS:
        engine.set_k1_variant(variant)
        engine.usage_reduce_dev(d_ns, d_req, 0, d_used)
        torch.cuda.synchronize()
        results[variant] = d_used.cpu().numpy().reshape(R, NS).copy()
        assert np.array_equal(results[variant], want), (variant, engine.last_k1_variant())
    # K2 on the candidate set
    hard, hm = make_hard(want, SEEDS["C3"])
    engine.quota_table_set(hard, hm, UNITS[R])
    engine.usage_set(want)
    sp1 = bgc_params(engine, SEEDS["C3"], 1, NS, dist, 0, n, a, b, d_cdf if dist else None, KINDS[R])
    engine.synth_pods_dev(sp1, 0, n, d_ns, d_req)
    d_v = torch.empty(n, dtype=torch.uint8, device="cuda")
    engine.quota_eval_dev(d_ns, d_req, None, d_v)
    whole = engine.checksum_u8_dev(d_v)
    p1 = oracle.gen_params(SEEDS["C3"], 1, NS, dist, 0, n, a, b, cdf if dist else None, KINDS[R])
    acc = 0
    for first in range(0, n, step):
        ns, req, _ = oracle.gen_fill(p1, first, step, threads=8)
        v = oracle.quota_eval(ns, req, None, NS, want, hard, hm, 6, threads=8)
        acc = (acc + oracle.checksum_u8(v, first)) & (2**64 - 1)
        if first in (0, step * 3):
            assert np.array_equal(d_v[first:first + step].cpu().numpy(), v)
    assert acc == whole
    hist = engine.histogram_u8_dev(d_v)
    assert int(hist.sum()) == n and hist[0] > 0 and hist[1:8].sum() > 0


@pytest.mark.parametrize("R", [1, 2, 3, 4])
@pytest.mark.parametrize("variant,NS", [(4, 1), (4, 9), (4, 1000), (4, 70_000), (5, 17), (5, 70_000), (5, 150_000)])
def test_k1_cluster_distributed_table(bgc, engine, oracle, R, variant, NS):
    """K1 with the table spread over a thread-block cluster's shared memory (DSMEM atomics)."""
    rng = np.random.default_rng(R * 31 + NS + variant)
    for n in (1, 4099, 300_007):
        ns = rng.integers(0, NS, n, dtype=np.uint32)
        req = [rng.integers(-2**63, 2**63 - 1, n, dtype=np.int64) for _ in range(R)]
        if n > 50:
            ns[3] = NS
            ns[n // 2] = 0xFFFFFFFF
            req[0][::3] = rng.integers(0, 9, req[0][::3].shape[0])
        want = oracle.usage_reduce(ns, req, NS, threads=4)
        engine.quota_table_set(np.zeros((R, NS), np.int64), np.zeros(NS, np.uint8), UNITS[R])
        engine.set_k1_variant(variant)
        got = engine.usage_reduce(ns, req)
        assert np.array_equal(got, want), (engine.last_k1_variant(), n)
    assert "cluster" in engine.last_k1_variant()


def test_k1_cluster_rejects_tables_that_do_not_fit(bgc, engine):
    R, NS = 3, 2_000_000
    engine.quota_table_set(np.zeros((R, NS), np.int64), np.zeros(NS, np.uint8), UNITS[R])
    engine.set_k1_variant(bgc.K1_CLUSTER16)
    with pytest.raises(bgc.BgcError) as e:
        engine.usage_reduce(np.zeros(8, np.uint32), [np.zeros(8, np.int64)] * R)
    assert e.value.code == bgc.E_BADARG


def test_two_contexts_in_two_threads_do_not_interfere(bgc, oracle):
    """One bgc_ctx per thread (the reference handler is stateless per request, admission.rs:145-147):
    two engines driven concurrently from two host threads must each produce their own exact result."""
    import threading
    R, NS, n = 3, 4000, 1_500_000
    rng = np.random.default_rng(99)
    cases = []
    for t in range(2):
        ns = rng.integers(0, NS, n, dtype=np.uint32)
        req = [rng.integers(0, 1 << 34, n, dtype=np.int64) for _ in range(R)]
        used = oracle.usage_reduce(ns, req, NS, threads=4)
        hard, hm = make_hard(used, 40 + t)
        cases.append((ns, req, used, hard, hm, oracle.quota_eval(ns, req, None, NS, used, hard, hm, 6, threads=4)))
    errors = []

    def work(t):
        try:
            ns, req, used, hard, hm, want = cases[t]
            eng = bgc.Engine(0)
            eng.quota_table_set(hard, hm, UNITS[R])
            for _ in range(5):
                assert np.array_equal(eng.usage_reduce(ns, req), used)
                assert np.array_equal(eng.quota_eval(ns, req), want)
                assert np.array_equal(eng.quota_eval(ns[:1000], [c[:1000] for c in req]), want[:1000])   # latency path
            eng.close()
        except Exception as e:  # noqa: BLE001
            errors.append((t, repr(e)))

    threads = [threading.Thread(target=work, args=(t,)) for t in range(2)]
    for th in threads:
        th.start()
    for th in threads:
        th.join()
    assert not errors, errors


def test_host_entry_point_infers_the_grouping_hint(engine, oracle):
    R, NS, n = 3, 2000, 400_000
    rng = np.random.default_rng(12)
    req = [rng.integers(0, 1 << 30, n, dtype=np.int64) for _ in range(R)]
    grouped = np.repeat(rng.permutation(NS).astype(np.uint32), n // NS)
    shuffled = rng.permutation(grouped)
    engine.quota_table_set(np.zeros((R, NS), np.int64), np.zeros(NS, np.uint8), UNITS[R])
    for ns, expect in ((grouped, "segmented_global_red"), (shuffled, "scatter_smem_direct")):
        assert np.array_equal(engine.usage_reduce(ns, req), oracle.usage_reduce(ns, req, NS, threads=4))
        assert engine.last_k1_variant() == expect


@pytest.mark.parametrize("R", [1, 2])
@pytest.mark.parametrize("with_mask", [False, True])
@pytest.mark.parametrize("NS,misalign", [(13_000, 0), (5, 0), (9_999, 1)])
def test_k2_shared_memory_table_variant(bgc, engine, oracle, R, with_mask, NS, misalign):
    """Large batches with one or two resources take k2_quota_eval_smem (table copied into shared memory):
    device pointers, 1.6 M pods, aligned and misaligned columns, unknown namespaces, extreme values."""
    import torch
    rng = np.random.default_rng(R * 100 + NS)
    n = 1_600_003
    engine.use_torch_stream()
    ns = rng.integers(0, NS, n, dtype=np.uint32)
    req = [rng.integers(0, 1 << 36, n, dtype=np.int64) for _ in range(R)]
    used, hard, hm = _table(engine, oracle, (ns, req), NS, R)
    hm[::5] = rng.integers(0, 1 << R, hm[::5].shape[0], dtype=np.uint8)
    hard[:, ::7] = np.iinfo(np.int64).max
    hard[:, 1::11] = np.iinfo(np.int64).min
    engine.quota_table_set(hard, hm, UNITS[R])
    engine.usage_set(used)
    cns = rng.integers(0, NS, n, dtype=np.uint32)
    creq = [rng.integers(-5, 1 << 34, n, dtype=np.int64) for _ in range(R)]
    creq[0][::1001] = np.iinfo(np.int64).max
    cns[::53] = NS + 1
    cns[1::97] = 0xFFFFFFFF
    mask = rng.integers(0, 256, n, dtype=np.uint8) if with_mask else None
    want = oracle.quota_eval(cns, creq, mask, NS, used, hard, hm, _compute_mask(UNITS[R]))

    def dev(a):
        t = torch.from_numpy(a.view(np.int32) if a.dtype == np.uint32 else a)
        buf = torch.empty(n + 8, dtype=t.dtype, device="cuda")
        buf[misalign:misalign + n] = t.cuda()
        return buf[misalign:misalign + n]
    d_v = torch.empty(n + 8, dtype=torch.uint8, device="cuda")[misalign:misalign + n]
    engine.quota_eval_dev(dev(cns), [dev(c) for c in creq], dev(mask) if with_mask else None, d_v)
    torch.cuda.synchronize()
    assert np.array_equal(d_v.cpu().numpy(), want)


def test_pageable_host_batches_take_the_bounce_pipeline(engine, oracle):
    """Host batches in ordinary (pageable) numpy memory larger than one 4 Mi-pod piece go through the pinned
    bounce ring (helper thread copying piece j+1 while piece j is in the pipeline): K0, K1 (pieces after the
    first accumulate) and K2 (with and without masks) must still equal the oracle, ragged tail included."""
    rng = np.random.default_rng(99)
    n, NS, R = (4 << 20) * 2 + 54_321, 9_000, 3
    ns = rng.integers(0, NS + 2, n, dtype=np.uint32)
    req = [rng.integers(0, 1 << 33, n, dtype=np.int64) for _ in range(R)]
    mask = rng.integers(0, 256, n, dtype=np.uint8)
    engine.quota_table_set(np.full((R, NS), 1 << 50, np.int64), np.full(NS, 7, np.uint8), UNITS[R])
    want_used = oracle.usage_reduce(ns, req, NS, threads=8)
    assert np.array_equal(engine.usage_reduce(ns, req), want_used)
    hard = want_used + rng.integers(-(1 << 33), 1 << 33, want_used.shape)
    hm = np.full(NS, 7, np.uint8)
    engine.quota_table_set(hard, hm, UNITS[R])
    engine.usage_set(want_used)
    for mk in (None, mask):
        assert np.array_equal(engine.quota_eval(ns, req, mk), oracle.quota_eval(ns, req, mk, NS, want_used, hard, hm, _compute_mask(UNITS[R]), threads=8))
    op, fl = oracle.gen_policy_fill(5, 0, n, threads=8)
    code, pm = engine.policy_eval(op, fl)
    rc, rp = oracle.policy_bits(op, fl, threads=8)
    assert np.array_equal(code, rc) and np.array_equal(pm, rp)
