"""GPU tests of the multi-GPU path on ONE device (row E and row N5).

* the shard generator: rank g's shard of one global stream (mix32(ns) % G == g, stream order, dense
  per-shard ids) equals the oracle's stream filtered on the host, for every rank, and the shards'
  K1/K2 results, gathered on the host, equal the unsharded evaluation (namespace-hash sharding needs
  no exchange);
* the N5 sequence: K1 on each half of a batch -> bgc_usage_get_dev -> sum of the tables (what the
  int64 all-reduce computes) -> bgc_usage_set_dev -> K2 on each half, against the oracle on the
  whole batch.
Reference anchors: one namespace per UserBootstrap (src/controller.rs:55-63), its ResourceQuota
(src/controller.rs:89-110), key set src/synchronizer.rs:249-281.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _stream(oracle, bgc, seed, set_, NS, dist, n, R):
    from oracle import quota_ref as Q
    from workloads import KINDS
    a, b = Q.scramble_params(NS)
    cdf = Q.zipf_cdf(NS, 1.1) if dist else None
    return (a, b, cdf, oracle.gen_params(seed, set_, NS, dist, 0, n, a, b, cdf, KINDS[R]))


@pytest.mark.parametrize("world", [1, 2, 8])
@pytest.mark.parametrize("dist,NS,n", [(1, 5000, 300_001), (0, 70_000, 150_000)])
def test_shard_of_global_stream_matches_filtered_oracle_stream(bgc, engine, oracle, world, dist, NS, n):
    import torch
    from workloads import KINDS, UNITS, make_hard
    R, seed = 3, 0xC5
    S = bgc.shard
    engine.use_torch_stream()
    a, b, cdf, p0 = _stream(oracle, bgc, seed, 0, NS, dist, n, R)
    _, _, _, p1 = _stream(oracle, bgc, seed, 1, NS, dist, n, R)
    d_cdf = torch.from_numpy(cdf.view(np.int64)).cuda() if dist else None
    run = oracle.gen_fill(p0, 0, n)
    cand = oracle.gen_fill(p1, 0, n, want_mask=True)
    used_all = oracle.usage_reduce(run[0], run[1], NS)
    hard, hm = make_hard(used_all, seed)
    want = oracle.quota_eval(cand[0], cand[1], cand[2], NS, used_all, hard, hm, 0b110)
    gathered = np.full(n, 0xEE, np.uint8)
    sizes = []
    for g in range(world):
        owned, local_of = S.local_ids(NS, world, g)
        d_local = torch.from_numpy(local_of.view(np.int32)).cuda()
        sets = []
        for set_ in (0, 1):
            sp = bgc.synth_params(seed, set_, NS, dist, 0, n, a, b, d_cdf, KINDS[R])
            # count in two slices (the shard is appended to), then emit
            k0 = engine.synth_pods_shard_dev(sp, 0, n // 3, world, g)
            k1 = engine.synth_pods_shard_dev(sp, n // 3, n - n // 3, world, g)
            m = k0 + k1
            d_ns = torch.empty(max(m, 1), dtype=torch.int32, device="cuda")
            d_req = [torch.empty(max(m, 1), dtype=torch.int64, device="cuda") for _ in range(R)]
            d_mask = torch.empty(max(m, 1), dtype=torch.uint8, device="cuda")
            d_src = torch.empty(max(m, 1), dtype=torch.int64, device="cuda")
            assert engine.synth_pods_shard_dev(sp, 0, n // 3, world, g, d_local, 0, d_ns, d_req, d_mask, d_src) == k0
            assert engine.synth_pods_shard_dev(sp, n // 3, n - n // 3, world, g, d_local, k0, d_ns, d_req, d_mask, d_src) == k1
            torch.cuda.synchronize()
            sets.append((m, d_ns, d_req, d_mask, d_src))
            # the same shard on the host: the oracle's stream, filtered
            src = run if set_ == 0 else cand
            keep = np.nonzero(S.shard_of(src[0], world) == g)[0]
            assert m == keep.shape[0]
            assert np.array_equal(d_src[:m].cpu().numpy(), keep)
            assert np.array_equal(d_ns[:m].cpu().numpy().view(np.uint32), local_of[src[0][keep]])
            for r in range(R):
                assert np.array_equal(d_req[r][:m].cpu().numpy(), src[1][r][keep])
            if set_ == 1:
                assert np.array_equal(d_mask[:m].cpu().numpy(), src[2][keep])
        sizes.append(sets[0][0])
        # the shard's own tables: only its owned namespaces, dense ids
        NSg = owned.shape[0]
        engine.quota_table_set(np.ascontiguousarray(hard[:, owned]), np.ascontiguousarray(hm[owned]), UNITS[R])
        m0, r_ns, r_req, _, _ = sets[0]
        m1, c_ns, c_req, c_mask, c_src = sets[1]
        engine.usage_reduce_dev(r_ns, r_req, 0, None, n=m0)
        torch.cuda.synchronize()
        assert np.array_equal(engine.usage_get(), used_all[:, owned]), "shard usage != owned slice of the global usage"
        d_v = torch.empty(max(m1, 1), dtype=torch.uint8, device="cuda")
        engine.quota_eval_dev(c_ns, c_req, c_mask, d_v, n=m1)
        torch.cuda.synchronize()
        gathered[c_src[:m1].cpu().numpy()] = d_v[:m1].cpu().numpy()      # the host-side gather
        assert NSg == int((S.shard_of(np.arange(NS, dtype=np.uint32), world) == g).sum())
    assert sum(sizes) == n
    assert np.array_equal(gathered, want), "gathered shard verdicts differ from the unsharded oracle"


def test_shard_generator_refuses_columns_that_are_too_small(bgc, engine):
    import torch
    from workloads import KINDS
    engine.use_torch_stream()
    sp = bgc.synth_params(1, 0, 100, 0, 0, 10_000, 1, 0, None, KINDS[2])
    m = engine.synth_pods_shard_dev(sp, 0, 10_000, 2, 0)
    d_ns = torch.empty(m - 1, dtype=torch.int32, device="cuda")
    d_req = [torch.empty(m - 1, dtype=torch.int64, device="cuda") for _ in range(2)]
    with pytest.raises(bgc.BgcError):
        engine.synth_pods_shard_dev(sp, 0, 10_000, 2, 0, None, 0, d_ns, d_req)
    with pytest.raises(bgc.BgcError):
        engine.synth_pods_shard_dev(sp, 0, 10_000, 2, 2)


@pytest.mark.parametrize("NS,dist", [(10_000, 1), (200_000, 0)])
def test_index_sharded_sequence_on_one_device(bgc, engine, oracle, NS, dist):
    """Row N5 without a second GPU: two index halves, each reduced into a full table, the tables summed
    (the all-reduce), the sum installed with bgc_usage_set_dev, each half evaluated against it."""
    import torch
    from workloads import KINDS, UNITS, make_hard
    R, seed, n = 3, 0x5EED, 400_003
    engine.use_torch_stream()
    a, b, cdf, p0 = _stream(oracle, bgc, seed, 0, NS, dist, n, R)
    _, _, _, p1 = _stream(oracle, bgc, seed, 1, NS, dist, n, R)
    d_cdf = torch.from_numpy(cdf.view(np.int64)).cuda() if dist else None
    run = oracle.gen_fill(p0, 0, n)
    cand = oracle.gen_fill(p1, 0, n, want_mask=True)
    used_all = oracle.usage_reduce(run[0], run[1], NS)
    hard, hm = make_hard(used_all, seed)
    want = oracle.quota_eval(cand[0], cand[1], cand[2], NS, used_all, hard, hm, 0b110)
    engine.quota_table_set(hard, hm, UNITS[R])
    halves = [bgc.shard.index_slice(n, 2, g) for g in range(2)]
    tables, cols = [], []
    for lo, hi in halves:
        m = hi - lo
        d_ns = torch.empty(m, dtype=torch.int32, device="cuda")
        d_req = [torch.empty(m, dtype=torch.int64, device="cuda") for _ in range(R)]
        engine.synth_pods_dev(bgc.synth_params(seed, 0, NS, dist, 0, n, a, b, d_cdf, KINDS[R]), lo, m, d_ns, d_req)
        engine.usage_reduce_dev(d_ns, d_req, 0)
        t = torch.empty(R * NS, dtype=torch.int64, device="cuda")
        engine.usage_get_dev(t)
        tables.append(t)
        c_ns = torch.empty(m, dtype=torch.int32, device="cuda")
        c_req = [torch.empty(m, dtype=torch.int64, device="cuda") for _ in range(R)]
        c_mask = torch.empty(m, dtype=torch.uint8, device="cuda")
        engine.synth_pods_dev(bgc.synth_params(seed, 1, NS, dist, 0, n, a, b, d_cdf, KINDS[R]), lo, m, c_ns, c_req, c_mask)
        cols.append((c_ns, c_req, c_mask))
    total = tables[0] + tables[1]                        # what ncclAllReduce(sum, int64) leaves on every rank
    torch.cuda.synchronize()
    assert np.array_equal(total.cpu().numpy().reshape(R, NS), used_all)
    got = []
    for (c_ns, c_req, c_mask) in cols:
        engine.usage_set_dev(total)
        v = torch.empty(c_ns.numel(), dtype=torch.uint8, device="cuda")
        engine.quota_eval_dev(c_ns, c_req, c_mask, v)
        torch.cuda.synchronize()
        got.append(v.cpu().numpy())
    assert np.array_equal(np.concatenate(got), want)
