"""CPU tests for the N>1 path (row E): namespace-hash sharding with no data-path collective.
world_size-2 gloo processes stand in for two GPUs' ranks; the per-rank evaluation is done by the
oracle here (the device path is covered by the -m gpu tests and bench.py --gpus N)."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_partition_is_disjoint_and_complete(bgc):
    S = bgc.shard
    rng = np.random.default_rng(0)
    ns = rng.integers(0, 1000, 20_000, dtype=np.uint32)
    cols = [rng.integers(0, 100, 20_000, dtype=np.int64)]
    for world in (1, 2, 4, 8):
        parts = S.partition(ns, cols, world)
        idx = np.concatenate([p[0] for p in parts])
        assert np.array_equal(np.sort(idx), np.arange(ns.shape[0]))
        owners = [set(np.unique(p[1]).tolist()) for p in parts]
        for a in range(world):
            for b in range(a + 1, world):
                assert not owners[a] & owners[b]        # a namespace lives on exactly one rank
        own = [set(S.owned_namespaces(1000, world, g).tolist()) for g in range(world)]
        assert all(owners[g] <= own[g] for g in range(world))
        sizes = np.array([len(p[0]) for p in parts])
        assert sizes.max() / sizes.mean() < 1.25        # dense ids are spread evenly


def test_sharded_eval_equals_unsharded(bgc, oracle):
    from workloads import make_hard
    S = bgc.shard
    rng = np.random.default_rng(1)
    n, NS, R = 30_000, 500, 3
    ns = rng.integers(0, NS, n, dtype=np.uint32)
    req = [rng.integers(0, 1 << 30, n, dtype=np.int64) for _ in range(R)]
    cns = rng.integers(0, NS, n, dtype=np.uint32)
    creq = [rng.integers(0, 1 << 30, n, dtype=np.int64) for _ in range(R)]
    used = oracle.usage_reduce(ns, req, NS)
    hard, hm = make_hard(used, 9)
    want = oracle.quota_eval(cns, creq, None, NS, used, hard, hm, 6)
    for world in (2, 4):
        run = S.partition(ns, req, world)
        cand = S.partition(cns, creq, world)
        parts = []
        for g in range(world):
            u = oracle.usage_reduce(run[g][1], run[g][2], NS)       # rank-local usage: only owned namespaces non-zero
            v = oracle.quota_eval(cand[g][1], cand[g][2], None, NS, u, hard, hm, 6)
            parts.append((cand[g][0], v))
        assert np.array_equal(S.gather_verdicts(n, parts), want)


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import importlib
    import torch
    import torch.distributed as dist
    import oracle
    from workloads import make_hard
    bgc = importlib.import_module("bacchus-gpu-controller_b200")
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.default_rng(5)                      # every rank regenerates the same batch
    n, NS, R = 8_000, 200, 2
    ns = rng.integers(0, NS, n, dtype=np.uint32)
    req = [rng.integers(0, 1 << 20, n, dtype=np.int64) for _ in range(R)]
    used_all = oracle.usage_reduce(ns, req, NS)
    hard, hm = make_hard(used_all, 3)
    mine = bgc.shard.shard_of(ns, world) == rank
    u = oracle.usage_reduce(ns[mine], [c[mine] for c in req], NS)
    v = oracle.quota_eval(ns[mine], [c[mine] for c in req], None, NS, u, hard, hm, 2)
    # the only cross-rank step is the timing/result gather bench.py performs: max over ranks and a count
    cnt = torch.tensor([int(mine.sum()), int((v != 0).sum())], dtype=torch.int64)
    dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
    t = torch.tensor([float(rank + 1)])
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    want = oracle.quota_eval(ns, req, None, NS, used_all, hard, hm, 2)
    ok = bool(np.array_equal(v, want[mine])) and int(cnt[0]) == n and int(cnt[1]) == int((want != 0).sum()) and t.item() == world
    dist.barrier()
    dist.destroy_process_group()
    q.put((rank, ok))


def _worker_index(rank, world, port, q):
    """Row N5: pods sharded by index, one int64 all-reduce of the usage table, then local evaluation."""
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    import importlib
    import torch
    import torch.distributed as dist
    import oracle
    from workloads import make_hard
    bgc = importlib.import_module("bacchus-gpu-controller_b200")
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.default_rng(6)
    n, NS, R = 9_001, 150, 3
    ns = rng.integers(0, NS, n, dtype=np.uint32)
    req = [rng.integers(0, 1 << 33, n, dtype=np.int64) for _ in range(R)]
    cns = rng.integers(0, NS, n, dtype=np.uint32)
    creq = [rng.integers(0, 1 << 33, n, dtype=np.int64) for _ in range(R)]
    used_all = oracle.usage_reduce(ns, req, NS)
    hard, hm = make_hard(used_all, 8)
    lo, hi = bgc.shard.index_slice(n, world, rank)
    local = torch.from_numpy(oracle.usage_reduce(ns[lo:hi], [c[lo:hi] for c in req], NS))
    bgc.shard.allreduce_usage(local)                               # the one exchange step
    ok = bool(np.array_equal(local.numpy(), used_all))
    v = oracle.quota_eval(cns[lo:hi], [c[lo:hi] for c in creq], None, NS, local.numpy(), hard, hm, 6)
    want = oracle.quota_eval(cns, creq, None, NS, used_all, hard, hm, 6)
    ok = ok and bool(np.array_equal(v, want[lo:hi]))
    dist.barrier()
    dist.destroy_process_group()
    q.put((rank, ok))


def test_two_rank_gloo_index_sharding_with_one_allreduce():
    import torch.multiprocessing as mp
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker_index, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=180) for _ in procs)
    for p in procs:
        p.join(60)
    assert res == [(0, True), (1, True)]
    import importlib
    S = importlib.import_module("bacchus-gpu-controller_b200").shard
    assert [S.index_slice(10, 3, r) for r in range(3)] == [(0, 4), (4, 8), (8, 10)]
    assert S.index_slice(2, 4, 3) == (2, 2)


def test_two_rank_gloo_shards_need_no_collective_on_the_data_path():
    import torch.multiprocessing as mp
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=180) for _ in procs)
    for p in procs:
        p.join(60)
    assert res == [(0, True), (1, True)]


def test_local_ids_are_dense_and_cover_every_namespace_once(bgc):
    """Row E: every rank's dense per-shard ids (shard.local_ids) — what bench.py's one-job C5 run and
    bgc_synth_pods_shard_dev use — partition the namespace space: each global id is owned by exactly one rank and
    maps to a distinct local id below that rank's table size."""
    S = bgc.shard
    for NS, world in ((1000, 1), (1000, 2), (12_345, 8)):
        seen = np.zeros(NS, np.int64)
        for g in range(world):
            owned, local_of = S.local_ids(NS, world, g)
            assert np.array_equal(owned, np.sort(owned)) and np.array_equal(S.shard_of(owned, world), np.full(owned.shape, g, np.uint32))
            assert np.array_equal(local_of[owned], np.arange(owned.shape[0], dtype=np.uint32))
            others = np.setdiff1d(np.arange(NS), owned)
            assert np.all(local_of[others] == 0xFFFFFFFF)
            seen[owned] += 1
        assert np.all(seen == 1)
