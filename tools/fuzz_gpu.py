#!/usr/bin/env python
"""Randomised soak test: random shapes, value ranges, K1 variants, hints, masks and pointer offsets,
every result compared bit for bit with the CPU oracle.  Usage: python tools/fuzz_gpu.py [--seconds S] [--seed N]"""
import argparse, importlib, os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle
from workloads import UNITS, make_hard
bgc = importlib.import_module("bacchus-gpu-controller_b200")

ap = argparse.ArgumentParser(); ap.add_argument("--seconds", type=float, default=60); ap.add_argument("--seed", type=int, default=1)
args = ap.parse_args()
eng = bgc.Engine(0); eng.use_torch_stream()
t_end = time.time() + args.seconds
it = 0
stats = {}
while time.time() < t_end:
    seed = args.seed * 1_000_003 + it
    rng = np.random.default_rng(seed)
    it += 1
    R = int(rng.integers(1, 9))
    NS = int(np.exp(rng.uniform(0, np.log(300_000))))
    n = int(np.exp(rng.uniform(0, np.log(400_000))))
    mode = rng.integers(0, 4)
    if mode == 0:
        ns = rng.integers(0, NS, n, dtype=np.uint32)
    elif mode == 1:      # skewed
        ns = (rng.zipf(1.3, n) % NS).astype(np.uint32)
    elif mode == 2:      # runs of random length
        lens = rng.geometric(1.0 / max(1, int(rng.integers(1, 600))), n)
        ns = np.repeat(rng.integers(0, NS, n, dtype=np.uint32), lens)[:n]
    else:                # few hot namespaces
        ns = rng.choice(rng.integers(0, NS, 3, dtype=np.uint32), n)
    if rng.random() < 0.3 and n > 4:
        ns[rng.integers(0, n, max(1, n // 50))] = rng.choice(np.array([NS, NS + 7, 0xFFFFFFFF], np.uint32))
    vmode = rng.integers(0, 4)
    def col():
        if vmode == 0: return rng.integers(0, 1 << 12, n, dtype=np.int64)
        if vmode == 1: return rng.integers(0, 1 << 40, n, dtype=np.int64)
        if vmode == 2: return rng.integers(-2**63, 2**63 - 1, n, dtype=np.int64)
        return rng.integers(-1000, 1 << 33, n, dtype=np.int64)
    req = [col() for _ in range(R)]
    for c in req:
        c[rng.random(n) < 0.3] = 0
    mask = rng.integers(0, 256, n, dtype=np.uint8) if rng.random() < 0.5 else None
    units = list(rng.integers(0, 3, R))
    cm = sum(1 << r for r, u in enumerate(units) if u)
    want = oracle.usage_reduce(ns, req, NS, threads=4)
    hard, hm = make_hard(want, seed)
    if rng.random() < 0.5:
        hm = rng.integers(0, 1 << R, NS, dtype=np.uint8)
    eng.quota_table_set(hard, hm, units)
    variant = int(rng.choice([0, 1, 2, 3, 2 | (1 << 8), 2 | (2 << 8), 2 | (3 << 8), 3 | (1 << 8), 3 | (3 << 8), 4, 5]))
    if R > 4 and (variant >> 8):
        variant &= 0xFF
    flags = int(rng.integers(0, 2))
    eng.set_k1_variant(variant)
    key = None
    try:
        off = int(rng.integers(0, 4)) if n > 8 else 0
        if rng.random() < 0.5:   # host path
            got = eng.usage_reduce(ns, req, flags=flags)
        else:                    # device path with an element offset (misaligned columns)
            d_ns = torch.from_numpy(np.concatenate([np.zeros(off, np.uint32), ns]).view(np.int32)).cuda()
            d_req = [torch.from_numpy(np.concatenate([np.zeros(off, np.int64), c])).cuda() for c in req]
            d_used = torch.empty(R * NS, dtype=torch.int64, device="cuda")
            eng.usage_reduce_dev(d_ns[off:], [c[off:] for c in d_req], flags, d_used)
            torch.cuda.synchronize()
            got = d_used.cpu().numpy().reshape(R, NS)
        key = eng.last_k1_variant()
    except bgc.BgcError as e:
        if "does not fit" in str(e):
            eng.set_k1_variant(0)
            got = eng.usage_reduce(ns, req, flags=flags)
            key = eng.last_k1_variant()
        else:
            raise
    assert np.array_equal(got, want), f"K1 mismatch seed={seed} R={R} NS={NS} n={n} variant={variant} flags={flags} {key}"
    stats[key] = stats.get(key, 0) + 1
    w16 = oracle.quota_eval16(ns, req, mask, NS, want, hard, hm, cm, threads=4)
    assert np.array_equal(eng.quota_eval(ns, req, mask, wide=True), w16), f"K2/16 mismatch seed={seed}"
    if R <= 4:
        w8 = oracle.quota_eval(ns, req, mask, NS, want, hard, hm, cm, threads=4)
        assert np.array_equal(eng.quota_eval(ns, req, mask), w8), f"K2/8 mismatch seed={seed}"
        if it % 3 == 0:
            h2 = hard + np.abs(hard) // 4 + 100
            eng.quota_table_set(h2, hm, units); eng.usage_set(want)
            gv, gu = oracle.admit_greedy(ns, req, mask, NS, want, h2, hm, cm)
            assert np.array_equal(eng.admit_greedy(ns, req, mask), gv), f"K3 verdict mismatch seed={seed} R={R} NS={NS} n={n}"
            assert np.array_equal(eng.usage_get(), gu), f"K3 usage mismatch seed={seed}"
            stats["k3"] = stats.get("k3", 0) + 1
    elif it % 3 == 0:      # R = 5..8: the u16 form of the greedy walkers
        h2 = hard + np.abs(hard) // 4 + 100
        eng.quota_table_set(h2, hm, units); eng.usage_set(want)
        gv, gu = oracle.admit_greedy16(ns, req, mask, NS, want, h2, hm, cm)
        assert np.array_equal(eng.admit_greedy(ns, req, mask, wide=True), gv), f"K3/16 verdict mismatch seed={seed} R={R} NS={NS} n={n}"
        assert np.array_equal(eng.usage_get(), gu), f"K3/16 usage mismatch seed={seed}"
        stats["k3_u16"] = stats.get("k3_u16", 0) + 1
    if R <= 4 and it % 2 == 0:   # narrow columns: random element types and shifts, u16 ids when they fit
        kinds = [str(rng.choice(["u8", "u16", "u32", "i32", "i64"])) for _ in range(R)]
        npt = {"u8": np.uint8, "u16": np.uint16, "u32": np.uint32, "i32": np.int32, "i64": np.int64}
        shifts = [int(rng.integers(0, 64)) if rng.random() < 0.3 else int(rng.integers(0, 24)) for _ in range(R)]
        tcols = [c.astype(npt[k]) for c, k in zip(req, kinds)]          # truncating casts: any bit pattern is a valid input
        tns = ns.astype(np.uint16) if NS <= 65536 and rng.random() < 0.7 else ns
        eng.quota_table_set(hard, hm, units)
        eng.set_k1_variant(int(rng.choice([0, 3])))
        tu = oracle.usage_reduce_typed(tns, tcols, shifts, NS, threads=4)
        assert np.array_equal(eng.usage_reduce_typed(tns, tcols, shifts), tu), f"typed K1 mismatch seed={seed} kinds={kinds} shifts={shifts}"
        tv = oracle.quota_eval_typed(tns, tcols, shifts, mask, NS, tu, hard, hm, cm, threads=4)
        assert np.array_equal(eng.quota_eval_typed(tns, tcols, shifts, mask), tv), f"typed K2 mismatch seed={seed} kinds={kinds} shifts={shifts}"
        stats["typed"] = stats.get("typed", 0) + 1
    if it % 7 == 0:        # one shard of a global stream = the oracle's stream filtered on the host
        from oracle import quota_ref as Q
        world = int(rng.choice([1, 2, 3, 8])); g = int(rng.integers(0, world))
        Rg = min(R, 4); kinds_g = [int(x) for x in rng.integers(0, 4, Rg)]
        dist = int(rng.integers(0, 2)); m = int(rng.integers(1, 150_000))
        a, b = Q.scramble_params(NS)
        cdf = Q.zipf_cdf(NS, 1.1) if dist else None
        d_cdf = torch.from_numpy(cdf.view(np.int64)).cuda() if dist else None
        first = int(rng.integers(0, 1 << 40))
        src = oracle.gen_fill(oracle.gen_params(seed, 1, NS, dist, 0, first + m, a, b, cdf, kinds_g), first, m, want_mask=True)
        owned, local_of = bgc.shard.local_ids(NS, world, g)
        sp = bgc.synth_params(seed, 1, NS, dist, 0, first + m, a, b, d_cdf, kinds_g)
        kept = eng.synth_pods_shard_dev(sp, first, m, world, g)
        keep = np.nonzero(bgc.shard.shard_of(src[0], world) == g)[0]
        assert kept == keep.shape[0], f"shard count mismatch seed={seed}"
        if kept:
            d_ns = torch.empty(kept, dtype=torch.int32, device="cuda"); d_rq = [torch.empty(kept, dtype=torch.int64, device="cuda") for _ in range(Rg)]
            d_mk = torch.empty(kept, dtype=torch.uint8, device="cuda"); d_sx = torch.empty(kept, dtype=torch.int64, device="cuda")
            eng.synth_pods_shard_dev(sp, first, m, world, g, torch.from_numpy(local_of.view(np.int32)).cuda(), 0, d_ns, d_rq, d_mk, d_sx)
            torch.cuda.synchronize()
            assert np.array_equal(d_ns.cpu().numpy().view(np.uint32), local_of[src[0][keep]]) and np.array_equal(d_sx.cpu().numpy(), keep + first)
            assert all(np.array_equal(d_rq[r].cpu().numpy(), src[1][r][keep]) for r in range(Rg)) and np.array_equal(d_mk.cpu().numpy(), src[2][keep])
        stats["shard_gen"] = stats.get("shard_gen", 0) + 1
    if it % 11 == 0:       # long segments: the chained-tile walker (hundreds of 1024-pod tiles per namespace)
        Rl = int(rng.integers(1, 5)); NSl = int(rng.integers(1, 6)); nl = int(rng.integers(40_000, 1_500_000))
        nsl = rng.integers(0, NSl, nl, dtype=np.uint32) if rng.random() < 0.7 else np.sort(rng.integers(0, NSl, nl, dtype=np.uint32))
        sizes = rng.choice(np.array([0, 0, 1, 2, 3, 7, 50, 250, 1000, 4000, 1 << 20, 1 << 33, 1 << 40], np.int64), int(rng.integers(2, 9)))
        reql = [rng.choice(sizes, nl) for _ in range(Rl)]
        if rng.random() < 0.3:      # a few requests outside the monotone range
            for c in reql:
                c[rng.integers(0, nl, int(rng.integers(1, 40)))] = rng.choice(np.array([-1, -5000, (1 << 40) + 1, -(1 << 50)], np.int64))
        maskl = rng.integers(0, 256, nl, dtype=np.uint8) | np.uint8(rng.integers(0, 16)) if rng.random() < 0.6 else None
        tot = oracle.usage_reduce(nsl, reql, NSl, threads=4)
        frac = float(rng.choice([0.0, 0.01, 0.3, 0.9, 1.0, 3.0]))
        hardl = (np.abs(tot).astype(np.float64) * frac).astype(np.int64) + int(rng.integers(0, 3))
        hml = rng.integers(0, 1 << Rl, NSl, dtype=np.uint8) if rng.random() < 0.5 else np.full(NSl, (1 << Rl) - 1, np.uint8)
        used0 = (hardl * float(rng.choice([0.0, 0.5, 0.999, 1.2]))).astype(np.int64)
        unitsl = [int(x) for x in rng.integers(0, 3, Rl)]
        eng.quota_table_set(hardl, hml, unitsl); eng.usage_set(used0)
        cml = sum(1 << r for r, u in enumerate(unitsl) if u)
        gv, gu = oracle.admit_greedy(nsl, reql, maskl, NSl, used0, hardl, hml, cml)
        assert np.array_equal(eng.admit_greedy(nsl, reql, maskl), gv), f"K3 long-segment verdict mismatch seed={seed} R={Rl} NS={NSl} n={nl}"
        assert np.array_equal(eng.usage_get(), gu), f"K3 long-segment usage mismatch seed={seed}"
        stats["k3_long"] = stats.get("k3_long", 0) + 1
    if it % 5 == 0:
        m = int(rng.integers(1, 200_000))
        op = rng.integers(0, 256, m, dtype=np.uint8); fl = rng.integers(0, 65536, m, dtype=np.uint16)
        c, p = eng.policy_eval(op, fl)
        rc, rp = oracle.policy_bits(np.minimum(op, 3).astype(np.uint8), (fl & 1023).astype(np.uint16), 4)
        assert np.array_equal(c, rc) and np.array_equal(p, rp), f"K0 mismatch seed={seed}"
        stats["k0"] = stats.get("k0", 0) + 1
print(f"fuzz ok: {it} random cases in {args.seconds:.0f} s, seed base {args.seed}; coverage {stats}")
