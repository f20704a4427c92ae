"""GPU parity tests of the narrow-column entry points (bgc_usage_reduce_typed / bgc_quota_eval_typed and
their _dev forms): for every element type, shifts, u16/u32 namespace ids, aligned and misaligned columns,
ragged sizes and multi-chunk host batches, the result equals the CPU oracle's typed form AND the oracle's
plain literal loop on the widened int64 columns, and the int64 GPU entry points on the same values.
Seam: the pod columns behind src/admission.rs:184-204; key set src/synchronizer.rs:249-281."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

NPT = {"u8": np.uint8, "u16": np.uint16, "u32": np.uint32, "i32": np.int32, "i64": np.int64}


def _draw(rng, kind, n):
    if kind == "u8":
        return rng.integers(0, 256, n).astype(np.uint8)
    if kind == "u16":
        return rng.integers(0, 65536, n).astype(np.uint16)
    if kind == "u32":
        return rng.integers(0, 1 << 32, n, dtype=np.uint64).astype(np.uint32)
    if kind == "i32":
        return rng.integers(-(1 << 31), 1 << 31, n).astype(np.int32)
    return rng.integers(-(1 << 62), 1 << 62, n).astype(np.int64)


def _widen(cols, shifts):
    out = []
    for c, s in zip(cols, shifts):
        with np.errstate(over="ignore"):
            out.append((c.astype(np.int64).view(np.uint64) << np.uint64(s)).view(np.int64))
    return out


CASES = [
    (["u8", "u16", "u32"], [0, 0, 20], np.uint16, 9_000, 200_003),      # the C3 key set in 9 bytes per pod
    (["u8", "u32", "i64"], [0, 0, 0], np.uint32, 10_000, 131_072),
    (["i32", "i64", "u8", "u16"], [3, 1, 40, 7], np.uint32, 30_000, 77_777),   # hash sink, odd shifts, negatives
    (["u16"], [63], np.uint16, 300, 5),
    (["i64", "i64"], [0, 0], np.uint16, 65_536, 50_001),                # every u16 id is a valid namespace
]


@pytest.mark.parametrize("kinds,shifts,ns_t,NS,n", CASES)
def test_typed_host_entry_points_vs_oracle(engine, oracle, kinds, shifts, ns_t, NS, n):
    from workloads import make_hard
    rng = np.random.default_rng(n)
    R = len(kinds)
    hi = min(NS + 3, np.iinfo(ns_t).max + 1)
    ns = rng.integers(0, hi, n).astype(ns_t)
    cols = [_draw(rng, k, n) for k in kinds]
    cns = rng.integers(0, hi, n).astype(ns_t)
    ccols = [_draw(rng, k, n) for k in kinds]
    mask = rng.integers(0, 256, n).astype(np.uint8)
    units = [[0, 1, 2, 0][r % 4] for r in range(R)]
    engine.quota_table_set(np.zeros((R, NS), np.int64), np.zeros(NS, np.uint8), units)
    wide, cwide = _widen(cols, shifts), _widen(ccols, shifts)
    want_used = oracle.usage_reduce(ns.astype(np.uint32), wide, NS)
    assert np.array_equal(oracle.usage_reduce_typed(ns, cols, shifts, NS, threads=3), want_used)
    got = engine.usage_reduce_typed(ns, cols, shifts)
    assert np.array_equal(got, want_used), engine.last_k1_variant()
    hard, hm = make_hard(np.abs(want_used) // 2, 5)
    engine.quota_table_set(hard, hm, units)
    engine.usage_set(want_used)
    cm = sum(1 << r for r in range(R) if units[r] != 0)
    want_v = oracle.quota_eval(cns.astype(np.uint32), cwide, mask, NS, want_used, hard, hm, cm)
    assert np.array_equal(oracle.quota_eval_typed(cns, ccols, shifts, mask, NS, want_used, hard, hm, cm, threads=2), want_v)
    assert np.array_equal(engine.quota_eval_typed(cns, ccols, shifts, mask), want_v)
    assert np.array_equal(engine.quota_eval_typed(cns, ccols, shifts, None),
                          oracle.quota_eval(cns.astype(np.uint32), cwide, None, NS, want_used, hard, hm, cm))
    # and the int64 entry points on the widened values say the same
    assert np.array_equal(engine.quota_eval(cns.astype(np.uint32), cwide, mask), want_v)


def test_typed_multichunk_host_batch_and_accumulate(engine, oracle):
    """More than one 4 Mi-pod staging chunk through the pipelined host path; ACCUMULATE adds a second batch."""
    rng = np.random.default_rng(7)
    n, NS = (4 << 20) * 2 + 12_345, 9_999
    ns = rng.integers(0, NS, n).astype(np.uint16)
    cols = [rng.integers(0, 9, n).astype(np.uint8), rng.integers(0, 16001, n).astype(np.uint16), rng.integers(0, 70_000, n).astype(np.uint32)]
    shifts = [0, 0, 20]
    engine.quota_table_set(np.full((3, NS), 1 << 40, np.int64), np.full(NS, 7, np.uint8), [0, 1, 2])
    want = oracle.usage_reduce_typed(ns, cols, shifts, NS, threads=8)
    assert np.array_equal(engine.usage_reduce_typed(ns, cols, shifts), want)
    got2 = engine.usage_reduce_typed(ns[:1000], [c[:1000] for c in cols], shifts, flags=2)
    with np.errstate(over="ignore"):
        assert np.array_equal(got2, want + oracle.usage_reduce_typed(ns[:1000], [c[:1000] for c in cols], shifts, NS))
    v = engine.quota_eval_typed(ns, cols, shifts)
    assert np.array_equal(v, oracle.quota_eval_typed(ns, cols, shifts, None, NS, got2, np.full((3, NS), 1 << 40, np.int64),
                                                     np.full(NS, 7, np.uint8), 6, threads=8))


@pytest.mark.parametrize("misalign", [0, 1, 3])
def test_typed_device_pointers_aligned_and_misaligned(bgc, engine, oracle, misalign):
    import torch
    rng = np.random.default_rng(11 + misalign)
    n, NS = 300_001, 12_000
    engine.use_torch_stream()
    ns = rng.integers(0, NS + 2, n).astype(np.uint16)
    cols = [rng.integers(0, 9, n).astype(np.uint8), rng.integers(0, 16001, n).astype(np.uint16), rng.integers(0, 1 << 31, n).astype(np.int32)]
    shifts = [0, 0, 5]

    def dev(a):  # a device copy that starts `misalign` elements into an allocation
        t = torch.from_numpy(a.view(np.int16) if a.dtype == np.uint16 else a)
        buf = torch.empty(n + 8, dtype=t.dtype, device="cuda")
        buf[misalign:misalign + n] = t.cuda()
        return buf[misalign:misalign + n]
    d_ns, d_cols = dev(ns), [dev(c) for c in cols]
    types = [bgc.T_U8, bgc.T_U16, bgc.T_I32]
    engine.quota_table_set(np.full((3, NS), 1 << 33, np.int64), np.full(NS, 7, np.uint8), [0, 1, 2])
    d_used = torch.empty(3 * NS, dtype=torch.int64, device="cuda")
    engine.usage_reduce_typed_dev(d_ns, d_cols, shifts, 0, d_used, types=types)
    torch.cuda.synchronize()
    want = oracle.usage_reduce_typed(ns, cols, shifts, NS)
    assert np.array_equal(d_used.cpu().numpy().reshape(3, NS), want)
    d_v = torch.empty(n + 8, dtype=torch.uint8, device="cuda")[misalign:misalign + n]
    engine.quota_eval_typed_dev(d_ns, d_cols, shifts, None, d_v, types=types)
    torch.cuda.synchronize()
    assert np.array_equal(d_v.cpu().numpy(), oracle.quota_eval_typed(ns, cols, shifts, None, NS, want, np.full((3, NS), 1 << 33, np.int64),
                                                                      np.full(NS, 7, np.uint8), 6))


def test_typed_argument_errors(bgc, engine):
    engine.quota_table_set(np.zeros((2, 10), np.int64), np.zeros(10, np.uint8), [0, 1])
    ns = np.zeros(4, np.uint32)
    cols = [np.zeros(4, np.uint8), np.zeros(4, np.uint16)]
    with pytest.raises(bgc.BgcError):
        engine.usage_reduce_typed(ns, cols, [0, 64])                       # shift out of range
    arr = bgc.make_cols([(cols[0].ctypes.data, 9, 0), (cols[1].ctypes.data, bgc.T_U16, 0)])
    with pytest.raises(bgc.BgcError):
        engine.usage_reduce_typed_ptr(ns.ctypes.data, bgc.T_U32, arr, 4)   # unknown element type
    arr = bgc.make_cols([(cols[0].ctypes.data, bgc.T_U8, 0), (cols[1].ctypes.data, bgc.T_U16, 0)])
    with pytest.raises(bgc.BgcError):
        engine.usage_reduce_typed_ptr(ns.ctypes.data, bgc.T_U8, arr, 4)    # ns ids are u16 or u32 only
    assert engine.usage_reduce_typed(ns[:0], [c[:0] for c in cols]).sum() == 0   # empty batch
