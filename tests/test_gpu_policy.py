"""GPU parity tests for K0 policy_eval (row A1), through the C ABI, against the committed truth
table and the oracle.  Bit-exact."""
import csv
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden", "policy_truth_table.csv")


def _golden():
    rows = list(csv.DictReader(open(GOLD)))
    return (np.array([int(r["op"]) for r in rows], np.uint8), np.array([int(r["flags"]) for r in rows], np.uint16),
            np.array([int(r["code"]) for r in rows], np.uint8), np.array([int(r["patch_mask"]) for r in rows], np.uint8))


def test_k0_exhaustive_truth_table(engine):
    op, fl, code, pm = _golden()
    c, m = engine.policy_eval(op, fl)
    assert np.array_equal(c, code) and np.array_equal(m, pm)


@pytest.mark.parametrize("n", [0, 1, 15, 16, 17, 255, 4097, 1_000_003])
def test_k0_ragged_sizes_vs_oracle(engine, oracle, n):
    op, fl = oracle.gen_policy_fill(0xB2000001, 0, n)
    c, m = engine.policy_eval(op, fl)
    rc, rm = oracle.policy_bits(op, fl, threads=4)
    assert np.array_equal(c, rc) and np.array_equal(m, rm)


def test_k0_unknown_operations_and_high_flag_bits(engine, oracle):
    rng = np.random.default_rng(0)
    n = 100_000
    op = rng.integers(0, 256, n, dtype=np.uint8)          # anything outside 0..2 is the `_` arm (admission.rs:305)
    fl = rng.integers(0, 65536, n, dtype=np.uint16)       # bits above 9 are ignored
    c, m = engine.policy_eval(op, fl)
    rc, rm = oracle.policy_bits(np.minimum(op, 3).astype(np.uint8), (fl & 1023).astype(np.uint16))
    assert np.array_equal(c, rc) and np.array_equal(m, rm)


def test_k0_device_pointers_aligned_and_misaligned(engine, oracle):
    import torch
    n = 300_001
    op, fl = oracle.gen_policy_fill(5, 0, n + 8)
    rc, rm = oracle.policy_bits(op, fl)
    d_op = torch.from_numpy(op).cuda()
    d_fl = torch.from_numpy(fl.view(np.int16)).cuda()
    engine.use_torch_stream()
    for off in (0, 1, 3, 8):
        d_c = torch.zeros(n + 8, dtype=torch.uint8, device="cuda")
        d_m = torch.zeros(n + 8, dtype=torch.uint8, device="cuda")
        engine.policy_eval_dev(d_op[off:off + n], d_fl[off:off + n], d_c[off:off + n], d_m[off:off + n])
        torch.cuda.synchronize()
        assert np.array_equal(d_c[off:off + n].cpu().numpy(), rc[off:off + n])
        assert np.array_equal(d_m[off:off + n].cpu().numpy(), rm[off:off + n])
        assert d_c[:off].sum().item() == 0 and d_c[off + n:].sum().item() == 0   # nothing written outside the slice


def test_k0_device_generator_matches_oracle_generator(engine, oracle):
    import torch
    n = 1_000_000
    d_op = torch.empty(n, dtype=torch.uint8, device="cuda")
    d_fl = torch.empty(n, dtype=torch.int16, device="cuda")
    engine.use_torch_stream()
    engine.synth_policy_dev(0xB2000001, 12345, n, d_op, d_fl)
    torch.cuda.synchronize()
    op, fl = oracle.gen_policy_fill(0xB2000001, 12345, n, threads=4)
    assert np.array_equal(d_op.cpu().numpy(), op) and np.array_equal(d_fl.cpu().numpy().view(np.uint16), fl)


def test_k0_full_size_checksum(engine, oracle):
    """2^28 requests generated on the device; code stream checked by slice checksums against the oracle
    regenerating the same slices (size-independent property: checksums of slices add up)."""
    import torch
    n = 1 << 28
    d_op = torch.empty(n, dtype=torch.uint8, device="cuda")
    d_fl = torch.empty(n, dtype=torch.int16, device="cuda")
    d_c = torch.empty(n, dtype=torch.uint8, device="cuda")
    d_m = torch.empty(n, dtype=torch.uint8, device="cuda")
    engine.use_torch_stream()
    engine.synth_policy_dev(3, 0, n, d_op, d_fl)
    engine.policy_eval_dev(d_op, d_fl, d_c, d_m)
    whole = engine.checksum_u8_dev(d_c)
    parts = 0
    step = 1 << 26
    for first in range(0, n, step):
        parts = (parts + engine.checksum_u8_dev(d_c[first:first + step], first)) & (2**64 - 1)
    assert parts == whole
    # oracle on three slices (start, middle, ragged end)
    for first, cnt in ((0, 1 << 20), (n // 2 + 7, 1 << 20), (n - 999_983, 999_983)):
        op, fl = oracle.gen_policy_fill(3, first, cnt, threads=4)
        rc, rm = oracle.policy_bits(op, fl, threads=4)
        assert oracle.checksum_u8(rc, first) == engine.checksum_u8_dev(d_c[first:first + cnt], first)
        assert np.array_equal(d_m[first:first + cnt].cpu().numpy(), rm)


def test_mutate_batch_end_to_end_strings(bgc, engine, oracle):
    """Host mirror of mutate(): strings in, responses out, decision made on the GPU (C1a)."""
    from oracle import policy_ref as P
    A = bgc.admission
    cfg_py, cfg = P.Config(), A.AdmissionConfig()
    reviews = [r for op in range(4) for fl in range(1024) if (r := P.request_from_bits(op, fl, cfg_py)) is not None]
    reviews = (reviews * 17)[:10_000]
    res = A.mutate_batch(engine, cfg, [A.AdmissionRequest.from_review(r) for r in reviews])
    import json
    for rv, got in zip(reviews, res):
        d = P.mutate(rv, cfg_py)
        assert (got.klass, got.reason, got.patch_mask) == (d.klass, d.reason, d.patch_mask)
        assert got.allowed == (d.klass in (P.ALLOW, P.ALLOW_PATCH))
        if d.reason not in (0, 8, 12, 13):
            assert got.message == d.message
        if d.patch_mask:
            assert json.loads(got.patch) == d.patches
