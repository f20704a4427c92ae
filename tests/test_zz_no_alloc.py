"""GPU test for bgc_usage_reduce_static_dev, the entry point for callers that cannot allocate.

Some users of libbgc.so run where device allocation at call time is off the table: the host
process gave every byte of HBM to its own pool at start-up (a library's cudaMalloc would fail or
steal budgeted memory), or the call sits on a latency-critical path where cudaMalloc/cudaFree would
stall all streams with their device-wide synchronisation.  For them the library offers one entry
point that accumulates into a fixed 512 KiB static device table instead of allocated memory.

What is checked here, through the C ABI: the static-table result is bit-identical to the regular
bgc_usage_reduce_dev on the same device columns AND to the CPU oracle's literal loop, on two calls
in a row with different inputs (the second call must not see the first one's sums, since the table
persists between calls); the device's free memory is the same before and after a call (the point
of the entry point: it allocates nothing once its one-time set-up is done); and oversized requests
are refused with a message instead of being launched.  This file sorts last on
purpose, so a `-x` run has already executed every other test when it gets here.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

STATIC_MAX_NS = 16384


def _static_reduce(bgc, d_ns, d_req, R, n, NS, flags, d_out, stream=0):
    fn = bgc.lib.bgc_usage_reduce_static_dev
    fn.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_size_t, C.c_uint32, C.c_uint, C.c_void_p, C.c_void_p,
                   C.c_char_p, C.c_size_t]
    fn.restype = C.c_int
    ptrs = (C.c_void_p * max(1, len(d_req)))(*[t.data_ptr() for t in d_req])
    err = C.create_string_buffer(256)
    rc = fn(C.c_void_p(d_ns.data_ptr() if d_ns is not None else 0), ptrs, R, n, NS, flags,
            C.c_void_p(d_out.data_ptr() if d_out is not None else 0), C.c_void_p(stream), err, 256)
    return rc, err.value.decode()


@pytest.mark.parametrize("NS,R,grouped", [(10_000, 3, 0), (STATIC_MAX_NS, 4, 0), (300, 2, 0), (10_000, 3, 1), (STATIC_MAX_NS, 1, 0), (12_000, 3, 0)])
def test_static_table_matches_regular_entry_point_twice_in_a_row(bgc, engine, oracle, NS, R, grouped):
    import torch
    from workloads import KINDS, UNITS
    n = 700_003
    engine.use_torch_stream()
    engine.quota_table_set(np.zeros((R, NS), np.int64), np.zeros(NS, np.uint8), UNITS[R])
    a = max(1, int(NS * 0.618))
    while np.gcd(a, NS) != 1:
        a += 1
    for call, seed in enumerate((101, 202)):                       # two calls in a row, different data
        d_ns = torch.empty(n, dtype=torch.int32, device="cuda")
        d_req = [torch.empty(n, dtype=torch.int64, device="cuda") for _ in range(R)]
        engine.synth_pods_dev(bgc.synth_params(seed, 0, NS, 0, grouped, n, a, 3, None, KINDS[R]), 0, n, d_ns, d_req)
        if call == 1:                                              # negative and huge values on the second call
            d_req[0][::7] = -(1 << 40)
            d_req[-1][::11] = (1 << 62)
        want = torch.empty(R * NS, dtype=torch.int64, device="cuda")
        engine.usage_reduce_dev(d_ns, d_req, grouped, want)        # existing entry point
        got = torch.full((R * NS,), -1, dtype=torch.int64, device="cuda")
        if call == 1:
            # the first call did the one-time set-up; from here on a call must leave the device's free
            # memory exactly as it found it (no cudaMalloc, no lazily loaded module, nothing)
            torch.cuda.synchronize()
            free_before = torch.cuda.mem_get_info()[0]
        rc, msg = _static_reduce(bgc, d_ns, d_req, R, n, NS, grouped, got)
        assert rc == 0, msg
        torch.cuda.synchronize()
        if call == 1:
            assert torch.cuda.mem_get_info()[0] == free_before, "the allocation-free entry point changed the free device memory"
        assert torch.equal(got, want), f"call {call}: static-table result differs from bgc_usage_reduce_dev"
        ref = oracle.usage_reduce(d_ns.cpu().numpy().view(np.uint32), [c.cpu().numpy() for c in d_req], NS)
        assert np.array_equal(got.cpu().numpy().reshape(R, NS), ref), f"call {call}: static-table result differs from the oracle"
    # an empty batch after two full ones must read back all zeros (nothing left over in the table)
    got = torch.full((R * NS,), -1, dtype=torch.int64, device="cuda")
    rc, msg = _static_reduce(bgc, d_ns, d_req, R, 0, NS, 0, got)
    assert rc == 0, msg
    torch.cuda.synchronize()
    assert int(got.abs().sum().item()) == 0


def test_static_table_rejects_what_it_cannot_hold(bgc, engine):
    import torch
    d_ns = torch.zeros(8, dtype=torch.int32, device="cuda")
    d_req = [torch.zeros(8, dtype=torch.int64, device="cuda")]
    out = torch.zeros(STATIC_MAX_NS + 1, dtype=torch.int64, device="cuda")
    rc, msg = _static_reduce(bgc, d_ns, d_req, 1, 8, STATIC_MAX_NS + 1, 0, out)
    assert rc == bgc.E_BADARG and "BGC_STATIC_MAX_NS" in msg
    rc, msg = _static_reduce(bgc, d_ns, d_req, 5, 8, 10, 0, out)
    assert rc == bgc.E_BADARG and "R" in msg
    rc, msg = _static_reduce(bgc, d_ns, d_req, 1, 8, 10, 2, out)          # ACCUMULATE makes no sense on a shared table
    assert rc == bgc.E_BADARG
    rc, msg = _static_reduce(bgc, d_ns, d_req, 1, 8, 10, 0, None)
    assert rc == bgc.E_BADARG
