"""CPU tests: the C++ host layer of libbgc (rows A2, A5, and the response renderer) against the
oracle's restatements.  No GPU calls."""
import json

import numpy as np
import pytest

from oracle import policy_ref as P


def test_flatten_requests_matches_oracle_on_every_constructible_request(bgc, oracle):
    A = bgc.admission
    cfg_py = P.Config()
    cfg = A.AdmissionConfig()
    reviews, expect = [], []
    for op in range(4):
        for fl in range(1024):
            req = P.request_from_bits(op, fl, cfg_py)
            if req is not None:
                reviews.append(req)
                expect.append((op, fl))
    reqs = [A.AdmissionRequest.from_review(r) for r in reviews]
    op, flags = A.flatten_requests(cfg, reqs)
    assert [(int(o), int(f)) for o, f in zip(op, flags)] == expect


@pytest.mark.parametrize("prefix,groups", [("oidc:", ["gpu", "admin"]), ("", ["gpu"]), ("oidc:", [""]), ("sso/", ["a", " b"])])
def test_flatten_requests_string_edge_cases(bgc, oracle, prefix, groups):
    A = bgc.admission
    cfg = A.AdmissionConfig(prefix, "edit", groups)
    names = [None, "", "alice", "Alice", "oidc:alice", "sso/al"]
    users = [None, "", prefix, prefix + "alice", "alice", "admin", "sso/al"]
    glists = [None, [], ["gpu"], [""], [" b"], ["x", "admin"]]
    reqs, expect = [], []
    for u in users:
        for nm in names:
            for g in glists:
                for op in (0, 1, 2, 3, 7):
                    for ku in (None, "", "bob"):
                        q = dict(username=u, groups=g, operation=op, has_object=True, name=nm, parse_ok=True,
                                 spec_kube_username=ku, has_quota=bool(op & 1), has_rolebinding=bool(op & 2))
                        reqs.append(A.AdmissionRequest(**q))
                        expect.append(oracle.policy_flatten_strings((prefix, "edit", groups), **q))
    op, flags = A.flatten_requests(cfg, reqs)
    assert [(int(o), int(f)) for o, f in zip(op, flags)] == expect


def test_split_groups_matches_reference_semantics(bgc, oracle):
    for s in ["", "gpu,admin", "gpu, admin,", ",", "a,,b", "한글,gpu"]:
        assert bgc.admission.split_groups(s) == s.split(",") == oracle.split_groups(s)


def test_reason_messages_are_the_reference_literals(bgc):
    for reason, (_, msg, _line) in P.REASONS.items():
        got = bgc.lib.bgc_reason_message(reason)
        if reason != 8:
            assert (None if got is None else got.decode()) == msg


def test_render_patch_equals_python_oracle_patches(bgc):
    A = bgc.admission
    cfg_py, cfg = P.Config(), A.AdmissionConfig()
    seen = 0
    for op in range(2):
        for fl in range(1024):
            rv = P.request_from_bits(op, fl, cfg_py)
            if rv is None:
                continue
            d = P.mutate(rv, cfg_py)
            txt = A.render_patch(cfg, A.AdmissionRequest.from_review(rv), d.patch_mask)
            if d.patch_mask:
                assert json.loads(txt) == d.patches
                assert list(json.loads(txt)[-1]["value"].keys()) == ["role_ref", "subjects"] or not (d.patch_mask & 2)
                seen += 1
            else:
                assert txt is None
    assert seen > 10
    # JSON escaping of odd user names
    q = A.AdmissionRequest('oidc:a"b\\c\n', ["gpu"], 0, True, 'a"b\\c\n')
    assert json.loads(A.render_patch(cfg, q, 7))[0]["value"] == 'a"b\\c\n'


QUANT = [
    ("0", 0, 0), ("1", 0, 1), ("500m", 1, 500), ("2", 1, 2000), ("0.5", 1, 500), ("1.5", 0, 2), ("100m", 0, 1),
    ("1Ki", 2, 1024), ("4Gi", 2, 4 << 30), ("1.5Gi", 2, 1610612736), ("1e3", 0, 1000), ("1E3", 0, 1000),
    ("1e-3", 1, 1), ("129M", 2, 129_000_000), ("1k", 0, 1000), ("5n", 1, 1), ("250u", 1, 1), ("12345678901", 0, 12345678901),
    ("8Ei", 2, None), ("+3", 0, 3), ("-1", 0, -1), ("2Ti", 2, 2 << 40), ("0.1Ki", 2, 103), ("16", 1, 16000),
]


@pytest.mark.parametrize("s,unit,want", QUANT)
def test_parse_quantity(bgc, s, unit, want):
    if want is None:
        with pytest.raises(ValueError):
            bgc.pods.parse_quantity(s, unit)
    else:
        assert bgc.pods.parse_quantity(s, unit) == want


@pytest.mark.parametrize("bad", ["", "abc", "1x", "1Kii", "1e", "--1", "1.2.3", "Gi", "1 Gi"])
def test_parse_quantity_rejects(bgc, bad):
    with pytest.raises(ValueError):
        bgc.pods.parse_quantity(bad, 0)


def _pod(ns, containers, init=None, overhead=None, phase="Running"):
    def c(req, lim=None):
        return {"resources": {"requests": req, **({"limits": lim} if lim else {})}}
    return {"metadata": {"namespace": ns}, "status": {"phase": phase},
            "spec": {"containers": [c(*x) if isinstance(x, tuple) else c(x) for x in containers],
                     **({"initContainers": [c(x) for x in init]} if init else {}),
                     **({"overhead": overhead} if overhead else {})}}


def test_flatten_pods_usage_rules(bgc):
    Pd = bgc.pods
    keys = ["requests.nvidia.com/gpu", "requests.cpu", "limits.memory"]
    m = Pd.NamespaceMap()
    pods = [
        _pod("alice", [({"cpu": "500m", "nvidia.com/gpu": "1"}, {"memory": "1Gi"}), ({"cpu": "1"}, {"memory": "512Mi"})]),
        _pod("bob", [{"cpu": "100m"}], init=[{"cpu": "2"}, {"cpu": "1"}], overhead={"cpu": "250m"}),
        _pod("alice", [{"cpu": "8"}], phase="Succeeded"),          # terminal → skipped
        _pod("carol", [{"memory": "1Gi"}]),                        # states no cpu → mask bit clear
        _pod("alice", [{"cpu": "1"}], phase="Failed"),
    ]
    ns, req, mask, unit = Pd.flatten_pods(pods, keys, m)
    assert ns.tolist() == [0, 1, 2] and len(m) == 3 and m.name(1) == "bob" and m.lookup("nobody") == 0xFFFFFFFF
    assert req[0].tolist() == [1, 0, 0]
    assert req[1].tolist() == [1500, 2250, 0]                      # max(Σ app, max init) + overhead
    assert req[2].tolist() == [(1 << 30) + (512 << 20), 0, 0]
    assert mask.tolist() == [0b110, 0b010, 0b000]                  # gpu only on one container; bob has no limits
    assert unit.tolist() == [0, 1, 2]
    # without interning, unknown namespaces become BAD ids
    ns2, *_ = Pd.flatten_pods([_pod("zed", [{"cpu": "1"}])], keys, m, intern=False)
    assert ns2.tolist() == [0xFFFFFFFF]
    with pytest.raises(bgc.BgcError):
        Pd.flatten_pods([_pod("x", [{"cpu": "lots"}])], keys, m)


def test_hard_from_sheet_row_follows_synchronizer(bgc):
    # synchronizer.rs:249-281: cpu → "<n>" cores, memory/storage → "<n>Gi", gpu/mig → "<n>"
    keys = ["requests.cpu", "requests.memory", "limits.cpu", "limits.memory"]
    hard, unit = bgc.pods.hard_from_sheet_row(16, 64, 2, 500, 1, keys)
    assert hard.tolist() == [16000, 64 << 30, 16000, 64 << 30] and unit.tolist() == [1, 2, 1, 2]
    hard, unit = bgc.pods.hard_from_sheet_row(16, 64, 2, 500, 1, ["requests.nvidia.com/gpu", "requests.nvidia.com/mig-1g.10gb", "requests.storage"])
    assert hard.tolist() == [2, 1, 500 << 30] and unit.tolist() == [0, 0, 0]
    # the same numbers through the quantity parser, as Kubernetes would read the strings the reference writes
    assert bgc.pods.parse_quantity("16", 1) == 16000 and bgc.pods.parse_quantity("64Gi", 2) == 64 << 30
    with pytest.raises(bgc.BgcError):
        bgc.pods.hard_from_sheet_row(1, 1, 1, 1, 1, ["requests.foo"])


@pytest.mark.parametrize("value,unit,text", [
    (0, 1, "0"), (500, 1, "500m"), (2000, 1, "2"), (16000, 1, "16"), (-250, 1, "-250m"), (1500, 1, "1500m"),
    (0, 2, "0"), (64 << 30, 2, "64Gi"), (1536 << 20, 2, "1536Mi"), (1000, 2, "1000"), (1 << 60, 2, "1Ei"), (3 << 10, 2, "3Ki"),
    (7, 0, "7"), (-1, 0, "-1"), (1 << 40, 0, "1099511627776"),
])
def test_format_quantity_round_trips(bgc, value, unit, text):
    assert bgc.pods.format_quantity(value, unit) == text
    assert bgc.pods.parse_quantity(text, unit) == value


def test_quota_denial_messages(bgc):
    Pd = bgc.pods
    keys = ["requests.nvidia.com/gpu", "requests.cpu", "requests.memory"]
    req, used, hard = [1, 1500, 2 << 30], [2, 7000, 3 << 30], [2, 8000, 4 << 30]
    assert Pd.quota_denial_message(0, "alice", keys, req, used, hard) is None
    assert Pd.quota_denial_message(0b011, "alice", keys, req, used, hard) == (
        "exceeded quota: alice, requested: requests.nvidia.com/gpu=1,requests.cpu=1500m, "
        "used: requests.nvidia.com/gpu=2,requests.cpu=7, limited: requests.nvidia.com/gpu=2,requests.cpu=8")
    assert Pd.quota_denial_message(0b100, "alice", keys, req, used, hard) == (
        "exceeded quota: alice, requested: requests.memory=2Gi, used: requests.memory=3Gi, limited: requests.memory=4Gi")
    # missing wins over exceeded; u8 layout has missing in bits 4..7, u16 in bits 8..15
    assert Pd.quota_denial_message((0b110 << 4) | 0b001, "bob", keys, req, used, hard) == "failed quota: bob: must specify requests.cpu,requests.memory"
    assert Pd.quota_denial_message((0b010 << 8), "bob", keys, req, used, hard, wide=True) == "failed quota: bob: must specify requests.cpu"
    with pytest.raises(bgc.BgcError):
        Pd.quota_denial_message(0xFF, "bob", keys, req, used, hard)          # unknown namespace has no quota to name


def test_parse_quantity_property_vs_exact_fractions(bgc):
    """Random quantity strings in every suffix family against exact rational arithmetic (ceil)."""
    from fractions import Fraction
    from hypothesis import given, settings, strategies as st
    suffixes = {"": 1, "n": Fraction(1, 10**9), "u": Fraction(1, 10**6), "m": Fraction(1, 1000), "k": 10**3, "M": 10**6, "G": 10**9,
                "T": 10**12, "Ki": 1 << 10, "Mi": 1 << 20, "Gi": 1 << 30, "Ti": 1 << 40}

    @settings(max_examples=400, deadline=None)
    @given(st.integers(0, 10**9), st.integers(0, 999999), st.integers(0, 6), st.sampled_from(sorted(suffixes)), st.integers(0, 2), st.booleans())
    def check(whole, frac, frac_digits, suffix, unit, negative):
        text = str(whole)
        value = Fraction(whole)
        if frac_digits:
            f = frac % (10 ** frac_digits)
            text += "." + str(f).zfill(frac_digits)
            value += Fraction(f, 10 ** frac_digits)
        value *= suffixes[suffix]
        if negative:
            text, value = "-" + text, -value
        text += suffix
        scaled = value * (1000 if unit == 1 else 1)
        want = -((-scaled.numerator) // scaled.denominator)         # ceil, also for negatives (rounds toward zero)
        if abs(want) >= 2**63:
            return
        assert bgc.pods.parse_quantity(text, unit) == want, text

    check()
    # decimal exponents
    for text, unit, want in [("1e3", 0, 1000), ("12e-1", 0, 2), ("5e-1", 1, 500), ("1.5e3", 2, 1500), ("25e-4", 1, 3)]:
        assert bgc.pods.parse_quantity(text, unit) == want


def test_compact_columns_plan_and_round_trip(bgc, oracle):
    """bgc_compact_plan / bgc_compact_column / bgc_compact_ns_ids (the host step between the flattener and the
    narrow-column entry points): the narrowest exact type and the largest common shift per column, lossless round
    trip, and the oracle's typed forms decide on the compact columns what they decide on the int64 ones."""
    rng = np.random.default_rng(3)
    n, NS = 100_003, 5_000
    cols = [rng.choice(np.array([0, 1, 2, 4, 8], np.int64), n),                     # gpu counts      -> u8, shift 0
            250 << rng.integers(0, 7, n).astype(np.int64),                          # cpu millicores  -> u16 << 1 (250 = 125 * 2)
            np.int64(1) << rng.integers(28, 37, n).astype(np.int64),                # memory bytes    -> u16 << 28
            rng.integers(0, 1 << 31, n).astype(np.int64) * 3,                       #                 -> i64 or u32
            rng.integers(-1000, 1000, n).astype(np.int64) << 4,                     # negatives       -> i32 << 4
            np.zeros(n, np.int64),                                                  # all zero        -> u8, 0
            rng.integers(-(1 << 62), 1 << 62, n).astype(np.int64) | 1]              # full range, odd -> i64, 0
    ns = rng.integers(0, NS, n).astype(np.uint32)
    ns[::97] = 0xFFFFFFFF
    ns[1::101] = NS + 3
    ns2, ccols, shifts = bgc.pods.compact_columns(ns, cols, NS)
    want = [(np.uint8, 0), (np.uint16, 1), (np.uint16, 28), None, (np.int32, 4), (np.uint8, 0), (np.int64, 0)]
    for c, d, sh, w in zip(cols, ccols, shifts, want):
        assert np.array_equal(d.astype(np.int64) << sh, c)
        if w is not None:
            assert (d.dtype, sh) == (np.dtype(w[0]), w[1]), (d.dtype, sh, w)
    assert ccols[3].dtype in (np.dtype(np.uint32), np.dtype(np.int64))
    assert ns2.dtype == np.uint16 and np.array_equal(ns2 == 0xFFFF, ns >= NS) and np.array_equal(ns2[ns < NS], ns[ns < NS])
    assert bgc.pods.compact_columns(ns, cols[:1], 70_000)[0].dtype == np.uint32
    # a representation that does not hold the values is refused, not silently truncated
    import ctypes as C
    out = np.empty(n, np.uint8)
    assert bgc.lib.bgc_compact_column(C.c_void_p(cols[2].ctypes.data), n, 3, 20, C.c_void_p(out.ctypes.data)) == bgc.E_BADARG
    R = 4
    used_a = oracle.usage_reduce(np.where(ns < NS, ns, 0xFFFFFFFF).astype(np.uint32), cols[:R], NS)
    used_b = oracle.usage_reduce_typed(ns2, ccols[:R], shifts[:R], NS)
    assert np.array_equal(used_a, used_b)
