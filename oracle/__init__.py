"""oracle/ — CPU oracle package.  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / `--impl reference` legs may
import this package.  The product (bacchus-gpu-controller_b200/) never does, and fails loudly when
its CUDA library is missing instead of falling back to anything here.

PARITY UNPINNED (see policy_ref.h / quota_ref.h): the reference (Rust) cannot be built in this
image and ships no tests or golden vectors, so there is no oracle/_ref; the oracle follows the
reference source text (policy) and this repo's own spec (quota).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "liboracle.so")
_lib = None


def build(force: bool = False) -> str:
    srcs = [os.path.join(_HERE, f) for f in ("policy_ref.c", "quota_ref.c", "policy_ref.h", "quota_ref.h", "Makefile")]
    stale = (not os.path.exists(_SO)) or any(os.path.getmtime(s) > os.path.getmtime(_SO) for s in srcs)
    if force or stale:
        subprocess.run(["make", "-C", _HERE] + (["-B"] if force else []), check=True, capture_output=True)
    return _SO


def use_native_build() -> str:
    """Rebuild the C oracle with -march=native ON THIS MACHINE into _build/liboracle_native.so and load that
    one from now on.  For the timed CPU baseline only (bench.py): the default build targets x86-64-v3 so that
    the same .so runs on whatever CPU the GPU box has."""
    global _lib, _SO
    out = os.path.join(_HERE, "_build", "liboracle_native.so")
    os.makedirs(os.path.dirname(out), exist_ok=True)
    cc = os.environ.get("ORACLE_CC", "/usr/bin/gcc")
    subprocess.run([cc, "-O3", "-march=native", "-fopenmp", "-fPIC", "-std=c11", "-shared", "-o", out,
                    os.path.join(_HERE, "policy_ref.c"), os.path.join(_HERE, "quota_ref.c")], check=True, capture_output=True)
    _SO, _lib = out, None
    return out


class PolicyRequest(C.Structure):
    _fields_ = [("username", C.c_char_p), ("groups", C.POINTER(C.c_char_p)), ("n_groups", C.c_size_t),
                ("operation", C.c_int), ("has_object", C.c_int), ("name", C.c_char_p),
                ("parse_ok", C.c_int), ("spec_kube_username", C.c_char_p), ("has_quota", C.c_int),
                ("has_rolebinding", C.c_int)]


class PolicyConfig(C.Structure):
    _fields_ = [("oidc_username_prefix", C.c_char_p), ("default_role_name", C.c_char_p),
                ("authorized_group_names", C.POINTER(C.c_char_p)), ("n_authorized", C.c_size_t)]


class PolicyDecision(C.Structure):
    _fields_ = [("klass", C.c_uint8), ("reason", C.c_uint8), ("patch_mask", C.c_uint8),
                ("message", C.c_char_p), ("kube_username_patch", C.c_char * 256),
                ("subject_name", C.c_char * 256)]


class GenParams(C.Structure):
    _fields_ = [("seed", C.c_uint64), ("set", C.c_uint32), ("ns_count", C.c_uint32), ("dist", C.c_uint32),
                ("grouped", C.c_uint32), ("n_total", C.c_uint64), ("scramble_a", C.c_uint64),
                ("scramble_b", C.c_uint64), ("cdf", C.c_void_p), ("R", C.c_uint32), ("kind", C.c_uint32 * 8)]


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_SO)
        L.policy_ref_message.restype = C.c_char_p
        L.gen_ref_hash.restype = C.c_uint64
        L.gen_ref_hash.argtypes = [C.c_uint64, C.c_uint32, C.c_uint64]
        L.gen_ref_checksum_u8.restype = C.c_uint64
        L.gen_ref_checksum_u8.argtypes = [C.c_void_p, C.c_uint64, C.c_size_t]
        L.policy_ref_split_groups.restype = C.c_size_t
        _lib = L
    return _lib


def _p(a):
    return C.c_void_p(a.ctypes.data) if a is not None else C.c_void_p(0)


def _ptrs(cols):
    arr = (C.c_void_p * len(cols))(*[c.ctypes.data for c in cols])
    return arr


def _strs(items):
    arr = (C.c_char_p * max(1, len(items)))(*[s.encode() for s in items])
    return arr


# ------------------------------------------------------------------ policy
def policy_bits(op: np.ndarray, flags: np.ndarray, threads: int = 1):
    n = op.shape[0]
    code = np.empty(n, np.uint8)
    pm = np.empty(n, np.uint8)
    lib().policy_ref_bits_batch(_p(op), _p(flags), C.c_size_t(n), _p(code), _p(pm), C.c_int(threads))
    return code, pm


def _mk_req(username, groups, operation, has_object, name, parse_ok, spec_kube_username, has_quota, has_rolebinding):
    keep = []
    r = PolicyRequest()
    r.username = None if username is None else username.encode()
    if groups is None:
        r.groups, r.n_groups = None, 0
    else:
        g = _strs(groups)
        keep.append(g)
        r.groups, r.n_groups = C.cast(g, C.POINTER(C.c_char_p)), len(groups)
    r.operation, r.has_object = operation, int(has_object)
    r.name = None if name is None else name.encode()
    r.parse_ok = int(parse_ok)
    r.spec_kube_username = None if spec_kube_username is None else spec_kube_username.encode()
    r.has_quota, r.has_rolebinding = int(has_quota), int(has_rolebinding)
    return r, keep


def _mk_cfg(prefix, role, authorized):
    c = PolicyConfig()
    a = _strs(authorized)
    c.oidc_username_prefix, c.default_role_name = prefix.encode(), role.encode()
    c.authorized_group_names, c.n_authorized = C.cast(a, C.POINTER(C.c_char_p)), len(authorized)
    return c, [a]


def policy_mutate_strings(cfg: tuple, **req):
    """policy_ref_mutate on real strings.  cfg = (prefix, default_role, [authorized groups])."""
    c, k1 = _mk_cfg(*cfg)
    r, k2 = _mk_req(**req)
    d = PolicyDecision()
    lib().policy_ref_mutate(C.byref(r), C.byref(c), C.byref(d))
    return dict(klass=d.klass, reason=d.reason, patch_mask=d.patch_mask,
                message=None if d.message is None else d.message.decode(),
                kube_username_patch=d.kube_username_patch.decode(), subject_name=d.subject_name.decode())


def policy_flatten_strings(cfg: tuple, **req):
    c, k1 = _mk_cfg(*cfg)
    r, k2 = _mk_req(**req)
    op, fl = C.c_uint8(), C.c_uint16()
    lib().policy_ref_flatten(C.byref(r), C.byref(c), C.byref(op), C.byref(fl))
    return op.value, fl.value


def split_groups(s: str) -> list:
    scratch = C.create_string_buffer(len(s.encode()) + 1)
    out = (C.c_char_p * 64)()
    n = lib().policy_ref_split_groups(s.encode(), scratch, out, C.c_size_t(64))
    return [out[i].decode() for i in range(min(n, 64))]


# ------------------------------------------------------------------ quota
def usage_reduce(ns_id, req, NS, threads: int = 1):
    R = len(req)
    used = np.empty((R, NS), np.int64)
    lib().quota_ref_usage_reduce(_p(ns_id), _ptrs(req), C.c_size_t(ns_id.shape[0]), C.c_uint32(NS),
                                 C.c_uint32(R), _p(used), C.c_int(threads))
    return used


def quota_eval(ns_id, req, req_mask, NS, used, hard, hard_mask, compute_mask, threads: int = 1):
    R = len(req)
    n = ns_id.shape[0]
    verdict = np.empty(n, np.uint8)
    lib().quota_ref_quota_eval(_p(ns_id), _ptrs(req), _p(req_mask), C.c_size_t(n), C.c_uint32(NS),
                               C.c_uint32(R), _p(np.ascontiguousarray(used)), _p(np.ascontiguousarray(hard)),
                               _p(hard_mask), C.c_uint8(compute_mask), _p(verdict), C.c_int(threads))
    return verdict


def admit_greedy(ns_id, req, req_mask, NS, used, hard, hard_mask, compute_mask):
    """Returns (verdict, used_after); `used` is not modified."""
    R = len(req)
    n = ns_id.shape[0]
    verdict = np.empty(n, np.uint8)
    u = np.ascontiguousarray(used, dtype=np.int64).copy()
    lib().quota_ref_admit_greedy(_p(ns_id), _ptrs(req), _p(req_mask), C.c_size_t(n), C.c_uint32(NS), C.c_uint32(R), _p(u),
                                 _p(np.ascontiguousarray(hard)), _p(hard_mask), C.c_uint8(compute_mask), _p(verdict))
    return verdict, u


def admit_greedy16(ns_id, req, req_mask, NS, used, hard, hard_mask, compute_mask):
    """u16 verdicts (missing << 8 | exceeded), any R <= 8.  Returns (verdict, used_after)."""
    R = len(req)
    n = ns_id.shape[0]
    verdict = np.empty(n, np.uint16)
    u = np.ascontiguousarray(used, dtype=np.int64).copy()
    lib().quota_ref_admit_greedy16(_p(ns_id), _ptrs(req), _p(req_mask), C.c_size_t(n), C.c_uint32(NS), C.c_uint32(R), _p(u),
                                   _p(np.ascontiguousarray(hard)), _p(hard_mask), C.c_uint8(compute_mask), _p(verdict))
    return verdict, u


def quota_eval16(ns_id, req, req_mask, NS, used, hard, hard_mask, compute_mask, threads: int = 1):
    R = len(req)
    n = ns_id.shape[0]
    verdict = np.empty(n, np.uint16)
    lib().quota_ref_quota_eval16(_p(ns_id), _ptrs(req), _p(req_mask), C.c_size_t(n), C.c_uint32(NS), C.c_uint32(R),
                                 _p(np.ascontiguousarray(used)), _p(np.ascontiguousarray(hard)), _p(hard_mask),
                                 C.c_uint8(compute_mask), _p(verdict), C.c_int(threads))
    return verdict


T_I64, T_U32, T_U16, T_U8, T_I32 = 0, 1, 2, 3, 4
_NP_OF_T = {T_I64: np.int64, T_U32: np.uint32, T_U16: np.uint16, T_U8: np.uint8, T_I32: np.int32}
_T_OF_NP = {np.dtype(v): k for k, v in _NP_OF_T.items()}


def _typed_args(ns_id, cols, shifts):
    ns_type = _T_OF_NP[ns_id.dtype]
    assert ns_type in (T_U32, T_U16)
    types = np.array([_T_OF_NP[c.dtype] for c in cols], np.uint8)
    sh = np.array(shifts if shifts is not None else [0] * len(cols), np.uint8)
    return ns_type, types, sh


def usage_workspace(NS, R, threads: int = 1):
    """Private per-thread tables for usage_reduce_typed / usage_reduce_ws, allocated once by the caller."""
    return np.zeros(max(1, threads) * R * NS, np.uint64)


def usage_reduce_typed(ns_id, cols, shifts, NS, threads: int = 1, workspace=None):
    """Narrow columns: ns_id u16/u32, cols any of u8/u16/u32/i32/i64, value = element << shift."""
    R = len(cols)
    ns_type, types, sh = _typed_args(ns_id, cols, shifts)
    used = np.empty((R, NS), np.int64)
    lib().quota_ref_usage_reduce_typed(_p(ns_id), C.c_uint32(ns_type), _ptrs(cols), _p(types), _p(sh), C.c_size_t(ns_id.shape[0]),
                                       C.c_uint32(NS), C.c_uint32(R), _p(used), C.c_int(threads), _p(workspace))
    return used


def quota_eval_typed(ns_id, cols, shifts, req_mask, NS, used, hard, hard_mask, compute_mask, threads: int = 1):
    R = len(cols)
    ns_type, types, sh = _typed_args(ns_id, cols, shifts)
    n = ns_id.shape[0]
    verdict = np.empty(n, np.uint8)
    lib().quota_ref_quota_eval_typed(_p(ns_id), C.c_uint32(ns_type), _ptrs(cols), _p(types), _p(sh), _p(req_mask), C.c_size_t(n),
                                     C.c_uint32(NS), C.c_uint32(R), _p(np.ascontiguousarray(used)), _p(np.ascontiguousarray(hard)),
                                     _p(hard_mask), C.c_uint8(compute_mask), _p(verdict), C.c_int(threads))
    return verdict


def usage_reduce_ws(ns_id, req, NS, threads, workspace):
    """usage_reduce with a caller-owned workspace (no allocation inside the timed call)."""
    R = len(req)
    used = np.empty((R, NS), np.int64)
    lib().quota_ref_usage_reduce_ws(_p(ns_id), _ptrs(req), C.c_size_t(ns_id.shape[0]), C.c_uint32(NS), C.c_uint32(R), _p(used),
                                    C.c_int(threads), _p(workspace))
    return used


def gen_params(seed, set_, NS, dist, grouped, n_total, a, b, cdf, kinds):
    p = GenParams()
    p.seed, p.set, p.ns_count, p.dist, p.grouped, p.n_total = seed, set_, NS, dist, int(grouped), n_total
    p.scramble_a, p.scramble_b = a, b
    p.cdf = cdf.ctypes.data if cdf is not None else None
    p.R = len(kinds)
    for i, k in enumerate(kinds):
        p.kind[i] = k
    return p


def gen_fill(params: GenParams, first: int, count: int, want_mask: bool = False, threads: int = 1):
    ns = np.empty(count, np.uint32)
    req = [np.empty(count, np.int64) for _ in range(params.R)]
    mask = np.empty(count, np.uint8) if want_mask else None
    lib().gen_ref_fill(C.byref(params), C.c_uint64(first), C.c_size_t(count), _p(ns), _ptrs(req), _p(mask),
                       C.c_int(threads))
    return ns, req, mask


def gen_policy_fill(seed: int, first: int, count: int, threads: int = 1):
    op = np.empty(count, np.uint8)
    fl = np.empty(count, np.uint16)
    lib().gen_ref_policy_fill(C.c_uint64(seed), C.c_uint64(first), C.c_size_t(count), _p(op), _p(fl), C.c_int(threads))
    return op, fl


def checksum_u8(v: np.ndarray, first: int = 0) -> int:
    return lib().gen_ref_checksum_u8(_p(v), C.c_uint64(first), C.c_size_t(v.shape[0]))
