"""Row A5: pod specs → SoA resource columns, through libbgc's C++ flattener (bgc_flatten_pods).

The reference never reads pods; the column set comes from the quota keys it writes
(reference src/synchronizer.rs:249-281) and the rules are Kubernetes' pod-usage semantics.
"""
from __future__ import annotations

import ctypes as C
from typing import Sequence

import numpy as np

from . import lib, OK, BgcError, UNIT_COUNT, UNIT_MILLI, UNIT_BYTES

# pod-relevant subset of the 7 keys of src/synchronizer.rs:249-281, with canonical units
QUOTA_KEYS = {
    "requests.cpu": UNIT_MILLI, "limits.cpu": UNIT_MILLI,
    "requests.memory": UNIT_BYTES, "limits.memory": UNIT_BYTES,
    "requests.nvidia.com/gpu": UNIT_COUNT, "requests.nvidia.com/mig-1g.10gb": UNIT_COUNT,
}
PHASES = {"Pending": 0, "Running": 1, "Succeeded": 2, "Failed": 3, "Unknown": 4}


class _RL(C.Structure):
    _fields_ = [("names", C.POINTER(C.c_char_p)), ("quantities", C.POINTER(C.c_char_p)), ("n", C.c_size_t)]


class _CS(C.Structure):
    _fields_ = [("requests", _RL), ("limits", _RL)]


class _Pod(C.Structure):
    _fields_ = [("ns", C.c_char_p), ("phase", C.c_int), ("containers", C.POINTER(_CS)), ("n_containers", C.c_size_t),
                ("init_containers", C.POINTER(_CS)), ("n_init_containers", C.c_size_t), ("overhead", C.POINTER(_RL))]


lib.bgc_nsmap_create.restype = C.c_void_p
lib.bgc_nsmap_destroy.argtypes = [C.c_void_p]
lib.bgc_nsmap_intern.argtypes = [C.c_void_p, C.c_char_p]
lib.bgc_nsmap_intern.restype = C.c_uint32
lib.bgc_nsmap_lookup.argtypes = [C.c_void_p, C.c_char_p]
lib.bgc_nsmap_lookup.restype = C.c_uint32
lib.bgc_nsmap_size.argtypes = [C.c_void_p]
lib.bgc_nsmap_size.restype = C.c_uint32
lib.bgc_nsmap_name.argtypes = [C.c_void_p, C.c_uint32]
lib.bgc_nsmap_name.restype = C.c_char_p
lib.bgc_parse_quantity.argtypes = [C.c_char_p, C.c_int, C.POINTER(C.c_int64)]


def parse_quantity(s: str, unit: int) -> int:
    out = C.c_int64()
    rc = lib.bgc_parse_quantity(s.encode(), unit, C.byref(out))
    if rc != OK:
        raise ValueError(f"bad quantity {s!r}")
    return out.value


lib.bgc_format_quantity.restype = C.c_long
lib.bgc_format_quantity.argtypes = [C.c_int64, C.c_int, C.c_char_p, C.c_size_t]
lib.bgc_render_quota_denial.restype = C.c_long


def format_quantity(value: int, unit: int) -> str:
    buf = C.create_string_buffer(64)
    if lib.bgc_format_quantity(C.c_int64(value), unit, buf, 64) < 0:
        raise ValueError("bad unit")
    return buf.value.decode()


def quota_denial_message(verdict: int, quota_name: str, quota_keys: Sequence[str], req, used, hard, wide: bool = False):
    """Message for one rejected pod (None when the pod is allowed); req/used/hard are the R values of
    this pod and of its namespace."""
    R = len(quota_keys)
    keys = (C.c_char_p * R)(*[k.encode() for k in quota_keys])
    unit = np.array([QUOTA_KEYS[k] for k in quota_keys], np.uint8)
    a = [np.ascontiguousarray(x, dtype=np.int64) for x in (req, used, hard)]
    buf = C.create_string_buffer(2048)
    n = lib.bgc_render_quota_denial(C.c_uint(int(verdict)), C.c_int(int(wide)), quota_name.encode(), keys, C.c_void_p(unit.ctypes.data),
                                    C.c_uint32(R), C.c_void_p(a[0].ctypes.data), C.c_void_p(a[1].ctypes.data),
                                    C.c_void_p(a[2].ctypes.data), buf, C.c_size_t(2048))
    if n < 0:
        raise BgcError(int(n), "bgc_render_quota_denial")
    return buf.value.decode() if n else None


class NamespaceMap:
    def __init__(self):
        self._h = C.c_void_p(lib.bgc_nsmap_create())

    def intern(self, name: str) -> int:
        return lib.bgc_nsmap_intern(self._h, name.encode())

    def lookup(self, name: str) -> int:
        return lib.bgc_nsmap_lookup(self._h, name.encode())

    def name(self, i: int):
        s = lib.bgc_nsmap_name(self._h, i)
        return None if s is None else s.decode()

    def __len__(self):
        return lib.bgc_nsmap_size(self._h)

    def __del__(self):
        if getattr(self, "_h", None):
            lib.bgc_nsmap_destroy(self._h)
            self._h = None


def _rl(d: dict, keep: list) -> _RL:
    d = d or {}
    n = len(d)
    names = (C.c_char_p * max(1, n))(*[k.encode() for k in d])
    quants = (C.c_char_p * max(1, n))(*[str(v).encode() for v in d.values()])
    keep += [names, quants]
    return _RL(C.cast(names, C.POINTER(C.c_char_p)), C.cast(quants, C.POINTER(C.c_char_p)), n)


def _containers(lst, keep):
    lst = lst or []
    arr = (_CS * max(1, len(lst)))()
    for i, c in enumerate(lst):
        res = c.get("resources") or {}
        arr[i].requests = _rl(res.get("requests"), keep)
        arr[i].limits = _rl(res.get("limits"), keep)
    keep.append(arr)
    return C.cast(arr, C.POINTER(_CS)), len(lst)


def flatten_pods(pods: Sequence[dict], quota_keys: Sequence[str], nsmap: NamespaceMap, intern: bool = True):
    """`pods` are Kubernetes Pod JSON objects (metadata.namespace, spec.containers[].resources,
    spec.initContainers, spec.overhead, status.phase).  Returns (ns_id, req[R], req_mask)."""
    keep: list = []
    n = len(pods)
    arr = (_Pod * max(1, n))()
    for i, p in enumerate(pods):
        spec = p.get("spec") or {}
        arr[i].ns = ((p.get("metadata") or {}).get("namespace") or "default").encode()
        arr[i].phase = PHASES.get((p.get("status") or {}).get("phase", "Pending"), 4)
        arr[i].containers, arr[i].n_containers = _containers(spec.get("containers"), keep)
        arr[i].init_containers, arr[i].n_init_containers = _containers(spec.get("initContainers"), keep)
        if spec.get("overhead"):
            o = _rl(spec["overhead"], keep)
            keep.append(o)
            arr[i].overhead = C.pointer(o)
    R = len(quota_keys)
    keys = (C.c_char_p * R)(*[k.encode() for k in quota_keys])
    unit = np.array([QUOTA_KEYS[k] for k in quota_keys], np.uint8)
    ns_id = np.empty(n, np.uint32)
    req = [np.empty(n, np.int64) for _ in range(R)]
    mask = np.empty(n, np.uint8)
    ptrs = (C.c_void_p * R)(*[c.ctypes.data for c in req])
    n_out = C.c_size_t()
    rc = lib.bgc_flatten_pods(arr, C.c_size_t(n), keys, C.c_void_p(unit.ctypes.data), C.c_uint32(R), nsmap._h,
                              C.c_int(int(intern)), C.c_void_p(ns_id.ctypes.data), ptrs,
                              C.c_void_p(mask.ctypes.data), C.byref(n_out))
    if rc != OK:
        raise BgcError(rc, "bgc_flatten_pods: bad quantity or argument")
    m = n_out.value
    return ns_id[:m].copy(), [c[:m].copy() for c in req], mask[:m].copy(), unit


def hard_from_sheet_row(cpu: int, memory_gi: int, gpu: int, storage_gi: int, mig: int, quota_keys: Sequence[str]):
    """The `hard` values src/synchronizer.rs:249-281 writes for one sheet row, in canonical units."""
    R = len(quota_keys)
    keys = (C.c_char_p * R)(*[k.encode() for k in quota_keys])
    hard = np.empty(R, np.int64)
    unit = np.empty(R, np.uint8)
    rc = lib.bgc_hard_from_sheet_row(C.c_int64(cpu), C.c_int64(memory_gi), C.c_int64(gpu), C.c_int64(storage_gi),
                                     C.c_int64(mig), keys, C.c_uint32(R), C.c_void_p(hard.ctypes.data),
                                     C.c_void_p(unit.ctypes.data))
    if rc != OK:
        raise BgcError(rc, "bgc_hard_from_sheet_row: unknown quota key")
    return hard, unit


# ---- compact columns for the narrow-column entry points (bgc_compact_*, include/bgc_host.h) -----------------
lib.bgc_compact_plan.argtypes = [C.c_void_p, C.c_size_t, C.POINTER(C.c_uint8), C.POINTER(C.c_uint8)]
lib.bgc_compact_column.argtypes = [C.c_void_p, C.c_size_t, C.c_uint8, C.c_uint8, C.c_void_p]
lib.bgc_compact_ns_ids.argtypes = [C.c_void_p, C.c_size_t, C.c_uint32, C.c_void_p]
_NP_OF_T = {0: np.int64, 1: np.uint32, 2: np.uint16, 3: np.uint8, 4: np.int32}


def compact_columns(ns_id: np.ndarray, req: Sequence[np.ndarray], ns_count: int):
    """int64 columns (what flatten_pods returns) -> (ns ids u16/u32, [narrow columns], [shifts]) for
    Engine.usage_reduce_typed / quota_eval_typed: per column the largest common shift and the narrowest exact type."""
    n = ns_id.shape[0]
    cols, shifts = [], []
    for c in req:
        c = np.ascontiguousarray(c, dtype=np.int64)
        t, sh = C.c_uint8(), C.c_uint8()
        if lib.bgc_compact_plan(C.c_void_p(c.ctypes.data), n, C.byref(t), C.byref(sh)) != OK:
            raise BgcError(-1, "bgc_compact_plan")
        out = np.empty(n, _NP_OF_T[t.value])
        if lib.bgc_compact_column(C.c_void_p(c.ctypes.data), n, t, sh, C.c_void_p(out.ctypes.data)) != OK:
            raise BgcError(-1, "bgc_compact_column: a value does not fit the planned representation")
        cols.append(out); shifts.append(int(sh.value))
    ns_id = np.ascontiguousarray(ns_id, dtype=np.uint32)
    if ns_count > 65535:          # ids stay u32
        return ns_id, cols, shifts
    ns16 = np.empty(n, np.uint16)
    lib.bgc_compact_ns_ids(C.c_void_p(ns_id.ctypes.data), n, ns_count, C.c_void_p(ns16.ctypes.data))
    return ns16, cols, shifts
