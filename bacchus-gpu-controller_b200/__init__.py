"""bacchus-gpu-controller_b200 — batched admission / quota evaluation on B200 (sm_100a).

Python face of libbgc.so (include/bgc.h): a thin ctypes binding plus a host-side mirror of the
reference's handler interface (reference src/admission.rs:184-431).  All compute happens in the
hand-written CUDA kernels of csrc/; there is NO CPU fallback — importing this package without the
built library, or creating an Engine without a CUDA device, raises.

The directory name carries the reference's hyphens, so import it with
``importlib.import_module("bacchus-gpu-controller_b200")`` (see tests/conftest.py).
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("BGC_LIB_PATH") or os.path.join(_HERE, "libbgc.so")   # override: development builds only

if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
        "or `make -C bacchus-gpu-controller_b200/csrc`. There is no CPU fallback."
    )

lib = C.CDLL(LIB_PATH)

OK, E_BADARG, E_CUDA, E_OOM, E_STATE = 0, -1, -2, -3, -4
MAX_R = 8
MAX_R8 = 4
UNIT_COUNT, UNIT_MILLI, UNIT_BYTES = 0, 1, 2
REDUCE_GROUPED, REDUCE_ACCUMULATE = 1, 2
K1_AUTO, K1_GLOBAL, K1_SMEM_DIRECT, K1_SMEM_HASH, K1_CLUSTER8, K1_CLUSTER16 = 0, 1, 2, 3, 4, 5
VERDICT_BAD_NS = 0xFF
VERDICT16_BAD_NS = 0xFFFF
ALLOW, ALLOW_PATCH, DENY, INVALID = 0, 1, 2, 3
GEN_GPU, GEN_CPU_MILLI, GEN_MEM_BYTES, GEN_MIG = 0, 1, 2, 3
DIST_UNIFORM, DIST_ZIPF_CDF = 0, 1

# every symbol include/*.h declares (tests check the library exports all of them)
EXPORTS = [
    "bgc_ctx_create", "bgc_ctx_destroy", "bgc_last_error", "bgc_ctx_set_stream", "bgc_ctx_reset_stream", "bgc_ctx_sync",
    "bgc_timings_get", "bgc_policy_eval", "bgc_policy_eval_dev", "bgc_quota_table_set",
    "bgc_usage_reduce", "bgc_usage_reduce_dev", "bgc_usage_get", "bgc_usage_set", "bgc_usage_get_dev", "bgc_usage_set_dev", "bgc_quota_eval", "bgc_quota_eval16", "bgc_quota_eval16_dev",
    "bgc_quota_eval_dev", "bgc_usage_reduce_typed", "bgc_usage_reduce_typed_dev", "bgc_quota_eval_typed", "bgc_quota_eval_typed_dev",
    "bgc_ctx_set_k1_variant", "bgc_last_k1_variant", "bgc_version",
    "bgc_synth_pods_dev", "bgc_synth_pods_shard_dev", "bgc_synth_policy_dev", "bgc_checksum_u8_dev", "bgc_histogram_u8_dev",
    "bgc_split_groups", "bgc_flatten_requests", "bgc_reason_message", "bgc_mutate_batch",
    "bgc_render_patch", "bgc_parse_quantity", "bgc_nsmap_create", "bgc_nsmap_destroy",
    "bgc_nsmap_intern", "bgc_nsmap_lookup", "bgc_nsmap_size", "bgc_nsmap_name", "bgc_flatten_pods",
    "bgc_hard_from_sheet_row", "bgc_compact_plan", "bgc_compact_column", "bgc_compact_ns_ids", "bgc_usage_reduce_static_dev", "bgc_admit_greedy", "bgc_admit_greedy_dev", "bgc_admit_greedy16", "bgc_admit_greedy16_dev",
    "bgc_sheet_infer_header", "bgc_sheet_parse_csv", "bgc_sheet_destroy", "bgc_sheet_row_count", "bgc_sheet_skipped_rows",
    "bgc_sheet_get_row", "bgc_sheet_match", "bgc_sheet_row_quota",
    "bgc_review_parse", "bgc_review_destroy", "bgc_review_has_request", "bgc_review_uid", "bgc_review_kind",
    "bgc_review_request", "bgc_review_pod", "bgc_render_review_response", "bgc_handle_reviews", "bgc_handle_pod_reviews",
    "bgc_format_quantity", "bgc_render_quota_denial",
    "bgc_reconcile_plan", "bgc_plan_destroy", "bgc_plan_count", "bgc_plan_item", "bgc_plan_requeue_seconds",
]


class Timings(C.Structure):
    _fields_ = [("h2d_ns", C.c_uint64), ("kernel_ns", C.c_uint64), ("d2h_ns", C.c_uint64), ("launches", C.c_uint64)]


class SynthParams(C.Structure):
    _fields_ = [("seed", C.c_uint64), ("set", C.c_uint32), ("ns_count", C.c_uint32), ("dist", C.c_uint32),
                ("grouped", C.c_uint32), ("n_total", C.c_uint64), ("scramble_a", C.c_uint64),
                ("scramble_b", C.c_uint64), ("d_cdf", C.c_void_p), ("R", C.c_uint32), ("kind", C.c_uint32 * MAX_R)]


lib.bgc_last_error.restype = C.c_char_p
lib.bgc_last_error.argtypes = [C.c_void_p]
lib.bgc_last_k1_variant.restype = C.c_char_p
lib.bgc_last_k1_variant.argtypes = [C.c_void_p]
lib.bgc_version.restype = C.c_char_p
lib.bgc_reason_message.restype = C.c_char_p
lib.bgc_ctx_create.argtypes = [C.c_int, C.POINTER(C.c_void_p)]
lib.bgc_ctx_destroy.argtypes = [C.c_void_p]
lib.bgc_ctx_set_stream.argtypes = [C.c_void_p, C.c_void_p]
lib.bgc_ctx_reset_stream.argtypes = [C.c_void_p]
lib.bgc_ctx_sync.argtypes = [C.c_void_p]
lib.bgc_timings_get.argtypes = [C.c_void_p, C.POINTER(Timings)]
lib.bgc_ctx_set_k1_variant.argtypes = [C.c_void_p, C.c_int]
_P, _Z = C.c_void_p, C.c_size_t
lib.bgc_policy_eval.argtypes = [_P, _P, _P, _Z, _P, _P]
lib.bgc_policy_eval_dev.argtypes = [_P, _P, _P, _Z, _P, _P]
lib.bgc_quota_table_set.argtypes = [_P, C.c_uint32, C.c_uint32, _P, _P, _P]
lib.bgc_usage_reduce.argtypes = [_P, _P, _P, _Z, C.c_uint, _P]
lib.bgc_usage_reduce_dev.argtypes = [_P, _P, _P, _Z, C.c_uint, _P]
lib.bgc_usage_get.argtypes = [_P, _P]
lib.bgc_usage_set.argtypes = [_P, _P]
lib.bgc_usage_get_dev.argtypes = [_P, _P]
lib.bgc_usage_set_dev.argtypes = [_P, _P]
lib.bgc_quota_eval.argtypes = [_P, _P, _P, _P, _Z, _P]
lib.bgc_quota_eval_dev.argtypes = [_P, _P, _P, _P, _Z, _P]
lib.bgc_quota_eval16.argtypes = [_P, _P, _P, _P, _Z, _P]
lib.bgc_quota_eval16_dev.argtypes = [_P, _P, _P, _P, _Z, _P]
lib.bgc_admit_greedy.argtypes = [_P, _P, _P, _P, _Z, _P]
lib.bgc_admit_greedy_dev.argtypes = [_P, _P, _P, _P, _Z, _P]
lib.bgc_admit_greedy16.argtypes = [_P, _P, _P, _P, _Z, _P]
lib.bgc_admit_greedy16_dev.argtypes = [_P, _P, _P, _P, _Z, _P]
lib.bgc_synth_pods_dev.argtypes = [_P, C.POINTER(SynthParams), C.c_uint64, _Z, _P, _P, _P]
lib.bgc_synth_pods_shard_dev.argtypes = [_P, C.POINTER(SynthParams), C.c_uint64, _Z, C.c_uint32, C.c_uint32, _P, C.c_uint64, _Z,
                                         _P, _P, _P, _P, C.POINTER(C.c_uint64)]
lib.bgc_synth_policy_dev.argtypes = [_P, C.c_uint64, C.c_uint64, _Z, _P, _P]
lib.bgc_checksum_u8_dev.argtypes = [_P, _P, C.c_uint64, _Z, C.POINTER(C.c_uint64)]
lib.bgc_histogram_u8_dev.argtypes = [_P, _P, _Z, _P]


T_I64, T_U32, T_U16, T_U8, T_I32 = 0, 1, 2, 3, 4
_T_OF_NP = {np.dtype(np.int64): T_I64, np.dtype(np.uint32): T_U32, np.dtype(np.uint16): T_U16, np.dtype(np.uint8): T_U8,
            np.dtype(np.int32): T_I32}
_T_WIDTH = {T_I64: 8, T_U32: 4, T_U16: 2, T_U8: 1, T_I32: 4}


class Col(C.Structure):
    """bgc_col: one request column of the typed entry points (value = element << shift)."""
    _fields_ = [("data", C.c_void_p), ("type", C.c_uint8), ("shift", C.c_uint8), ("reserved", C.c_uint8 * 6)]


lib.bgc_usage_reduce_typed.argtypes = [_P, _P, C.c_int, C.POINTER(Col), _Z, C.c_uint, _P]
lib.bgc_usage_reduce_typed_dev.argtypes = [_P, _P, C.c_int, C.POINTER(Col), _Z, C.c_uint, _P]
lib.bgc_quota_eval_typed.argtypes = [_P, _P, C.c_int, C.POINTER(Col), _P, _Z, _P]
lib.bgc_quota_eval_typed_dev.argtypes = [_P, _P, C.c_int, C.POINTER(Col), _P, _Z, _P]


def _torch_type(t):
    import torch
    return {torch.int64: T_I64, torch.int32: T_I32, torch.uint8: T_U8, torch.int16: T_U16, torch.uint16: T_U16,
            torch.uint32: T_U32}[t.dtype]


def make_cols(ptrs_types_shifts):
    """[(address, BGC_T_*, shift), ...] -> ctypes array of bgc_col"""
    arr = (Col * max(1, len(ptrs_types_shifts)))()
    for i, (ptr, typ, sh) in enumerate(ptrs_types_shifts):
        arr[i].data, arr[i].type, arr[i].shift = ptr, typ, sh
    return arr


class BgcError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"bgc error {code}: {msg}")
        self.code = code


def _np_ptr(a):
    return None if a is None else C.c_void_p(a.ctypes.data)


def _dev_ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _check_np(a, dtype, n=None, name="array"):
    if not isinstance(a, np.ndarray) or a.dtype != dtype or not a.flags["C_CONTIGUOUS"]:
        raise TypeError(f"{name} must be a C-contiguous numpy array of {np.dtype(dtype).name}")
    if n is not None and a.shape[0] != n:
        raise ValueError(f"{name} has {a.shape[0]} rows, expected {n}")


class Engine:
    """One bgc_ctx bound to one CUDA device (replaces the reference's Arc<AppState>,
    src/admission.rs:145-147).  Not thread-safe: use one Engine per thread."""

    def __init__(self, device: int = 0):
        h = C.c_void_p()
        rc = lib.bgc_ctx_create(device, C.byref(h))
        if rc != OK:
            raise BgcError(rc, "bgc_ctx_create failed: no usable CUDA device (there is no CPU fallback)")
        self._h = h
        self.device = device
        self.NS = 0
        self.R = 0

    def close(self):
        if getattr(self, "_h", None):
            lib.bgc_ctx_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc):
        if rc != OK:
            raise BgcError(rc, lib.bgc_last_error(self._h).decode())

    # -- plumbing
    def set_stream(self, cuda_stream_handle: int | None):
        """A cudaStream_t handle (0 = legacy default stream); None = back to the engine's own stream."""
        if cuda_stream_handle is None:
            self._ck(lib.bgc_ctx_reset_stream(self._h))
        else:
            self._ck(lib.bgc_ctx_set_stream(self._h, C.c_void_p(cuda_stream_handle)))

    def use_torch_stream(self):
        import torch
        self.set_stream(torch.cuda.current_stream(self.device).cuda_stream)

    def sync(self):
        self._ck(lib.bgc_ctx_sync(self._h))

    def timings(self) -> dict:
        t = Timings()
        self._ck(lib.bgc_timings_get(self._h, C.byref(t)))
        return dict(h2d_ns=t.h2d_ns, kernel_ns=t.kernel_ns, d2h_ns=t.d2h_ns, launches=t.launches)

    def set_k1_variant(self, v: int):
        self._ck(lib.bgc_ctx_set_k1_variant(self._h, v))

    def last_k1_variant(self) -> str:
        return lib.bgc_last_k1_variant(self._h).decode()

    # -- K0 (row A1)
    def policy_eval(self, op: np.ndarray, flags: np.ndarray):
        n = op.shape[0]
        _check_np(op, np.uint8, name="op")
        _check_np(flags, np.uint16, n, "flags")
        code = np.empty(n, np.uint8)
        pm = np.empty(n, np.uint8)
        self._ck(lib.bgc_policy_eval(self._h, _np_ptr(op), _np_ptr(flags), n, _np_ptr(code), _np_ptr(pm)))
        return code, pm

    def policy_eval_dev(self, d_op, d_flags, d_code, d_pm):
        self._ck(lib.bgc_policy_eval_dev(self._h, _dev_ptr(d_op), _dev_ptr(d_flags), d_op.numel(),
                                         _dev_ptr(d_code), _dev_ptr(d_pm)))

    # -- quota table
    def quota_table_set(self, hard: np.ndarray, hard_mask: np.ndarray, unit):
        hard = np.ascontiguousarray(hard, dtype=np.int64)
        R, NS = hard.shape
        _check_np(hard_mask, np.uint8, NS, "hard_mask")
        u = np.asarray(unit, dtype=np.uint8)
        if u.shape[0] != R:
            raise ValueError("unit must have R entries")
        self._ck(lib.bgc_quota_table_set(self._h, NS, R, _np_ptr(hard), _np_ptr(hard_mask), _np_ptr(u)))
        self.NS, self.R = NS, R

    def usage_get(self) -> np.ndarray:
        out = np.empty((self.R, self.NS), np.int64)
        self._ck(lib.bgc_usage_get(self._h, _np_ptr(out)))
        return out

    def usage_set(self, used: np.ndarray):
        used = np.ascontiguousarray(used, dtype=np.int64)
        if used.shape != (self.R, self.NS):
            raise ValueError("used must be [R][NS]")
        self._ck(lib.bgc_usage_set(self._h, _np_ptr(used)))

    def usage_get_dev(self, d_used_out):
        self._ck(lib.bgc_usage_get_dev(self._h, _dev_ptr(d_used_out)))

    def usage_set_dev(self, d_used):
        self._ck(lib.bgc_usage_set_dev(self._h, _dev_ptr(d_used)))

    # -- K1 (row A3)
    def usage_reduce(self, ns_id: np.ndarray, req: list, flags: int = 0, want_used: bool = True):
        n = ns_id.shape[0]
        _check_np(ns_id, np.uint32, name="ns_id")
        for r, c in enumerate(req):
            _check_np(c, np.int64, n, f"req[{r}]")
        if len(req) != self.R:
            raise ValueError(f"expected {self.R} request columns")
        ptrs = (C.c_void_p * len(req))(*[c.ctypes.data for c in req])
        out = np.empty((self.R, self.NS), np.int64) if want_used else None
        self._ck(lib.bgc_usage_reduce(self._h, _np_ptr(ns_id), ptrs, n, flags, _np_ptr(out)))
        return out

    def usage_reduce_dev(self, d_ns_id, d_req: list, flags: int = 0, d_used_out=None, n: int | None = None):
        if len(d_req) != self.R:
            raise ValueError(f"expected {self.R} request columns")
        ptrs = (C.c_void_p * len(d_req))(*[t.data_ptr() for t in d_req])
        self._ck(lib.bgc_usage_reduce_dev(self._h, _dev_ptr(d_ns_id), ptrs, d_ns_id.numel() if n is None else n,
                                          flags, _dev_ptr(d_used_out)))

    # -- K2 (row A4)
    def quota_eval(self, ns_id: np.ndarray, req: list, req_mask: np.ndarray | None = None, wide: bool = False) -> np.ndarray:
        n = ns_id.shape[0]
        _check_np(ns_id, np.uint32, name="ns_id")
        for r, c in enumerate(req):
            _check_np(c, np.int64, n, f"req[{r}]")
        if req_mask is not None:
            _check_np(req_mask, np.uint8, n, "req_mask")
        if len(req) != self.R:
            raise ValueError(f"expected {self.R} request columns")
        ptrs = (C.c_void_p * len(req))(*[c.ctypes.data for c in req])
        if wide:      # u16 verdicts: missing << 8 | exceeded, any R <= 8
            verdict = np.empty(n, np.uint16)
            self._ck(lib.bgc_quota_eval16(self._h, _np_ptr(ns_id), ptrs, _np_ptr(req_mask), n, _np_ptr(verdict)))
        else:
            verdict = np.empty(n, np.uint8)
            self._ck(lib.bgc_quota_eval(self._h, _np_ptr(ns_id), ptrs, _np_ptr(req_mask), n, _np_ptr(verdict)))
        return verdict

    def quota_eval_dev(self, d_ns_id, d_req: list, d_req_mask, d_verdict, n: int | None = None):
        """d_verdict: uint8 tensor (R <= 4) or int16/uint16 tensor (u16 form, R <= 8)."""
        ptrs = (C.c_void_p * len(d_req))(*[t.data_ptr() for t in d_req])
        fn = lib.bgc_quota_eval16_dev if d_verdict.element_size() == 2 else lib.bgc_quota_eval_dev
        self._ck(fn(self._h, _dev_ptr(d_ns_id), ptrs, _dev_ptr(d_req_mask), d_ns_id.numel() if n is None else n, _dev_ptr(d_verdict)))

    # -- rows A3/A4 over narrow columns (bgc_*_typed)
    def usage_reduce_typed(self, ns_id: np.ndarray, cols: list, shifts=None, flags: int = 0, want_used: bool = True):
        """ns_id u16/u32, cols any of u8/u16/u32/i32/i64 numpy arrays; value = element << shift."""
        n = ns_id.shape[0]
        shifts = shifts or [0] * len(cols)
        arr = make_cols([(c.ctypes.data, _T_OF_NP[c.dtype], s) for c, s in zip(cols, shifts)])
        out = np.empty((self.R, self.NS), np.int64) if want_used else None
        self._ck(lib.bgc_usage_reduce_typed(self._h, _np_ptr(ns_id), _T_OF_NP[ns_id.dtype], arr, n, flags, _np_ptr(out)))
        return out

    def quota_eval_typed(self, ns_id: np.ndarray, cols: list, shifts=None, req_mask: np.ndarray | None = None) -> np.ndarray:
        n = ns_id.shape[0]
        shifts = shifts or [0] * len(cols)
        arr = make_cols([(c.ctypes.data, _T_OF_NP[c.dtype], s) for c, s in zip(cols, shifts)])
        verdict = np.empty(n, np.uint8)
        self._ck(lib.bgc_quota_eval_typed(self._h, _np_ptr(ns_id), _T_OF_NP[ns_id.dtype], arr, _np_ptr(req_mask), n, _np_ptr(verdict)))
        return verdict

    def usage_reduce_typed_dev(self, d_ns_id, d_cols: list, shifts=None, flags: int = 0, d_used_out=None, n: int | None = None, types=None):
        """torch device tensors; int16 tensors are read as u16, int32 as i32 unless `types` says otherwise."""
        shifts = shifts or [0] * len(d_cols)
        types = types or [_torch_type(t) for t in d_cols]
        arr = make_cols([(t.data_ptr(), ty, s) for t, ty, s in zip(d_cols, types, shifts)])
        ns_t = T_U16 if d_ns_id.element_size() == 2 else T_U32
        self._ck(lib.bgc_usage_reduce_typed_dev(self._h, _dev_ptr(d_ns_id), ns_t, arr, d_ns_id.numel() if n is None else n, flags,
                                                _dev_ptr(d_used_out)))

    def quota_eval_typed_dev(self, d_ns_id, d_cols: list, shifts, d_req_mask, d_verdict, n: int | None = None, types=None):
        shifts = shifts or [0] * len(d_cols)
        types = types or [_torch_type(t) for t in d_cols]
        arr = make_cols([(t.data_ptr(), ty, s) for t, ty, s in zip(d_cols, types, shifts)])
        ns_t = T_U16 if d_ns_id.element_size() == 2 else T_U32
        self._ck(lib.bgc_quota_eval_typed_dev(self._h, _dev_ptr(d_ns_id), ns_t, arr, _dev_ptr(d_req_mask),
                                              d_ns_id.numel() if n is None else n, _dev_ptr(d_verdict)))

    def usage_reduce_typed_ptr(self, ns_ptr: int, ns_type: int, cols, n: int, flags: int = 0, used_out_ptr: int = 0):
        """cols = make_cols([...]) over raw (pinned) host addresses"""
        self._ck(lib.bgc_usage_reduce_typed(self._h, C.c_void_p(ns_ptr), ns_type, cols, n, flags, C.c_void_p(used_out_ptr)))

    def quota_eval_typed_ptr(self, ns_ptr: int, ns_type: int, cols, mask_ptr: int, n: int, verdict_ptr: int):
        self._ck(lib.bgc_quota_eval_typed(self._h, C.c_void_p(ns_ptr), ns_type, cols, C.c_void_p(mask_ptr), n, C.c_void_p(verdict_ptr)))

    # -- K3 (row N1): sequential admission inside the batch; updates the usage table
    def admit_greedy(self, ns_id: np.ndarray, req: list, req_mask: np.ndarray | None = None, wide: bool = False) -> np.ndarray:
        n = ns_id.shape[0]
        _check_np(ns_id, np.uint32, name="ns_id")
        for r, c in enumerate(req):
            _check_np(c, np.int64, n, f"req[{r}]")
        if req_mask is not None:
            _check_np(req_mask, np.uint8, n, "req_mask")
        if len(req) != self.R:
            raise ValueError(f"expected {self.R} request columns")
        ptrs = (C.c_void_p * len(req))(*[c.ctypes.data for c in req])
        verdict = np.empty(n, np.uint16 if wide else np.uint8)      # wide: u16 verdicts, any R <= 8
        fn = lib.bgc_admit_greedy16 if wide else lib.bgc_admit_greedy
        self._ck(fn(self._h, _np_ptr(ns_id), ptrs, _np_ptr(req_mask), n, _np_ptr(verdict)))
        return verdict

    def admit_greedy_dev(self, d_ns_id, d_req: list, d_req_mask, d_verdict, n: int | None = None):
        ptrs = (C.c_void_p * len(d_req))(*[t.data_ptr() for t in d_req])
        fn = lib.bgc_admit_greedy16_dev if d_verdict.element_size() == 2 else lib.bgc_admit_greedy_dev
        self._ck(fn(self._h, _dev_ptr(d_ns_id), ptrs, _dev_ptr(d_req_mask), d_ns_id.numel() if n is None else n, _dev_ptr(d_verdict)))

    # -- raw-pointer variants (host pinned buffers, slices of larger allocations)
    def usage_reduce_ptr(self, ns_ptr: int, req_ptrs: list, n: int, flags: int = 0, used_out_ptr: int = 0):
        ptrs = (C.c_void_p * len(req_ptrs))(*req_ptrs)
        self._ck(lib.bgc_usage_reduce(self._h, C.c_void_p(ns_ptr), ptrs, n, flags, C.c_void_p(used_out_ptr)))

    def quota_eval_ptr(self, ns_ptr: int, req_ptrs: list, mask_ptr: int, n: int, verdict_ptr: int):
        ptrs = (C.c_void_p * len(req_ptrs))(*req_ptrs)
        self._ck(lib.bgc_quota_eval(self._h, C.c_void_p(ns_ptr), ptrs, C.c_void_p(mask_ptr), n, C.c_void_p(verdict_ptr)))

    def policy_eval_ptr(self, op_ptr: int, flags_ptr: int, n: int, code_ptr: int, pm_ptr: int):
        self._ck(lib.bgc_policy_eval(self._h, C.c_void_p(op_ptr), C.c_void_p(flags_ptr), n, C.c_void_p(code_ptr), C.c_void_p(pm_ptr)))

    # -- synthetic inputs on the device (include/bgc_synth.h)
    def synth_pods_dev(self, params: SynthParams, first: int, count: int, d_ns_id, d_req: list, d_req_mask=None):
        ptrs = (C.c_void_p * len(d_req))(*[t.data_ptr() for t in d_req])
        self._ck(lib.bgc_synth_pods_dev(self._h, C.byref(params), first, count, _dev_ptr(d_ns_id), ptrs, _dev_ptr(d_req_mask)))

    def synth_pods_shard_dev(self, params: SynthParams, first: int, count: int, world: int, rank: int, d_local_of=None,
                             out_pos: int = 0, d_ns_id=None, d_req: list | None = None, d_req_mask=None, d_src_index=None) -> int:
        """One shard (mix32(ns) % world == rank) of pods [first, first+count) of ONE global stream, in stream
        order, appended at out_pos of the shard columns; d_ns_id=None counts only.  Returns the pods kept."""
        kept = C.c_uint64()
        ptrs = (C.c_void_p * len(d_req))(*[t.data_ptr() for t in d_req]) if d_req else None
        cap = d_ns_id.numel() if d_ns_id is not None else 0
        self._ck(lib.bgc_synth_pods_shard_dev(self._h, C.byref(params), first, count, world, rank, _dev_ptr(d_local_of), out_pos, cap,
                                              _dev_ptr(d_ns_id), ptrs, _dev_ptr(d_req_mask), _dev_ptr(d_src_index), C.byref(kept)))
        return kept.value

    def synth_policy_dev(self, seed: int, first: int, count: int, d_op, d_flags):
        self._ck(lib.bgc_synth_policy_dev(self._h, seed, first, count, _dev_ptr(d_op), _dev_ptr(d_flags)))

    def checksum_u8_dev(self, d_v, first: int = 0, count: int | None = None) -> int:
        out = C.c_uint64()
        self._ck(lib.bgc_checksum_u8_dev(self._h, _dev_ptr(d_v), first, d_v.numel() if count is None else count, C.byref(out)))
        return out.value

    def histogram_u8_dev(self, d_v, count: int | None = None) -> np.ndarray:
        out = np.zeros(256, np.uint64)
        self._ck(lib.bgc_histogram_u8_dev(self._h, _dev_ptr(d_v), d_v.numel() if count is None else count, _np_ptr(out)))
        return out


def synth_params(seed, set_, NS, dist, grouped, n_total, a, b, d_cdf, kinds) -> SynthParams:
    p = SynthParams()
    p.seed, p.set, p.ns_count, p.dist, p.grouped, p.n_total = seed, set_, NS, dist, int(grouped), n_total
    p.scramble_a, p.scramble_b = a, b
    p.d_cdf = d_cdf.data_ptr() if d_cdf is not None else None
    p.R = len(kinds)
    for i, k in enumerate(kinds):
        p.kind[i] = k
    return p


def version() -> str:
    return lib.bgc_version().decode()


from . import admission, pods, shard, sheet, review, reconcile  # noqa: E402,F401
