"""Row N2: the quota-table producer — batch form of the reference synchronizer's sheet logic
(reference src/synchronizer.rs:63-169, 208-281), through libbgc's C++ host layer (bgc_sheet_*)."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import lib, OK, BgcError


class _Row(C.Structure):
    _fields_ = [("name", C.c_char_p), ("department", C.c_char_p), ("id_username", C.c_char_p), ("gpu_server", C.c_char_p),
                ("authorized", C.c_char_p), ("gpu_request", C.c_int64), ("cpu_request", C.c_int64),
                ("memory_request", C.c_int64), ("storage_request", C.c_int64), ("mig_request", C.c_int64)]


lib.bgc_sheet_infer_header.restype = C.c_char_p
lib.bgc_sheet_infer_header.argtypes = [C.c_char_p]
lib.bgc_sheet_parse_csv.restype = C.c_void_p
lib.bgc_sheet_parse_csv.argtypes = [C.c_char_p, C.c_size_t, C.c_char_p, C.c_size_t]
lib.bgc_sheet_destroy.argtypes = [C.c_void_p]
lib.bgc_sheet_row_count.restype = C.c_size_t
lib.bgc_sheet_row_count.argtypes = [C.c_void_p]
lib.bgc_sheet_skipped_rows.restype = C.c_size_t
lib.bgc_sheet_skipped_rows.argtypes = [C.c_void_p]
lib.bgc_sheet_get_row.argtypes = [C.c_void_p, C.c_size_t, C.POINTER(_Row)]
lib.bgc_sheet_match.argtypes = [C.c_void_p, C.c_char_p, C.c_void_p, C.c_size_t, C.c_void_p]
lib.bgc_sheet_row_quota.argtypes = [C.POINTER(_Row), C.c_void_p, C.c_void_p, C.c_void_p, C.c_char_p, C.c_size_t]


def infer_header(header: str):
    r = lib.bgc_sheet_infer_header(header.encode())
    return None if r is None else r.decode()


class Sheet:
    def __init__(self, content: str | bytes):
        raw = content.encode() if isinstance(content, str) else content
        err = C.create_string_buffer(512)
        self._h = lib.bgc_sheet_parse_csv(raw, len(raw), err, 512)
        if not self._h:
            raise ValueError(err.value.decode() or "parse failed")

    def __len__(self):
        return lib.bgc_sheet_row_count(self._h)

    @property
    def skipped(self) -> int:
        return lib.bgc_sheet_skipped_rows(self._h)

    def _raw_row(self, i: int) -> _Row:
        r = _Row()
        if lib.bgc_sheet_get_row(self._h, i, C.byref(r)) != OK:
            raise IndexError(i)
        return r

    def row(self, i: int) -> dict:
        r = self._raw_row(i)
        return {k: (getattr(r, k).decode() if isinstance(getattr(r, k), bytes) else getattr(r, k)) for k, _ in _Row._fields_}

    def match(self, server_name: str, names: list) -> list:
        arr = (C.c_char_p * max(1, len(names)))(*[n.encode() for n in names])
        out = (C.c_long * max(1, len(names)))()
        if lib.bgc_sheet_match(self._h, server_name.encode(), arr, len(names), out) != OK:
            raise BgcError(-1, "bgc_sheet_match")
        return [out[i] for i in range(len(names))]

    def quota(self, i: int):
        """({key: quantity string} in BTreeMap order, {key: canonical int})."""
        r = self._raw_row(i)
        keys = (C.c_char_p * 7)()
        quants = (C.c_char_p * 7)()
        canon = (C.c_int64 * 7)()
        scratch = C.create_string_buffer(512)
        if lib.bgc_sheet_row_quota(C.byref(r), keys, quants, canon, scratch, 512) != OK:
            raise BgcError(-1, "bgc_sheet_row_quota")
        return ({keys[k].decode(): quants[k].decode() for k in range(7)}, {keys[k].decode(): canon[k] for k in range(7)})

    def hard_table(self, server_name: str, names: list, quota_keys: list):
        """Rows → the [R][NS] `hard` table and hard_mask for bgc_quota_table_set: namespace i is the
        UserBootstrap names[i]; names without an authorised row get hard_mask 0 (no quota object)."""
        idx = self.match(server_name, names)
        hard = np.zeros((len(quota_keys), len(names)), np.int64)
        mask = np.zeros(len(names), np.uint8)
        for i, ri in enumerate(idx):
            if ri < 0:
                continue
            _, canon = self.quota(ri)
            for r, k in enumerate(quota_keys):
                hard[r, i] = canon[k]
            mask[i] = (1 << len(quota_keys)) - 1
        return hard, mask, idx

    def __del__(self):
        if getattr(self, "_h", None):
            lib.bgc_sheet_destroy(self._h)
            self._h = None
