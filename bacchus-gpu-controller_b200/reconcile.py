"""Row N3: the reconcile plan — which objects the reference controller would server-side-apply for a
UserBootstrap (reference src/controller.rs:50-155), as (kind, namespace, name, path, body) tuples.
Wire format only: nothing is sent anywhere."""
from __future__ import annotations

import ctypes as C
import json

from . import lib

lib.bgc_reconcile_plan.restype = C.c_void_p
lib.bgc_reconcile_plan.argtypes = [C.c_char_p, C.c_size_t, C.c_char_p, C.c_size_t]
lib.bgc_plan_destroy.argtypes = [C.c_void_p]
lib.bgc_plan_count.restype = C.c_size_t
lib.bgc_plan_count.argtypes = [C.c_void_p]
lib.bgc_plan_item.argtypes = [C.c_void_p, C.c_size_t] + [C.POINTER(C.c_char_p)] * 5
lib.bgc_plan_requeue_seconds.argtypes = [C.c_int]


def plan(userbootstrap: dict | str | bytes) -> list:
    raw = userbootstrap if isinstance(userbootstrap, bytes) else (userbootstrap if isinstance(userbootstrap, str) else json.dumps(userbootstrap)).encode()
    err = C.create_string_buffer(256)
    h = lib.bgc_reconcile_plan(raw, len(raw), err, 256)
    if not h:
        raise ValueError(err.value.decode())
    try:
        out = []
        for i in range(lib.bgc_plan_count(h)):
            f = [C.c_char_p() for _ in range(5)]
            lib.bgc_plan_item(h, i, *[C.byref(x) for x in f])
            kind, ns, name, path, body = [None if x.value is None else x.value.decode() for x in f]
            out.append(dict(kind=kind, namespace=ns, name=name, path=path, body=body))
        return out
    finally:
        lib.bgc_plan_destroy(h)


def requeue_seconds(failed: bool) -> int:
    return lib.bgc_plan_requeue_seconds(int(failed))
