"""Row N4: AdmissionReview v1 JSON in and out — the batched form of the reference's `mutate_handler`
(reference src/admission.rs:184-204): wire bytes → AdmissionRequest fields (C++ parser in libbgc) →
flatten (row A2) → decision on the GPU (K0, row A1) → AdmissionReview response JSON with the
reference's messages and base64 JSONPatch.  Byte-level parity with kube-core/serde_json is unpinned."""
from __future__ import annotations

import ctypes as C
from typing import Sequence

import numpy as np

from . import lib, OK, BgcError, Engine
from .admission import AdmissionConfig, _CRequest, _CResponse, _CConfig
from .pods import _Pod

lib.bgc_review_parse.restype = C.c_void_p
lib.bgc_review_parse.argtypes = [C.c_char_p, C.c_size_t, C.c_char_p, C.c_size_t]
lib.bgc_review_destroy.argtypes = [C.c_void_p]
lib.bgc_review_has_request.argtypes = [C.c_void_p]
lib.bgc_review_uid.restype = C.c_char_p
lib.bgc_review_uid.argtypes = [C.c_void_p]
lib.bgc_review_kind.restype = C.c_char_p
lib.bgc_review_kind.argtypes = [C.c_void_p]
lib.bgc_review_request.argtypes = [C.c_void_p, C.POINTER(_CRequest)]
lib.bgc_review_pod.argtypes = [C.c_void_p, C.POINTER(_Pod)]
lib.bgc_render_review_response.restype = C.c_long
lib.bgc_handle_pod_reviews.restype = C.c_int
lib.bgc_handle_pod_reviews.argtypes = [C.c_void_p, C.POINTER(C.c_char_p), C.POINTER(C.c_size_t), C.c_size_t, C.POINTER(C.c_char_p), C.c_void_p, C.c_uint32,
                                       C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_uint32, C.c_char_p, C.c_size_t, C.POINTER(C.c_size_t),
                                       C.POINTER(C.c_uint64)]
lib.bgc_handle_reviews.restype = C.c_int
lib.bgc_handle_reviews.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(C.c_char_p), C.POINTER(C.c_size_t), C.c_size_t, C.c_char_p, C.c_size_t,
                                   C.POINTER(C.c_size_t), C.POINTER(C.c_uint64)]
lib.bgc_render_review_response.argtypes = [C.c_char_p, C.POINTER(_CResponse), C.c_char_p, C.c_char_p, C.c_char_p, C.c_size_t]
lib.bgc_render_patch.argtypes = [C.POINTER(_CConfig), C.POINTER(_CRequest), C.c_uint8, C.c_char_p, C.c_size_t]


class Review:
    """One parsed AdmissionReview document (owns the strings the request view points into)."""

    def __init__(self, body: bytes | str):
        raw = body.encode() if isinstance(body, str) else body
        err = C.create_string_buffer(256)
        self._h = lib.bgc_review_parse(raw, len(raw), err, 256)
        if not self._h:
            raise ValueError(err.value.decode() or "malformed AdmissionReview")

    @property
    def has_request(self) -> bool:
        return bool(lib.bgc_review_has_request(self._h))

    @property
    def uid(self) -> str:
        return lib.bgc_review_uid(self._h).decode()

    @property
    def kind(self) -> str:
        return lib.bgc_review_kind(self._h).decode()

    def c_request(self) -> _CRequest:
        r = _CRequest()
        if lib.bgc_review_request(self._h, C.byref(r)) != OK:
            raise BgcError(-1, "review has no request")
        return r

    def request_fields(self) -> dict:
        r = self.c_request()
        dec = lambda b: None if b is None else b.decode()
        return dict(username=dec(r.username), groups=None if not r.groups else [r.groups[i].decode() for i in range(r.n_groups)],
                    operation=r.operation, has_object=bool(r.has_object), name=dec(r.name), parse_ok=bool(r.parse_ok),
                    spec_kube_username=dec(r.spec_kube_username), has_quota=bool(r.has_quota), has_rolebinding=bool(r.has_rolebinding))

    def c_pod(self) -> _Pod:
        p = _Pod()
        if lib.bgc_review_pod(self._h, C.byref(p)) != OK:
            raise BgcError(-1, "review carries no object")
        return p

    def __del__(self):
        if getattr(self, "_h", None):
            lib.bgc_review_destroy(self._h)
            self._h = None


def _render(uid: str, resp: _CResponse, message_override, patch) -> str:
    buf = C.create_string_buffer(8192)
    n = lib.bgc_render_review_response(uid.encode(), C.byref(resp), None if message_override is None else message_override.encode(),
                                       None if patch is None else patch, buf, 8192)
    if n < 0:
        raise BgcError(int(n), "bgc_render_review_response")
    return buf.value.decode()


def _parse_all(bodies):
    """One Review per body, None where the body is not a well-formed AdmissionReview: like the reference's
    handler, which rejects only the offending request (src/admission.rs:189-197), a bad body must not cost
    the other requests of the batch their answers."""
    out = []
    for b in bodies:
        try:
            out.append(Review(b))
        except ValueError:
            out.append(None)
    return out


def _invalid(message: str) -> str:
    return _render("", _CResponse(0, 3, 0, 0, None), message, None)


def handle_reviews(engine: Engine, config: AdmissionConfig, bodies: Sequence[bytes | str], stage_ms: dict | None = None) -> list:
    """Batched `mutate_handler` (src/admission.rs:184-204): one response JSON string per request body, through
    ONE C call (bgc_handle_reviews: parse, flatten, one K0 launch, render).  Reviews without a `request` and
    bodies that do not parse are answered `invalid` on their own like :189-197.  `stage_ms`, when given, is
    filled with the host time of the four stages."""
    n = len(bodies)
    if n == 0:
        return []
    blob, ptrs, lens = _marshal_bodies(bodies)
    cap = len(blob) // 2 + 768 * n + 64
    off = np.zeros(n + 1, np.uint64)
    stages = (C.c_uint64 * 4)()
    cc, keep = config._c()
    for _ in range(2):
        buf = np.empty(cap, np.uint8)
        rc = lib.bgc_handle_reviews(engine._h, C.byref(cc), ptrs.ctypes.data_as(C.POINTER(C.c_char_p)), lens.ctypes.data_as(C.POINTER(C.c_size_t)),
                                    C.c_size_t(n), buf.ctypes.data_as(C.c_char_p), C.c_size_t(cap), off.ctypes.data_as(C.POINTER(C.c_size_t)), stages)
        if rc == 0 or int(off[n]) <= cap:
            break
        cap = int(off[n])      # the call reported the size it needs
    engine._ck(rc)
    if stage_ms is not None:
        stage_ms.update(zip(("parse_flatten_release", "decide", "render", "copy"), (s / 1e6 for s in stages)))
    return buf[:int(off[n]) - 1].tobytes().decode().split("\0")   # responses are JSON text: no NUL inside one


def handle_reviews_stepwise(engine: Engine, config: AdmissionConfig, bodies: Sequence[bytes | str]) -> list:
    """The same handler spelled out call by call (bgc_review_parse → bgc_review_request → bgc_mutate_batch →
    bgc_render_patch → bgc_render_review_response); kept for the tests that compare the two."""
    reviews = _parse_all(bodies)
    live = [i for i, r in enumerate(reviews) if r is not None and r.has_request]
    creqs = (_CRequest * max(1, len(live)))()
    for k, i in enumerate(live):
        lib.bgc_review_request(reviews[i]._h, C.byref(creqs[k]))
    cc, keep = config._c()
    out = (_CResponse * max(1, len(live)))()
    engine._ck(lib.bgc_mutate_batch(engine._h, C.byref(cc), creqs, C.c_size_t(len(live)), out))
    res = [None] * len(reviews)
    pbuf = C.create_string_buffer(4096)
    for k, i in enumerate(live):
        patch = None
        if out[k].patch_mask:
            n = lib.bgc_render_patch(C.byref(cc), C.byref(creqs[k]), out[k].patch_mask, pbuf, 4096)
            if n < 0:
                raise BgcError(int(n), "bgc_render_patch")
            patch = pbuf.value
        res[i] = _render(reviews[i].uid, out[k], None, patch)
    for i, r in enumerate(reviews):
        if res[i] is None:   # only the offending request is answered `invalid`; the rest of the batch is served
            res[i] = _invalid("malformed AdmissionReview" if r is None else "invalid request: AdmissionReview carries no request")
    return res


def flatten_pod_reviews(bodies: Sequence[bytes | str], quota_keys: Sequence[str], nsmap, intern: bool = False, reviews=None):
    """Pod-create AdmissionReviews → (ns_id, req[R], req_mask) columns through bgc_flatten_pods.
    `reviews`: already parsed Review objects for `bodies` (so a caller that parsed them does not parse twice)."""
    from .pods import QUOTA_KEYS
    if reviews is None:
        reviews = [Review(b) for b in bodies]
    n = len(reviews)
    arr = (_Pod * max(1, n))()
    for i, r in enumerate(reviews):
        arr[i] = r.c_pod()
    R = len(quota_keys)
    keys = (C.c_char_p * R)(*[k.encode() for k in quota_keys])
    unit = np.array([QUOTA_KEYS[k] for k in quota_keys], np.uint8)
    ns_id = np.empty(n, np.uint32)
    req = [np.empty(n, np.int64) for _ in range(R)]
    mask = np.empty(n, np.uint8)
    ptrs = (C.c_void_p * R)(*[c.ctypes.data for c in req])
    n_out = C.c_size_t()
    rc = lib.bgc_flatten_pods(arr, C.c_size_t(n), keys, C.c_void_p(unit.ctypes.data), C.c_uint32(R), nsmap._h, C.c_int(int(intern)),
                              C.c_void_p(ns_id.ctypes.data), ptrs, C.c_void_p(mask.ctypes.data), C.byref(n_out))
    if rc != OK:
        raise BgcError(rc, "bgc_flatten_pods")
    m = n_out.value
    return ns_id[:m].copy(), [c[:m].copy() for c in req], mask[:m].copy(), unit


def _marshal_bodies(bodies):
    """bodies -> (blob kept alive, pointer array, length array) for the one-call handlers"""
    raw = [b.encode() if isinstance(b, str) else b for b in bodies]
    n = len(raw)
    blob = b"".join(raw)                                   # one buffer, pointers computed from the lengths
    lens = np.fromiter(map(len, raw), np.uint64, n)
    base = C.cast(C.c_char_p(blob), C.c_void_p).value or 0
    ptrs = base + np.concatenate([np.zeros(1, np.uint64), np.cumsum(lens[:-1], dtype=np.uint64)])
    return blob, ptrs, lens


def handle_pod_reviews(engine, bodies: Sequence[bytes | str], quota_keys: Sequence[str], nsmap, used=None, hard=None, wide: bool | None = None,
                       stage_ms: dict | None = None) -> list:
    """Pod-create AdmissionReviews, wire bytes to wire bytes, through ONE C call (bgc_handle_pod_reviews): parse ->
    flatten (row A5) -> snapshot verdicts on the GPU (K2) -> AdmissionReview responses, denied pods carrying the
    message kube-apiserver's ResourceQuota plugin would give.  `used` / `hard` ([R][NS] arrays) are only needed for
    the denial text; pass None to get bare allow/deny.  An `engine` that is not a real Engine (the CPU tests' stand-in)
    takes the call-by-call form.  Not a reference path: the reference's webhook never sees pods."""
    from .pods import QUOTA_KEYS
    if not isinstance(engine, Engine):
        return handle_pod_reviews_stepwise(engine, bodies, quota_keys, nsmap, used, hard, wide)
    R = len(quota_keys)
    if wide is None:
        wide = R > 4
    n = len(bodies)
    if n == 0:
        return []
    blob, ptrs, lens = _marshal_bodies(bodies)
    keys = (C.c_char_p * R)(*[k.encode() for k in quota_keys])
    unit = np.array([QUOTA_KEYS[k] for k in quota_keys], np.uint8)
    u = h = None
    ns_count = 0
    if used is not None and hard is not None:
        u, h = np.ascontiguousarray(used, dtype=np.int64), np.ascontiguousarray(hard, dtype=np.int64)
        if u.shape != h.shape or u.ndim != 2 or u.shape[0] != R:
            raise ValueError("used / hard must be [R][NS] arrays")
        ns_count = u.shape[1]
    cap = len(blob) // 4 + 1024 * n + 64
    off = np.zeros(n + 1, np.uint64)
    stages = (C.c_uint64 * 5)()
    for _ in range(2):
        buf = np.empty(cap, np.uint8)
        rc = lib.bgc_handle_pod_reviews(engine._h, ptrs.ctypes.data_as(C.POINTER(C.c_char_p)), lens.ctypes.data_as(C.POINTER(C.c_size_t)), C.c_size_t(n),
                                        keys, C.c_void_p(unit.ctypes.data), C.c_uint32(R), nsmap._h, C.c_int(int(bool(wide))),
                                        C.c_void_p(u.ctypes.data if u is not None else 0), C.c_void_p(h.ctypes.data if h is not None else 0),
                                        C.c_uint32(ns_count), buf.ctypes.data_as(C.c_char_p), C.c_size_t(cap),
                                        off.ctypes.data_as(C.POINTER(C.c_size_t)), stages)
        if rc == 0 or int(off[n]) <= cap:
            break
        cap = int(off[n])
    if rc == -1 and int(off[n]) == 0:
        raise ValueError("bgc_handle_pod_reviews: bad argument (a terminal pod in a create review, or R / wide do not match)")
    engine._ck(rc)
    if stage_ms is not None:
        stage_ms.update(zip(("parse_flatten_release", "gather", "verdicts", "render", "copy"), (s / 1e6 for s in stages)))
    return buf[:int(off[n]) - 1].tobytes().decode().split("\0")


def handle_pod_reviews_stepwise(engine, bodies: Sequence[bytes | str], quota_keys: Sequence[str], nsmap, used=None, hard=None, wide: bool | None = None) -> list:
    """The same handler call by call (bgc_review_parse / bgc_review_pod / bgc_flatten_pods / engine.quota_eval /
    bgc_render_quota_denial / bgc_render_review_response): the cross-check of the one-call form, and the form the CPU
    tests drive with a stand-in engine."""
    from .pods import quota_denial_message
    R = len(quota_keys)
    if wide is None:
        wide = R > 4
    parsed = _parse_all(bodies)
    good = []
    for i, r in enumerate(parsed):   # a body that is malformed, or carries no pod object, is answered `invalid` on its own
        if r is None:
            continue
        try:
            r.c_pod()
            good.append(i)
        except BgcError:
            pass
    reviews = [parsed[i] for i in good]
    ns_id, req, mask, _unit = flatten_pod_reviews(None, quota_keys, nsmap, intern=False, reviews=reviews)
    if ns_id.shape[0] != len(reviews):
        raise ValueError("terminal pods cannot appear in create reviews")
    verdict = engine.quota_eval(ns_id, req, mask, wide=wide) if wide else engine.quota_eval(ns_id, req, mask)
    bad = 0xFFFF if wide else 0xFF
    answers = {}
    for i, r in enumerate(reviews):
        v = int(verdict[i])
        if v == 0:
            resp, msg = _CResponse(1, 0, 0, 0, None), None
        elif v == bad:
            resp, msg = _CResponse(0, 2, 0, 0, None), "namespace is not known to the quota engine"
        else:
            ns = int(ns_id[i])
            msg = "exceeded quota" if used is None or hard is None else quota_denial_message(
                v, nsmap.name(ns), quota_keys, [int(c[i]) for c in req], [int(used[k][ns]) for k in range(R)],
                [int(hard[k][ns]) for k in range(R)], wide=wide)
            resp = _CResponse(0, 2, 0, 0, None)
        answers[good[i]] = _render(r.uid, resp, msg, None)
    return [answers.get(i) or _invalid("malformed AdmissionReview" if parsed[i] is None else "invalid request: AdmissionReview carries no pod object")
            for i in range(len(parsed))]
