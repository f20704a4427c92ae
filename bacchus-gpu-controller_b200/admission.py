"""Host-side mirror of the reference's admission handler interface.

`AdmissionConfig` ~ `struct Config` (reference src/admission.rs:22-39), `AdmissionRequest` ~ the
fields `mutate()` reads from `AdmissionRequest<DynamicObject>` (src/admission.rs:241-431), and
`mutate_batch()` ~ `mutate(req, state)` applied to a batch: strings are flattened on the host by
libbgc's C++ layer (bgc_flatten_requests, row A2), the decision runs on the GPU (K0, row A1), and
responses carry the reference's own message literals and JSON patches.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field
from typing import Optional, Sequence

import numpy as np

from . import lib, OK, BgcError, Engine

OPERATIONS = {"CREATE": 0, "UPDATE": 1, "DELETE": 2, "CONNECT": 3}
CLASS_NAMES = ("ALLOW", "ALLOW_PATCH", "DENY", "INVALID")


class _CConfig(C.Structure):
    _fields_ = [("oidc_username_prefix", C.c_char_p), ("default_role_name", C.c_char_p),
                ("authorized_group_names", C.POINTER(C.c_char_p)), ("n_authorized", C.c_size_t)]


class _CRequest(C.Structure):
    _fields_ = [("username", C.c_char_p), ("groups", C.POINTER(C.c_char_p)), ("n_groups", C.c_size_t),
                ("operation", C.c_int), ("has_object", C.c_int), ("name", C.c_char_p), ("parse_ok", C.c_int),
                ("spec_kube_username", C.c_char_p), ("has_quota", C.c_int), ("has_rolebinding", C.c_int)]


class _CResponse(C.Structure):
    _fields_ = [("allowed", C.c_uint8), ("klass", C.c_uint8), ("reason", C.c_uint8), ("patch_mask", C.c_uint8),
                ("message", C.c_char_p)]


lib.bgc_split_groups.restype = C.c_size_t
lib.bgc_render_patch.restype = C.c_long


@dataclass
class AdmissionConfig:
    oidc_username_prefix: str = "oidc:"      # chart default, values.yaml:85
    default_role_name: str = "edit"          # values.yaml:86
    authorized_group_names: list = field(default_factory=lambda: ["gpu", "admin"])  # values.yaml:87-89

    @classmethod
    def from_env(cls, env: dict) -> "AdmissionConfig":
        """envy::prefixed("CONF_") + comma_separated_deserialize (src/admission.rs:41-50,138)."""
        return cls(env["CONF_OIDC_USERNAME_PREFIX"], env["CONF_DEFAULT_ROLE_NAME"],
                   split_groups(env["CONF_AUTHORIZED_GROUP_NAMES"]))

    def _c(self):
        arr = (C.c_char_p * max(1, len(self.authorized_group_names)))(*[g.encode() for g in self.authorized_group_names])
        c = _CConfig(self.oidc_username_prefix.encode(), self.default_role_name.encode(),
                     C.cast(arr, C.POINTER(C.c_char_p)), len(self.authorized_group_names))
        return c, arr


@dataclass
class AdmissionRequest:
    username: Optional[str]
    groups: Optional[Sequence[str]]
    operation: str | int
    has_object: bool = True
    name: Optional[str] = None
    parse_ok: bool = True
    spec_kube_username: Optional[str] = None
    has_quota: bool = False
    has_rolebinding: bool = False

    @classmethod
    def from_review(cls, request: dict) -> "AdmissionRequest":
        """From the `request` member of an AdmissionReview v1 JSON document."""
        ui = request.get("userInfo") or {}
        obj = request.get("object")
        spec = obj.get("spec") if isinstance(obj, dict) else None
        ok = _parses_as_userbootstrap(obj) if obj is not None else False
        return cls(
            username=ui.get("username"), groups=ui.get("groups"),
            operation=str(request.get("operation", "")).upper(),
            has_object=obj is not None,
            name=(obj.get("metadata") or {}).get("name") if obj is not None else None,
            parse_ok=ok,
            spec_kube_username=spec.get("kube_username") if ok else None,
            has_quota=bool(ok and spec.get("quota") is not None),
            has_rolebinding=bool(ok and spec.get("rolebinding") is not None),
        )


def _parses_as_userbootstrap(obj) -> bool:
    """Whether `obj` deserialises as UserBootstrap (src/admission.rs:341; crd.rs:19-42): `spec` must
    be an object whose optional members have the right JSON types and whose rolebinding, when
    present, carries role_ref{apiGroup,kind,name}.  (serde behaviour restated from memory — unpinned.)"""
    if not isinstance(obj, dict) or not isinstance(obj.get("spec"), dict):
        return False
    spec = obj["spec"]
    if not isinstance(spec.get("kube_username", ""), (str, type(None))):
        return False
    for k in ("quota", "role"):
        if not isinstance(spec.get(k, {}), (dict, type(None))):
            return False
    rb = spec.get("rolebinding")
    if rb is not None:
        if not isinstance(rb, dict):
            return False
        rr = rb.get("role_ref")
        if not isinstance(rr, dict) or not all(isinstance(rr.get(k), str) for k in ("apiGroup", "kind", "name")):
            return False
        subs = rb.get("subjects")
        if subs is not None:
            if not isinstance(subs, list):
                return False
            if not all(isinstance(x, dict) and isinstance(x.get("kind"), str) and isinstance(x.get("name"), str) for x in subs):
                return False             # k8s-openapi Subject: kind and name are required
    quota = spec.get("quota")
    if isinstance(quota, dict) and quota.get("hard") is not None:
        hard = quota["hard"]
        if not isinstance(hard, dict) or not all(isinstance(v, (str, int, float)) and not isinstance(v, bool) for v in hard.values()):
            return False                 # a Quantity deserialises from a string or a number
    return True


@dataclass
class AdmissionResponse:
    allowed: bool
    klass: int
    reason: int
    patch_mask: int
    message: Optional[str]
    patch: Optional[str] = None  # JSONPatch text (src/admission.rs:429)

    @property
    def class_name(self) -> str:
        return CLASS_NAMES[self.klass]


def split_groups(value: str) -> list:
    raw = value.encode()
    scratch = C.create_string_buffer(len(raw) + 1)
    cap = raw.count(b",") + 1
    out = (C.c_char_p * cap)()
    n = lib.bgc_split_groups(raw, scratch, out, C.c_size_t(cap))
    return [out[i].decode() for i in range(n)]


def _pack_requests(reqs: Sequence[AdmissionRequest]):
    arr = (_CRequest * max(1, len(reqs)))()
    keep = []
    for i, q in enumerate(reqs):
        c = arr[i]
        c.username = None if q.username is None else q.username.encode()
        if q.groups is None:
            c.groups, c.n_groups = None, 0
        else:
            g = (C.c_char_p * max(1, len(q.groups)))(*[x.encode() for x in q.groups])
            keep.append(g)
            c.groups, c.n_groups = C.cast(g, C.POINTER(C.c_char_p)), len(q.groups)
        c.operation = OPERATIONS.get(q.operation, 3) if isinstance(q.operation, str) else int(q.operation)
        c.has_object = int(q.has_object)
        c.name = None if q.name is None else q.name.encode()
        c.parse_ok = int(q.parse_ok)
        c.spec_kube_username = None if q.spec_kube_username is None else q.spec_kube_username.encode()
        c.has_quota, c.has_rolebinding = int(q.has_quota), int(q.has_rolebinding)
    return arr, keep


def flatten_requests(config: AdmissionConfig, reqs: Sequence[AdmissionRequest]):
    """Row A2: strings → (op u8[n], flags u16[n])."""
    n = len(reqs)
    cc, keep_c = config._c()
    arr, keep_r = _pack_requests(reqs)
    op = np.empty(n, np.uint8)
    flags = np.empty(n, np.uint16)
    rc = lib.bgc_flatten_requests(C.byref(cc), arr, C.c_size_t(n), C.c_void_p(op.ctypes.data), C.c_void_p(flags.ctypes.data))
    if rc != OK:
        raise BgcError(rc, "bgc_flatten_requests: bad argument")
    return op, flags


def render_patch(config: AdmissionConfig, req: AdmissionRequest, patch_mask: int) -> Optional[str]:
    cc, keep_c = config._c()
    arr, keep_r = _pack_requests([req])
    buf = C.create_string_buffer(4096)
    n = lib.bgc_render_patch(C.byref(cc), arr, C.c_uint8(patch_mask), buf, C.c_size_t(4096))
    if n < 0:
        raise BgcError(int(n), "bgc_render_patch failed")
    return buf.value.decode() if n else None


def mutate_batch(engine: Engine, config: AdmissionConfig, reqs: Sequence[AdmissionRequest], with_patch: bool = True):
    """Batched `mutate(req, state)`: one AdmissionResponse per request, decided on the GPU."""
    n = len(reqs)
    cc, keep_c = config._c()
    arr, keep_r = _pack_requests(reqs)
    out = (_CResponse * max(1, n))()
    rc = lib.bgc_mutate_batch(engine._h, C.byref(cc), arr, C.c_size_t(n), out)
    engine._ck(rc)
    res = []
    for i in range(n):
        o = out[i]
        r = AdmissionResponse(bool(o.allowed), o.klass, o.reason, o.patch_mask,
                              None if o.message is None else o.message.decode())
        if with_patch and o.patch_mask:
            r.patch = render_patch(config, reqs[i], o.patch_mask)
        res.append(r)
    return res
