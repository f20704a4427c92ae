"""oracle/policy_ref.py — TEST INFRASTRUCTURE ONLY (second, independent CPU restatement).

Restates /root/reference/src/admission.rs:184-431 at the level of the AdmissionReview JSON the
webhook receives (a Python dict), so that it shares no code or data model with oracle/policy_ref.c
(which works on flattened C strings).  PARITY UNPINNED: the reference has no tests and cannot be
built here; both restatements follow the source text and are diffed against each other and against
tests/golden/policy_truth_table.csv.

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may import this module.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Any, Optional

ALLOW, ALLOW_PATCH, DENY, INVALID = 0, 1, 2, 3
CLASS_NAMES = ("ALLOW", "ALLOW_PATCH", "DENY", "INVALID")

# reason → (class, message literal, reference line)
REASONS = {
    0: (None, None, "admission.rs:427/429"),
    1: (INVALID, "cannot get requester's username from request", "admission.rs:251"),
    2: (DENY, "user is not in authorized group", "admission.rs:277"),
    3: (DENY, "normal user is not allowed to delete resource", "admission.rs:287"),
    4: (DENY, "normal user is not allowed to update resource", "admission.rs:298"),
    5: (INVALID, "invalid operation", "admission.rs:306"),
    6: (INVALID, "cannot get resource name from request", "admission.rs:324"),
    7: (DENY, "username not match with resource name", "admission.rs:333"),
    8: (INVALID, None, "admission.rs:345"),  # serde error text, unpinned
    9: (DENY, "kube_username field is empty. you are an admin, so fill it", "admission.rs:369"),
    10: (DENY, "quota field is not empty. you are a normal user, so leave it empty", "admission.rs:379"),
    11: (DENY, "rolebinding field is not empty. you are a normal user, so leave it empty", "admission.rs:420"),
    12: (ALLOW, None, "admission.rs:293"),
    13: (ALLOW, None, "admission.rs:317"),
}

P_USER, P_RB, P_SUBJ_ORIG = 1, 2, 4
OPS = {"CREATE": 0, "UPDATE": 1, "DELETE": 2, "CONNECT": 3}

F_HAS_USERNAME, F_IS_NORMAL, F_IN_GROUP, F_HAS_OBJECT, F_HAS_NAME = 1, 2, 4, 8, 16
F_NAME_EQ, F_PARSE_OK, F_KUBE_USERNAME_NONEMPTY, F_HAS_QUOTA, F_HAS_ROLEBINDING = 32, 64, 128, 256, 512


@dataclass
class Config:
    """admission.rs:22-39 (only the fields mutate() reads)."""
    oidc_username_prefix: str = "oidc:"          # values.yaml:85
    default_role_name: str = "edit"              # values.yaml:86
    authorized_group_names: list = field(default_factory=lambda: ["gpu", "admin"])  # values.yaml:87-89

    @staticmethod
    def groups_from_env(value: str) -> list:
        """comma_separated_deserialize, admission.rs:41-50: split(',') without trimming."""
        return value.split(",")


@dataclass
class Decision:
    klass: int
    reason: int
    patch_mask: int = 0
    message: Optional[str] = None
    patches: list = field(default_factory=list)  # JSON-patch operations in push order

    @property
    def code(self) -> int:
        return (self.klass << 4) | self.reason


def _parses_as_userbootstrap(obj: dict) -> bool:
    """Approximation of DynamicObject::try_parse::<UserBootstrap>() (admission.rs:341).

    [from memory — unpinned] kube-derive makes `spec` a required field; every field of
    UserBootstrapSpec is Option (crd.rs:19-31); RoleBinding.role_ref is required (crd.rs:40) and
    RoleRef needs apiGroup/kind/name strings (k8s-openapi 0.18).  Type mismatches fail serde.
    """
    spec = obj.get("spec")
    if not isinstance(spec, dict):
        return False
    ku = spec.get("kube_username")
    if ku is not None and not isinstance(ku, str):
        return False
    for key in ("quota", "role"):
        v = spec.get(key)
        if v is not None and not isinstance(v, dict):
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
            for sub in subs:   # k8s-openapi 0.18 Subject: `kind` and `name` are required
                if not isinstance(sub, dict) or not isinstance(sub.get("kind"), str) or not isinstance(sub.get("name"), str):
                    return False
    quota = spec.get("quota")
    if isinstance(quota, dict) and quota.get("hard") is not None:   # BTreeMap<String, Quantity>; Quantity <- string | number
        hard = quota["hard"]
        if not isinstance(hard, dict):
            return False
        for v in hard.values():
            if isinstance(v, bool) or not isinstance(v, (str, int, float)):
                return False
    return True


def mutate(request: dict, config: Config) -> Decision:
    """`mutate(req, state)` on the `request` member of an AdmissionReview v1 (admission.rs:241-431)."""
    user_info = request.get("userInfo") or {}
    username = user_info.get("username")
    if username is None:                                             # :248-255
        return Decision(INVALID, 1, message=REASONS[1][1])

    prefix = config.oidc_username_prefix
    if username.startswith(prefix):                                  # :219
        normal, kube_username = True, username[len(prefix):]         # :222-228
    else:
        normal, kube_username = False, username                      # :231-236
    original_username = username

    groups = user_info.get("groups") or []                           # :263-267
    is_in_group = any(g in config.authorized_group_names for g in groups)  # :268-270

    op = str(request.get("operation", "")).upper()
    if op == "CREATE":                                               # :273-283
        if normal and not is_in_group:
            return Decision(DENY, 2, message=REASONS[2][1])
    elif op == "DELETE":                                             # :284-294
        if normal:
            return Decision(DENY, 3, message=REASONS[3][1])
        return Decision(ALLOW, 12)
    elif op == "UPDATE":                                             # :295-304
        if normal:
            return Decision(DENY, 4, message=REASONS[4][1])
    else:                                                            # :305-309
        return Decision(INVALID, 5, message=REASONS[5][1])

    obj = request.get("object")
    if obj is None:                                                  # :312-318
        return Decision(ALLOW, 13)

    name = (obj.get("metadata") or {}).get("name")
    if name is None:                                                 # :321-328
        return Decision(INVALID, 6, message=REASONS[6][1])

    if normal and kube_username != name:                             # :331-338
        return Decision(DENY, 7, message=REASONS[7][1])

    if not _parses_as_userbootstrap(obj):                            # :341-347
        return Decision(INVALID, 8, message="<serde error text>")
    spec = obj["spec"]

    patches: list = []
    mask = 0
    if normal:                                                       # :352-358
        patches.append({"op": "add", "path": "/spec/kube_username", "value": kube_username})
        mask |= P_USER
    elif not (spec.get("kube_username") or ""):                      # :359-373
        return Decision(DENY, 9, message=REASONS[9][1])

    if spec.get("quota") is not None and normal:                     # :376-383
        return Decision(DENY, 10, message=REASONS[10][1])

    if spec.get("rolebinding") is None:                              # :386-416
        patches.append({"op": "add", "path": "/spec/rolebinding", "value": {}})
        subject = original_username if normal else spec["kube_username"]   # :392-397
        mask |= P_RB | (P_SUBJ_ORIG if normal else 0)
        patches.append({
            "op": "add", "path": "/spec/rolebinding",
            "value": {
                "role_ref": {"apiGroup": "rbac.authorization.k8s.io", "kind": "ClusterRole",
                             "name": config.default_role_name},
                "subjects": [{"apiGroup": "rbac.authorization.k8s.io", "kind": "User", "name": subject}],
            },
        })
    elif normal:                                                     # :417-424
        return Decision(DENY, 11, message=REASONS[11][1])

    if not patches:                                                  # :426-427
        return Decision(ALLOW, 0)
    return Decision(ALLOW_PATCH, 0, patch_mask=mask, patches=patches)  # :429


def flatten(request: dict, config: Config) -> tuple:
    """Row A2: the predicate bits of one request.  Returns (op, flags)."""
    f = 0
    user_info = request.get("userInfo") or {}
    username = user_info.get("username")
    kube_username = None
    if username is not None:
        f |= F_HAS_USERNAME
        if username.startswith(config.oidc_username_prefix):
            f |= F_IS_NORMAL
            kube_username = username[len(config.oidc_username_prefix):]
        else:
            kube_username = username
    if any(g in config.authorized_group_names for g in (user_info.get("groups") or [])):
        f |= F_IN_GROUP
    obj = request.get("object")
    if obj is not None:
        f |= F_HAS_OBJECT
        name = (obj.get("metadata") or {}).get("name")
        if name is not None:
            f |= F_HAS_NAME
            if kube_username is not None and kube_username == name:
                f |= F_NAME_EQ
        if _parses_as_userbootstrap(obj):
            f |= F_PARSE_OK
            spec = obj["spec"]
            if spec.get("kube_username"):
                f |= F_KUBE_USERNAME_NONEMPTY
            if spec.get("quota") is not None:
                f |= F_HAS_QUOTA
            if spec.get("rolebinding") is not None:
                f |= F_HAS_ROLEBINDING
    op = OPS.get(str(request.get("operation", "")).upper(), 3)
    return op, f


def decide_bits(op: int, flags: int) -> tuple:
    """Row A1 batched form on bits.  Returns (code, patch_mask); code = class<<4 | reason.

    Written as a priority list (first match wins) rather than nested control flow, so that it is
    structurally different from policy_ref.c::policy_ref_bits.
    """
    b = lambda m: bool(flags & m)
    normal = b(F_IS_NORMAL)
    continuing = op in (0, 1)  # Create / Update fall through to the object checks
    rules = [
        (not b(F_HAS_USERNAME), INVALID, 1),
        (op == 0 and normal and not b(F_IN_GROUP), DENY, 2),
        (op == 2 and normal, DENY, 3),
        (op == 2, ALLOW, 12),
        (op == 1 and normal, DENY, 4),
        (not continuing, INVALID, 5),
        (not b(F_HAS_OBJECT), ALLOW, 13),
        (not b(F_HAS_NAME), INVALID, 6),
        (normal and not b(F_NAME_EQ), DENY, 7),
        (not b(F_PARSE_OK), INVALID, 8),
        (not normal and not b(F_KUBE_USERNAME_NONEMPTY), DENY, 9),
        (normal and b(F_HAS_QUOTA), DENY, 10),
        (normal and b(F_HAS_ROLEBINDING), DENY, 11),
    ]
    for cond, klass, reason in rules:
        if cond:
            return (klass << 4) | reason, 0
    mask = (P_USER if normal else 0)
    if not b(F_HAS_ROLEBINDING):
        mask |= P_RB | (P_SUBJ_ORIG if normal else 0)
    return ((ALLOW_PATCH if mask else ALLOW) << 4), mask


def request_from_bits(op: int, flags: int, config: Config) -> Optional[dict]:
    """Build a concrete AdmissionReview `request` whose predicates equal (op, flags), or None if the
    combination is self-contradictory at the string level (e.g. has_name without has_object)."""
    b = lambda m: bool(flags & m)
    if not b(F_HAS_USERNAME) and (b(F_IS_NORMAL) or b(F_NAME_EQ)):
        return None
    if not b(F_HAS_OBJECT) and (flags & (F_HAS_NAME | F_NAME_EQ | F_PARSE_OK | F_KUBE_USERNAME_NONEMPTY | F_HAS_QUOTA | F_HAS_ROLEBINDING)):
        return None
    if not b(F_HAS_NAME) and b(F_NAME_EQ):
        return None
    if not b(F_PARSE_OK) and (flags & (F_KUBE_USERNAME_NONEMPTY | F_HAS_QUOTA | F_HAS_ROLEBINDING)):
        return None
    req: dict = {"uid": "u-%d-%d" % (op, flags), "operation": ["CREATE", "UPDATE", "DELETE", "CONNECT"][op], "userInfo": {}}
    stripped = "alice"
    if b(F_HAS_USERNAME):
        req["userInfo"]["username"] = (config.oidc_username_prefix + stripped) if b(F_IS_NORMAL) else stripped
        if not b(F_IS_NORMAL) and stripped.startswith(config.oidc_username_prefix):
            return None  # empty prefix makes everyone Normal (admission.rs:219)
    req["userInfo"]["groups"] = ["system:authenticated"] + ([config.authorized_group_names[0]] if b(F_IN_GROUP) else [])
    if b(F_HAS_OBJECT):
        obj: dict = {"apiVersion": "bacchus.io/v1", "kind": "UserBootstrap", "metadata": {}}
        if b(F_HAS_NAME):
            obj["metadata"]["name"] = stripped if b(F_NAME_EQ) else "Alice"  # case differs → mismatch
        if b(F_PARSE_OK):
            spec: dict = {}
            if b(F_KUBE_USERNAME_NONEMPTY):
                spec["kube_username"] = "bob"
            if b(F_HAS_QUOTA):
                spec["quota"] = {"hard": {"requests.cpu": "8"}}
            if b(F_HAS_ROLEBINDING):
                spec["rolebinding"] = {"role_ref": {"apiGroup": "rbac.authorization.k8s.io", "kind": "ClusterRole", "name": "view"}}
            obj["spec"] = spec
        else:
            obj["spec"] = "not-an-object"
        req["object"] = obj
    return req
