"""oracle/reconcile_ref.py — TEST INFRASTRUCTURE ONLY.  Restates `reconcile()`,
/root/reference/src/controller.rs:50-155, as the list of objects it would apply.  PARITY UNPINNED
(no reference tests; body field order is k8s-openapi's, not restated here — bodies are compared as
parsed JSON)."""
from __future__ import annotations

PATCH_MANAGER = "bacchus-gpu-controller.bacchus.io"  # controller.rs:22


def plan(ub: dict) -> list:
    meta = ub.get("metadata") or {}
    if meta.get("name") is None or meta.get("uid") is None:
        raise ValueError("no controller owner reference")           # :52 unwrap / :55-62
    oref = {"apiVersion": "bacchus.io/v1", "kind": "UserBootstrap", "name": meta["name"], "uid": meta["uid"],
            "controller": True, "blockOwnerDeletion": True}
    name = meta["name"].lower()                                      # :63
    q = f"?fieldManager={PATCH_MANAGER}&force=true"                  # :67
    md = {"name": name, "ownerReferences": [oref]}
    spec = ub.get("spec") or {}
    out = [dict(kind="Namespace", namespace=None, name=name, path=f"/api/v1/namespaces/{name}{q}",
                body={"apiVersion": "v1", "kind": "Namespace", "metadata": md})]                      # :69-87
    if spec.get("quota") is not None:                                                                  # :89-110
        out.append(dict(kind="ResourceQuota", namespace=name, name=name, path=f"/api/v1/namespaces/{name}/resourcequotas/{name}{q}",
                        body={"apiVersion": "v1", "kind": "ResourceQuota", "metadata": md, "spec": spec["quota"]}))
    if spec.get("role") is not None:                                                                   # :112-124
        role = spec["role"]
        rmeta = dict(role.get("metadata") or {})
        rmeta["ownerReferences"] = [oref]
        body = {"apiVersion": "rbac.authorization.k8s.io/v1", "kind": "Role", "metadata": rmeta}
        if "rules" in role:
            body["rules"] = role["rules"]
        out.append(dict(kind="Role", namespace=name, name=name, path=f"/apis/rbac.authorization.k8s.io/v1/namespaces/{name}/roles/{name}{q}", body=body))
    rb = spec.get("rolebinding")
    if rb is not None and (ub.get("status") or {}).get("synchronized_with_sheet") is True:             # :126-152
        body = {"apiVersion": "rbac.authorization.k8s.io/v1", "kind": "RoleBinding", "metadata": md, "roleRef": rb.get("role_ref")}
        if rb.get("subjects") is not None:
            body["subjects"] = rb["subjects"]
        out.append(dict(kind="RoleBinding", namespace=name, name=name,
                        path=f"/apis/rbac.authorization.k8s.io/v1/namespaces/{name}/rolebindings/{name}{q}", body=body))
    return out
