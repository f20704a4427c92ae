"""Small builders shared by tests."""


def ub_review(groups=("gpu",), spec=None, username="oidc:alice", name="alice", op="CREATE"):
    """An AdmissionReview v1 for a UserBootstrap write (the only kind the reference's webhook sees)."""
    user = {"username": username}
    if groups is not None:
        user["groups"] = list(groups) if isinstance(groups, (list, tuple)) else groups
    return {"apiVersion": "admission.k8s.io/v1", "kind": "AdmissionReview",
            "request": {"uid": "u-1", "operation": op, "userInfo": user,
                        "kind": {"group": "bacchus.io", "version": "v1", "kind": "UserBootstrap"},
                        "object": {"apiVersion": "bacchus.io/v1", "kind": "UserBootstrap", "metadata": {"name": name}, "spec": spec or {}}}}
