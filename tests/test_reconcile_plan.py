"""CPU tests for row N3: libbgc's reconcile plan (C++) vs the Python restatement of
src/controller.rs:50-155."""
import itertools
import json

import pytest

from oracle import reconcile_ref as Rf

QUOTA = {"hard": {"requests.cpu": "16", "requests.memory": "64Gi", "requests.nvidia.com/gpu": "2"}}
ROLE = {"metadata": {"name": "extra", "labels": {"a": "b"}, "ownerReferences": [{"name": "stale"}]},
        "rules": [{"apiGroups": [""], "resources": ["pods"], "verbs": ["get", "list"]}]}
RB = {"role_ref": {"apiGroup": "rbac.authorization.k8s.io", "kind": "ClusterRole", "name": "edit"},
      "subjects": [{"apiGroup": "rbac.authorization.k8s.io", "kind": "User", "name": "oidc:alice"}]}


def _ub(name="Alice", uid="1234-abcd", quota=None, role=None, rb=None, status=None):
    spec = {"kube_username": "alice"}
    if quota is not None:
        spec["quota"] = quota
    if role is not None:
        spec["role"] = role
    if rb is not None:
        spec["rolebinding"] = rb
    ub = {"apiVersion": "bacchus.io/v1", "kind": "UserBootstrap", "metadata": {"name": name, "uid": uid}, "spec": spec}
    if status is not None:
        ub["status"] = status
    return ub


@pytest.mark.parametrize("quota,role,rb,status", list(itertools.product(
    [None, QUOTA], [None, ROLE], [None, RB, {"role_ref": RB["role_ref"]}],
    [None, {"synchronized_with_sheet": False}, {"synchronized_with_sheet": True}])))
def test_plan_matches_reference_restatement(bgc, quota, role, rb, status):
    ub = _ub(quota=quota, role=role, rb=rb, status=status)
    got = bgc.reconcile.plan(ub)
    want = Rf.plan(ub)
    assert [(g["kind"], g["namespace"], g["name"], g["path"]) for g in got] == [(w["kind"], w["namespace"], w["name"], w["path"]) for w in want]
    for g, w in zip(got, want):
        assert json.loads(g["body"]) == w["body"]
    kinds = [g["kind"] for g in got]
    assert kinds[0] == "Namespace" and ("ResourceQuota" in kinds) == (quota is not None) and ("Role" in kinds) == (role is not None)
    assert ("RoleBinding" in kinds) == (rb is not None and status == {"synchronized_with_sheet": True})
    assert got[0]["name"] == "alice"                                    # lower-cased (controller.rs:63)
    assert json.loads(got[0]["body"])["metadata"]["ownerReferences"][0]["name"] == "Alice"   # the owner keeps its own name


def test_plan_errors_and_requeue(bgc):
    for bad in (_ub(name=None), _ub(uid=None)):
        with pytest.raises(ValueError):
            Rf.plan(bad)
        with pytest.raises(ValueError):
            bgc.reconcile.plan(bad)
    with pytest.raises(ValueError):
        bgc.reconcile.plan("{not json")
    assert bgc.reconcile.requeue_seconds(False) == 30 and bgc.reconcile.requeue_seconds(True) == 3
    # quota: null is "None" (no ResourceQuota), as serde reads Option
    ub = _ub()
    ub["spec"]["quota"] = None
    assert [g["kind"] for g in bgc.reconcile.plan(ub)] == ["Namespace"]
