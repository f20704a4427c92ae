"""CPU tests: the policy oracle (reference src/admission.rs:206-431 restated) against the committed
truth table, its second restatement, and string-level edge cases."""
import csv
import os

import numpy as np
import pytest

from oracle import policy_ref as P

GOLD = os.path.join(os.path.dirname(__file__), "golden", "policy_truth_table.csv")
CFG = ("oidc:", "edit", ["gpu", "admin"])


def load_golden():
    with open(GOLD) as f:
        rows = list(csv.DictReader(f))
    assert len(rows) == 4096
    op = np.array([int(r["op"]) for r in rows], np.uint8)
    fl = np.array([int(r["flags"]) for r in rows], np.uint16)
    code = np.array([int(r["code"]) for r in rows], np.uint8)
    pm = np.array([int(r["patch_mask"]) for r in rows], np.uint8)
    return op, fl, code, pm


def test_c_oracle_matches_golden_table(oracle):
    op, fl, code, pm = load_golden()
    c, m = oracle.policy_bits(op, fl)
    assert np.array_equal(c, code) and np.array_equal(m, pm)
    c8, m8 = oracle.policy_bits(op, fl, threads=4)
    assert np.array_equal(c8, code) and np.array_equal(m8, pm)


def test_python_bits_matches_golden_table():
    op, fl, code, pm = load_golden()
    for o, f, c, m in zip(op, fl, code, pm):
        assert P.decide_bits(int(o), int(f)) == (int(c), int(m))


def test_golden_table_is_exhaustive_and_cites_reference():
    with open(GOLD) as f:
        rows = list(csv.DictReader(f))
    assert {(int(r["op"]), int(r["flags"])) for r in rows} == {(o, f) for o in range(4) for f in range(1024)}
    assert all(r["reference_line"].startswith("admission.rs:") for r in rows)
    # every reason of mutate() is reachable
    assert {int(r["reason"]) for r in rows} == set(range(14))


def test_dict_level_mutate_agrees_with_bits_on_every_constructible_request():
    cfg = P.Config()
    built = 0
    for op in range(4):
        for fl in range(1024):
            req = P.request_from_bits(op, fl, cfg)
            if req is None:
                continue
            built += 1
            assert P.flatten(req, cfg) == (op, fl)
            d = P.mutate(req, cfg)
            assert (d.code, d.patch_mask) == P.decide_bits(op, fl)
    assert built == 600


def _c_req(**kw):
    base = dict(username="oidc:alice", groups=["gpu"], operation=0, has_object=True, name="alice", parse_ok=True,
                spec_kube_username=None, has_quota=False, has_rolebinding=False)
    base.update(kw)
    return base


def test_string_level_c_oracle_agrees_with_dict_level_python(oracle):
    """The two restatements share no code; drive both with the same concrete requests."""
    cfg_py = P.Config()
    n = 0
    for op in range(4):
        for fl in range(1024):
            req = P.request_from_bits(op, fl, cfg_py)
            if req is None:
                continue
            ui = req["userInfo"]
            obj = req.get("object")
            ok = obj is not None and isinstance(obj.get("spec"), dict)
            spec = obj["spec"] if ok else {}
            cre = dict(username=ui.get("username"), groups=ui.get("groups"), operation=op, has_object=obj is not None,
                       name=(obj or {}).get("metadata", {}).get("name") if obj else None, parse_ok=ok,
                       spec_kube_username=spec.get("kube_username"), has_quota="quota" in spec,
                       has_rolebinding="rolebinding" in spec)
            d_c = oracle.policy_mutate_strings(CFG, **cre)
            d_py = P.mutate(req, cfg_py)
            assert (d_c["klass"], d_c["reason"], d_c["patch_mask"]) == (d_py.klass, d_py.reason, d_py.patch_mask)
            if d_py.reason not in (0, 8, 12, 13):
                assert d_c["message"] == d_py.message
            assert oracle.policy_flatten_strings(CFG, **cre) == (op, fl)
            if d_py.patch_mask & P.P_USER:
                assert d_c["kube_username_patch"] == d_py.patches[0]["value"]
            if d_py.patch_mask & P.P_RB:
                assert d_c["subject_name"] == d_py.patches[-1]["value"]["subjects"][0]["name"]
            n += 1
    assert n == 600


def test_edge_empty_prefix_makes_everyone_normal(oracle):
    # admission.rs:219 — "".starts_with("") is true, so an empty prefix classifies admins as Normal
    d = oracle.policy_mutate_strings(("", "edit", ["gpu"]), **_c_req(username="root", groups=[], name="root"))
    assert (d["klass"], d["reason"]) == (P.DENY, 2)
    d = oracle.policy_mutate_strings(("", "edit", ["gpu"]), **_c_req(username="root", groups=["gpu"], name="root"))
    assert (d["klass"], d["patch_mask"]) == (P.ALLOW_PATCH, 7)
    assert d["kube_username_patch"] == "root" and d["subject_name"] == "root"


def test_edge_username_equal_to_prefix(oracle):
    # kube_username == "" after strip_prefix (admission.rs:222-227); name "" matches, anything else denies
    d = oracle.policy_mutate_strings(CFG, **_c_req(username="oidc:", name=""))
    assert (d["klass"], d["kube_username_patch"]) == (P.ALLOW_PATCH, "")
    d = oracle.policy_mutate_strings(CFG, **_c_req(username="oidc:", name="x"))
    assert (d["klass"], d["reason"]) == (P.DENY, 7)


def test_edge_group_list_from_empty_env(oracle):
    # admission.rs:45-49: "" splits to [""], so only a literally empty group name is authorised
    assert oracle.split_groups("") == [""]
    assert oracle.split_groups("gpu,admin") == ["gpu", "admin"]
    assert oracle.split_groups("gpu, admin,") == ["gpu", " admin", ""]
    assert P.Config.groups_from_env("gpu, admin,") == ["gpu", " admin", ""]
    d = oracle.policy_mutate_strings(("oidc:", "edit", [""]), **_c_req(groups=["gpu"]))
    assert (d["klass"], d["reason"]) == (P.DENY, 2)
    d = oracle.policy_mutate_strings(("oidc:", "edit", [""]), **_c_req(groups=[""]))
    assert d["klass"] == P.ALLOW_PATCH
    # " admin" (untrimmed) does not authorise "admin"
    d = oracle.policy_mutate_strings(("oidc:", "edit", ["gpu", " admin"]), **_c_req(groups=["admin"]))
    assert (d["klass"], d["reason"]) == (P.DENY, 2)


def test_edge_case_differing_name_is_denied(oracle):
    # admission.rs:332 compares case-sensitively although controller.rs:63 lowercases later
    d = oracle.policy_mutate_strings(CFG, **_c_req(name="Alice"))
    assert (d["klass"], d["reason"], d["message"]) == (P.DENY, 7, "username not match with resource name")


def test_edge_quota_wins_over_rolebinding_for_normal_user(oracle):
    d = oracle.policy_mutate_strings(CFG, **_c_req(has_quota=True, has_rolebinding=True))
    assert (d["klass"], d["reason"]) == (P.DENY, 10)


def test_edge_admin_update_with_rolebinding_is_plain_allow(oracle):
    d = oracle.policy_mutate_strings(CFG, **_c_req(username="admin", groups=None, operation=1, name="bob",
                                                   spec_kube_username="bob", has_quota=True, has_rolebinding=True))
    assert (d["klass"], d["reason"], d["patch_mask"], d["message"]) == (P.ALLOW, 0, 0, None)


def test_edge_admin_default_rolebinding_subject_is_spec_kube_username(oracle):
    d = oracle.policy_mutate_strings(CFG, **_c_req(username="admin", groups=None, name="bob", spec_kube_username="bob"))
    assert (d["klass"], d["patch_mask"], d["subject_name"]) == (P.ALLOW_PATCH, P.P_RB, "bob")


def test_edge_delete_with_object_present(oracle):
    d = oracle.policy_mutate_strings(CFG, **_c_req(username="admin", operation=2))
    assert (d["klass"], d["reason"]) == (P.ALLOW, 12)
    d = oracle.policy_mutate_strings(CFG, **_c_req(operation=2))
    assert (d["klass"], d["reason"]) == (P.DENY, 3)


def test_edge_missing_username_and_connect(oracle):
    d = oracle.policy_mutate_strings(CFG, **_c_req(username=None))
    assert (d["klass"], d["reason"], d["message"]) == (P.INVALID, 1, "cannot get requester's username from request")
    d = oracle.policy_mutate_strings(CFG, **_c_req(operation=3))
    assert (d["klass"], d["reason"], d["message"]) == (P.INVALID, 5, "invalid operation")


def test_message_literals_match_python_table(oracle):
    for reason, (_, msg, _line) in P.REASONS.items():
        got = oracle.lib().policy_ref_message(reason)
        if reason == 8:
            continue
        assert (None if got is None else got.decode()) == msg
