"""Regenerates tests/golden/policy_truth_table.csv — the exhaustive 4 ops × 2^10 flags truth table
of the batched form of admission.rs::mutate.

The reference (Rust, no tests, not buildable here) cannot produce these rows, so the table is
produced by oracle/policy_ref.py::decide_bits and each outcome carries the reference line that
returns it; tests/test_policy_oracle.py then checks the C oracle, the dict-level Python oracle and
the CUDA kernel against the committed file.  Run:  python tests/golden/make_policy_table.py
"""
import csv
import os
import sys

sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", ".."))
from oracle import policy_ref as P  # noqa: E402

out = os.path.join(os.path.dirname(__file__), "policy_truth_table.csv")
with open(out, "w", newline="") as f:
    w = csv.writer(f, lineterminator="\n")
    w.writerow(["op", "flags", "code", "patch_mask", "class", "reason", "reference_line"])
    for op in range(4):
        for flags in range(1024):
            code, mask = P.decide_bits(op, flags)
            w.writerow([op, flags, code, mask, P.CLASS_NAMES[code >> 4], code & 15, P.REASONS[code & 15][2]])
print("wrote", out)
