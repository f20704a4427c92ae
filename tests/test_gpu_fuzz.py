"""A short run of the randomised GPU-vs-oracle soak test (tools/fuzz_gpu.py): random R (1..8), namespace
counts, batch sizes, value ranges (incl. full-range and negative), K1 variants and launch shapes, grouped
hints, masks and misaligned device columns; K0/K1/K2 (u8 and u16)/K3 all compared bit for bit."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
def test_fuzz_15_seconds():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "fuzz_gpu.py"), "--seconds", "15", "--seed", "7"],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "fuzz ok" in r.stdout
