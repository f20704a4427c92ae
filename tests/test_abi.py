"""CPU tests: libbgc.so loads and exports every symbol include/*.h declares (no compute calls)."""
import ctypes
import glob
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    names = set()
    for h in glob.glob(os.path.join(ROOT, "include", "*.h")):
        text = re.sub(r"/\*.*?\*/", "", open(h).read(), flags=re.S)
        names |= set(re.findall(r"\b(bgc_[a-z0-9_]+)\s*\(", text))
    return names


def test_library_exports_every_declared_symbol(bgc):
    syms = declared_symbols()
    assert len(syms) >= 30
    for s in sorted(syms):
        assert hasattr(bgc.lib, s), f"libbgc.so does not export {s}"
    assert syms == set(bgc.EXPORTS), syms ^ set(bgc.EXPORTS)


def test_headers_cite_reference_lines():
    for h in ("bgc.h", "bgc_host.h"):
        text = open(os.path.join(ROOT, "include", h)).read()
        assert len(re.findall(r"src/(admission|synchronizer|controller)\.rs:\d+", text)) >= 10


def test_library_is_sm100a_only(bgc):
    import subprocess
    out = subprocess.run(["cuobjdump", "-lelf", bgc.LIB_PATH], capture_output=True, text=True).stdout
    assert "sm_100a" in out
    assert not re.search(r"sm_(?!100a)\d+", out)


def test_no_cpu_fallback_without_device(bgc):
    import torch
    if torch.cuda.is_available():
        return
    h = ctypes.c_void_p()
    assert bgc.lib.bgc_ctx_create(0, ctypes.byref(h)) == bgc.E_CUDA
    try:
        bgc.Engine(0)
    except bgc.BgcError as e:
        assert e.code == bgc.E_CUDA
    else:
        raise AssertionError("Engine() must fail without a CUDA device")


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "bacchus-gpu-controller_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")) or f == "Makefile":
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                assert not re.search(r"import\s+oracle|from\s+oracle|oracle/|liboracle|_ref\.(c|h|py)", text), \
                    f"{f} reaches into oracle/"


def test_headers_are_plain_c(tmp_path):
    """The boundary is a C ABI: every header must compile as C99 (not only as C++), and the C example links."""
    import subprocess
    src = tmp_path / "all_headers.c"
    src.write_text('#include "bgc.h"\n#include "bgc_host.h"\n#include "bgc_synth.h"\nint main(void) { return BGC_OK; }\n')
    r = subprocess.run(["/usr/bin/gcc", "-std=c99", "-Wall", "-Wextra", "-Werror", "-pedantic", "-fsyntax-only",
                        "-I", os.path.join(ROOT, "include"), str(src)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    exe = tmp_path / "mutate_batch"
    r = subprocess.run(["/usr/bin/gcc", "-std=c99", "-Wall", "-I", os.path.join(ROOT, "include"), os.path.join(ROOT, "examples", "mutate_batch.c"),
                        "-L", os.path.join(ROOT, "bacchus-gpu-controller_b200"), "-lbgc",
                        "-Wl,-rpath," + os.path.join(ROOT, "bacchus-gpu-controller_b200"), "-o", str(exe)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    import torch
    out = subprocess.run([str(exe)], capture_output=True, text=True)
    if torch.cuda.is_available():
        assert out.returncode == 0, out.stderr
        assert "request 0: allowed=1 class=1" in out.stdout and "user is not in authorized group" in out.stdout
        assert "verdicts = 0 1 1" in out.stdout
    else:
        assert out.returncode == 2          # no device: fails loudly, no CPU fallback


import pytest  # noqa: E402


@pytest.mark.gpu
def test_c_example_runs_on_the_device(tmp_path):
    import subprocess
    exe = tmp_path / "mutate_batch"
    r = subprocess.run(["/usr/bin/gcc", "-std=c99", "-I", os.path.join(ROOT, "include"), os.path.join(ROOT, "examples", "mutate_batch.c"),
                        "-L", os.path.join(ROOT, "bacchus-gpu-controller_b200"), "-lbgc",
                        "-Wl,-rpath," + os.path.join(ROOT, "bacchus-gpu-controller_b200"), "-o", str(exe)], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    out = subprocess.run([str(exe)], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    assert "request 0: allowed=1 class=1" in out.stdout and "user is not in authorized group" in out.stdout
    assert "request 2: allowed=1 class=0 reason=12" in out.stdout
    assert "used gpu=[1,1] cpu=[6000,1000]; verdicts = 0 1 1" in out.stdout
    assert "narrow columns: used gpu=[1,1] cpu=[6000,1000]; verdicts = 0 1 1" in out.stdout
    import json
    answers = [json.loads(line.split(" -> ", 1)[1])["response"] for line in out.stdout.splitlines() if line.startswith("review ")]
    assert [a["allowed"] for a in answers] == [True, False, False] and answers[0]["uid"] == "u-1" and "patch" in answers[0]
    assert answers[1]["status"]["message"] == "user is not in authorized group" and answers[2]["status"]["reason"] == "InvalidRequest"
