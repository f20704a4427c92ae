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
