"""The C-ABI library loads without a GPU and exports every function include/ccb.h declares; argument validation
that needs no device works; the Python binding refuses to run without the library (no CPU fallback)."""

import ctypes as C
import os
import re

import pytest

from common import ROOT
from coolchic_b200 import _native as N

HEADER = os.path.join(ROOT, "include", "ccb.h")


def declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    src = re.sub(r"//[^\n]*", "", src)
    return sorted(set(re.findall(r"\b(ccb_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_and_library_exports_the_same_set():
    names = declared_functions()
    assert len(names) >= 15
    lib = N.lib()
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/ccb.h but not exported by libccb.so"
    assert sorted(N.EXPORTS) == names, (set(N.EXPORTS) ^ set(names))


def test_header_cites_the_reference():
    src = open(HEADER).read()
    for cite in ("coolchic.py", "quantizer.py", "arm.py", "upsampling.py", "synthesis.py", "loss.py", "train.py",
                 "arm_cpu.cpp", "cc-contexts.h"):
        assert cite in src, cite
    assert "extern \"C\"" in src
    code = re.sub(r"/\*.*?\*/", "", src, flags=re.S)  # no torch types in the signatures
    assert "torch" not in code.lower() and "at::" not in code and "Tensor" not in code


def test_strerror_and_version():
    lib = N.lib()
    assert lib.ccb_version() > 0
    assert lib.ccb_strerror(0).decode().lower().startswith("ok") or lib.ccb_strerror(0).decode() != ""
    for rc in (-1, -2, -3, -4):
        assert lib.ccb_strerror(rc).decode() != ""


def test_plan_validation_needs_no_device_for_bad_arguments():
    lib = N.lib()
    h = C.c_void_p()
    bad = N.make_arch(1, 32, 32, 7, 12, 2, ["8-1-linear-relu", "3-1-linear-none"], 8, 7)  # dim_arm not in {8,16,24,32}
    rc = lib.ccb_plan_create(C.byref(h), C.byref(bad), 0)
    assert rc != 0 and not h.value
    bad2 = N.make_arch(1, 32, 32, 7, 16, 2, ["8-3-linear-relu", "3-1-linear-none"], 8, 7)  # first layer must be 1x1
    assert lib.ccb_plan_create(C.byref(h), C.byref(bad2), 0) != 0
    assert lib.ccb_plan_create(None, C.byref(bad), 0) != 0


def test_missing_library_raises(monkeypatch):
    monkeypatch.setattr(N, "_lib", None)
    monkeypatch.setattr(N, "LIB_PATH", os.path.join(ROOT, "does", "not", "exist.so"))
    with pytest.raises(N.CcbError):
        N.lib()


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "coolchic_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".sh")):
                s = open(os.path.join(dp, f)).read()
                assert "oracle" not in s.replace("# oracle", ""), os.path.join(dp, f)


def test_parse_synthesis_layers():
    assert N.parse_synthesis_layers(["48-1-linear-relu", "X-1-linear-none", "X-3-residual-none"]) == \
        [(48, 1, 0, 1), (3, 1, 0, 0), (3, 3, 1, 0)]
    with pytest.raises(ValueError):
        N.parse_synthesis_layers(["48-1-conv-relu"])
