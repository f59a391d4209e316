"""CPU-side checks of the drop-in boundary: libpmps.so loads, exports every symbol include/pmps.h declares, and the
ctypes mirrors of the ABI structs have the C layout.  No compute call is made (there is no GPU here)."""
import ctypes
import os
import re

import pytest

from common import ROOT
from polymps_b200 import engine
from polymps_b200.params import PmpsParams
from polymps_b200.state import PmpsHostArrays, PmpsScalars


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "pmps.h")).read()
    return sorted(set(re.findall(r"\b(pmps_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    lib = engine.load_library()
    syms = _declared_symbols()
    assert len(syms) >= 45
    missing = [s for s in syms if not hasattr(lib, s)]
    assert not missing, missing
    assert lib.pmps_abi_version() == 5


def test_struct_layouts_match_header():
    # sizes computed from the header: 30 int32 + (30 + 3 + 20 + 15) doubles; 2 int32 + 53 pointers
    assert ctypes.sizeof(PmpsParams) == 30 * 4 + 76 * 8
    assert ctypes.sizeof(PmpsHostArrays) == 8 + 8 * (7 + 1 + 18 + 15 + 13)
    assert ctypes.sizeof(PmpsScalars) == 56


def test_create_fails_loudly_without_gpu_or_with_bad_params():
    lib = engine.load_library()
    p = PmpsParams(); h = ctypes.c_void_p()
    rc = lib.pmps_create(ctypes.byref(p), 0, ctypes.byref(h))   # abi_version 0: rejected before any CUDA call
    assert rc < 0 and not h
    assert b"ABI" in lib.pmps_last_error(None)


def test_no_product_code_imports_the_oracle():
    bad = []
    for dp, dn, fn in os.walk(os.path.join(ROOT, "polymps_b200")):
        for f in fn:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".inc", ".cpp")):
                t = open(os.path.join(dp, f), errors="ignore").read()
                if re.search(r"pyoracle|mps_oracle|libmpsoracle|oracle/", t):
                    bad.append(os.path.join(dp, f))
    assert not bad, bad
