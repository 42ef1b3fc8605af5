"""CPU-only: the C-ABI library loads and exports every symbol of include/cbp_capi.h
(no compute calls without a GPU), and the product package never touches oracle/."""
import ctypes
import os
import re

from cbc_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    header = open(os.path.join(ROOT, "include", "cbp_capi.h")).read()
    declared = set(re.findall(r"CBP_API\s+[\w\s\*]+?\b(cbp_\w+)\s*\(", header))
    assert declared == set(capi.SYMBOLS), declared ^ set(capi.SYMBOLS)
    lib = ctypes.CDLL(capi.LIB_PATH)
    for s in declared:
        assert hasattr(lib, s), s


def test_product_code_never_reaches_the_oracle():
    for base, _, files in os.walk(os.path.join(ROOT, "cbc_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".hpp", ".h")):
                src = open(os.path.join(base, f), errors="ignore").read()
                assert "oracle_py" not in src and "cbp_oracle" not in src, f
    for f in os.listdir(os.path.join(ROOT, "include")):
        assert "cbp_oracle" not in open(os.path.join(ROOT, "include", f)).read()
