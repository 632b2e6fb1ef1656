"""The C-ABI library loads and exports every symbol include/vvb200.h declares (no compute calls:
this runs without a GPU), and the ctypes table in vibevoice_b200/_lib.py covers the same set."""
import ctypes
import os
import re

import pytest

from vibevoice_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "vvb200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"VV_API\s+[\w\s\*]+?\b(vv_\w+)\s*\(", src)))


def test_header_declares_the_expected_entry_points():
    names = _declared()
    for must in ("vv_head_create", "vv_head_set_schedule", "vv_head_forward", "vv_head_sample", "vv_dec_create",
                 "vv_dec_decode", "vv_dec_step", "vv_cache_create", "vv_cache_zero", "vv_last_error", "vv_version"):
        assert must in names
    assert len(names) >= 20


def test_ctypes_table_matches_header():
    assert sorted(_lib.SIGNATURES) == _declared()


def test_library_exports_every_declared_symbol():
    if not os.path.exists(_lib.LIB_PATH):
        pytest.fail(f"{_lib.LIB_PATH} missing: run `python -c 'import __graft_entry__ as g; g.build()'`")
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for name in _declared():
        assert hasattr(lib, name), name
    lib.vv_version.restype = ctypes.c_int
    assert lib.vv_version() > 0
    lib.vv_last_error.restype = ctypes.c_char_p
    assert isinstance(lib.vv_last_error(), bytes)


def test_no_torch_types_in_the_abi():
    src = open(os.path.join(ROOT, "include", "vvb200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)     # declarations only
    assert "torch" not in src.lower() and "at::" not in src and "Tensor" not in src


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "vibevoice_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle", txt, flags=re.M), f
