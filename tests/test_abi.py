"""The C-ABI library loads and exports every symbol include/stevi_b200/c_api.h declares (no compute without a GPU)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "stevi_b200", "c_api.h")
LIB = os.path.join(ROOT, "libstevi_b200", "libstevi_b200.so")


def declared_symbols():
    txt = open(HEADER).read()
    return sorted(set(re.findall(r"STEVI_API\s+[\w\s\*]+?\b(stevi_\w+)\s*\(", txt)))


@pytest.fixture(scope="module")
def lib():
    if not os.path.exists(LIB):
        import __graft_entry__ as g
        g.build()
    return ctypes.CDLL(LIB)


def test_header_declares_the_hot_path():
    names = declared_symbols()
    for needed in ("stevi_ctx_create", "stevi_census_transform", "stevi_cost_volume", "stevi_sgm", "stevi_wta",
                   "stevi_index_to_disp", "stevi_census_sgm_pipeline", "stevi_cost_volume_2d", "stevi_wta_2d", "stevi_index2d_to_disp",
                   "stevi_truncated_cost_volume_2d", "stevi_refine_disp_2d_cost_interp", "stevi_refine_disp_2d_cost_patch_interp",
                   "stevi_variable_cost_volume_aggregation"):
        assert needed in names


def test_every_declared_symbol_is_exported(lib):
    missing = [n for n in declared_symbols() if not hasattr(lib, n)]
    assert not missing, missing


def test_no_cpu_fallback(lib):
    """Without a CUDA device context creation fails loudly; with one it succeeds."""
    lib.stevi_version.restype = ctypes.c_char_p
    assert b"sm_100a" in lib.stevi_version()
    h = ctypes.c_void_p()
    rc = lib.stevi_ctx_create(0, None, ctypes.byref(h))
    if lib.stevi_device_count() == 0:
        assert rc == 4 and not h  # STEVI_E_CUDA
    else:
        assert rc == 0
        lib.stevi_ctx_destroy(h)


def test_product_never_touches_the_oracle():
    """libstevi_b200/ (python and csrc) must not import, link or load anything under oracle/."""
    pkg = os.path.join(ROOT, "libstevi_b200")
    for dirpath, _, files in os.walk(pkg):
        if "build" in dirpath or "__pycache__" in dirpath:
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                txt = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "import oracle" not in txt and "libstevi_oracle" not in txt and "libstevi_ref" not in txt, f
