"""The C-ABI library loads and exports every symbol include/pt_b200.h declares (no compute)."""
import ctypes as C
import re
from pathlib import Path

import pytest

from pathtracing_b200 import capi

ROOT = Path(__file__).resolve().parent.parent


def declared_symbols():
    text = (ROOT / "include" / "pt_b200.h").read_text()
    return sorted(set(re.findall(r"PT_API\s+[\w\s\*]+?\b(pt_\w+)\s*\(", text)))


def test_header_and_binding_agree():
    assert declared_symbols() == sorted(capi.EXPORTS)


def test_library_exports_every_declared_symbol():
    lib = capi.load_library()
    for name in declared_symbols():
        assert hasattr(lib, name), name


def test_struct_sizes_match_header():
    assert C.sizeof(capi.pt_trace_stats) == 48
    assert capi.RAY_DTYPE.itemsize == 32 and capi.HIT_DTYPE.itemsize == 16
    assert C.sizeof(capi.pt_render_stats) == 48


def test_no_device_is_a_loud_error():
    lib = capi.load_library()
    if lib.pt_device_count() > 0:
        pytest.skip("a CUDA device is present")
    assert not lib.pt_scene_create(0)
    assert b"no CPU fallback" in lib.pt_last_error()
