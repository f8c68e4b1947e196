"""The C-ABI library must load on a CPU-only box and export every symbol include/b2_api.h declares."""
import ctypes
import os
import re

import pytest

import b200_sensor as bs

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_in_header():
    src = open(os.path.join(ROOT, "include", "b2_api.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(b2_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    bs.build_native()
    lib = ctypes.CDLL(bs._ffi.LIB_PATH)
    names = _declared_in_header()
    assert len(names) >= 25
    for n in names:
        assert hasattr(lib, n), f"libb2sensor.so does not export {n}"


def test_python_prototypes_cover_header():
    assert _declared_in_header() == bs._ffi.declared_symbols()


def test_fails_loudly_without_gpu():
    import torch

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    lib = bs._ffi.lib()
    h = ctypes.c_void_p()
    rc = lib.b2_ctx_create(0, ctypes.byref(h))
    assert rc != 0 and lib.b2_last_error()
    with pytest.raises(bs._ffi.B2Error):
        bs._ffi.check(rc, "b2_ctx_create")


def test_struct_sizes_match_header():
    # sizes implied by include/b2_api.h (ints are 4 bytes, no padding surprises)
    f = bs._ffi
    assert ctypes.sizeof(f.AeDesc) == 4 + 9 * 4 + 8 * 4 + 4
    assert ctypes.sizeof(f.Norm) == 4 + 3 * 64 * 4 + 4   # + order
    assert ctypes.sizeof(f.CsvSpec) == 4 + 3 * 64 * 4 + 8
    assert ctypes.sizeof(f.AvroProg) == 8 + 64 * 4
    assert ctypes.sizeof(f.Adam) == 24
    assert ctypes.sizeof(f.Partition) == 40
