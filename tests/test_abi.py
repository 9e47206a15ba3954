"""CPU: the C-ABI library loads and exports every symbol include/b200llm.h declares, the ctypes
table covers the header, and argument errors are reported without touching a GPU."""
import ctypes
import os

import pytest

from building_llm_from_scratch_b200 import _lib


@pytest.fixture(scope="module")
def lib():
    if not os.path.exists(_lib.LIB_PATH):
        import __graft_entry__
        __graft_entry__.build()
    return _lib.load()


def test_header_symbols_exported(lib):
    names = _lib.header_symbols()
    assert len(names) >= 30
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/b200llm.h but not exported"


def test_ctypes_table_matches_header():
    assert sorted(_lib.SIGNATURES) == _lib.header_symbols()


def test_version_and_error_string(lib):
    assert lib.b200_version() == 100
    # argument validation happens before any CUDA call
    rc = lib.b200_rmsnorm_fwd_bf16(None, None, None, None, None, 4, 64, 1e-5, 1, None)
    assert rc == -1
    assert b"rmsnorm_fwd" in lib.b200_last_error()
    rc = lib.b200_attention_fwd_bf16(ctypes.c_void_p(16), 64, ctypes.c_void_p(16), 64, ctypes.c_void_p(16), 64,
                                     ctypes.c_void_p(16), 64, ctypes.c_void_p(16), 1, 8, 1, 1, 48, 1.0, 0.0, 0, None, None)
    assert rc == -1 and b"head_dim" in lib.b200_last_error()


def test_missing_library_fails_loudly(monkeypatch):
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", "/nonexistent/libb200llm.so")
    with pytest.raises(RuntimeError, match="no CPU or library fallback"):
        _lib.load()
