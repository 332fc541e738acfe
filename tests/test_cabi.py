"""The drop-in boundary: libb200cs.so loads without a GPU and exports every function that
include/b200cs.h declares; argument errors come back as status codes with a message (no
exceptions, no exit) -- SURVEY.md 8b.  No compute call is made here."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, "include", "b200cs.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b200cs_[a-z0-9_]+)\s*\(", text)))


@pytest.fixture(scope="module")
def lib():
    from fv3net_b200 import _lib

    if not os.path.exists(_lib.LIB_PATH):
        from fv3net_b200.csrc import build

        build.build()
    return _lib.load()


def test_header_symbols_are_exported(lib):
    names = _declared()
    assert len(names) >= 25
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, f"declared in include/b200cs.h but not exported: {missing}"


def test_python_binding_covers_the_header():
    from fv3net_b200 import _lib

    assert sorted(_lib.EXPORTS) == _declared()


def test_status_and_error_text_without_a_gpu(lib):
    assert lib.b200cs_version() == 100
    # NULL arguments are rejected before any CUDA call
    rc = lib.b200cs_block_upsample(None, None, 4, 1, 2, 2, 2, 2, 0, 0, None)
    assert rc == -1 and b"upsample" in lib.b200cs_last_error()
    rc = lib.b200cs_blend(None, None, None, None, 0, 4, None)
    assert rc == -1 and b"blend" in lib.b200cs_last_error()
    rc = lib.b200cs_elementwise(99, None, None, None, None, 0, 4, 0.0, None)
    assert rc == -1 and b"unknown op" in lib.b200cs_last_error()
    assert lib.b200cs_pressure_coarsen_wide_workspace_bytes(1, 79, 3072, 3072, 64) == 4 * (79 * 48 * 48 + 384 * 96 * 79 * 5)
    assert lib.b200cs_pressure_coarsen_gather_workspace_bytes(6, 384) == 4 * 6 * 48
    assert lib.b200cs_mappm_workspace_bytes(1000, 79) > 0


def test_missing_library_fails_loudly(monkeypatch):
    from fv3net_b200 import _lib

    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", os.path.join(ROOT, "fv3net_b200", "no_such_lib.so"))
    with pytest.raises(_lib.B200csError, match="no CPU fallback"):
        _lib.load()
