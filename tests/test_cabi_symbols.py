"""The C-ABI library builds, loads without a GPU and exports every symbol that
include/gaudimm_b200.h declares (no compute calls here)."""
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "gaudimm_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(gmm_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_exported(built_lib):
    from gaudi_b200 import _lib
    names = declared_symbols()
    assert len(names) >= 15
    for n in names:
        assert hasattr(built_lib, n), "libgaudimm_b200.so does not export %s" % n
        assert n in _lib.SIGNATURES, "ctypes binding misses %s" % n
    assert set(_lib.SIGNATURES) == set(names)


def test_error_reporting_without_gpu(built_lib):
    import ctypes as C
    # NULL context is rejected with a message, never dereferenced
    rc = built_lib.gmm_set_params(None, 5.0, 0.6, -0.4, 0.2, 2.1, 0.0, 1.0, 0)
    assert rc == -1
    assert b"ctx is NULL" in built_lib.gmm_last_error()
    assert built_lib.gmm_version() >= 100
    assert built_lib.gmm_destroy(None) == 0
