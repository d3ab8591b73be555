"""The C-ABI library loads here (no GPU) and exports every symbol include/jubann.h declares."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_functions():
    text = open(os.path.join(ROOT, "include", "jubann.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    names = re.findall(r"\b(?:int|const char\*)\s+(version|jubann_\w+)\s*\(", text)
    return sorted(set(names))


def test_header_declares_expected_surface():
    names = declared_functions()
    for must in ("jubann_create", "jubann_set_rows_csr", "jubann_query_sig", "jubann_query_csr",
                 "jubann_query_rows", "jubann_scan_local_dev", "jubann_merge_dev", "version"):
        assert must in names


def test_library_exports_every_declared_symbol():
    from jubatus_b200 import capi
    path = capi.lib_path()
    assert os.path.exists(path), "build first: python -c 'import __graft_entry__ as g; g.build()'"
    lib = ctypes.CDLL(path)
    for name in declared_functions():
        assert hasattr(lib, name), "missing export: " + name
    assert set(declared_functions()) == set(capi.EXPORTS)


def test_version_and_errors_without_gpu():
    from jubatus_b200 import capi
    lib = capi.load_library()
    assert lib.jubann_abi_version() == 2
    assert b"jubann" in lib.version()
    h = ctypes.c_void_p()
    # argument validation happens before any CUDA call
    assert lib.jubann_create(b"nope", 64, 0, 0, 1, ctypes.byref(h)) == -1
    assert b"unsupported_method" in lib.jubann_last_error()
    assert lib.jubann_create(b"lsh", 0, 0, 0, 1, ctypes.byref(h)) == -1
    assert b"hash_num" in lib.jubann_last_error()
    assert lib.jubann_create(b"lsh", 64, 0, 2, 2, ctypes.byref(h)) == -1


def test_no_cpu_fallback_without_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from jubatus_b200 import Engine, JubannError
    with pytest.raises(JubannError) as ei:
        Engine("lsh", 64)
    assert "no CUDA device" in str(ei.value) or ei.value.code == -2
