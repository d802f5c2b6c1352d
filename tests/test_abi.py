"""CPU: the C-ABI library loads without a GPU and exports every symbol the header declares."""
import ctypes as C
import os
import re

from nnstreamer_b200 import _lib as L

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_functions():
    src = open(os.path.join(ROOT, "include", "nnstreamer_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"NNSB_API\s+[\w\s\*]+?\b(nnsb_\w+)\s*\(", src)))


def test_every_declared_symbol_is_exported_and_bound():
    names = header_functions()
    assert len(names) >= 50
    lib = C.CDLL(L.LIB_PATH)
    for n in names:
        assert hasattr(lib, n), "header declares %s but the library does not export it" % n
    assert sorted(L.declared_symbols()) == names, "ctypes prototypes and header differ"


def test_no_oracle_or_cpu_path_in_the_product():
    """The product must not reference oracle/ (a CPU fallback would void parity)."""
    pkg = os.path.join(ROOT, "nnstreamer_b200")
    for dirpath, _, files in os.walk(pkg):
        if "build" in dirpath.split(os.sep):
            continue
        for f in files:
            if f.endswith((".py", ".cpp", ".cu", ".cuh", ".h", ".c")):
                text = open(os.path.join(dirpath, f), errors="replace").read()
                assert "oracle" not in text.lower(), \
                    "%s mentions the oracle" % os.path.join(dirpath, f)


def test_compute_calls_fail_loudly_without_a_gpu():
    import pytest
    lib = L.lib()
    n = C.c_int(0)
    if lib.nnsb_device_count(C.byref(n)) == 0 and n.value > 0:
        pytest.skip("a GPU is visible here")
    ctx = C.c_void_p()
    assert lib.nnsb_ctx_create(0, C.byref(ctx)) != 0
    assert b"cuda" in lib.nnsb_last_error().lower()
