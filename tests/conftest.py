"""pytest configuration: registers the `gpu` marker and loads the package + checkers.

`-m "not gpu"` covers the oracle against the reference build / golden vectors, the host logic
and the C-ABI symbol table; `-m gpu` holds the parity tests proper (through the C ABI and
through the reference's own test-backend-ops harness against the plugin)."""
import ctypes as C
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as graft  # noqa: E402

graft.load_package()
REF_DIR = os.path.join(ROOT, "oracle", "_ref")
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def orc():
    """The C restatement (oracle/liboracle.so), built on demand with gcc."""
    so = os.path.join(ROOT, "oracle", "liboracle.so")
    srcs = [os.path.join(ROOT, "oracle", f) for f in ("oracle_quants.c", "oracle_ops.c", "oracle.h")]
    if not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
        subprocess.check_call(["make", "liboracle.so"], cwd=os.path.join(ROOT, "oracle"))
    L = C.CDLL(so)
    L.orc_nmse.restype = C.c_double
    L.orc_fp16_to_fp32.restype = C.c_float
    L.orc_fp16_to_fp32.argtypes = [C.c_uint16]
    L.orc_fp32_to_fp16.restype = C.c_uint16
    L.orc_fp32_to_fp16.argtypes = [C.c_float]
    L.orc_row_size.restype = C.c_size_t
    L.orc_row_size.argtypes = [C.c_int, C.c_int64]
    return L


@pytest.fixture(scope="session")
def refbase():
    """The reference's own libggml-base.so (oracle/_ref), or skip when it was not built."""
    so = os.path.join(REF_DIR, "libggml-base.so")
    if not os.path.exists(so):
        pytest.skip("oracle/_ref not built (needs /root/reference at build time)")
    L = C.CDLL(so)
    L.ggml_quantize_chunk.restype = C.c_size_t
    L.ggml_quantize_chunk.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_void_p]
    L.ggml_row_size.restype = C.c_size_t
    L.ggml_row_size.argtypes = [C.c_int, C.c_int64]
    return L


def ptr(a):
    return a.ctypes.data_as(C.c_void_p)
