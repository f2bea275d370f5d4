"""Host logic of the plugin's row-split buffer type (csrc/backend/split_buffer.cpp): the partition arithmetic against a
restatement of the reference's get_row_split (ggml-cuda/ggml-cuda.cu:778-804, tensor_split normalisation :1078-1090) and
against the row split the multi-process decode chain uses (llama-mi50.cpp_b200/session.py:row_split)."""
import ctypes as C
import os

import numpy as np
import pytest

from conftest import REF_DIR, ROOT

PKG = os.path.join(ROOT, "llama-mi50.cpp_b200")


@pytest.fixture(scope="module")
def plugin():
    so = os.path.join(PKG, "libggml-b200.so")
    base = os.path.join(REF_DIR, "libggml-base.so")
    if not (os.path.exists(so) and os.path.exists(base)):
        pytest.skip("plugin / oracle/_ref not built (needs /root/reference at build time)")
    C.CDLL(base, mode=C.RTLD_GLOBAL)
    lib = C.CDLL(so)
    lib.ggml_backend_b200_split_rows.restype = C.c_int
    lib.ggml_backend_b200_split_rows.argtypes = [C.POINTER(C.c_float), C.c_int, C.c_longlong, C.c_int,
                                                 C.POINTER(C.c_longlong), C.POINTER(C.c_longlong)]
    return lib


def ref_row_split(tensor_split, n, nrows, dev, rounding=128):
    """get_row_split with the cumulative fractions of ggml_backend_cuda_split_buffer_type (float32 arithmetic as in the C code)."""
    ts = np.asarray(tensor_split if tensor_split is not None else [0.0] * n, np.float32)
    if not ts.any():
        cum = np.array([np.float32(i / n) for i in range(n)], np.float32)
    else:
        cum = np.zeros(n, np.float32)
        s = np.float32(0)
        for i in range(n):
            cum[i] = s
            s = np.float32(s + ts[i])
        cum = (cum / s).astype(np.float32)
    lo = 0 if dev == 0 else int(np.float32(nrows) * cum[dev])
    lo -= lo % rounding
    if dev == n - 1:
        hi = nrows
    else:
        hi = int(np.float32(nrows) * cum[dev + 1])
        hi -= hi % rounding
    return lo, max(hi, lo)


def call(lib, ts, n, nrows, dev):
    lo, hi = C.c_longlong(), C.c_longlong()
    arr = None if ts is None else (C.c_float * n)(*ts)
    assert lib.ggml_backend_b200_split_rows(arr, n, nrows, dev, C.byref(lo), C.byref(hi)) == 0
    return lo.value, hi.value


@pytest.mark.parametrize("n", [1, 2, 3, 4, 8])
@pytest.mark.parametrize("nrows", [0, 1, 127, 128, 1024, 4096, 14336, 28672, 128256, 152064])
@pytest.mark.parametrize("ts", [None, "skew", "zero_dev"])
def test_partition_matches_reference_formula(plugin, n, nrows, ts):
    split = None
    if ts == "skew":
        split = [float(i + 1) for i in range(n)]
    elif ts == "zero_dev":
        split = [0.0 if i == 1 else 1.0 for i in range(n)] if n > 1 else [1.0]
    covered = 0
    for dev in range(n):
        lo, hi = call(plugin, split, n, nrows, dev)
        assert (lo, hi) == ref_row_split(split, n, nrows, dev)
        assert lo == covered and hi >= lo            # slices tile [0, nrows) in device order: concatenation, never a sum
        assert lo % 128 == 0 and (hi % 128 == 0 or dev == n - 1)
        covered = hi
    assert covered == nrows


def test_equal_split_agrees_with_the_decode_chain_sessions(plugin):
    import __graft_entry__ as graft
    graft.load_package()
    from llama_mi50_b200.session import row_split
    for n in (2, 4, 8):
        for nrows in (1024, 4096, 14336, 128256):
            for dev in range(n):
                assert call(plugin, None, n, nrows, dev) == row_split(nrows, n, dev)


def test_bad_arguments_are_rejected(plugin):
    lo, hi = C.c_longlong(), C.c_longlong()
    f = plugin.ggml_backend_b200_split_rows
    assert f(None, 0, 10, 0, C.byref(lo), C.byref(hi)) == -1
    assert f(None, 2, 10, 2, C.byref(lo), C.byref(hi)) == -1
    assert f(None, 2, -1, 0, C.byref(lo), C.byref(hi)) == -1
    assert f(None, 2, 10, 0, None, C.byref(hi)) == -1
