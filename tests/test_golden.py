"""Golden vectors generated from the reference itself (tests/golden/make_golden.py -> quant_golden.npz: ggml_quantize_chunk,
quantize_row_q8_1_ref, dequantize_row_* of /root/reference/ggml/src/ggml-quants.c on seeded rows).
The C restatement (oracle/) reproduces them bit for bit without oracle/_ref being present; tests/test_gpu_golden.py holds the
CUDA kernels against the same vectors."""
import ctypes as C
import os

import numpy as np
import pytest

from conftest import GOLDEN, ptr

TYPES = {"q4_0": 2, "q8_0": 8, "q4_K": 12, "q6_K": 14, "mxfp4": 39}


@pytest.fixture(scope="module")
def gold():
    return np.load(os.path.join(GOLDEN, "quant_golden.npz"))


@pytest.mark.parametrize("name", ["q4_0", "q8_0", "mxfp4"])
def test_oracle_quantizers_match_golden(orc, gold, name):
    x = gold["x"]
    for r in range(x.shape[0]):
        want = gold[f"{name}_bytes"][r]
        got = np.zeros_like(want)
        getattr(orc, f"orc_quantize_row_{name}")(ptr(np.ascontiguousarray(x[r])), ptr(got), C.c_int64(x.shape[1]))
        assert np.array_equal(got, want), f"row {r}"


def test_oracle_q8_1_matches_golden(orc, gold):
    x = gold["x"]
    for r in range(x.shape[0]):
        want = gold["q8_1_bytes"][r]
        got = np.zeros_like(want)
        orc.orc_quantize_row_q8_1(ptr(np.ascontiguousarray(x[r])), ptr(got), C.c_int64(x.shape[1]))
        assert np.array_equal(got, want), f"row {r}"


@pytest.mark.parametrize("name", list(TYPES))
def test_oracle_dequantizers_match_golden(orc, gold, name):
    q, d = gold[f"{name}_bytes"], gold[f"{name}_deq"]
    for r in range(q.shape[0]):
        got = np.zeros(d.shape[1], np.float32)
        assert orc.orc_dequantize_row(TYPES[name], ptr(np.ascontiguousarray(q[r])), ptr(got), C.c_int64(d.shape[1])) == 0
        assert np.array_equal(got.view(np.uint32), d[r].view(np.uint32)), f"row {r}"


