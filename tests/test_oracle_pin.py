"""Pins the C restatement (oracle/) against the reference itself (oracle/_ref/libggml-base.so,
built from /root/reference by oracle/Makefile): bit-exact for every quantizer/dequantizer.
Runs on CPU; on the GPU box oracle/_ref travels with the snapshot so it runs there too."""
import ctypes as C

import numpy as np
import pytest

from conftest import ptr

TYPES = {"q4_0": 2, "q8_0": 8, "q8_1": 9, "q4_K": 12, "q6_K": 14, "mxfp4": 39}


def _data(seed, n, kind):
    rng = np.random.default_rng(seed)
    if kind == "normal":
        return rng.standard_normal(n).astype(np.float32)
    if kind == "uniform":
        return rng.uniform(-1, 1, n).astype(np.float32)
    if kind == "ties":  # many exact .5 quotients and repeated maxima
        return (rng.integers(-254, 255, n) / 2.0).astype(np.float32)
    if kind == "zeros":
        x = np.zeros(n, np.float32)
        x[n // 2:] = rng.standard_normal(n - n // 2).astype(np.float32) * 1e-30
        return x
    if kind == "wide":
        return (rng.standard_normal(n) * np.exp(rng.uniform(-20, 20, n))).astype(np.float32)
    raise ValueError(kind)


def test_fp16_roundtrip_exhaustive(orc):
    h = np.arange(65536, dtype=np.uint16)
    f = h.view(np.float16).astype(np.float32)
    ok = ~np.isnan(f)
    got = np.array([orc.orc_fp16_to_fp32(int(v)) for v in h[ok]], np.float32)
    assert np.array_equal(got.view(np.uint32), f[ok].view(np.uint32))
    back = np.array([orc.orc_fp32_to_fp16(float(v)) for v in f[ok]], np.uint16)
    assert np.array_equal(back, h[ok])


def test_fp32_to_fp16_rounding(orc):
    rng = np.random.default_rng(1)
    x = np.concatenate([rng.standard_normal(20000) * np.exp(rng.uniform(-30, 15, 20000)),
                        [65504.0, 65519.9, 65520.0, 1e9, 2.0 ** -24, 2.0 ** -25, 2.0 ** -25 * 1.0001, 6.1e-5, 0.0, -0.0]]).astype(np.float32)
    want = x.astype(np.float16).view(np.uint16)
    got = np.array([orc.orc_fp32_to_fp16(float(v)) for v in x], np.uint16)
    assert np.array_equal(got, want)


@pytest.mark.parametrize("kind", ["normal", "uniform", "ties", "zeros", "wide"])
@pytest.mark.parametrize("name", ["q4_0", "q8_0", "q8_1", "mxfp4"])
def test_quantize_row_bit_exact(orc, refbase, name, kind):
    k = 32 * 257
    x = _data(hash((name, kind)) % 2 ** 31, k, kind)
    if name == "mxfp4" and kind == "wide":
        x = np.clip(x, -1e30, 1e30)
    nbytes = orc.orc_row_size(TYPES[name], k)
    a = np.zeros(nbytes, np.uint8)
    b = np.zeros(nbytes, np.uint8)
    getattr(orc, f"orc_quantize_row_{name}")(ptr(x), ptr(a), C.c_int64(k))
    getattr(refbase, f"quantize_row_{name}_ref")(ptr(x), ptr(b), C.c_int64(k))
    assert np.array_equal(a, b)


@pytest.mark.parametrize("name", ["q4_0", "q8_0", "mxfp4", "q4_K", "q6_K"])
def test_dequantize_row_bit_exact(orc, refbase, name):
    k = 256 * 33
    t = TYPES[name]
    x = _data(7 + t, k, "normal")
    nbytes = refbase.ggml_row_size(t, k)
    assert nbytes == orc.orc_row_size(t, k)
    q = np.zeros(nbytes, np.uint8)
    assert refbase.ggml_quantize_chunk(t, ptr(x), ptr(q), 0, 1, k, None) == nbytes
    a = np.zeros(k, np.float32)
    b = np.zeros(k, np.float32)
    assert orc.orc_dequantize_row(t, ptr(q), ptr(a), C.c_int64(k)) == 0
    getattr(refbase, f"dequantize_row_{name}")(ptr(q), ptr(b), C.c_int64(k))
    assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
    # and on arbitrary bytes (every bit pattern of scales/quants the kernels may meet)
    rng = np.random.default_rng(t)
    q2 = rng.integers(0, 256, nbytes, dtype=np.uint8)
    if name in ("q4_0", "q8_0"):   # keep f16 scales finite
        q2.reshape(-1, nbytes // (k // 32))[:, 1] &= 0x3F
    if name == "q4_K":
        q2.reshape(-1, 144)[:, [1, 3]] &= 0x3F
    if name == "q6_K":
        q2.reshape(-1, 210)[:, 209] &= 0x3F
    orc.orc_dequantize_row(t, ptr(q2), ptr(a), C.c_int64(k))
    getattr(refbase, f"dequantize_row_{name}")(ptr(q2), ptr(b), C.c_int64(k))
    assert np.array_equal(a.view(np.uint32), b.view(np.uint32))


@pytest.mark.parametrize("name", ["q4_0", "q8_0", "mxfp4", "q4_K", "q6_K"])
def test_mul_mat_oracles_agree(orc, refbase, name):
    """q8_1-activation oracle vs f64 ground truth stays far inside the reference's 5e-4 NMSE gate."""
    t = TYPES[name]
    M, K, N = 48, 512, 3
    rng = np.random.default_rng(11 + t)
    wf = rng.standard_normal(M * K).astype(np.float32)
    rb = refbase.ggml_row_size(t, K)
    w = np.zeros(rb * M, np.uint8)
    refbase.ggml_quantize_chunk(t, ptr(wf), ptr(w), 0, M, K, None)
    x = rng.uniform(-1, 1, N * K).astype(np.float32)
    a = np.zeros(M * N, np.float32)
    b = np.zeros(M * N, np.float32)
    assert orc.orc_mul_mat_f64(t, ptr(w), C.c_int64(rb), C.c_int64(M), C.c_int64(K), ptr(x), C.c_int64(N), ptr(a)) == 0
    assert orc.orc_mul_mat_q8_1(t, ptr(w), C.c_int64(rb), C.c_int64(M), C.c_int64(K), ptr(x), C.c_int64(N), ptr(b)) == 0
    assert orc.orc_nmse(ptr(a), ptr(b), C.c_size_t(M * N)) < 5e-5
