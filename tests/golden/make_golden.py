#!/usr/bin/env python3
"""Generates tests/golden/quant_golden.npz from the REFERENCE itself (oracle/_ref/libggml-base.so, built from
/root/reference by oracle/Makefile): seeded f32 rows, their reference-quantised blocks (ggml_quantize_chunk for
q4_0 / q8_0 / q4_K / q6_K / mxfp4, quantize_row_q8_1_ref for the q8_1 activation blocks; ggml/src/ggml-quants.c) and the
reference's dequantised values (dequantize_row_*).  The fixture is small (a few hundred KB) and committed, so the oracle and
the CUDA kernels stay pinned to the reference's bits even where oracle/_ref is not present.
usage: python tests/golden/make_golden.py      (in the build container, after `make -C oracle ref`)"""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
TYPES = {"q4_0": 2, "q8_0": 8, "q4_K": 12, "q6_K": 14, "mxfp4": 39}
K, ROWS = 1024, 6


def ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def rows(seed):
    rng = np.random.default_rng(seed)
    x = np.zeros((ROWS, K), np.float32)
    x[0] = rng.standard_normal(K)
    x[1] = rng.uniform(-1, 1, K)
    x[2] = rng.integers(-254, 255, K) / 2.0                       # exact .5 quotients and repeated maxima
    x[3, K // 2:] = rng.standard_normal(K // 2) * 1e-3            # half a row of zeros
    x[4] = np.clip(rng.standard_normal(K) * np.exp(rng.uniform(-12, 12, K)), -1e20, 1e20)
    x[5] = rng.standard_normal(K) * 0.02                          # weight-like
    return x.astype(np.float32)


def main():
    L = C.CDLL(os.path.join(ROOT, "oracle", "_ref", "libggml-base.so"))
    L.ggml_quantize_chunk.restype = C.c_size_t
    L.ggml_quantize_chunk.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_void_p]
    L.ggml_row_size.restype = C.c_size_t
    L.ggml_row_size.argtypes = [C.c_int, C.c_int64]
    x = rows(20260101)
    out = {"x": x}
    for name, t in TYPES.items():
        rb = L.ggml_row_size(t, K)
        q = np.zeros((ROWS, rb), np.uint8)
        assert L.ggml_quantize_chunk(t, ptr(x), ptr(q), 0, ROWS, K, None) == rb * ROWS
        d = np.zeros((ROWS, K), np.float32)
        for r in range(ROWS):
            getattr(L, f"dequantize_row_{name}")(ptr(q[r]), ptr(d[r]), C.c_int64(K))
        out[f"{name}_bytes"], out[f"{name}_deq"] = q, d
    q81 = np.zeros((ROWS, K // 32 * 36), np.uint8)
    for r in range(ROWS):
        L.quantize_row_q8_1_ref(ptr(x[r]), ptr(q81[r]), C.c_int64(K))
    out["q8_1_bytes"] = q81
    np.savez_compressed(os.path.join(HERE, "quant_golden.npz"), **out)
    print("wrote", os.path.join(HERE, "quant_golden.npz"), {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
