"""GPU parity through the C ABI: b200_quantize_q8_1 (bit-exact vs oracle) and b200_mul_mat_vec_q
(vs the oracle's q8_1-semantics product and the f64 ground truth with the reference's NMSE gate,
tests/test-backend-ops.cpp:3718-3720: 5e-4)."""
import ctypes as C

import numpy as np
import pytest

from conftest import ptr

pytestmark = pytest.mark.gpu

TYPES = {"q4_0": 2, "q8_0": 8, "q4_K": 12, "q6_K": 14, "mxfp4": 39}


def _mods():
    import torch
    from llama_mi50_b200 import capi, gguf_synth
    return torch, capi, gguf_synth


def _quant_weights(refbase, orc, gguf_synth, rng, t, K, M):
    """Properly quantized weights via the reference quantizer when available, else sampled blocks."""
    if refbase is not None:
        wf = (rng.standard_normal(M * K) * 0.5).astype(np.float32)
        rb = refbase.ggml_row_size(t, K)
        w = np.zeros(rb * M, np.uint8)
        refbase.ggml_quantize_chunk(t, ptr(wf), ptr(w), 0, M, K, None)
        return w
    return gguf_synth.make_tensor_bytes(rng, t, (K, M), sigma=0.5)


@pytest.mark.parametrize("shape", [(1, 32), (3, 96), (1, 4096), (5, 14336), (2, 100), (7, 8 * 33), (1, 1000)])
@pytest.mark.parametrize("kind", ["normal", "ties", "zeros"])
def test_quantize_q8_1_bit_exact(orc, shape, kind):
    torch, capi, _ = _mods()
    rows, K = shape
    rng = np.random.default_rng(rows * 131 + K)
    if kind == "normal":
        x = rng.standard_normal((rows, K)).astype(np.float32)
    elif kind == "ties":
        x = (rng.integers(-254, 255, (rows, K)) / 2.0).astype(np.float32)
    else:
        x = np.zeros((rows, K), np.float32)
        x[:, K // 2:] = rng.standard_normal((rows, K - K // 2)).astype(np.float32) * 1e-20
    Kp = -(-K // 512) * 512
    got = capi.quantize_q8_1(torch.from_numpy(x).cuda(), Kp).cpu().numpy()
    xp = np.zeros((rows, Kp), np.float32)
    xp[:, :K] = x
    want = np.zeros((rows, Kp // 32 * 36), np.uint8)
    for r in range(rows):
        orc.orc_quantize_row_q8_1(ptr(xp[r]), ptr(want[r]), C.c_int64(Kp))
    assert np.array_equal(got, want)


@pytest.mark.parametrize("name", list(TYPES))
@pytest.mark.parametrize("M,K,N", [(16, 256, 1), (16, 256, 8), (33, 512, 3), (257, 1024, 1), (64, 4096, 2), (5, 14336, 1), (1, 256, 5)])
def test_mmvq_vs_oracle(orc, request, name, M, K, N):
    torch, capi, gguf_synth = _mods()
    try:
        refbase = request.getfixturevalue("refbase")
    except pytest.skip.Exception:
        refbase = None
    t = TYPES[name]
    rng = np.random.default_rng(M * 7 + K + N)
    w = _quant_weights(refbase, orc, gguf_synth, rng, t, K, M)
    x = rng.uniform(-1, 1, (N, K)).astype(np.float32)
    got = capi.mul_mat_vec_q(torch.from_numpy(w).cuda(), t, [K, M, 1, 1], torch.from_numpy(x).cuda()).cpu().numpy().reshape(N, M)
    rb = gguf_synth.row_bytes(t, K)
    want = np.zeros((N, M), np.float32)
    truth = np.zeros((N, M), np.float32)
    assert orc.orc_mul_mat_q8_1(t, ptr(w), C.c_int64(rb), C.c_int64(M), C.c_int64(K), ptr(x), C.c_int64(N), ptr(want)) == 0
    assert orc.orc_mul_mat_f64(t, ptr(w), C.c_int64(rb), C.c_int64(M), C.c_int64(K), ptr(x), C.c_int64(N), ptr(truth)) == 0
    scale = np.abs(want).max() + 1e-30
    assert np.abs(got - want).max() <= 2e-5 * scale + 1e-6, "differs from the q8_1-semantics oracle beyond f32 summation order"
    assert orc.orc_nmse(ptr(truth), ptr(np.ascontiguousarray(got)), C.c_size_t(N * M)) < 5e-4


@pytest.mark.parametrize("name", list(TYPES))
def test_mmvq_batch_broadcast(orc, name):
    """ne02/ne03 batches with broadcast ratios (reference test cases bs=[..], nr=[..], test-backend-ops.cpp:7650-7700)."""
    torch, capi, gguf_synth = _mods()
    t = TYPES[name]
    K, M, N, ne02, ne03, r2, r3 = 256, 16, 2, 2, 2, 3, 2
    rng = np.random.default_rng(5)
    w = gguf_synth.make_tensor_bytes(rng, t, (K, M, ne02, ne03), sigma=0.5)
    x = rng.uniform(-1, 1, (ne03 * r3, ne02 * r2, N, K)).astype(np.float32)
    got = capi.mul_mat_vec_q(torch.from_numpy(w).cuda(), t, [K, M, ne02, ne03], torch.from_numpy(x).cuda()).cpu().numpy()
    rb = gguf_synth.row_bytes(t, K)
    for s in range(ne03 * r3):
        for c in range(ne02 * r2):
            wi = (s // r3) * ne02 + (c // r2)
            wm = w[wi * M * rb:(wi + 1) * M * rb]
            want = np.zeros((N, M), np.float32)
            orc.orc_mul_mat_q8_1(t, ptr(wm), C.c_int64(rb), C.c_int64(M), C.c_int64(K), ptr(np.ascontiguousarray(x[s, c])), C.c_int64(N), ptr(want))
            assert np.abs(got[s, c] - want).max() <= 2e-5 * np.abs(want).max() + 1e-6


@pytest.mark.parametrize("name", list(TYPES))
@pytest.mark.parametrize("n_tok,n_used,bcast", [(1, 4, False), (3, 2, True), (4, 4, False)])
def test_mmvq_ids(orc, name, n_tok, n_used, bcast):
    """MUL_MAT_ID: dst[:,e,t] = W[ids[e,t]] . x[:, e or 0, t] (reference ggml.c:3217-3242)."""
    torch, capi, gguf_synth = _mods()
    t = TYPES[name]
    K, M, n_exp = 256, 32, 8
    rng = np.random.default_rng(n_tok * 10 + n_used)
    w = gguf_synth.make_tensor_bytes(rng, t, (K, M, n_exp), sigma=0.5)
    ne11 = 1 if bcast else n_used
    x = rng.uniform(-1, 1, (n_tok, ne11, K)).astype(np.float32)
    ids = np.stack([rng.permutation(n_exp)[:n_used] for _ in range(n_tok)]).astype(np.int32)
    got = capi.mul_mat_vec_q(torch.from_numpy(w).cuda(), t, [K, M, n_exp, 1], torch.from_numpy(x).cuda(),
                             ids=torch.from_numpy(ids).cuda()).cpu().numpy()
    rb = gguf_synth.row_bytes(t, K)
    for tk in range(n_tok):
        for e in range(n_used):
            wm = w[ids[tk, e] * M * rb:(ids[tk, e] + 1) * M * rb]
            want = np.zeros(M, np.float32)
            orc.orc_mul_mat_q8_1(t, ptr(wm), C.c_int64(rb), C.c_int64(M), C.c_int64(K), ptr(np.ascontiguousarray(x[tk, e % ne11])), C.c_int64(1), ptr(want))
            assert np.abs(got[tk, e] - want).max() <= 2e-5 * np.abs(want).max() + 1e-6


@pytest.mark.parametrize("name", ["q4_0", "q4_K", "mxfp4"])
@pytest.mark.parametrize("glu", [2, 1, 3])
def test_mmvq_fused_gate_glu_bias(orc, name, glu):
    """Fused epilogue: glu(gate.x + gb) * (up.x + b) (reference mmvq.cu:321-350, test-backend-ops.cpp:5420-5549)."""
    torch, capi, gguf_synth = _mods()
    t = TYPES[name]
    K, M = 512, 64
    rng = np.random.default_rng(glu)
    wu = gguf_synth.make_tensor_bytes(rng, t, (K, M), sigma=0.05)
    wg = gguf_synth.make_tensor_bytes(rng, t, (K, M), sigma=0.05)
    x = rng.uniform(-1, 1, (1, K)).astype(np.float32)
    b = rng.standard_normal(M).astype(np.float32)
    gb = rng.standard_normal(M).astype(np.float32)
    got = capi.mul_mat_vec_q(torch.from_numpy(wu).cuda(), t, [K, M, 1, 1], torch.from_numpy(x).cuda(), bias=torch.from_numpy(b).cuda(),
                             gate_bytes=torch.from_numpy(wg).cuda(), gate_bias=torch.from_numpy(gb).cuda(), glu_op=glu).cpu().numpy().reshape(M)
    rb = gguf_synth.row_bytes(t, K)
    up = np.zeros(M, np.float32)
    gate = np.zeros(M, np.float32)
    orc.orc_mul_mat_q8_1(t, ptr(wu), C.c_int64(rb), C.c_int64(M), C.c_int64(K), ptr(x), C.c_int64(1), ptr(up))
    orc.orc_mul_mat_q8_1(t, ptr(wg), C.c_int64(rb), C.c_int64(M), C.c_int64(K), ptr(x), C.c_int64(1), ptr(gate))
    up += b
    gate += gb
    want = np.zeros(M, np.float32)
    orc.orc_glu(glu, ptr(gate), ptr(up), ptr(want), C.c_int64(M), C.c_float(1.702), C.c_float(7.0))
    assert np.abs(got - want).max() <= 1e-4 * np.abs(want).max() + 1e-5
