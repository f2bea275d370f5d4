"""GPU parity of the streaming fused GEMV (b200_gemv_fused, SURVEY rows A2-A8 + the fusions of
ggml-cuda.cu:3737-3771, 3844-3854) through the C ABI against the pinned C oracle.

 * the q8_1 activation blocks the prologue produces are bit-exact vs quantize_row_q8_1_ref;
 * each product equals the oracle's q8_1-semantics product up to f32 summation order (2e-5 of the
   output scale) and passes the reference's NMSE gate (5e-4, tests/test-backend-ops.cpp:3718-3720)
   against the f64 ground truth."""
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


def _weights(request, gs, rng, t, K, M):
    try:
        refbase = request.getfixturevalue("refbase")
    except pytest.skip.Exception:
        refbase = None
    if refbase is not None:
        wf = (rng.standard_normal(M * K) * 0.5).astype(np.float32)
        w = np.zeros(refbase.ggml_row_size(t, K) * M, np.uint8)
        refbase.ggml_quantize_chunk(t, ptr(wf), ptr(w), 0, M, K, None)
        return w
    return gs.make_tensor_bytes(rng, t, (K, M), sigma=0.5)


def _dev(torch, a):
    """device copy with >= 16 bytes of slack after the payload (the kernel's bulk copies round up to 16 B)"""
    t = torch.zeros(a.nbytes + 64, dtype=torch.uint8, device="cuda")
    t[:a.nbytes] = torch.from_numpy(a.view(np.uint8).reshape(-1))
    return t


def _oracle(orc, gs, t, w, M, K, x):
    want = np.zeros(M, np.float32); truth = np.zeros(M, np.float32)
    rb = gs.row_bytes(t, K)
    assert orc.orc_mul_mat_q8_1(t, ptr(w), C.c_int64(rb), C.c_int64(M), C.c_int64(K), ptr(x), C.c_int64(1), ptr(want)) == 0
    assert orc.orc_mul_mat_f64(t, ptr(w), C.c_int64(rb), C.c_int64(M), C.c_int64(K), ptr(x), C.c_int64(1), ptr(truth)) == 0
    return want, truth


def _check(orc, got, want, truth):
    scale = np.abs(want).max() + 1e-30
    assert np.abs(got - want).max() <= 2e-5 * scale + 1e-6, "differs from the q8_1-semantics oracle beyond f32 summation order"
    assert orc.orc_nmse(ptr(truth), ptr(np.ascontiguousarray(got)), C.c_size_t(got.size)) < 5e-4


SHAPES = [(16, 256), (33, 512), (257, 1024), (1000, 4096), (4096, 4096), (300, 14336), (7, 2048), (1, 256), (149, 8192), (40, 28672)]


@pytest.mark.parametrize("name", list(TYPES))
@pytest.mark.parametrize("M,K", SHAPES)
def test_single_segment(orc, request, name, M, K):
    torch, capi, gs = _mods()
    t = TYPES[name]
    rng = np.random.default_rng(M * 31 + K)
    w = _weights(request, gs, rng, t, K, M)
    x = rng.uniform(-1, 1, K).astype(np.float32)
    dst = torch.full((M,), float("nan"), dtype=torch.float32, device="cuda")
    xq = torch.zeros(K // 32 * 36, dtype=torch.uint8, device="cuda")
    capi.gemv_fused(torch.from_numpy(x).cuda(), [(_dev(torch, w), t, M, None, dst)], xq_out=xq)
    got = dst.cpu().numpy()
    want, truth = _oracle(orc, gs, t, w, M, K, x)
    _check(orc, got, want, truth)
    ref_q = np.zeros(K // 32 * 36, np.uint8)
    orc.orc_quantize_row_q8_1(ptr(x), ptr(ref_q), C.c_int64(K))
    assert np.array_equal(xq.cpu().numpy(), ref_q), "prologue q8_1 blocks are not bit-exact vs quantize_row_q8_1_ref"


@pytest.mark.parametrize("name", ["q8_0", "mxfp4", "q4_0"])
@pytest.mark.parametrize("M,K", [(2880, 2880), (513, 2880), (5, 5632), (64, 96)])
def test_rows_not_16_byte_multiples(orc, request, name, M, K):
    """row sizes that are not multiples of 16 B (gpt-oss K=2880, TinyLlama F=5632): bulk copies start mid-16B"""
    torch, capi, gs = _mods()
    t = TYPES[name]
    rng = np.random.default_rng(M + K)
    w = _weights(request, gs, rng, t, K, M)
    x = rng.uniform(-1, 1, K).astype(np.float32)
    dst = torch.zeros(M, dtype=torch.float32, device="cuda")
    capi.gemv_fused(torch.from_numpy(x).cuda(), [(_dev(torch, w), t, M, None, dst)])
    want, truth = _oracle(orc, gs, t, w, M, K, x)
    _check(orc, dst.cpu().numpy(), want, truth)


@pytest.mark.parametrize("types", [("q4_K", "q4_K", "q6_K"), ("q4_K", "q4_K", "q4_K"), ("q4_0", "q4_0", "q4_0"), ("q8_0", "q8_0", "q8_0")])
@pytest.mark.parametrize("K,Ms", [(4096, (4096, 1024, 1024)), (2048, (2048, 256, 256)), (512, (100, 3, 50))])
def test_qkv_with_rms_norm_and_bias(orc, request, types, K, Ms):
    """three segments sharing one RMS-normalised activation (wq/wk/wv of a layer), optional bias"""
    torch, capi, gs = _mods()
    rng = np.random.default_rng(K + sum(Ms))
    x = rng.standard_normal(K).astype(np.float32)
    nw = (1.0 + 0.1 * rng.standard_normal(K)).astype(np.float32)
    eps = 1e-5
    scale = np.float32(1.0) / np.sqrt(np.float32((x.astype(np.float64) ** 2).mean()) + np.float32(eps))
    xn = (x * np.float32(scale)) * nw
    segs, ws, dsts, biases = [], [], [], []
    for name, M in zip(types, Ms):
        t = TYPES[name]
        w = _weights(request, gs, rng, t, K, M)
        b = rng.standard_normal(M).astype(np.float32)
        d = torch.zeros(M, dtype=torch.float32, device="cuda")
        ws.append((t, w, M)); dsts.append(d); biases.append(b)
        segs.append((_dev(torch, w), t, M, torch.from_numpy(b).cuda(), d))
    capi.gemv_fused(torch.from_numpy(x).cuda(), segs, norm_w=torch.from_numpy(nw).cuda(), eps=eps)
    for (t, w, M), d, b in zip(ws, dsts, biases):
        want, truth = _oracle(orc, gs, t, w, M, K, xn.astype(np.float32))
        got = d.cpu().numpy() - b
        # the normalisation scale may differ in the last ulp (summation order), which can flip a q8 rounding:
        # compare with the NMSE gate and a loose max-error bound instead of the summation-order bound
        assert orc.orc_nmse(ptr(truth), ptr(np.ascontiguousarray(got)), C.c_size_t(M)) < 5e-4
        assert np.abs(got - want).max() <= 2e-3 * (np.abs(want).max() + 1e-30)


@pytest.mark.parametrize("name", ["q4_K", "q4_0", "q8_0", "mxfp4", "q6_K"])
@pytest.mark.parametrize("glu", [0, 1, 2, 3, 4, 5])
def test_gate_up_glu(orc, request, name, glu):
    torch, capi, gs = _mods()
    t = TYPES[name]
    K, M = 1024, 700
    rng = np.random.default_rng(glu * 7 + t)
    wg, wu = _weights(request, gs, rng, t, K, M), _weights(request, gs, rng, t, K, M)
    x = rng.uniform(-1, 1, K).astype(np.float32)
    dst = torch.zeros(M, dtype=torch.float32, device="cuda")
    capi.gemv_fused(torch.from_numpy(x).cuda(), [(_dev(torch, wg), t, M, None, None), (_dev(torch, wu), t, M, None, dst)], glu_op=glu)
    g, _ = _oracle(orc, gs, t, wg, M, K, x)
    u, _ = _oracle(orc, gs, t, wu, M, K, x)
    want = np.zeros(M, np.float32)
    orc.orc_glu(glu, ptr(g), ptr(u), ptr(want), C.c_int64(M), C.c_float(1.702), C.c_float(7.0))
    got = dst.cpu().numpy()
    assert np.abs(got - want).max() <= 1e-4 * (np.abs(want).max() + 1e-30) + 1e-5


def test_residual_epilogue_and_determinism(orc, request):
    """dst = W.x + residual; repeated launches and different SM counts give bit-identical results"""
    torch, capi, gs = _mods()
    t, K, M = 12, 14336, 4096
    rng = np.random.default_rng(5)
    w = _weights(request, gs, rng, t, K, M)
    x = rng.uniform(-1, 1, K).astype(np.float32)
    res = rng.standard_normal(M).astype(np.float32)
    wd, xd, rd = _dev(torch, w), torch.from_numpy(x).cuda(), torch.from_numpy(res).cuda()
    outs = []
    for n_sm in (0, 0, 37, 1, 148, 296):
        d = torch.zeros(M, dtype=torch.float32, device="cuda")
        capi.gemv_fused(xd, [(wd, t, M, rd, d)], n_sm=n_sm)
        outs.append(d.cpu().numpy())
    for o in outs[1:]:
        assert np.array_equal(o, outs[0])
    want, truth = _oracle(orc, gs, t, w, M, K, x)
    _check(orc, outs[0] - res, want, truth)
