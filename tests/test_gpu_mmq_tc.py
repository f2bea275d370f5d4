"""GPU parity of the tensor-core prefill GEMM (b200_mul_mat_q, csrc/kernels/mmq_tc.cu; replaces reference
ggml-cuda/mmq.cu:71-218) against a float64 product of the oracle-dequantised weights with the f32 activations.
Tolerance: NMSE <= 1e-4 (the reference harness allows 5e-4 for MUL_MAT, tests/test-backend-ops.cpp:3718-3720; here the
only roundings are f16 weight expansion and 16-bit activations)."""
import ctypes as C

import numpy as np
import pytest

from conftest import ptr

pytestmark = pytest.mark.gpu

Q4_0, Q8_0, Q4_K, Q6_K, MXFP4 = 2, 8, 12, 14, 39


def _dequant(orc, gs, raw, t, K, M):
    out = np.zeros((M, K), np.float32)
    rb = gs.row_bytes(t, K)
    for m in range(M):
        row = np.ascontiguousarray(raw[m * rb:(m + 1) * rb])
        assert orc.orc_dequantize_row(t, ptr(row), ptr(out[m]), C.c_int64(K)) == 0
    return out.astype(np.float64)


@pytest.mark.parametrize("act_fmt", [0])
@pytest.mark.parametrize("t", [Q4_K, Q6_K, Q4_0, Q8_0, MXFP4])
@pytest.mark.parametrize("M,K,N", [(256, 512, 300), (128, 256, 9), (384, 1024, 512), (200, 768, 64), (256, 2048, 40), (4096, 4096, 512),
                                   (256, 4096, 40), (128, 14336, 64)])      # the attn_v / ffn_down shapes of Llama-3-8B (Q6_K in Q4_K_M)
def test_mul_mat_q_vs_float64(orc, t, M, K, N, act_fmt):
    import torch
    from llama_mi50_b200 import capi, gguf_synth as gs
    if gs.row_bytes(t, K) % 16 != 0:
        pytest.skip("rows of this type/K are not 16-byte multiples: not a tensor-map shape (the plugin keeps the GEMV-window path)")
    if M * K > 4e6 and t != Q4_K:
        pytest.skip("large case once")
    if t == MXFP4:
        act_fmt = 1          # bf16 operands: e2m1 x 2^(e-127) does not fit f16 (csrc/kernels/mmq_tc.cu)
    rng = np.random.default_rng(M * 7 + K + N + t)
    raw = gs.make_tensor_bytes(rng, t, (K, M), 0.02)
    x = rng.standard_normal((N, K)).astype(np.float32)
    x[0, :8] = [3e4, -7e4, 1e-6, 0.0, 5.0, -5.0, 123.456, 1e5]      # beyond f16 range: the conversion clamps to +-65504
    wd = torch.zeros(raw.nbytes + 64, dtype=torch.uint8, device="cuda"); wd[:raw.nbytes] = torch.from_numpy(raw).cuda()
    xd = torch.from_numpy(x).cuda()
    dst = torch.full((N, M), float("nan"), device="cuda")
    capi.mul_mat_q(wd, t, K, M, xd, dst, act_fmt=act_fmt)
    torch.cuda.synchronize()
    got = dst.cpu().numpy().astype(np.float64)
    W = _dequant(orc, gs, raw, t, K, M)
    xr = x.astype(np.float64)
    if act_fmt == 0:
        xr = np.clip(xr, -65504.0, 65504.0)
    want = xr @ W.T
    assert np.isfinite(got).all()
    nmse = ((got - want) ** 2).sum() / (want ** 2).sum()
    assert nmse <= 1e-4, f"NMSE {nmse}"
    # per-row check too (a wrong row inside a big product would hide in the global NMSE)
    row_nmse = ((got - want) ** 2).sum(axis=1) / ((want ** 2).sum(axis=1) + 1e-30)
    assert row_nmse.max() <= 1e-3, f"worst column NMSE {row_nmse.max()} at {row_nmse.argmax()}"


def test_mul_mat_q_residual_epilogue(orc):
    """{MUL_MAT, ADD} of a prefill graph: the residual stream is added in the GEMM epilogue, also when it aliases dst."""
    import torch
    from llama_mi50_b200 import capi, gguf_synth as gs
    M, K, N, t = 512, 1024, 300, Q4_K
    rng = np.random.default_rng(5)
    raw = gs.make_tensor_bytes(rng, t, (K, M), 0.02)
    x = rng.standard_normal((N, K)).astype(np.float32)
    res = rng.standard_normal((N, M)).astype(np.float32)
    wd = torch.zeros(raw.nbytes + 64, dtype=torch.uint8, device="cuda"); wd[:raw.nbytes] = torch.from_numpy(raw).cuda()
    xd = torch.from_numpy(x).cuda()
    plain = torch.empty((N, M), device="cuda")
    xb = capi.mul_mat_q(wd, t, K, M, xd, plain, act_fmt=0)
    fused = torch.from_numpy(res).cuda()                       # in place: addend == dst
    capi.mul_mat_q(wd, t, K, M, xd, fused, act_fmt=0, xb=xb, addend=fused)
    torch.cuda.synchronize()
    assert torch.equal(fused, plain + torch.from_numpy(res).cuda())
