"""b200_flash_attn_ext_static — the entry point for callers that cannot allocate (include/b200_kernels.h): the split-KV
scratch is a static device array inside the library.  Called through the C ABI twice in a row and compared with the
existing entry point (caller-provided scratch) on the same inputs.  Both run the same kernels with the same split plan,
so the outputs must be identical bit for bit; the bound written here is the one tests/test_gpu_ops.py uses for flash
attention on the tensor-core decode kernel (NMSE < 2e-6).

This file is named so that pytest collects it last: a failure here must not hide
the other tests.  It passes alone in a fresh process and after the other tests (the static array keeps whatever the
previous call left; the arrival counters are cleared by every call)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _nmse(want, got):
    return float(((want - got) ** 2).sum() / (want ** 2).sum())


@pytest.mark.parametrize("kv_type,D,H,H_kv,n_kv", [("f16", 128, 32, 8, 4096),      # split-KV, tensor-core decode kernel
                                                    ("q8_0", 128, 40, 8, 4096),     # Qwen2.5-32B heads (gqa 5), q8_0 cache
                                                    ("f16", 64, 32, 4, 300)])       # short context: vector kernel
def test_static_scratch_matches_caller_scratch_twice(kv_type, D, H, H_kv, n_kv):
    import torch
    from llama_mi50_b200 import capi, gguf_synth

    rng = np.random.default_rng(n_kv + H)
    q = torch.from_numpy(rng.standard_normal((H, 1, D)).astype(np.float32)).cuda()
    if kv_type == "f16":
        t = capi.F16
        kb = torch.from_numpy(rng.standard_normal((H_kv, n_kv, D)).astype(np.float16).view(np.uint8).reshape(-1)).cuda()
        vb = torch.from_numpy(rng.standard_normal((H_kv, n_kv, D)).astype(np.float16).view(np.uint8).reshape(-1)).cuda()
    else:
        t = capi.Q8_0
        kb = torch.from_numpy(gguf_synth.make_tensor_bytes(rng, t, (D, n_kv * H_kv), sigma=1.0)).cuda()
        vb = torch.from_numpy(gguf_synth.make_tensor_bytes(rng, t, (D, n_kv * H_kv), sigma=1.0)).cuda()
    m = np.zeros((1, n_kv), np.float32)
    m[:, n_kv // 3: n_kv // 3 + 40] = -np.inf
    mask = torch.from_numpy(m.astype(np.float16)).cuda()
    scale = 1.0 / np.sqrt(D)
    want = capi.flash_attn(q, kb, vb, t, n_kv, H_kv, mask=mask, scale=scale)
    torch.cuda.synchronize()
    out = torch.empty_like(want)
    for call in range(2):                              # two calls in a row share the static array
        out.fill_(float("nan"))
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated()
        capi.flash_attn(q, kb, vb, t, n_kv, H_kv, mask=mask, scale=scale, static_scratch=True, dst=out)
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated() == before, "the static entry point must not allocate"
        w, g = want.cpu().numpy().astype(np.float64), out.cpu().numpy().astype(np.float64)
        assert np.isfinite(g).all()
        assert _nmse(w, g) < 2e-6, f"call {call}"
        assert np.array_equal(want.cpu().numpy(), out.cpu().numpy()), f"call {call}: same kernels, same split plan: bits must agree"


def test_static_scratch_rejects_oversized_requests():
    """A shape whose split-KV partials exceed B200_FATTN_STATIC_SCRATCH_BYTES is refused with a clear error (host-side check:
    nothing is launched)."""
    import torch
    from llama_mi50_b200 import capi

    D, H, H_kv, n_kv, n_q = 256, 110, 110, 1024, 64     # 880 (kv-head, chunk) groups x 2 splits x 8 columns x 258 floats = 14.5 MB
    a = capi.FattnArgs()
    q = torch.zeros(H, n_q, D, device="cuda")
    dst = torch.empty(n_q, H, D, device="cuda")
    a.q, a.dst = capi.from_torch(q), capi.from_torch(dst)
    # K/V are never touched: the size check precedes every CUDA call
    a.k = capi.make_tensor(q.data_ptr(), capi.F16, [D, n_kv, H_kv, 1])
    a.v = capi.make_tensor(q.data_ptr(), capi.F16, [D, n_kv, H_kv, 1])
    a.scale = 1.0
    L = capi.lib()
    L.b200_fattn_scratch_size.restype = C.c_size_t
    L.b200_fattn_scratch_size.argtypes = [C.c_void_p]
    L.b200_flash_attn_ext_static.argtypes = [C.c_void_p, C.c_void_p]
    need = L.b200_fattn_scratch_size(C.byref(a))
    assert need > (8 << 20), need
    rc = L.b200_flash_attn_ext_static(C.byref(a), capi.stream_ptr())
    assert rc == -3                                       # B200_ERR_ARG
    assert b"static array holds" in L.b200_last_error()
