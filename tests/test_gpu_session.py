"""GPU parity of the whole decode step (session.DecodeSession = the op sequence of reference
src/models/llama.cpp:3-165 issued through the C ABI) against a float64 numpy restatement of the same
graph over oracle-dequantized weights.  Tolerance: NMSE of the logits <= 2e-3 after several layers
and tokens (each quantized matmul is allowed 5e-4 by tests/test-backend-ops.cpp:3718-3720; the
q8_1 activation rounding and the f16 KV cache compound over layers)."""
import ctypes as C

import numpy as np
import pytest

from conftest import ptr

pytestmark = pytest.mark.gpu


def _dequant(orc, gs, data, t, ne):
    if t == gs.F32:
        return data.view(np.float32).astype(np.float64).reshape(ne[::-1])
    K, M = ne
    out = np.zeros((M, K), np.float32)
    rb = gs.row_bytes(t, K)
    raw = np.ascontiguousarray(data)
    for m in range(M):
        row = raw[m * rb:(m + 1) * rb]
        assert orc.orc_dequantize_row(t, ptr(row), ptr(out[m]), C.c_int64(K)) == 0
    return out.astype(np.float64)


def _rope(x, pos, n_head, hd, base, neox):
    x = x.reshape(n_head, hd).copy()
    half = hd // 2
    i = np.arange(half)
    theta = pos * base ** (-2.0 * i / hd)
    c, s = np.cos(theta), np.sin(theta)
    if neox:
        a, b = x[:, :half].copy(), x[:, half:].copy()
        x[:, :half], x[:, half:] = a * c - b * s, a * s + b * c
    else:
        a, b = x[:, 0::2].copy(), x[:, 1::2].copy()
        x[:, 0::2], x[:, 1::2] = a * c - b * s, a * s + b * c
    return x.reshape(-1)


def reference_decode(orc, gs, shape, weights, tokens):
    """float64 forward of the llama decode graph, one token at a time; returns logits per token."""
    W = {n: _dequant(orc, gs, d, t, ne) for n, (d, t, ne) in weights.items()}
    s = shape
    hd, H, Hkv = s.head_dim, s.n_head, s.n_head_kv
    neox = s.arch == "qwen2"

    def norm(x, w):
        return x / np.sqrt((x * x).mean() + s.rms_eps) * w

    kc = [[] for _ in range(s.n_layer)]
    vc = [[] for _ in range(s.n_layer)]
    outs = []
    for pos, tok in enumerate(tokens):
        x = W["token_embd.weight"][tok].copy()
        for il in range(s.n_layer):
            p = f"blk.{il}."
            h = norm(x, W[p + "attn_norm.weight"])
            q, k, v = W[p + "attn_q.weight"] @ h, W[p + "attn_k.weight"] @ h, W[p + "attn_v.weight"] @ h
            if p + "attn_q.bias" in W:
                q, k, v = q + W[p + "attn_q.bias"], k + W[p + "attn_k.bias"], v + W[p + "attn_v.bias"]
            q, k = _rope(q, pos, H, hd, s.rope_base, neox), _rope(k, pos, Hkv, hd, s.rope_base, neox)
            kc[il].append(k.astype(np.float16).astype(np.float64)); vc[il].append(v.astype(np.float16).astype(np.float64))
            K = np.stack(kc[il]).reshape(-1, Hkv, hd); V = np.stack(vc[il]).reshape(-1, Hkv, hd)
            att = np.zeros(H * hd)
            for hh in range(H):
                g = hh // (H // Hkv)
                sc = K[:, g, :] @ q[hh * hd:(hh + 1) * hd] / np.sqrt(hd)
                pr = np.exp(sc - sc.max()); pr /= pr.sum()
                att[hh * hd:(hh + 1) * hd] = pr @ V[:, g, :]
            x = x + W[p + "attn_output.weight"] @ att
            h = norm(x, W[p + "ffn_norm.weight"])
            g_, u_ = W[p + "ffn_gate.weight"] @ h, W[p + "ffn_up.weight"] @ h
            x = x + W[p + "ffn_down.weight"] @ (g_ / (1 + np.exp(-g_)) * u_)
        outs.append(W["output.weight"] @ norm(x, W["output_norm.weight"]))
    return outs


@pytest.mark.parametrize("key", ["tiny-q4_0", "tiny-q4_k_m", "tiny-q8_0"])
@pytest.mark.parametrize("graph,fused", [(False, True), (True, True), (False, False), (False, "chain"), (True, "chain")])
def test_decode_session_vs_numpy(orc, key, graph, fused):
    import torch
    from llama_mi50_b200 import gguf_synth as gs
    from llama_mi50_b200.session import DecodeSession
    shape = gs.SHAPES[key]
    sess = DecodeSession(shape, n_ctx=256, fused=fused)
    weights = {n: (d.cpu().numpy().view(np.uint8).reshape(-1), t, ne) for n, (d, t, ne) in sess.w.items()}
    tokens = [3, 77, 5, 130, 9, 250]
    want = reference_decode(orc, gs, shape, weights, tokens)
    got = []
    if graph:
        sess.resident_tokens = False
        sess.capture(256)
        for t in tokens:
            got.append(sess.decode(t).cpu().numpy().reshape(-1).astype(np.float64))
    else:
        sess.resident_tokens = False
        for i, t in enumerate(tokens):
            got.append(sess.decode_eager(t, i).cpu().numpy().reshape(-1).astype(np.float64))
    torch.cuda.synchronize()
    for i, (g, w) in enumerate(zip(got, want)):
        nmse = ((g - w) ** 2).sum() / (w ** 2).sum()
        assert nmse <= 2e-3, f"token {i}: logits NMSE {nmse}"


def test_graph_replay_matches_eager():
    """CUDA-graph replay and eager issue of the same step produce bit-identical logits."""
    import torch
    from llama_mi50_b200 import gguf_synth as gs
    from llama_mi50_b200.session import DecodeSession
    shape = gs.SHAPES["tiny-q4_k_m"]
    a, b = DecodeSession(shape, n_ctx=256), DecodeSession(shape, n_ctx=256)
    a.resident_tokens = b.resident_tokens = False
    b.capture(256)
    for i, t in enumerate([1, 2, 3, 4]):
        la = a.decode_eager(t, i).clone()
        lb = b.decode(t).clone()
        torch.cuda.synchronize()
        assert torch.equal(la, lb), f"token {i}"


class _OneRankDist:
    """torch.distributed stand-in for a single-rank peer group (handle exchange and barrier are trivial)."""

    @staticmethod
    def all_gather_object(out, obj):
        out[0] = obj

    @staticmethod
    def barrier():
        pass


@pytest.mark.parametrize("key", ["tiny-q4_k_m", "tiny-q8_0"])
@pytest.mark.parametrize("graph", [False, True])
def test_peer_chain_single_rank_matches_chain(key, graph):
    """The row-split kernel variant (system-scope barriers on never-reset counters, symmetric arena) with one rank
    must reproduce the single-GPU chain bit for bit, over many launches (the counters accumulate per launch)."""
    import torch
    from llama_mi50_b200 import gguf_synth as gs
    from llama_mi50_b200.session import DecodeSession
    shape = gs.SHAPES[key]
    a = DecodeSession(shape, n_ctx=256, fused="chain")
    b = DecodeSession(shape, n_ctx=256, fused="chain", peer_dist=_OneRankDist())
    assert b.arena is not None and b.arena.total_ctas == torch.cuda.get_device_properties(0).multi_processor_count
    a.resident_tokens = b.resident_tokens = False
    if graph:
        b.capture(256)
    for i, t in enumerate([1, 2, 3, 4, 250, 17, 99, 5, 6, 7]):
        la = a.decode_eager(t, i).clone()
        lb = (b.decode(t) if graph else b.decode_eager(t, i)).clone()
        torch.cuda.synchronize()
        assert torch.equal(la, lb), f"token {i}"


def test_peer_chain_rejects_dst_outside_arena():
    import ctypes as C
    import torch
    from llama_mi50_b200 import capi, gguf_synth as gs
    from llama_mi50_b200.session import DecodeSession
    b = DecodeSession(gs.SHAPES["tiny-q4_k_m"], n_ctx=256, fused="chain", peer_dist=_OneRankDist())
    b.logits = torch.zeros_like(b.logits)          # an ordinary allocation: other ranks could not reach it
    with pytest.raises(RuntimeError, match="outside the symmetric arena"):
        b._build_chain(256)
