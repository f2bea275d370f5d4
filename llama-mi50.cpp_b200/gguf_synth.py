"""Synthetic random-init GGUF models of the BASELINE.json shapes (no network, no checkpoints).

Own GGUF v3 writer plus numpy generators for the quantized block formats the hot path reads
(wire formats: reference ggml/src/ggml-common.h:170-175 q4_0, :219-224 q8_0, :190-195 mxfp4,
:295-306 q4_K, :330-336 q6_K).  Nothing here touches /root/reference or oracle/: the GPU box
has neither gguf-py nor the reference tree, and bench inputs must not come from the oracle.

Q4_0 / Q8_0 / MXFP4 tensors are *quantized* from N(0, sigma) with a straightforward numpy
restatement of the block rule (they are cheap).  Q4_K / Q6_K tensors are *sampled directly in
the packed domain*: uniformly random quants and 6-bit/8-bit sub-scales with the super-block
scale chosen so the dequantized weights have std ~= sigma and (for Q4_K) zero mean.  Both are
"random-init weights of the named architecture"; throughput does not depend on the values.

Tensor type mix for Q4_K_M follows reference src/llama-quant.cpp:185-187 (use_more_bits),
:207-228 (output -> Q6_K), :302-303 (attn_v), :358-364 (ffn_down).
"""
from __future__ import annotations

import struct
from dataclasses import dataclass, field

import numpy as np

# ggml type ids (reference ggml/include/ggml.h:390-430)
F32, F16, Q4_0, Q8_0, Q4_K, Q6_K, MXFP4 = 0, 1, 2, 8, 12, 14, 39
TYPE_NAME = {F32: "f32", F16: "f16", Q4_0: "q4_0", Q8_0: "q8_0", Q4_K: "q4_K", Q6_K: "q6_K", MXFP4: "mxfp4"}
TYPE_ID = {v: k for k, v in TYPE_NAME.items()}
# (elements per block, bytes per block)
BLOCK = {F32: (1, 4), F16: (1, 2), Q4_0: (32, 18), Q8_0: (32, 34), Q4_K: (256, 144), Q6_K: (256, 210), MXFP4: (32, 17)}

# llama_ftype (reference include/llama.h:118-154)
FTYPE = {"q4_0": 2, "q8_0": 7, "q4_k_m": 15, "mxfp4_moe": 38}


def row_bytes(ttype: int, k: int) -> int:
    be, bb = BLOCK[ttype]
    assert k % be == 0, (ttype, k)
    return k // be * bb


def tensor_nbytes(ttype: int, shape_ne) -> int:
    n_rows = int(np.prod(shape_ne[1:])) if len(shape_ne) > 1 else 1
    return row_bytes(ttype, int(shape_ne[0])) * n_rows


# ----------------------------------------------------------------------------- block generators

def quantize_q8_0(x: np.ndarray) -> np.ndarray:
    """f32 [..., K] -> packed q8_0 bytes. d = amax/127, q = round(x/d) (ggml-quants.c:197-223)."""
    xb = x.reshape(-1, 32).astype(np.float32)
    amax = np.abs(xb).max(axis=1)
    d = (amax / np.float32(127.0)).astype(np.float32)
    idd = np.where(d != 0, np.float32(1.0) / np.where(d != 0, d, 1), 0).astype(np.float32)
    q = np.rint(xb * idd[:, None]).astype(np.int8)  # ties-to-even vs roundf ties-away: synthetic data only
    out = np.empty((xb.shape[0], 34), np.uint8)
    out[:, 0:2] = d.astype(np.float16).view(np.uint8).reshape(-1, 2)
    out[:, 2:] = q.view(np.uint8)
    return out.reshape(-1)


def quantize_q4_0(x: np.ndarray) -> np.ndarray:
    """f32 -> packed q4_0 (ggml-quants.c:43-78): d = signed-max / -8, q = min(15, x/d + 8.5)."""
    xb = x.reshape(-1, 32).astype(np.float32)
    idx = np.abs(xb).argmax(axis=1)
    mx = xb[np.arange(xb.shape[0]), idx]
    d = (mx / np.float32(-8.0)).astype(np.float32)
    idd = np.where(d != 0, np.float32(1.0) / np.where(d != 0, d, 1), 0).astype(np.float32)
    q = np.minimum(15, (xb * idd[:, None] + np.float32(8.5)).astype(np.int32)).astype(np.uint8)
    out = np.empty((xb.shape[0], 18), np.uint8)
    out[:, 0:2] = d.astype(np.float16).view(np.uint8).reshape(-1, 2)
    out[:, 2:] = q[:, :16] | (q[:, 16:] << 4)
    return out.reshape(-1)


_E2M1 = np.array([0, 1, 2, 3, 4, 6, 8, 12, 0, -1, -2, -3, -4, -6, -8, -12], np.float32)  # kvalues_mxfp4 (2x e2m1)


def quantize_mxfp4(x: np.ndarray) -> np.ndarray:
    """f32 -> packed mxfp4 {e8m0 e; nibbles}: value = kvalues[q] * 2^(e-128) (ggml-common.h:190-195,1094)."""
    xb = x.reshape(-1, 32).astype(np.float32)
    amax = np.abs(xb).max(axis=1)
    with np.errstate(divide="ignore"):
        e = np.where(amax > 0, np.floor(np.log2(np.where(amax > 0, amax, 1))) - 2 + 127, 0).astype(np.int32)
    e = np.clip(e, 0, 254).astype(np.uint8)
    d = np.ldexp(np.float32(1.0), e.astype(np.int32) - 128).astype(np.float32)  # "e8m0 half" scale
    y = xb / d[:, None]
    q = np.abs(y[:, :, None] - _E2M1[None, None, :]).argmin(axis=2).astype(np.uint8)
    out = np.empty((xb.shape[0], 17), np.uint8)
    out[:, 0] = e
    out[:, 1:] = q[:, :16] | (q[:, 16:] << 4)
    return out.reshape(-1)


def sample_q4_k(rng: np.random.Generator, n_blocks: int, sigma: float) -> np.ndarray:
    """Random valid q4_K super-blocks: w = d*sc*q - dmin*m, sc,m in [0,63], q in [0,15]."""
    out = rng.integers(0, 256, size=(n_blocks, 144), dtype=np.uint8)
    # E[w] = d*31.5*7.5 - dmin*31.5 = 0  ->  dmin = 7.5 d ;  std(w) ~= 178 d
    d = (sigma / 178.0) * (0.75 + 0.5 * rng.random(n_blocks, dtype=np.float32))
    out[:, 0:2] = d.astype(np.float16).view(np.uint8).reshape(-1, 2)
    out[:, 2:4] = (7.5 * d).astype(np.float16).view(np.uint8).reshape(-1, 2)
    return out.reshape(-1)


def sample_q6_k(rng: np.random.Generator, n_blocks: int, sigma: float) -> np.ndarray:
    """Random valid q6_K super-blocks: w = d*sc*(q-32), sc int8, q in [0,63]."""
    out = rng.integers(0, 256, size=(n_blocks, 210), dtype=np.uint8)
    d = (sigma / 1369.0) * (0.75 + 0.5 * rng.random(n_blocks, dtype=np.float32))
    out[:, 208:210] = d.astype(np.float16).view(np.uint8).reshape(-1, 2)
    return out.reshape(-1)


def sample_q8_0(rng: np.random.Generator, n_blocks: int, sigma: float) -> np.ndarray:
    """Random valid q8_0 blocks (large tensors only: drawing and quantising 35 G normals took 9 minutes for the Qwen2.5-32B
    shape): quants uniform in [-127, 127] (std 73), d such that std(w) = sigma."""
    out = rng.integers(0, 256, size=(n_blocks, 34), dtype=np.uint8)
    q = out[:, 2:].view(np.int8)
    q[q == -128] = 0
    d = (sigma / 73.0) * (0.75 + 0.5 * rng.random(n_blocks, dtype=np.float32))
    out[:, 0:2] = d.astype(np.float16).view(np.uint8).reshape(-1, 2)
    return out.reshape(-1)


def sample_mxfp4(rng: np.random.Generator, n_blocks: int, sigma: float) -> np.ndarray:
    """Random valid mxfp4 blocks: nibbles uniform (|kvalue| mean-square 34.25 -> std 5.85 in units of 2^(e-128)), e chosen per
    tensor so that std(w) ~ sigma, +-1 per block."""
    out = rng.integers(0, 256, size=(n_blocks, 17), dtype=np.uint8)
    e0 = int(round(128 + np.log2(max(sigma, 1e-12) / 5.85)))
    out[:, 0] = np.clip(e0 + rng.integers(-1, 2, size=n_blocks), 1, 254).astype(np.uint8)
    return out.reshape(-1)


SAMPLE_FROM = 1 << 26      # elements: tensors at least this large are sampled in the packed domain (throughput does not depend on
                           # the values; the exact quantizers stay the reference for tests and small tensors)


def make_tensor_bytes(rng: np.random.Generator, ttype: int, ne, sigma: float = 0.02, value: float | None = None) -> np.ndarray:
    """Packed bytes of a random-init tensor with ggml shape `ne` (ne[0] fastest)."""
    n = int(np.prod(ne))
    if ttype == F32:
        a = np.full(n, value, np.float32) if value is not None else (rng.standard_normal(n, dtype=np.float32) * sigma)
        return a.view(np.uint8)
    if ttype == F16:
        return (rng.standard_normal(n, dtype=np.float32) * sigma).astype(np.float16).view(np.uint8)
    be, bb = BLOCK[ttype]
    nb = n // be
    if ttype == Q4_K:
        return sample_q4_k(rng, nb, sigma)
    if ttype == Q6_K:
        return sample_q6_k(rng, nb, sigma)
    if n >= SAMPLE_FROM and ttype in (Q8_0, MXFP4):
        # a pool of 4 M random valid blocks, laid down repeatedly at random rotations (the generator delivers ~0.3 GB/s; a
        # 63 GB model would take minutes of GPU-box time for values nothing depends on)
        fn = sample_q8_0 if ttype == Q8_0 else sample_mxfp4
        pool_blocks = min(nb, 1 << 22)
        pool = fn(rng, pool_blocks, sigma).reshape(pool_blocks, bb)
        out = np.empty((nb, bb), np.uint8)
        for b0 in range(0, nb, pool_blocks):
            b1 = min(nb, b0 + pool_blocks)
            r = int(rng.integers(0, pool_blocks))
            out[b0:b1] = np.roll(pool, r, axis=0)[:b1 - b0] if r else pool[:b1 - b0]
        return out.reshape(-1)
    out = np.empty(nb * bb, np.uint8)
    step = 1 << 20  # blocks per chunk, bounds the f32 temporary
    fn = {Q4_0: quantize_q4_0, Q8_0: quantize_q8_0, MXFP4: quantize_mxfp4}[ttype]
    for b0 in range(0, nb, step):
        b1 = min(nb, b0 + step)
        x = rng.standard_normal((b1 - b0) * be, dtype=np.float32) * np.float32(sigma)
        out[b0 * bb:b1 * bb] = fn(x)
    return out


# ----------------------------------------------------------------------------- model shapes

@dataclass
class ModelShape:
    name: str
    arch: str
    n_layer: int
    n_embd: int
    n_ff: int
    n_head: int
    n_head_kv: int
    head_dim: int
    n_vocab: int
    ftype: str
    n_ctx_train: int = 8192
    rope_base: float = 500000.0
    rms_eps: float = 1e-5
    n_expert: int = 0
    n_expert_used: int = 0
    swa: int = 0
    qkv_bias: bool = False
    extra: dict = field(default_factory=dict)


# Deep random-init models with N(0, 0.02) weights everywhere are CHAOTIC: every block adds more than the residual stream
# carries, so a perturbation of one quantisation step grows layer by layer (measured on the reference CPU backend alone:
# Llama-3-8B shape, f16 vs q8_0 KV cache -> logits NMSE 0.25-0.48; oracle/plugin_parity.cpp "cpu" probe).  A parity test on
# such a model measures nothing.  The deep shapes therefore use block weights of sigma 0.008 under an embedding of sigma
# 0.4 (same probe: 5e-4..8e-4), which is how a trained model behaves: each block is a small update of the stream.
# Throughput does not depend on the values.
_DEEP = {"sigma": 0.008, "sigma_embd": 0.4}

SHAPES = {
    # BASELINE.json configs[0..4]; hparams per SURVEY.md section 8
    "tinyllama-1.1b-q4_0": ModelShape("tinyllama-1.1b", "llama", 22, 2048, 5632, 32, 4, 64, 32000, "q4_0", 2048, 10000.0, extra=_DEEP),
    "llama3-8b-q4_k_m": ModelShape("llama3-8b", "llama", 32, 4096, 14336, 32, 8, 128, 128256, "q4_k_m", extra=_DEEP),
    "qwen2.5-32b-q8_0": ModelShape("qwen2.5-32b", "qwen2", 64, 5120, 27648, 40, 8, 128, 152064, "q8_0", 32768, 1000000.0, 1e-6, qkv_bias=True, extra=_DEEP),
    "llama3-70b-q4_k_m": ModelShape("llama3-70b", "llama", 80, 8192, 28672, 64, 8, 128, 128256, "q4_k_m", extra=_DEEP),
    # gpt-oss-120b (reference src/models/openai-moe-iswa.cpp, src/llama-model.cpp:2262-2280, 6570-6609): 36 layers, 128 experts / 4
    # used, expert FFN 2880, sliding window 128 on every other layer, attention sinks, biases everywhere; experts MXFP4, the rest
    # Q8_0 (ftype MXFP4_MOE).  YaRN (factor 32 over 4096) as the released model has it.
    "gpt-oss-120b-mxfp4": ModelShape("gpt-oss-120b", "gpt-oss", 36, 2880, 2880, 64, 8, 64, 201088, "mxfp4_moe", 131072, 150000.0, 1e-5,
                                      n_expert=128, n_expert_used=4, swa=128, qkv_bias=True,
                                      extra={**_DEEP, "yarn_factor": 32.0, "yarn_orig_ctx": 4096}),
    # test-size models (same graph topology, seconds on a CPU)
    "tiny-moe-mxfp4": ModelShape("tiny-moe", "gpt-oss", 2, 256, 256, 4, 2, 64, 512, "mxfp4_moe", 512, 10000.0, 1e-5,
                                  n_expert=8, n_expert_used=2, swa=64, qkv_bias=True, extra={"yarn_factor": 4.0, "yarn_orig_ctx": 128}),
    "tiny-q4_0": ModelShape("tiny", "llama", 2, 256, 512, 4, 2, 64, 512, "q4_0", 512, 10000.0),
    "tiny-q4_k_m": ModelShape("tiny-k", "llama", 4, 512, 1024, 8, 2, 64, 1024, "q4_k_m", 512, 10000.0),
    "tiny-q8_0": ModelShape("tiny-q8", "qwen2", 2, 256, 512, 4, 2, 64, 512, "q8_0", 512, 10000.0, 1e-6, qkv_bias=True),
}


def use_more_bits(i_layer: int, n_layers: int) -> bool:
    return i_layer < n_layers // 8 or i_layer >= 7 * n_layers // 8 or (i_layer - n_layers // 8) % 3 == 2


def tensor_plan(s: ModelShape):
    """[(name, ggml_type, ne)] in file order; the weight-type mix of the ftype."""
    if s.arch == "gpt-oss":
        return tensor_plan_gpt_oss(s)
    base = {"q4_0": Q4_0, "q8_0": Q8_0, "q4_k_m": Q4_K}[s.ftype]
    out_t = {"q4_0": Q6_K, "q8_0": Q8_0, "q4_k_m": Q6_K}[s.ftype]
    E, F, KVD, QD = s.n_embd, s.n_ff, s.n_head_kv * s.head_dim, s.n_head * s.head_dim
    if E % 256 != 0 and base in (Q4_K, Q6_K):
        raise ValueError("k-quants need n_embd % 256 == 0")
    if E % 256 != 0 and out_t == Q6_K:
        out_t = Q8_0
    plan = [("token_embd.weight", base, (E, s.n_vocab))]
    for i in range(s.n_layer):
        more = s.ftype == "q4_k_m" and use_more_bits(i, s.n_layer)
        p = f"blk.{i}."
        plan += [(p + "attn_norm.weight", F32, (E,)),
                 (p + "attn_q.weight", base, (E, QD)),
                 (p + "attn_k.weight", base, (E, KVD)),
                 (p + "attn_v.weight", Q6_K if more else base, (E, KVD))]
        if s.qkv_bias:
            plan += [(p + "attn_q.bias", F32, (QD,)), (p + "attn_k.bias", F32, (KVD,)), (p + "attn_v.bias", F32, (KVD,))]
        plan += [(p + "attn_output.weight", base, (QD, E)),
                 (p + "ffn_norm.weight", F32, (E,)),
                 (p + "ffn_gate.weight", base, (E, F)),
                 (p + "ffn_up.weight", base, (E, F)),
                 (p + "ffn_down.weight", Q6_K if more else base, (F, E))]
    plan += [("output_norm.weight", F32, (E,)), ("output.weight", out_t, (E, s.n_vocab))]
    return plan


def tensor_plan_gpt_oss(s: ModelShape):
    """gpt-oss (reference src/llama-model.cpp:6570-6609): expert tensors MXFP4 [K, M, n_expert], everything else Q8_0 / F32."""
    E, F, KVD, QD, X = s.n_embd, s.n_ff, s.n_head_kv * s.head_dim, s.n_head * s.head_dim, s.n_expert
    plan = [("token_embd.weight", Q8_0, (E, s.n_vocab))]
    for i in range(s.n_layer):
        p = f"blk.{i}."
        plan += [(p + "attn_norm.weight", F32, (E,)),
                 (p + "attn_q.weight", Q8_0, (E, QD)), (p + "attn_q.bias", F32, (QD,)),
                 (p + "attn_k.weight", Q8_0, (E, KVD)), (p + "attn_k.bias", F32, (KVD,)),
                 (p + "attn_v.weight", Q8_0, (E, KVD)), (p + "attn_v.bias", F32, (KVD,)),
                 (p + "attn_output.weight", Q8_0, (QD, E)), (p + "attn_output.bias", F32, (E,)),
                 (p + "attn_sinks.weight", F32, (s.n_head,)),
                 (p + "post_attention_norm.weight", F32, (E,)),
                 (p + "ffn_gate_inp.weight", F32, (E, X)), (p + "ffn_gate_inp.bias", F32, (X,)),
                 (p + "ffn_gate_exps.weight", MXFP4, (E, F, X)), (p + "ffn_gate_exps.bias", F32, (F, X)),
                 (p + "ffn_up_exps.weight", MXFP4, (E, F, X)), (p + "ffn_up_exps.bias", F32, (F, X)),
                 (p + "ffn_down_exps.weight", MXFP4, (F, E, X)), (p + "ffn_down_exps.bias", F32, (E, X))]
    plan += [("output_norm.weight", F32, (E,)), ("output.weight", Q8_0, (E, s.n_vocab))]
    return plan


def matmul_weight_bytes(s: ModelShape) -> int:
    """Algorithmic weight bytes read per generated token: every MUL_MAT weight once."""
    total = 0
    for name, t, ne in tensor_plan(s):
        if name.endswith(".weight") and len(ne) == 2 and name != "token_embd.weight":
            total += tensor_nbytes(t, ne)
        elif name.endswith("_exps.weight"):                 # MUL_MAT_ID: the n_expert_used selected experts of n_expert
            total += tensor_nbytes(t, ne) // ne[2] * s.n_expert_used
    return total


def matmul_weight_params(s: ModelShape) -> int:
    """Weight elements of every MUL_MAT of one token (2 flop each): the prefill GEMM work per prompt token."""
    total = 0
    for name, t, ne in tensor_plan(s):
        if name.endswith(".weight") and len(ne) == 2 and name != "token_embd.weight":
            total += int(ne[0]) * int(ne[1])
        elif name.endswith("_exps.weight"):
            total += int(ne[0]) * int(ne[1]) * s.n_expert_used
    return total


def decode_bytes_per_token(s: ModelShape, n_ctx_avg: float, kv_type: int = F16) -> dict:
    """Algorithmic HBM bytes for one generated token."""
    w = matmul_weight_bytes(s)
    kv_row = row_bytes(kv_type, s.head_dim) * s.n_head_kv
    kv_read = int(2 * s.n_layer * kv_row * n_ctx_avg)
    kv_write = 2 * s.n_layer * kv_row
    emb = row_bytes(tensor_plan(s)[0][1], s.n_embd)
    logits = s.n_vocab * 4
    return {"weights": w, "kv_read": kv_read, "kv_write": kv_write, "embd_row": emb, "logits": logits,
            "total": w + kv_read + kv_write + emb + logits}


# ----------------------------------------------------------------------------- GGUF v3 writer

_GGUF_U32, _GGUF_F32, _GGUF_STR, _GGUF_ARR, _GGUF_U64 = 4, 6, 8, 9, 10


def _s(b: str) -> bytes:
    e = b.encode()
    return struct.pack("<Q", len(e)) + e


def _kv(key: str, val) -> bytes:
    if isinstance(val, str):
        return _s(key) + struct.pack("<I", _GGUF_STR) + _s(val)
    if isinstance(val, float):
        return _s(key) + struct.pack("<If", _GGUF_F32, val)
    if isinstance(val, int):
        return _s(key) + struct.pack("<II", _GGUF_U32, val)
    raise TypeError(type(val))


def model_kv(s: ModelShape) -> dict:
    a = s.arch
    kv = {
        "general.architecture": a,
        "general.name": s.name + "-random-init",
        "general.file_type": FTYPE[s.ftype],
        "general.quantization_version": 2,
        f"{a}.context_length": s.n_ctx_train,
        f"{a}.embedding_length": s.n_embd,
        f"{a}.block_count": s.n_layer,
        f"{a}.feed_forward_length": s.n_ff,
        f"{a}.attention.head_count": s.n_head,
        f"{a}.attention.head_count_kv": s.n_head_kv,
        f"{a}.attention.layer_norm_rms_epsilon": float(s.rms_eps),
        f"{a}.rope.dimension_count": s.head_dim,
        f"{a}.rope.freq_base": float(s.rope_base),
        f"{a}.vocab_size": s.n_vocab,
        "tokenizer.ggml.model": "no_vocab",  # reference src/llama-vocab.cpp:1706-1727
    }
    if s.n_head * s.head_dim != s.n_embd:
        kv[f"{a}.attention.key_length"] = s.head_dim
        kv[f"{a}.attention.value_length"] = s.head_dim
    if s.n_expert:
        kv[f"{a}.expert_count"] = s.n_expert
        kv[f"{a}.expert_used_count"] = s.n_expert_used
        kv[f"{a}.expert_feed_forward_length"] = s.n_ff
    if s.swa:
        kv[f"{a}.attention.sliding_window"] = s.swa
    if s.extra.get("yarn_factor"):
        kv[f"{a}.rope.scaling.type"] = "yarn"
        kv[f"{a}.rope.scaling.factor"] = float(s.extra["yarn_factor"])
        kv[f"{a}.rope.scaling.original_context_length"] = int(s.extra["yarn_orig_ctx"])
    return kv


def write_gguf(path: str, s: ModelShape, seed: int = 1234, sigma: float = 0.02, alignment: int = 32) -> dict:
    """Stream a random-init GGUF of shape `s` to `path`. Returns {'bytes','tensors','weight_bytes'}."""
    rng = np.random.default_rng(seed)
    plan = tensor_plan(s)
    kv = model_kv(s)
    hdr = struct.pack("<4sIQQ", b"GGUF", 3, len(plan), len(kv))
    for k, v in kv.items():
        hdr += _kv(k, v)
    off = 0
    infos = b""
    offs = []
    for name, t, ne in plan:
        infos += _s(name) + struct.pack("<I", len(ne)) + b"".join(struct.pack("<Q", d) for d in ne)
        infos += struct.pack("<IQ", t, off)
        offs.append(off)
        off += -(-tensor_nbytes(t, ne) // alignment) * alignment
    head = hdr + infos
    pad = (-len(head)) % alignment
    with open(path, "wb") as f:
        f.write(head + b"\0" * pad)
        base = f.tell()
        for (name, t, ne), o in zip(plan, offs):
            assert f.tell() == base + o, name
            val = 1.0 if name.endswith("_norm.weight") else None
            sg = 0.0 if name.endswith(".bias") else float(s.extra.get("sigma", sigma))
            if name == "token_embd.weight":
                sg = float(s.extra.get("sigma_embd", sg))
            if val is None and t == F32 and sg == 0.0:
                val = 0.0
            data = make_tensor_bytes(rng, t, ne, sg if sg else sigma, val)
            assert data.nbytes == tensor_nbytes(t, ne), (name, data.nbytes)
            f.write(memoryview(data))
            f.write(b"\0" * ((-data.nbytes) % alignment))
        total = f.tell()
    return {"bytes": total, "tensors": len(plan), "weight_bytes": matmul_weight_bytes(s)}


if __name__ == "__main__":
    import argparse, json, time

    ap = argparse.ArgumentParser(description=__doc__.splitlines()[0])
    ap.add_argument("shape", choices=sorted(SHAPES))
    ap.add_argument("out")
    ap.add_argument("--seed", type=int, default=1234)
    a = ap.parse_args()
    t0 = time.time()
    r = write_gguf(a.out, SHAPES[a.shape], a.seed)
    r["seconds"] = round(time.time() - t0, 2)
    print(json.dumps(r))
