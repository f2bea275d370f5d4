"""Host logic of the gpt-oss (MoE) model writer: tensor plan, active bytes per token, packed-domain samplers.  No GPU."""
import ctypes as C

import numpy as np

from conftest import ptr


def _gs():
    from llama_mi50_b200 import gguf_synth
    return gguf_synth


def test_gpt_oss_plan_matches_the_reference_loader_shapes():
    """reference src/llama-model.cpp:6570-6609: expert tensors [n_embd, n_ff_exp, n_expert] / [n_ff_exp, n_embd, n_expert] in MXFP4,
    biases per expert, sinks per head, router f32."""
    gs = _gs()
    s = gs.SHAPES["gpt-oss-120b-mxfp4"]
    plan = {n: (t, ne) for n, t, ne in gs.tensor_plan(s)}
    assert plan["blk.0.ffn_gate_exps.weight"] == (gs.MXFP4, (2880, 2880, 128))
    assert plan["blk.35.ffn_down_exps.weight"] == (gs.MXFP4, (2880, 2880, 128))
    assert plan["blk.0.ffn_up_exps.bias"] == (gs.F32, (2880, 128))
    assert plan["blk.0.ffn_gate_inp.weight"] == (gs.F32, (2880, 128))
    assert plan["blk.0.attn_sinks.weight"] == (gs.F32, (64,))
    assert plan["blk.0.attn_q.weight"] == (gs.Q8_0, (2880, 4096)) and plan["blk.0.attn_k.weight"] == (gs.Q8_0, (2880, 512))
    assert plan["blk.0.attn_output.weight"] == (gs.Q8_0, (4096, 2880))
    kv = gs.model_kv(s)
    assert kv["gpt-oss.expert_count"] == 128 and kv["gpt-oss.expert_used_count"] == 4 and kv["gpt-oss.attention.sliding_window"] == 128
    assert kv["gpt-oss.rope.scaling.type"] == "yarn"


def test_gpt_oss_active_bytes_per_token():
    """SURVEY.md section 8(d) / BASELINE.md: about 3.59 GB of weights are read per generated token (4 of 128 experts)."""
    gs = _gs()
    s = gs.SHAPES["gpt-oss-120b-mxfp4"]
    w = gs.matmul_weight_bytes(s)
    assert 3.55e9 < w < 3.62e9, w
    total = sum(gs.tensor_nbytes(t, ne) for _, t, ne in gs.tensor_plan(s))
    assert 62e9 < total < 65e9, total
    # dense models: every 2-D matmul weight once
    assert abs(gs.matmul_weight_bytes(gs.SHAPES["llama3-8b-q4_k_m"]) - 4.616e9) < 5e6


def test_packed_domain_samplers_give_valid_blocks(orc):
    """Large Q8_0 / MXFP4 tensors are sampled as random valid blocks: they must dequantise (with the oracle, i.e. the pinned
    restatement of the reference's dequantize_row_*) to finite values of about the requested standard deviation."""
    gs = _gs()
    rng = np.random.default_rng(7)
    for t, fn in ((gs.Q8_0, gs.sample_q8_0), (gs.MXFP4, gs.sample_mxfp4)):
        nb = 4096
        blocks = np.ascontiguousarray(fn(rng, nb, 0.02))
        out = np.zeros(nb * 32, np.float32)
        assert orc.orc_dequantize_row(t, ptr(blocks), ptr(out), C.c_int64(nb * 32)) == 0
        assert np.isfinite(out).all()
        assert 0.008 < out.std() < 0.05, (t, out.std())
