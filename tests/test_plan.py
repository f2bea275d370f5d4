"""Host logic of the plugin's decode-chain planner (csrc/backend/chain_plan.cpp) on the decode graph the
UNMODIFIED reference really builds: oracle/_ref/plan-check runs the reference's CPU backend on a synthetic GGUF,
rebuilds the executed single-token graph (same node order, real ggml-alloc addresses) and prints the plan.
Expectation (reference graph: src/models/llama.cpp:3-165, llama-graph.cpp:1675-1900): per layer the 5 phases
qkv / attention / wo+residual / gate,up+GLU / down+residual, plus the logits matmul; plain nodes only for the
embedding lookup, the first layer's mask cast, the last layer's output-row selection and the (OUTPUT-flagged)
final norm."""
import os
import re
import subprocess

import pytest

from conftest import REF_DIR, ROOT


@pytest.fixture(scope="module")
def plan_check():
    exe = os.path.join(REF_DIR, "plan-check")
    if not os.path.exists(exe):
        if not os.path.isdir("/root/reference"):
            pytest.skip("oracle/_ref/plan-check not built (needs /root/reference at build time)")
        subprocess.check_call(["make", "harness"], cwd=os.path.join(ROOT, "oracle"))
    return exe


@pytest.mark.parametrize("key", ["tiny-q4_0", "tiny-q4_k_m", "tiny-q8_0"])
def test_decode_graph_is_planned_as_chain(plan_check, key, tmp_path):
    from llama_mi50_b200 import gguf_synth as gs
    shape = gs.SHAPES[key]
    path = str(tmp_path / f"{key}.gguf")
    gs.write_gguf(path, shape)
    out = subprocess.run([plan_check, path, REF_DIR], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = out.stdout.splitlines()
    assert lines[0].endswith("same_order 1"), lines[0]
    m = re.search(r"summary phases (\d+) attn (\d+) plain_compute_nodes (\d+) scratch_floats (\d+)", lines[-1])
    phases, attn, plain, scratch = map(int, m.groups())
    L = shape.n_layer
    assert attn == L
    assert phases == 5 * L + 1
    assert scratch == shape.head_dim * (shape.n_head + 2 * shape.n_head_kv)
    plain_ops = [ln.split()[1] for ln in lines if ln.startswith("NODE")]
    # embedding lookup, mask cast, final norm (+ the last layer's row selection when ggml-alloc made fusing it a hazard)
    assert set(plain_ops) <= {"GET_ROWS", "CPY", "RMS_NORM", "MUL", "ADD"}, plain_ops
    assert plain <= 7
    # every projection phase of a layer fuses the RMS norm and has 3 segments; the GLU phase pairs gate and up
    qkv = [ln for ln in lines if ln.startswith("PHASE gemv") and "segs=3" in ln]
    glu = [ln for ln in lines if ln.startswith("PHASE gemv") and "segs=2" in ln and "glu=-1" not in ln]
    assert len(qkv) == L and all("norm=1" in ln for ln in qkv)
    assert len(glu) == L and all("norm=1" in ln for ln in glu)
    if shape.qkv_bias:
        assert all("bias=1" in ln for ln in qkv)


def test_gpt_oss_attention_block_is_planned_as_chain(plan_check, tmp_path):
    """gpt-oss (reference src/models/openai-moe-iswa.cpp): attention sinks, YaRN RoPE, q/k/v/o biases and sliding-window masks do
    not keep the attention block out of the chain: per layer one qkv phase (norm fused, 3 biased segments), one attention
    phase and the wo phase; the MoE part of the layer stays plain nodes (fused in csrc/backend/exec.cpp)."""
    from llama_mi50_b200 import gguf_synth as gs
    shape = gs.SHAPES["tiny-moe-mxfp4"]
    path = str(tmp_path / "tiny-moe.gguf")
    gs.write_gguf(path, shape)
    out = subprocess.run([plan_check, path, REF_DIR], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = out.stdout.splitlines()
    m = re.search(r"summary phases (\d+) attn (\d+) plain_compute_nodes (\d+) scratch_floats (\d+)", lines[-1])
    phases, attn, plain, scratch = map(int, m.groups())
    L = shape.n_layer
    assert attn == L
    assert phases == 3 * L + 1                      # qkv, attention, wo per layer + the logits matmul
    assert scratch == shape.head_dim * (shape.n_head + 2 * shape.n_head_kv)
    qkv = [ln for ln in lines if ln.startswith("PHASE gemv") and "segs=3" in ln]
    assert len(qkv) == L and all("norm=1" in ln and "bias=1" in ln for ln in qkv)
    plain_ops = {ln.split()[1] for ln in lines if ln.startswith("NODE")}
    assert {"MUL_MAT_ID", "ADD_ID", "ARGSORT", "SOFT_MAX", "GLU"} <= plain_ops      # the MoE part: plain nodes for exec.cpp's fusions
    assert "ROPE" not in plain_ops and "FLASH_ATTN_EXT" not in plain_ops and "SET_ROWS" not in plain_ops
