"""End-to-end parity of the ggml backend plugin THROUGH the reference's own public API (llama.h): the same synthetic
GGUF decodes the same tokens on the reference CPU backend and on the plugin (all layers offloaded; decode graphs run
as decode chains, csrc/backend/chain_plan.cpp).  oracle/_ref/plugin-parity compares the logits of every step;
tolerance NMSE <= 2e-3 (see tests/test_gpu_session.py for the budget)."""
import os
import subprocess

import pytest

from conftest import REF_DIR, ROOT

pytestmark = pytest.mark.gpu
PLUGIN = os.path.join(ROOT, "llama-mi50.cpp_b200", "libggml-b200.so")


@pytest.mark.parametrize("chain", [True, False])
@pytest.mark.parametrize("key", ["tiny-q4_0", "tiny-q4_k_m", "tiny-q8_0"])
def test_plugin_matches_cpu_backend(key, chain, tmp_path):
    exe = os.path.join(REF_DIR, "plugin-parity")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/plugin-parity not built")
    from llama_mi50_b200 import gguf_synth as gs
    path = str(tmp_path / f"{key}.gguf")
    gs.write_gguf(path, gs.SHAPES[key])
    env = dict(os.environ)
    if not chain:
        env["GGML_B200_DISABLE_CHAIN"] = "1"
    out = subprocess.run([exe, path, REF_DIR, PLUGIN, "5", "12", "2e-3"], capture_output=True, text=True, timeout=600, env=env)
    assert out.returncode == 0, out.stdout[-3000:] + out.stderr[-3000:]
    assert "plugin-parity: PASS" in out.stdout


@pytest.mark.parametrize("key", ["tiny-q4_k_m", "tiny-q8_0"])
def test_plugin_row_split_matches_cpu_backend(key, tmp_path):
    """--split-mode row: every 2-D matmul weight lives in the backend's split buffer type (csrc/backend/split_buffer.cpp;
    reference ggml-cuda.cu:776-1107) and is multiplied slice by slice (exec.cpp run_mul_mat_split).  On one GPU the single
    device owns every row; on a multi-GPU box the slices run on their devices and store into dst over peer memory."""
    exe = os.path.join(REF_DIR, "plugin-parity")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/plugin-parity not built")
    from llama_mi50_b200 import gguf_synth as gs
    path = str(tmp_path / f"{key}.gguf")
    gs.write_gguf(path, gs.SHAPES[key])
    env = dict(os.environ, B200_PARITY_SPLIT_ROW="1", GGML_B200_PRINT_STATS="1")
    out = subprocess.run([exe, path, REF_DIR, PLUGIN, "20", "8", "2e-3"], capture_output=True, text=True, timeout=600, env=env)
    assert out.returncode == 0, out.stdout[-3000:] + out.stderr[-3000:]
    assert "plugin-parity: PASS" in out.stdout
