"""Row split over NVLink peer memory: on every rank the logits of the N-GPU decode chain must
equal those of the single-GPU chain BIT FOR BIT (slices are concatenated, never summed).  Runs tools/peer_chain_check.py
under torchrun on 2 GPUs (eager launches and CUDA-graph replay, including ranks that own 0 rows of a small matrix);
skipped on a box with fewer than 2 GPUs.  The host-side split logic is covered on CPU by tests/test_rowsplit_gloo.py."""
import os
import subprocess
import sys

import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu


def _n_gpus():
    import torch
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


@pytest.mark.parametrize("world", [2])
def test_peer_chain_bit_identical(world):
    if _n_gpus() < world:
        pytest.skip(f"needs {world} GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
           "--master-port", "29731", os.path.join(ROOT, "tools", "peer_chain_check.py"), "tiny,graph"]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=900, cwd=ROOT)
    assert out.returncode == 0, out.stdout[-3000:] + out.stderr[-3000:]
    assert "peer-chain-check: PASS" in out.stdout


def test_plugin_row_split_two_devices(tmp_path):
    """--split-mode row through the plugin on a box with >= 2 GPUs: the slices live on different devices."""
    if _n_gpus() < 2:
        pytest.skip("needs 2 GPUs")
    from conftest import REF_DIR
    from llama_mi50_b200 import gguf_synth as gs
    exe = os.path.join(REF_DIR, "plugin-parity")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/plugin-parity not built")
    path = str(tmp_path / "tiny-q4_k_m.gguf")
    gs.write_gguf(path, gs.SHAPES["tiny-q4_k_m"])
    env = dict(os.environ, B200_PARITY_SPLIT_ROW="1")
    plugin = os.path.join(ROOT, "llama-mi50.cpp_b200", "libggml-b200.so")
    out = subprocess.run([exe, path, REF_DIR, plugin, "20", "12", "2e-3"], capture_output=True, text=True, timeout=600, env=env)
    assert out.returncode == 0, out.stdout[-3000:] + out.stderr[-3000:]
    assert "plugin-parity: PASS" in out.stdout
