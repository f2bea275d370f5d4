#!/usr/bin/env python3
"""A few whole-token decode-chain launches of the Llama-3-8B Q4_K_M shape, for ncu (-k regex:gemv_stream_kernel)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as graft
graft.load_package()
import torch
from llama_mi50_b200 import gguf_synth as gs
from llama_mi50_b200.session import DecodeSession

key = sys.argv[1] if len(sys.argv) > 1 else "llama3-8b-q4_k_m"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 6
sess = DecodeSession(gs.SHAPES[key], n_ctx=256, fused="chain")
for i in range(n):
    sess.decode_eager(5 + i, i)
torch.cuda.synchronize()
print("done", n, "tokens")
