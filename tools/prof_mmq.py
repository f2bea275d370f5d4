#!/usr/bin/env python3
"""A few tensor-core prefill GEMM launches (ffn_up shape of Llama-3-8B, N=512) for ncu (-k regex:mmq_tc_kernel)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as graft
graft.load_package()
import torch
from llama_mi50_b200 import capi, gguf_synth as gs

M, K, N, t = 14336, 4096, 512, 12
raw = gs.make_tensor_bytes(np.random.default_rng(0), t, (K, M), 0.02)
wd = torch.zeros(raw.nbytes + 64, dtype=torch.uint8, device="cuda"); wd[:raw.nbytes] = torch.from_numpy(raw).cuda()
x = torch.randn(N, K, device="cuda"); dst = torch.zeros(N, M, device="cuda")
xb = capi.mul_mat_q(wd, t, K, M, x, dst, act_fmt=0)
for _ in range(4):
    capi.mul_mat_q(wd, t, K, M, x, dst, act_fmt=0, xb=xb)
torch.cuda.synchronize()
print("done")
