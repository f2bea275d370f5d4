#!/usr/bin/env python3
"""Micro-benchmark of the tensor-core prefill GEMM on the Llama-3-8B pp512 shapes (CUDA events, weights rotate past L2)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as graft
graft.load_package()
import torch
from llama_mi50_b200 import capi, gguf_synth as gs

Q4_K, Q6_K = 12, 14
CASES = [("wq_q4k", Q4_K, 4096, 4096), ("wk_q4k", Q4_K, 1024, 4096), ("ffn_up_q4k", Q4_K, 14336, 4096), ("ffn_down_q4k", Q4_K, 4096, 14336),
         ("ffn_down_q6k", Q6_K, 4096, 14336)]
N = int(sys.argv[1]) if len(sys.argv) > 1 else 512
fmt = int(sys.argv[2]) if len(sys.argv) > 2 else 1
rng = np.random.default_rng(0)
for name, t, M, K in CASES:
    raw = gs.make_tensor_bytes(rng, t, (K, M), 0.02)
    copies = max(2, int(200e6 // raw.nbytes) + 1)
    ws = []
    for c in range(copies):
        wd = torch.zeros(raw.nbytes + 64, dtype=torch.uint8, device="cuda"); wd[:raw.nbytes] = torch.from_numpy(raw).cuda(); ws.append(wd)
    x = torch.randn(N, K, device="cuda"); dst = torch.zeros(N, M, device="cuda")
    xb = capi.mul_mat_q(ws[0], t, K, M, x, dst, act_fmt=fmt)
    torch.cuda.synchronize()
    reps = 20
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for r in range(reps):
        capi.mul_mat_q(ws[r % copies], t, K, M, x, dst, act_fmt=fmt, xb=xb)
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / reps
    print(f"{name:14s} M={M:6d} K={K:6d} N={N}: {us:9.1f} us  {2.0 * M * K * N / us / 1e6:8.1f} TFLOP/s", flush=True)
