#!/usr/bin/env python3
"""Per-chunk timeline of CTA 0 of the streaming GEMV (flags bit 2): clock64 stamps of producer / consumer warp 0 / finalizer."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as graft
graft.load_package()
import torch
from llama_mi50_b200 import capi, gguf_synth as gs

case = sys.argv[1] if len(sys.argv) > 1 else "out"
flags = int(sys.argv[2], 0) if len(sys.argv) > 2 else 0
K, t, M = {"out": (4096, 14, 128256), "wo": (4096, 12, 4096), "gu": (4096, 12, 14336)}[case]
rng = np.random.default_rng(0)
w = gs.make_tensor_bytes(rng, t, (K, M), 0.02)
wd = torch.zeros(w.nbytes + 64, dtype=torch.uint8, device="cuda"); wd[:w.nbytes] = torch.from_numpy(w).cuda()
x = torch.randn(K, device="cuda"); dst = torch.zeros(M, device="cuda")
dbg = torch.zeros(3 * 4096 * 4, dtype=torch.int64, device="cuda")
for i in range(3):
    capi.gemv_fused(x, [(wd, t, M, None, dst)], xq_out=dbg, flags=flags | 4)
torch.cuda.synchronize()
d = dbg.cpu().numpy().reshape(3, 4096, 4)
n = int((d[0, :, 0] != 0).sum())
t0 = d[0, 0, 0]
print("chunks", n)
print(" n | prod: wait-done issue-done | cons0: full-seen done arrived | fin: rfull-seen done   (clk since first producer stamp)")
for i in list(range(min(n, 14))) + list(range(max(14, n - 4), n)):
    r = lambda a: int(a - t0) if a else -1
    print(f"{i:3d} | {r(d[0,i,0]):7d} {r(d[0,i,1]):7d} | {r(d[1,i,0]):7d} {r(d[1,i,1]):7d} {r(d[1,i,2]):7d} | {r(d[2,i,0]):7d} {r(d[2,i,1]):7d}")
