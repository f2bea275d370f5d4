#!/usr/bin/env python3
"""Per-phase timeline of the decode chain (flags bit 2): globaltimer stamps of CTA 0 and the last CTA."""
import ctypes as C
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as graft
graft.load_package()
import torch
from llama_mi50_b200 import capi, gguf_synth as gs
from llama_mi50_b200.session import DecodeSession

key = sys.argv[1] if len(sys.argv) > 1 else "llama3-8b-q4_k_m"
flags = int(sys.argv[2], 0) if len(sys.argv) > 2 else 0
sess = DecodeSession(gs.SHAPES[key], n_ctx=256, fused="chain")
sess.chain_flags = flags | 4
for i in range(4):
    sess.decode_eager(5, i)
torch.cuda.synchronize()
n = sess.chain.n
out = np.zeros(n * 16, np.uint64)
L = capi.lib()
L.b200_chain_debug_read.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
assert L.b200_chain_debug_read(sess.chain.h, out.ctypes.data_as(C.c_void_p), n) == 0
d = out.reshape(n, 2, 8).astype(np.int64)
t0 = d[0, 0, 0]
names = ["qkv", "attn", "wo", "gateup", "down"]
print("phase      | cta0: barrier-passed  +prologue  +consume  +publish | last cta: same      (ns; deltas within phase)   phase period")
prev = t0
tot = {}
for i in range(n):
    nm = names[i % 5] if i < n - 1 else "logits"
    a, b = d[i, 0], d[i, 1]
    period = a[0] - prev; prev = a[0]
    tot.setdefault(nm, []).append((period, a[1] - a[0], a[2] - a[1], a[3] - a[2], b[3] - a[3], a[6] - a[2], a[4] - a[6], a[5] - a[4], a[3] - a[5]))
    if i < 12 or i > n - 4:
        print(f"{i:3d} {nm:7s}| {a[0]-t0:9d} {a[1]-a[0]:8d} {a[2]-a[1]:8d} {a[3]-a[2]:8d} | {b[0]-t0:9d} {b[1]-b[0]:8d} {b[2]-b[1]:8d} {b[3]-b[2]:8d}   {period:8d}")
print("total ns", d[n - 1, 0, 3] - t0)
for nm, v in tot.items():
    v = np.array(v[1:] if len(v) > 1 else v)
    print(f"{nm:7s} mean: wait-for-barrier-since-prev-phase-start {v[:,0].mean():8.0f}  prologue {v[:,1].mean():7.0f}  consume {v[:,2].mean():7.0f}  publish {v[:,3].mean():7.0f}  last-vs-first cta publish skew {v[:,4].mean():7.0f} | last consumer warp after warp0 {v[:,5].mean():6.0f}  finalizer done after that {v[:,6].mean():6.0f}  fence {v[:,7].mean():6.0f}  atomic {v[:,8].mean():6.0f}")
