#!/usr/bin/env python3
"""Per-CTA view of the decode chain (flags bit 2): when every CTA entered each phase and when its rows were consumed
(%globaltimer).  Prints, per phase kind, the spread of the finish times across the grid: the phase ends for everybody when
the slowest CTA is done."""
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
sess = DecodeSession(gs.SHAPES[key], n_ctx=256, fused="chain")
sess.chain_flags = 4
for i in range(4):
    sess.decode_eager(5, i)
torch.cuda.synchronize()
n = sess.chain.n
L = capi.lib()
L.b200_chain_grid.restype = C.c_int
grid = L.b200_chain_grid()
out = np.zeros(n * grid * 2, np.uint64)
L.b200_chain_debug_read_ctas.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
assert L.b200_chain_debug_read_ctas(sess.chain.h, out.ctypes.data_as(C.c_void_p), n) == 0
d = out.reshape(n, grid, 2).astype(np.int64)
names = ["qkv", "attn", "wo", "gateup", "down"]
t0 = d[0, :, 0].min()
print(f"token {(d[n - 1, :, 1].max() - t0) / 1e3:.1f} us, grid {grid}")
print("phase    entry-spread  busy(min/med/max)      end-spread   slowest CTAs (end - median end, ns)")
acc = {}
for i in range(5, n - 1):          # skip the first layer
    nm = names[i % 5]
    e, f = d[i, :, 0], d[i, :, 1]
    busy = f - e
    acc.setdefault(nm, []).append((e.max() - e.min(), busy.min(), np.median(busy), busy.max(), f.max() - f.min(), f.max() - np.median(f)))
for nm, v in acc.items():
    m = np.array(v, dtype=np.float64).mean(axis=0)
    print(f"{nm:7s} {m[0]:12.0f}  {m[1]:7.0f}/{m[2]:7.0f}/{m[3]:7.0f}  {m[4]:12.0f}   {m[5]:8.0f}")
# which CTAs finish the gate/up phase last, on average
gu = [i for i in range(5, n - 1) if i % 5 == 3]
late = np.zeros(grid)
for i in gu:
    f = d[i, :, 1]
    late += f - np.median(f)
late /= len(gu)
order = np.argsort(-late)
print("gateup: mean lateness vs the median CTA (ns):", ", ".join(f"cta{c}:{late[c]:.0f}" for c in order[:12]), "...", ", ".join(f"cta{c}:{late[c]:.0f}" for c in order[-6:]))
dn = [i for i in range(5, n - 1) if i % 5 == 4]
late = np.zeros(grid)
for i in dn:
    f = d[i, :, 1]
    late += f - np.median(f)
late /= len(dn)
order = np.argsort(-late)
print("down  : mean lateness vs the median CTA (ns):", ", ".join(f"cta{c}:{late[c]:.0f}" for c in order[:12]), "...", ", ".join(f"cta{c}:{late[c]:.0f}" for c in order[-6:]))
