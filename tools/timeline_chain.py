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
# stamps are SM cycle counters (clock64): converted to ns at the SM clock the bench reports under load (1965 MHz)
d = (out.reshape(n, 2, 8).astype(np.float64) / 1.965).astype(np.int64)
# stamps per phase and CTA (0 = CTA 0, 1 = last CTA): [0] phase entered (descriptor staged, every warp of the CTA has left the
# previous phase), [5] inputs present (tagged words settled: the dependency wait + the L2 round trip), [6] normalised + quantised,
# [1] prologue done (CTA barrier), [2] rows consumed (warp 0)
names = ["qkv", "attn", "wo", "gateup", "down"]
for cta in (0, 1):
    t0 = d[0, cta, 0]
    rows = {}
    for i in range(n):
        nm = names[i % 5] if i < n - 1 else "logits"
        a = d[i, cta]
        a1 = a[1] if a[1] else a[0]
        s5 = a[5] if a[5] else a[0]
        s6 = a[6] if a[6] else a1
        if nm == "attn" and a[5] and i == 6: print("attn stamps rel. to entry:", [int(x - a[0]) for x in (a[5], a[3], a[4], a[6], a[2])])
        if nm == "attn" and a[5]:      # attention: [5] q/k/v present and rotated, [6] warp 0 through its positions
            a1 = a[6]; s6 = a[5]       # columns: wait+load = inputs, norm+quant = positions, consume = combine + store
        nxt = d[i + 1, cta, 0] if i + 1 < n else a[2]
        # [3] ns warp 0 waited for weight bytes, [4] chunks it consumed, [7] when the producer requested the phase's first chunk
        lead = (a[0] - a[7]) if a[7] else 0
        rows.setdefault(nm, []).append((s5 - a[0], s6 - s5, a1 - s6, a[2] - a1, nxt - a[2], a[3], a[4], lead))
    print(f"== CTA {'0' if cta == 0 else 'last'}: token {(d[n-1, cta, 2] - t0) / 1e3:.1f} us; mean ns per phase kind (first layer skipped)")
    print("phase   wait+load  norm+quant  barrier   consume  to-next     total | of consume: data-wait  chunks | first chunk requested this long before phase entry")
    layer = 0.0
    for nm, v in rows.items():
        v = np.array(v[1:] if len(v) > 1 else v, dtype=np.float64)
        m = v.mean(axis=0)
        if nm != "logits": layer += m[:5].sum()
        print(f"{nm:7s} {m[0]:9.0f} {m[1]:11.0f} {m[2]:8.0f} {m[3]:9.0f} {m[4]:8.0f} {m[:5].sum():9.0f} | {m[5]:21.0f} {m[6]:7.1f} | {m[7]:10.0f}")
    print(f"per layer {layer / 1e3:.2f} us")
