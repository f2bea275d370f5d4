#!/usr/bin/env python3
"""Micro-benchmark of flash_attn_ext decode (b200_flash_attn_ext, csrc/kernels/fattn.cu) on the BASELINE.json shapes.
KV caches of several "layers" are rotated so that every launch streams from HBM (one layer's K+V at 32k is 71-134 MB,
L2 is 126 MB).  Roofline: algorithmic bytes = K + V bytes of one layer / time vs the measured
HBM copy peak.  usage: bench_fattn.py [n_kv ...]"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as graft
graft.load_package()
import torch
from llama_mi50_b200 import capi

peak = 6548.8
try:
    peak = float(json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    pass
CASES = [("llama3-8b f16", 128, 32, 8, capi.F16), ("llama3-8b q8_0", 128, 32, 8, capi.Q8_0),
         ("qwen2.5-32b q8_0 (C3)", 128, 40, 8, capi.Q8_0), ("qwen2.5-32b f16", 128, 40, 8, capi.F16),
         ("llama3-70b f16", 128, 64, 8, capi.F16), ("gpt-oss q8_0", 64, 64, 8, capi.Q8_0)]
kvs = [int(a) for a in sys.argv[1:]] or [4096, 32768]
if os.environ.get("FATTN_CASE"):
    CASES = [c for c in CASES if os.environ["FATTN_CASE"] in c[0]]
for n_kv in kvs:
    for name, D, H, Hkv, t in CASES:
        row = D * 2 if t == capi.F16 else D // 32 * 34
        nbytes = Hkv * n_kv * row
        layers = max(2, int(400e6 // (2 * nbytes)) + 1)
        ks = [torch.randint(0, 60, (nbytes,), dtype=torch.uint8, device="cuda") for _ in range(layers)]
        vs = [torch.randint(0, 60, (nbytes,), dtype=torch.uint8, device="cuda") for _ in range(layers)]
        q = torch.randn(H, 1, D, device="cuda")
        mask = torch.zeros(1, n_kv, dtype=torch.float16, device="cuda")
        for i in range(3):
            capi.flash_attn(q, ks[i % layers], vs[i % layers], t, n_kv, Hkv, mask=mask, scale=D ** -0.5)
        torch.cuda.synchronize()
        reps = 20
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(reps):
            capi.flash_attn(q, ks[i % layers], vs[i % layers], t, n_kv, Hkv, mask=mask, scale=D ** -0.5)
        e1.record(); torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / reps          # includes the scratch allocation of the ctypes wrapper
        gbs = 2 * nbytes / us / 1e3
        print(f"n_kv {n_kv:6d} {name:24s} D={D} H={H} H_kv={Hkv}: {us:8.1f} us  {gbs:7.1f} GB/s  {gbs / peak:5.3f} of measured HBM peak", flush=True)
        del ks, vs
