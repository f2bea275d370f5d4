#!/usr/bin/env python3
"""Micro-benchmark of the streaming fused GEMV (b200_gemv_fused) on the Llama-3-8B decode shapes.
Weights rotate through enough copies to exceed L2 (126 MB); CUDA events on the launching stream.
usage: python tools/bench_gemv.py [--reps 50] [--case NAME] [--n-sm N]"""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as graft  # noqa: E402

graft.load_package()
import torch  # noqa: E402
from llama_mi50_b200 import capi, gguf_synth as gs  # noqa: E402

Q4_K, Q6_K, Q4_0, Q8_0, MXFP4 = 12, 14, 2, 8, 39
# name: (K, [(type, M)...], norm, glu, bias)
CASES = {
    "qkv_q4k":      (4096, [(Q4_K, 4096), (Q4_K, 1024), (Q4_K, 1024)], True, False, False),
    "qkv_q4k_q6k":  (4096, [(Q4_K, 4096), (Q4_K, 1024), (Q6_K, 1024)], True, False, False),
    "wo_q4k":       (4096, [(Q4_K, 4096)], False, False, True),
    "gateup_q4k":   (4096, [(Q4_K, 14336), (Q4_K, 14336)], True, True, False),
    "down_q4k":     (14336, [(Q4_K, 4096)], False, False, True),
    "down_q6k":     (14336, [(Q6_K, 4096)], False, False, True),
    "out_q6k":      (4096, [(Q6_K, 128256)], True, False, False),
    "ffn_q4_0":     (2048, [(Q4_0, 5632), (Q4_0, 5632)], True, True, False),
    "ffn_q8_0":     (5120, [(Q8_0, 27648), (Q8_0, 27648)], True, True, False),
    "moe_mxfp4":    (2880, [(MXFP4, 2880)], False, False, True),
}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=50)
    ap.add_argument("--case", default=None)
    ap.add_argument("--n-sm", type=int, default=0)
    ap.add_argument("--copies", type=int, default=0)
    ap.add_argument("--flags", type=lambda x: int(x, 0), default=0)
    a = ap.parse_args()
    L = capi.lib()
    L.b200_gemv_fused.argtypes = [capi.C.c_void_p, capi.C.c_void_p]
    rng = np.random.default_rng(0)
    for name, (K, segs, norm, glu, bias) in CASES.items():
        if a.case and name != a.case:
            continue
        nbytes = sum(gs.row_bytes(t, K) * M for t, M in segs)
        copies = a.copies or max(2, int(300e6 // nbytes) + 1)
        x = torch.randn(K, device="cuda")
        nw = torch.ones(K, device="cuda") if norm else None
        keep, calls, base = [], [], []
        for c in range(copies):
            cs = []
            for i, (t, M) in enumerate(segs):
                w = gs.make_tensor_bytes(rng, t, (K, M), 0.02) if c == 0 else None
                if c == 0:
                    wd = torch.zeros(w.nbytes + 64, dtype=torch.uint8, device="cuda"); wd[:w.nbytes] = torch.from_numpy(w).cuda()
                    base.append(wd)
                else:
                    wd = base[i].clone(); keep.append(wd)
                dst = None if (glu and i == 0) else torch.zeros(M, device="cuda")
                b = torch.randn(M, device="cuda") if bias else None
                keep += [dst, b]
                cs.append((wd, t, M, b, dst))
            calls.append(capi.gemv_fused_args(x, cs, norm_w=nw, eps=1e-5, glu_op=capi.GLU_SWIGLU if glu else capi.GLU_NONE,
                                              n_sm=a.n_sm, flags=a.flags, keep=keep))
        st = torch.cuda.current_stream().cuda_stream
        for i in range(copies):
            capi.check(L.b200_gemv_fused(capi.C.byref(calls[i]), st), name)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for r in range(a.reps):
            L.b200_gemv_fused(capi.C.byref(calls[r % copies]), st)
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / a.reps
        print(f"{name:14s} K={K:6d} bytes={nbytes/1e6:8.2f} MB  {us:8.2f} us/launch  {nbytes/us/1e3:8.1f} GB/s  ({copies} copies)", flush=True)


if __name__ == "__main__":
    main()
