#!/usr/bin/env python3
"""Same-box A/B of decode-chain tuning flags: one session (weights loaded once), one chain per flags value.
usage: ab_chain.py <shape-key> <flags> [<flags> ...]   (flags as in b200_chain_create: see gemv_stream.cu)
Prints tok/s per flags value over two alternating rounds and checks the logits are bit-identical across values."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as graft
graft.load_package()
import torch
from llama_mi50_b200 import gguf_synth as gs
from llama_mi50_b200.session import DecodeSession

key = sys.argv[1]
flag_list = [int(a, 0) for a in sys.argv[2:]] or [0]
N_TOK, WARM = 64, 8
sess = DecodeSession(gs.SHAPES[key], n_ctx=256, fused="chain")
ref = None
for rnd in range(2):
    for fl in flag_list:
        sess.chain_flags, sess.chain = fl, None
        sess.reset()
        for i in range(WARM):
            sess.decode_eager(5 + i, i)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for i in range(WARM, WARM + N_TOK):
            sess.decode_eager(5 + i, i)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / N_TOK
        lg = sess.logits.clone()
        same = "first" if ref is None else ("bit-identical" if torch.equal(lg, ref) else "DIFFERENT")
        if ref is None:
            ref = lg
        print(f"round {rnd} flags {fl:#010x}: {1e3 / ms:8.1f} tok/s  {ms * 1e3:8.1f} us/token  logits {same}", flush=True)
