#!/usr/bin/env python3
"""Staged multi-GPU diagnostic (torchrun): every stage logs to gpurun_out/mgpu_diag_rank<r>.log before and after."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as graft
graft.load_package()
import torch, torch.distributed as dist
from llama_mi50_b200 import gguf_synth as gs
from llama_mi50_b200.rowsplit import make_gather
from llama_mi50_b200.session import DecodeSession

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
os.makedirs("gpurun_out", exist_ok=True)
logf = open(f"gpurun_out/mgpu_diag_rank{rank}.log", "w")
def log(*a):
    print(f"[{time.strftime('%H:%M:%S')}] r{rank}:", *a, file=logf, flush=True)
    print(f"r{rank}:", *a, flush=True)

torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
log("pg up")
stages = sys.argv[1].split(",") if len(sys.argv) > 1 else ["eager", "graph", "big"]
tokens = [3, 77, 5, 130, 9, 250]
try:
    if "eager" in stages or "graph" in stages:
        shape = gs.SHAPES["tiny-q4_k_m"]
        for fused in (True, False):
            sp = DecodeSession(shape, n_ctx=256, rank=rank, world=world, gather=make_gather(dist, world, rank), fused=fused)
            sp.resident_tokens = False
            log(f"tiny session built fused={fused}")
            got = [sp.decode_eager(t, i).clone() for i, t in enumerate(tokens)]
            torch.cuda.synchronize(); log("eager split decode done")
            one = DecodeSession(shape, n_ctx=256, fused=fused); one.resident_tokens = False
            want = [one.decode_eager(t, i).clone() for i, t in enumerate(tokens)]
            torch.cuda.synchronize()
            log("eager parity:", [bool(torch.equal(a, b)) for a, b in zip(want, got)], "maxdiff", max((a - b).abs().max().item() for a, b in zip(want, got)))
            if "graph" in stages:
                sp2 = DecodeSession(shape, n_ctx=256, rank=rank, world=world, gather=make_gather(dist, world, rank), fused=fused)
                sp2.resident_tokens = False
                log("capturing"); sp2.capture(256); log("captured")
                got2 = [sp2.decode(t).clone() for t in tokens]
                torch.cuda.synchronize()
                log("graph parity:", [bool(torch.equal(a, b)) for a, b in zip(want, got2)])
    if "big" in stages:
        shape = gs.SHAPES["llama3-8b-q4_k_m"]
        sp = DecodeSession(shape, n_ctx=256, rank=rank, world=world, gather=make_gather(dist, world, rank), fused=True)
        log("8B session built")
        for i in range(3): sp.decode_eager(5, i)
        torch.cuda.synchronize(); log("8B eager 3 tokens done")
        t0 = time.time()
        for i in range(20): sp.decode_eager(5, 3 + i)
        torch.cuda.synchronize(); log(f"8B eager: {20 / (time.time() - t0):.1f} tok/s")
        log("capturing 8B"); sp.capture(256); log("captured 8B")
        dist.barrier(); torch.cuda.synchronize()
        t0 = time.time()
        for i in range(64): sp.graph.replay()
        torch.cuda.synchronize(); log(f"8B graph replay: {64 / (time.time() - t0):.1f} tok/s")
except Exception as e:
    import traceback
    log("EXCEPTION", repr(e)); log(traceback.format_exc())
    raise
finally:
    log("exiting")
dist.barrier(); torch.cuda.synchronize(); logf.flush(); sys.stdout.flush()
os._exit(0)
