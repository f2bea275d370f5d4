#!/usr/bin/env python3
"""Row split over NVLink peer memory (torchrun, N GPUs of one node): parity and throughput of the decode chain.

Every rank runs the chain over its row slices; the GEMV epilogues store into every rank's activation buffers and the
in-kernel barriers span the ranks (include/b200_kernels.h, b200_chain_create_peer).  Criterion:
outputs are concatenated, never summed, so on EVERY rank the logits equal those of the single-GPU chain BIT FOR BIT.
Stages log to gpurun_out/peer_rank<r>.log before and after, so a stuck stage is visible from the files.

usage: python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/peer_chain_check.py [stages]
stages: comma list of tiny, graph, big (default all)"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as graft  # noqa: E402

graft.load_package()
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
from llama_mi50_b200 import gguf_synth as gs  # noqa: E402
from llama_mi50_b200.session import DecodeSession  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
os.makedirs("gpurun_out", exist_ok=True)
logf = open(f"gpurun_out/peer_rank{rank}.log", "w")


def log(*a):
    print(f"[{time.strftime('%H:%M:%S')}] r{rank}:", *a, file=logf, flush=True)
    print(f"r{rank}:", *a, flush=True)


def main():
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    log("process group up, world", world)
    stages = sys.argv[1].split(",") if len(sys.argv) > 1 else ["tiny", "graph", "big"]
    tokens = [3, 77, 5, 130, 9, 250, 17, 4]
    ok = True
    for key in (["tiny-q4_k_m", "tiny-q8_0"] if "tiny" in stages else []):
        shape = gs.SHAPES[key]
        sp = DecodeSession(shape, n_ctx=256, rank=rank, world=world, fused="chain", peer_dist=dist)
        sp.resident_tokens = False
        log(key, "peer session built; arena deltas", [hex(d) for d in sp.arena.delta[:world]], "total CTAs", sp.arena.total_ctas)
        got = [sp.decode_eager(t, i).clone() for i, t in enumerate(tokens)]
        torch.cuda.synchronize()
        log(key, "eager peer decode done")
        one = DecodeSession(shape, n_ctx=256, fused="chain")
        one.resident_tokens = False
        want = [one.decode_eager(t, i).clone() for i, t in enumerate(tokens)]
        torch.cuda.synchronize()
        same = [bool(torch.equal(a, b)) for a, b in zip(want, got)]
        ok = ok and all(same)
        log(key, "eager parity vs single-GPU chain:", same, "max |diff|", max((a - b).abs().max().item() for a, b in zip(want, got)))
        if "graph" in stages:
            dist.barrier()
            sp2 = DecodeSession(shape, n_ctx=256, rank=rank, world=world, fused="chain", peer_dist=dist)
            sp2.resident_tokens = False
            log(key, "capturing"); sp2.capture(256); log(key, "captured")
            got2 = [sp2.decode(t).clone() for t in tokens]
            torch.cuda.synchronize()
            same = [bool(torch.equal(a, b)) for a, b in zip(want, got2)]
            ok = ok and all(same)
            log(key, "graph parity:", same)
    if "big" in stages:
        key = "llama3-8b-q4_k_m"
        shape = gs.SHAPES[key]
        dist.barrier()
        sp = DecodeSession(shape, n_ctx=256, rank=rank, world=world, fused="chain", peer_dist=dist)
        sp.resident_tokens = False
        sp.chain_flags = int(os.environ.get("CHAIN_FLAGS", "0"), 0)
        log("8B peer session built, chain flags", sp.chain_flags)
        got = [sp.decode_eager(t, i).clone() for i, t in enumerate(tokens[:4])]
        torch.cuda.synchronize(); log("8B eager peer decode done")
        one = DecodeSession(shape, n_ctx=256, fused="chain")
        one.resident_tokens = False
        want = [one.decode_eager(t, i).clone() for i, t in enumerate(tokens[:4])]
        torch.cuda.synchronize()
        same = [bool(torch.equal(a, b)) for a, b in zip(want, got)]
        ok = ok and all(same)
        log("8B parity vs single-GPU chain:", same)
        one.capture(256)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for _ in range(16): one.graph.replay()
        torch.cuda.synchronize(); e0.record()
        for _ in range(128): one.graph.replay()
        e1.record(); torch.cuda.synchronize()
        log(f"8B single-GPU chain on this rank alone (other ranks busy too): {128 / (e0.elapsed_time(e1) / 1e3):.1f} tok/s")
        del one
        dist.barrier()
        sp.resident_tokens = True
        sp.capture(256); log("8B peer chain captured")
        for _ in range(16): sp.graph.replay()
        dist.barrier(); torch.cuda.synchronize()
        for rep in range(3):
            sp.reset()
            e0.record()
            for _ in range(128): sp.graph.replay()
            e1.record(); torch.cuda.synchronize()
            t = torch.tensor([e0.elapsed_time(e1)], device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            log(f"8B peer chain x{world}: {128 / (t.item() / 1e3):.1f} tok/s (tg128, max over ranks)")
    log("peer-chain-check:", "PASS" if ok else "FAIL")
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    rc = 0 if flag.item() == 1 else 1
    torch.cuda.synchronize(); sys.stdout.flush(); logf.flush()
    os._exit(rc)      # no communicator teardown: it was observed to hang with captured NCCL kernels alive


if __name__ == "__main__":
    try:
        main()
    except Exception as e:  # noqa: BLE001
        import traceback
        log("EXCEPTION", repr(e)); log(traceback.format_exc())
        logf.flush()
        os._exit(3)
