#!/usr/bin/env python3
"""Row-split parity on N GPUs (torchrun): every rank decodes the same tokens with its row slice of every matmul weight
and an all-gather per matmul; rank 0 also decodes them unsplit.  Criterion: the combine is a
concatenation, so the N-GPU logits equal the 1-GPU logits BIT FOR BIT.
usage: python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/rowsplit_parity.py [shape]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as graft  # noqa: E402

graft.load_package()
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
from llama_mi50_b200 import gguf_synth as gs  # noqa: E402
from llama_mi50_b200.rowsplit import make_gather  # noqa: E402
from llama_mi50_b200.session import DecodeSession  # noqa: E402


def main():
    key = sys.argv[1] if len(sys.argv) > 1 else "tiny-q4_k_m"
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    shape = gs.SHAPES[key]
    tokens = [3, 77, 5, 130, 9, 250, 17, 4]
    split = DecodeSession(shape, n_ctx=256, rank=rank, world=world, gather=make_gather(dist, world, rank), fused=True)
    split.resident_tokens = False
    got = [split.decode_eager(t, i).clone() for i, t in enumerate(tokens)]
    torch.cuda.synchronize()
    ok = True
    if rank == 0:
        one = DecodeSession(shape, n_ctx=256, fused=True)
        one.resident_tokens = False
        for i, t in enumerate(tokens):
            want = one.decode_eager(t, i)
            same = torch.equal(want, got[i])
            ok = ok and same
            print(f"token {i}: row-split x{world} == single GPU: {same}  (max |diff| {(want - got[i]).abs().max().item():.3e})", flush=True)
        print("rowsplit-parity:", "PASS" if ok else "FAIL", flush=True)
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.broadcast(flag, 0)
    rc = 0 if flag.item() == 1 else 1
    torch.cuda.synchronize(); sys.stdout.flush()
    os._exit(rc)      # no communicator teardown: it was observed to hang with captured NCCL kernels alive


if __name__ == "__main__":
    main()
