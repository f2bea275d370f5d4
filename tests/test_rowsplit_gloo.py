"""Row split host logic on CPU: world_size-2 gloo processes.
 * split_bounds / row_split follow the reference's get_row_split (ggml-cuda.cu:791-804): cumulative equal fractions,
   boundaries rounded DOWN to 128 rows, the last rank takes the remainder;
 * the combine is a concatenation (all-gather) — equal slices directly, unequal slices through the padded staging —
   so the gathered vector is bit-identical on every rank to the unsplit one."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import conftest  # noqa: F401  (registers the package alias; also in the spawned workers)
from llama_mi50_b200.rowsplit import make_gather, split_bounds
from llama_mi50_b200.session import row_split


def test_split_bounds_match_reference_rule():
    # reference: row_low = round_down(nrows * split[id], 128) (0 for id 0), row_high = next rank's low or nrows
    for n_rows in (128, 1024, 4096, 14336, 128256, 1000, 32000):
        for world in (1, 2, 4, 8):
            b = split_bounds(n_rows, world)
            assert b[0] == 0 and b[-1] == n_rows and len(b) == world + 1
            for r in range(1, world):
                exact = n_rows * r // world
                assert b[r] == exact - exact % 128
            assert all(b[i] <= b[i + 1] for i in range(world))
            for r in range(world):
                assert row_split(n_rows, world, r) == (b[r], b[r + 1])


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close()
    return p


def _worker(rank, world, port, sizes, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        gather = make_gather(dist, world, rank)
        ok = True
        for n in sizes:
            ref = torch.from_numpy(np.random.default_rng(n).standard_normal(n).astype(np.float32))
            lo, hi = row_split(n, world, rank)
            local = torch.zeros(max(hi - lo, 1) + 7)       # the session's slice buffer is larger than the slice
            local[:hi - lo] = ref[lo:hi]
            full = torch.full((n,), float("nan"))
            gather(local, full, lo, hi)
            ok = ok and torch.equal(full, ref)
            gather(local, full, lo, hi)                    # staging buffers are reused
            ok = ok and torch.equal(full, ref)
        out[rank] = ok
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2])
def test_gather_is_a_bit_exact_concat(world):
    # 4096/14336: equal 128-aligned slices; 1000/32000/128256: unequal slices (padded staging path)
    sizes = [4096, 14336, 1000, 32000, 128256]
    ctx = mp.get_context("spawn")
    out = ctx.Manager().dict()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, sizes, out)) for r in range(world)]
    for p in procs: p.start()
    for p in procs: p.join(120)
    assert all(p.exitcode == 0 for p in procs)
    assert dict(out) == {r: True for r in range(world)}
