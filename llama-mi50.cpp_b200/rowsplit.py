"""Row-split (tensor-parallel) plumbing for the decode session: SURVEY.md section 8(e).

The reference combines the row slices of a split matmul with peer copies into the main device's
dst (ggml-cuda.cu:1726-1745): a concatenation, never a sum.  Here every rank needs the full
activation for the next op, so the combine is an all-gather of f32 row slices over NCCL/NVLink.
Slices follow get_row_split (ggml-cuda.cu:791-804): boundaries rounded down to 128 rows, the last
rank takes the remainder, so slices may be unequal; unequal slices are gathered through a padded
staging buffer.
"""
from __future__ import annotations


def split_bounds(n_rows: int, world: int, round_to: int = 128):
    b = []
    for r in range(world + 1):
        if r >= world:
            b.append(n_rows)
        else:
            x = n_rows * r // world
            b.append(x - x % round_to)
    return b


def make_gather(dist, world: int, rank: int, group=None):
    """Returns gather(local[n_loc], full[n], lo, hi): full[lo_r:hi_r] = local_r for every rank r."""
    import torch
    staging = {}

    def gather(local, full, lo, hi):
        n = full.numel()
        b = split_bounds(n, world)
        sizes = [b[r + 1] - b[r] for r in range(world)]
        assert (lo, hi) == (b[rank], b[rank + 1]) and local.numel() >= hi - lo
        if len(set(sizes)) == 1:
            dist.all_gather_into_tensor(full, local[:hi - lo], group=group)
            return
        mx = max(sizes)
        key = (n, full.device)
        if key not in staging:
            staging[key] = (torch.zeros(mx, dtype=full.dtype, device=full.device),
                            torch.zeros(world * mx, dtype=full.dtype, device=full.device))
        send, recv = staging[key]
        send[:hi - lo].copy_(local[:hi - lo])
        dist.all_gather_into_tensor(recv, send, group=group)
        for r in range(world):
            full[b[r]:b[r + 1]].copy_(recv[r * mx:r * mx + sizes[r]])

    return gather
