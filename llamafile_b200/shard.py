"""Row (N) sharding of one weight matrix over the GPUs of a box: rank r owns a contiguous range of
output rows, rounded to 128-row tiles like the reference's split buffers (llama.cpp/ggml/src/ggml-cuda.cu:646-674);
activations are replicated; the output slices are all-gathered (NCCL over NVLink on the GPUs, gloo in CPU tests).
Because a row's result does not depend on which other rows are computed with it, the gathered result is bit-identical
to the single-GPU one."""
from __future__ import annotations

TILE = 128


def shard_rows(N: int, rank: int, world: int, tile: int = TILE) -> tuple[int, int]:
    tiles = (N + tile - 1) // tile
    lo = (tiles * rank // world) * tile
    hi = min(N, (tiles * (rank + 1) // world) * tile)
    return lo, max(lo, hi)


def max_shard(N: int, world: int, tile: int = TILE) -> int:
    return max(shard_rows(N, r, world, tile)[1] - shard_rows(N, r, world, tile)[0] for r in range(world))


def gather_rows(local, N: int, rank: int, world: int, group=None, out=None, staging=None):
    """local: [T, hi-lo] tensor of this rank's output rows -> [T, N] on every rank. Uneven splits (vocab 128256 = 1002
    tiles) are padded to the widest shard for the collective and trimmed after."""
    import torch
    import torch.distributed as dist
    T = local.shape[0]
    width = max_shard(N, world)
    lo, hi = shard_rows(N, rank, world)
    if staging is None:
        staging = torch.zeros(T, width, dtype=local.dtype, device=local.device)
    staging[:, : hi - lo] = local
    flat = torch.empty(world * T, width, dtype=local.dtype, device=local.device)   # concatenation form (gloo and nccl)
    dist.all_gather_into_tensor(flat, staging, group=group)
    buf = flat.view(world, T, width)
    if out is None:
        out = torch.empty(T, N, dtype=local.dtype, device=local.device)
    for r in range(world):
        l, h = shard_rows(N, r, world)
        out[:, l:h] = buf[r, :, : h - l]
    return out
