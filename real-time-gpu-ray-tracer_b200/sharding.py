"""Image-tile sharding of a frame over N ranks — the host-side arithmetic shared by bench.py and
the multi-process tests.  Mirrors compute_local_rows / global_row in csrc/api.cu and csrc/kernels_wave.cu:
tile t (tile_rows scan-lines) belongs to rank t % world; a rank's planes are packed (local tile k = global tile
k * world + rank) and padded to whole tiles.

Frame sharding of the particle animation (config 3): frame i belongs to rank i % world; a rank reads,
rebuilds and renders its own frames and nothing is exchanged per frame."""
from __future__ import annotations

import numpy as np


def local_rows(height: int, rank: int, world: int, tile_rows: int) -> int:
    if world == 1:
        return height
    tiles = (height + tile_rows - 1) // tile_rows
    local_tiles = (tiles - rank + world - 1) // world if rank < tiles else 0
    return local_tiles * tile_rows


def global_row(ly: int, rank: int, world: int, tile_rows: int) -> int:
    return ((ly // tile_rows) * world + rank) * tile_rows + (ly % tile_rows)


def owned_rows(height: int, rank: int, world: int, tile_rows: int) -> list[int]:
    """Global scan-lines rendered by `rank`, in the order of its packed plane (rows past the image are dropped)."""
    rows = [global_row(ly, rank, world, tile_rows) for ly in range(local_rows(height, rank, world, tile_rows))]
    return [gy for gy in rows if gy < height]


def assemble(planes, height: int, row_bytes: int, world: int, tile_rows: int):
    """De-interleaves the gathered packed planes (one 1-D uint8 tensor/array per rank, padded to equal length) into
    the full frame [height, row_bytes].  Works on torch tensors (device or host) and numpy arrays alike."""
    if world == 1:                                               # a single rank's plane is the frame itself (not padded to whole tiles)
        return planes[0][:height * row_bytes].reshape(height, row_bytes)
    if type(planes[0]).__module__.split(".")[0] == "torch":      # (numpy 2 arrays have .device too)
        import torch
        stack = torch.stack(list(planes))
        return stack.view(world, -1, tile_rows, row_bytes).permute(1, 0, 2, 3).reshape(-1, row_bytes)[:height].contiguous()
    stack = np.stack(list(planes))
    return np.ascontiguousarray(stack.reshape(world, -1, tile_rows, row_bytes).transpose(1, 0, 2, 3).reshape(-1, row_bytes)[:height])


def padded_plane_bytes(height: int, row_bytes: int, world: int, tile_rows: int) -> int:
    """Bytes of the largest rank's packed plane (rank 0): gather buffers are padded to this size."""
    return local_rows(height, 0, world, tile_rows) * row_bytes


def frames_of_rank(n_frames: int, rank: int, world: int) -> list[int]:
    """Animation frames rendered by `rank`, in order: frame i -> rank i % world."""
    return list(range(rank, n_frames, world))


def interleave_frames(per_rank, n_frames: int):
    """Inverse of frames_of_rank: per_rank[q][j] is rank q's j-th result; returns the results in frame order."""
    world = len(per_rank)
    out = [per_rank[i % world][i // world] for i in range(n_frames)]
    assert all(len(per_rank[q]) >= len(frames_of_rank(n_frames, q, world)) for q in range(world))
    return out
