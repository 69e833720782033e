"""Multi-GPU plumbing: one process per GPU (torch.distributed, NCCL over NVLink; gloo in CPU tests).

The path shards by image tiles (rays are independent):
  * the flattened scene blob is produced once on rank 0 and replicated with ONE broadcast;
  * every rank renders the 8x4-pixel tiles it owns (static chunked interleave, rt_render_params
    tile_rank/tile_world/tile_chunk) — no data-path collective during traversal or shading;
  * the framebuffer is gathered to rank 0 (the only real exchange step of the path).
Everything here is device-agnostic (works on CPU tensors with gloo) so that the N > 1 host logic is
covered by world_size-2 CPU tests.
"""
from __future__ import annotations

import torch
import torch.distributed as dist


def broadcast_blob(blob: bytes | None, device, src: int = 0) -> torch.Tensor:
    """Rank `src` passes the exported blob bytes; every rank returns a uint8 tensor on `device`."""
    rank = dist.get_rank()
    n = torch.zeros(1, dtype=torch.int64, device=device)
    if rank == src:
        n[0] = len(blob)
    dist.broadcast(n, src)
    if rank == src:
        t = torch.frombuffer(bytearray(blob), dtype=torch.uint8).to(device)
    else:
        t = torch.empty(int(n.item()), dtype=torch.uint8, device=device)
    dist.broadcast(t, src)
    return t


def default_chunk(world: int) -> int:
    """Tiles per ownership run.  One image row of 8x4 tiles at 1920 px is 240 tiles; runs of 16 tiles
    keep each rank's rays coherent while interleaving finely enough to balance cheap and expensive
    image regions across ranks."""
    return 16


def tile_share(rank: int, world: int, chunk: int | None = None):
    return (rank, world, chunk or default_chunk(world))


def gather_frame(local: torch.Tensor, scratch: torch.Tensor | None = None, dst: int = 0) -> torch.Tensor:
    """Framebuffer gather to rank `dst`.  `local` holds this rank's tiles and zeros elsewhere, so a
    byte-wise SUM over ranks is the union (each pixel is written by exactly one rank).  Returns the
    full frame on `dst` (in `scratch` when given, so `local` keeps its zeros for the next frame)."""
    if dist.get_rank() == dst:
        out = scratch if scratch is not None else torch.empty_like(local)
        out.copy_(local)
        dist.reduce(out, dst, op=dist.ReduceOp.SUM)
        return out
    dist.reduce(local, dst, op=dist.ReduceOp.SUM)
    return local


def max_over_ranks(value: float, device) -> float:
    t = torch.tensor([value], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())
