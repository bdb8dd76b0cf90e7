"""Multi-GPU plumbing: one process per GPU (torch.distributed), no data-path collective except the
gather of finished framebuffer pieces to rank 0 (NCCL over NVLink / NVSwitch on GPUs, gloo in the CPU tests).

The reference's pixel loop is a pure function of the pixel (simple_raytracer.cpp:511-523) and its frames
are independent (:534), so both shardings are exact — N-GPU output is byte-identical to 1-GPU output:

  * one frame   -> interleaved blocks of `row_block` pixel rows: block b goes to rank b % world
                   (srt_render_params.row_block / shard_index / shard_count; interleaving balances the
                   sky-heavy top against the geometry-heavy middle of the frame);
  * image series -> frame f goes to rank f % world; the scene (a few MB) is replicated.

torch is plumbing here: device buffers, the process group, the gather.
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist

DEFAULT_ROW_BLOCK = 16


def owned_blocks(height: int, row_block: int, rank: int, world: int) -> List[Tuple[int, int]]:
    """Row ranges [r0, r1) of the blocks rank `rank` renders (same rule as the kernel's tile_row_to_py)."""
    if row_block <= 0 or row_block % 4:
        raise ValueError("row_block must be a positive multiple of 4 (the tile height)")
    n_blocks = (height + row_block - 1) // row_block
    return [(b * row_block, min((b + 1) * row_block, height)) for b in range(rank, n_blocks, world)]


def frames_of_rank(n_frames: int, rank: int, world: int) -> List[int]:
    """Image-series sharding: frame f -> rank f % world."""
    return list(range(rank, n_frames, world))


def padded_height(height: int, row_block: int, world: int) -> int:
    """Height rounded up so that every rank owns the same number of whole blocks (gather needs equal sizes)."""
    per_round = row_block * world
    return (height + per_round - 1) // per_round * per_round


def gather_row_blocks(frame: torch.Tensor, height: int, row_block: int, group=None, dst: int = 0) -> Optional[torch.Tensor]:
    """`frame` is [padded_height, W, C] on every rank, with this rank's own blocks filled in.  Gathers the
    blocks to rank `dst` and returns the assembled [height, W, C] frame there (None elsewhere)."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    hp = frame.shape[0]
    if hp % (row_block * world):
        raise ValueError("frame height must be padded with padded_height()")
    blocks = frame.view(hp // row_block, row_block, *frame.shape[1:])
    mine = blocks[rank::world].contiguous()
    if world == 1:
        return frame[:height]
    parts = [torch.empty_like(mine) for _ in range(world)] if rank == dst else None
    dist.gather(mine, parts, dst=dst, group=group)
    if rank != dst:
        return None
    for r in range(world):
        blocks[r::world] = parts[r]
    return frame[:height]


def gather_row_blocks_in_place(frame: torch.Tensor, height: int, row_block: int, group=None, dst: int = 0) -> Optional[torch.Tensor]:
    """`frame` is [>= height, W, C] and contiguous on every rank, with this rank's own row blocks rendered in place.  Every
    block travels straight from the owner's frame into the same rows of rank `dst`'s frame: one send / receive per block,
    all of them in one batch (one NCCL group: a single launch per rank), no packing on the sender, no unpacking on the
    receiver — the render kernels wrote the bytes the collective sends.  Returns frame[:height] on `dst`, None elsewhere."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    if world == 1:
        return frame[:height]
    if not frame.is_contiguous():
        raise ValueError("frame must be contiguous (row blocks are sent as views)")
    n_blocks = (height + row_block - 1) // row_block
    ops = []
    for b in range(n_blocks):
        owner = b % world
        if owner == dst or rank not in (owner, dst):
            continue
        rows = frame[b * row_block:min((b + 1) * row_block, height)]
        peer = dist.get_global_rank(group, dst if rank == owner else owner) if group is not None else (dst if rank == owner else owner)
        ops.append(dist.P2POp(dist.isend if rank == owner else dist.irecv, rows, peer, group))
    if ops:
        for req in dist.batch_isend_irecv(ops):
            req.wait()
    return frame[:height] if rank == dst else None


def gather_frames(frame: torch.Tensor, group=None, dst: int = 0) -> Optional[List[torch.Tensor]]:
    """Image series: every rank contributes one finished frame; rank `dst` receives all of them in rank order."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    if world == 1:
        return [frame]
    parts = [torch.empty_like(frame) for _ in range(world)] if rank == dst else None
    dist.gather(frame, parts, dst=dst, group=group)
    return parts


class ShardedFrameRenderer:
    """One frame over all ranks of the process group.  Every rank holds the (replicated) scene."""

    def __init__(self, ctx, width: int, height: int, row_block: int = DEFAULT_ROW_BLOCK, group=None):
        self.ctx, self.group = ctx, group
        self.width, self.height, self.row_block = width, height, row_block
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        hp = padded_height(height, row_block, self.world)
        self.frame = torch.zeros((hp, width, 3), dtype=torch.uint8, device=torch.device("cuda", ctx.device))

    def render(self, scene, params) -> Optional[torch.Tensor]:
        """Renders this rank's row blocks in place and gathers; returns the [H, W, 3] frame on rank 0."""
        params.row_begin, params.row_end = 0, self.height
        params.row_block, params.shard_index, params.shard_count = self.row_block, self.rank, self.world
        self.ctx.render_raw(scene, params, rgb=self.frame.data_ptr())
        if getattr(self.ctx, "stream", None) is None:
            self.ctx.sync()               # context on its own stream: order the gather after the render
        if self.world == 1:
            return self.frame[: self.height]
        return gather_row_blocks_in_place(self.frame, self.height, self.row_block, self.group)
