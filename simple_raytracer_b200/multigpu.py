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


class _DeviceBuffer:
    """A raw device allocation seen by torch through __cuda_array_interface__ (uint8, 1-D)."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (int(nbytes),), "typestr": "|u1", "data": (int(ptr), False), "version": 3, "strides": None}


class ShardedFrameRenderer:
    """One frame over all ranks of the process group.  Every rank holds the (replicated) scene.

    mode "peer" (default on GPUs): rank 0 owns the frame (a cudaMalloc buffer exported over CUDA IPC); every other rank maps
    it and renders its interleaved row blocks STRAIGHT into rank 0's memory — the shade kernel's stores travel over NVLink /
    NVSwitch, there is no gather and no staging copy; the join is one 4-byte all-reduce on each rank's render stream.
    mode "gather": every rank renders into its own full-size frame and the row blocks are sent from there into the same
    rows of rank 0's frame (NCCL batched send / receive, gather_row_blocks_in_place) — also what the gloo CPU tests use."""

    def __init__(self, ctx, width: int, height: int, row_block: int = DEFAULT_ROW_BLOCK, group=None, mode: Optional[str] = None):
        self.ctx, self.group = ctx, group
        self.width, self.height, self.row_block = width, height, row_block
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.mode = mode or ("peer" if self.world > 1 else "gather")
        self._own_ptr = self._peer_ptr = None
        dev = torch.device("cuda", ctx.device)
        nbytes = height * width * 3
        if self.mode == "peer" and self.world > 1:
            info = [None]
            if self.rank == 0:
                self._own_ptr = ctx.device_alloc(nbytes)
                info[0] = ctx.ipc_export(self._own_ptr)
            dist.broadcast_object_list(info, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
            ok = 1
            if self.rank != 0:
                try:
                    self._peer_ptr = ctx.ipc_open(info[0])
                except Exception:
                    ok = 0                                  # no peer access to rank 0's GPU from here
            flag = torch.tensor([ok], dtype=torch.int32, device=dev)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)   # every rank takes the same path
            if int(flag.item()) == 1:
                if self.rank == 0:
                    self.target = self._own_ptr
                    self.frame = torch.as_tensor(_DeviceBuffer(self._own_ptr, nbytes), device=dev).view(height, width, 3)
                else:
                    self.target = self._peer_ptr
                    self.frame = None
                self._token = torch.zeros(1, dtype=torch.int32, device=dev)
                return
            self._release_shared()
        self.mode = "gather"
        self.frame = torch.zeros((height, width, 3), dtype=torch.uint8, device=dev)
        self.target = self.frame.data_ptr()

    def _release_shared(self):
        if self._peer_ptr is not None:
            self.ctx.ipc_close(self._peer_ptr)
            self._peer_ptr = None
        if dist.is_initialized() and self.world > 1:
            dist.barrier(group=self.group)        # nobody still maps the frame when its owner frees it
        if self._own_ptr is not None:
            self.frame = None
            self.ctx.sync()
            self.ctx.device_free(self._own_ptr)
            self._own_ptr = None

    def render(self, scene, params) -> Optional[torch.Tensor]:
        """Renders this rank's row blocks and joins; returns the [H, W, 3] frame on rank 0 (valid in stream order)."""
        params.row_begin, params.row_end = 0, self.height
        if self.world > 1:
            params.row_block, params.shard_index, params.shard_count = self.row_block, self.rank, self.world
        self.ctx.render_raw(scene, params, rgb=self.target)
        if getattr(self.ctx, "stream", None) is None:
            self.ctx.sync()               # context on its own stream: order the join after the render
        if self.world == 1:
            return self.frame[: self.height]
        if self.mode == "peer":
            # every rank's all-reduce kernel runs after its render kernels (stream order) and none completes before all have
            # started: when it is done on rank 0, every rank's stores into the frame have been performed
            dist.all_reduce(self._token, group=self.group)
            return self.frame if self.rank == 0 else None
        return gather_row_blocks_in_place(self.frame, self.height, self.row_block, self.group)

    def close(self):
        if self.mode == "peer" and self.world > 1:
            self._release_shared()
