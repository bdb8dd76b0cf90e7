"""CPU tests (gloo, world_size 2) of the multi-GPU plumbing: row-block / frame sharding rules and the
gathers to rank 0.  The render itself is replaced by a deterministic pattern — what is tested is that the
union of the shards, gathered, is byte-identical to the whole."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from simple_raytracer_b200 import multigpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _pattern(h, w):
    y, x = np.mgrid[0:h, 0:w]
    return torch.from_numpy(np.stack([(x * 7 + y * 3) % 251, (x + 5 * y) % 241, (x * y) % 239], -1).astype(np.uint8))


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        results = {}
        for (H, W, rb) in ((1080, 64, 16), (37, 9, 4), (100, 16, 8)):
            full = _pattern(H, W)
            hp = multigpu.padded_height(H, rb, world)
            frame = torch.zeros((hp, W, 3), dtype=torch.uint8)
            for r0, r1 in multigpu.owned_blocks(H, rb, rank, world):        # "render" my blocks
                frame[r0:r1] = full[r0:r1]
            out = multigpu.gather_row_blocks(frame, H, rb)
            if rank == 0:
                results[(H, W, rb)] = bool(torch.equal(out, full))
            else:
                assert out is None
            # the zero-copy form: blocks travel from the owner's frame into the same rows of rank 0's frame (no padding)
            frame2 = torch.zeros((H, W, 3), dtype=torch.uint8)
            for r0, r1 in multigpu.owned_blocks(H, rb, rank, world):
                frame2[r0:r1] = full[r0:r1]
            out2 = multigpu.gather_row_blocks_in_place(frame2, H, rb)
            if rank == 0:
                results[(H, W, rb, "in place")] = bool(torch.equal(out2, full)) and out2.data_ptr() == frame2.data_ptr()
            else:
                assert out2 is None
        # image series: frame f -> rank f % world, gathered in rank order each step
        n_frames = 6
        mine = multigpu.frames_of_rank(n_frames, rank, world)
        got = []
        for step in range(n_frames // world):
            f = mine[step]
            parts = multigpu.gather_frames(torch.full((4, 5, 3), f, dtype=torch.uint8))
            if rank == 0:
                got += [int(p[0, 0, 0]) for p in parts]
        if rank == 0:
            results["series"] = got == list(range(n_frames))
            q.put(results)
    finally:
        dist.destroy_process_group()


def test_gloo_world2_shards_gather_to_identical_frame():
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    res = q.get()
    assert res and all(res.values()), res


def test_owned_blocks_match_kernel_row_mapping():
    """owned_blocks() (host) == tile_row_to_py() of srt_kernels.cuh (device) for every local tile row."""
    def tile_row_to_py(row_begin, btr, si, sc, ty):
        blk, off = divmod(ty, btr)
        return row_begin + ((blk * sc + si) * btr + off) * 4
    for H, rb, world in ((1080, 16, 8), (1080, 16, 3), (2160, 32, 8), (37, 4, 2)):
        seen = np.zeros(H, np.int32)
        for rank in range(world):
            blocks = multigpu.owned_blocks(H, rb, rank, world)
            rows = [r for r0, r1 in blocks for r in range(r0, r1)]
            n_blocks = -(-(-(-H // 4)) // (rb // 4))
            mine = (n_blocks - rank + world - 1) // world if n_blocks > rank else 0
            krows = []
            for ty in range(mine * (rb // 4)):
                py = tile_row_to_py(0, rb // 4, rank, world, ty)
                krows += [r for r in range(py, py + 4) if r < H]
            assert rows == krows
            seen[rows] += 1
        assert np.all(seen == 1)          # a partition of the frame
    with pytest.raises(ValueError):
        multigpu.owned_blocks(100, 6, 0, 2)
