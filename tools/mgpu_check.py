"""Multi-GPU check (run under torchrun, one rank per GPU):
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/mgpu_check.py
Row-block-sharded single frame gathered to rank 0 over NCCL == the same frame rendered by rank 0 alone,
byte for byte; the same for frames of the image series gathered by frame."""
import os
import sys
from pathlib import Path

import numpy as np
import torch
import torch.distributed as dist

REPO = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(REPO))
from simple_raytracer_b200 import _abi, capi, host, multigpu, scenes

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
ctx = capi.Context(local, stream=torch.cuda.current_stream().cuda_stream)
cfg = scenes.CONFIGS["C3"]
W, H, focal = scenes.scaled(cfg, 540)
hs = host.HostScene(); view = cfg.build(hs, 0.0); light = scenes.light_view_space(hs, view); arrays = hs.flatten()
sc = ctx.upload(arrays)
ok = True
for mode in ("peer", "gather"):
    for block in (16, 32):
        r = multigpu.ShardedFrameRenderer(ctx, W, H, row_block=block, mode=mode)
        p = _abi.default_params(W, H, focal, light, cfg.light_count, True)
        for rep in range(3):
            out = r.render(sc, p)
            torch.cuda.synchronize()
            dist.barrier()
        if rank == 0:
            p2 = _abi.default_params(W, H, focal, light, cfg.light_count, True)
            want = ctx.render(sc, p2, ids=False, want_t=False)["rgb"]
            same = np.array_equal(out.cpu().numpy(), want)
            print(f"{mode}, row_block {block}: {world}-rank sharded frame == single-GPU frame: {same}")
            ok &= same
        r.close()
# series: frame f -> rank f % world
frames = multigpu.frames_of_rank(2 * world, rank, world)
rgb = torch.empty((H, W, 3), dtype=torch.uint8, device="cuda")
for step, f in enumerate(frames):
    hs = host.HostScene(); view = cfg.build(hs, 1.5 * f); light = scenes.light_view_space(hs, view); sc.update(hs.flatten())
    p = _abi.default_params(W, H, focal, light, cfg.light_count, True)
    ctx.render_raw(sc, p, rgb=rgb.data_ptr())
    parts = multigpu.gather_frames(rgb)
    torch.cuda.synchronize()
    if rank == 0:
        for r_, part in enumerate(parts):
            fr = step * world + r_
            hs2 = host.HostScene(); v2 = cfg.build(hs2, 1.5 * fr); l2 = scenes.light_view_space(hs2, v2)
            sc2 = ctx.upload(hs2.flatten())
            want = ctx.render(sc2, _abi.default_params(W, H, focal, l2, cfg.light_count, True), ids=False, want_t=False)["rgb"]
            sc2.destroy()
            same = np.array_equal(part.cpu().numpy(), want)
            print(f"series frame {fr} from rank {r_}: identical to rank 0's own render: {same}")
            ok &= same
dist.barrier()
if rank == 0:
    print("MGPU CHECK", "PASSED" if ok else "FAILED")
dist.destroy_process_group()
sys.exit(0 if ok else 1)
