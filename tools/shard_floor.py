"""Where does the time of ONE shard of a sharded frame go?  (the floor of the single-frame strong scaling)
Renders shard 0 of R (interleaved row blocks) of C3 at the given angle on one GPU: per-pass device times (one piece), the
two-band span, and the cost of the longest primary tiles (variant 102: per-tile cycle counts in the t plane).
    python tools/shard_floor.py [R=8] [angle=0]"""
import sys
from pathlib import Path
import numpy as np
REPO = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(REPO))
from simple_raytracer_b200 import _abi, capi, host, scenes

R = int(sys.argv[1]) if len(sys.argv) > 1 else 8
angle = float(sys.argv[2]) if len(sys.argv) > 2 else 0.0
cfg = scenes.CONFIGS["C3"]
hs = host.HostScene(); view = cfg.build(hs, angle); light = scenes.light_view_space(hs, view); arrays = hs.flatten(); hs.close()
W, H, focal = cfg.width, cfg.height, cfg.focal
ctx = capi.Context(0); sc = ctx.upload(arrays)
dev = ctx.device_alloc(W * H * 3)


def params(shard, variant=0):
    p = _abi.default_params(W, H, focal, light, cfg.light_count, True)
    p.variant = variant
    if R > 1:
        p.row_block, p.shard_index, p.shard_count = 16, shard, R
    return p


for bands in (1, 2):
    ctx.set_bands(bands)
    ctx.set_stage_timing(bands == 1)
    for shard in range(min(R, 3)):
        ms = []
        for it in range(6):
            ctx.render_raw(sc, params(shard), rgb=dev)
            ctx.sync()
            if it >= 2:
                ms.append(ctx.last_stats()["kernel_ms"])
        extra = ""
        if bands == 1:
            extra = " passes " + ", ".join(f"{k} {v:.3f}" for k, v in ctx.last_stage_ms().items())
        print(f"C3 @ {angle:g} deg, shard {shard} of {R}, bands {bands}: {np.mean(ms):.3f} ms (min {np.min(ms):.3f}){extra}")
ctx.set_bands(1)
ctx.set_stage_timing(False)
t = np.zeros((H, W), np.float32); obj = np.zeros((H, W), np.int32); tri = np.zeros((H, W), np.int32)
ctx.render_raw(sc, params(0, 102), None, obj.ctypes.data, tri.ctypes.data, t.ctypes.data)
ctx.sync()
cyc = t[::4, ::8]
mine = np.zeros(cyc.shape[0], bool)
for r0 in range(0, H, 16 * R):
    mine[r0 // 4:(r0 + 16) // 4] = True
c = cyc[mine]
srt = np.sort(c.ravel())[::-1]
print(f"primary tiles of shard 0: {c.size}; cycles mean {c.mean():.0f}, median {np.median(c):.0f}, p99 {np.percentile(c, 99):.0f}, "
      f"five longest {[int(x) for x in srt[:5]]} = {[round(float(x) / 1.965e3, 1) for x in srt[:5]]} us at 1965 MHz")
ctx.device_free(dev); sc.destroy(); ctx.destroy()
