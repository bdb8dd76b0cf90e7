"""Renders a few frames of one view of one config through the C ABI — the command ncu wraps.
    python tools/profile_frame.py --config C3 --angle 180 --bands 1 --frames 3 [--lights N] [--variant V]
Prints per-frame kernel time and (one piece, default pipeline) the per-pass times."""
import argparse
import sys
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from simple_raytracer_b200 import _abi, capi, host, scenes  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--config", default="C3")
ap.add_argument("--angle", type=float, default=180.0)
ap.add_argument("--bands", type=int, default=1)
ap.add_argument("--frames", type=int, default=3)
ap.add_argument("--lights", type=int, default=-1)
ap.add_argument("--variant", type=int, default=0)
a = ap.parse_args()
cfg = scenes.CONFIGS[a.config]
W, H, focal = cfg.width, cfg.height, cfg.focal
N = cfg.light_count if a.lights < 0 else a.lights
hs = host.HostScene()
view = cfg.build(hs, a.angle)
light = np.asarray(scenes.light_view_space(hs, view) if view is not None else scenes.LIGHT_WORLD, np.float32)
arrays = hs.flatten(with_flat_lists=not cfg.use_bvh)
hs.close()
ctx = capi.Context(0)
ctx.set_bands(a.bands)
if a.bands == 1 and a.variant == 0:
    ctx.set_stage_timing(True)
sc = ctx.upload(arrays)
p = _abi.default_params(W, H, focal, light, N, cfg.use_bvh)
p.variant = a.variant
dev = capi.lib().srt_device_alloc(ctx.h, W * H * 3)
for i in range(a.frames):
    ctx.render_raw(sc, p, rgb=dev)
    ctx.sync()
    st = ctx.last_stats()
    extra = ""
    if a.bands == 1 and a.variant == 0 and cfg.use_bvh:
        try:
            extra = " passes " + ", ".join(f"{k} {v:.3f}" for k, v in ctx.last_stage_ms().items())
        except capi.SrtError:
            pass
    rays = W * H + (N * st["hit_pixels"] if len(arrays.objects) >= 2 else 0)
    print(f"{a.config} @ {a.angle:g} deg {W}x{H} N={N} bands={a.bands}: frame {i}: {st['kernel_ms']:.3f} ms, {rays / st['kernel_ms'] / 1e3:.0f} Mrays/s, "
          f"{st['kernel_launches']} launches{extra}")
capi.lib().srt_device_free(ctx.h, dev)
sc.destroy()
ctx.destroy()
