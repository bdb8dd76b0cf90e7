"""Kernel time of the C3 scene over the camera orbit (which views are expensive)."""
import sys
from pathlib import Path
import numpy as np
REPO = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(REPO))
from simple_raytracer_b200 import _abi, capi, host, scenes

cfg = scenes.CONFIGS["C3"]
W, H, focal = cfg.width, cfg.height, cfg.focal
ctx = capi.Context(0)
host.set_asset_cache(True)
variants = [int(v) for v in sys.argv[1:]] or [0]
for ang in range(0, 360, 20):
    hs = host.HostScene(); view = cfg.build(hs, float(ang)); light = scenes.light_view_space(hs, view)
    sc = ctx.upload(hs.flatten())
    row = []
    for v in variants:
        p = _abi.default_params(W, H, focal, light, cfg.light_count, True); p.variant = v
        ms = []
        for _ in range(4):
            ctx.render_raw(sc, p, None, None, None, None); ms.append(ctx.last_stats()["kernel_ms"])
        st = ctx.last_stats()
        row.append("v%d %.3f ms" % (v, min(ms)))
    p.variant = 104
    ctx.render_raw(sc, p, None, None, None, None); g = ctx.last_stats()
    p.variant = 100
    ctx.render_raw(sc, p, None, None, None, None); c = ctx.last_stats()
    print("angle %3d: %s | hit %.2f M, box tests %.0f M, tri tests %.0f M, exact fallbacks %d; GPU shadow box tests: occluded rays %.0f M, free rays %.0f M" % (
        ang, ", ".join(row), st["hit_pixels"] / 1e6, c["node_visits"] / 1e6, c["tri_tests"] / 1e6, st["exact_slab_tests"], g["node_visits"] / 1e6, g["tri_tests"] / 1e6))
    sc.destroy()
