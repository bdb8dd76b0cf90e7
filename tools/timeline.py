"""Debug tool: per-tile cost map of the render kernel (variant 101).  Writes gpurun_out/timeline.npz."""
import sys
from pathlib import Path
import numpy as np
REPO = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(REPO))
from simple_raytracer_b200 import _abi, capi, host, scenes

cfg = scenes.CONFIGS["C3"]
hs = host.HostScene(); view = cfg.build(hs, 0.0); light = scenes.light_view_space(hs, view); arrays = hs.flatten()
W, H, focal = cfg.width, cfg.height, cfg.focal
p = _abi.default_params(W, H, focal, light, cfg.light_count, True)
ctx = capi.Context(0); sc = ctx.upload(arrays)
for _ in range(3):
    ctx.render(sc, p)
p.variant = int(sys.argv[1]) if len(sys.argv) > 1 else 101
got = ctx.render(sc, p)
cyc = got["t"][::4, ::8]; sm = got["obj_id"][::4, ::8]; ns = got["tri_id"][::4, ::8].astype(np.int64)
ns = ns - ns.min()
print("kernel ms", got["stats"]["kernel_ms"], "tiles", cyc.size)
print("tile cycles: mean %.0f median %.0f p99 %.0f max %.0f" % (cyc.mean(), np.median(cyc), np.percentile(cyc, 99), cyc.max()))
end = ns + cyc / 1.9
print("last start us %.1f, last end us %.1f" % (ns.max() / 1e3, end.max() / 1e3))
order = np.argsort(cyc.ravel())[::-1][:20]
for o in order:
    ty, tx = divmod(int(o), cyc.shape[1])
    print("tile (%d,%d) px (%d,%d): %.0f cycles = %.1f us, start %.1f us, sm %d" % (tx, ty, tx * 8, ty * 4, cyc[ty, tx], cyc[ty, tx] / 1.9e3, ns[ty, tx] / 1e3, sm[ty, tx]))
rows = cyc.reshape(cyc.shape[0], -1).sum(axis=1)
print("row-band (40 px) cost share:", np.round(100 * rows.reshape(-1, 10).sum(axis=1) / rows.sum(), 1).tolist())
np.savez_compressed(REPO / "gpurun_out" / "timeline.npz", cyc=cyc, sm=sm, ns=ns)
