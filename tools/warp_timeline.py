"""Debug tool: per-warp timestamps of the dataflow kernel (variant 103): when each warp ran out of primary
tiles and when it finished, and how many items of each kind it processed."""
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
p.variant = 103
got = ctx.render(sc, p)
nw = ctx.info()["sm_count"] * 8 * 4
raw = got["t"].reshape(-1).view(np.uint64)[: nw * 4].reshape(nw, 4)   # 8 u64 per warp = 16 floats -> view gives 8 u64 per 16 floats
raw = got["t"].reshape(-1)[: nw * 16].view(np.uint64).reshape(nw, 8)
ok = raw[:, 0] > 0
raw = raw[ok]
t0 = raw[:, 0].min()
start, pend, end = [(raw[:, k] - t0) / 1e3 for k in range(3)]
print("kernel ms", got["stats"]["kernel_ms"], "warps", len(raw))
print("warp start us: max %.1f" % start.max())
print("primary queue empty (per warp) us: min %.1f median %.1f max %.1f" % (pend.min(), np.median(pend), pend.max()))
print("warp end us: min %.1f median %.1f max %.1f" % (end.min(), np.median(end), end.max()))
print("items per warp: primary mean %.1f, shadow mean %.1f (max %d), shade-from-memory mean %.1f" % (raw[:, 3].mean(), raw[:, 4].mean(), raw[:, 4].max(), raw[:, 5].mean()))
h, edges = np.histogram(pend, bins=12)
print("histogram of primary-queue-empty times:", list(zip(np.round(edges[:-1]).astype(int).tolist(), h.tolist())))
h, edges = np.histogram(end, bins=12)
print("histogram of end times:", list(zip(np.round(edges[:-1]).astype(int).tolist(), h.tolist())))
