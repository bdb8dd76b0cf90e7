"""Image-series driver (BASELINE configs[4]; the reference's frame loop, simple_raytracer.cpp:534-794).

The reference rebuilds everything per frame — OBJ parse, transforms (the camera is baked into the vertices,
:558), BVH build — on one thread, ~0.1-0.4 s per frame, while the GPU renders a frame in about a millisecond.
This driver keeps the per-frame work bit-identical and takes it off the critical path:

  * assets are parsed / decoded once (host.set_asset_cache: loading is deterministic);
  * frames are prepared `depth` ahead on a pool of host threads (the C++ scene calls release the GIL);
  * the prepared, flattened scene is uploaded straight from the C++-owned arrays (no Python copy), rendered,
    and the frame is read back — on the context's stream, one frame after the other.

With several GPUs each rank runs this driver on its own frames (frame f -> rank f % world, multigpu.py).
"""
from __future__ import annotations

import os
import time
from collections import deque
from concurrent.futures import ThreadPoolExecutor
from typing import Callable, Iterable, Iterator, Optional, Tuple

import numpy as np

from . import _abi, capi, host, scenes


def prepare_frame(cfg: scenes.SceneConfig, angle_deg: float):
    """Host prep of one frame through the scene surface.  Returns (HostScene, desc pointer, light)."""
    hs = host.HostScene()
    view = cfg.build(hs, angle_deg)
    light = scenes.light_view_space(hs, view) if view is not None else np.array(scenes.LIGHT_WORLD, np.float32)
    desc = hs.flatten_desc(with_flat_lists=not cfg.use_bvh)
    return hs, desc, np.asarray(light, np.float32)


class SeriesRenderer:
    def __init__(self, ctx: capi.Context, cfg: scenes.SceneConfig, width: int, height: int, focal: float,
                 light_count: int, workers: Optional[int] = None, depth: Optional[int] = None, variant: int = 0):
        self.ctx, self.cfg = ctx, cfg
        self.width, self.height, self.focal, self.light_count, self.variant = width, height, focal, light_count, variant
        self.workers = workers or max(1, len(os.sched_getaffinity(0)))
        self.depth = depth or 2 * self.workers
        self.scene: Optional[capi.Scene] = None
        host.set_asset_cache(True)

    def frames(self, angles: Iterable[float], out: Optional[np.ndarray] = None) -> Iterator[Tuple[int, np.ndarray, dict]]:
        """Yields (index, rgb[H][W][3] host array, timing) per frame, in order.  `out`: reuse one host buffer
        (the yielded array is then only valid until the next iteration)."""
        angles = list(angles)
        pool = ThreadPoolExecutor(max_workers=self.workers)
        pending = deque()
        nxt = 0
        try:
            while nxt < len(angles) and len(pending) < self.depth:
                pending.append(pool.submit(prepare_frame, self.cfg, angles[nxt])); nxt += 1
            for i in range(len(angles)):
                t0 = time.perf_counter()
                hs, desc, light = pending.popleft().result()
                t1 = time.perf_counter()
                if nxt < len(angles):
                    pending.append(pool.submit(prepare_frame, self.cfg, angles[nxt])); nxt += 1
                if self.scene is None:
                    h = capi.C.c_void_p()
                    capi._check(capi.lib().srt_scene_upload(self.ctx.h, desc, capi.C.byref(h)))
                    self.scene = capi.Scene(self.ctx, h)
                else:
                    self.scene.update(desc.contents)
                p = _abi.default_params(self.width, self.height, self.focal, light, self.light_count, self.cfg.use_bvh)
                p.variant = self.variant
                rgb = out if out is not None else np.empty((self.height, self.width, 3), np.uint8)
                self.ctx.render_raw(self.scene, p, rgb=rgb.ctypes.data)
                self.ctx.sync()                      # the C++-owned arrays of `hs` may go away now
                pool.submit(hs.close)                # freeing 70 k triangles + the Node tree takes milliseconds: off the critical path
                t2 = time.perf_counter()
                yield i, rgb, dict(wait_prep_ms=1e3 * (t1 - t0), gpu_ms=1e3 * (t2 - t1))
        finally:
            for f in pending:
                f.cancel()
            pool.shutdown(wait=True, cancel_futures=True)

    def close(self):
        if self.scene is not None:
            self.scene.destroy()
            self.scene = None
