"""Image-series driver (BASELINE configs[4]; the reference's frame loop, simple_raytracer.cpp:534-794).

The reference rebuilds everything per frame — OBJ parse, transforms (the camera is baked into the vertices,
:558), BVH build — on one thread, ~0.1-0.4 s per frame, while the GPU renders a frame in about a millisecond.
This driver keeps the per-frame work bit-identical and takes it off the critical path:

  * assets are parsed / decoded once (host.set_asset_cache: loading is deterministic);
  * frames are prepared `depth` ahead on a pool of host threads (the C++ scene calls release the GIL);
  * the prepared, flattened scene is uploaded straight from the C++-owned arrays (no Python copy), rendered,
    and the frame is read back into pinned memory — asynchronously, on two contexts (streams) that alternate, so
    the GPU leg of frame i runs while the driver waits for the prep of frame i+1.

With `prep="device"` the per-frame host work disappears altogether: the meshes stay on the GPU in object space and
every frame is one recorded scene script (a few matrices) + srt_scene_prep (transform, BVH build, flatten as CUDA
kernels, bit-identical to the host result, prep.py) + render + download.  A frame the device path cannot take
(non-finite coordinates) is redone through the host path.

With several GPUs each rank runs this driver on its own frames (frame f -> rank f % world, multigpu.py).
"""
from __future__ import annotations

import os
import time
from collections import deque
from concurrent.futures import ThreadPoolExecutor
from typing import Callable, Iterable, Iterator, Optional, Tuple

import numpy as np

from . import _abi, capi, host, prep as _prep, scenes


def prepare_frame(cfg: scenes.SceneConfig, angle_deg: float):
    """Host prep of one frame through the scene surface.  Returns (HostScene, desc pointer, light)."""
    hs = host.HostScene()
    view = cfg.build(hs, angle_deg)
    light = scenes.light_view_space(hs, view) if view is not None else np.array(scenes.LIGHT_WORLD, np.float32)
    desc = hs.flatten_desc(with_flat_lists=not cfg.use_bvh)
    return hs, desc, np.asarray(light, np.float32)


class _Lane:
    """One context + device scene + pinned host frame.  Two lanes alternate so that the upload / render / download
    of frame i overlaps the wait for the host prep of frame i+1."""

    def __init__(self, ctx: capi.Context, width: int, height: int, own_ctx: bool):
        self.ctx, self.own_ctx = ctx, own_ctx
        self.scene: Optional[capi.Scene] = None
        self.frame = capi.PinnedArray((height, width, 3), np.uint8)

    def close(self):
        if self.scene is not None:
            self.scene.destroy()
            self.scene = None
        self.frame.free()
        if self.own_ctx:
            self.ctx.destroy()


class SeriesRenderer:
    def __init__(self, ctx: capi.Context, cfg: scenes.SceneConfig, width: int, height: int, focal: float,
                 light_count: int, workers: Optional[int] = None, depth: Optional[int] = None, variant: int = 0,
                 save_pattern: Optional[str] = None, prep: str = "host", lanes: Optional[int] = None):
        """save_pattern: e.g. "images/generation/output{index}.bmp" — every frame is also written as the BMP the
        reference's drawImage saves (:488-494), on the worker pool, off the critical path."""
        self.ctx, self.cfg = ctx, cfg
        self.save_pattern = save_pattern
        self.width, self.height, self.focal, self.light_count, self.variant = width, height, focal, light_count, variant
        self.workers = workers or max(1, len(os.sched_getaffinity(0)))
        self.depth = depth or 2 * self.workers
        # lanes = contexts (streams) that take frames in turn.  Two are enough to overlap copies with renders; the device
        # prep is a chain of short, latency-bound launches that fills few SMs, so more frames in flight overlap better.
        n_lanes = lanes or (6 if prep == "device" else 2)
        self.lanes = [_Lane(ctx, width, height, False)] + [_Lane(capi.Context(ctx.device), width, height, True) for _ in range(n_lanes - 1)]
        host.set_asset_cache(True)
        if prep not in ("host", "device"):
            raise ValueError("prep must be 'host' or 'device'")
        self.prep = prep
        self.mesh_cache = _prep.MeshCache(ctx) if prep == "device" else None
        self.host_fallbacks = 0

    def _frames_device(self, angles, out):
        """prep='device': record the scene script, prepare + render + download on the lane's stream."""
        inflight = deque()                           # (index, lane, timing)
        pool = ThreadPoolExecutor(max_workers=2) if self.save_pattern else None

        def retire():
            i, lane, tm = inflight.popleft()
            t0 = time.perf_counter()
            try:
                lane.scene.prep_status()             # synchronises the lane
            except capi.SrtError as e:
                if e.code != _abi.SRT_ERR_UNSUPPORTED:
                    raise
                self.host_fallbacks += 1             # redo this frame through the host path
                hs, desc, light = prepare_frame(self.cfg, angles[i])
                lane.scene.destroy()
                lane.scene = None
                self._issue(lane, desc, light)
                lane.ctx.sync()
                hs.close()
                lane.scene.destroy()
                lane.scene = None
            rgb = lane.frame.array
            if pool is not None:
                pool.submit(host.write_image, self.save_pattern.format(index=i, angle=angles[i]), rgb.copy())
            if out is not None:
                np.copyto(out, rgb)
                rgb = out
            tm["gpu_ms"] += 1e3 * (time.perf_counter() - t0)
            return i, rgb, tm

        try:
            for i in range(len(angles)):
                if len(inflight) == len(self.lanes):
                    yield retire()
                t0 = time.perf_counter()
                lane = self.lanes[i % len(self.lanes)]
                ds = _prep.DeviceScene(self.mesh_cache)
                view = self.cfg.build(ds, angles[i])
                light = scenes.light_view_space(ds, view) if view is not None else np.array(scenes.LIGHT_WORLD, np.float32)
                lane.scene = ds.commit(scene=lane.scene, with_flat_lists=not self.cfg.use_bvh, ctx=lane.ctx)
                ds.close()
                p = _abi.default_params(self.width, self.height, self.focal, np.asarray(light, np.float32), self.light_count, self.cfg.use_bvh)
                p.variant = self.variant
                lane.ctx.render_raw(lane.scene, p, rgb=lane.frame.ptr)
                inflight.append((i, lane, dict(wait_prep_ms=0.0, gpu_ms=0.0, issue_ms=1e3 * (time.perf_counter() - t0))))
            while inflight:
                yield retire()
        finally:
            for _i, lane, _tm in inflight:
                lane.ctx.sync()
            if pool is not None:
                pool.shutdown(wait=True)

    def _issue(self, lane: _Lane, desc, light):
        """Upload + render + download of one frame on the lane's stream, asynchronous."""
        if lane.scene is None:
            h = capi.C.c_void_p()
            capi._check(capi.lib().srt_scene_upload(lane.ctx.h, desc, capi.C.byref(h)))
            lane.scene = capi.Scene(lane.ctx, h)
        else:
            lane.scene.update(desc.contents)
        p = _abi.default_params(self.width, self.height, self.focal, light, self.light_count, self.cfg.use_bvh)
        p.variant = self.variant
        lane.ctx.render_raw(lane.scene, p, rgb=lane.frame.ptr)

    def frames(self, angles: Iterable[float], out: Optional[np.ndarray] = None) -> Iterator[Tuple[int, np.ndarray, dict]]:
        """Yields (index, rgb[H][W][3] host array, timing) per frame, in order.  The yielded array is the lane's
        pinned frame, or `out` if given — either way valid only until the next iteration: the next frame is issued onto the
        lane that was just retired and starts overwriting its buffer at once.  Copy what you keep."""
        angles = list(angles)
        # several lanes already keep frames in flight concurrently: a second row band per frame (a caller may have asked for
        # it on the shared context) would only add launches
        before = [lane.ctx.set_bands(1) for lane in self.lanes] if len(self.lanes) > 1 else []
        try:
            yield from self._frames(angles, out)
        finally:
            for lane, b in zip(self.lanes, before):
                lane.ctx.set_bands(b)

    def _frames(self, angles, out):
        if self.prep == "device":
            yield from self._frames_device(angles, out)
            return
        pool = ThreadPoolExecutor(max_workers=self.workers)
        pending = deque()
        inflight = deque()                           # (index, lane, host scene, timing)
        nxt = 0

        def retire():
            i, lane, hs, tm = inflight.popleft()
            t0 = time.perf_counter()
            lane.ctx.sync()                          # frame i is in the lane's pinned buffer; `hs` may go away now
            pool.submit(hs.close)                    # freeing 70 k triangles + the Node tree takes milliseconds: off the critical path
            rgb = lane.frame.array
            if self.save_pattern:
                pool.submit(host.write_image, self.save_pattern.format(index=i, angle=angles[i]), rgb.copy())
            if out is not None:
                np.copyto(out, rgb)
                rgb = out
            tm["gpu_ms"] += 1e3 * (time.perf_counter() - t0)
            return i, rgb, tm

        try:
            while nxt < len(angles) and len(pending) < self.depth:
                pending.append(pool.submit(prepare_frame, self.cfg, angles[nxt])); nxt += 1
            for i in range(len(angles)):
                t0 = time.perf_counter()
                hs, desc, light = pending.popleft().result()
                t1 = time.perf_counter()
                if nxt < len(angles):
                    pending.append(pool.submit(prepare_frame, self.cfg, angles[nxt])); nxt += 1
                if len(inflight) == len(self.lanes):
                    yield retire()                   # frees the lane frame i is about to use
                t2 = time.perf_counter()
                lane = self.lanes[i % len(self.lanes)]
                self._issue(lane, desc, light)
                inflight.append((i, lane, hs, dict(wait_prep_ms=1e3 * (t1 - t0), gpu_ms=1e3 * (time.perf_counter() - t2))))
            while inflight:
                yield retire()
        finally:
            for _i, lane, hs, _tm in inflight:
                lane.ctx.sync()
                hs.close()
            for f in pending:
                f.cancel()
            pool.shutdown(wait=True)                 # (queued frame writes and scene frees still run)

    def close(self):
        for lane in self.lanes:
            lane.close()
        self.lanes = []
        if self.mesh_cache is not None:
            self.mesh_cache.close()
            self.mesh_cache = None
