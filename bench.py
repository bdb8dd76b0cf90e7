#!/usr/bin/env python
"""bench.py — the headline measurement: Mrays/s (and frames/s) of the per-pixel render loop on the
bunny + ground, 16-light soft-shadow scene at 1920x1080 (BASELINE.json configs[2] = SURVEY.md C3,
frames of the configs[4] image series C5 when several GPUs shard the series by frame).

    python bench.py --gpus 1 --steps K --warmup W                 # one B200
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
           --master-port P bench.py --gpus N --steps K --warmup W # N B200s, one rank per GPU
    python bench.py --impl reference ...                          # the reference's CPU code, same metric

A "step" is one 1080p frame per rank: the N*(K+W) frames of a run are an image series over one full camera
orbit (frame g at angle 360*g/(N*(K+W)) degrees, frame 0 == C3), frame g = step*N + rank — so the mix of
cheap and expensive views is the same whatever N is.  `value` = rays of all ranks / max-over-ranks device
time, scenes resident in HBM; `e2e` = the same through the C ABI from pinned HOST buffers (scene upload +
render + download of the rank's frame; --gather adds the NCCL gather of all frames to rank 0).  Ray-count rule:
rays = W*H + light_count * hit_pixels * [scene has >= 2 objects].

torch is plumbing here (device buffers, streams, events, torch.distributed); the render path is
simple_raytracer_b200/lib/libsrt_b200.so through include/srt.h and raises if it is missing.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

REPO = Path(__file__).resolve().parent
if str(REPO) not in sys.path:
    sys.path.insert(0, str(REPO))

METRIC = "Mrays/s (bunny + ground, 16-light soft shadows, 1920x1080)"
UNIT = "Mrays/s"
WORKLOAD = ("C3/C5: stanford-bunny.obj (x150) + ground cube, 16 staircase lights, 1920x1080, focal 1080, BVH; "
            "image series over one camera orbit, frame step*N+rank of N*(steps+warmup), frame 0 = C3")
L2_FLUSH_BYTES = 256 << 20          # twice the 126 MB L2


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="C3", help="scene config key (bench lines use C3; others are for profiling)")
    ap.add_argument("--height", type=int, default=0, help="override the image height (same field of view)")
    ap.add_argument("--lights", type=int, default=-1)
    ap.add_argument("--variant", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--cpu-seconds", type=float, default=20.0, help="CPU work (core-seconds) of the cpu_baseline sample")
    ap.add_argument("--still", action="store_true", help="render frame 0 on every step (profiling)")
    ap.add_argument("--e2e-lanes", type=int, default=3, help="contexts / streams that take the e2e leg's frames in turn")
    ap.add_argument("--e2e-bands", type=int, default=0, help="row bands of the two e2e lanes (0: the library default)")
    ap.add_argument("--gather", action="store_true", help="e2e: gather every rank's frame to rank 0 over NCCL before the download")
    ap.add_argument("--no-single-frame", action="store_true", help="skip the single-frame strong-scaling leg")
    ap.add_argument("--row-block", type=int, default=16, help="rows per block of the single-frame sharding")
    ap.add_argument("--series-lanes", type=int, default=6, help="frames in flight (contexts / streams) of the device-prep series leg")
    ap.add_argument("--series-frames", type=int, default=-1,
                    help="frames of the pipelined image-series leg incl. host prep (-1: 3 x host threads, max 96; 0: skip)")
    return ap.parse_args()


# ------------------------------------------------------------------------------------------------
def frame_setup(args):
    from simple_raytracer_b200 import scenes
    cfg = scenes.CONFIGS[args.config]
    if args.height > 0:
        W, H, focal = scenes.scaled(cfg, args.height)
    else:
        W, H, focal = cfg.width, cfg.height, cfg.focal
    N = cfg.light_count if args.lights < 0 else args.lights
    return cfg, W, H, focal, N


def frame_angle(g, total, still=False):
    """Camera angle of global series frame g of `total` views spread over one orbit (frame 0 = the config itself)."""
    return 0.0 if still or total <= 0 else 360.0 * (g % total) / total


def build_frame(backend_factory, cfg, angle):
    """Host prep of one series frame through the scene surface: load, transform, BVH, flatten."""
    from simple_raytracer_b200 import scenes
    b = backend_factory()
    view = cfg.build(b, angle)
    light = scenes.light_view_space(b, view) if view is not None else np.array(scenes.LIGHT_WORLD, np.float32)
    arrays = b.flatten(with_flat_lists=not cfg.use_bvh)
    return b, arrays, np.asarray(light, np.float32)


def run_shape(args, world):
    """(n_unique, total): distinct frames kept per rank, and the number of views the orbit is divided into — shared by both
    arms so that they time the same views."""
    n_unique = min(args.steps + args.warmup, 32)
    return n_unique, world * n_unique


def make_config(args, cfg, W, H, focal, N, world, n_tris, n_nodes):
    """The `config` object of the JSON line: identical in both arms (the driver compares them)."""
    n_unique, total = run_shape(args, world)
    return {"workload": WORKLOAD if args.config == "C3" and not args.still else f"{cfg.key}: {cfg.description}",
            "width": W, "height": H, "focal": focal, "lights": N, "triangles": int(n_tris), "bvh_nodes": int(n_nodes),
            "parallelism": f"frame-sharded x{world}",
            "views": ("frame 0 on every step" if args.still else
                      f"{total} views spread over one camera orbit (view g at {360.0 / total:g}*g degrees); step s of rank r is view "
                      f"(s mod {n_unique})*{world}+r; timed steps {args.warmup}..{args.warmup + args.steps - 1}")}


def ray_count(W, H, N, hit_pixels, n_objects):
    return W * H + (N * hit_pixels if n_objects >= 2 else 0)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 100 ms while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.thread = [], None, None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def stop(self, t0, t1):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        rows = [r for (t, r) in self.rows if t0 - 0.05 <= t <= t1 + 0.15] or [r for (_t, r) in self.rows]
        for r in rows:
            f = [x.strip() for x in r.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1])); pw.append(float(f[2]))
            except ValueError:
                continue
            for k, n in enumerate(names):
                if f[3 + k].lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}


# ------------------------------------------------------------------------------------------------
def cpu_reference_sample(cfg, W, H, focal, N, core_seconds, angles, threads=0, per_step_wall=None):
    """Times the reference's CPU implementation of the path (oracle/_ref when present, else the C oracle port) on a bounded
    column sample of the frame.  Step i samples the view at angles[i % len(angles)] (built on first use).
    Returns (fn(step), info, scene_info)."""
    from oracle import pyoracle
    from simple_raytracer_b200 import _abi, host
    ncores = threads or len(os.sched_getaffinity(0)) or (os.cpu_count() or 1)
    kind = "reference" if pyoracle.ref_available() else "port"
    factory = pyoracle.RefBackend if kind == "reference" else host.HostScene
    views = {}

    def view(i):
        k = i % len(angles)
        if k not in views:
            views[k] = build_frame(factory, cfg, angles[k])
        return views[k]

    def run(col_begin, col_step, th, i=0):
        b, arrays, light = view(i)
        if kind == "reference":
            r = b.render(W, H, focal, light, N, cfg.use_bvh, faithful=True, threads=th, col_begin=col_begin, col_step=col_step, ids=False)
            cols = np.arange(col_begin, W, col_step)
            return r["traced"], int(np.count_nonzero(r["obj_id"][:, cols] >= 0)), r["seconds"]
        p = _abi.default_params(W, H, focal, light, N, cfg.use_bvh)
        r = pyoracle.oracle_render(arrays, p, threads=th, col_begin=col_begin, col_step=col_step)
        return r["traced"], r["counters"]["hit_pixels"], r["seconds"]

    arrays0 = view(0)[1]
    n_obj = len(arrays0.objects)
    # calibrate on ~24 columns spread over the frame, all threads
    cal_step = max(1, W // 24)
    run(cal_step // 2, cal_step, ncores)              # untimed: first touch of the scene, thread start-up
    traced, hits, sec = run(cal_step // 2, cal_step, ncores)
    ncal = len(range(cal_step // 2, W, cal_step))
    core_s_per_col = sec * min(ncores, ncal) / max(1, ncal)
    want_cols = int((per_step_wall * ncores if per_step_wall is not None else core_seconds) / max(core_s_per_col, 1e-9))
    want_cols = max(min(want_cols, W), min(W, ncores))
    col_step = max(1, W // want_cols)
    ncols = len(range(0, W, col_step))
    sample = (f"{'unmodified reference (oracle/_ref, re-shading on every closer hit like :429-449)' if kind == 'reference' else 'C oracle port'}"
              f", {ncols} of {W} pixel columns (every {col_step}th) x {H} rows of each step's view, {ncores} threads; "
              f"{len(angles)} view(s): " + ", ".join(f"{a:g}" for a in angles[:6]) + (" ... degrees" if len(angles) > 6 else " degrees"))

    def step(i):
        traced, hits, sec = run(0, col_step, ncores, i)
        return ray_count(1, traced, N, hits, n_obj), sec, traced   # W*H replaced by the traced pixel count

    return step, dict(kind=kind, cores=ncores, sample=sample), dict(triangles=len(arrays0.geom), bvh_nodes=len(arrays0.nodes))


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return 0
    cfg, W, H, focal, N = frame_setup(args)
    n_unique, total = run_shape(args, world)
    # the GPU arm's views: step s of rank r renders view (s mod n_unique)*world + r.  Step s here samples that step's view
    # of rank s mod world — at N = 1 exactly the GPU arm's frames, at N > 1 one of the N views of every step.
    angles = [frame_angle((s % n_unique) * world + (s % world), total, args.still) for s in range(args.warmup + args.steps)]
    step, info, scene_info = cpu_reference_sample(cfg, W, H, focal, N, 0.0, angles, per_step_wall=1.0)
    for i in range(args.warmup):
        step(i)
    rays = 0
    sec = 0.0
    pixels = 0
    for i in range(args.steps):
        r, s, px = step(args.warmup + i)
        rays += r; sec += s; pixels += px
    v = rays / sec / 1e6
    line = {
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * sec / max(args.steps, 1), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": make_config(args, cfg, W, H, focal, N, world, scene_info["triangles"], scene_info["bvh_nodes"]),
        "note": "each step is a bounded column sample of that step's view; Mrays/s is per ray, so it extrapolates linearly",
        "frames_per_s_extrapolated": (pixels / (W * H)) / sec if sec > 0 else None,
        "cpu_baseline": {"value": v, "unit": UNIT, **info},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))
    return 0


# ------------------------------------------------------------------------------------------------
def single_frame_leg(args, ctx, rank, world, local, dev, barrier):
    """ONE frame over all N GPUs (strong scaling): interleaved blocks of --row-block rows, block b ->
    rank b % N, every rank renders its blocks in place into a full-size device frame, and the blocks travel from there
    into the same rows of rank 0's frame over NCCL (one batched send/recv group, no packing).  Rank 0 then renders the
    whole frame alone and compares the bytes.  At N = 1 it is the plain one-GPU frame (two row bands).  Times are CUDA
    events on each rank's stream around render + gather, max over ranks, mean over frames; L2 flushed between frames."""
    import torch
    import torch.distributed as dist
    from simple_raytracer_b200 import _abi, capi, host, multigpu, scenes
    out = {}
    flush = torch.empty(L2_FLUSH_BYTES, dtype=torch.uint8, device=dev)
    cases = [(k_, a_, m_) for (k_, a_) in (("C3", 0.0), ("C3", 180.0), ("C4", 0.0)) for m_ in (("peer", "gather") if world > 1 else ("peer",))]
    for key, angle, mode in cases:
        cfg = scenes.CONFIGS[key]
        W, H, focal, N = cfg.width, cfg.height, cfg.focal, cfg.light_count
        hs = host.HostScene()
        view = cfg.build(hs, angle)
        light = np.asarray(scenes.light_view_space(hs, view), np.float32)
        arrays = hs.flatten()
        hs.close()
        sc = ctx.upload(arrays)
        sharded = multigpu.ShardedFrameRenderer(ctx, W, H, row_block=args.row_block, mode=mode) if world > 1 else None
        frame1 = torch.empty((H, W, 3), dtype=torch.uint8, device=dev)

        def one_frame():
            p = _abi.default_params(W, H, focal, light, N, True)
            if sharded is None:
                ctx.render_raw(sc, p, rgb=frame1.data_ptr())
                return frame1
            return sharded.render(sc, p)

        iters, warm = 12, 3
        evs, render_ms = [], []
        res = None
        for it in range(warm + iters):
            flush.zero_()
            barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            res = one_frame()
            e1.record()
            torch.cuda.synchronize()
            if it >= warm:
                evs.append(e0.elapsed_time(e1))
                render_ms.append(ctx.last_stats()["kernel_ms"])
        t = torch.tensor([float(np.mean(evs)), float(np.mean(render_ms)), float(np.min(render_ms))], dtype=torch.float64, device=dev)
        tmin = t.clone()
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dist.all_reduce(tmin, op=dist.ReduceOp.MIN)
        identical = None
        hits = ctx.last_stats()["hit_pixels"]
        if rank == 0:
            p = _abi.default_params(W, H, focal, light, N, True)
            ctx.render_raw(sc, p, rgb=frame1.data_ptr())
            hits = ctx.last_stats()["hit_pixels"]
            identical = bool(torch.equal(res, frame1)) if world > 1 else True
            name = f"{key}_{W}x{H}_{N}lights_{angle:g}deg"
            rays = ray_count(W, H, N, hits, len(arrays.objects))
            entry = {"ms_per_frame": float(t[0]), "render_ms_slowest_rank": float(t[1]), "render_ms_fastest_rank": float(tmin[1]),
                     "join_ms": float(t[0] - t[1]), "frames_per_s": 1e3 / float(t[0]), "Mrays_per_s": rays / float(t[0]) / 1e3,
                     "identical": identical, "bytes_from_other_ranks": int(W * H * 3 * (world - 1) / world)}
            if mode == "peer":
                out[name] = entry
            else:
                out[name]["nccl_gather"] = entry
        if sharded is not None:
            sharded.close()
        sc.destroy()
        barrier()
    if rank == 0:
        out["what"] = (f"one frame over {world} GPU(s), one process per GPU: interleaved {args.row_block}-row blocks, block b -> rank b mod {world}; every rank's "
                       "kernels store their blocks straight into rank 0's frame (CUDA IPC mapping, NVLink stores, no gather, no staging copy), "
                       "joined by one 4-byte NCCL all-reduce on the render streams; `nccl_gather` = the same frame with every rank rendering into "
                       "its own frame and the blocks sent into rank 0's frame by batched NCCL send/recv; `identical` = bytes equal to rank 0's own "
                       "full render; ms = CUDA events around render + join, max over ranks, mean of 12 frames after 3 warm-up, L2 flushed and "
                       "ranks barriered before every frame") if world > 1 else \
                      "one frame on one GPU (default pipeline, two row bands), same timing method as the N-GPU lines"
    return out if rank == 0 else None


def single_frame_c_abi_leg(args, world, local):
    """The same single frame through the C entry points (srt_mgpu_*, one host thread, no Python / torch on the data path):
    rank 0 drives all N GPUs of the box; GPUs 1.. write their row blocks straight into GPU 0's frame through peer-mapped
    memory (NVLink stores from the shade kernel), joined by events.  Run while the other ranks wait at a barrier."""
    from simple_raytracer_b200 import _abi, capi, host, scenes
    out = {}
    m = capi.MultiGpu(list(range(world)))
    single = capi.Context(0)
    try:
        info = m.info()
        for key, angle in (("C3", 0.0), ("C3", 180.0), ("C4", 0.0)):
            cfg = scenes.CONFIGS[key]
            W, H, focal, N = cfg.width, cfg.height, cfg.focal, cfg.light_count
            hs = host.HostScene()
            view = cfg.build(hs, angle)
            light = np.asarray(scenes.light_view_space(hs, view), np.float32)
            arrays = hs.flatten()
            hs.close()
            m.upload(arrays)
            p = _abi.default_params(W, H, focal, light, N, True)
            dev_rgb = capi.lib().srt_device_alloc(single.h, W * H * 3)
            ms = []
            for it in range(15):
                m.render_frame_raw(p, rgb=dev_rgb, row_block=args.row_block)
                m.sync()
                if it >= 3:
                    ms.append(m.last_stats()["kernel_ms"])
            got = m.render_frame(p, row_block=args.row_block, ids=False, want_t=False)["rgb"]
            sc = single.upload(arrays)
            want = single.render(sc, p, ids=False, want_t=False)["rgb"]
            sc.destroy()
            capi.lib().srt_device_free(single.h, dev_rgb)
            out[f"{key}_{W}x{H}_{N}lights_{angle:g}deg"] = {"ms_per_frame": float(np.mean(ms)), "ms_best": float(np.min(ms)),
                                                            "frames_per_s": 1e3 / float(np.mean(ms)), "identical": bool(np.array_equal(got, want))}
        out["peer_writes"] = info["peer_writes"]
        out["what"] = (f"srt_mgpu_render_frame over {world} GPU(s) from one host thread: fork on GPU 0's stream, every GPU renders its "
                       f"interleaved {args.row_block}-row blocks into GPU 0's device frame ({info['peer_writes']} of {world} through direct / peer-mapped stores), "
                       "join by events; ms = CUDA events on GPU 0's stream from fork to join, mean of 12 frames after 3 warm-up (L2 not flushed)")
    finally:
        m.destroy()
        single.destroy()
    return out


# ------------------------------------------------------------------------------------------------
def run_b200_arm(args):
    import torch
    import torch.distributed as dist
    from simple_raytracer_b200 import _abi, capi, host

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available() or capi.device_count() == 0:
        raise RuntimeError("bench.py: no CUDA device — the render path has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    stream = torch.cuda.current_stream()
    ctx = capi.Context(local, stream=stream.cuda_stream)
    info = ctx.info()

    cfg, W, H, focal, N = frame_setup(args)
    K, Wm = args.steps, args.warmup
    n_unique, total_frames = run_shape(args, world)   # distinct series frames kept resident per rank; views of the orbit

    # ---- host prep (outside every timed region; reported) ----
    t_prep0 = time.time()
    frames = []
    for s in range(n_unique):
        f = s * world + rank
        _b, arrays, light = build_frame(host.HostScene, cfg, frame_angle(f, total_frames, args.still))
        frames.append((arrays, light))
        del _b
    prep_ms = 1e3 * (time.time() - t_prep0) / max(n_unique, 1)
    scenes_dev = [ctx.upload(a) for (a, _l) in frames]
    params = []
    for (_a, light) in frames:
        p = _abi.default_params(W, H, focal, light, N, cfg.use_bvh)
        p.variant = args.variant
        params.append(p)
    n_obj = len(frames[0][0].objects)
    scene_bytes = frames[0][0].nbytes()

    rgb = torch.empty((H, W, 3), dtype=torch.uint8, device=dev)
    flush = torch.empty(L2_FLUSH_BYTES, dtype=torch.uint8, device=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def render_step(s):
        k = s % n_unique
        ctx.render_raw(scenes_dev[k], params[k], rgb=rgb.data_ptr())

    hit_cache = {}

    def hits_of(s):
        k = s % n_unique
        if k not in hit_cache:
            render_step(s)
            hit_cache[k] = ctx.last_stats()["hit_pixels"]
        return hit_cache[k]

    # ---- device-timed region: scenes resident in HBM ----
    for s in range(Wm):
        flush.zero_()
        render_step(s)
    rays_local = sum(ray_count(W, H, N, hits_of(Wm + s), n_obj) for s in range(K))
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    kernel_ms = []
    barrier()
    sampler = ClockSampler(local) if rank == 0 else None
    t0 = time.time()
    for s in range(K):
        flush.zero_()                               # L2 flush between timed iterations (256 MiB = 2 x the 126 MB L2)
        ev[s][0].record()
        render_step(Wm + s)
        ev[s][1].record()
    barrier()
    t1 = time.time()
    clocks = sampler.stop(t0, t1) if sampler else None
    step_ms = [a.elapsed_time(b) for a, b in ev]
    total_ms = float(sum(step_ms))
    # the render kernel alone (events inside srt_render, same stream) for the roofline
    for s in range(K):                                   # the same K frames as `value`
        flush.zero_()
        render_step(Wm + s)
        kernel_ms.append(ctx.last_stats()["kernel_ms"])
    kern_ms = float(np.mean(kernel_ms)) if kernel_ms else float("nan")
    launches_per_step = ctx.last_stats()["kernel_launches"]
    # per-pass device times of the same K frames, rendered in one piece (with two bands the passes of the halves interleave)
    stages = None
    if rank == 0 and args.variant == 0 and cfg.use_bvh:
        bands_before = ctx.set_bands(1)
        ctx.set_stage_timing(True)
        try:
            acc = dict(classify=0.0, primary=0.0, shadow=0.0, shade=0.0)
            for s in range(K):
                flush.zero_()
                render_step(Wm + s)
                for k_, v_ in ctx.last_stage_ms().items():
                    acc[k_] += v_ / K
            stages = acc
        except capi.SrtError:
            stages = None
        finally:
            ctx.set_stage_timing(False)
            ctx.set_bands(bands_before)
    # the metric's own configuration (C3 = frame 0 of the orbit) on its own, device-timed the same way
    c3_still = None
    if rank == 0 and not args.still:
        k0 = 0                                            # frames[0] is global frame `rank` = 0 on rank 0: angle 0
        evs = []
        for it in range(3 + 8):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); ctx.render_raw(scenes_dev[k0], params[k0], rgb=rgb.data_ptr()); e1.record()
            evs.append((e0, e1))
        torch.cuda.synchronize()
        ms0 = float(np.mean([a.elapsed_time(b) for a, b in evs[3:]]))
        rays0 = ray_count(W, H, N, hits_of(k0), n_obj)
        c3_still = {"ms_per_frame": ms0, "value": rays0 / (ms0 * 1e-3) / 1e6, "unit": UNIT, "rays_per_frame": rays0,
                    "what": "frame 0 of the orbit = BASELINE configs[2] (C3) itself, same timing method"}
    # the same K frames rendered in one piece (srt_context_set_bands(ctx, 1); the default is two row bands per frame)
    one_piece = None
    if rank == 0 and not args.still and args.variant == 0:
        bands_before = ctx.set_bands(1)
        try:
            for s in range(3):
                flush.zero_(); render_step(Wm + s)
            evb = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
            for s in range(K):
                flush.zero_()
                evb[s][0].record(); render_step(Wm + s); evb[s][1].record()
            torch.cuda.synchronize()
            msb = float(sum(a.elapsed_time(b) for a, b in evb))
            one_piece = {"value": rays_local / (msb * 1e-3) / 1e6, "unit": UNIT, "ms_per_step": msb / K,
                         "what": "rank 0, the same K frames and timing as `value`, after srt_context_set_bands(ctx, 1): every frame in "
                                 "one piece on one stream (the default renders the upper and lower half on two streams; same bytes)"}
        finally:
            ctx.set_bands(bands_before)
    # algorithmic work of frame Wm (the first timed one) under the REFERENCE's traversal rules, counted by the
    # instrumented CUDA kernel (SRT_VARIANT_COUNT; tests check its counters against the oracle's)
    work = None
    if rank == 0 and cfg.use_bvh:
        tot = dict(node_visits=0, tri_tests=0, hit_pixels=0, light_evals=0)
        nk = K                                           # the same frames `value` and kernel_ms cover
        for s in range(nk):
            k0 = (Wm + s) % n_unique
            pc = _abi.default_params(W, H, focal, frames[k0][1], N, cfg.use_bvh)
            pc.variant = _abi.VARIANT_COUNT
            ctx.render_raw(scenes_dev[k0], pc, rgb=rgb.data_ptr())
            st = ctx.last_stats()
            tot["node_visits"] += st["node_visits"]; tot["tri_tests"] += st["tri_tests"]
            tot["hit_pixels"] += st["hit_pixels"]; tot["light_evals"] += N * st["hit_pixels"]
        work = {k: v / nk for k, v in tot.items()}

    # ---- end to end through the C ABI from pinned host buffers ----
    e2e = None
    if not args.no_e2e:
        pinned = []
        # one copy of the decoded textures for all frames, like a caller with a texture cache holds (the per-frame scenes
        # of this bench were flattened from separate ObjectManagers, each with its own copy of the same pixels):
        # srt_scene_update keeps a resident texture whose host pointer, size and sampled fingerprint are unchanged
        tex0 = frames[0][0].textures
        for (a, _l) in frames:
            if a.textures is not tex0 and len(a.textures) == len(tex0) and all(np.array_equal(x, y) for x, y in zip(a.textures, tex0)):
                a.textures = tex0
            keep = {}
            for name in ("nodes", "geom", "attr", "objects", "flat_geom", "flat_attr"):
                src = getattr(a, name)
                if src is None:
                    continue
                pa = capi.PinnedArray(src.shape, src.dtype)
                pa.array[...] = src
                keep[name] = pa
            pinned.append((_abi.FlatSceneArrays(keep["nodes"].array, keep["geom"].array, keep["attr"].array,
                                                keep["objects"].array, a.textures,
                                                keep["flat_geom"].array if "flat_geom" in keep else None,
                                                keep["flat_attr"].array if "flat_attr" in keep else None), keep))
        do_gather = args.gather and world > 1
        host_rgb = torch.empty((world if (rank == 0 and do_gather) else 1, H, W, 3), dtype=torch.uint8).pin_memory()
        gathered = [torch.empty_like(rgb) for _ in range(world)] if (rank == 0 and do_gather) else None
        h2d = sum(getattr(pinned[0][0], n).nbytes for n in ("nodes", "geom", "attr", "objects", "flat_geom", "flat_attr")
                  if getattr(pinned[0][0], n) is not None)
        d2h = host_rgb.numel() if (rank == 0 or not do_gather) else 0
        # Two contexts on two streams, frames alternating between them: the upload of frame k+1 and the download of
        # frame k-1 run on the copy engines while frame k renders (frames of a series are independent).  With
        # --gather the frames go through NCCL on one stream instead.
        n_lanes = 1 if do_gather else max(1, args.e2e_lanes)
        extra = []
        for _i in range(n_lanes - 1):
            st_ = torch.cuda.Stream(device=dev)
            extra.append((capi.Context(local, stream=st_.cuda_stream), st_))
        e2e_bands_before = ctx.set_bands(args.e2e_bands) if args.e2e_bands else None
        lanes = [dict(ctx=ctx, stream=stream, scene=ctx.upload(pinned[0][0]), rgb=rgb, host=host_rgb)]
        for cx, st_ in extra:
            if args.e2e_bands:
                cx.set_bands(args.e2e_bands)
            lanes.append(dict(ctx=cx, stream=st_, scene=cx.upload(pinned[0][0]), rgb=torch.empty_like(rgb),
                              host=torch.empty((1, H, W, 3), dtype=torch.uint8).pin_memory()))

        def e2e_step(s):
            k = s % n_unique
            ln = lanes[s % len(lanes)]
            with torch.cuda.stream(ln["stream"]):
                flush.zero_()                                            # L2 flush between iterations
                ln["scene"].update(pinned[k][0])                         # H2D of this frame's scene (async, pinned)
                ln["ctx"].render_raw(ln["scene"], params[k], rgb=ln["rgb"].data_ptr())   # the render kernels
                if do_gather:
                    dist.gather(ln["rgb"], gathered, dst=0)              # framebuffers -> rank 0 over NVLink (NCCL)
                    if rank == 0:
                        for r in range(world):
                            host_rgb[r].copy_(gathered[r], non_blocking=True)
                else:
                    ln["host"][0].copy_(ln["rgb"], non_blocking=True)    # D2H of this rank's finished frame

        for s in range(Wm):
            e2e_step(s)
        barrier()
        e_start, e_end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e_join = torch.cuda.Event()
        e_start.record(stream)
        for _cx, st_ in extra:
            st_.wait_event(e_start)
        for s in range(K):
            e2e_step(Wm + s)
        for _cx, st_ in extra:
            ej = torch.cuda.Event()
            ej.record(st_)
            stream.wait_event(ej)
        e_end.record(stream)
        barrier()
        e2e_ms = float(e_start.elapsed_time(e_end))                      # includes the K L2 flushes
        e2e = (e2e_ms, h2d, d2h)
        for ln in lanes:
            ln["scene"].destroy()
        for cx, _st in extra:
            cx.destroy()
        if e2e_bands_before:
            ctx.set_bands(e2e_bands_before)

    # ---- the reference's own result format (ImageData: sparse float lists, column-major) through the C ABI ----
    imagedata = None
    if not args.no_e2e and rank == 0 and not args.still:
        fp = capi.C.POINTER(capi.C.c_float)
        pts = capi.PinnedArray((W * H, 2), np.float32)
        col = capi.PinnedArray((W * H, 3), np.float32)
        sc2 = ctx.upload(pinned[0][0])
        n_out = capi.C.c_uint64(0)

        def imagedata_step(s):
            k = s % n_unique
            sc2.update(pinned[k][0])
            capi._check(capi.lib().srt_render_sparse(ctx.h, sc2.h, capi.C.byref(params[k]), capi.C.byref(n_out)))
            capi._check(capi.lib().srt_fetch_sparse(ctx.h, pts.array.ctypes.data_as(fp), col.array.ctypes.data_as(fp), n_out.value))
            return n_out.value

        for s in range(Wm):
            imagedata_step(s)
        torch.cuda.synchronize()
        t_a = time.perf_counter()
        entries = sum(imagedata_step(Wm + s) for s in range(K))
        torch.cuda.synchronize()
        ms = 1e3 * (time.perf_counter() - t_a) / K
        # the same lists with two frames in flight: two contexts take frames in turn; while the host waits for one lane's
        # lists (27 MB per frame over PCIe), the other lane renders (srt_render_sparse_begin / _count / srt_fetch_sparse)
        il = [dict(ctx=ctx, scene=sc2, pts=pts, col=col, busy=False)]
        for _i in range(1):                              # two lanes: a third measured slower here (1.53 vs 1.48 ms; the lists' 27 MB per frame bound the lane)
            ctx_c = capi.Context(local)
            il.append(dict(ctx=ctx_c, scene=ctx_c.upload(pinned[0][0]), pts=capi.PinnedArray((W * H, 2), np.float32),
                           col=capi.PinnedArray((W * H, 3), np.float32), busy=False))

        def retire(ln):
            if not ln["busy"]:
                return 0
            capi._check(capi.lib().srt_render_sparse_count(ln["ctx"].h, capi.C.byref(n_out)))
            # (a consumer reads the lane's previous lists HERE: the count's return means the copies enqueued before this frame are done)
            capi._check(capi.lib().srt_fetch_sparse_async(ln["ctx"].h, ln["pts"].array.ctypes.data_as(fp), ln["col"].array.ctypes.data_as(fp), n_out.value))
            ln["busy"] = False
            return n_out.value

        def piped_step(s):
            k = s % n_unique
            ln = il[s % len(il)]
            got = retire(ln)
            ln["scene"].update(pinned[k][0])
            capi._check(capi.lib().srt_render_sparse_begin(ln["ctx"].h, ln["scene"].h, capi.C.byref(params[k])))
            ln["busy"] = True
            return got

        for s in range(Wm):
            piped_step(s)
        for ln in il:
            retire(ln)
        torch.cuda.synchronize()
        t_b = time.perf_counter()
        entries_p = sum(piped_step(Wm + s) for s in range(K)) + sum(retire(ln) for ln in il)
        torch.cuda.synchronize()
        ms_p = 1e3 * (time.perf_counter() - t_b) / K
        imagedata = {"ms_per_frame": ms_p, "frames_per_s": 1e3 / ms_p, "Mrays_per_s": rays_local / K / ms_p / 1e3,
                     "entries_per_frame": entries_p / K, "d2h_bytes_per_frame": 20 * entries_p / K,
                     "h2d_bytes_per_frame": int(sum(getattr(pinned[0][0], n).nbytes for n in ("nodes", "geom", "attr", "objects"))),
                     "synchronous": {"ms_per_frame": ms, "frames_per_s": 1e3 / ms, "entries_per_frame": entries / K},
                     "what": "rank 0, the reference's own return type end to end: per frame srt_scene_update from pinned host arrays + render + "
                             "device-side compaction into the ImageData layout (simple_raytracer.cpp:505-525: non-black pixels, column-major, "
                             f"floats) + download of both lists into pinned host memory; {len(il)} contexts take frames in turn "
                             "(srt_render_sparse_begin / _count / srt_fetch_sparse_async: nothing blocks the host but the wait for a frame's count), host wall clock over the K frames incl. the final sync; `synchronous` = one "
                             "context, one blocking srt_render_sparse + srt_fetch_sparse per frame (what srt_dropin.hpp's single call does)"}
        for ln in il[1:]:
            ln["scene"].destroy(); ln["pts"].free(); ln["col"].free()
            ln["ctx"].destroy()
        sc2.destroy()
        pts.free(); col.free()

    # ---- image series including host prep (parse cache + prep thread pool + upload + render + download) ----
    series = None
    if args.series_frames != 0 and not args.still:
        from simple_raytracer_b200 import scenes as _sc, series as _series
        threads = max(1, len(os.sched_getaffinity(0)) // max(1, int(os.environ.get("LOCAL_WORLD_SIZE", world))))
        n_fr = args.series_frames if args.series_frames > 0 else min(96, 3 * threads)
        sr = _series.SeriesRenderer(ctx, cfg, W, H, focal, N, workers=threads, variant=args.variant)
        my = [360.0 * (i * world + rank) / (n_fr * world) for i in range(n_fr)]
        list(sr.frames(my[:2]))                                        # warm-up: fills the asset cache
        barrier()
        t_s0 = time.time()
        tms = [tm for (_i, _rgb, tm) in sr.frames(my)]                 # frames arrive in the lanes' pinned host buffers
        waits = [tm["wait_prep_ms"] for tm in tms]
        gpu_leg = float(np.mean([tm["gpu_ms"] for tm in tms]))
        torch.cuda.synchronize()
        t_s1 = time.time()
        sr.close()
        series = (t_s1 - t_s0, n_fr, threads, float(np.mean(waits)), gpu_leg)
        # the same series with the per-frame prep on the GPU (srt_scene_prep): no host pool at all
        sd = _series.SeriesRenderer(ctx, cfg, W, H, focal, N, workers=1, variant=args.variant, prep="device", lanes=args.series_lanes)
        n_fd = max(2 * n_fr, 96)
        myd = [360.0 * (i * world + rank) / (n_fd * world) for i in range(n_fd)]
        list(sd.frames(myd[:2 * args.series_lanes]))                   # warm-up: meshes parsed + uploaded once, every lane's scene allocated
        barrier()
        t_d0 = time.time()
        tmd = [tm for (_i, _rgb, tm) in sd.frames(myd)]
        torch.cuda.synchronize()
        t_d1 = time.time()
        series_dev = (t_d1 - t_d0, n_fd, float(np.mean([tm["issue_ms"] for tm in tmd])), sd.host_fallbacks)
        sd.close()

    # ---- one frame over all GPUs (strong scaling) ----
    single_frame = single_frame_c = None
    if not args.no_single_frame and not args.still and args.config == "C3":
        single_frame = single_frame_leg(args, ctx, rank, world, local, dev, barrier)
        barrier()
        # rank 0 drives all GPUs alone; the other ranks wait on the HOST (a key of the rendezvous store) — an NCCL barrier
        # would spin in a kernel on their GPUs and time-slice against rank 0's render kernels there
        store = None
        if world > 1:
            try:
                store = dist.distributed_c10d._get_default_store()
            except Exception:
                store = None                     # no store to wait on: the leg is skipped on every rank alike
        if world > 1 and store is None:
            single_frame_c = {"skipped": "no rendezvous store to park the other ranks on"} if rank == 0 else None
        elif rank == 0:
            try:
                single_frame_c = single_frame_c_abi_leg(args, world, local)
            except Exception as e:               # reporting leg: never takes the bench line down
                single_frame_c = {"failed": str(e)}
            if store is not None:
                store.set("srt_single_frame_c_abi_done", "1")
        elif store is not None:
            store.wait(["srt_single_frame_c_abi_done"])
        barrier()

    # ---- max over ranks, totals ----
    agg = torch.tensor([total_ms, e2e[0] if e2e else 0.0, kern_ms, prep_ms, series[0] if series else 0.0,
                        series_dev[0] if series else 0.0], dtype=torch.float64, device=dev)
    tot = torch.tensor([float(rays_local)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(agg, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    total_ms, e2e_ms, kern_ms_max, prep_ms_max, series_s, series_dev_s = [float(x) for x in agg.tolist()]
    rays_total = float(tot.item())

    line = None
    if rank == 0:
        value = rays_total / (total_ms * 1e-3) / 1e6
        peaks = {}
        try:
            peaks = json.loads((REPO / "MEASURED_PEAKS.json").read_text())
        except Exception:
            pass
        hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
        # algorithmic HBM bytes of one launch: one read of the scene + one write of the RGB8 frame
        alg_bytes = scene_bytes + W * H * 3
        ach = alg_bytes / (kern_ms * 1e-3) / 1e9
        # measured traffic / issue figures of the three render kernels: the committed ncu capture (tools/ncu_summary.py --json)
        ncu = None
        ncu_path = REPO / "profiles" / "r02_stage_metrics.json"
        try:
            if args.config == "C3" and args.variant == 0:
                ncu = json.loads(ncu_path.read_text())
        except Exception:
            ncu = None
        ncu_stage = {}
        if ncu:
            for kr in ncu["kernels"]:
                for tag, name in (("primary", "render_phase<0"), ("shadow", "render_phase<1"), ("shade", "render_phase<2")):
                    if name in kr["kernel"] and tag not in ncu_stage:
                        ncu_stage[tag] = kr["metrics"]
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": Wm,
            "ms_per_step": total_ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": make_config(args, cfg, W, H, focal, N, world, len(frames[0][0].geom), len(frames[0][0].nodes)),
            "timing": {"l2": f"flushed between timed iterations ({L2_FLUSH_BYTES >> 20} MiB device memset, 2 x the 126 MB L2)", "variant": args.variant,
                       "sm_count": info["sm_count"], "row_bands": 2 if args.variant == 0 else 1,
                       "events": "CUDA events on the launching stream around every srt_render, max over ranks"},
            "frames_per_s": world * K / (total_ms * 1e-3),
            "rays_per_frame": rays_total / (K * world),
            "kernel_ms": kern_ms,
            "kernel_ms_what": "mean over the same K frames of the span first launch -> last launch end inside srt_render (both row bands)",
            "host_prep_ms_per_frame": prep_ms_max,
            "gpu_launches": K * world * launches_per_step,
            "clocks": clocks,
            "roofline_hbm": {"bound": "hbm", "achieved": ach, "peak": hbm_peak, "unit": "GB/s", "frac": ach / hbm_peak,
                             "traffic": (sum(m.get("dram__bytes_read.sum", 0.0) + m.get("dram__bytes_write.sum", 0.0) for m in ncu_stage.values())
                                         if len(ncu_stage) == 3 else None),
                             "traffic_source": f"ncu --set full of one frame's three render kernels, caches cold per kernel: profiles/{ncu_path.name}" if ncu_stage else None,
                             "peak_source": "measured (MEASURED_PEAKS.json)" if peaks else "fallback",
                             "note": "the scene is L2-resident by design: compulsory HBM bytes = one scene read + one RGB8 frame write; "
                                     "this object is here because the contract asks for an hbm|tensor roofline — the binding bound is `roofline`"},
        }
        if stages:
            line["stage_ms_one_piece"] = dict(stages, what="mean device time of each pass over the same K frames, rendered in one piece "
                                                          "(srt_context_set_stage_timing: events between the launches)")
        if work:
            # F_alg = 12 V + 45 T + 75 S + 30 hit flops, B_alg,L2 = 32 V + 48 T + 48 hit bytes per frame
            V, T, S, hp = work["node_visits"], work["tri_tests"], work["light_evals"], work["hit_pixels"]
            flops = 12.0 * V + 45.0 * T + 75.0 * S + 30.0 * hp
            l2b = 32.0 * V + 48.0 * T + 48.0 * hp
            sm_hz = (clocks or {}).get("sm_mhz") or float(peaks.get("sm_max_mhz", 1965.0))
            fp32_peak = info["sm_count"] * 128 * sm_hz * 1e6 / 1e12          # non-FMA: the parity arithmetic may not contract
            try:                                                              # pinned by a read microbenchmark on this box
                l2_peak = max(ctx.measure_read_bandwidth(32 << 20, 40) for _ in range(3))
                l2_src = "measured in this run: all SMs streaming a 32 MiB L2-resident buffer with 16-byte loads (srt_measure_read_bandwidth), best of 3"
            except Exception:
                l2_peak = 6300.0 * sm_hz * 1e6 / 1e9
                l2_src = "6300 B/cycle (microarchitecture guide, measured on B300) x SM clock"
            line["algorithmic_work_per_frame"] = dict(work, flops=flops, l2_bytes=l2b,
                                                      rule="reference traversal rules (S3/S5/S11), own-object shadow traversal excluded")
            line["roofline_fp32"] = {"bound": "fp32 issue (non-FMA)", "achieved": flops / (kern_ms * 1e-3) / 1e12, "peak": fp32_peak,
                                     "unit": "TFLOP/s", "frac": flops / (kern_ms * 1e-3) / 1e12 / fp32_peak,
                                     "peak_source": "148 SMs x 128 lanes x SM clock under load"}
            l2_measured = (sum(m.get("l1tex__m_xbar2l1tex_read_bytes.sum", 0.0) + m.get("l1tex__m_l1tex2xbar_write_bytes.sum", 0.0) for m in ncu_stage.values())
                           if len(ncu_stage) == 3 else None)
            line["roofline_l2"] = {"bound": "l2", "achieved": l2b / (kern_ms * 1e-3) / 1e9, "peak": l2_peak, "unit": "GB/s",
                                   "frac": l2b / (kern_ms * 1e-3) / 1e9 / l2_peak, "algorithmic_bytes": l2b,
                                   "measured_l2_bytes": l2_measured,
                                   "measured_frac": (l2_measured / (kern_ms * 1e-3) / 1e9 / l2_peak) if l2_measured else None,
                                   "measured_source": f"l1tex__m_xbar2l1tex_read_bytes + l1tex__m_l1tex2xbar_write_bytes (what crosses the crossbar between L2 and the SMs' L1) of the three render kernels of ONE frame (the capture's frame, not the K-frame mean): profiles/{ncu_path.name}" if l2_measured else None,
                                   "note": "`achieved` is the ALGORITHMIC byte rate (every box / triangle fetched from L2 once per test); L1 absorbs most of "
                                           "those fetches, so the bytes that really cross L2 -> SM are `measured_l2_bytes`",
                                   "peak_source": l2_src}
            # The binding bound: the kernels are issue-bound with half-empty warps.  Achieved = issue-slot utilisation x active
            # threads per issued instruction / 32, time-weighted over the three render kernels of the ncu capture: the
            # fraction of the machine's thread-instruction slots that carry a live thread.
            if len(ncu_stage) == 3:
                tot_t = sum(m.get("gpu__time_duration.sum", 0.0) for m in ncu_stage.values())
                per = {}
                acc = 0.0
                for tag, m in ncu_stage.items():
                    iss = m.get("smsp__issue_active.avg.pct_of_peak_sustained_active", 0.0) / 100.0
                    act = m.get("smsp__thread_inst_executed_per_inst_executed.ratio", 0.0) / 32.0
                    busy = (m.get("sm__cycles_active.avg", 0.0) / m.get("sm__cycles_elapsed.avg", 1.0)) if m.get("sm__cycles_elapsed.avg") else 1.0
                    per[tag] = {"time_share": m.get("gpu__time_duration.sum", 0.0) / tot_t if tot_t else None, "issue_slots_busy": iss,
                                "sm_active_share": busy, "active_threads_of_32": act * 32.0, "thread_slots_used": iss * act * busy,
                                "l1_hit_pct": m.get("l1tex__t_sector_hit_rate.pct"), "l2_hit_pct": m.get("lts__t_sector_hit_rate.pct"),
                                "l2_to_sm_bytes": m.get("l1tex__m_xbar2l1tex_read_bytes.sum"), "registers": m.get("launch__registers_per_thread")}
                    acc += iss * act * busy * (m.get("gpu__time_duration.sum", 0.0) / tot_t if tot_t else 0.0)
                line["roofline"] = {"bound": "issue (thread-instruction slots)", "achieved": acc, "peak": 1.0, "unit": "fraction of thread-instruction issue slots",
                                    "frac": acc, "traffic": line["roofline_hbm"]["traffic"], "per_kernel": per,
                                    "source": f"profiles/{ncu_path.name} (ncu --set full, one heavy mid-orbit frame, one piece)",
                                    "note": "issue-slot utilisation x SM-active share x active threads per instruction / 32, time-weighted over primary / shadow / shade; "
                                            "the contract's hbm object is `roofline_hbm`, the L2 and FP32 figures `roofline_l2` / `roofline_fp32`"}
        # the second bound of the dominant (shadow) kernel: L1 data-pipe wavefronts, from the end-of-round capture that holds the LSU metrics
        try:
            if args.config == "C3" and args.variant == 0:
                l1_path = REPO / "profiles" / "r02f_c3_stage_metrics.json"
                for kr in json.loads(l1_path.read_text())["kernels"]:
                    m = kr["metrics"]
                    if "render_phase<1" in kr["kernel"] and "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed" in m:
                        f = m["l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed"] / 100.0
                        rq = m.get("l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum")
                        line["roofline_l1"] = {"bound": "l1 data pipe (LSU wavefronts) of the shadow pass", "achieved": f, "peak": 1.0,
                                               "unit": "wavefronts per clock per SM", "frac": f,
                                               "frac_of_active_cycles": m.get("l1tex__throughput.avg.pct_of_peak_sustained_active", 0.0) / 100.0,
                                               "issue_slots_busy_same_capture": m.get("smsp__issue_active.avg.pct_of_peak_sustained_active", 0.0) / 100.0,
                                               "global_load_requests": rq, "source": f"profiles/{l1_path.name} (ncu --set full, the 180-degree frame, one piece)",
                                               "note": "a 16-byte load is four data-pipe wavefronts whatever the number of active lanes: the pass is within 15 % of "
                                                       "this bound and of the issue bound at once"}
                        break
        except Exception:
            pass
        if "roofline" not in line:
            line["roofline"] = dict(line["roofline_hbm"], note="no ncu capture for this configuration: the hbm object stands in; see roofline_l2 / roofline_fp32")
        if imagedata:
            line["e2e_imagedata"] = imagedata
        if e2e:
            line["e2e"] = {"value": rays_total / (e2e_ms * 1e-3) / 1e6, "unit": UNIT, "h2d_bytes_per_step": int(e2e[1]) * world,
                           "d2h_bytes_per_step": int(e2e[2]) * (1 if args.gather else world), "ms_per_step": e2e_ms / K,
                           "frames_per_s": world * K / (e2e_ms * 1e-3),
                           "what": f"per rank: srt_scene_update from pinned host arrays + srt_render + D2H of the RGB8 frame to pinned host memory, frames taken in turn by {max(1, args.e2e_lanes)} contexts / streams so that copies overlap renders; the L2 flushes between frames are inside this time"
                                   + (" (all frames gathered to rank 0 over NCCL first)" if args.gather else "")}
        if single_frame:
            line["single_frame"] = single_frame
        if single_frame_c:
            line["single_frame_c_abi"] = single_frame_c
        if c3_still:
            line["c3_frame0"] = c3_still
        if one_piece:
            line["value_one_piece"] = one_piece
        if series:
            line["series_with_host_prep"] = {
                "frames_per_s": world * series[1] / series_s, "frames": world * series[1], "host_threads_per_rank": series[2],
                "mean_wait_for_prep_ms": series[3], "mean_upload_render_download_ms": series[4],
                "what": "whole image-series pipeline per rank: cached assets, per-frame transform + BVH build + flatten on a host "
                        "thread pool (bit-identical to rebuilding from files), upload, render, download to pinned host memory on two "
                        "alternating contexts (the GPU leg overlaps the wait for the next frame's prep)"}
            line["series_with_device_prep"] = {
                "frames_per_s": world * series_dev[1] / series_dev_s, "frames": world * series_dev[1], "host_threads_per_rank": 1,
                "mean_issue_ms": series_dev[2], "host_fallback_frames": series_dev[3], "lanes": args.series_lanes,
                "what": "the same pipeline with the per-frame prep on the GPU: meshes resident in object space, per frame the "
                        "recorded scene script (matrices) -> srt_scene_prep (transformTriangles + createBoundingHierarchy + "
                        "flatten as CUDA kernels, bit-identical to the host result) -> render -> download; one host thread"}
        if world == 1 and not args.no_cpu_baseline:
            try:
                timed = [frame_angle(((Wm + s_) % n_unique) * world + rank, total_frames, args.still) for s_ in range(K)]
                pick = [timed[0]] if args.still else [timed[(i * len(timed)) // 4] for i in range(4)]   # four of the timed views
                nv = len(pick)
                step, cinfo, _si = cpu_reference_sample(cfg, W, H, focal, N, args.cpu_seconds / nv, pick)
                rs = [step(i) for i in range(nv)]
                r, s = sum(x[0] for x in rs), sum(x[1] for x in rs)
                line["cpu_baseline"] = {"value": r / s / 1e6, "unit": UNIT, **cinfo, "seconds": s}
                # also the reference as shipped — one thread — on a smaller sample of the same views
                step1, cinfo1, _si = cpu_reference_sample(cfg, W, H, focal, N, min(args.cpu_seconds, 4.0) / nv, pick, threads=1)
                rs1 = [step1(i) for i in range(nv)]
                r1, s1 = sum(x[0] for x in rs1), sum(x[1] for x in rs1)
                line["cpu_baseline"]["one_core"] = {"value": r1 / s1 / 1e6, "unit": UNIT, "cores": 1, "sample": cinfo1["sample"], "seconds": s1}
            except Exception as e:      # the CPU baseline is reporting, never the product path
                line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": 0, "kind": "port", "sample": f"failed: {e}"}
        print(json.dumps(line))
    for s_ in scenes_dev:
        s_.destroy()
    ctx.destroy()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def main():
    args = parse_args()
    if args.impl == "reference":
        return run_reference_arm(args)
    return run_b200_arm(args)


if __name__ == "__main__":
    sys.exit(main())
