"""Frames/s of the image-series driver (C5 workload: C3 scene over a camera orbit, 1080p, 16 lights) for a number of
frames in flight, device prep against host prep.   python tools/series_bench.py [--frames 96] [--lanes 1 2 3 4 6]"""
import argparse
import json
import sys
import time
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))

from simple_raytracer_b200 import capi, scenes, series  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=96)
    ap.add_argument("--lanes", type=int, nargs="+", default=[1, 2, 3, 4, 6])
    ap.add_argument("--config", default="C3")
    args = ap.parse_args()
    cfg = scenes.CONFIGS[args.config]
    ctx = capi.Context(0)
    angles = [360.0 * i / args.frames for i in range(args.frames)]
    out = {}
    for lanes in args.lanes:
        sr = series.SeriesRenderer(ctx, cfg, cfg.width, cfg.height, cfg.focal, cfg.light_count, workers=1, prep="device", lanes=lanes)
        list(sr.frames(angles[:4]))
        t0 = time.perf_counter()
        n = sum(1 for _ in sr.frames(angles))
        dt = time.perf_counter() - t0
        sr.close()
        out[f"device_prep_lanes_{lanes}"] = n / dt
    print(json.dumps(out))
    ctx.destroy()


if __name__ == "__main__":
    main()
