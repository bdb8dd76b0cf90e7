"""Times the per-frame scene prep of the C3 / C5 series: device path (srt_scene_prep) against the host mirror.

    python tools/prep_bench.py [--frames 48] [--config C3]
Device time = CUDA-side duration of one srt_scene_prep (events are not exposed for it: K calls back to back on one
stream, one sync, wall clock / K — the calls are asynchronous, so this is the GPU's rate unless the host loop is
slower, which is reported separately).  Host time = HostScene build + flatten on one core.
"""
import argparse
import json
import sys
import time
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))

from simple_raytracer_b200 import capi, host, prep, scenes  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", type=int, default=48)
    ap.add_argument("--config", default="C3")
    args = ap.parse_args()
    cfg = scenes.CONFIGS[args.config]
    ctx = capi.Context(0)
    cache = prep.MeshCache(ctx)
    host.set_asset_cache(True)
    angles = [360.0 * f / args.frames for f in range(args.frames)]

    # warm-up: parse + upload meshes, allocate
    ds = prep.DeviceScene(cache)
    cfg.build(ds, 0.0)
    sc = ds.commit()
    sc.prep_status()
    ds.close()

    # host-side recording cost alone (no GPU work): the matrices of a frame
    t0 = time.perf_counter()
    recs = []
    for a in angles:
        ds = prep.DeviceScene(cache)
        cfg.build(ds, a)
        recs.append(ds)
    t_record = (time.perf_counter() - t0) / len(angles)

    ctx.sync()
    t0 = time.perf_counter()
    for ds in recs:
        sc = ds.commit(scene=sc)
    t_issue = (time.perf_counter() - t0) / len(angles)
    sc.prep_status()
    t_dev = (time.perf_counter() - t0) / len(angles)
    for ds in recs:
        ds.close()

    t0 = time.perf_counter()
    n_host = min(len(angles), 8)
    for a in angles[:n_host]:
        hs = host.HostScene()
        cfg.build(hs, a)
        hs.flatten_desc()
        hs.close()
    t_host = (time.perf_counter() - t0) / n_host
    c = sc.counts()
    print(json.dumps(dict(config=args.config, frames=len(angles), triangles=c["n_tris"], nodes=c["n_nodes"],
                          device_prep_ms=1e3 * t_dev, device_issue_ms=1e3 * t_issue, record_ms=1e3 * t_record,
                          host_prep_ms_one_core=1e3 * t_host, speedup_vs_one_core=t_host / t_dev)))
    sc.destroy()
    cache.close()
    ctx.destroy()


if __name__ == "__main__":
    main()
