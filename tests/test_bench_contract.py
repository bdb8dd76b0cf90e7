"""bench.py's contract, checked where it can be without a GPU: the reference arm (the unmodified reference's CPU code,
oracle/_ref) prints ONE JSON line with the required keys, loads nothing of the product, and describes its run with the very
`config` object the GPU arm prints for the same arguments (the driver compares the two)."""
import argparse
import json
import os
import subprocess
import sys
from pathlib import Path

import pytest

REPO = Path(__file__).resolve().parent.parent


@pytest.mark.ref
def test_reference_arm_line_and_independence():
    code = (
        "import sys, runpy\n"
        "sys.argv = ['bench.py', '--impl', 'reference', '--steps', '2', '--warmup', '1']\n"
        "try:\n    runpy.run_path('bench.py', run_name='__main__')\nexcept SystemExit:\n    pass\n"
        "maps = open('/proc/self/maps').read()\n"
        "print('LIBS', sorted({l.split('/')[-1] for l in maps.splitlines() if 'libsrt' in l or 'liboracle' in l}))\n")
    r = subprocess.run([sys.executable, "-c", code], cwd=REPO, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    for k in ("impl", "metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
              "vs_baseline", "dtype", "data", "config", "cpu_baseline", "e2e", "gpu_launches"):
        assert k in d, k
    assert d["impl"] == "reference" and d["unit"] == "Mrays/s" and d["higher_is_better"] is True and d["gpu_launches"] == 0
    assert d["value"] > 0 and d["e2e"]["value"] == d["value"] and d["e2e"]["h2d_bytes_per_step"] == 0
    assert d["cpu_baseline"]["kind"] == "reference" and d["cpu_baseline"]["cores"] >= 1
    assert d["config"]["width"] == 1920 and d["config"]["height"] == 1080 and d["config"]["lights"] == 16
    assert d["config"]["triangles"] == 69463 and d["config"]["bvh_nodes"] == 24216
    libs = [l for l in r.stdout.splitlines() if l.startswith("LIBS")][0]
    assert "libsrt_ref.so" in libs and "libsrt_b200" not in libs and "dropin" not in libs, libs


def test_both_arms_describe_the_run_with_one_function():
    sys.path.insert(0, str(REPO))
    import bench
    from simple_raytracer_b200 import scenes
    args = argparse.Namespace(steps=20, warmup=5, config="C3", still=False)
    cfg = scenes.CONFIGS["C3"]
    a = bench.make_config(args, cfg, 1920, 1080, 1080.0, 16, 8, 69463, 24216)
    b = bench.make_config(args, cfg, 1920, 1080, 1080.0, 16, 8, 69463, 24216)
    assert a == b and a["parallelism"] == "frame-sharded x8" and "200 views" in a["views"]
    n_unique, total = bench.run_shape(args, 8)
    assert (n_unique, total) == (25, 200)
    # the reference arm's step s samples a view the GPU arm times in step s: view (s mod n_unique) * N + (s mod N)
    views = {(s % n_unique) * 8 + r for s in range(5, 25) for r in range(8)}
    assert all(((s % n_unique) * 8 + (s % 8)) in views for s in range(5, 25))
    assert bench.frame_angle(0, total) == 0.0 and abs(bench.frame_angle(100, total) - 180.0) < 1e-9
