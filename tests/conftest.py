"""pytest configuration: markers, paths, shared scene fixtures.

`-m "not gpu"` runs everywhere (oracle vs golden vectors, host logic, ABI surface);
`-m gpu` are the parity tests proper: CUDA path through the C ABI vs the oracle.
"""
import os
import sys
from pathlib import Path

import numpy as np
import pytest

REPO = Path(__file__).resolve().parent.parent
if str(REPO) not in sys.path:
    sys.path.insert(0, str(REPO))

GOLDEN = REPO / "tests" / "golden"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu on the GPU box")
    config.addinivalue_line("markers", "ref: needs oracle/_ref (the reference compiled from /root/reference)")


def _has_gpu():
    try:
        from simple_raytracer_b200 import capi
        return capi.device_count() > 0
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    from oracle import pyoracle
    have_ref = pyoracle.ref_available()
    for item in items:
        if "ref" in item.keywords and not have_ref:
            item.add_marker(pytest.mark.skip(reason="oracle/_ref not built (needs /root/reference)"))


@pytest.fixture(scope="session")
def gpu_ctx():
    from simple_raytracer_b200 import capi
    if capi.device_count() == 0:
        pytest.fail("gpu test selected but no CUDA device is visible — the CUDA path has no fallback")
    ctx = capi.Context(0)
    yield ctx
    ctx.destroy()


def scene_from_reference(key, angle=0.0):
    """Build config `key` with the UNMODIFIED reference and flatten it.  Returns (arrays, light)."""
    from oracle import pyoracle
    from simple_raytracer_b200 import scenes
    cfg = scenes.CONFIGS[key]
    b = pyoracle.RefBackend()
    view = cfg.build(b, angle)
    light = scenes.light_view_space(b, view) if view is not None else np.array(scenes.LIGHT_WORLD, np.float32)
    arrays = b.flatten(with_flat_lists=not cfg.use_bvh)
    return b, arrays, np.asarray(light, np.float32)


_scene_cache = {}


@pytest.fixture(scope="session")
def ref_scene():
    """Cached (backend, arrays, light) per config key, built through oracle/_ref."""
    def get(key, angle=0.0):
        k = (key, float(angle))
        if k not in _scene_cache:
            _scene_cache[k] = scene_from_reference(key, angle)
        return _scene_cache[k]
    return get


def flat_from_fixture(fx):
    """FlatSceneArrays from the arrays embedded in a golden frame fixture."""
    from simple_raytracer_b200 import _abi
    return _abi.FlatSceneArrays(fx["nodes"], fx["geom"], fx["attr"], fx["objects"], (),
                                fx["flat_geom"] if "flat_geom" in fx.files else None,
                                fx["flat_attr"] if "flat_attr" in fx.files else None,
                                [str(x) for x in fx["object_names"]])


_fixture_scene_cache = {}


def scene_for_fixture(key, fx):
    """(arrays, light) for golden frame `key`: the embedded scene when the fixture carries one, else built
    with this repo's host mirror from the staged assets (byte-identical to the reference-built scene, textured C4
    included: tests/test_host_scene.py)."""
    from oracle import pyoracle
    from simple_raytracer_b200 import host, scenes
    angle = float(fx["angle"]) if "angle" in fx.files else 0.0
    key, cfg_key = (key, angle), key
    if key in _fixture_scene_cache:
        return _fixture_scene_cache[key]
    if "nodes" in fx.files:
        out = (flat_from_fixture(fx), np.asarray(fx["light"], np.float32))
    else:
        cfg = scenes.CONFIGS[cfg_key]
        try:
            host.assets_dir()
        except FileNotFoundError:
            pytest.skip("staged assets not present")
        b = host.HostScene()
        view = cfg.build(b, angle)
        light = scenes.light_view_space(b, view) if view is not None else np.array(scenes.LIGHT_WORLD, np.float32)
        out = (b.flatten(with_flat_lists=not cfg.use_bvh), np.asarray(light, np.float32))
    _fixture_scene_cache[key] = out
    return out


def compare_images(a, b):
    """max abs error and PSNR between two uint8 images."""
    d = a.astype(np.int32) - b.astype(np.int32)
    mx = int(np.abs(d).max()) if d.size else 0
    mse = float((d.astype(np.float64) ** 2).mean()) if d.size else 0.0
    psnr = float("inf") if mse == 0 else 10.0 * np.log10(255.0 ** 2 / mse)
    return mx, psnr
