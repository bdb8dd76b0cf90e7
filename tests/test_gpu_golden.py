"""GPU tests against the COMMITTED golden frames (outputs of the unmodified reference, tests/golden/):
the CUDA path through the C ABI, no oracle and no reference library involved at run time except for
building the textured scene."""
import numpy as np
import pytest

from conftest import GOLDEN, compare_images, scene_for_fixture
from simple_raytracer_b200 import _abi

pytestmark = pytest.mark.gpu

KEYS = ["C1", "C1b", "cubes4", "spheres", "shipped", "C2", "C3", "C4"]
ORBIT_KEYS = ["C3_a36", "C3_a90", "C3_a135", "C3_a180", "C3_a200", "C3_a270", "C3_a358p5"]      # the views bench.py times


@pytest.mark.parametrize("key", KEYS + ORBIT_KEYS)
def test_cuda_reproduces_reference_frames(gpu_ctx, key):
    fx = np.load(GOLDEN / f"frame_{key}.npz")
    arrays, _light = scene_for_fixture(key.split("_a")[0], fx)
    W, H, focal, N, use_bvh = [float(x) for x in fx["params"]]
    p = _abi.default_params(int(W), int(H), focal, fx["light"], int(N), bool(use_bvh))
    sc = gpu_ctx.upload(arrays)
    got = gpu_ctx.render(sc, p)
    sc.destroy()
    npx = int(W * H)
    assert np.count_nonzero(got["obj_id"] == fx["obj_id"]) == npx
    assert np.count_nonzero(got["tri_id"] == fx["tri_id"]) == npx
    assert np.array_equal(got["t"], fx["t"])
    mx, psnr = compare_images(got["rgb"], fx["rgb"])
    assert mx <= 1 and psnr >= 50.0          # tolerance of BASELINE.json north_star (powf is not bit-defined)


@pytest.mark.parametrize("prep_mode", ["host", "device"])
def test_series_driver_reproduces_reference_orbit_frames(gpu_ctx, prep_mode):
    """The image-series driver (BASELINE configs[4]) against frames rendered by the UNMODIFIED reference: host prep pool and
    device prep (srt_scene_prep), scene updated in place per lane, frames in flight on several streams — every frame within
    1 LSB of the reference's (the only tolerance: powf), not merely equal to another GPU frame."""
    from simple_raytracer_b200 import scenes, series
    fxs = [np.load(GOLDEN / f"frame_{k}.npz") for k in ORBIT_KEYS]
    W, H, focal, N, _bvh = [float(x) for x in fxs[0]["params"]]
    angles = [float(fx["angle"]) for fx in fxs]
    sr = series.SeriesRenderer(gpu_ctx, scenes.CONFIGS["C3"], int(W), int(H), focal, int(N), workers=3, prep=prep_mode)
    try:
        n = 0
        for i, rgb, _tm in sr.frames(angles):
            mx, psnr = compare_images(rgb, fxs[i]["rgb"])
            differing = int(np.count_nonzero((rgb != fxs[i]["rgb"]).any(axis=2)))
            assert mx <= 1 and psnr >= 50.0 and differing <= 3, (prep_mode, angles[i], mx, psnr, differing)
            n += 1
        assert n == len(angles) and sr.host_fallbacks == 0
    finally:
        sr.close()
