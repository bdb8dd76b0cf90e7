"""CPU tests: the oracle (oracle/srt_oracle.c, the plain-C restatement) against the committed golden
vectors, which are outputs of the UNMODIFIED reference (tests/golden/make_golden.py), and — where
oracle/_ref exists — against the reference itself, live.  Everything here is bit-exact."""
import ctypes as C
from pathlib import Path

import numpy as np
import pytest

from conftest import GOLDEN, flat_from_fixture, scene_for_fixture
from oracle import pyoracle
from simple_raytracer_b200 import _abi

f32 = np.float32
_f32p = C.POINTER(C.c_float)
FRAME_KEYS = ["C1", "C1b", "cubes4", "spheres", "shipped", "C2", "C3", "C4"]
# views of the C3 / C5 scene around the camera orbit (tests/golden/make_golden.py ORBIT_ANGLES): series frames 24, 120, 239,
# the quarter turns and the two heaviest views — the frames bench.py times
ORBIT_KEYS = ["C3_a36", "C3_a90", "C3_a135", "C3_a180", "C3_a200", "C3_a270", "C3_a358p5"]


def fp(a):
    return np.ascontiguousarray(a, dtype=f32).ctypes.data_as(_f32p)


@pytest.fixture(scope="module")
def kat():
    return np.load(GOLDEN / "leaf_kat.npz")


def tri_edges(tri):
    """The flattener's p0 / e1 / e2 (rayTriangleIntersection :45-51) from 3 x vec4, in fp32."""
    p = tri.reshape(-1, 3, 4)
    c = (p[:, :, :3] / p[:, :, 3:4]).astype(f32)
    return c[:, 0], (c[:, 1] - c[:, 0]).astype(f32), (c[:, 2] - c[:, 0]).astype(f32)


def same_bits(a, b):
    a = np.ascontiguousarray(a, f32).view(np.uint32)
    b = np.ascontiguousarray(b, f32).view(np.uint32)
    nan = np.isnan(a.view(f32)) & np.isnan(b.view(f32))
    return bool(np.all((a == b) | nan | ((a.view(f32) == 0) & (b.view(f32) == 0))))


def test_moller_trumbore_kat(kat):
    lib = pyoracle.oracle_lib()
    p0, e1, e2 = tri_edges(kat["tri"])
    got = np.array([lib.srt_oracle_moller_trumbore(fp(kat["o"][i]), fp(kat["d"][i]), fp(p0[i]), fp(e1[i]), fp(e2[i]))
                    for i in range(len(p0))], f32)
    assert same_bits(got, kat["t"])
    assert np.isfinite(kat["t"]).sum() > 500            # the vectors do exercise hits


def test_slab_kat(kat):
    lib = pyoracle.oracle_lib()
    n = len(kat["o"])
    got = np.array([lib.srt_oracle_slab(fp(kat["o"][i]), fp(kat["d"][i]), fp(kat["bmin"][i]), fp(kat["bmax"][i])) for i in range(n)])
    assert np.array_equal(got != 0, kat["slab"] != 0)
    zero = np.zeros(3, f32)                             # intersectRayAabb :204-248 == origin (0,0,0)
    got0 = np.array([lib.srt_oracle_slab(fp(zero), fp(kat["d"][i]), fp(kat["bmin"][i]), fp(kat["bmax"][i])) for i in range(n)])
    assert np.array_equal(got0 != 0, kat["slab0"] != 0)


def test_barycentric_and_phong_kat(kat):
    lib = pyoracle.oracle_lib()
    p0, e1, e2 = tri_edges(kat["tri"])
    n = len(p0)
    white = np.ones(3, f32)
    bary = np.zeros((n, 3), f32)
    ph = np.zeros((n, 3), f32)
    for i in range(n):
        out = np.zeros(3, f32)
        lib.srt_oracle_barycentric(fp(p0[i]), fp(e1[i]), fp(e2[i]), fp(kat["point"][i]), fp(out))
        bary[i] = out
        out2 = np.zeros(3, f32)
        a, s, sh = [float(x) for x in kat["props"][i]]
        lib.srt_oracle_phong(fp(kat["normal"][i]), fp(kat["d"][i]), fp(kat["light"][i]), fp(white), fp(kat["color"][i]),
                             a, s, sh, float(kat["tdist"][i]), fp(out2))
        ph[i] = out2
    assert same_bits(bary, kat["bary"])
    assert same_bits(ph, kat["phong"])


def test_tonemap_quantise_kat(kat):
    lib = pyoracle.oracle_lib()
    m = len(kat["tone_in"])
    for i in range(m):
        out = np.zeros(3, f32)
        q = np.zeros(3, np.int32)
        lib.srt_oracle_tonemap_quantise(fp(kat["tone_in"][i]), 0.5, f32(1.1), fp(out), q.ctypes.data_as(C.POINTER(C.c_int32)))
        assert same_bits(out, kat["tone"][i]), i
        assert np.array_equal(q & 0xff, kat["q"][i] & 0xff), i


def test_light_staircase_matches_reference_sequence():
    """softShadow :363-383: +3 on x, y, z cyclically, sequential fp32 adds."""
    lib = pyoracle.oracle_lib()
    L0 = np.array([-200.00002, 8.493649, -514.7114], f32)
    out = np.zeros((16, 3), f32)
    lib.srt_oracle_light_positions(fp(L0), 16, 3.0, fp(out))
    want = np.zeros((16, 3), f32)
    L = L0.copy()
    for i in range(16):
        want[i] = L
        L[i % 3] = f32(L[i % 3] + f32(3.0))
    assert same_bits(out, want)


@pytest.mark.parametrize("key", FRAME_KEYS + ORBIT_KEYS)
def test_oracle_reproduces_reference_frames(key):
    """Whole frames rendered by the reference (committed) == the oracle on the same scene, every plane."""
    fx = np.load(GOLDEN / f"frame_{key}.npz")
    arrays, light = scene_for_fixture(key.split("_a")[0], fx)
    W, H, focal, N, use_bvh = [float(x) for x in fx["params"]]
    assert np.allclose(light[:3], fx["light"], rtol=0, atol=0)
    p = _abi.default_params(int(W), int(H), focal, fx["light"], int(N), bool(use_bvh))
    got = pyoracle.oracle_render(arrays, p)
    assert np.array_equal(got["obj_id"], fx["obj_id"])
    assert np.array_equal(got["tri_id"], fx["tri_id"])
    assert same_bits(got["t"], fx["t"])
    assert np.array_equal(got["rgb"], fx["rgb"])
    assert (fx["obj_id"] >= 0).sum() > 100


@pytest.mark.ref
def test_oracle_vs_live_reference_and_selfcheck(ref_scene):
    """Restated loops == the verbatim sendRaysAndIntersectPointsColors at the shipped constants, and the
    oracle == the reference's functions on a frame that is not among the committed ones."""
    b, arrays, light = ref_scene("shipped")
    l4 = np.array([light[0], light[1], light[2], 1.0], f32)
    assert b.selfcheck(60, 40, l4) == 0
    for (W, H, focal, N) in ((75, 50, 50.0, 3), (33, 21, 400.0, 1)):
        p = _abi.default_params(W, H, focal, light, N, True)
        want = b.render(W, H, focal, light, N, True, faithful=True)
        got = pyoracle.oracle_render(arrays, p)
        for k in ("rgb", "obj_id", "tri_id"):
            assert np.array_equal(want[k], got[k]), k
        assert same_bits(want["t"], got["t"])


def test_smooth_normal_kat():
    """interpolateNormal :132-140 at calculateBarycentricCoords :79-117: the oracle against the reference's own two
    functions (tests/golden/leaf_smooth_normal.npz, generated by make_golden.py through oracle/_ref)."""
    lib = pyoracle.oracle_lib()
    g = np.load(GOLDEN / "leaf_smooth_normal.npz")
    p0, e1, e2 = tri_edges(g["tri"])
    n = len(p0)
    got = np.zeros((n, 3), f32)
    for i in range(n):
        out = np.zeros(3, f32)
        lib.srt_oracle_smooth_normal(fp(p0[i]), fp(e1[i]), fp(e2[i]), fp(np.ascontiguousarray(g["normals"][i])),
                                     fp(np.ascontiguousarray(g["point"][i])), fp(out))
        got[i] = out
    assert same_bits(got, g["out"])
    assert np.isfinite(g["out"]).all(axis=1).sum() > 2000


def test_smooth_normal_frames_in_the_oracle():
    """smooth_normals = 1 changes the shading and nothing else."""
    from simple_raytracer_b200 import _abi, host, scenes
    cfg = scenes.CONFIGS["spheres"]
    hs = host.HostScene()
    cfg.build(hs, 0.0)
    arrays = hs.flatten()
    hs.close()
    assert arrays.vnormal is not None and len(arrays.vnormal) == len(arrays.geom)
    assert np.abs(np.linalg.norm(arrays.vnormal["n"][:, :, :3], axis=2) - 1.0).max() < 1e-3     # sphere.obj ships unit normals
    p = _abi.default_params(90, 60, 60.0, scenes.LIGHT_WORLD, 4, True)
    flat = pyoracle.oracle_render(arrays, p)
    p.smooth_normals = 1
    smooth = pyoracle.oracle_render(arrays, p)
    assert np.array_equal(flat["tri_id"], smooth["tri_id"]) and np.array_equal(flat["t"], smooth["t"])
    hit = flat["tri_id"] >= 0
    assert hit.sum() > 300 and (flat["rgb"][hit] != smooth["rgb"][hit]).any(axis=1).mean() > 0.5
    assert np.array_equal(flat["rgb"][~hit], smooth["rgb"][~hit])


def test_texture_sample_modes_of_the_oracle():
    """texture_mode NEAREST / BILINEAR (srt.h; not reference modes): wrapped addressing, exact at texel corners, the mean
    half way, constant on a constant image."""
    import ctypes as C
    from simple_raytracer_b200 import _abi
    lib = pyoracle.oracle_lib()
    rng = np.random.default_rng(3)
    Wt, Ht = 5, 3
    img = rng.integers(0, 256, (Ht, Wt, 3), dtype=np.uint8)

    def sample(x, y, mode, image=img):
        out = np.zeros(3, np.float32)
        lib.srt_oracle_texture_sample(image.ctypes.data_as(C.c_void_p), image.shape[1], image.shape[0], float(x), float(y), mode,
                                      out.ctypes.data_as(C.POINTER(C.c_float)))
        return out

    f = lambda v: (v.astype(np.float32) / np.float32(255.0))
    for x in range(-7, 9):
        for y in range(-5, 7):
            want = f(img[y % Ht, x % Wt])
            assert np.array_equal(sample(x, y, _abi.TEX_NEAREST), want)
            assert np.array_equal(sample(x + 0.75, y + 0.25, _abi.TEX_NEAREST), want)
            assert np.array_equal(sample(x, y, _abi.TEX_BILINEAR), want)           # weights (1, 0): the corner texel itself
            mid = sample(x + 0.5, y, _abi.TEX_BILINEAR)
            assert np.allclose(mid, (want + f(img[y % Ht, (x + 1) % Wt])) / 2, atol=1e-6)
            mid = sample(x, y + 0.5, _abi.TEX_BILINEAR)
            assert np.allclose(mid, (want + f(img[(y + 1) % Ht, x % Wt])) / 2, atol=1e-6)
    flat = np.full((4, 4, 3), 200, np.uint8)
    assert np.allclose(sample(1.3, 2.9, _abi.TEX_BILINEAR, flat), 200 / 255, atol=1e-6)
    for bad in (float("nan"), float("inf"), -float("inf"), 3e30):                  # never out of the image
        assert np.isfinite(sample(bad, 1.0, _abi.TEX_BILINEAR)).all() and np.isfinite(sample(1.0, bad, _abi.TEX_NEAREST)).all()


def test_quality_switches_in_the_oracle():
    """normalise_lights, shadow_segment: what they change and what they leave alone (all off = the golden-checked path)."""
    from simple_raytracer_b200 import _abi, host, scenes
    cfg = scenes.CONFIGS["spheres"]
    hs = host.HostScene()
    cfg.build(hs, 0.0)
    arrays = hs.flatten()
    hs.close()
    W, H, focal = 90, 60, 60.0
    p = _abi.default_params(W, H, focal, scenes.LIGHT_WORLD, 1, True)
    one = pyoracle.oracle_render(arrays, p)
    p.normalise_lights = 1
    assert np.array_equal(pyoracle.oracle_render(arrays, p)["rgb"], one["rgb"])      # a sum of one sample divided by one
    p.light_count, p.light_step = 6, 0.0                                               # six samples at the same place
    assert np.abs(pyoracle.oracle_render(arrays, p)["rgb"].astype(int) - one["rgb"].astype(int)).max() <= 1
    # a light in the middle of the scene: geometry behind it shadows as shipped, not with shadow_segment
    mid = arrays.objects["box_min"].min(axis=0) * 0.5 + arrays.objects["box_max"].max(axis=0) * 0.5
    p = _abi.default_params(W, H, focal, mid, 1, True)
    p.no_sky_fill = 1                                      # a hit that shades to black stays black (:518 turns it into sky)
    shipped = pyoracle.oracle_render(arrays, p)
    p.shadow_segment, p.shadow_eps = 1, 1e-4
    seg = pyoracle.oracle_render(arrays, p)
    assert np.array_equal(shipped["tri_id"], seg["tri_id"]) and np.array_equal(shipped["t"], seg["t"])
    hit = shipped["tri_id"] >= 0
    assert (seg["rgb"][hit].astype(int) >= shipped["rgb"][hit].astype(int)).all()
    assert (seg["rgb"][hit] != shipped["rgb"][hit]).any(axis=1).sum() > 20
