"""GPU parity on inputs the shipped scenes do not contain: seeded random triangle soups (degenerate and axis-aligned
triangles, flat boxes, tiny / huge scales, camera inside the geometry, many small objects), empty scenes, zero lights,
the maximum light count, one-pixel images — CUDA path through the C ABI against the oracle, ids and t bit-identical,
8-bit image within 1 LSB.  Scenes are written as OBJ files and built by the product's host mirror."""
import numpy as np
import pytest

from conftest import REPO  # noqa: F401  (sys.path)
from oracle import pyoracle
from simple_raytracer_b200 import _abi, capi, host

pytestmark = pytest.mark.gpu
f32 = np.float32


def write_obj(path, verts, faces):
    with open(path, "w") as f:
        for v in verts:
            f.write("v %.9g %.9g %.9g\n" % tuple(float(x) for x in v))
        for a, b, c in faces:
            f.write("f %d %d %d\n" % (a + 1, b + 1, c + 1))


def soup(rng, n_tris, centre, extent, kind):
    """Random triangles around `centre`; `kind` mixes in the nasty cases."""
    centre = np.asarray(centre, np.float64)
    verts, faces = [], []
    for _ in range(n_tris):
        c = centre + rng.uniform(-extent, extent, 3)
        size = extent * rng.choice([0.02, 0.1, 0.5])
        p = c + rng.uniform(-size, size, (3, 3))
        r = rng.random()
        if kind == "nasty":
            if r < 0.10:
                p[2] = p[1]                                    # zero-area triangle (|det| < 1e-12, :57)
            elif r < 0.25:
                p[:, rng.integers(0, 3)] = c[0]                # axis-aligned triangle: a flat leaf box
            elif r < 0.30:
                p = np.round(p)                                # integer coordinates: exact ties, +-0
        base = len(verts)
        verts.extend(p)
        faces.append((base, base + 1, base + 2))
    return verts, faces


def build_scene(tmp_path, rng, n_objects, tris_per_object, kind, scale=1.0):
    hs = host.HostScene(assets=tmp_path)
    for o in range(n_objects):
        name = "obj%d.obj" % o
        centre = np.array([rng.uniform(-30, 30), rng.uniform(-20, 20), rng.uniform(40, 120)]) * scale
        n = int(tris_per_object if np.isscalar(tris_per_object) else tris_per_object[o])
        verts, faces = soup(rng, n, centre, 15.0 * scale, kind)
        write_obj(tmp_path / name, verts, faces)
        assert hs.load_obj(name) >= 0
        hs.set_color(name, tuple(rng.uniform(0.1, 1.0, 3)))
        hs.set_properties(name, float(rng.uniform(0.1, 0.9)), float(rng.uniform(0, 1)), float(rng.choice([2, 8, 32, 100])))
        hs.create_bounding_hierarchy(name)
    return hs


def check(gpu_ctx, arrays, p, variants=(0, 2, 4, 1)):
    want = pyoracle.oracle_render(arrays, p)
    sc = gpu_ctx.upload(arrays)
    try:
        for v in variants:
            p.variant = v
            got = gpu_ctx.render(sc, p)
            assert np.array_equal(got["obj_id"], want["obj_id"]), v
            assert np.array_equal(got["tri_id"], want["tri_id"]), v
            assert np.array_equal(got["t"], want["t"]), v
            assert np.abs(got["rgb"].astype(np.int32) - want["rgb"].astype(np.int32)).max() <= 1, v
    finally:
        p.variant = 0
        sc.destroy()
    return want


@pytest.mark.parametrize("seed,n_objects,tris,kind,scale,lights", [
    (1, 3, 300, "plain", 1.0, 3),
    (2, 2, 800, "nasty", 1.0, 2),
    (3, 8, 40, "nasty", 1.0, 4),            # many small objects: the root loops and the own-object rule
    (4, 2, 500, "nasty", 1e-3, 2),          # tiny scene
    (5, 2, 500, "nasty", 3e3, 2),           # huge coordinates: the filter margin grows with the box size
    (6, 5, (1, 2, 9, 17, 300), "plain", 1.0, 16),   # objects of 1 / 2 triangles (empty and one-triangle leaves), 16 lights
])
def test_random_soups_match_oracle(gpu_ctx, tmp_path, seed, n_objects, tris, kind, scale, lights):
    rng = np.random.default_rng(seed)
    hs = build_scene(tmp_path, rng, n_objects, tris, kind, scale)
    arrays = hs.flatten()
    light = np.array([500.0, -300.0, -200.0], f32) * f32(scale)
    p = _abi.default_params(96, 64, 80.0, light, lights, True)
    p.light_step = float(3.0 * scale)
    want = check(gpu_ctx, arrays, p)
    assert (want["tri_id"] >= 0).mean() > 0.05                  # the camera does see the soup
    hs.close()


def test_camera_inside_the_geometry(gpu_ctx, tmp_path):
    """Boxes that contain the origin (negative entry parameters, S3: no t >= 0 clipping in the slab test)."""
    rng = np.random.default_rng(11)
    hs = host.HostScene(assets=tmp_path)
    for o, centre in enumerate(((0, 0, 0), (5, 3, 20))):
        verts, faces = soup(rng, 400, centre, 25.0, "nasty")
        write_obj(tmp_path / ("in%d.obj" % o), verts, faces)
        hs.load_obj("in%d.obj" % o)
        hs.create_bounding_hierarchy("in%d.obj" % o)
    p = _abi.default_params(80, 60, 40.0, np.array([30, -40, -10], f32), 3, True)
    check(gpu_ctx, hs.flatten(), p)
    hs.close()


def test_zero_lights_and_max_lights_and_one_pixel(gpu_ctx, tmp_path):
    rng = np.random.default_rng(5)
    hs = build_scene(tmp_path, rng, 2, 200, "plain")
    arrays = hs.flatten()
    light = np.array([500.0, -300.0, -200.0], f32)
    p0 = _abi.default_params(64, 48, 60.0, light, 0, True)      # no light: every hit pixel is (0,0,0) -> sky (:518, :476-487)
    w0 = check(gpu_ctx, arrays, p0)
    assert (w0["tri_id"] >= 0).any() and np.all(w0["rgb"] == np.array([173, 216, 230], np.uint8))
    p64 = _abi.default_params(48, 32, 40.0, light, 64, True)    # SRT_MAX_LIGHTS
    check(gpu_ctx, arrays, p64)
    for W, H in ((1, 1), (1, 9), (33, 1)):
        check(gpu_ctx, arrays, _abi.default_params(W, H, 10.0, light, 2, True))
    hs.close()


def test_empty_scene_and_error_codes(gpu_ctx, tmp_path):
    """No object at all: every pixel is sky, nothing is hit.  Invalid requests fail with an error code and a message,
    where the reference would crash or loop (S17)."""
    empty = _abi.FlatSceneArrays(np.zeros(0, _abi.NODE_DTYPE), np.zeros(0, _abi.GEOM_DTYPE), np.zeros(0, _abi.ATTR_DTYPE),
                                 np.zeros(0, _abi.OBJECT_DTYPE), ())
    p = _abi.default_params(40, 30, 30.0, np.array([1, 2, 3], f32), 2, True)
    sc = gpu_ctx.upload(empty)
    got = gpu_ctx.render(sc, p)
    assert np.all(got["tri_id"] == -1) and np.all(got["rgb"] == np.array([173, 216, 230], np.uint8))
    assert got["stats"]["hit_pixels"] == 0
    for field, value in (("width", 0), ("height", -1), ("light_count", 65), ("light_count", -1), ("row_block", 3)):
        q = _abi.default_params(40, 30, 30.0, np.array([1, 2, 3], f32), 2, True)
        setattr(q, field, value)
        buf = np.zeros((30, 40, 3), np.uint8)
        with pytest.raises(capi.SrtError) as e:
            gpu_ctx.render_raw(sc, q, rgb=buf.ctypes.data)
            gpu_ctx.sync()
        assert e.value.code == _abi.SRT_ERR_INVALID_ARG, field
    sc.destroy()
    # a node that points outside the arrays is rejected at upload
    rng = np.random.default_rng(9)
    hs = build_scene(tmp_path, rng, 1, 50, "plain")
    arrays = hs.flatten()
    arrays.nodes["a"][0] = 10 ** 6
    with pytest.raises(capi.SrtError) as e:
        gpu_ctx.upload(arrays)
    assert e.value.code == _abi.SRT_ERR_SCENE
    hs.close()
