"""Device scene prep (srt_scene_prep: transformTriangles + createBoundingHierarchy + flatten on the GPU) against the
host path — this repository's mirror of the reference's ObjectManager (itself checked byte for byte against the
unmodified reference in test_host_scene.py) and, where oracle/_ref is present, the reference itself.

The bar is bit-exactness of every array the render kernels read: nodes (boxes + topology), leaf-ordered geometry,
attributes and object records.  That includes the permutation libstdc++'s std::sort leaves among equal keys
(Object.cpp:241-247), which decides the contents of the leaves.
"""
import numpy as np
import pytest

from conftest import REPO  # noqa: F401  (sys.path)
from simple_raytracer_b200 import _abi, capi, host, prep, scenes

pytestmark = pytest.mark.gpu
f32 = np.float32


def first_diff(a, b):
    a8 = np.frombuffer(a.tobytes(), np.uint8).reshape(len(a), -1)
    b8 = np.frombuffer(b.tobytes(), np.uint8).reshape(len(b), -1)
    rows = np.nonzero((a8 != b8).any(axis=1))[0]
    return len(rows), (int(rows[0]) if len(rows) else -1)


def assert_same_scene(want: _abi.FlatSceneArrays, got: _abi.FlatSceneArrays, want_tex_names=None, got_tex_names=None, what=""):
    assert len(want.nodes) == len(got.nodes) and len(want.geom) == len(got.geom) and len(want.objects) == len(got.objects), what
    # topology first (a, b), then boxes: separates "wrong tree shape" from "wrong sort / bounds"
    assert np.array_equal(want.nodes["a"], got.nodes["a"]) and np.array_equal(want.nodes["b"], got.nodes["b"]), what + ": node topology"
    n, at = first_diff(want.nodes, got.nodes)
    assert n == 0, f"{what}: {n} of {len(want.nodes)} nodes differ, first {at}: want {want.nodes[at]} got {got.nodes[at]}"
    n, at = first_diff(want.objects, got.objects)
    assert n == 0, f"{what}: object record {at} differs: want {want.objects[at]} got {got.objects[at]}"
    g_want, g_got = want.geom.copy(), got.geom.copy()
    if want_tex_names is not None:
        # texture slot numbers are the caller's: compare names
        wn = np.array(list(want_tex_names) + [""], dtype=object)[g_want["tex"]]
        gn = np.array(list(got_tex_names) + [""], dtype=object)[g_got["tex"]]
        assert np.array_equal(wn, gn), what + ": texture of a triangle"
        g_want["tex"] = 0
        g_got["tex"] = 0
    n, at = first_diff(g_want, g_got)
    assert n == 0, f"{what}: {n} of {len(g_want)} triangles differ (leaf order), first {at}: want {g_want[at]} got {g_got[at]}"
    n, at = first_diff(want.attr, got.attr)
    assert n == 0, f"{what}: {n} attribute records differ, first {at}"
    if want.vnormal is not None and got.vnormal is not None:
        n, at = first_diff(want.vnormal, got.vnormal)
        assert n == 0, f"{what}: {n} vertex-normal records differ, first {at}"
    if want.flat_geom is not None:
        assert got.flat_geom is not None and len(got.flat_geom) == len(want.flat_geom), what + ": load-order lists"
        fw, fg = want.flat_geom.copy(), got.flat_geom.copy()
        fw["tex"] = 0
        fg["tex"] = 0
        assert first_diff(fw, fg)[0] == 0 and first_diff(want.flat_attr, got.flat_attr)[0] == 0, what + ": load-order lists"
        assert np.array_equal(want.objects["flat_first"], got.objects["flat_first"])


def both(gpu_ctx, cache, key, angle):
    cfg = scenes.CONFIGS[key]
    hs = host.HostScene(cache.assets)
    view = cfg.build(hs, angle)
    want = hs.flatten(with_flat_lists=not cfg.use_bvh)
    ds = prep.DeviceScene(cache)
    view_d = cfg.build(ds, angle)
    assert (view is None and view_d is None) or np.array_equal(view, view_d)
    assert ds.object_names() == hs.object_names()
    sc = ds.commit(with_flat_lists=not cfg.use_bvh)
    sc.prep_status()
    got = sc.download()
    light = scenes.light_view_space(hs, view) if view is not None else np.array(scenes.LIGHT_WORLD, f32)
    hs.close()
    ds.close()
    return cfg, want, got, sc, np.asarray(light, f32)


@pytest.mark.parametrize("key,angles", [
    ("C3", (0.0, 37.5, 180.0, 301.5)),
    ("C2", (12.0,)),
    ("C1", (0.0,)),
    ("shipped", (0.0,)),
    ("spheres", (0.0,)),
    ("cubes4", (40.0,)),
])
def test_device_prep_equals_host_prep(gpu_ctx, key, angles):
    cache = prep.MeshCache(gpu_ctx)
    try:
        for angle in angles:
            cfg, want, got, sc, light = both(gpu_ctx, cache, key, angle)
            assert_same_scene(want, got, what=f"{key} @ {angle}")
            assert got.max_depth == want.max_depth
            sc.destroy()
    finally:
        cache.close()


def test_device_prepared_scene_renders_the_same_frame(gpu_ctx):
    cache = prep.MeshCache(gpu_ctx)
    try:
        cfg, want, got, sc, light = both(gpu_ctx, cache, "C3", 75.0)
        W, H, focal = scenes.scaled(cfg, 270)
        p = _abi.default_params(W, H, focal, light, cfg.light_count, True)
        up = gpu_ctx.upload(want)
        for variant in (0, 2, 1):
            p.variant = variant
            a = gpu_ctx.render(up, p)
            b = gpu_ctx.render(sc, p)
            for plane in ("rgb", "obj_id", "tri_id", "t"):
                assert np.array_equal(a[plane], b[plane]), (variant, plane)
            assert a["stats"]["hit_pixels"] == b["stats"]["hit_pixels"] > 0
        up.destroy()
        sc.destroy()
    finally:
        cache.close()


def test_device_prepared_scene_renders_smooth_normals(gpu_ctx):
    """Vertex normals travel with the mesh and are re-ordered with the leaves: a device-prepared scene renders the
    smooth-normal mode like the uploaded one."""
    cache = prep.MeshCache(gpu_ctx)
    try:
        cfg, want, got, sc, light = both(gpu_ctx, cache, "spheres", 0.0)
        assert want.vnormal is not None and got.vnormal is not None
        p = _abi.default_params(180, 120, 120.0, light, 4, True)
        p.smooth_normals = 1
        up = gpu_ctx.upload(want)
        a = gpu_ctx.render(up, p)
        b = gpu_ctx.render(sc, p)
        for plane in ("rgb", "obj_id", "tri_id", "t"):
            assert np.array_equal(a[plane], b[plane]), plane
        assert a["stats"]["hit_pixels"] > 0
        up.destroy()
        sc.destroy()
    finally:
        cache.close()


def test_device_prep_reuses_the_scene_across_frames(gpu_ctx):
    """The image series: one device scene, re-prepared in place for every frame (asynchronous; nothing but the
    matrices crosses the bus)."""
    cache = prep.MeshCache(gpu_ctx)
    cfg = scenes.CONFIGS["C3"]
    sc = None
    try:
        for f in (0, 7, 100, 239):
            angle = scenes.SERIES_DEG_PER_FRAME * f
            hs = host.HostScene()
            cfg.build(hs, angle)
            want = hs.flatten()
            hs.close()
            ds = prep.DeviceScene(cache)
            cfg.build(ds, angle)
            sc = ds.commit(scene=sc)
            ds.close()
            sc.prep_status()
            assert_same_scene(want, sc.download(), what=f"frame {f}")
    finally:
        if sc is not None:
            sc.destroy()
        cache.close()


@pytest.mark.parametrize("angle", [0.0, 141.0])
def test_device_prep_textured_multi_object(gpu_ctx, angle):
    """C4 layout (four textured meshes, house.obj with its ear-clipped faces) through the product's own OBJ reader."""
    cache = prep.MeshCache(gpu_ctx)
    try:
        cfg = scenes.CONFIGS["C4"]
        hs = host.HostScene()
        cfg.build(hs, angle)
        want = hs.flatten()
        ds = prep.DeviceScene(cache)
        cfg.build(ds, angle)
        assert ds.object_names() == hs.object_names()
        sc = ds.commit()
        sc.prep_status()
        got = sc.download()
        assert_same_scene(want, got, want.texture_names, cache.texture_names, what=f"C4 @ {angle}")
        assert sorted(want.texture_names) == sorted(cache.texture_names)
        hs.close()
        ds.close()
        sc.destroy()
    finally:
        cache.close()


def test_series_driver_with_device_prep_equals_host_prep(gpu_ctx):
    """The image-series driver end to end: frames prepared on the GPU are byte-identical to frames prepared on the host."""
    from simple_raytracer_b200 import series
    cfg = scenes.CONFIGS["C3"]
    W, H, focal = scenes.scaled(cfg, 180)
    angles = [0.0, 1.5, 91.5, 200.0, 358.5]
    frames = {}
    for mode in ("host", "device"):
        sr = series.SeriesRenderer(gpu_ctx, cfg, W, H, focal, 4, workers=2, prep=mode)
        frames[mode] = [rgb.copy() for (_i, rgb, _tm) in sr.frames(angles)]
        assert sr.host_fallbacks == 0
        sr.close()
    for a, b in zip(frames["host"], frames["device"]):
        assert np.array_equal(a, b)
    assert len({f.tobytes() for f in frames["device"]}) == len(angles)      # and they are five different frames


def test_device_prep_on_a_context_bound_to_the_default_stream():
    """The legacy default stream cannot be captured into a CUDA graph: such a context launches kernel by kernel."""
    ctx = capi.Context(0, stream=0)
    cache = prep.MeshCache(ctx)
    try:
        for angle in (10.0, 20.0, 30.0):
            cfg, want, got, sc, light = both(ctx, cache, "C3", angle)
            assert_same_scene(want, got, what=f"default stream @ {angle}")
            sc.destroy()
    finally:
        cache.close()
        ctx.destroy()


def write_obj(path, tris):
    with open(path, "w") as f:
        for t in tris.reshape(-1, 3):
            f.write("v %.9g %.9g %.9g\n" % tuple(float(x) for x in t))
        for i in range(len(tris)):
            f.write("f %d %d %d\n" % (3 * i + 1, 3 * i + 2, 3 * i + 3))


def key_pattern(rng, n, kind):
    """x of every triangle's first vertex: the sort key of the root (x is made the longest axis)."""
    i = np.arange(n, dtype=np.float64)
    if kind == "random":
        return rng.uniform(-1000, 1000, n)
    if kind == "few":                       # eight distinct keys: long runs of ties
        return rng.integers(0, 8, n).astype(np.float64) * 100.0
    if kind == "ascending":
        return i
    if kind == "descending_zeros":          # exhausts introsort's depth limit: the heap-sort fallback
        return (n - i) * (rng.integers(0, 4, n) != 0)
    if kind == "shared":                    # a mesh: every first vertex shared by a few triangles
        return np.repeat(rng.uniform(-1000, 1000, n // 5 + 1), 5)[:n]
    if kind == "zeros":                     # all keys equal, +0 and -0 mixed: box corners with a signed zero
        return np.where(rng.integers(0, 2, n) == 0, 0.0, -0.0)
    raise ValueError(kind)


@pytest.mark.parametrize("kind,n", [
    ("random", 5000), ("few", 3000), ("ascending", 9001), ("descending_zeros", 12000), ("shared", 20000), ("zeros", 700),
    ("random", 17), ("random", 16), ("random", 9), ("random", 8), ("random", 3), ("random", 1), ("descending_zeros", 70000),
    ("few", 70000), ("ascending", 50000), ("shared", 150000), ("random", 12289), ("random", 4097), ("shared", 400000),
])
def test_sort_permutation_on_key_patterns(gpu_ctx, tmp_path, kind, n):
    """One object whose root sorts on a prescribed key sequence; no transform, so keys are exactly these."""
    rng = np.random.default_rng(n * 31 + len(kind))
    tris = rng.uniform(-5.0, 5.0, (n, 3, 3))
    if kind == "zeros":
        tris[:, :, 0] = np.where(rng.integers(0, 2, (n, 3)) == 0, 0.0, -0.0)
        tris[:, :, 1] *= 100.0
    else:
        tris[:, :, 0] = 0.0
    tris[:, 0, 0] = key_pattern(rng, n, kind)
    tris[:, 1, 0] = tris[:, 0, 0]
    tris[:, 2, 0] = tris[:, 0, 0]
    write_obj(tmp_path / "soup.obj", tris)
    cache = prep.MeshCache(gpu_ctx, assets=tmp_path)
    try:
        hs = host.HostScene(assets=tmp_path)
        hs.load_obj("soup.obj")
        hs.create_bounding_hierarchy("soup.obj")
        want = hs.flatten(with_flat_lists=True)
        ds = prep.DeviceScene(cache)
        ds.load_obj("soup.obj")
        ds.create_bounding_hierarchy("soup.obj")
        sc = ds.commit(with_flat_lists=True)
        sc.prep_status()
        assert_same_scene(want, sc.download(), what=f"{kind} n={n}")
        hs.close()
        ds.close()
        sc.destroy()
    finally:
        cache.close()


def test_empty_scene_and_non_finite_coordinates(gpu_ctx, tmp_path):
    cache = prep.MeshCache(gpu_ctx, assets=tmp_path)
    try:
        ds = prep.DeviceScene(cache)
        sc = ds.commit()
        sc.prep_status()
        assert sc.counts()["n_objects"] == 0
        p = _abi.default_params(64, 32, 400.0, (0, 0, 0), 1, True)
        out = gpu_ctx.render(sc, p)
        assert (out["rgb"] == np.array([173, 216, 230], np.uint8)).all()
        sc.destroy()
        ds.close()
        # a transform that overflows: reported, not silently mis-sorted
        rng = np.random.default_rng(5)
        write_obj(tmp_path / "s.obj", rng.uniform(1.0, 2.0, (100, 3, 3)))
        ds = prep.DeviceScene(cache)
        ds.load_obj("s.obj")
        ds.transform("s.obj", ds.xf_scale(3e38, 3e38, 3e38))
        ds.transform("s.obj", ds.xf_scale(10.0, 10.0, 10.0))
        ds.create_bounding_hierarchy("s.obj")
        sc = ds.commit()
        with pytest.raises(capi.SrtError) as e:
            sc.prep_status()
        assert e.value.code == _abi.SRT_ERR_UNSUPPORTED
        sc.destroy()
        ds.close()
    finally:
        cache.close()


@pytest.mark.ref
def test_device_prep_equals_the_unmodified_reference(gpu_ctx, ref_scene):
    """Straight against oracle/_ref: the reference's own ObjectManager, transformTriangles and createBoundingHierarchy."""
    cache = prep.MeshCache(gpu_ctx)
    try:
        for key, angle in (("C3", 0.0), ("C3", 90.0)):
            _b, want, _light = ref_scene(key, angle)
            ds = prep.DeviceScene(cache)
            scenes.CONFIGS[key].build(ds, angle)
            sc = ds.commit()
            sc.prep_status()
            assert_same_scene(want, sc.download(), what=f"reference {key} @ {angle}")
            sc.destroy()
            ds.close()
    finally:
        cache.close()


class _Split:
    """Scene backend that sends the load / colour / clone calls to `loaded` and `full`, the per-frame calls
    (transformTriangles, createBoundingHierarchy) to `full` only, and records the transforms."""

    def __init__(self, loaded, full):
        self.loaded, self.full, self.calls = loaded, full, []

    def transform(self, name, m16):
        self.calls.append((name, np.asarray(m16, f32).copy()))
        self.full.transform(name, m16)

    def create_bounding_hierarchy(self, name):
        self.full.create_bounding_hierarchy(name)

    def __getattr__(self, item):
        fa, fb = getattr(self.loaded, item), getattr(self.full, item)
        if item in ("load_obj", "set_color", "set_properties", "clone"):
            def both_(*args):
                fa(*args)
                return fb(*args)
            return both_
        return fb


@pytest.mark.ref
def test_cpp_device_prep_header_against_the_reference_types(gpu_ctx):
    """include/srt_prep.hpp compiled against the reference's ObjectManager: the scene it builds on the GPU equals the
    flattened result of the reference's own transformTriangles + createBoundingHierarchy, and renderPrepared returns
    the ImageData of the verbatim sendRaysAndIntersectPointsColors (:505)."""
    from oracle import pyoracle
    cfg = scenes.CONFIGS["shipped"]
    loaded, full = pyoracle.RefBackend(), pyoracle.RefBackend()
    b = _Split(loaded, full)
    view = cfg.build(b, 20.0)
    light = scenes.light_view_space(full, view)
    want = full.flatten()
    got, pts, col = loaded.device_prep(b.calls, len(want.nodes), len(want.geom), cfg.width, cfg.height, light)
    assert_same_scene(want, got, what="srt::DevicePrep on the reference's types")
    wp, wc = full.sparse(cfg.width, cfg.height, light, dropin=False)
    assert np.array_equal(wp, pts)
    assert np.abs(wc - col).max() <= 1.0
    loaded.close()
    full.close()


def test_device_prep_is_repeatable_under_load(gpu_ctx):
    """Race hunt without a sanitizer: the same frames prepared 60 times each, alternating between two compositions (so the
    cached launch graphs are swapped every time) while a second context renders on another stream — every download is
    byte-identical to the host build.  A missing fence or barrier in the multi-CTA sort shows up here as a rare mismatch."""
    cache = prep.MeshCache(gpu_ctx)
    other = capi.Context(0)
    cases = []
    try:
        for key, angle in (("C3", 33.0), ("C4", 141.0)):
            cfg = scenes.CONFIGS[key]
            hs = host.HostScene()
            view = cfg.build(hs, angle)
            want = hs.flatten()
            light = scenes.light_view_space(hs, view)
            hs.close()
            cases.append((cfg, angle, want, light))
        busy = other.upload(cases[0][2])
        pb = _abi.default_params(480, 270, 270.0, cases[0][3], 16, True)
        sc = {0: None, 1: None}
        for rep in range(60):
            for i, (cfg, angle, want, _light) in enumerate(cases):
                if rep % 3 == 0:
                    other.render(busy, pb, ids=False, want_t=False)
                ds = prep.DeviceScene(cache)
                cfg.build(ds, angle)
                sc[i] = ds.commit(scene=sc[i])
                ds.close()
                sc[i].prep_status()
                assert_same_scene(want, sc[i].download(), want.texture_names, cache.texture_names, what=f"repeat {rep}, scene {i}")
        busy.destroy()
        for s in sc.values():
            s.destroy()
    finally:
        other.destroy()
        cache.close()
