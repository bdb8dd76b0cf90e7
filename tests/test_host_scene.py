"""CPU tests of the host-side scene mirror (csrc/host): same names, argument meaning and error behaviour
as the reference's ObjectManager / Transformation (Object.h:59-89, Transformation.h:10-20), and — where
oracle/_ref exists — byte-identical flattened scenes."""
import numpy as np
import pytest

from oracle import pyoracle
from simple_raytracer_b200 import _abi, capi, host, scenes


def need_assets():
    try:
        host.assets_dir()
    except FileNotFoundError:
        pytest.skip("staged assets not present")


def test_obj_number_parser():
    """tinyobjloader's tryParseDouble rules that decide vertex bits."""
    assert host.parse_obj_number("1.5") == 1.5
    assert host.parse_obj_number("-0.033") == -0.033
    assert host.parse_obj_number("1e3") == 1000.0
    assert host.parse_obj_number("+2.5E-2") == 0.025
    assert host.parse_obj_number(".5") == 0.5
    assert host.parse_obj_number("abc") is None


def test_load_defaults_and_errors():
    need_assets()
    s = host.HostScene()
    n = s.load_obj("cube.obj")
    assert n == 12
    assert s.get_properties("cube.obj") == (np.float32(0.2), np.float32(0.5), np.float32(15.0))   # Object.cpp:29-34
    tris, flags = s.triangles("cube.obj")
    assert tris.shape == (12, 30) and not flags.any()
    assert np.all(tris[:, [3, 7, 11]] == 1.0)                                                     # w = 1
    with pytest.raises(KeyError):
        s.clone("nope.obj", "x.obj", False)                                                       # std::out_of_range, Object.cpp:174
    with pytest.raises(KeyError):
        s.triangles("nope.obj")
    # an object without a bounding hierarchy is what the reference would dereference as a null Node* (:298)
    with pytest.raises(capi.SrtError) as e:
        s.flatten()
    assert e.value.code == _abi.SRT_ERR_SCENE
    s.create_bounding_hierarchy("cube.obj")
    a = s.flatten(with_flat_lists=True)
    assert len(a.geom) == 12 and len(a.flat_geom) == 12 and len(a.nodes) == 3 and a.max_depth == 1
    # missing file: the reference prints and continues with an empty object (Object.cpp:35-39)
    s2 = host.HostScene()
    with pytest.raises(FileNotFoundError):              # ... which this mirror reports loudly unless asked not to
        s2.load_obj("does_not_exist.obj")
    s2.strict = False
    assert s2.load_obj("does_not_exist.obj") <= 0 and "does_not_exist" in s2.warning()


def test_concurrent_loads_do_not_depend_on_the_process_cwd(tmp_path):
    """Regression for the intermittent series mismatch of round 1: load_obj used to chdir() around a call that releases
    the GIL, so prep threads swapped the CWD under each other and a frame silently lost an object.  Eight threads load
    the bunny at once, with the cache cold, from a CWD that holds no assets; every load must see all its triangles."""
    import os
    from concurrent.futures import ThreadPoolExecutor
    need_assets()
    host.set_asset_cache(False)
    old = os.getcwd()
    os.chdir(tmp_path)
    try:
        def load(_i):
            hs = host.HostScene()
            n = (hs.load_obj("./obj/stanford-bunny.obj"), hs.load_obj("./obj/cube.obj"))
            hs.close()
            return n, os.getcwd()
        with ThreadPoolExecutor(8) as pool:
            out = list(pool.map(load, range(24)))
    finally:
        os.chdir(old)
    assert all(n == (69451, 12) for n, _cwd in out)
    assert all(os.path.samefile(cwd, tmp_path) for _n, cwd in out)


def test_flat_layout_invariants():
    need_assets()
    s = host.HostScene()
    view = scenes.build_bunny_ground(s, 0.0)
    a = s.flatten()
    assert a.object_names == ["./obj/stanford-bunny.obj", "./obj/cube.obj"]          # unordered_map iteration order (S2)
    assert len(a.geom) == 69451 + 12
    leaf = (a.nodes["b"] & _abi.SRT_LEAF_BIT) != 0
    cnt = a.nodes["b"][leaf] & 0x7fffffff
    assert cnt.max() <= 8 and cnt.sum() == len(a.geom)                               # leaf size rule, Object.cpp:261
    assert np.array_equal(np.sort(a.nodes["a"][leaf]), np.concatenate([[0], np.cumsum(cnt[np.argsort(a.nodes["a"][leaf])])[:-1]]))
    # every triangle lies inside its leaf's box (boxes are min/max of the vertices, Object.cpp:205-221)
    first = a.nodes["a"][leaf]; lo = a.nodes["bmin"][leaf]; hi = a.nodes["bmax"][leaf]
    for f, c, l, h in list(zip(first, cnt, lo, hi))[::97]:
        g = a.geom[f:f + c]
        pts = np.concatenate([g["p0"], g["p0"] + g["e1"], g["p0"] + g["e2"]])
        assert np.all(pts >= l - 1e-4) and np.all(pts <= h + 1e-4)
    light = scenes.light_view_space(s, view)
    assert np.allclose(light[:3], [-200.00002, 8.493649, -514.7114], atol=1e-3)


@pytest.mark.ref
@pytest.mark.parametrize("key", ["C1", "C3", "shipped", "spheres", "cubes4"])
def test_mirror_flattens_byte_identical_to_reference(key):
    cfg = scenes.CONFIGS[key]
    rb = pyoracle.RefBackend(); rv = cfg.build(rb, 0.0); ra = rb.flatten(with_flat_lists=not cfg.use_bvh)
    hb = host.HostScene(); hv = cfg.build(hb, 0.0); ha = hb.flatten(with_flat_lists=not cfg.use_bvh)
    assert ra.object_names == ha.object_names
    for k in ("nodes", "geom", "attr", "objects"):
        assert np.array_equal(getattr(ra, k).view(np.uint8), getattr(ha, k).view(np.uint8)), k
    if not cfg.use_bvh:
        assert np.array_equal(ra.flat_geom.view(np.uint8), ha.flat_geom.view(np.uint8))
    if rv is not None:
        assert np.array_equal(np.asarray(rv).view(np.uint32), np.asarray(hv).view(np.uint32))


@pytest.mark.ref
def test_mirror_series_frame_and_textured_scene():
    """A rotated series frame (C5) and the textured scene (C4) are byte-identical — including house.obj, whose
    7 faces of more than four vertices go through the ear clipper (srt_polygon.cpp, tiny_obj_loader.h:1618)."""
    cfg = scenes.CONFIGS["C3"]
    rb = pyoracle.RefBackend(); cfg.build(rb, 37.5); ra = rb.flatten()
    hb = host.HostScene(); cfg.build(hb, 37.5); ha = hb.flatten()
    for k in ("nodes", "geom", "attr", "objects"):
        assert np.array_equal(getattr(ra, k).view(np.uint8), getattr(ha, k).view(np.uint8)), k
    cfg = scenes.CONFIGS["C4"]
    rb = pyoracle.RefBackend(); cfg.build(rb, 0.0); ra = rb.flatten()
    hb = host.HostScene(); cfg.build(hb, 0.0); ha = hb.flatten()
    assert ra.object_names == ha.object_names and ra.texture_names == ha.texture_names
    assert all(np.array_equal(a, b) for a, b in zip(ra.textures, ha.textures))       # sidecars == stb_image pixels
    assert np.array_equal(ra.objects.view(np.uint8), ha.objects.view(np.uint8))
    for k in ("nodes", "geom", "attr"):
        assert np.array_equal(getattr(ra, k).view(np.uint8), getattr(ha, k).view(np.uint8)), k


def test_transformation_matrices():
    """Transformation.cpp:6-90: column-major mat4, translation in column 3, view = T*Rz*Ry*Rx."""
    s = host.HostScene()
    T = s.xf_translate(1, 2, 3).reshape(4, 4)
    assert T[3].tolist() == [1, 2, 3, 1]
    S = s.xf_scale(2, 3, 4).reshape(4, 4)
    assert np.diag(S).tolist() == [2, 3, 4, 1]
    R = s.xf_rotate_y(s.radians(90.0)).reshape(4, 4)
    v = s.mat_vec(R.reshape(16), (1, 0, 0, 1))
    assert np.allclose(np.abs(v[:3]), [0, 0, 1], atol=1e-6)
    M = s.xf_view((1, 2, 3), (0.1, 0.2, 0.3))
    I = s.mat_mul(M, s.mat_inverse(M)).reshape(4, 4)
    assert np.allclose(I, np.eye(4), atol=1e-5)
    Mi = s.xf_mirror(1, 0, 0).reshape(4, 4)
    assert np.diag(Mi).tolist() == [-1, 1, 1, 1]


@pytest.mark.ref
def test_write_bmp_equals_reference_draw_image(tmp_path):
    """host.write_bmp of a dense sky-filled frame == the file the reference's drawImage (:461-498, CImg save_bmp) writes
    from the corresponding ImageData lists, byte for byte — at widths with and without row padding."""
    from oracle import pyoracle
    rng = np.random.default_rng(4)
    for W, H in ((37, 21), (64, 48), (150, 5), (1, 1)):
        rgb = rng.integers(0, 256, (H, W, 3), dtype=np.uint8)
        rgb[rng.random((H, W)) < 0.3] = 0                                  # not drawn -> sky in both
        ys, xs = np.nonzero(rgb.any(axis=2))
        pts = np.stack([xs, ys], axis=1).astype(np.float32)
        col = rgb[ys, xs].astype(np.float32)
        want = pyoracle.ref_draw_image(tmp_path, W, H, pts, col, angle=W)
        dense = rgb.copy()
        dense[~rgb.any(axis=2)] = (173, 216, 230)                          # what srt_render returns (sky fill :476-487)
        path = tmp_path / f"ours{W}.bmp"
        host.write_bmp(path, dense)
        assert path.read_bytes() == want, (W, H)


# ---- host side of the device scene prep (prep.py): everything that does not need the GPU -----------------------
def test_object_order_is_the_reference_maps_iteration_order():
    """S2: the object order of a recorded scene = iteration order of objTriangles for the same insertion sequence."""
    from simple_raytracer_b200 import host, scenes
    for key in ("C3", "C4", "spheres", "cubes4"):
        hs = host.HostScene()
        log = []

        class Rec:
            def __getattr__(self, item):
                f = getattr(hs, item)
                if item in ("load_obj", "clone"):
                    def wrapped(*args):
                        log.append(args[1] if item == "clone" else args[0])
                        return f(*args)
                    return wrapped
                return f

        scenes.CONFIGS[key].build(Rec(), 30.0)
        assert host.object_order(log) == hs.object_names(), key
        hs.close()


def test_triangle_texture_slots_and_pixels():
    from simple_raytracer_b200 import host
    hs = host.HostScene()
    hs.load_obj("./obj/house/house.obj")
    slots, names = hs.triangle_textures("./obj/house/house.obj")
    tris, flags = hs.triangles("./obj/house/house.obj")
    assert len(slots) == len(tris) and len(names) >= 2
    assert np.array_equal(slots >= 0, flags != 0)
    assert set(np.unique(slots[slots >= 0])) == set(range(len(names)))
    first_use = [int(np.nonzero(slots == k)[0][0]) for k in range(len(names))]
    assert first_use == sorted(first_use)                       # slots numbered by first appearance in load order
    px = hs.texture(names[0])
    assert px.ndim == 3 and px.shape[2] == 3 and px.dtype == np.uint8
    hs.close()


def test_recorded_scene_script_matches_the_host_calls():
    """prep.DeviceScene records what the host backend executes: same objects, same matrices in the same order, the
    clone idiom copies the transforms made so far, colours / properties follow the reference's defaults."""
    from simple_raytracer_b200 import host, prep, scenes

    class NoGpuCache:                      # MeshCache without a device: meshes are placeholders
        assets = None
        textures = []

        class M:
            n_tris = 0

        def mesh(self, path):
            return self.M()

    ds = prep.DeviceScene(NoGpuCache())
    calls = []
    hs = host.HostScene()

    class Tee:
        def __getattr__(self, item):
            f = getattr(hs, item)
            if item == "transform":
                def wrapped(name, m):
                    calls.append((name, np.asarray(m, np.float32).copy()))
                    return f(name, m)
                return wrapped
            return f

    scenes.CONFIGS["cubes4"].build(ds, 40.0)
    scenes.CONFIGS["cubes4"].build(Tee(), 40.0)
    assert ds.object_names() == hs.object_names()
    # cube1..3 were cloned after the first scale of cube.obj: they inherit exactly that one matrix, then their own
    own = {}
    for name, m in calls:
        own.setdefault(name, []).append(m)
    for name in ("cube1.obj", "cube2.obj", "cube3.obj"):
        got = ds.objects[name].matrices
        want = [own["cube.obj"][0]] + own[name]
        assert len(got) == len(want) and all(np.array_equal(a, b) for a, b in zip(got, want)), name
    assert all(np.array_equal(a, b) for a, b in zip(ds.objects["cube.obj"].matrices, own["cube.obj"]))
    assert ds.props["cube.obj"] == prep.LOAD_PROPERTIES and "cube1.obj" not in ds.props      # clone(..., False), :731
    assert ds.colors["cube1.obj"] == (1.0, 0.0, 1.0)
    with pytest.raises(KeyError):
        ds.clone("nope.obj", "x.obj", False)
    ds.close()
    hs.close()


def _polygon_faces(rng):
    """Faces of 5-120 vertices in random planes (80 and more make the library index its nodes along a z-order curve): convex, star-shaped, combs, with repeated / collinear points, slightly
    non-planar, self-crossing (the ear clipper's fallbacks), in a plane facing x, and collinear (dropped)."""
    faces = []

    def place(xy, wobble=0.0, facing_x=False):
        xy = np.asarray(xy, np.float64)
        if facing_x:
            u, v, o = np.array([0, 1.0, 0]), np.array([0, 0, 1.0]), np.array([2.0, 0, 0])
        else:
            q, _ = np.linalg.qr(rng.normal(size=(3, 3)))
            u, v, o = q[0], q[1], rng.normal(size=3) * 3
        p = o + xy[:, :1] * u + xy[:, 1:] * v
        if wobble:
            p = p + rng.normal(size=p.shape) * wobble
        faces.append(p.astype(np.float32))

    for n in (5, 6, 7, 8, 11, 16, 33, 60, 80, 120):
        ang = np.linspace(0, 2 * np.pi, n, endpoint=False)
        place(np.c_[np.cos(ang), np.sin(ang)] * 2.0)                                     # convex
        place(np.c_[np.cos(-ang), np.sin(-ang)] * 2.0, facing_x=True)                    # other winding, normal along x
        r = np.where(np.arange(n) % 2 == 0, 2.0, 0.7)
        place(np.c_[np.cos(ang) * r, np.sin(ang) * r])                                   # star
        r = rng.uniform(0.3, 2.0, n)
        place(np.c_[np.cos(ang) * r, np.sin(ang) * r], wobble=0.01)                      # star-shaped, off-plane
        place(rng.uniform(-2, 2, (n, 2)))                                                # self-crossing
    for teeth in (2, 3, 7):                                                              # comb: many reflex corners, collinear base
        xs = np.arange(2 * teeth + 1, dtype=np.float64)
        top = np.c_[xs, np.where(np.arange(2 * teeth + 1) % 2 == 0, 2.0, 0.5)]
        base = np.c_[xs[::-1], np.zeros(2 * teeth + 1)]
        place(np.r_[top, base])
    sq = np.array([[0, 0], [1, 0], [2, 0], [2, 0], [2, 2], [1, 2], [0, 2], [0, 1]], np.float64)
    place(sq); place(sq[::-1])                                                           # repeated + collinear points
    place(np.array([[0, 0], [2, 2], [2, 0], [0, 2], [1, 3]], np.float64))                # bow tie with a cap
    place(np.c_[np.arange(6.0), np.arange(6.0) * 0.5])                                   # collinear: zero normal
    place(np.zeros((5, 2)))                                                              # one point five times
    return faces


@pytest.mark.ref
def test_polygon_faces_triangulate_like_the_reference_loader(tmp_path):
    """srt_polygon.cpp against tinyobjloader + mapbox earcut as the reference builds them (simple_raytracer.cpp:15-16):
    same triangles in the same order for every face, bit for bit."""
    rng = np.random.default_rng(20240611)
    for rep in range(4):
        faces = _polygon_faces(rng)
        path = tmp_path / f"polys{rep}.obj"
        with open(path, "w") as f:
            base = 1
            for p in faces:
                for x, y, z in p:
                    f.write(f"v {float(x)!r} {float(y)!r} {float(z)!r}\n")
                f.write("f " + " ".join(str(base + k) for k in range(len(p))) + "\n")
                base += len(p)
        rb = pyoracle.RefBackend(); hb = host.HostScene()
        n_ref = rb.load_obj(str(path)); n_ours = hb.load_obj(str(path))
        assert n_ref == n_ours and n_ours > 0
        rt, rf = rb.triangles(str(path)); ht, hf = hb.triangles(str(path))
        assert rt.shape == ht.shape and len(rt) > 500
        assert np.array_equal(rt.view(np.uint32), ht.view(np.uint32)) and np.array_equal(rf, hf)


def _ref_decode(path):
    import ctypes as C
    lib = pyoracle.ref_lib()
    lib.ref_decode_texture.argtypes = [C.c_char_p, C.c_void_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    w, h = C.c_int(0), C.c_int(0)
    if lib.ref_decode_texture(str(path).encode(), None, 0, C.byref(w), C.byref(h)) != 0:
        return None
    buf = np.zeros((h.value, w.value, 3), np.uint8)
    assert lib.ref_decode_texture(str(path).encode(), buf.ctypes.data, buf.nbytes, C.byref(w), C.byref(h)) == 0
    return buf


@pytest.mark.ref
def test_texture_files_decode_like_the_reference_stb_image():
    """Every texture of the reference's assets (PNG RGB / RGBA; JPEG baseline 4:4:4 and 4:2:0 with restart intervals,
    one progressive) through our own decoders (srt_png.cpp, srt_jpeg.cpp; no sidecar) == stbi_load(..., 3) as the
    reference calls it (Object.cpp:57), bit for bit."""
    files = sorted(list(host.assets_dir().glob("obj/*/*.png")) + list(host.assets_dir().glob("obj/*/*.jpg")))
    assert len(files) >= 10
    for f in files:
        want = _ref_decode(f)
        assert want is not None, f
        got = host.decode_image(f, use_sidecar=False)
        assert got.shape == want.shape and np.array_equal(got, want), f.name


def _write_png(path, samples, depth, ctype, interlace, rng, palette=None):
    """Minimal PNG writer for the decoder test: samples [H, W, C] integers of `depth` bits, random filter per row."""
    import struct, zlib
    H, W, Cn = samples.shape

    def pack_rows(img):                                     # -> list of byte rows
        h, w, _ = img.shape
        flat = img.reshape(h, w * Cn).astype(np.uint32)
        if depth == 16:
            return [np.stack([r >> 8, r & 255], axis=1).astype(np.uint8).reshape(-1) for r in flat]
        if depth == 8:
            return [r.astype(np.uint8) for r in flat]
        per = 8 // depth
        rows = []
        for r in flat:
            pad = (-len(r)) % per
            r = np.concatenate([r, np.zeros(pad, np.uint32)]).reshape(-1, per)
            rows.append(sum(r[:, k] << (8 - depth * (k + 1)) for k in range(per)).astype(np.uint8))
        return rows

    def filtered(img):
        bpp = max(1, Cn * depth // 8)
        out, prev = bytearray(), None
        for row in pack_rows(img):
            cur = row.astype(np.int32)
            up = prev if prev is not None else np.zeros_like(cur)
            left = np.concatenate([np.zeros(bpp, np.int32), cur[:-bpp]]) if len(cur) > bpp else np.zeros_like(cur)
            ul = np.concatenate([np.zeros(bpp, np.int32), up[:-bpp]]) if len(cur) > bpp else np.zeros_like(cur)
            f = int(rng.integers(0, 5))
            p = left + up - ul
            pa, pb, pc = np.abs(p - left), np.abs(p - up), np.abs(p - ul)
            paeth = np.where((pa <= pb) & (pa <= pc), left, np.where(pb <= pc, up, ul))
            pred = [0, left, up, (left + up) >> 1, paeth][f]
            out.append(f)
            out += ((cur - pred) & 255).astype(np.uint8).tobytes()
            prev = cur
        return bytes(out)

    if interlace:
        raw = b""
        for x0, y0, dx, dy in ((0, 0, 8, 8), (4, 0, 8, 8), (0, 4, 4, 8), (2, 0, 4, 4), (0, 2, 2, 4), (1, 0, 2, 2), (0, 1, 1, 2)):
            sub = samples[y0::dy, x0::dx]
            if sub.shape[0] and sub.shape[1]:
                raw += filtered(sub)
    else:
        raw = filtered(samples)

    def chunk(tag, body):
        return struct.pack(">I", len(body)) + tag + body + struct.pack(">I", zlib.crc32(tag + body))

    comp = zlib.compressobj(int(rng.integers(0, 10)))
    z = comp.compress(raw) + comp.flush()
    cut = len(z) // 3
    blob = b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", W, H, depth, ctype, 0, 0, int(interlace)))
    if palette is not None:
        blob += chunk(b"PLTE", palette.astype(np.uint8).tobytes())
    blob += chunk(b"tEXt", b"k\0v") + chunk(b"IDAT", z[:cut]) + chunk(b"IDAT", z[cut:]) + chunk(b"IEND", b"")
    path.write_bytes(blob)


@pytest.mark.ref
def test_generated_png_variants_decode_like_the_reference(tmp_path):
    """Every colour type x bit depth the format allows, with and without Adam7, odd sizes, all five filters, stored /
    fixed / dynamic deflate blocks, IDAT split in two: srt_png.cpp == the reference's stb_image."""
    rng = np.random.default_rng(7)
    kinds = [(0, d) for d in (1, 2, 4, 8, 16)] + [(2, 8), (2, 16)] + [(3, d) for d in (1, 2, 4, 8)] + [(4, 8), (4, 16), (6, 8), (6, 16)]
    n = 0
    for ctype, depth in kinds:
        Cn = {0: 1, 2: 3, 3: 1, 4: 2, 6: 4}[ctype]
        for W, H in ((1, 1), (5, 3), (9, 17), (33, 8), (64, 40)):
            for interlace in (False, True):
                smooth = rng.integers(0, 2)
                img = rng.integers(0, 1 << depth, (H, W, Cn))
                if smooth:                                   # compressible content: dynamic blocks with long matches
                    img = (np.add.outer(np.arange(H), np.arange(W))[:, :, None] // 3 + np.arange(Cn)) % (1 << depth)
                palette = rng.integers(0, 256, (1 << min(depth, 8), 3)) if ctype == 3 else None
                path = tmp_path / f"t{ctype}_{depth}_{W}x{H}_{int(interlace)}.png"
                _write_png(path, img, depth, ctype, interlace, rng, palette)
                want = _ref_decode(path)
                assert want is not None, path.name
                got = host.decode_image(path, use_sidecar=False)
                assert got.shape == want.shape and np.array_equal(got, want), path.name
                n += 1
    assert n == len(kinds) * 10
    bad = tmp_path / "bad.png"
    bad.write_bytes(path.read_bytes()[:60])
    with pytest.raises(ValueError):
        host.decode_image(bad, use_sidecar=False)


@pytest.mark.ref
def test_generated_jpeg_variants_decode_like_the_reference(tmp_path):
    """JPEG files written by Pillow's libjpeg — grey / 4:4:4 / 4:2:2 / 4:2:0 / 4:1:1, baseline and progressive, optimised
    tables, restart intervals, sizes that are not multiples of the MCU, several qualities: srt_jpeg.cpp == stb_image."""
    Image = pytest.importorskip("PIL.Image")
    rng = np.random.default_rng(11)
    n = 0
    for W, H in ((1, 1), (7, 5), (16, 16), (17, 33), (100, 3), (131, 67)):
        yy, xx = np.mgrid[0:H, 0:W]
        base = np.stack([(xx * 255 // max(W - 1, 1)), (yy * 255 // max(H - 1, 1)), ((xx + yy) * 5) % 256], axis=2)
        img = np.clip(base + rng.integers(-40, 40, (H, W, 3)), 0, 255).astype(np.uint8)
        for mode, sub in (("L", None), ("RGB", "4:4:4"), ("RGB", "4:2:2"), ("RGB", "4:2:0"), ("RGB", "4:1:1")):
            for progressive in (False, True):
                for quality, extra in ((30, {}), (75, {"optimize": True}), (95, {"restart_marker_blocks": 3}), (100, {"restart_marker_rows": 1})):
                    im = Image.fromarray(img if mode == "RGB" else img[:, :, 0], mode)
                    path = tmp_path / f"j_{W}x{H}_{mode}_{(sub or 'g').replace(':', '')}_{int(progressive)}_{quality}.jpg"
                    kw = dict(quality=quality, progressive=progressive, **extra)
                    if sub:
                        kw["subsampling"] = sub
                    try:
                        im.save(path, "JPEG", **kw)
                    except (TypeError, ValueError, KeyError):
                        continue
                    want = _ref_decode(path)
                    assert want is not None, path.name
                    got = host.decode_image(path, use_sidecar=False)
                    assert got.shape == want.shape and np.array_equal(got, want), path.name
                    n += 1
    assert n >= 100


def test_damaged_texture_files_fail_cleanly(tmp_path):
    """Truncated and corrupted PNG / JPEG files: decode_image raises or returns an image of the announced size — it never
    crashes, hangs or over-allocates (the decoders were also run over 21,000 such files under ASan / UBSan)."""
    rng = np.random.default_rng(5)
    files = []
    p = tmp_path / "a.png"
    _write_png(p, rng.integers(0, 256, (21, 34, 3)), 8, 2, False, rng)
    files.append(p)
    p = tmp_path / "b.png"
    _write_png(p, rng.integers(0, 4, (19, 31, 1)), 2, 3, True, rng, rng.integers(0, 256, (4, 3)))
    files.append(p)
    try:
        from PIL import Image
        img = rng.integers(0, 256, (37, 53, 3), dtype=np.uint8)
        Image.fromarray(img).save(tmp_path / "c.jpg", quality=80, subsampling="4:2:0")
        Image.fromarray(img).save(tmp_path / "d.jpg", quality=60, progressive=True)
        files += [tmp_path / "c.jpg", tmp_path / "d.jpg"]
    except ImportError:
        pass
    decoded = failed = 0
    for f in files:
        good = f.read_bytes()
        assert host.decode_image(f, use_sidecar=False).ndim == 3
        for it in range(150):
            m = bytearray(good)
            if it % 3 == 0:
                m = m[:int(rng.integers(0, len(m)))]
            elif it % 3 == 1:
                for _ in range(int(rng.integers(1, 8))):
                    m[int(rng.integers(0, len(m)))] ^= 1 << int(rng.integers(0, 8))
            else:
                at = int(rng.integers(0, min(len(m), 400)))
                m[at] = int(rng.integers(0, 256))
            bad = tmp_path / ("m" + f.suffix)
            bad.write_bytes(bytes(m))
            try:
                out = host.decode_image(bad, use_sidecar=False)
                assert out.ndim == 3 and out.shape[2] == 3 and out.size <= 3 << 28
                decoded += 1
            except ValueError:
                failed += 1
    assert failed > 50 and decoded + failed == 150 * len(files)


def test_png_output_round_trips(tmp_path):
    """host.write_png (8-bit RGB, fixed-Huffman deflate with LZ77 matches, per-row filter choice): our own decoder, Pillow and
    — where it is built — the reference's stb_image all return the frame bytes."""
    rng = np.random.default_rng(9)
    frames = [rng.integers(0, 256, (1, 1, 3), dtype=np.uint8), rng.integers(0, 256, (37, 53, 3), dtype=np.uint8),
              np.zeros((64, 300, 3), np.uint8), np.full((5, 7, 3), 255, np.uint8)]
    yy, xx = np.mgrid[0:270, 0:480]
    smooth = np.stack([xx * 255 // 479, yy * 255 // 269, (xx + yy) % 256], axis=2).astype(np.uint8)
    smooth[100:140, 200:300] = (173, 216, 230)              # a flat sky-coloured patch: long matches
    frames.append(smooth)
    for i, f in enumerate(frames):
        path = tmp_path / f"o{i}.png"
        host.write_png(path, f)
        assert np.array_equal(host.decode_image(path, use_sidecar=False), f)
        try:
            from PIL import Image
            assert np.array_equal(np.asarray(Image.open(path).convert("RGB")), f)
        except ImportError:
            pass
        if pyoracle.ref_available():
            assert np.array_equal(_ref_decode(path), f)
    assert (tmp_path / "o4.png").stat().st_size < smooth.nbytes // 3      # it does compress
    host.write_image(tmp_path / "x.bmp", smooth)
    assert (tmp_path / "x.bmp").read_bytes()[:2] == b"BM"
