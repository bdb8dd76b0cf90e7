"""oracle/pyoracle.py — TEST INFRASTRUCTURE, NOT PRODUCT CODE.

ctypes front ends for
  * oracle/liboracle.so       — the plain-C restatement (srt_oracle.c), and
  * oracle/_ref/libsrt_ref.so — the UNMODIFIED reference compiled from /root/reference (ref_shim.cpp).
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
"""
from __future__ import annotations

import contextlib
import ctypes as C
import os
import sys
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
REPO = HERE.parent
if str(REPO) not in sys.path:
    sys.path.insert(0, str(REPO))

from simple_raytracer_b200 import _abi  # noqa: E402

ORACLE_SO = HERE / "liboracle.so"
REF_SO = HERE / "_ref" / "libsrt_ref.so"
ASSETS = HERE.parent / "assets"        # staged reference DATA (tools/stage_assets.py); shared with the product's host mirror

_f32p = C.POINTER(C.c_float)


def _fp(a):
    return a.ctypes.data_as(_f32p)


class OracleCounters(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in (
        "primary_rays", "hit_pixels", "shadow_rays", "node_visits_primary", "tri_tests_primary",
        "node_visits_shadow", "tri_tests_shadow", "lights_evaluated", "texel_fetches", "occluded")]

    def as_dict(self):
        return {n: int(getattr(self, n)) for n, _ in self._fields_}


_oracle = None


def oracle_lib():
    global _oracle
    if _oracle is None:
        if not ORACLE_SO.exists():
            raise FileNotFoundError(f"{ORACLE_SO} missing: run `make -C oracle oracle` (or __graft_entry__.build())")
        lib = C.CDLL(str(ORACLE_SO))
        lib.srt_oracle_render.restype = C.c_long
        lib.srt_oracle_render.argtypes = [C.POINTER(_abi.SceneDesc), C.POINTER(_abi.RenderParams),
                                          C.POINTER(_abi.Frame), C.c_int, C.c_int, C.c_int,
                                          C.POINTER(OracleCounters), C.POINTER(C.c_double)]
        lib.srt_oracle_moller_trumbore.restype = C.c_float
        lib.srt_oracle_moller_trumbore.argtypes = [_f32p] * 5
        lib.srt_oracle_slab.restype = C.c_int
        lib.srt_oracle_slab.argtypes = [_f32p] * 4
        lib.srt_oracle_barycentric.argtypes = [_f32p] * 5
        lib.srt_oracle_phong.argtypes = [_f32p] * 5 + [C.c_float] * 4 + [_f32p]
        lib.srt_oracle_tonemap_quantise.argtypes = [_f32p, C.c_float, C.c_float, _f32p, C.POINTER(C.c_int32)]
        lib.srt_oracle_light_positions.argtypes = [_f32p, C.c_int, C.c_float, _f32p]
        lib.srt_oracle_light_positions_mode.argtypes = [C.POINTER(_abi.RenderParams), _f32p]
        lib.srt_oracle_light_positions_mode.restype = None
        lib.srt_oracle_smooth_normal.argtypes = [_f32p, _f32p, _f32p, _f32p, _f32p, _f32p]
        lib.srt_oracle_smooth_normal.restype = None
        lib.srt_oracle_texture_sample.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_float, C.c_float, C.c_int32, _f32p]
        lib.srt_oracle_texture_sample.restype = None
        _oracle = lib
    return _oracle


def oracle_render(scene: _abi.FlatSceneArrays, params: _abi.RenderParams, *, ids=True, want_t=True,
                  threads=0, col_begin=0, col_step=1):
    """Render with the C restatement.  Returns dict(rgb, obj_id, tri_id, t, counters, seconds, traced)."""
    lib = oracle_lib()
    H, W = params.height, params.width
    rgb = np.zeros((H, W, 3), np.uint8)
    obj = np.full((H, W), -1, np.int32) if ids else None
    tri = np.full((H, W), -1, np.int32) if ids else None
    t = np.full((H, W), np.inf, np.float32) if want_t else None
    fr = _abi.Frame()
    fr.rgb = rgb.ctypes.data
    fr.obj_id = obj.ctypes.data if ids else None
    fr.tri_id = tri.ctypes.data if ids else None
    fr.t = t.ctypes.data if want_t else None
    cnt = OracleCounters()
    sec = C.c_double(0)
    desc = scene.desc()
    n = lib.srt_oracle_render(C.byref(desc), C.byref(params), C.byref(fr), threads, col_begin, col_step,
                              C.byref(cnt), C.byref(sec))
    if n < 0:
        raise RuntimeError("srt_oracle_render failed (bad arguments)")
    return dict(rgb=rgb, obj_id=obj, tri_id=tri, t=t, counters=cnt.as_dict(), seconds=sec.value, traced=int(n))


# -------------------------------------------------------------------------------------------------
class RefRenderArgs(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("focal", C.c_float), ("light_pos", C.c_float * 3),
                ("light_count", C.c_int32), ("use_bvh", C.c_int32), ("faithful", C.c_int32), ("threads", C.c_int32),
                ("col_begin", C.c_int32), ("col_step", C.c_int32), ("sky", C.c_uint8 * 3), ("pad", C.c_uint8)]


_ref = None
_ref_dropin = None
REF_DROPIN_SO = HERE / "_ref" / "libsrt_ref_dropin.so"


def ref_dropin_lib():
    """libsrt_ref_dropin.so: the product's drop-in headers compiled against the reference's types.  The ONLY test library
    that links the product (libsrt_b200.so); loaded by the drop-in tests alone, never by the oracle / CPU-baseline legs."""
    global _ref_dropin
    if _ref_dropin is None:
        if not REF_DROPIN_SO.exists():
            raise FileNotFoundError(f"{REF_DROPIN_SO} missing: needs /root/reference and `make -C oracle ref`")
        lib = C.CDLL(str(REF_DROPIN_SO))
        vp, cp, i32 = C.c_void_p, C.c_char_p, C.c_int
        lib.ref_dropin_render.restype = C.c_long
        lib.ref_dropin_render.argtypes = [vp, i32, i32, _f32p, _f32p, _f32p, C.c_long, C.c_char_p, i32]
        lib.ref_device_prep.restype = C.c_long
        lib.ref_device_prep.argtypes = [vp, i32, C.POINTER(cp), _f32p, i32, i32, _f32p, _f32p, _f32p, C.c_long, vp, vp, vp, vp,
                                        C.c_char_p, i32]
        _ref_dropin = lib
    return _ref_dropin


def ref_available() -> bool:
    return REF_SO.exists() and ASSETS.is_dir()


def ref_lib():
    global _ref
    if _ref is None:
        if not REF_SO.exists():
            raise FileNotFoundError(f"{REF_SO} missing: needs /root/reference and `make -C oracle ref`")
        lib = C.CDLL(str(REF_SO))
        vp, cp, fl, i32 = C.c_void_p, C.c_char_p, C.c_float, C.c_int
        lib.ref_scene_create.restype = vp
        lib.ref_scene_destroy.argtypes = [vp]
        lib.ref_last_error.restype = cp
        lib.ref_last_error.argtypes = [vp]
        lib.ref_load_obj.argtypes = [vp, cp]
        lib.ref_set_color.argtypes = [vp, cp, fl, fl, fl]
        lib.ref_set_properties.argtypes = [vp, cp, fl, fl, fl]
        lib.ref_get_properties.argtypes = [vp, cp, _f32p]
        lib.ref_clone.argtypes = [vp, cp, cp, i32]
        lib.ref_transform.argtypes = [vp, cp, _f32p]
        lib.ref_create_bounding_hierarchy.argtypes = [vp, cp]
        lib.ref_object_count.argtypes = [vp]
        lib.ref_object_name.argtypes = [vp, i32, cp, i32]
        lib.ref_triangle_count.argtypes = [vp, cp]
        lib.ref_get_triangles.argtypes = [vp, cp, _f32p, C.POINTER(C.c_uint8)]
        lib.ref_xf_scale.argtypes = [fl, fl, fl, _f32p]
        for n in ("ref_xf_rotate_x", "ref_xf_rotate_y", "ref_xf_rotate_z"):
            getattr(lib, n).argtypes = [fl, _f32p]
        lib.ref_xf_mirror.argtypes = [i32, i32, i32, _f32p]
        lib.ref_xf_shear.argtypes = [_f32p, _f32p]
        lib.ref_xf_translate.argtypes = [fl, fl, fl, _f32p]
        lib.ref_xf_view.argtypes = [_f32p, _f32p, _f32p]
        lib.ref_mat_inverse.argtypes = [_f32p, _f32p]
        lib.ref_mat_mul.argtypes = [_f32p, _f32p, _f32p]
        lib.ref_mat_vec.argtypes = [_f32p, _f32p, _f32p]
        lib.ref_radians.restype = fl
        lib.ref_radians.argtypes = [fl]
        lib.ref_flatten.argtypes = [vp, i32]
        lib.ref_flat_desc.restype = C.POINTER(_abi.SceneDesc)
        lib.ref_flat_desc.argtypes = [vp]
        lib.ref_flat_max_depth.argtypes = [vp]
        lib.ref_flat_texture_name.argtypes = [vp, i32, cp, i32]
        lib.ref_render.restype = C.c_long
        lib.ref_render.argtypes = [vp, C.POINTER(RefRenderArgs), vp, vp, vp, vp, C.POINTER(C.c_double)]
        lib.ref_render_verbatim.restype = C.c_long
        lib.ref_render_verbatim.argtypes = [vp, i32, i32, _f32p, vp, C.POINTER(C.c_double)]
        lib.ref_selfcheck.argtypes = [vp, i32, i32, _f32p]
        lib.ref_verbatim_sparse.restype = C.c_long
        lib.ref_verbatim_sparse.argtypes = [vp, i32, i32, _f32p, _f32p, _f32p, C.c_long]
        lib.ref_draw_image.argtypes = [i32, i32, _f32p, _f32p, C.c_long, i32]
        lib.ref_leaf_moller_trumbore.argtypes = [i32, _f32p, _f32p, _f32p, _f32p]
        lib.ref_leaf_slab.argtypes = [i32, _f32p, _f32p, _f32p, _f32p, C.POINTER(C.c_uint8)]
        lib.ref_leaf_slab_origin0.argtypes = [i32, _f32p, _f32p, _f32p, C.POINTER(C.c_uint8)]
        lib.ref_leaf_barycentric.argtypes = [i32, _f32p, _f32p, _f32p]
        lib.ref_leaf_smooth_normal.argtypes = [i32, _f32p, _f32p, _f32p, _f32p]
        lib.ref_leaf_normal.argtypes = [i32, _f32p, _f32p]
        lib.ref_leaf_phong.argtypes = [i32] + [_f32p] * 7
        lib.ref_leaf_tonemap.argtypes = [i32, _f32p, _f32p, C.POINTER(C.c_int32)]
        _ref = lib
    return _ref


_libm = C.CDLL("libm.so.6")
_libm.cosf.restype = C.c_float
_libm.cosf.argtypes = [C.c_float]
_libm.sinf.restype = C.c_float
_libm.sinf.argtypes = [C.c_float]


def ref_draw_image(workdir, width, height, points, colors, angle=0) -> bytes:
    """The reference's drawImage(..., saveImage=True) (:461-498) on the given ImageData lists; returns the bytes of the
    BMP it wrote (images/generation/output<angle>.bmp under `workdir`)."""
    import os
    from pathlib import Path as _P
    lib = ref_lib()
    out_dir = _P(workdir) / "images" / "generation"
    out_dir.mkdir(parents=True, exist_ok=True)
    pts = np.ascontiguousarray(points, np.float32)
    col = np.ascontiguousarray(colors, np.float32)
    old = os.getcwd()
    os.chdir(workdir)
    try:
        if lib.ref_draw_image(int(width), int(height), _fp(pts), _fp(col), len(pts), int(angle)) != 0:
            raise RuntimeError("reference drawImage failed")
    finally:
        os.chdir(old)
    return (out_dir / f"output{int(angle)}.bmp").read_bytes()


@contextlib.contextmanager
def in_assets_dir():
    """The reference resolves OBJ and texture paths against the CWD (Object.cpp:55-57)."""
    old = os.getcwd()
    os.chdir(ASSETS)
    try:
        yield
    finally:
        os.chdir(old)


def _m16(m):
    a = np.ascontiguousarray(np.asarray(m, dtype=np.float32).reshape(16))
    return a


class RefBackend:
    """Scene backend over the UNMODIFIED reference (see simple_raytracer_b200/scenes.py for the protocol)."""

    def __init__(self):
        self.lib = ref_lib()
        self.h = self.lib.ref_scene_create()

    def close(self):
        if self.h:
            self.lib.ref_scene_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- scene surface
    def load_obj(self, path):
        with in_assets_dir():
            return self.lib.ref_load_obj(self.h, path.encode())

    def set_color(self, name, rgb):
        self.lib.ref_set_color(self.h, name.encode(), *[float(x) for x in rgb])

    def set_properties(self, name, a, s, sh):
        self.lib.ref_set_properties(self.h, name.encode(), float(a), float(s), float(sh))

    def get_properties(self, name):
        out = np.zeros(3, np.float32)
        if self.lib.ref_get_properties(self.h, name.encode(), _fp(out)) != 0:
            raise KeyError(name)
        return tuple(out)

    def clone(self, src, dst, copy_properties):
        if self.lib.ref_clone(self.h, src.encode(), dst.encode(), 1 if copy_properties else 0) != 0:
            raise KeyError(src)

    def transform(self, name, m16):
        m = _m16(m16)
        self.lib.ref_transform(self.h, name.encode(), _fp(m))

    def create_bounding_hierarchy(self, name):
        self.lib.ref_create_bounding_hierarchy(self.h, name.encode())

    def object_names(self):
        n = self.lib.ref_object_count(self.h)
        out = []
        buf = C.create_string_buffer(1024)
        for i in range(n):
            self.lib.ref_object_name(self.h, i, buf, 1024)
            out.append(buf.value.decode())
        return out

    def triangles(self, name):
        n = self.lib.ref_triangle_count(self.h, name.encode())
        if n < 0:
            raise KeyError(name)
        out = np.zeros((n, 30), np.float32)
        flags = np.zeros(n, np.uint8)
        self.lib.ref_get_triangles(self.h, name.encode(), _fp(out), flags.ctypes.data_as(C.POINTER(C.c_uint8)))
        return out, flags

    # -- matrices
    def _mat(self, fn, *args):
        out = np.zeros(16, np.float32)
        fn(*args, _fp(out))
        return out

    def xf_scale(self, sx, sy, sz):
        return self._mat(self.lib.ref_xf_scale, float(sx), float(sy), float(sz))

    def xf_rotate_x(self, rad):
        return self._mat(self.lib.ref_xf_rotate_x, float(rad))

    def xf_rotate_y(self, rad):
        return self._mat(self.lib.ref_xf_rotate_y, float(rad))

    def xf_rotate_z(self, rad):
        return self._mat(self.lib.ref_xf_rotate_z, float(rad))

    def xf_mirror(self, mx, my, mz):
        return self._mat(self.lib.ref_xf_mirror, int(mx), int(my), int(mz))

    def xf_shear(self, s6):
        s = np.asarray(s6, np.float32)
        return self._mat(self.lib.ref_xf_shear, _fp(s))

    def xf_translate(self, x, y, z):
        return self._mat(self.lib.ref_xf_translate, float(x), float(y), float(z))

    def xf_view(self, pos, rot):
        p = np.asarray(pos, np.float32)
        r = np.asarray(rot, np.float32)
        return self._mat(self.lib.ref_xf_view, _fp(p), _fp(r))

    def mat_inverse(self, m):
        return self._mat(self.lib.ref_mat_inverse, _fp(_m16(m)))

    def mat_mul(self, a, b):
        return self._mat(self.lib.ref_mat_mul, _fp(_m16(a)), _fp(_m16(b)))

    def mat_vec(self, m, v4):
        v = np.asarray(v4, np.float32)
        out = np.zeros(4, np.float32)
        self.lib.ref_mat_vec(_fp(_m16(m)), _fp(v), _fp(out))
        return out

    def radians(self, deg):
        return np.float32(self.lib.ref_radians(float(deg)))

    def cosf(self, x):
        return np.float32(_libm.cosf(float(x)))

    def sinf(self, x):
        return np.float32(_libm.sinf(float(x)))

    # -- flatten through the product flattener applied to the real ObjectManager
    def flatten(self, with_flat_lists=False) -> _abi.FlatSceneArrays:
        rc = self.lib.ref_flatten(self.h, 1 if with_flat_lists else 0)
        if rc != 0:
            raise RuntimeError("flatten failed: " + self.lib.ref_last_error(self.h).decode())
        d = self.lib.ref_flat_desc(self.h)
        tex_names = []
        buf = C.create_string_buffer(1024)
        for i in range(d.contents.n_textures):
            self.lib.ref_flat_texture_name(self.h, i, buf, 1024)
            tex_names.append(buf.value.decode())
        return _abi.FlatSceneArrays.from_desc_pointer(d, self.object_names(), tex_names,
                                                      self.lib.ref_flat_max_depth(self.h))

    # -- render with the reference's own functions
    def render(self, width, height, focal, light_pos, light_count=1, use_bvh=True, faithful=False, threads=0,
               col_begin=0, col_step=1, ids=True):
        a = RefRenderArgs()
        a.width, a.height, a.focal = int(width), int(height), float(focal)
        a.light_pos[:] = [float(x) for x in light_pos[:3]]
        a.light_count = int(light_count)
        a.use_bvh = 1 if use_bvh else 0
        a.faithful = 1 if faithful else 0
        a.threads = threads
        a.col_begin, a.col_step = col_begin, col_step
        a.sky[:] = [173, 216, 230]
        rgb = np.zeros((height, width, 3), np.uint8)
        obj = np.full((height, width), -1, np.int32)
        tri = np.full((height, width), -1, np.int32)
        t = np.full((height, width), np.inf, np.float32)
        sec = C.c_double(0)
        n = self.lib.ref_render(self.h, C.byref(a), rgb.ctypes.data, obj.ctypes.data,
                                tri.ctypes.data if ids else None, t.ctypes.data, C.byref(sec))
        if n < 0:
            raise RuntimeError("ref_render failed: " + self.lib.ref_last_error(self.h).decode())
        return dict(rgb=rgb, obj_id=obj, tri_id=tri if ids else None, t=t, seconds=sec.value, traced=int(n))

    def render_verbatim(self, width, height, light4):
        l4 = np.asarray(light4, np.float32)
        rgb = np.zeros((height, width, 3), np.uint8)
        sec = C.c_double(0)
        n = self.lib.ref_render_verbatim(self.h, width, height, _fp(l4), rgb.ctypes.data, C.byref(sec))
        return dict(rgb=rgb, seconds=sec.value, points=int(n))

    def sparse(self, width, height, light4, dropin):
        """ImageData (points, colours) of the verbatim :505 function, or of include/srt_dropin.hpp's
        srt::Renderer compiled against the reference's own types (needs a GPU)."""
        l4 = np.asarray(light4, np.float32)
        cap = width * height
        pts = np.zeros((cap, 2), np.float32)
        col = np.zeros((cap, 3), np.float32)
        if dropin:
            err = C.create_string_buffer(1024)
            n = ref_dropin_lib().ref_dropin_render(self.h, width, height, _fp(l4), _fp(pts), _fp(col), cap, err, 1024)
            if n < 0:
                raise RuntimeError("srt_dropin: " + err.value.decode(errors="replace"))
        else:
            n = self.lib.ref_verbatim_sparse(self.h, width, height, _fp(l4), _fp(pts), _fp(col), cap)
        return pts[:n].copy(), col[:n].copy()

    def device_prep(self, calls, n_nodes, n_tris, width, height, light4):
        """include/srt_prep.hpp (srt::DevicePrep) + srt::Renderer::renderPrepared compiled against the reference's own
        types, on this scene's meshes AS LOADED: `calls` = [(name, m16), ...] are the transformTriangles calls of one
        frame.  Returns (FlatSceneArrays built on the GPU, ImageData points, colours).  Needs a GPU."""
        l4 = np.asarray(light4, np.float32)
        names = (C.c_char_p * max(len(calls), 1))(*[n.encode() for n, _m in calls])
        mats = np.ascontiguousarray(np.stack([_m16(m) for _n, m in calls]) if calls else np.zeros((0, 16), np.float32))
        n_obj = self.lib.ref_object_count(self.h)
        nodes = np.zeros(n_nodes, _abi.NODE_DTYPE)
        geom = np.zeros(n_tris, _abi.GEOM_DTYPE)
        attr = np.zeros(n_tris, _abi.ATTR_DTYPE)
        objs = np.zeros(n_obj, _abi.OBJECT_DTYPE)
        cap = width * height
        pts = np.zeros((cap, 2), np.float32)
        col = np.zeros((cap, 3), np.float32)
        err = C.create_string_buffer(1024)
        n = ref_dropin_lib().ref_device_prep(self.h, len(calls), names, _fp(mats), width, height, _fp(l4), _fp(pts), _fp(col), cap,
                                     nodes.ctypes.data, geom.ctypes.data, attr.ctypes.data, objs.ctypes.data, err, 1024)
        if n < 0:
            raise RuntimeError("srt_prep: " + err.value.decode(errors="replace"))
        return _abi.FlatSceneArrays(nodes, geom, attr, objs), pts[:n].copy(), col[:n].copy()

    def selfcheck(self, width, height, light4):
        l4 = np.asarray(light4, np.float32)
        return self.lib.ref_selfcheck(self.h, width, height, _fp(l4))
