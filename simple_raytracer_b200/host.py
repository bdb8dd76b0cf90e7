"""Python plumbing over the host-side scene surface (include/srt_host.h, csrc/host/*).

`HostScene` is this repository's ObjectManager + Transformation (C++), exposed call for call the
way the reference's main() uses them.  It implements the backend protocol of `scenes.py`.
No arithmetic happens in Python: matrices, transforms, BVH build and flattening are C++.
"""
from __future__ import annotations

import contextlib
import ctypes as C
import os
from pathlib import Path
from typing import Optional

import numpy as np

from . import _abi, capi

_f32p = C.POINTER(C.c_float)

# Mesh / texture data staged by __graft_entry__.build() from the reference checkout (data, not
# source; git-ignored).  Override with SRT_ASSETS.
DEFAULT_ASSETS = Path(__file__).resolve().parent.parent / "assets"


def assets_dir() -> Path:
    p = Path(os.environ.get("SRT_ASSETS", str(DEFAULT_ASSETS)))
    if not p.is_dir():
        raise FileNotFoundError(
            f"asset directory {p} not found: run __graft_entry__.build() where /root/reference exists, "
            "or set SRT_ASSETS to a directory holding cube.obj, sphere.obj and obj/")
    return p


_bound = False


def _lib():
    global _bound
    L = capi.lib()
    if not _bound:
        vp, cp, fl, i32 = C.c_void_p, C.c_char_p, C.c_float, C.c_int
        L.srt_host_scene_create.restype = vp
        L.srt_host_scene_destroy.argtypes = [vp]
        L.srt_host_scene_destroy.restype = None
        L.srt_host_last_error.restype = cp
        L.srt_host_last_error.argtypes = [vp]
        L.srt_host_last_warning.restype = cp
        L.srt_host_last_warning.argtypes = [vp]
        L.srt_host_scene_set_base.argtypes = [vp, cp]
        L.srt_host_scene_set_base.restype = None
        L.srt_host_load_obj.argtypes = [vp, cp]
        L.srt_host_set_color.argtypes = [vp, cp, fl, fl, fl]
        L.srt_host_set_color.restype = None
        L.srt_host_set_properties.argtypes = [vp, cp, fl, fl, fl]
        L.srt_host_set_properties.restype = None
        L.srt_host_get_properties.argtypes = [vp, cp, _f32p]
        L.srt_host_clone.argtypes = [vp, cp, cp, i32]
        L.srt_host_transform.argtypes = [vp, cp, _f32p]
        L.srt_host_transform.restype = None
        L.srt_host_create_bounding_hierarchy.argtypes = [vp, cp]
        L.srt_host_create_bounding_hierarchy.restype = None
        L.srt_host_object_count.argtypes = [vp]
        L.srt_host_object_name.argtypes = [vp, i32, cp, i32]
        L.srt_host_triangle_count.argtypes = [vp, cp]
        L.srt_host_get_triangles.argtypes = [vp, cp, _f32p, C.POINTER(C.c_uint8)]
        L.srt_host_set_texture.argtypes = [vp, cp, vp, i32, i32]
        L.srt_host_xf_scale.argtypes = [fl, fl, fl, _f32p]
        for n in ("srt_host_xf_rotate_x", "srt_host_xf_rotate_y", "srt_host_xf_rotate_z"):
            getattr(L, n).argtypes = [fl, _f32p]
        L.srt_host_xf_mirror.argtypes = [i32, i32, i32, _f32p]
        L.srt_host_xf_shear.argtypes = [_f32p, _f32p]
        L.srt_host_xf_translate.argtypes = [fl, fl, fl, _f32p]
        L.srt_host_xf_view.argtypes = [_f32p, _f32p, _f32p]
        L.srt_host_mat_inverse.argtypes = [_f32p, _f32p]
        L.srt_host_mat_mul.argtypes = [_f32p, _f32p, _f32p]
        L.srt_host_mat_vec.argtypes = [_f32p, _f32p, _f32p]
        L.srt_host_radians.restype = fl
        L.srt_host_radians.argtypes = [fl]
        L.srt_host_flatten.argtypes = [vp, i32]
        L.srt_host_flat_desc.restype = C.POINTER(_abi.SceneDesc)
        L.srt_host_flat_desc.argtypes = [vp]
        L.srt_host_flat_max_depth.argtypes = [vp]
        L.srt_host_flat_texture_name.argtypes = [vp, i32, cp, i32]
        L.srt_host_parse_double.argtypes = [cp, C.POINTER(C.c_double)]
        L.srt_host_write_bmp.argtypes = [cp, C.c_void_p, i32, i32]
        L.srt_host_write_png.argtypes = [cp, C.c_void_p, i32, i32]
        L.srt_host_set_asset_cache.argtypes = [i32]
        L.srt_host_set_asset_cache.restype = None
        L.srt_host_triangle_textures.argtypes = [vp, cp, C.POINTER(C.c_int32), C.c_char_p, i32]
        L.srt_host_texture_data.argtypes = [vp, cp, C.POINTER(vp), C.POINTER(i32), C.POINTER(i32)]
        L.srt_host_decode_image.argtypes = [cp, i32, vp, C.c_size_t, C.POINTER(i32), C.POINTER(i32)]
        L.srt_host_object_order.argtypes = [C.POINTER(cp), i32, C.POINTER(i32)]
        _bound = True
    return L


_libm = C.CDLL("libm.so.6")
_libm.cosf.restype = C.c_float
_libm.cosf.argtypes = [C.c_float]
_libm.sinf.restype = C.c_float
_libm.sinf.argtypes = [C.c_float]


def _fp(a):
    return a.ctypes.data_as(_f32p)


def _m16(m):
    return np.ascontiguousarray(np.asarray(m, dtype=np.float32).reshape(16))


class HostScene:
    """One srt::ObjectManager.  OBJ and texture paths resolve against `assets` (the reference resolves
    them against the process CWD, Object.cpp:55-57; scene scripts use the reference's relative names)."""

    def __init__(self, assets: Optional[Path] = None):
        self.lib = _lib()
        self.h = self.lib.srt_host_scene_create()
        self.assets = Path(assets) if assets is not None else None
        self.strict = True       # False: loader diagnostics only go to warning(), like the reference's std::cerr

    def close(self):
        if getattr(self, "h", None):
            self.lib.srt_host_scene_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- ObjectManager
    def load_obj(self, path: str) -> int:
        # Paths resolve against the asset directory inside the C++ scene.  (This used to chdir() around the call: the CWD is
        # process-wide and the call releases the GIL, so two prep threads of the series driver could swap it under each
        # other — an OBJ then failed to open, the loader "printed and continued" like the reference (Object.cpp:35-39), and
        # the frame silently lacked that object.  DESIGN.md §4, "the intermittent series mismatch".)
        base = self.assets if self.assets is not None else assets_dir()
        self.lib.srt_host_scene_set_base(self.h, str(Path(base).resolve()).encode())
        n = self.lib.srt_host_load_obj(self.h, path.encode())
        w = self.warning()
        if w and self.strict:
            # the reference prints this to std::cerr and renders on without the object (Object.cpp:35-39, :63-65)
            raise FileNotFoundError(f"loadObjFile({path!r}) under {base}: {w.strip()}")
        return n

    def warning(self) -> str:
        return self.lib.srt_host_last_warning(self.h).decode(errors="replace")

    def set_color(self, name, rgb):
        self.lib.srt_host_set_color(self.h, name.encode(), *[float(x) for x in rgb])

    def set_properties(self, name, a, s, sh):
        self.lib.srt_host_set_properties(self.h, name.encode(), float(a), float(s), float(sh))

    def get_properties(self, name):
        out = np.zeros(3, np.float32)
        if self.lib.srt_host_get_properties(self.h, name.encode(), _fp(out)) != 0:
            raise KeyError(name)
        return tuple(out)

    def clone(self, src, dst, copy_properties):
        if self.lib.srt_host_clone(self.h, src.encode(), dst.encode(), 1 if copy_properties else 0) != 0:
            raise KeyError(src)      # std::out_of_range from getTriangles (Object.cpp:174)

    def transform(self, name, m16):
        self.lib.srt_host_transform(self.h, name.encode(), _fp(_m16(m16)))

    def create_bounding_hierarchy(self, name):
        self.lib.srt_host_create_bounding_hierarchy(self.h, name.encode())

    def set_texture(self, name, rgb):
        a = np.ascontiguousarray(rgb, dtype=np.uint8)
        if self.lib.srt_host_set_texture(self.h, name.encode(), a.ctypes.data, a.shape[1], a.shape[0]) != 0:
            raise ValueError("bad texture")

    def object_names(self):
        out, buf = [], C.create_string_buffer(1024)
        for i in range(self.lib.srt_host_object_count(self.h)):
            self.lib.srt_host_object_name(self.h, i, buf, 1024)
            out.append(buf.value.decode())
        return out

    def triangles(self, name):
        n = self.lib.srt_host_triangle_count(self.h, name.encode())
        if n < 0:
            raise KeyError(name)
        out = np.zeros((n, 30), np.float32)
        flags = np.zeros(n, np.uint8)
        self.lib.srt_host_get_triangles(self.h, name.encode(), _fp(out), flags.ctypes.data_as(C.POINTER(C.c_uint8)))
        return out, flags

    def triangle_textures(self, name):
        """(slot per triangle [n] int32, slot names) — slots numbered by first appearance in load order, -1 = untextured."""
        n = self.lib.srt_host_triangle_count(self.h, name.encode())
        if n < 0:
            raise KeyError(name)
        slots = np.full(n, -1, np.int32)
        buf = C.create_string_buffer(1 << 16)
        k = self.lib.srt_host_triangle_textures(self.h, name.encode(), slots.ctypes.data_as(C.POINTER(C.c_int32)), buf, len(buf))
        if k < 0:
            raise RuntimeError("srt_host_triangle_textures failed")
        names = [x for x in buf.value.decode().split("\n") if x]
        assert len(names) == k
        return slots, names

    def texture(self, texture_name) -> np.ndarray:
        """Decoded RGB8 pixels [H][W][3] of a loaded texture (a copy)."""
        ptr, w, h = C.c_void_p(), C.c_int(), C.c_int()
        if self.lib.srt_host_texture_data(self.h, texture_name.encode(), C.byref(ptr), C.byref(w), C.byref(h)) != 0:
            raise KeyError(texture_name)
        buf = (C.c_uint8 * (w.value * h.value * 3)).from_address(ptr.value)
        return np.frombuffer(buf, np.uint8).reshape(h.value, w.value, 3).copy()

    # -- Transformation / glm
    def _mat(self, fn, *args):
        out = np.zeros(16, np.float32)
        fn(*args, _fp(out))
        return out

    def xf_scale(self, sx, sy, sz):
        return self._mat(self.lib.srt_host_xf_scale, float(sx), float(sy), float(sz))

    def xf_rotate_x(self, rad):
        return self._mat(self.lib.srt_host_xf_rotate_x, float(rad))

    def xf_rotate_y(self, rad):
        return self._mat(self.lib.srt_host_xf_rotate_y, float(rad))

    def xf_rotate_z(self, rad):
        return self._mat(self.lib.srt_host_xf_rotate_z, float(rad))

    def xf_mirror(self, mx, my, mz):
        return self._mat(self.lib.srt_host_xf_mirror, int(mx), int(my), int(mz))

    def xf_shear(self, s6):
        return self._mat(self.lib.srt_host_xf_shear, _fp(np.asarray(s6, np.float32)))

    def xf_translate(self, x, y, z):
        return self._mat(self.lib.srt_host_xf_translate, float(x), float(y), float(z))

    def xf_view(self, pos, rot):
        return self._mat(self.lib.srt_host_xf_view, _fp(np.asarray(pos, np.float32)), _fp(np.asarray(rot, np.float32)))

    def mat_inverse(self, m):
        return self._mat(self.lib.srt_host_mat_inverse, _fp(_m16(m)))

    def mat_mul(self, a, b):
        return self._mat(self.lib.srt_host_mat_mul, _fp(_m16(a)), _fp(_m16(b)))

    def mat_vec(self, m, v4):
        out = np.zeros(4, np.float32)
        self.lib.srt_host_mat_vec(_fp(_m16(m)), _fp(np.asarray(v4, np.float32)), _fp(out))
        return out

    def radians(self, deg):
        return np.float32(self.lib.srt_host_radians(float(deg)))

    def cosf(self, x):
        return np.float32(_libm.cosf(float(x)))

    def sinf(self, x):
        return np.float32(_libm.sinf(float(x)))

    # -- flatten
    def flatten_desc(self, with_flat_lists=False):
        """Flatten in C++ and return a pointer to the C-owned srt_scene_desc (zero-copy; valid until
        the scene is changed, flattened again or closed)."""
        rc = self.lib.srt_host_flatten(self.h, 1 if with_flat_lists else 0)
        if rc != 0:
            raise capi.SrtError(rc, self.lib.srt_host_last_error(self.h).decode(errors="replace"))
        return self.lib.srt_host_flat_desc(self.h)

    def flatten(self, with_flat_lists=False) -> _abi.FlatSceneArrays:
        d = self.flatten_desc(with_flat_lists)
        names, buf = [], C.create_string_buffer(1024)
        for i in range(d.contents.n_textures):
            self.lib.srt_host_flat_texture_name(self.h, i, buf, 1024)
            names.append(buf.value.decode())
        return _abi.FlatSceneArrays.from_desc_pointer(d, self.object_names(), names,
                                                      self.lib.srt_host_flat_max_depth(self.h))


def object_order(names):
    """Iteration order of the reference's objTriangles map for keys first touched in the order `names` (S2)."""
    uniq = list(dict.fromkeys(names))
    arr = (C.c_char_p * max(len(uniq), 1))(*[n.encode() for n in uniq])
    out = (C.c_int * max(len(uniq), 1))()
    k = _lib().srt_host_object_order(arr, len(uniq), out)
    return [uniq[out[i]] for i in range(k)]


def write_bmp(path, rgb: np.ndarray):
    """drawImage(..., saveImage=True) (simple_raytracer.cpp:461-498): the BMP file CImg writes, from a dense sky-filled
    RGB8 frame [H][W][3] — byte-identical to the reference's file."""
    a = np.ascontiguousarray(rgb, np.uint8)
    if a.ndim != 3 or a.shape[2] != 3:
        raise ValueError("rgb must be [H][W][3] uint8")
    if _lib().srt_host_write_bmp(str(path).encode(), a.ctypes.data, a.shape[1], a.shape[0]) != 0:
        raise OSError(f"cannot write {path}")


def write_png(path, rgb: np.ndarray):
    """The same dense RGB8 frame as an 8-bit RGB PNG (lossless; not a reference format — the reference saves BMP)."""
    a = np.ascontiguousarray(rgb, np.uint8)
    if a.ndim != 3 or a.shape[2] != 3:
        raise ValueError("rgb must be [H][W][3] uint8")
    if _lib().srt_host_write_png(str(path).encode(), a.ctypes.data, a.shape[1], a.shape[0]) != 0:
        raise OSError(f"cannot write {path}")


def write_image(path, rgb: np.ndarray):
    """write_png for a path ending in .png, else the reference's BMP."""
    (write_png if str(path).lower().endswith(".png") else write_bmp)(path, rgb)


def decode_image(path, use_sidecar: bool = False) -> np.ndarray:
    """A texture file as the [H, W, 3] uint8 pixels stbi_load(path, ..., 3) returns (Object.cpp:57): PNG / JPEG through the
    library's own decoders; other formats fail (use_sidecar is ignored: the sidecar fallback of round 1 is gone)."""
    w, h = C.c_int32(0), C.c_int32(0)
    if _lib().srt_host_decode_image(str(path).encode(), int(use_sidecar), None, 0, C.byref(w), C.byref(h)) != 0:
        raise ValueError(f"cannot decode {path}")
    out = np.zeros((h.value, w.value, 3), np.uint8)
    if _lib().srt_host_decode_image(str(path).encode(), int(use_sidecar), out.ctypes.data, out.nbytes, C.byref(w), C.byref(h)) != 0:
        raise ValueError(f"cannot decode {path}")
    return out


def set_asset_cache(enabled: bool):
    """Parse each OBJ/MTL and decode each texture once per process instead of once per frame (bit-identical)."""
    _lib().srt_host_set_asset_cache(1 if enabled else 0)


def parse_obj_number(text: str):
    out = C.c_double(0)
    ok = _lib().srt_host_parse_double(text.encode(), C.byref(out))
    return out.value if ok else None
