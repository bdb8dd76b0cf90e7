"""ctypes mirror of include/srt.h (the C ABI).  Plumbing only — no arithmetic happens here.

Every structure below must stay byte-identical to its C counterpart; tests/test_abi.py checks the
sizes against the constants the shared library reports.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

SRT_MAX_LIGHTS = 64
SRT_LEAF_BIT = 0x80000000
SRT_NO_INDEX = 0xFFFFFFFF

SRT_OK = 0
SRT_ERR_INVALID_ARG = -1
SRT_ERR_NO_DEVICE = -2
SRT_ERR_CUDA = -3
SRT_ERR_OOM = -4
SRT_ERR_SCENE = -5
SRT_ERR_UNSUPPORTED = -6

LIGHTS_STAIRCASE = 0      # the reference's sequence (parity)
LIGHTS_JITTERED = 1       # counter-based jittered samples in the staircase's box (not a reference mode)
TEX_TRUNCATE, TEX_NEAREST, TEX_BILINEAR = 0, 1, 2     # srt_render_params::texture_mode

VARIANT_DEFAULT = 0
VARIANT_SIMPLE = 1
VARIANT_FUSED = 2
VARIANT_OVERLAP = 4
VARIANT_COUNT = 100
VARIANT_TIMELINE = 101

# numpy record layouts of the flattened scene (include/srt.h)
NODE_DTYPE = np.dtype([("bmin", "<f4", 3), ("a", "<u4"), ("bmax", "<f4", 3), ("b", "<u4")])
GEOM_DTYPE = np.dtype([("p0", "<f4", 3), ("obj", "<i4"), ("e1", "<f4", 3), ("tex", "<i4"),
                       ("e2", "<f4", 3), ("reserved", "<u4")])
ATTR_DTYPE = np.dtype([("n", "<f4", 3), ("pad0", "<f4"), ("uv", "<f4", (3, 2)), ("pad1", "<f4", 2)])
OBJECT_DTYPE = np.dtype([("color", "<f4", 3), ("ambient", "<f4"), ("specular", "<f4"), ("shininess", "<f4"),
                         ("root", "<u4"), ("tri_first", "<u4"), ("tri_count", "<u4"), ("flat_first", "<u4"),
                         ("box_min", "<f4", 3), ("box_max", "<f4", 3)])
VNORMAL_DTYPE = np.dtype([("n", "<f4", (3, 4))])
assert VNORMAL_DTYPE.itemsize == 48
assert NODE_DTYPE.itemsize == 32 and GEOM_DTYPE.itemsize == 48 and ATTR_DTYPE.itemsize == 48
assert OBJECT_DTYPE.itemsize == 64


class TextureDesc(C.Structure):
    _fields_ = [("rgb", C.c_void_p), ("width", C.c_int32), ("height", C.c_int32)]


class SceneDesc(C.Structure):
    _fields_ = [
        ("nodes", C.c_void_p), ("n_nodes", C.c_uint64),
        ("geom", C.c_void_p), ("attr", C.c_void_p), ("n_tris", C.c_uint64),
        ("flat_geom", C.c_void_p), ("flat_attr", C.c_void_p), ("n_flat", C.c_uint64),
        ("objects", C.c_void_p), ("n_objects", C.c_uint32),
        ("textures", C.POINTER(TextureDesc)), ("n_textures", C.c_uint32),
        ("vnormal", C.c_void_p), ("flat_vnormal", C.c_void_p),
    ]


class RenderParams(C.Structure):
    _fields_ = [
        ("width", C.c_int32), ("height", C.c_int32), ("focal", C.c_float),
        ("light_pos", C.c_float * 3), ("light_count", C.c_int32), ("light_step", C.c_float),
        ("light_color", C.c_float * 3), ("shadow_div", C.c_float), ("tone_k", C.c_float),
        ("gamma", C.c_float), ("sky", C.c_uint8 * 3), ("pad0", C.c_uint8),
        ("use_bvh", C.c_int32), ("row_begin", C.c_int32), ("row_end", C.c_int32),
        ("variant", C.c_int32), ("no_sky_fill", C.c_int32), ("row_block", C.c_int32),
        ("shard_index", C.c_int32), ("shard_count", C.c_int32),
        ("light_mode", C.c_int32), ("light_seed", C.c_uint32), ("smooth_normals", C.c_int32),
        ("normalise_lights", C.c_int32), ("shadow_segment", C.c_int32), ("shadow_eps", C.c_float), ("texture_mode", C.c_int32),
    ]


class MeshDesc(C.Structure):
    _fields_ = [("points", C.c_void_p), ("uv", C.c_void_p), ("tex", C.c_void_p), ("n_tris", C.c_uint64),
                ("normals", C.c_void_p)]


class PrepObject(C.Structure):
    _fields_ = [("mesh", C.c_void_p), ("matrices", C.c_void_p), ("n_matrices", C.c_uint32),
                ("color", C.c_float * 3), ("ambient", C.c_float), ("specular", C.c_float), ("shininess", C.c_float)]


class Frame(C.Structure):
    _fields_ = [("rgb", C.c_void_p), ("obj_id", C.c_void_p), ("tri_id", C.c_void_p), ("t", C.c_void_p)]


class Stats(C.Structure):
    _fields_ = [
        ("primary_rays", C.c_uint64), ("hit_pixels", C.c_uint64), ("shadow_rays", C.c_uint64),
        ("node_visits", C.c_uint64), ("tri_tests", C.c_uint64), ("exact_slab_tests", C.c_uint64),
        ("kernel_ms", C.c_float), ("kernel_launches", C.c_int32),
    ]


def default_params(width: int, height: int, focal: float, light_pos, light_count: int = 1,
                   use_bvh: bool = True) -> RenderParams:
    """The reference's literals with the per-config values filled in."""
    p = RenderParams()
    p.width, p.height, p.focal = int(width), int(height), float(focal)
    p.light_pos[:] = [float(x) for x in light_pos[:3]]
    p.light_count = int(light_count)
    p.light_step = 3.0            # simple_raytracer.cpp:372-382
    p.light_color[:] = [1.0, 1.0, 1.0]   # :433
    p.shadow_div = 5.0            # :369
    p.tone_k = 0.5                # :391
    p.gamma = 1.1                 # :396
    p.sky[:] = [173, 216, 230]    # :476
    p.use_bvh = 1 if use_bvh else 0
    p.row_begin = 0
    p.row_end = 0
    p.variant = 0
    p.no_sky_fill = 0
    return p


class FlatSceneArrays:
    """Owner of the numpy arrays behind a SceneDesc (keeps them alive while C reads them)."""

    def __init__(self, nodes, geom, attr, objects, textures=(), flat_geom=None, flat_attr=None,
                 object_names=(), texture_names=(), max_depth=0, vnormal=None, flat_vnormal=None):
        self.nodes = np.ascontiguousarray(nodes, dtype=NODE_DTYPE)
        self.geom = np.ascontiguousarray(geom, dtype=GEOM_DTYPE)
        self.attr = np.ascontiguousarray(attr, dtype=ATTR_DTYPE)
        self.objects = np.ascontiguousarray(objects, dtype=OBJECT_DTYPE)
        self.textures = [np.ascontiguousarray(t, dtype=np.uint8) for t in textures]   # each [H][W][3]
        self.flat_geom = None if flat_geom is None else np.ascontiguousarray(flat_geom, dtype=GEOM_DTYPE)
        self.flat_attr = None if flat_attr is None else np.ascontiguousarray(flat_attr, dtype=ATTR_DTYPE)
        self.vnormal = None if vnormal is None else np.ascontiguousarray(vnormal, dtype=VNORMAL_DTYPE)
        self.flat_vnormal = None if flat_vnormal is None else np.ascontiguousarray(flat_vnormal, dtype=VNORMAL_DTYPE)
        self.object_names = list(object_names)
        self.texture_names = list(texture_names)
        self.max_depth = int(max_depth)
        self._tex_descs = None

    def desc(self) -> SceneDesc:
        d = SceneDesc()
        d.nodes = self.nodes.ctypes.data
        d.n_nodes = len(self.nodes)
        d.geom = self.geom.ctypes.data
        d.attr = self.attr.ctypes.data
        d.n_tris = len(self.geom)
        if self.flat_geom is not None and len(self.flat_geom):
            d.flat_geom = self.flat_geom.ctypes.data
            d.flat_attr = self.flat_attr.ctypes.data
            d.n_flat = len(self.flat_geom)
        d.objects = self.objects.ctypes.data
        d.n_objects = len(self.objects)
        if self.vnormal is not None and len(self.vnormal) == len(self.geom) and len(self.geom):
            d.vnormal = self.vnormal.ctypes.data
        if self.flat_vnormal is not None and d.n_flat and len(self.flat_vnormal) == d.n_flat:
            d.flat_vnormal = self.flat_vnormal.ctypes.data
        if self.textures:
            arr = (TextureDesc * len(self.textures))()
            for i, t in enumerate(self.textures):
                arr[i].rgb = t.ctypes.data
                arr[i].height, arr[i].width = t.shape[0], t.shape[1]
            self._tex_descs = arr
            d.textures = C.cast(arr, C.POINTER(TextureDesc))
            d.n_textures = len(self.textures)
        return d

    def nbytes(self) -> int:
        n = self.nodes.nbytes + self.geom.nbytes + self.attr.nbytes + self.objects.nbytes
        n += sum(t.nbytes for t in self.textures)
        if self.flat_geom is not None:
            n += self.flat_geom.nbytes + self.flat_attr.nbytes
        return n

    @staticmethod
    def from_desc_pointer(desc_ptr, object_names=(), texture_names=(), max_depth=0) -> "FlatSceneArrays":
        """Deep-copy a C-owned srt_scene_desc into numpy arrays."""
        d = desc_ptr.contents if hasattr(desc_ptr, "contents") else desc_ptr

        def grab(ptr, n, dt):
            if not ptr or n == 0:
                return np.zeros(0, dtype=dt)
            buf = (C.c_char * (n * dt.itemsize)).from_address(ptr)
            return np.frombuffer(buf, dtype=dt, count=n).copy()

        nodes = grab(d.nodes, d.n_nodes, NODE_DTYPE)
        geom = grab(d.geom, d.n_tris, GEOM_DTYPE)
        attr = grab(d.attr, d.n_tris, ATTR_DTYPE)
        objects = grab(d.objects, d.n_objects, OBJECT_DTYPE)
        flat_geom = grab(d.flat_geom, d.n_flat, GEOM_DTYPE) if d.n_flat else None
        flat_attr = grab(d.flat_attr, d.n_flat, ATTR_DTYPE) if d.n_flat else None
        texs = []
        for i in range(d.n_textures):
            t = d.textures[i]
            buf = (C.c_uint8 * (t.width * t.height * 3)).from_address(t.rgb)
            texs.append(np.frombuffer(buf, dtype=np.uint8).reshape(t.height, t.width, 3).copy())
        vn = grab(d.vnormal, d.n_tris, VNORMAL_DTYPE) if d.vnormal else None
        fvn = grab(d.flat_vnormal, d.n_flat, VNORMAL_DTYPE) if (d.flat_vnormal and d.n_flat) else None
        return FlatSceneArrays(nodes, geom, attr, objects, texs, flat_geom, flat_attr, object_names,
                               texture_names, max_depth, vn, fvn)
