"""ctypes binding of the C ABI (include/srt.h) — the same entry points a C++ host would link.

Plumbing only.  The render path is the CUDA library `lib/libsrt_b200.so`; if it is missing or there
is no CUDA device, everything here raises — there is no CPU fallback of any kind.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path
from typing import Optional

import numpy as np

from . import _abi

import os as _os

LIB_PATH = Path(_os.environ.get("SRT_LIB") or (Path(__file__).resolve().parent / "lib" / "libsrt_b200.so"))   # SRT_LIB: A/B builds

EXPORTS = [
    "srt_abi_version", "srt_device_count", "srt_context_create", "srt_context_destroy",
    "srt_context_set_stream", "srt_context_set_bands", "srt_context_get_info", "srt_scene_upload", "srt_scene_update",
    "srt_scene_destroy", "srt_render", "srt_sync", "srt_last_stats", "srt_default_params",
    "srt_light_positions", "srt_host_alloc", "srt_host_free", "srt_device_alloc", "srt_device_free",
    "srt_last_error", "srt_eval_leaf", "srt_render_sparse", "srt_fetch_sparse",
    "srt_mesh_upload", "srt_mesh_destroy", "srt_mesh_info", "srt_scene_prep", "srt_scene_prep_status",
    "srt_scene_counts", "srt_scene_download", "srt_scene_download_vnormals", "srt_measure_read_bandwidth",
    "srt_sparse_from_frame_static",
    "srt_render_sparse_begin", "srt_render_sparse_count", "srt_fetch_sparse_async",
    "srt_context_set_stage_timing", "srt_last_stage_ms",
    "srt_ipc_export", "srt_ipc_open", "srt_ipc_close",
    "srt_mgpu_create", "srt_mgpu_destroy", "srt_mgpu_info", "srt_mgpu_scene_upload", "srt_mgpu_render_frame",
    "srt_mgpu_render_series", "srt_mgpu_sync", "srt_mgpu_last_stats",
]


class SrtError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"srt error {code}: {message}")
        self.code = code


_lib = None


def lib():
    """Load libsrt_b200.so (built by `make` / __graft_entry__.build()).  Raises if absent."""
    global _lib
    if _lib is None:
        if not LIB_PATH.exists():
            raise FileNotFoundError(
                f"{LIB_PATH} not built — run `make` (or __graft_entry__.build()); there is no fallback path")
        L = C.CDLL(str(LIB_PATH))
        vp = C.c_void_p
        L.srt_abi_version.restype = C.c_int
        L.srt_device_count.argtypes = [C.POINTER(C.c_int)]
        L.srt_context_create.argtypes = [C.c_int, C.POINTER(vp)]
        L.srt_context_destroy.argtypes = [vp]
        L.srt_context_destroy.restype = None
        L.srt_context_set_stream.argtypes = [vp, vp]
        L.srt_context_set_bands.argtypes = [vp, C.c_int]
        L.srt_context_get_info.argtypes = [vp, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int),
                                           C.POINTER(C.c_size_t)]
        L.srt_scene_upload.argtypes = [vp, C.POINTER(_abi.SceneDesc), C.POINTER(vp)]
        L.srt_scene_update.argtypes = [vp, vp, C.POINTER(_abi.SceneDesc)]
        L.srt_scene_destroy.argtypes = [vp]
        L.srt_scene_destroy.restype = None
        L.srt_render.argtypes = [vp, vp, C.POINTER(_abi.RenderParams), C.POINTER(_abi.Frame)]
        L.srt_sync.argtypes = [vp]
        L.srt_last_stats.argtypes = [vp, C.POINTER(_abi.Stats)]
        L.srt_default_params.argtypes = [C.POINTER(_abi.RenderParams)]
        L.srt_default_params.restype = None
        L.srt_light_positions.argtypes = [C.POINTER(_abi.RenderParams), C.POINTER(C.c_float)]
        L.srt_render_sparse.argtypes = [vp, vp, C.POINTER(_abi.RenderParams), C.POINTER(C.c_uint64)]
        L.srt_render_sparse_begin.argtypes = [vp, vp, C.POINTER(_abi.RenderParams)]
        L.srt_render_sparse_count.argtypes = [vp, C.POINTER(C.c_uint64)]
        L.srt_fetch_sparse_async.argtypes = [vp, C.POINTER(C.c_float), C.POINTER(C.c_float), C.c_uint64]
        L.srt_fetch_sparse.argtypes = [vp, C.POINTER(C.c_float), C.POINTER(C.c_float), C.c_uint64]
        L.srt_sparse_from_frame_static.argtypes = [vp, vp, C.c_int32, C.c_int32, C.c_int32, C.c_int32, vp, vp, vp]
        L.srt_context_set_stage_timing.argtypes = [vp, C.c_int]
        L.srt_last_stage_ms.argtypes = [vp, C.POINTER(C.c_float)]
        L.srt_ipc_export.argtypes = [vp, vp, vp]
        L.srt_ipc_open.argtypes = [vp, vp, C.POINTER(vp)]
        L.srt_ipc_close.argtypes = [vp, vp]
        L.srt_mgpu_create.argtypes = [C.POINTER(C.c_int), C.c_int, C.POINTER(vp)]
        L.srt_mgpu_destroy.argtypes = [vp]
        L.srt_mgpu_destroy.restype = None
        L.srt_mgpu_info.argtypes = [vp, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.srt_mgpu_scene_upload.argtypes = [vp, C.POINTER(_abi.SceneDesc)]
        L.srt_mgpu_render_frame.argtypes = [vp, C.POINTER(_abi.RenderParams), C.POINTER(_abi.Frame), C.c_int]
        L.srt_mgpu_render_series.argtypes = [vp, C.POINTER(C.POINTER(_abi.SceneDesc)), C.POINTER(_abi.RenderParams), C.c_uint32,
                                             C.POINTER(vp)]
        L.srt_mgpu_sync.argtypes = [vp]
        L.srt_mgpu_last_stats.argtypes = [vp, C.POINTER(_abi.Stats)]
        L.srt_measure_read_bandwidth.argtypes = [C.c_void_p, C.c_size_t, C.c_int, C.POINTER(C.c_double)]
        L.srt_eval_leaf.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_float), C.POINTER(C.c_float), C.c_size_t]
        L.srt_host_alloc.argtypes = [C.c_size_t]
        L.srt_host_alloc.restype = vp
        L.srt_host_free.argtypes = [vp]
        L.srt_host_free.restype = None
        L.srt_device_alloc.argtypes = [vp, C.c_size_t]
        L.srt_device_alloc.restype = vp
        L.srt_device_free.argtypes = [vp, vp]
        L.srt_device_free.restype = None
        L.srt_last_error.restype = C.c_char_p
        L.srt_mesh_upload.argtypes = [vp, C.POINTER(_abi.MeshDesc), C.POINTER(vp)]
        L.srt_mesh_destroy.argtypes = [vp]
        L.srt_mesh_destroy.restype = None
        L.srt_mesh_info.argtypes = [vp, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), C.POINTER(C.c_uint32)]
        L.srt_scene_prep.argtypes = [vp, C.POINTER(_abi.PrepObject), C.c_uint32, C.POINTER(_abi.TextureDesc), C.c_uint32,
                                     C.c_int, C.POINTER(vp)]
        L.srt_scene_prep_status.argtypes = [vp, vp]
        L.srt_scene_counts.argtypes = [vp, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), C.POINTER(C.c_uint64),
                                       C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]
        L.srt_scene_download.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp]
        L.srt_scene_download_vnormals.argtypes = [vp, vp, vp, vp]
        _lib = L
    return _lib


def _check(rc: int):
    if rc != 0:
        raise SrtError(rc, lib().srt_last_error().decode(errors="replace"))


def device_count() -> int:
    n = C.c_int(0)
    rc = lib().srt_device_count(C.byref(n))
    return n.value if rc == 0 else 0


class PinnedArray:
    """numpy view over pinned host memory from srt_host_alloc (for asynchronous H2D/D2H copies)."""

    def __init__(self, shape, dtype):
        self.dtype = np.dtype(dtype)
        self.shape = tuple(int(s) for s in (shape if isinstance(shape, (tuple, list)) else (shape,)))
        n = int(np.prod(self.shape)) * self.dtype.itemsize
        self.ptr = lib().srt_host_alloc(max(n, 1))
        if not self.ptr:
            raise SrtError(_abi.SRT_ERR_OOM, "srt_host_alloc failed")
        buf = (C.c_char * max(n, 1)).from_address(self.ptr)
        self.array = np.frombuffer(buf, dtype=self.dtype, count=int(np.prod(self.shape))).reshape(self.shape)

    def free(self):
        if self.ptr:
            self.array = None
            lib().srt_host_free(self.ptr)
            self.ptr = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class Scene:
    def __init__(self, ctx: "Context", handle):
        self.ctx = ctx
        self.h = handle

    def update(self, arrays_or_desc):
        """Re-upload new contents (next frame of a series) into the same device buffers, async."""
        d = arrays_or_desc.desc() if hasattr(arrays_or_desc, "desc") else arrays_or_desc
        self._keep = (arrays_or_desc, d)
        _check(lib().srt_scene_update(self.ctx.h, self.h, C.byref(d)))

    def counts(self) -> dict:
        nn, nt, nf, no, md = C.c_uint64(), C.c_uint64(), C.c_uint64(), C.c_uint32(), C.c_uint32()
        _check(lib().srt_scene_counts(self.h, C.byref(nn), C.byref(nt), C.byref(nf), C.byref(no), C.byref(md)))
        return dict(n_nodes=nn.value, n_tris=nt.value, n_flat=nf.value, n_objects=no.value, max_depth=md.value)

    def download(self) -> _abi.FlatSceneArrays:
        """The device scene as host arrays (srt_scene_download; tests and debugging)."""
        c = self.counts()
        nodes = np.zeros(c["n_nodes"], _abi.NODE_DTYPE)
        geom = np.zeros(c["n_tris"], _abi.GEOM_DTYPE)
        attr = np.zeros(c["n_tris"], _abi.ATTR_DTYPE)
        fgeom = np.zeros(c["n_flat"], _abi.GEOM_DTYPE)
        fattr = np.zeros(c["n_flat"], _abi.ATTR_DTYPE)
        objs = np.zeros(c["n_objects"], _abi.OBJECT_DTYPE)
        _check(lib().srt_scene_download(self.ctx.h, self.h, nodes.ctypes.data, geom.ctypes.data, attr.ctypes.data,
                                        fgeom.ctypes.data, fattr.ctypes.data, objs.ctypes.data))
        vn = np.zeros(c["n_tris"], _abi.VNORMAL_DTYPE)
        fvn = np.zeros(c["n_flat"], _abi.VNORMAL_DTYPE)
        have_vn = lib().srt_scene_download_vnormals(self.ctx.h, self.h, vn.ctypes.data, fvn.ctypes.data) == 0
        return _abi.FlatSceneArrays(nodes, geom, attr, objs, (), fgeom if len(fgeom) else None,
                                    fattr if len(fattr) else None, max_depth=c["max_depth"],
                                    vnormal=vn if have_vn else None, flat_vnormal=fvn if (have_vn and len(fvn)) else None)

    def prep_status(self):
        """Synchronises; raises SrtError(SRT_ERR_UNSUPPORTED) if the last srt_scene_prep met a non-finite coordinate."""
        _check(lib().srt_scene_prep_status(self.ctx.h, self.h))

    def destroy(self):
        if self.h:
            lib().srt_scene_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.destroy()
        except Exception:
            pass


class Mesh:
    """Object-space triangles resident on the GPU (srt_mesh): what loadObjFile put into objTriangles[name]."""

    def __init__(self, ctx: "Context", points, uv=None, tex=None, normals=None):
        pts = np.ascontiguousarray(points, np.float32).reshape(-1, 12)
        d = _abi.MeshDesc()
        d.points = pts.ctypes.data
        d.n_tris = len(pts)
        if uv is not None:
            uv = np.ascontiguousarray(uv, np.float32).reshape(-1, 6)
            assert len(uv) == len(pts)
            d.uv = uv.ctypes.data
        if tex is not None:
            tex = np.ascontiguousarray(tex, np.int32).reshape(-1)
            assert len(tex) == len(pts)
            d.tex = tex.ctypes.data
        if normals is not None:
            normals = np.ascontiguousarray(normals, np.float32).reshape(-1, 9)
            assert len(normals) == len(pts)
            d.normals = normals.ctypes.data
        h = C.c_void_p()
        _check(lib().srt_mesh_upload(ctx.h, C.byref(d), C.byref(h)))
        self.ctx, self.h, self.n_tris = ctx, h, len(pts)

    def info(self) -> dict:
        nt, nn, md = C.c_uint64(), C.c_uint64(), C.c_uint32()
        _check(lib().srt_mesh_info(self.h, C.byref(nt), C.byref(nn), C.byref(md)))
        return dict(n_tris=nt.value, n_nodes=nn.value, max_depth=md.value)

    def destroy(self):
        if self.h:
            lib().srt_mesh_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.destroy()
        except Exception:
            pass


class Context:
    """One per GPU (srt_context).  Renders are asynchronous on the context stream."""

    def __init__(self, device: int = 0, stream: Optional[int] = None):
        h = C.c_void_p()
        _check(lib().srt_context_create(int(device), C.byref(h)))
        self.h = h
        self.device = int(device)
        self.stream = stream              # None: the context's own stream
        if stream is not None:
            _check(lib().srt_context_set_stream(self.h, C.c_void_p(stream)))

    def set_bands(self, bands: int) -> int:
        """Row bands per frame of the default pipeline (2, the default: one frame's halves overlap on two streams, for
        callers that render one frame at a time; 1: in one piece).  Returns the previous value."""
        before = lib().srt_context_set_bands(self.h, int(bands))
        if before < 0:
            _check(before)
        return before

    def set_stage_timing(self, enabled: bool):
        _check(lib().srt_context_set_stage_timing(self.h, 1 if enabled else 0))

    def last_stage_ms(self) -> dict:
        """{classify, primary, shadow, shade} device ms of the last one-piece default-pipeline render (stage timing on)."""
        ms = (C.c_float * 4)()
        _check(lib().srt_last_stage_ms(self.h, ms))
        return dict(classify=ms[0], primary=ms[1], shadow=ms[2], shade=ms[3])

    def info(self):
        sm, ma, mi, l2 = C.c_int(), C.c_int(), C.c_int(), C.c_size_t()
        _check(lib().srt_context_get_info(self.h, C.byref(sm), C.byref(ma), C.byref(mi), C.byref(l2)))
        return dict(sm_count=sm.value, cc=(ma.value, mi.value), l2_bytes=l2.value)

    def upload(self, arrays: _abi.FlatSceneArrays) -> Scene:
        d = arrays.desc()
        h = C.c_void_p()
        _check(lib().srt_scene_upload(self.h, C.byref(d), C.byref(h)))
        return Scene(self, h)

    def prep_scene(self, objects, textures=None, with_flat_lists: bool = False, scene: Optional[Scene] = None) -> Scene:
        """srt_scene_prep: transform + BVH build + flatten on the device, asynchronous.  objects: sequence of
        (Mesh, matrices [k][16] float32, color, (ambient, specular, shininess)) in objTriangles iteration order;
        textures: list of [H][W][3] uint8 arrays (None keeps the textures of an existing scene)."""
        n = len(objects)
        arr = (_abi.PrepObject * max(n, 1))()
        keep = []
        for i, (mesh, mats, color, props) in enumerate(objects):
            m = np.ascontiguousarray(mats, np.float32).reshape(-1, 16)
            keep.append(m)
            arr[i].mesh = mesh.h
            arr[i].matrices = m.ctypes.data if len(m) else None
            arr[i].n_matrices = len(m)
            arr[i].color[:] = [float(x) for x in color]
            arr[i].ambient, arr[i].specular, arr[i].shininess = [float(x) for x in props]
        tex_arr, n_tex = None, 0
        if textures is not None:
            tex_keep = [np.ascontiguousarray(t, np.uint8) for t in textures]
            keep.append(tex_keep)
            n_tex = len(tex_keep)
            tex_arr = (_abi.TextureDesc * max(n_tex, 1))()
            for i, t in enumerate(tex_keep):
                tex_arr[i].rgb = t.ctypes.data
                tex_arr[i].height, tex_arr[i].width = t.shape[0], t.shape[1]
        h = scene.h if scene is not None else C.c_void_p()
        _check(lib().srt_scene_prep(self.h, arr, n, tex_arr, n_tex, 1 if with_flat_lists else 0, C.byref(h)))
        if scene is None:
            scene = Scene(self, h)
        return scene

    def render_raw(self, scene: Scene, params: _abi.RenderParams, rgb=None, obj_id=None, tri_id=None, t=None):
        """Pointers (ints) may be host or device addresses; asynchronous."""
        fr = _abi.Frame()
        fr.rgb, fr.obj_id, fr.tri_id, fr.t = rgb, obj_id, tri_id, t
        _check(lib().srt_render(self.h, scene.h, C.byref(params), C.byref(fr)))

    def render(self, scene: Scene, params: _abi.RenderParams, ids: bool = True, want_t: bool = True):
        """Render into fresh host arrays (synchronous convenience wrapper)."""
        H, W = params.height, params.width
        rgb = np.zeros((H, W, 3), np.uint8)
        obj = np.full((H, W), -1, np.int32) if ids else None
        tri = np.full((H, W), -1, np.int32) if ids else None
        tt = np.full((H, W), np.inf, np.float32) if want_t else None
        self.render_raw(scene, params, rgb.ctypes.data, obj.ctypes.data if ids else None,
                        tri.ctypes.data if ids else None, tt.ctypes.data if want_t else None)
        self.sync()
        return dict(rgb=rgb, obj_id=obj, tri_id=tri, t=tt, stats=self.last_stats())

    def render_sparse(self, scene: Scene, params: _abi.RenderParams):
        """The reference's ImageData: (points[n][2], colors[n][3]) of the drawn pixels, column-major order."""
        n = C.c_uint64(0)
        _check(lib().srt_render_sparse(self.h, scene.h, C.byref(params), C.byref(n)))
        pts = np.zeros((n.value, 2), np.float32)
        col = np.zeros((n.value, 3), np.float32)
        fp = C.POINTER(C.c_float)
        _check(lib().srt_fetch_sparse(self.h, pts.ctypes.data_as(fp), col.ctypes.data_as(fp), n.value))
        return pts, col

    def measure_read_bandwidth(self, nbytes: int, repeats: int = 20) -> float:
        """GB/s at which all SMs stream an `nbytes` device buffer (<= L2 size: L2 -> SM bandwidth; >> L2: HBM)."""
        out = C.c_double(0.0)
        _check(lib().srt_measure_read_bandwidth(self.h, int(nbytes), int(repeats), C.byref(out)))
        return out.value

    def sync(self):
        _check(lib().srt_sync(self.h))

    def device_alloc(self, nbytes: int) -> int:
        p = lib().srt_device_alloc(self.h, int(nbytes))
        if not p:
            raise SrtError(_abi.SRT_ERR_OOM, "srt_device_alloc failed")
        return int(p)

    def device_free(self, ptr: int):
        lib().srt_device_free(self.h, C.c_void_p(ptr))

    def ipc_export(self, ptr: int) -> bytes:
        buf = C.create_string_buffer(64)
        _check(lib().srt_ipc_export(self.h, C.c_void_p(ptr), buf))
        return buf.raw

    def ipc_open(self, handle: bytes) -> int:
        out = C.c_void_p()
        _check(lib().srt_ipc_open(self.h, C.create_string_buffer(handle, 64), C.byref(out)))
        return int(out.value)

    def ipc_close(self, ptr: int):
        _check(lib().srt_ipc_close(self.h, C.c_void_p(ptr)))

    def last_stats(self) -> dict:
        s = _abi.Stats()
        _check(lib().srt_last_stats(self.h, C.byref(s)))
        return dict(primary_rays=int(s.primary_rays), hit_pixels=int(s.hit_pixels), shadow_rays=int(s.shadow_rays),
                    node_visits=int(s.node_visits), tri_tests=int(s.tri_tests), exact_slab_tests=int(s.exact_slab_tests),
                    kernel_ms=float(s.kernel_ms), kernel_launches=int(s.kernel_launches))

    def destroy(self):
        if self.h:
            lib().srt_context_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.destroy()
        except Exception:
            pass


class MultiGpu:
    """srt_mgpu: several GPUs of one box from one host thread (include/srt.h).  devices: CUDA ordinals, may repeat."""

    def __init__(self, devices):
        devices = [int(d) for d in devices]
        arr = (C.c_int * len(devices))(*devices)
        h = C.c_void_p()
        _check(lib().srt_mgpu_create(arr, len(devices), C.byref(h)))
        self.h, self.devices = h, devices
        self._keep = None

    def info(self) -> dict:
        n, pw = C.c_int(), C.c_int()
        _check(lib().srt_mgpu_info(self.h, C.byref(n), C.byref(pw)))
        return dict(n_devices=n.value, peer_writes=pw.value)

    def upload(self, arrays_or_desc):
        d = arrays_or_desc.desc() if hasattr(arrays_or_desc, "desc") else arrays_or_desc
        self._keep = (arrays_or_desc, d)
        _check(lib().srt_mgpu_scene_upload(self.h, C.byref(d)))

    def render_frame_raw(self, params: _abi.RenderParams, rgb=None, obj_id=None, tri_id=None, t=None, row_block: int = 0):
        fr = _abi.Frame()
        fr.rgb, fr.obj_id, fr.tri_id, fr.t = rgb, obj_id, tri_id, t
        _check(lib().srt_mgpu_render_frame(self.h, C.byref(params), C.byref(fr), int(row_block)))

    def render_frame(self, params: _abi.RenderParams, row_block: int = 0, ids: bool = True, want_t: bool = True):
        H, W = params.height, params.width
        rgb = np.zeros((H, W, 3), np.uint8)
        obj = np.full((H, W), -1, np.int32) if ids else None
        tri = np.full((H, W), -1, np.int32) if ids else None
        tt = np.full((H, W), np.inf, np.float32) if want_t else None
        self.render_frame_raw(params, rgb.ctypes.data, obj.ctypes.data if ids else None, tri.ctypes.data if ids else None,
                              tt.ctypes.data if want_t else None, row_block)
        self.sync()
        return dict(rgb=rgb, obj_id=obj, tri_id=tri, t=tt, stats=self.last_stats())

    def render_series(self, scenes, params, out_frames):
        """scenes: FlatSceneArrays (or descs) per frame; params: RenderParams per frame; out_frames: host arrays [H][W][3] uint8."""
        n = len(scenes)
        descs = [s.desc() if hasattr(s, "desc") else s for s in scenes]
        dp = (C.POINTER(_abi.SceneDesc) * max(n, 1))(*[C.pointer(d) for d in descs])
        pa = (_abi.RenderParams * max(n, 1))(*params)
        op = (C.c_void_p * max(n, 1))(*[int(o.ctypes.data) if hasattr(o, "ctypes") else int(o) for o in out_frames])
        self._series_keep = (scenes, descs, dp, pa, op, out_frames)
        _check(lib().srt_mgpu_render_series(self.h, dp, pa, n, op))

    def sync(self):
        _check(lib().srt_mgpu_sync(self.h))

    def last_stats(self) -> dict:
        s = _abi.Stats()
        _check(lib().srt_mgpu_last_stats(self.h, C.byref(s)))
        return dict(primary_rays=int(s.primary_rays), hit_pixels=int(s.hit_pixels), shadow_rays=int(s.shadow_rays),
                    exact_slab_tests=int(s.exact_slab_tests), kernel_ms=float(s.kernel_ms), kernel_launches=int(s.kernel_launches))

    def destroy(self):
        if self.h:
            lib().srt_mgpu_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.destroy()
        except Exception:
            pass


LEAF_POWF, LEAF_SLAB, LEAF_TRIANGLE, LEAF_PHONG, LEAF_TONE, LEAF_SMOOTH_NORMAL, LEAF_TRIANGLE_FILTERED = range(7)
_LEAF_WIDTHS = {LEAF_POWF: (2, 1), LEAF_SLAB: (12, 3), LEAF_TRIANGLE: (15, 1), LEAF_PHONG: (19, 3), LEAF_TONE: (3, 1),
                LEAF_SMOOTH_NORMAL: (21, 3), LEAF_TRIANGLE_FILTERED: (15, 3)}


def eval_leaf(ctx: "Context", op: int, cases) -> np.ndarray:
    """srt_eval_leaf: one device leaf function of the render kernels on n cases ([n][in_width] float32)."""
    wi, wo = _LEAF_WIDTHS[op]
    x = np.ascontiguousarray(cases, np.float32).reshape(-1, wi)
    out = np.zeros((len(x), wo), np.float32)
    fp = C.POINTER(C.c_float)
    _check(lib().srt_eval_leaf(ctx.h, op, x.ctypes.data_as(fp), out.ctypes.data_as(fp), len(x)))
    return out


def light_positions(params: _abi.RenderParams) -> np.ndarray:
    out = np.zeros((max(params.light_count, 0), 3), np.float32)
    if params.light_count > 0:
        _check(lib().srt_light_positions(C.byref(params), out.ctypes.data_as(C.POINTER(C.c_float))))
    return out
