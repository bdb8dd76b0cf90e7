"""Device scene prep: a scene backend (scenes.py protocol) that records the reference's scene calls and replays
the per-frame part — transformTriangles, createBoundingHierarchy, flatten — on the GPU (srt_scene_prep, srt.h).

The reference's frame loop (simple_raytracer.cpp:534-794) re-parses every OBJ, re-transforms every vertex and
rebuilds every hierarchy on the host for each frame.  Here a mesh is parsed once (the host mirror's OBJ reader,
through its asset cache), uploaded once in object space, and each frame costs one small upload (the matrices) plus
a few kernel launches whose result is bit-identical to what the host path produces for the same calls
(tests/test_gpu_prep.py).  No arithmetic happens in Python.
"""
from __future__ import annotations

from typing import Dict, List, Optional

import numpy as np

from . import capi, host

LOAD_COLOR = (1.0, 0.0, 0.0)             # loadObjFile, Object.cpp:29
LOAD_PROPERTIES = (0.2, 0.5, 15.0)       # Object.cpp:31-34


class MeshCache:
    """Per context: OBJ path -> device mesh, plus the table of textures those meshes reference.  Texture slots are
    global to the cache (a mesh's per-triangle slots are fixed when it is uploaded)."""

    def __init__(self, ctx: capi.Context, assets=None):
        self.ctx = ctx
        self.assets = assets
        self.meshes: Dict[str, capi.Mesh] = {}
        self.texture_names: List[str] = []
        self.textures: List[np.ndarray] = []

    def mesh(self, path: str) -> capi.Mesh:
        m = self.meshes.get(path)
        if m is not None:
            return m
        hs = host.HostScene(self.assets)
        try:
            hs.load_obj(path)
            tris, _flags = hs.triangles(path)
            slots, names = hs.triangle_textures(path)
            remap = np.full(len(names) + 1, -1, np.int32)          # local slot -> cache slot (last entry: untextured)
            for k, name in enumerate(names):
                if name not in self.texture_names:
                    self.texture_names.append(name)
                    self.textures.append(hs.texture(name))         # KeyError: textured triangle without texture (:354)
                remap[k] = self.texture_names.index(name)
            tex = remap[slots] if len(names) else None
            uv = tris[:, 21:27] if len(names) else None
            m = capi.Mesh(self.ctx, tris[:, 0:12], uv, tex, normals=tris[:, 12:21])
        finally:
            hs.close()
        self.meshes[path] = m
        return m

    def close(self):
        for m in self.meshes.values():
            m.destroy()
        self.meshes.clear()


class _Object:
    __slots__ = ("mesh", "matrices", "built_at")

    def __init__(self, mesh, matrices=()):
        self.mesh = mesh
        self.matrices = list(matrices)
        self.built_at = None               # number of matrices applied when createBoundingHierarchy was called


class DeviceScene:
    """One frame's ObjectManager, recorded.  Implements the backend protocol of scenes.py; the matrix helpers are
    the host mirror's (Transformation / glm, C++)."""

    def __init__(self, cache: MeshCache):
        self.cache = cache
        self._math = host.HostScene(cache.assets)
        self.objects: Dict[str, _Object] = {}      # insertion order = first touch of objTriangles[name]
        self.colors: Dict[str, tuple] = {}
        self.props: Dict[str, tuple] = {}

    def close(self):
        self._math.close()

    # -- ObjectManager
    def load_obj(self, path: str) -> int:
        mesh = self.cache.mesh(path)
        self.colors[path] = LOAD_COLOR
        self.props[path] = LOAD_PROPERTIES
        if path in self.objects:
            self.objects[path].__init__(mesh)      # objTriangles[name] = triangles (:169): key keeps its place
        else:
            self.objects[path] = _Object(mesh)
        return mesh.n_tris

    def set_color(self, name, rgb):
        self.colors[name] = tuple(float(x) for x in rgb)

    def set_properties(self, name, a, s, sh):
        self.props[name] = (float(a), float(s), float(sh))

    def get_properties(self, name):
        if name not in self.props:
            raise KeyError(name)
        return tuple(np.float32(x) for x in self.props[name])

    def clone(self, src, dst, copy_properties):
        if src not in self.objects:
            raise KeyError(src)                    # std::out_of_range from getTriangles (Object.cpp:174)
        o = self.objects[src]
        if dst in self.objects:
            self.objects[dst].__init__(o.mesh, o.matrices)
        else:
            self.objects[dst] = _Object(o.mesh, o.matrices)
        if copy_properties:
            self.props[dst] = self.props.get(src, (0.0, 0.0, 0.0))

    def transform(self, name, m16):
        if name not in self.objects:
            raise KeyError(f"{name}: transformTriangles on a name that was never loaded is not supported on the device path")
        self.objects[name].matrices.append(np.asarray(m16, np.float32).reshape(16).copy())

    def create_bounding_hierarchy(self, name):
        if name not in self.objects:
            raise KeyError(name)
        self.objects[name].built_at = len(self.objects[name].matrices)

    def object_names(self):
        return host.object_order(list(self.objects.keys()))

    # -- Transformation / glm: the host mirror's
    def __getattr__(self, item):
        if item.startswith("xf_") or item in ("mat_inverse", "mat_mul", "mat_vec", "radians", "cosf", "sinf"):
            return getattr(self._math, item)
        raise AttributeError(item)

    # -- the device part
    def commit(self, scene: Optional[capi.Scene] = None, with_flat_lists: bool = False,
               ctx: Optional[capi.Context] = None) -> capi.Scene:
        """Transform + build + flatten on the GPU, asynchronous on the stream of `ctx` (default: the scene's context,
        else the cache's; any context of the device the meshes live on will do)."""
        ctx = ctx or (scene.ctx if scene is not None else self.cache.ctx)
        recipe = []
        for name in self.object_names():
            o = self.objects[name]
            if o.built_at is None:
                raise capi.SrtError(-5, f"object '{name}' has no bounding hierarchy (null Node* at simple_raytracer.cpp:298)")
            if o.built_at != len(o.matrices):
                raise capi.SrtError(-6, f"object '{name}': transformTriangles after createBoundingHierarchy leaves the hierarchy "
                                        "stale in the reference; not supported on the device path")
            mats = np.stack(o.matrices) if o.matrices else np.zeros((0, 16), np.float32)
            recipe.append((o.mesh, mats, self.colors.get(name, (0.0, 0.0, 0.0)), self.props.get(name, (0.0, 0.0, 0.0))))
        new_textures = scene is None or getattr(scene, "_n_textures", -1) != len(self.cache.textures)
        out = ctx.prep_scene(recipe, self.cache.textures if new_textures else None, with_flat_lists, scene)
        out._n_textures = len(self.cache.textures)
        return out
