"""tests/golden/make_golden.py — generates the committed golden vectors from the UNMODIFIED reference.

Run in the container that has /root/reference (after `make ref`):   python tests/golden/make_golden.py
The reference ships no tests or reproducible fixtures for this path, so the pins are
outputs of the reference's own functions executed here through oracle/_ref/libsrt_ref.so:

  leaf_kat.npz     inputs + outputs of rayTriangleIntersection :42-75, intersectRayAabbNoOrigin :252-293,
                   intersectRayAabb :204-248, calculateBarycentricCoords :79-117, calculateTriangleNormal :32-37,
                   phongIllumination :144-200 and the tone-map / gamma / quantise lines :391-398, :447-449
                   on hand-made edge cases (axis-parallel rays, shared edge / vertex, t = 0, degenerate
                   triangle, box behind the origin, 0/0) and seeded random cases;
  frame_<key>.npz  whole small frames rendered by the reference (faithful mode: re-shading on every closer
                   hit like :429-449): rgb, obj_id, tri_id, t, the parameters, and — for the scenes small
                   enough to commit — the flattened scene itself, so the oracle can be checked without assets;
  frame_C3_a<angle>.npz  the same for views of the C3 / C5 scene around the camera orbit (the frames bench.py times).
"""
import ctypes as C
import sys
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
REPO = HERE.parent.parent
sys.path.insert(0, str(REPO))

from oracle import pyoracle  # noqa: E402
from simple_raytracer_b200 import scenes  # noqa: E402

f32 = np.float32
_f32p = C.POINTER(C.c_float)


def fp(a):
    return a.ctypes.data_as(_f32p)


def leaf_cases(rng):
    """Rays, triangles, boxes: edge cases first, then random."""
    n_rand = 3000
    o = [(0, 0, 0)] * 12 + [(1, 2, 3), (0, 0, 0), (0.5, 0.5, -1), (0, 0, 0)]
    d = [(0, 0, 400), (0, 5, 400), (5, 0, 400), (0, 0, 1), (1, 1, 1), (0, 0, 1), (0, 0, 1), (0, 0, 1),
         (-3, 7, 400), (0, 0, -1), (1e-20, 0, 1), (0, 0, 1), (0, 0, 1), (0, 1, 0), (0, 0, 1), (3, -4, 400)]
    tri = [
        (-1, -1, 5, 1, 1, -1, 5, 1, 0, 1, 5, 1),          # centre ray, plain hit
        (-1, -1, 5, 1, 1, -1, 5, 1, 0, 1, 5, 1),
        (-1, -1, 5, 1, 1, -1, 5, 1, 0, 1, 5, 1),
        (0, 0, 5, 1, 1, 0, 5, 1, 0, 1, 5, 1),             # ray through a vertex
        (0, 0, 0, 1, 1, 0, 0, 1, 0, 1, 0, 1),             # t = 0
        (0, -1, 5, 1, 0, 1, 5, 1, 1, 0, 5, 1),            # ray through an edge
        (0, -1, 5, 1, 0, 1, 5, 1, -1, 0, 5, 1),           # the neighbour sharing that edge
        (0, 0, 5, 1, 1, 1, 5, 1, 2, 2, 5, 1),             # degenerate (collinear)
        (-2, -2, 9, 2, 2, -2, 9, 2, 0, 2, 9, 2),          # w != 1
        (-1, -1, 5, 1, 1, -1, 5, 1, 0, 1, 5, 1),          # behind
        (-1, -1, 5, 1, 1, -1, 5, 1, 0, 1, 5, 1),
        (-1, -1, 0, 1, 1, -1, 1e-7, 1, 0, 1, 0, 1),       # nearly parallel
        (-1, -1, 5, 1, 1, -1, 5, 1, 0, 1, 5, 1),
        (-1, 3, -1, 1, 1, 3, -1, 1, 0, 3, 1, 1),
        (0, 0, 2, 1, 1, 0, 2, 1, 0, 1, 2, 1),
        (-50, -50, 600, 1, 50, -50, 600, 1, 0, 80, 640, 1),
    ]
    bmin = [(-1, -1, 4), (-1, 0, 4), (0, -1, 4), (0, 0, 4), (1, 1, 1), (-1, -1, -5), (-1, -1, 4), (0.5, 0.5, 4),
            (-1, -1, 4), (-1, -1, 4), (-1, -1, 4), (0, -1, 4), (0, 2, 3), (0, 3, -1), (0.5, 0.5, 1), (-50, -50, 600)]
    bmax = [(1, 1, 6), (1, 1, 6), (1, 1, 6), (1, 1, 6), (2, 2, 2), (1, 1, -4), (0, 1, 6), (1, 1, 6),
            (1, 1, 6), (1, 1, 6), (1, 1, 6), (0, 1, 6), (1, 4, 3), (1, 3, 1), (0.5, 0.5, 3), (50, 80, 640)]
    O = np.array(o, f32); D = np.array(d, f32); T = np.array(tri, f32)
    B0 = np.array(bmin, f32); B1 = np.array(bmax, f32)
    # random: integer-valued primary-like directions and arbitrary shadow-like rays
    Dr = np.concatenate([np.stack([rng.integers(-960, 960, n_rand // 2), rng.integers(-540, 540, n_rand // 2),
                                   np.full(n_rand // 2, 1080)], 1).astype(f32),
                         rng.normal(0, 200, (n_rand - n_rand // 2, 3)).astype(f32)])
    Or = np.concatenate([np.zeros((n_rand // 2, 3), f32), rng.normal(0, 30, (n_rand - n_rand // 2, 3)).astype(f32)])
    c = rng.normal(0, 20, (n_rand, 1, 3)).astype(f32) + np.array([0, 0, 60], f32)
    P = (c + rng.normal(0, 6, (n_rand, 3, 3)).astype(f32))
    Tr = np.concatenate([P, np.ones((n_rand, 3, 1), f32)], 2).reshape(n_rand, 12)
    lo = (c[:, 0] - np.abs(rng.normal(0, 8, (n_rand, 3)))).astype(f32)
    hi = (c[:, 0] + np.abs(rng.normal(0, 8, (n_rand, 3)))).astype(f32)
    # aim half of the random rays at their triangle / box so hits are common
    aim = rng.random(n_rand) < 0.6
    target = (P.mean(axis=1) + rng.normal(0, 2, (n_rand, 3))).astype(f32)
    Dr[aim] = (target - Or)[aim]
    return (np.ascontiguousarray(np.concatenate([O, Or])), np.ascontiguousarray(np.concatenate([D, Dr])),
            np.ascontiguousarray(np.concatenate([T, Tr])), np.ascontiguousarray(np.concatenate([B0, lo])),
            np.ascontiguousarray(np.concatenate([B1, hi])))


def make_leaf_kat():
    lib = pyoracle.ref_lib()
    rng = np.random.default_rng(20240607)
    O, D, T, B0, B1 = leaf_cases(rng)
    n = len(O)
    t = np.zeros(n, f32)
    lib.ref_leaf_moller_trumbore(n, fp(O), fp(D), fp(T), fp(t))
    slab = np.zeros(n, np.uint8)
    lib.ref_leaf_slab(n, fp(O), fp(D), fp(B0), fp(B1), slab.ctypes.data_as(C.POINTER(C.c_uint8)))
    slab0 = np.zeros(n, np.uint8)
    lib.ref_leaf_slab_origin0(n, fp(D), fp(B0), fp(B1), slab0.ctypes.data_as(C.POINTER(C.c_uint8)))
    # barycentrics / normal / phong at the hit points of the rays that hit (others at an arbitrary point)
    tt = np.where(np.isfinite(t), t, f32(0.5)).astype(f32)
    Pt = np.ascontiguousarray((O + tt[:, None] * D).astype(f32))
    bary = np.zeros((n, 3), f32)
    lib.ref_leaf_barycentric(n, fp(T), fp(Pt), fp(bary))
    normal = np.zeros((n, 3), f32)
    lib.ref_leaf_normal(n, fp(T), fp(normal))
    light = np.ascontiguousarray((rng.normal(0, 300, (n, 3)) + np.array([500, -300, -200])).astype(f32))
    color = np.ascontiguousarray(rng.random((n, 3)).astype(f32))
    props = np.ascontiguousarray(np.stack([np.full(n, 0.2), rng.choice([0.0, 0.5], n), rng.choice([15.0, 1.0, 7.5], n)], 1).astype(f32))
    phong = np.zeros((n, 3), f32)
    lib.ref_leaf_phong(n, fp(T), fp(D), fp(light), fp(color), fp(props), fp(tt), fp(phong))
    tone_in = np.ascontiguousarray(np.concatenate([np.abs(phong), rng.random((n, 3)).astype(f32) * 8,
                                                   np.array([[0, 0, 0], [1e-30, 0.5, 1e9]], f32)]).astype(f32))
    m = len(tone_in)
    tone = np.zeros((m, 3), f32)
    q = np.zeros((m, 3), np.int32)
    lib.ref_leaf_tonemap(m, fp(tone_in), fp(tone), q.ctypes.data_as(C.POINTER(C.c_int32)))
    np.savez_compressed(HERE / "leaf_kat.npz", o=O, d=D, tri=T, bmin=B0, bmax=B1, t=t, slab=slab, slab0=slab0, point=Pt,
                        bary=bary, normal=normal, light=light, color=color, props=props, tdist=tt, phong=phong,
                        tone_in=tone_in, tone=tone, q=q)
    print("leaf_kat.npz:", n, "cases;", int(np.isfinite(t).sum()), "triangle hits,", int(slab.sum()), "box hits")


def make_smooth_normal_kat():
    """leaf_smooth_normal.npz: the reference's interpolateNormal :132-140 at calculateBarycentricCoords :79-117 of the
    committed leaf cases' points, with seeded vertex normals (unit, unnormalised, zero and degenerate ones)."""
    lib = pyoracle.ref_lib()
    kat = np.load(HERE / "leaf_kat.npz")
    T, Pt = np.ascontiguousarray(kat["tri"]), np.ascontiguousarray(kat["point"])
    n = len(T)
    rng = np.random.default_rng(20240611)
    N = rng.normal(0, 1, (n, 3, 3))
    unit = rng.random(n) < 0.7
    N[unit] /= np.linalg.norm(N[unit], axis=2, keepdims=True)
    N[rng.random(n) < 0.02] = 0.0                       # a mesh without vn lines: normalize(0) = NaN, like the reference
    same = rng.random(n) < 0.05
    N[same, 1] = N[same, 0]; N[same, 2] = N[same, 0]    # flat-shaded export: three equal normals
    N = np.ascontiguousarray(N.reshape(n, 9).astype(f32))
    out = np.zeros((n, 3), f32)
    lib.ref_leaf_smooth_normal(n, fp(T), fp(N), fp(Pt), fp(out))
    np.savez_compressed(HERE / "leaf_smooth_normal.npz", tri=T, point=Pt, normals=N, out=out)
    print("leaf_smooth_normal.npz:", n, "cases;", int(np.isfinite(out).all(axis=1).sum()), "finite")


# (key, width, height, focal, lights, embed the flattened scene?)
FRAMES = [
    ("C1", 96, 96, 75.0, 1, True),
    ("C1b", 96, 96, 75.0, 1, True),
    ("cubes4", 90, 60, 60.0, 4, True),
    ("spheres", 90, 60, 60.0, 8, True),
    ("shipped", 120, 80, 80.0, 1, False),
    ("C2", 128, 72, 72.0, 1, False),
    ("C3", 128, 72, 72.0, 16, False),
    ("C4", 128, 72, 72.0, 2, False),
]


def make_frames():
    for key, W, H, focal, N, embed in FRAMES:
        cfg = scenes.CONFIGS[key]
        b = pyoracle.RefBackend()
        view = cfg.build(b, 0.0)
        light = scenes.light_view_space(b, view) if view is not None else np.array(scenes.LIGHT_WORLD, f32)
        light = np.asarray(light, f32)
        r = b.render(W, H, focal, light, N, cfg.use_bvh, faithful=True, threads=0)
        out = dict(rgb=r["rgb"], obj_id=r["obj_id"], tri_id=r["tri_id"], t=r["t"],
                   params=np.array([W, H, focal, N, 1 if cfg.use_bvh else 0], np.float64), light=light[:3])
        if embed:
            a = b.flatten(with_flat_lists=not cfg.use_bvh)
            out.update(nodes=a.nodes, geom=a.geom, attr=a.attr, objects=a.objects)
            if a.flat_geom is not None:
                out.update(flat_geom=a.flat_geom, flat_attr=a.flat_attr)
            out["object_names"] = np.array(a.object_names)
        np.savez_compressed(HERE / f"frame_{key}.npz", **out)
        print(f"frame_{key}.npz: {W}x{H} N={N} hit pixels {int((r['obj_id'] >= 0).sum())}")
        b.close()


# Camera-orbit views of the C3 / C5 scene (the views bench.py times; the image series of configs[4] is this orbit at
# 1.5 degrees per frame, simple_raytracer.cpp:534 with SURVEY.md C5's step): series frames 24, 120 and 239, the quarter
# turns, and the two heaviest views of the orbit (most shadow rays cross the bunny, profiles/r01_orbit_cost.txt).
ORBIT_ANGLES = [36.0, 90.0, 135.0, 180.0, 200.0, 270.0, 358.5]


def angle_tag(angle):
    return ("%g" % angle).replace(".", "p")


def make_orbit_frames():
    W, H, focal, N = 320, 180, 180.0, 16
    cfg = scenes.CONFIGS["C3"]
    for angle in ORBIT_ANGLES:
        b = pyoracle.RefBackend()
        view = cfg.build(b, angle)
        light = np.asarray(scenes.light_view_space(b, view), f32)
        r = b.render(W, H, focal, light, N, cfg.use_bvh, faithful=True, threads=0)
        np.savez_compressed(HERE / f"frame_C3_a{angle_tag(angle)}.npz", rgb=r["rgb"], obj_id=r["obj_id"], tri_id=r["tri_id"], t=r["t"],
                            params=np.array([W, H, focal, N, 1], np.float64), light=light[:3], angle=np.float64(angle))
        print(f"frame_C3_a{angle_tag(angle)}.npz: {W}x{H} N={N} angle {angle}: hit pixels {int((r['obj_id'] >= 0).sum())}, "
              f"bunny pixels {int((r['obj_id'] == 0).sum())}")
        b.close()


if __name__ == "__main__":
    import sys
    if "--smooth-normal-only" in sys.argv:
        make_smooth_normal_kat()
    elif "--orbit-only" in sys.argv:
        make_orbit_frames()
    else:
        make_orbit_frames()
        make_leaf_kat()
        make_smooth_normal_kat()
        make_frames()
