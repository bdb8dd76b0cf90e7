"""GPU tests of the DEVICE leaf functions (srt_eval_leaf, through the C ABI) against the known-answer vectors the
unmodified reference produced (tests/golden/leaf_kat.npz, generator tests/golden/make_golden.py): slab test (exact
and the kernels' filtered version), Moller-Trumbore, Phong, tone map + gamma + quantisation, and powf."""
import ctypes as C

import numpy as np
import pytest

from conftest import GOLDEN
from simple_raytracer_b200 import capi

pytestmark = pytest.mark.gpu
f32 = np.float32


@pytest.fixture(scope="module")
def kat():
    return np.load(GOLDEN / "leaf_kat.npz")


def tri_edges(tri):
    p = tri.reshape(-1, 3, 4)
    c = (p[:, :, :3] / p[:, :, 3:4]).astype(f32)
    return c[:, 0], (c[:, 1] - c[:, 0]).astype(f32), (c[:, 2] - c[:, 0]).astype(f32)


def same_bits(a, b):
    a = np.ascontiguousarray(a, f32)
    b = np.ascontiguousarray(b, f32)
    return (a.view(np.uint32) == b.view(np.uint32)) | (np.isnan(a) & np.isnan(b)) | ((a == 0) & (b == 0))


def test_slab_exact_and_filtered_equal_reference(gpu_ctx, kat):
    """intersectRayAabbNoOrigin :252-293 and intersectRayAabb :204-248 (origin 0): the exact device test AND the
    filtered test the traversal uses give the reference's decision on every vector (incl. zero / NaN-producing axes)."""
    for o, want in ((kat["o"], kat["slab"]), (np.zeros_like(kat["o"]), kat["slab0"])):
        got = capi.eval_leaf(gpu_ctx, capi.LEAF_SLAB, np.hstack([o, kat["d"], kat["bmin"], kat["bmax"]]))
        assert np.array_equal(got[:, 0] != 0, want != 0)
        assert np.array_equal(got[:, 1] != 0, want != 0)
    assert (got[:, 2] == 0).sum() > len(got) // 2        # the filter does decide most cases by itself


def test_filtered_slab_on_grazing_rays(gpu_ctx):
    """Rays aimed exactly at box corners / edges / faces, axis-parallel rays, zero components: the filtered decision
    must equal the exact one everywhere (its margin sends the uncertain cases to the exact test)."""
    rng = np.random.default_rng(7)
    n = 200000
    bmin = rng.uniform(-50, 50, (n, 3)).astype(f32)
    bmax = (bmin + rng.uniform(0, 30, (n, 3)).astype(f32) * (rng.random((n, 3)) > 0.1)).astype(f32)   # some flat boxes
    o = rng.uniform(-100, 100, (n, 3)).astype(f32)
    o[rng.random(n) < 0.3] = 0                                       # primary rays start at the origin
    corner = np.where(rng.random((n, 3)) < 0.5, bmin, bmax)
    d = (corner - o).astype(f32)                                     # through a corner: two or three slabs tie
    k = rng.integers(0, 4, n)
    d[k == 1, 0] = 0                                                 # axis-parallel
    d[k == 2] *= f32(1e-3)
    mid = ((bmin + bmax) * f32(0.5)).astype(f32)
    d[k == 3] = (mid - o)[k == 3]                                    # through the centre: clear hits
    got = capi.eval_leaf(gpu_ctx, capi.LEAF_SLAB, np.hstack([o, d, bmin, bmax]))
    assert np.array_equal(got[:, 0], got[:, 1])
    assert 0.05 < (got[:, 2] != 0).mean() < 0.9                      # both the filter and the fallback were exercised


def test_moller_trumbore_equals_reference(gpu_ctx, kat):
    p0, e1, e2 = tri_edges(kat["tri"])
    got = capi.eval_leaf(gpu_ctx, capi.LEAF_TRIANGLE, np.hstack([kat["o"], kat["d"], p0, e1, e2]))[:, 0]
    assert same_bits(got, kat["t"]).all()


def test_filtered_moller_trumbore_is_the_plain_function(gpu_ctx, kat):
    """moller_trumbore_filtered (the shadow pass: clear misses decided without 1/det) returns the bits of the plain
    function: on the reference's known-answer vectors, and on 600 k rays aimed AT the edges and vertices of their triangle
    (exactly, and displaced by a few ulp to 1e-3 of its size), at primary-ray and shadow-ray scales, both windings,
    including degenerate and tiny triangles.  Both outcomes of the filter must have been exercised."""
    p0, e1, e2 = tri_edges(kat["tri"])
    got = capi.eval_leaf(gpu_ctx, capi.LEAF_TRIANGLE_FILTERED, np.hstack([kat["o"], kat["d"], p0, e1, e2]))
    assert same_bits(got[:, 0], kat["t"]).all() and same_bits(got[:, 1], kat["t"]).all()
    rng = np.random.default_rng(77)
    n = 600_000
    scale = 10.0 ** rng.uniform(-3, 2.5, (n, 1))
    c = rng.normal(0, 60, (n, 3)) + np.array([0, 0, 200.0])
    P = c[:, None, :] + rng.normal(0, 1, (n, 3, 3)) * scale[:, None, :]
    flat = rng.random(n) < 0.03
    P[flat, 2] = P[flat, 0] + (P[flat, 1] - P[flat, 0]) * rng.random((int(flat.sum()), 1))        # collinear: det ~ 0
    swap = rng.random(n) < 0.5
    P[swap, 1], P[swap, 2] = P[swap, 2].copy(), P[swap, 1].copy()
    w = rng.random((n, 3))
    kind = rng.integers(0, 4, n)
    w[kind == 0, 2] = 0.0                                                # on the edge p0-p1
    w[kind == 1, 0] = 0.0                                                # on the edge p1-p2 (u + v = 1)
    w[kind == 2] = np.eye(3)[rng.integers(0, 3, int((kind == 2).sum()))]   # a vertex
    w /= w.sum(axis=1, keepdims=True)                                    # kind 3: interior
    target = (w[:, :, None] * P).sum(axis=1)
    o = np.where(rng.random((n, 1)) < 0.5, 0.0, rng.normal(0, 80, (n, 3)))
    d = target - o
    eps = 10.0 ** rng.uniform(-8, -3, (n, 1)) * (rng.random((n, 1)) < 0.7)
    d = d + rng.normal(0, 1, (n, 3)) * eps * np.linalg.norm(d, axis=1, keepdims=True)
    behind = rng.random(n) < 0.1
    d[behind] = -d[behind]                                               # t < 0
    P32 = P.astype(f32)
    x = np.hstack([o.astype(f32), d.astype(f32), P32[:, 0], P32[:, 1] - P32[:, 0], P32[:, 2] - P32[:, 0]])
    got = capi.eval_leaf(gpu_ctx, capi.LEAF_TRIANGLE_FILTERED, x)
    assert same_bits(got[:, 0], got[:, 1]).all()
    hit = np.isfinite(got[:, 1])
    tail = got[:, 2] != 0
    assert hit.mean() > 0.15 and (~hit).mean() > 0.3
    assert (tail | ~hit).all()                                           # every hit went through the reference's operations
    assert (~tail & ~hit).mean() > 0.25 and (tail & ~hit).sum() > 1000   # clear misses skipped them, close calls did not


def test_phong_equals_reference(gpu_ctx, kat):
    """phongIllumination :144-200: every operation is IEEE-defined except powf (glibc's is not correctly rounded:
    about 1e-3 of its results are one ulp off), so: bit-identical on >= 99 % of the vectors, one float ulp elsewhere."""
    n = len(kat["normal"])
    white = np.ones((n, 3), f32)
    x = np.hstack([kat["normal"], kat["d"], kat["light"], white, kat["color"], kat["props"], kat["tdist"][:, None]])
    got = capi.eval_leaf(gpu_ctx, capi.LEAF_PHONG, x)
    same = same_bits(got, kat["phong"])
    assert same.mean() > 0.99
    fin = np.isfinite(kat["phong"]) & ~same
    ulp = np.spacing(np.abs(kat["phong"][fin]))
    assert np.all(np.abs(got[fin] - kat["phong"][fin]) <= ulp)


def test_tone_gamma_quantise_equals_reference(gpu_ctx, kat):
    c = kat["tone_in"].reshape(-1)
    x = np.stack([c, np.full_like(c, 0.5), np.full_like(c, 1.1)], axis=1)
    got = capi.eval_leaf(gpu_ctx, capi.LEAF_TONE, x)[:, 0].astype(np.int64) & 0xff
    want = kat["q"].reshape(-1).astype(np.int64) & 0xff
    assert np.abs(got - want).max() <= 1
    assert (got == want).mean() > 0.999


def test_powf_is_correctly_rounded_and_agrees_with_libm(gpu_ctx):
    """srt_powf (fast double-precision path + pow() fallback) against the correctly rounded value (80-bit long double
    pow rounded once) and against the host's powf, the function the reference calls (:196, :398)."""
    rng = np.random.default_rng(3)
    n = 400000
    x = np.concatenate([rng.random(n, dtype=f32),                              # specular base / tone-mapped colour
                        (rng.random(n, dtype=f32) ** 8).astype(f32),
                        np.exp(rng.normal(0, 8, n)).astype(f32)])
    y = np.concatenate([rng.choice(np.array([1.1, 5, 15, 32, 50, 100, 0.5, 2, 3, 1], f32), n),
                        (rng.random(n, dtype=f32) * 300).astype(f32),
                        rng.normal(0, 4, n).astype(f32)])
    got = capi.eval_leaf(gpu_ctx, capi.LEAF_POWF, np.stack([x, y], axis=1))[:, 0]
    with np.errstate(all="ignore"):
        cr = np.power(x.astype(np.longdouble), y.astype(np.longdouble)).astype(f32)
    # (exact ties, e.g. a 12-bit x squared, are the known exceptions: any rounding error breaks a tie)
    assert (same_bits(got, cr)).mean() > 0.99998
    libm = C.CDLL("libm.so.6")
    libm.powf.restype = C.c_float
    libm.powf.argtypes = [C.c_float, C.c_float]
    idx = rng.choice(len(x), 60000, replace=False)
    host = np.array([libm.powf(float(a), float(b)) for a, b in zip(x[idx], y[idx])], f32)
    assert same_bits(got[idx], host).mean() > 0.995
    fin = np.isfinite(host)
    assert same_bits(got[idx][~fin], host[~fin]).all()
    assert np.all(np.abs(got[idx][fin].astype(np.float64) - host[fin]) <= np.spacing(np.abs(host[fin])).astype(np.float64))
    # special cases of C99 Annex F that the shading can produce
    inf, nan = f32(np.inf), f32(np.nan)
    sx = np.array([0, -0.0, 1, 1, 0, inf, 2, 0.5, -1, -2, -2, nan, 1e-45, 3e38, 1e-40, 0, 1], f32)
    sy = np.array([32, 32, nan, inf, 0, 2, inf, inf, 2, 3, 0.5, 1, 2, 2, 0.5, -1, 0], f32)
    sg = capi.eval_leaf(gpu_ctx, capi.LEAF_POWF, np.stack([sx, sy], axis=1))[:, 0]
    sh = np.array([libm.powf(float(a), float(b)) for a, b in zip(sx, sy)], f32)
    assert same_bits(sg, sh).all(), (sg, sh)


def test_smooth_normal_equals_reference(gpu_ctx):
    """interpolateNormal :132-140 at calculateBarycentricCoords :79-117 on the device against the reference's own two
    functions (tests/golden/leaf_smooth_normal.npz): bit-identical, NaN for a mesh without normals like the reference."""
    g = np.load(GOLDEN / "leaf_smooth_normal.npz")
    p0, e1, e2 = tri_edges(g["tri"])
    got = capi.eval_leaf(gpu_ctx, capi.LEAF_SMOOTH_NORMAL, np.hstack([p0, e1, e2, g["normals"], g["point"]]))
    assert same_bits(got, g["out"]).all()
