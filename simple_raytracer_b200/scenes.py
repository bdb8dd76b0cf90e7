"""The BASELINE.json configs as scene scripts.

Each script drives a *scene backend* through the reference's own scene surface, call for call the
way the reference's main() does (simple_raytracer.cpp:542-778).  Two backends implement the same
methods: this repo's host mirror (simple_raytracer_b200.host.HostScene — the product) and the
unmodified reference compiled under oracle/_ref (tests only).  Object names are the strings main()
uses, because they are unordered_map keys and their hash order decides the object order (S2).

Backend protocol (all matrices are 16 floats, column-major like glm):
    load_obj(path) set_color(name, rgb) set_properties(name, a, s, sh) get_properties(name)
    clone(src, dst, copy_properties) transform(name, m16) create_bounding_hierarchy(name)
    xf_scale xf_rotate_x xf_rotate_y xf_rotate_z xf_translate xf_view mat_inverse mat_vec
    radians cosf sinf
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Callable, Dict, Tuple

import numpy as np

LIGHT_WORLD = (500.0, -300.0, -200.0, 1.0)          # simple_raytracer.cpp:776

f32 = np.float32


@dataclass
class SceneConfig:
    key: str
    description: str
    width: int
    height: int
    focal: float
    light_count: int
    use_bvh: bool
    build: Callable            # build(backend, angle_deg) -> view matrix (16 floats)
    needs_textures: bool = False
    extra: Dict = field(default_factory=dict)


def _orbit_view(b, radius: float, height: float, pitch_deg: float, angle_deg: float):
    """Camera of main(): simple_raytracer.cpp:546-551 (complex scene) / :707-711 (cube)."""
    radians = b.radians(f32(angle_deg))
    cx = f32(f32(radius) * b.cosf(radians))
    cz = f32(f32(radius) * b.sinf(radians))
    rot = (b.radians(f32(pitch_deg)), b.radians(f32(f32(angle_deg) + f32(90.0))), b.radians(f32(0.0)))
    return b.xf_view((cx, f32(height), cz), rot)


def light_view_space(b, view):
    """lightPos = glm::inverse(viewMatrix) * lightPos   (simple_raytracer.cpp:776-778)."""
    return b.mat_vec(b.mat_inverse(view), LIGHT_WORLD)


def _place(b, name, view, *matrices):
    for m in matrices:
        b.transform(name, m)
    b.transform(name, b.mat_inverse(view))            # camera baked into the vertices, :558
    b.create_bounding_hierarchy(name)


def build_cube(b, angle_deg: float = 0.0):
    """C1 — "One Sample Cube for Testing", simple_raytracer.cpp:703-722."""
    view = _orbit_view(b, 100.0, 0.0, 0.0, angle_deg)
    b.load_obj("cube.obj")
    _place(b, "cube.obj", view, b.xf_scale(20.0, 20.0, 20.0), b.xf_rotate_y(b.radians(f32(25.0))))
    return view


def _ground_cube(b, view):
    """simple_raytracer.cpp:553-559."""
    b.load_obj("./obj/cube.obj")
    b.set_color("./obj/cube.obj", (0.0, 1.0, 0.0))
    _place(b, "./obj/cube.obj", view, b.xf_scale(35.0, 35.0, 35.0), b.xf_translate(0.0, 10.0, 0.0))


def _bunny(b, view, scale=150.0, pos=(10.0, -25.0 + 0.033 * 150.0, 0.0)):
    """Template simple_raytracer.cpp:582-591 with scale 50 -> 150 and x 25 -> 10."""
    n = "./obj/stanford-bunny.obj"
    b.load_obj(n)
    b.set_color(n, (0.9, 0.9, 0.9))
    _place(b, n, view, b.xf_scale(scale, scale, scale), b.xf_rotate_x(b.radians(f32(181.0))),
           b.xf_rotate_y(b.radians(f32(90.0))), b.xf_translate(*pos))


def build_bunny(b, angle_deg: float = 0.0):
    """C2 — bunny only, shipped camera (:546-551)."""
    view = _orbit_view(b, 50.0, -50.0, 30.0, angle_deg)
    _bunny(b, view)
    return view


def build_bunny_ground(b, angle_deg: float = 0.0):
    """C3 / C5 — ground cube (:553-559) + bunny; C5 frame f uses angle_deg = 1.5 * f."""
    view = _orbit_view(b, 50.0, -50.0, 30.0, angle_deg)
    _ground_cube(b, view)
    _bunny(b, view)
    return view


def build_shipped_cube_bunny(b, angle_deg: float = 0.0):
    """The part of the shipped main() whose meshes exist here: ground cube + bunny as shipped (:553-559, :582-591)."""
    view = _orbit_view(b, 50.0, -50.0, 30.0, angle_deg)
    _ground_cube(b, view)
    _bunny(b, view, scale=50.0, pos=(25.0, -23.0, 0.0))
    return view


def build_six_spheres(b, angle_deg: float = 0.0):
    """"6 Sphere Triangles transformed", simple_raytracer.cpp:622-673 (no view transform, light not transformed)."""
    b.load_obj("sphere.obj")
    b.transform("sphere.obj", b.xf_translate(0.0, 6.0, 30.0))
    for k, pos in enumerate([(6.0, 0.0, 0.0), (-6.0, 0.0, 0.0), (0.0, -12.0, 0.0), (6.0, -12.0, 0.0), (-6.0, -12.0, 0.0)], 1):
        name = "sphere%d.obj" % k
        b.clone("sphere.obj", name, False)
        b.set_color(name, (1.0, 0.0, 0.0))
        b.transform(name, b.xf_translate(*pos))
    for k in range(6):
        b.create_bounding_hierarchy("sphere.obj" if k == 0 else "sphere%d.obj" % k)
    return None          # identity view: light stays (500,-300,-200)


def build_four_cubes(b, angle_deg: float = 0.0):
    """"Scene with 4 Cubes in different colors", simple_raytracer.cpp:726-769."""
    view = _orbit_view(b, 100.0, 0.0, 0.0, angle_deg)
    b.load_obj("cube.obj")
    b.set_color("cube.obj", (1.0, 1.0, 0.0))
    b.transform("cube.obj", b.xf_scale(10.0, 10.0, 10.0))
    for name, col, pos in (("cube1.obj", (1.0, 0.0, 1.0), (0.0, -15.0, -15.0)),
                           ("cube2.obj", (1.0, 0.0, 0.0), (0.0, -15.0, 15.0)),
                           ("cube3.obj", (0.0, 1.0, 0.0), (0.0, 15.0, 15.0))):
        b.clone("cube.obj", name, False)
        b.set_color(name, col)
        b.transform(name, b.xf_translate(*pos))
    b.transform("cube.obj", b.xf_translate(0.0, 15.0, -15.0))
    inv = b.mat_inverse(view)
    for name in ("cube.obj", "cube1.obj", "cube2.obj", "cube3.obj"):
        b.transform(name, inv)
    for name in ("cube.obj", "cube1.obj", "cube2.obj", "cube3.obj"):
        b.create_bounding_hierarchy(name)
    return view


def _textured(b, path, view, *matrices, clones=()):
    """Textured-object idiom of main(): specular := 0 (:564, :595) and the clone idiom (:597-600)."""
    b.load_obj(path)
    a, _s, sh = b.get_properties(path)
    b.set_properties(path, a, 0.0, sh)
    for dst in clones:
        b.clone(path, dst, True)


def build_textured_multi(b, angle_deg: float = 0.0):
    """C4 — the shipped complex-scene layout (:553-618) with the meshes that exist in this checkout
    (house, 2 trees, grass, 2 horses, bird in place of the missing dog/cat — SURVEY.md §8d, H7)."""
    view = _orbit_view(b, 50.0, -50.0, 30.0, angle_deg)
    _ground_cube(b, view)
    rx = lambda deg: b.xf_rotate_x(b.radians(f32(deg)))
    ry = lambda deg: b.xf_rotate_y(b.radians(f32(deg)))
    s = lambda k: b.xf_scale(k, k, k)
    T = b.xf_translate
    tree = "./obj/tree/tree.obj"
    _textured(b, tree, view, clones=(tree + "1",))
    _place(b, tree, view, s(0.03), rx(-90.0), T(-6.0, -25.0, -25.0))           # :602-606
    _place(b, tree + "1", view, s(0.035), rx(-90.0), T(-6.0, -25.0, 25.0))     # :608-612
    horse = "./obj/horse/horse.obj"
    _textured(b, horse, view, clones=(horse + "1",))
    _place(b, horse, view, s(0.008), rx(-90.0), ry(125.0), T(25.0, -25.0, -14.0))   # cat 0 slot :568-573
    _place(b, horse + "1", view, s(0.008), rx(-90.0), ry(70.0), T(25.0, -25.0, 8.0))  # cat 1 slot :575-580
    bird = "./obj/bird/bird.obj"
    _textured(b, bird, view)
    _place(b, bird, view, s(0.25), rx(-90.0), ry(90.0), T(10.0, -25.0, 0.0))
    grass = "./obj/grass/grass.obj"
    _textured(b, grass, view)
    _place(b, grass, view, s(0.12), rx(-90.0), T(5.0, -25.2, 0.0))
    house = "./obj/house/house.obj"
    _textured(b, house, view)
    _place(b, house, view, s(0.02), rx(180.0), ry(90.0), T(-20.0, -25.0, 0.0))
    return view


CONFIGS: Dict[str, SceneConfig] = {
    "C1": SceneConfig("C1", "cube.obj, single point light, Phong, 512x512, legacy one-box path (no BVH)",
                      512, 512, 400.0, 1, False, build_cube),
    "C1b": SceneConfig("C1b", "cube.obj, single point light, Phong, 512x512, BVH path", 512, 512, 400.0, 1, True, build_cube),
    "C2": SceneConfig("C2", "stanford-bunny.obj, BVH slab-test traversal, Phong, 1920x1080, 1 light",
                      1920, 1080, 1080.0, 1, True, build_bunny),
    "C3": SceneConfig("C3", "stanford-bunny.obj + ground cube, 16-light soft shadows, 1920x1080, focal 1080, BVH",
                      1920, 1080, 1080.0, 16, True, build_bunny_ground),
    "C4": SceneConfig("C4", "textured multi-object scene (house, 2 trees, grass, 2 horses, bird, ground), 3840x2160",
                      3840, 2160, 2160.0, 1, True, build_textured_multi, needs_textures=True),
    "shipped": SceneConfig("shipped", "ground cube + bunny exactly as in the shipped main(), 600x400, focal 400, 1 light",
                           600, 400, 400.0, 1, True, build_shipped_cube_bunny),
    "spheres": SceneConfig("spheres", "6 spheres scene of main() :622-673, 600x400", 600, 400, 400.0, 1, True, build_six_spheres),
    "cubes4": SceneConfig("cubes4", "4 coloured cubes scene of main() :726-769, 600x400", 600, 400, 400.0, 1, True, build_four_cubes),
}

SERIES_FRAMES = 240            # C5: 240 frames at 1080p
SERIES_DEG_PER_FRAME = 1.5


def scaled(cfg: SceneConfig, height: int) -> Tuple[int, int, float]:
    """Same field of view at another resolution: focal scales with height."""
    w = int(round(cfg.width * height / cfg.height))
    return w, height, float(cfg.focal * height / cfg.height)
