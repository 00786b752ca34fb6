"""Workload scenes (BASELINE.json configs), built through the host mirror of the reference API.

* ``reflect_refract_scene()``  — the chapter-11 "reflection and refraction" room that the
  reference ships as the data file ``jamis.yml`` (13 shapes: 6 planes + 7 spheres, one light;
  values per ``jamis.yml:15-183``).  Configs C1, C2, C3, C5.
* ``bench_scene()``            — the criterion bench's scene (``benches/render.rs:11-67``).
* ``synthetic_scene()``        — C4: 9,992 random spheres + 8 planes, seed 0xB200.
* ``default_world()``          — ``World::default()`` (``src/shape.rs:784-795``).
* ``cube_scene()``             — a room of cubes (``src/shape.rs:500-630``), §8 f1.
"""
import math
import random as _random

from .camera import Camera
from .scene import (CheckerPattern, Color, Cube, GradientPattern, GridPattern, Light, Material, Plane, RingPattern, Sphere,
                    StripePattern, TestPattern, World)
from .transform import Matrix

_I = Matrix.identity


def default_world():
    return World.default()


def reflect_refract_camera(hsize=400, vsize=200, samples=1):
    cam = Camera.new_with_transform(
        hsize, vsize, 1.152,
        Matrix.make_view_transform((-2.6, 1.5, -3.9), (-0.6, 1.0, -0.8), (0.0, 1.0, 0.0)))
    cam.set_samples(samples)
    return cam


def reflect_refract_scene():
    """-> World.  Shape order = file order: floor, ceiling, west, east, north, south walls,
    4 background balls, red ball, blue glass ball, green glass ball."""
    w = World(Light(Color(1, 1, 1), (-4.9, 4.9, -1.0)))

    def wall_material():
        p = StripePattern(Color(0.45, 0.45, 0.45), Color(0.55, 0.55, 0.55))
        p.set_transform(_I().scaling(0.25, 0.25, 0.25).rotation_y(1.5708))
        return Material(ambient=0.0, diffuse=0.4, specular=0.0, reflectiveness=0.3, pattern=p)

    floor_m = Material(specular=0.0, reflectiveness=0.4,
                       pattern=CheckerPattern(Color(0.35, 0.35, 0.35), Color(0.65, 0.65, 0.65)))
    w.add_shape(Plane(_I().rotation_y(0.31415), floor_m))
    w.add_shape(Plane(_I().translation(0, 5, 0), Material(color=Color(0.8, 0.8, 0.8), ambient=0.3, specular=0.0)))
    w.add_shape(Plane(_I().rotation_y(1.5708).rotation_z(1.5708).translation(-5, 0, 0), wall_material()))
    w.add_shape(Plane(_I().rotation_y(1.5708).rotation_z(1.5708).translation(5, 0, 0), wall_material()))
    w.add_shape(Plane(_I().rotation_x(1.5708).translation(0, 0, 5), wall_material()))
    w.add_shape(Plane(_I().rotation_x(1.5708).translation(0, 0, -5), wall_material()))
    # background balls
    w.add_shape(Sphere(_I().scaling(0.4, 0.4, 0.4).translation(4.6, 0.4, 1), Material(color=Color(0.8, 0.5, 0.3), shininess=50)))
    w.add_shape(Sphere(_I().scaling(0.3, 0.3, 0.3).translation(4.7, 0.3, 0.4), Material(color=Color(0.9, 0.4, 0.5), shininess=50)))
    w.add_shape(Sphere(_I().scaling(0.5, 0.5, 0.5).translation(-1, 0.5, 4.5), Material(color=Color(0.4, 0.9, 0.6), shininess=50)))
    w.add_shape(Sphere(_I().scaling(0.3, 0.3, 0.3).translation(-1.7, 0.3, 4.7), Material(color=Color(0.4, 0.6, 0.9), shininess=50)))
    # foreground balls
    w.add_shape(Sphere(_I().translation(-0.6, 1, 0.6), Material(color=Color(1, 0.3, 0.2), specular=0.4, shininess=5)))
    glass = dict(ambient=0.0, diffuse=0.4, specular=0.9, shininess=300, reflectiveness=0.9, transparency=0.9, refractive_index=1.5)
    w.add_shape(Sphere(_I().scaling(0.7, 0.7, 0.7).translation(0.6, 0.7, -0.6), Material(color=Color(0, 0, 0.2), **glass)))
    w.add_shape(Sphere(_I().scaling(0.5, 0.5, 0.5).translation(-0.7, 0.5, -0.8), Material(color=Color(0, 0.2, 0), **glass)))
    return w


def bench_scene():
    """-> (World, Camera): three spheres on a checkered plane, 400x300 (benches/render.rs:11-67)."""
    w = World(Light())
    w.add_shape(Sphere(_I().translation(-0.5, 1., 0.5),
                       Material(color=Color.RED, diffuse=0.1, transparency=1.0, refractive_index=1.15,
                                specular=0.1, ambient=0.1)))
    w.add_shape(Sphere(_I().scaling(0.5, 0.5, 0.5).translation(1., 0.7, -3.5),
                       Material(color=Color.GREEN, diffuse=0.1, transparency=1.0, ambient=0.1,
                                refractive_index=1.5, specular=0.1)))
    w.add_shape(Sphere(_I().scaling(0.8, 0.8, 0.8).translation(-2.5, 0.53, -0.75).rotation_x(math.pi / 4.0),
                       Material(color=Color.BLUE, diffuse=0.7, specular=0.3)))
    w.add_shape(Plane(_I().translation(0., -3., 0.),
                      Material(color=Color.RED, diffuse=0.2, ambient=0.6, specular=0.3,
                               pattern=CheckerPattern(Color.WHITE, Color.BLACK))))
    cam = Camera.new_with_transform(400, 300, math.pi / 3.0,
                                    Matrix.make_view_transform((0., 7., 0.), (0., 0., 0.), (1., 0., 0.)))
    return w, cam


# ---- C4: synthetic 10k-shape scene ---------------------------------------------------------
_MASK = (1 << 64) - 1


def _splitmix64(x):
    x = (x + 0x9E3779B97F4A7C15) & _MASK
    z = x
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & _MASK
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & _MASK
    return z ^ (z >> 31)


def _u01(seed, i, k):
    """Counter-based uniform in [0,1): keyed by (shape index, draw index)."""
    return (_splitmix64(seed ^ _splitmix64(i * 64 + k)) >> 11) * (1.0 / 9007199254740992.0)


def synthetic_scene(n_spheres=9992, seed=0xB200):
    """-> World of ``n_spheres`` random spheres + 8 planes.

    Spheres: centre in [-20,20]x[0.4,12]x[-20,20], radius in [0.05,0.35]; every 10th is also
    non-uniformly scaled and rotated (exercises the stored inverse-transpose).  Materials: 70 %
    diffuse, 20 % reflective 0.3-0.9, 10 % glass.  Planes: checker floor, a thin transparent
    "water" sheet, three striped planes, a tilted sub-floor and two tilted mirrors on the left; the
    upper right of the view is open, so rays escaping there are genuine misses.  world_id wraps at 256 like the reference's release build.
    """
    w = World(Light(Color(1, 1, 1), (-10.0, 30.0, -10.0)))
    w.add_shape(Plane(_I(), Material(specular=0.0, reflectiveness=0.2,
                                     pattern=CheckerPattern(Color(0.3, 0.3, 0.3), Color(0.7, 0.7, 0.7)))))
    w.add_shape(Plane(_I().translation(0, 0.2, 0),
                      Material(color=Color(0.05, 0.1, 0.15), ambient=0.0, diffuse=0.2, specular=0.8, shininess=200,
                               reflectiveness=0.3, transparency=0.8, refractive_index=1.33)))

    def wall():
        p = StripePattern(Color(0.4, 0.45, 0.5), Color(0.6, 0.55, 0.5))
        p.set_transform(_I().scaling(0.5, 0.5, 0.5))
        return Material(ambient=0.05, diffuse=0.5, specular=0.0, reflectiveness=0.1, pattern=p)

    # The planes leave the upper right of the view open (no ceiling, no right or far wall, mirrors on the
    # left), so rays escaping there are genuine misses for the hit/miss mask.
    w.add_shape(Plane(_I().rotation_z(math.pi / 2).translation(-25, 0, 0), wall()))              # left wall
    w.add_shape(Plane(_I().rotation_x(math.pi / 2).translation(0, 0, -40), wall()))              # wall behind the camera
    w.add_shape(Plane(_I().rotation_x(0.06).translation(0, -6, 60), wall()))                      # ground rising gently behind the field
    mirror = dict(color=Color(0.05, 0.05, 0.05), ambient=0.0, diffuse=0.1, specular=0.9, shininess=300, reflectiveness=0.9)
    w.add_shape(Plane(_I().rotation_z(math.pi / 2).rotation_y(-0.5).translation(-22, 0, 5), Material(**mirror)))
    w.add_shape(Plane(_I().rotation_z(math.pi / 2).rotation_y(-0.3).rotation_z(0.15).translation(-24, 0, 18), Material(**mirror)))
    w.add_shape(Plane(_I().rotation_z(0.08).translation(0, -0.3, 0),
                      Material(color=Color(0.3, 0.25, 0.2), specular=0.0, diffuse=0.7)))               # slightly tilted sub-floor
    for i in range(n_spheres):
        u = [_u01(seed, i, k) for k in range(16)]
        cx, cy, cz = -20 + 40 * u[0], 0.4 + 11.6 * u[1], -20 + 40 * u[2]
        r = 0.05 + 0.30 * u[3]
        m = _I().scaling(r, r, r)
        if i % 10 == 9:
            m = _I().scaling(r * (0.6 + 0.8 * u[4]), r * (0.6 + 0.8 * u[5]), r * (0.6 + 0.8 * u[6]))
            m = m.rotation_x(6.283 * u[7]).rotation_y(6.283 * u[8]).rotation_z(6.283 * u[9])
        m = m.translation(cx, cy, cz)
        col = Color(0.15 + 0.8 * u[10], 0.15 + 0.8 * u[11], 0.15 + 0.8 * u[12])
        kind = u[13]
        if kind < 0.7:
            mat = Material(color=col, shininess=10 + 190 * u[14])
        elif kind < 0.9:
            mat = Material(color=col, diffuse=0.6, reflectiveness=0.3 + 0.6 * u[14], shininess=100)
        else:
            mat = Material(color=Color(col.red * 0.2, col.green * 0.2, col.blue * 0.2), ambient=0.0, diffuse=0.4,
                           specular=0.9, shininess=300, reflectiveness=0.9, transparency=0.9, refractive_index=1.5)
        w.add_shape(Sphere(m, mat))
    return w


def synthetic_camera(hsize=3840, vsize=2160, samples=1):
    cam = Camera.new_with_transform(hsize, vsize, math.pi / 3,
                                    Matrix.make_view_transform((0.0, 6.0, -30.0), (0.0, 3.0, 0.0), (0.0, 1.0, 0.0)))
    cam.set_samples(samples)
    return cam


# ---- Cube scenes -----------------------------------------------------------
def cube_scene(seed=7, n_cubes=9, with_spheres=True, hsize=200, vsize=140):
    """-> (World, Camera).  SURVEY.md §8 f1 (Cube, src/shape.rs:500-630).  A room of cubes: axis-aligned, rotated, sheared, scaled thin; opaque, mirror, glass (reflective and
    transparent), patterned; a cube nested in a glass cube; a sphere inside a glass cube; a plane floor."""
    rnd = _random.Random(seed)
    I = Matrix.identity
    w = World(Light(Color(1, 1, 1), (-6, 8, -7)))
    floor = Material(specular=0.0, reflectiveness=0.3, pattern=CheckerPattern(Color(0.3, 0.3, 0.3), Color(0.7, 0.7, 0.7)))
    w.add_shape(Plane(I().translation(0, -1, 0), floor))
    w.add_shape(Cube(I().scaling(12, 8, 12).translation(0, 6.9, 0), Material(color=Color(0.6, 0.7, 0.8), specular=0.1)))  # the room: camera inside a cube
    glass = dict(diffuse=0.2, specular=0.9, shininess=300, reflectiveness=0.7, transparency=0.9, refractive_index=1.5)
    w.add_shape(Cube(I().scaling(1.2, 1.2, 1.2).rotation_y(0.5).translation(-1.5, 0.21, 2), Material(color=Color(0.05, 0.1, 0.05), **glass)))
    w.add_shape(Cube(I().scaling(0.5, 0.5, 0.5).rotation_y(0.5).translation(-1.5, 0.0, 2), Material(color=Color(0.9, 0.2, 0.2))))  # nested
    w.add_shape(Cube(I().scaling(1.0, 0.7, 1.0).rotation_z(0.3).rotation_y(-0.4).translation(2.0, 0.4, 3), Material(color=Color(0.1, 0.05, 0.1), **glass)))
    if with_spheres:
        w.add_shape(Sphere(I().scaling(0.4, 0.4, 0.4).translation(2.0, 0.4, 3), Material(color=Color(0.2, 0.4, 0.9))))  # inside the glass cube
        w.add_shape(Sphere(I().scaling(0.7, 0.7, 0.7).translation(0.3, -0.3, 0.5), Material(color=Color(0.1, 0.1, 0.1), **glass)))
    pats = [StripePattern, CheckerPattern, GradientPattern, RingPattern, GridPattern, TestPattern]
    for k in range(n_cubes):
        m = I().scaling(rnd.uniform(0.2, 0.8), rnd.uniform(0.05, 0.9), rnd.uniform(0.2, 0.8))
        if k % 3 == 1:
            m = m.shearing(rnd.uniform(-.4, .4), 0, 0, rnd.uniform(-.4, .4), 0, 0)
        m = m.rotation_x(rnd.uniform(0, 3)).rotation_y(rnd.uniform(0, 3)).translation(rnd.uniform(-4, 4), rnd.uniform(-0.5, 2.5), rnd.uniform(0, 7))
        mat = Material(color=Color(rnd.random(), rnd.random(), rnd.random()), diffuse=rnd.uniform(0.3, 0.9), specular=rnd.uniform(0, 0.9),
                           shininess=rnd.uniform(5, 200), reflectiveness=rnd.choice([0, 0, 0.4, 0.8]))
        if k % 2 == 0:
            p = pats[k % len(pats)](Color(rnd.random(), rnd.random(), rnd.random()), Color(rnd.random(), rnd.random(), rnd.random()))
            p.set_transform(I().scaling(0.3, 0.3, 0.3).rotation_y(rnd.uniform(0, 3)))
            mat.set_pattern(p)
        w.add_shape(Cube(m, mat))
    cam = Camera.new_with_transform(hsize, vsize, 1.1, Matrix.make_view_transform((0.5, 2.2, -6.5), (0, 0.6, 3), (0, 1, 0)))
    return w, cam
