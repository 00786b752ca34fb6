"""Host-side mirror of the reference's scene API — ``Color``, ``Light``, ``Material``, the
patterns, ``Sphere``/``Plane``/``Cube`` and ``World`` — with the same names, argument meaning
and error behaviour as ``src/color.rs``, ``src/material.rs`` and ``src/shape.rs``.

This layer stays on the host, as in the reference; it does no ray tracing.  Its one job for
the hot path is ``World.flatten()``: the ``rtc_shape[]`` / ``rtc_light`` descriptors the C ABI
takes (what the Rust shim's ``flatten()`` produces).
"""
from . import ffi
from .transform import Matrix


class Color:
    """src/color.rs:5-24"""
    __slots__ = ("red", "green", "blue")

    def __init__(self, r, g, b):
        self.red, self.green, self.blue = float(r), float(g), float(b)

    def __eq__(self, o):
        return isinstance(o, Color) and (self.red, self.green, self.blue) == (o.red, o.green, o.blue)

    def __iter__(self):
        return iter((self.red, self.green, self.blue))

    def __repr__(self):
        return f"Color({self.red}, {self.green}, {self.blue})"

    @staticmethod
    def color_from_spec(s):
        """``#rrggbb`` -> Color (src/color.rs:26-41)."""
        if len(s) != 7 or not s.startswith("#"):
            raise ValueError("invalid color representaion")
        return Color(int(s[1:3], 16) / 255.0, int(s[3:5], 16) / 255.0, int(s[5:7], 16) / 255.0)


Color.BLACK = Color(0.0, 0.0, 0.0)
Color.WHITE = Color(1.0, 1.0, 1.0)
Color.RED = Color(1.0, 0.0, 0.0)
Color.GREEN = Color(0.0, 1.0, 0.0)
Color.BLUE = Color(0.0, 0.0, 1.0)


def _color(c):
    return c if isinstance(c, Color) else Color(*c)


class Light:
    """src/material.rs:10-31 — a single point light; default white at (-10, 10, -10)."""

    def __init__(self, intensity=None, position=(-10., 10., -10.)):
        self.intensity = _color(intensity) if intensity is not None else Color.WHITE
        self.position = tuple(float(v) for v in position)


class Pattern:
    """src/material.rs:33-46.  ``set_transform`` stores the inverse, like the reference."""
    kind = ffi.PATTERN_NONE

    def __init__(self, a=None, b=None):
        self.xf_inv = Matrix.identity()
        self.color_a = _color(a) if a is not None else Color.BLACK
        self.color_b = _color(b) if b is not None else Color.BLACK

    def set_transform(self, xf):
        self.xf_inv = xf.inverse()
        return self


class TestPattern(Pattern):        # src/material.rs:48-70
    kind = ffi.PATTERN_TEST
    __test__ = False


class StripePattern(Pattern):      # :72-104
    kind = ffi.PATTERN_STRIPE


class GradientPattern(Pattern):    # :106-136
    kind = ffi.PATTERN_GRADIENT


class RingPattern(Pattern):        # :138-171
    kind = ffi.PATTERN_RING


class CheckerPattern(Pattern):     # :173-206
    kind = ffi.PATTERN_CHECKER


class GridPattern(Pattern):        # :208-242  (a = base colour, b = grid-line colour)
    kind = ffi.PATTERN_GRID


class Material:
    """src/material.rs:244-369.  ``Material()`` is ``Default`` (white); ``Material.DEFAULT()`` is
    the const with RED (``:273-283``)."""

    def __init__(self, color=Color.WHITE, ambient=0.1, diffuse=0.9, specular=0.9, shininess=200.0,
                 reflectiveness=0.0, transparency=0.0, refractive_index=1.0, pattern=None):
        self.pattern = pattern
        self.color = _color(color) if color is not None else None
        self.ambient, self.diffuse, self.specular, self.shininess = ambient, diffuse, specular, shininess
        self.reflectiveness, self.transparency, self.refractive_index = reflectiveness, transparency, refractive_index

    @staticmethod
    def DEFAULT():
        return Material(color=Color.RED)

    @staticmethod
    def solid(color, ambient, diffuse, specular, shininess, reflectiveness, transparency, refractive_index):
        return Material(color, ambient, diffuse, specular, shininess, reflectiveness, transparency, refractive_index)

    @staticmethod
    def solid_with_defaults(color):
        return Material(color=color)

    @staticmethod
    def pattern_with_defaults(pattern):
        return Material(color=Color.RED, pattern=pattern)

    def set_pattern(self, pattern):
        self.pattern = pattern
        return self

    def copy(self):
        m = Material()
        m.__dict__.update(self.__dict__)
        return m


class Shape:
    """src/shape.rs:17-49.  ``transform`` holds the INVERSE and ``transform_transpose`` its
    transpose, exactly like the reference's fields (``:284-285``, ``:300-301``)."""
    kind = None

    def __init__(self, m=None, material=None):
        m = m if m is not None else Matrix.identity()
        self.transform = m.inverse()
        self.transform_transpose = m.inverse().transpose()
        self.material = material if material is not None else Material()
        self.world_id = 0

    @classmethod
    def new(cls):
        return cls()

    @classmethod
    def new_with_transform(cls, m):
        return cls(m)

    @classmethod
    def new_with_transform_and_material(cls, m, mat):
        return cls(m, mat)

    def get_material_mut(self):
        return self.material

    def get_material(self):
        return self.material


class Sphere(Shape):
    kind = ffi.SHAPE_SPHERE

    def set_transform(self, m):  # src/shape.rs:319-322
        self.transform = m.inverse()
        self.transform_transpose = self.transform.transpose()

    @staticmethod
    def glass_sphere():  # :324-329
        s = Sphere()
        s.material.transparency = 1.0
        s.material.refractive_index = 1.5
        return s


class Plane(Shape):
    kind = ffi.SHAPE_PLANE

    def set_transform(self, m):  # src/shape.rs:446-449 — leaves the inverse-transpose stale (kept)
        self.transform = m.inverse()
        self.transform_transpose = self.transform_transpose.transpose()


class Cube(Shape):
    kind = ffi.SHAPE_CUBE

    def set_transform(self, m):  # src/shape.rs:535-538 — same stale transpose
        self.transform = m.inverse()
        self.transform_transpose = self.transform_transpose.transpose()


class World:
    """src/shape.rs:633-683 (container part).  The shading methods are the GPU path."""

    def __init__(self, light=None):
        self.shapes = []
        self.light = light if light is not None else Light()
        self.extra_lights = []   # lights 1.. of a multi-light world (rtc_scene_upload_lights; the reference keeps one)
        self.last_world_id = 0

    @staticmethod
    def new(light):
        return World(light)

    @staticmethod
    def default():  # :784-795
        w = World(Light())
        m = Material.solid_with_defaults(Color(0.8, 1.0, 0.6))
        m.diffuse = 0.7
        m.specular = 0.2
        w.add_shape(Sphere.new_with_transform_and_material(Matrix.identity(), m))
        w.add_shape(Sphere.new_with_transform(Matrix.identity().scaling(0.5, 0.5, 0.5)))
        return w

    @staticmethod
    def with_(func):  # World::with, :646-650
        w = World.default()
        func(w)
        return w

    def add_shape(self, s):  # :661-667; u8 id, wraps like the release build
        new_id = (self.last_world_id + 1) & 0xFF
        s.world_id = new_id
        self.last_world_id = new_id
        self.shapes.append(s)
        return self

    def add_light(self, light):
        """A further point light: the 'FIXME -- multiple lights' of World::shade_hit (src/shape.rs:686).  Each light adds
        lighting(light, .., is_shadowed_by_light(over_point, light)) to the surface colour (include/rtc.h)."""
        if 1 + len(self.extra_lights) >= ffi.MAX_LIGHTS:
            raise ValueError("at most %d lights" % ffi.MAX_LIGHTS)
        self.extra_lights.append(light)
        return self

    def flatten_lights(self):
        """-> (RtcLight * n_lights): World.light first, then the lights added with add_light."""
        ls = [self.light] + list(self.extra_lights)
        arr = (ffi.RtcLight * len(ls))()
        for d, l in zip(arr, ls):
            d.intensity[:] = list(l.intensity)
            d.position[:] = list(l.position)
        return arr

    def get_shape(self, i):
        return self.shapes[i]

    get_shape_mut = get_shape

    def flatten(self):
        """-> (RtcLight, (RtcShape * n), n): the flat descriptors of include/rtc.h."""
        light = ffi.RtcLight()
        light.intensity[:] = list(self.light.intensity)
        light.position[:] = list(self.light.position)
        n = len(self.shapes)
        arr = (ffi.RtcShape * max(n, 1))()
        for d, s in zip(arr, self.shapes):
            mat = s.material
            d.kind = s.kind
            d.world_id = s.world_id
            d.inv[:] = s.transform.to_row_major()
            d.inv_t[:] = s.transform_transpose.to_row_major()
            d.has_color = 1 if mat.color is not None else 0
            d.color[:] = list(mat.color) if mat.color is not None else [0.0, 0.0, 0.0]
            d.ambient, d.diffuse, d.specular, d.shininess = mat.ambient, mat.diffuse, mat.specular, mat.shininess
            d.reflective, d.transparency, d.refractive_index = mat.reflectiveness, mat.transparency, mat.refractive_index
            p = mat.pattern
            d.pattern_kind = p.kind if p is not None else ffi.PATTERN_NONE
            d.pat_inv[:] = (p.xf_inv if p is not None else Matrix.identity()).to_row_major()
            d.pat_a[:] = list(p.color_a) if p is not None else [0.0, 0.0, 0.0]
            d.pat_b[:] = list(p.color_b) if p is not None else [0.0, 0.0, 0.0]
        return light, arr, n
