"""Scene front-ends: the two on-disk / in-script scene formats on the caller's side
of ``Camera::render``, built on the host mirror of the reference API (``scene.py``, ``camera.py``).

* ``load_yaml(text)`` — the YAML dialect of the book's scene files, the format of the reference's
  data file ``jamis.yml`` (the reference itself ships no loader for it): a list of
  ``add: camera | light | plane | sphere | cube`` items and ``define: name`` constants (``value:``,
  optional ``extend:``); ``transform:`` lists are applied in listed order (``[translate|scale x y z]``,
  ``[rotate-x|rotate-y|rotate-z r]``, ``[shear xy xz yx yz zx zy]``, or the name of a defined list);
  materials are dicts or names; defaults are ``Material::default()`` (src/material.rs:273-283).
* ``world_from_table(world)`` / ``camera_from_table(camera)`` — the table schema the reference's Lua
  scripts pass to ``Render(world, camera, file)`` (src/lua.rs:186-330), as Python dicts: same keys, same
  order of operations (``rotate_x``, ``rotate_y``, ``rotate_z``, uniform ``scale``, then ``position``),
  the first light only (:148-150), ``material`` keys as in :186-229 with the shape-level ``color`` /
  ``pattern`` override (:223-231), pattern types ``checks`` / ``stripes`` / ``grid`` (:107-138).

Errors are ``ValueError`` with the offending key, where the reference returns a ``LuaError``.
"""
import yaml

from .camera import Camera
from .scene import (CheckerPattern, Color, Cube, GradientPattern, GridPattern, Light, Material, Plane, RingPattern, Sphere,
                    StripePattern, World)
from .transform import Matrix

_SHAPES = {"sphere": Sphere, "plane": Plane, "cube": Cube}


# ---- YAML dialect ---------------------------------------------------------------------------------
def _color(v, what):
    if not (isinstance(v, (list, tuple)) and len(v) == 3):
        raise ValueError(f"{what}: a colour is a list of three numbers, got {v!r}")
    return Color(float(v[0]), float(v[1]), float(v[2]))


def _apply_ops(m, ops, defs, what):
    """Post-multiply nothing: every builder of the reference pre-multiplies (src/transform.rs:53-105), so
    chaining the listed operations in order applies them to a point in that order."""
    if isinstance(ops, str):
        if ops not in defs:
            raise ValueError(f"{what}: unknown transform {ops!r}")
        ops = defs[ops]
    for op in ops or []:
        if isinstance(op, str):                       # a defined list used inside a list
            m = _apply_ops(m, op, defs, what)
            continue
        name, args = op[0], [float(a) for a in op[1:]]
        if name == "translate" and len(args) == 3:
            m = m.translation(*args)
        elif name == "scale" and len(args) == 3:
            m = m.scaling(*args)
        elif name in ("rotate-x", "rotate-y", "rotate-z") and len(args) == 1:
            m = getattr(m, "rotation_" + name[-1])(args[0])
        elif name in ("shear", "shearing") and len(args) == 6:
            m = m.shearing(*args)
        else:
            raise ValueError(f"{what}: bad transform step {op!r}")
    return m


_PATTERNS = {"stripes": StripePattern, "checkers": CheckerPattern, "gradient": GradientPattern, "ring": RingPattern, "rings": RingPattern,
             "grid": GridPattern}
_MATERIAL_KEYS = {"ambient": "ambient", "diffuse": "diffuse", "specular": "specular", "shininess": "shininess",
                  "reflective": "reflectiveness", "transparency": "transparency", "refractive-index": "refractive_index"}


def _resolve(value, defs, what):
    """A defined constant: `value` dict merged over its `extend` base (dicts only)."""
    if isinstance(value, str):
        if value not in defs:
            raise ValueError(f"{what}: unknown name {value!r}")
        return defs[value]
    return value


def _material(spec, defs, what):
    spec = _resolve(spec, defs, what) or {}
    m = Material()
    for key, v in spec.items():
        if key == "color":
            m.color = _color(v, what)
        elif key == "pattern":
            kind = v.get("type")
            if kind not in _PATTERNS:
                raise ValueError(f"{what}: unknown pattern type {kind!r}")
            cols = v.get("colors")
            if not (isinstance(cols, (list, tuple)) and len(cols) == 2):
                raise ValueError(f"{what}: a pattern needs two colours")
            p = _PATTERNS[kind](_color(cols[0], what), _color(cols[1], what))
            if v.get("transform") is not None:          # no key: the pattern keeps its identity (no inverse is taken)
                p.set_transform(_apply_ops(Matrix.identity(), v["transform"], defs, what))
            m.set_pattern(p)
        elif key in _MATERIAL_KEYS:
            setattr(m, _MATERIAL_KEYS[key], float(v))
        else:
            raise ValueError(f"{what}: unknown material key {key!r}")
    return m


def load_yaml(text, all_lights=False):
    """-> (World, Camera).  `text` is the YAML document (a string or an open file).  The reference's World holds one
    point light, so by default the first `add: light` counts; all_lights=True keeps every light of the file
    (World.add_light -> rtc_scene_upload_lights, the book's multi-light shade_hit)."""
    doc = yaml.safe_load(text)
    if not isinstance(doc, list):
        raise ValueError("scene file: the top level is a list of `add:` / `define:` items")
    defs, cam, light, shapes, more_lights = {}, None, None, [], []
    for n, item in enumerate(doc):
        what = f"item {n}"
        if "define" in item:
            value = item.get("value")
            if "extend" in item:
                base = _resolve(item["extend"], defs, what)
                if not (isinstance(base, dict) and isinstance(value, dict)):
                    raise ValueError(f"{what}: `extend` needs two mappings")
                value = {**base, **value}
            defs[item["define"]] = value
        elif item.get("add") == "camera":
            cam = Camera.new_with_transform(int(item["width"]), int(item["height"]), float(item["field-of-view"]),
                                            Matrix.make_view_transform(tuple(map(float, item["from"])), tuple(map(float, item["to"])),
                                                                       tuple(map(float, item["up"]))))
        elif item.get("add") == "light":
            if light is None:                          # World holds one point light (src/shape.rs:635): the first one counts
                light = Light(_color(item["intensity"], what), tuple(map(float, item["at"])))
            elif all_lights:
                more_lights.append(Light(_color(item["intensity"], what), tuple(map(float, item["at"]))))
        elif item.get("add") in _SHAPES:
            shapes.append(_SHAPES[item["add"]](_apply_ops(Matrix.identity(), item.get("transform"), defs, what),
                                               _material(item.get("material"), defs, what)))
        else:
            raise ValueError(f"{what}: expected `define:` or `add: camera|light|plane|sphere|cube`, got {item!r}")
    if cam is None or light is None:
        raise ValueError("scene file: needs one `add: camera` and one `add: light`")
    world = World(light)
    for l in more_lights:
        world.add_light(l)
    for s in shapes:
        world.add_shape(s)
    return world, cam


# ---- Lua-table schema (src/lua.rs:186-330) ----------------------------------------------------------
def _xyz(t, what):
    try:
        return (float(t["x"]), float(t["y"]), float(t["z"]))
    except (KeyError, TypeError):
        raise ValueError(f"{what}: expected a table with x, y, z") from None


def _rgb(t, what):
    try:
        return Color(float(t["r"]), float(t["g"]), float(t["b"]))
    except (KeyError, TypeError):
        raise ValueError(f"{what}: expected a table with r, g, b") from None


def _transform_from_table(t):   # src/lua.rs:264-300
    m = Matrix.identity()
    for axis in ("x", "y", "z"):
        if "rotate_" + axis in t:
            m = getattr(m, "rotation_" + axis)(float(t["rotate_" + axis]))
    if "scale" in t:
        s = float(t["scale"])
        m = m.scaling(s, s, s)
    if t.get("position") is not None:
        m = m.translation(*_xyz(t["position"], "position"))
    return m


def _pattern_from_table(t):     # src/lua.rs:107-138
    m = _transform_from_table(t)
    kind = t.get("type")
    if kind == "checks":
        p = CheckerPattern(_rgb(t["color_a"], "color_a"), _rgb(t["color_b"], "color_b"))
    elif kind == "stripes":
        p = StripePattern(_rgb(t["color_a"], "color_a"), _rgb(t["color_b"], "color_b"))
    elif kind == "grid":
        p = GridPattern(Color(1., 1., 1.), Color(0., 0., 0.))
    else:
        raise ValueError(f"invalid pattern type: {kind}")
    p.set_transform(m)
    return p


_TABLE_MATERIAL_KEYS = ("ambient", "diffuse", "specular", "shininess", "reflectiveness", "transparency", "refractive_index")


def _material_from_table(t):    # src/lua.rs:186-232
    m = Material()
    for key, v in (t.get("material") or {}).items():
        if key in _TABLE_MATERIAL_KEYS:
            setattr(m, key, float(v))
        elif key == "color":
            m.color = _rgb(v, "material.color")
        elif key == "pattern":
            m.pattern = _pattern_from_table(v)          # Material.pattern is set directly, the colour stays (:209-214)
        else:
            raise ValueError(f"Invalid material property: {key}")
    if isinstance(t.get("color"), dict):
        m.color = _rgb(t["color"], "color")
    elif isinstance(t.get("pattern"), dict):
        m.pattern = _pattern_from_table(t["pattern"])
    return m


def camera_from_table(t):       # src/lua.rs:234-247
    cam = Camera.new_with_transform(int(t["screenwidth"]), int(t["screenheight"]), float(t["fov"]),
                                    Matrix.make_view_transform(_xyz(t["position"], "position"), _xyz(t["lookat"], "lookat"), _xyz(t["up"], "up")))
    if "samples" in t:
        cam.set_samples(int(t["samples"]))
    return cam


def world_from_table(t, all_lights=False):        # src/lua.rs:303-330
    """lights_from_table keeps the first light of the list (src/lua.rs:148-150); all_lights=True keeps them all."""
    lights = t["lights"]
    seq = list(lights) if isinstance(lights, (list, tuple)) else [lights[k] for k in sorted(lights)]   # Lua sequences start at 1
    first = seq[0]
    world = World(Light(_rgb(first["color"], "light.color"), _xyz(first["position"], "light.position")))
    if all_lights:
        for l in seq[1:]:
            world.add_light(Light(_rgb(l["color"], "light.color"), _xyz(l["position"], "light.position")))
    for shape in t["shapes"]:
        kind = shape.get("type")
        if kind not in _SHAPES:
            raise ValueError(f"Invalid shape type: {kind}")
        world.add_shape(_SHAPES[kind](_transform_from_table(shape), _material_from_table(shape)))
    return world
