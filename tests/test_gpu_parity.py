"""GPU parity tests proper: the CUDA path, called through the C ABI, against the oracle on the
same inputs.  f64 parity mode: hit/miss mask bit-exact, |delta| <= 1e-9 per channel (north star).
f32 fast mode: within 1/255 after the reference's truncating 8-bit rule on >= 99.9 % of pixels."""
import math
import os
import random

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
TOL64 = 1e-9
NCPU = os.cpu_count() or 1


def _check64(rtc, orc, renderer, world, cam, rect=None, max_depth=5, flat=None):
    rect = rtc.ffi.RECT_FULL if rect is None else rect
    if flat is not None:
        renderer.upload_flat(*flat)
        ow = orc.OracleWorld(flat=flat)
    else:
        renderer.upload(world)
        ow = orc.OracleWorld(world)
    want, want_hit, cnt = ow.render(cam, rect=rect, max_depth=max_depth, threads=NCPU)
    opts = rtc.Renderer.make_opts(precision=rtc.ffi.PRECISION_F64_PARITY, rect=rect, max_depth=max_depth)
    got, hit, stats = renderer.render(cam.flatten(), opts)
    assert np.array_equal(hit, want_hit), f"hit mask differs at {np.argwhere(hit != want_hit)[:5]}"
    err = np.abs(got - want).max()
    assert err <= TOL64, f"max |delta| = {err} at {np.unravel_index(np.abs(got - want).argmax(), got.shape)}"
    assert stats.rays == cnt.rays and stats.shadow_rays == cnt.shadow_rays and stats.shaded_hits == cnt.shaded_hits
    return got, want, stats


def test_render1_kat_through_the_c_abi(rtc, orc, renderer):
    """src/camera.rs:216-231: default world 11x11, pixel (5,5) = (0.38066, 0.47583, 0.2855)."""
    world = rtc.World.default()
    cam = rtc.Camera.new_with_transform(11, 11, math.pi / 2.,
                                        rtc.Matrix.make_view_transform((0., 0., -5.), (0., 0., 0.), (0., 1., 0.)))
    c = cam.render(world, renderer=renderer)
    p = c.get_pixel(5, 5)
    assert abs(p.red - 0.38066) < 1e-4 and abs(p.green - 0.47583) < 1e-4 and abs(p.blue - 0.2855) < 1e-4
    assert (c.pixels[10, :] == 0).all() and (c.pixels[:, 10] == 0).all()  # Camera::render's loop bounds
    assert c.hit_mask[5, 5] == 1 and c.hit_mask[0, 0] == 0
    _check64(rtc, orc, renderer, world, cam, rect=rtc.ffi.RECT_REFERENCE_RENDER)
    _check64(rtc, orc, renderer, world, cam, rect=rtc.ffi.RECT_FULL)


@pytest.mark.parametrize("rect", [0, 1])
def test_c1_reflect_refract_320x160(rtc, orc, renderer, rect):
    world = rtc.scenes.reflect_refract_scene()
    cam = rtc.scenes.reflect_refract_camera(320, 160)
    got, want, stats = _check64(rtc, orc, renderer, world, cam, rect=rect)
    if rect == 1:
        assert stats.rays == 650692 and stats.primary_hits == 320 * 160


@pytest.mark.parametrize("depth", [0, 1, 2, 3, 5, 8])
def test_depth_limits(rtc, orc, renderer, depth):
    world = rtc.scenes.reflect_refract_scene()
    cam = rtc.scenes.reflect_refract_camera(96, 48)
    _check64(rtc, orc, renderer, world, cam, max_depth=depth)


def test_supersample_2x2(rtc, orc, renderer):
    """antialiasing_samples = 0: the deterministic 4-sample path of render_pixel (src/camera.rs:96-113)."""
    world = rtc.scenes.reflect_refract_scene()
    cam = rtc.scenes.reflect_refract_camera(160, 80, samples=0)
    got, want, stats = _check64(rtc, orc, renderer, world, cam)
    assert stats.primary_hits == 160 * 80 * 4


@pytest.mark.parametrize("size", [(1, 1), (2, 2), (7, 3), (33, 17), (129, 65)])
def test_ragged_sizes(rtc, orc, renderer, size):
    world = rtc.scenes.reflect_refract_scene()
    for samples in (1, 0):
        cam = rtc.scenes.reflect_refract_camera(size[0], size[1], samples=samples)
        for rect in (0, 1):
            _check64(rtc, orc, renderer, world, cam, rect=rect)


def test_bench_scene(rtc, orc, renderer):
    world, cam = rtc.scenes.bench_scene()
    _check64(rtc, orc, renderer, world, cam)


def _random_world(rtc, seed, n=24):
    rnd = random.Random(seed)
    w = rtc.World(rtc.Light(rtc.Color(rnd.uniform(0.5, 1), rnd.uniform(0.5, 1), rnd.uniform(0.5, 1)),
                            (rnd.uniform(-8, 8), rnd.uniform(4, 10), rnd.uniform(-8, 8))))
    pats = [rtc.TestPattern, rtc.StripePattern, rtc.GradientPattern, rtc.RingPattern, rtc.CheckerPattern, rtc.GridPattern]

    def material():
        m = rtc.Material(color=rtc.Color(rnd.random(), rnd.random(), rnd.random()), ambient=rnd.uniform(0, 0.3),
                         diffuse=rnd.uniform(0.2, 0.9), specular=rnd.uniform(0, 0.9), shininess=rnd.uniform(1, 300))
        k = rnd.random()
        if k < 0.3:
            m.reflectiveness = rnd.uniform(0.1, 0.9)
        elif k < 0.55:
            m.reflectiveness, m.transparency, m.refractive_index = rnd.uniform(0.1, 0.9), rnd.uniform(0.1, 1.0), rnd.uniform(1.0, 2.5)
        elif k < 0.7:
            m.transparency, m.refractive_index = rnd.uniform(0.1, 1.0), rnd.uniform(1.0, 2.5)
        if rnd.random() < 0.5:
            p = rnd.choice(pats)(rtc.Color(rnd.random(), rnd.random(), rnd.random()), rtc.Color(rnd.random(), rnd.random(), rnd.random()))
            p.set_transform(rtc.Matrix.identity().scaling(rnd.uniform(0.2, 1), rnd.uniform(0.2, 1), rnd.uniform(0.2, 1))
                            .rotation_y(rnd.uniform(0, 3)).translation(rnd.uniform(-1, 1), 0, 0))
            m.set_pattern(p)
        return m

    w.add_shape(rtc.Plane(rtc.Matrix.identity().translation(0, -1, 0), material()))
    if rnd.random() < 0.5:
        w.add_shape(rtc.Plane(rtc.Matrix.identity().rotation_x(math.pi / 2).translation(0, 0, 12), material()))
    for _ in range(n):
        m = rtc.Matrix.identity().scaling(rnd.uniform(0.3, 1.5), rnd.uniform(0.3, 1.5), rnd.uniform(0.3, 1.5))
        m = m.rotation_z(rnd.uniform(0, 3)).rotation_x(rnd.uniform(0, 3)).shearing(rnd.uniform(-.3, .3), 0, 0, rnd.uniform(-.3, .3), 0, 0)
        m = m.translation(rnd.uniform(-5, 5), rnd.uniform(-0.5, 4), rnd.uniform(-2, 8))
        w.add_shape(rtc.Sphere(m, material()))
    # nested glass spheres (containers in compute_refractive, src/shape.rs:115-141)
    for r, ri in ((2.0, 1.5), (1.4, 2.0), (0.8, 1.2)):
        w.add_shape(rtc.Sphere(rtc.Matrix.identity().scaling(r, r, r).translation(0, 1.2, 3),
                               rtc.Material(color=rtc.Color(0.1, 0.1, 0.1), diffuse=0.2, reflectiveness=0.5, transparency=0.9, refractive_index=ri)))
    cam = rtc.Camera.new_with_transform(120, 90, 1.0, rtc.Matrix.make_view_transform((0, 2.0, -9.0), (0, 1, 3), (0, 1, 0)))
    return w, cam


@pytest.mark.parametrize("seed", [1, 2, 3, 4, 5, 6])
def test_random_worlds_all_patterns_nested_glass(rtc, orc, renderer, seed):
    w, cam = _random_world(rtc, seed)
    _check64(rtc, orc, renderer, w, cam)


def test_camera_inside_glass_and_behind_planes(rtc, orc, renderer):
    """Rays that start inside shapes: negative-t entries count as containers."""
    w, _ = _random_world(rtc, 9, n=6)
    cam = rtc.Camera.new_with_transform(80, 60, 1.3, rtc.Matrix.make_view_transform((0.2, 1.3, 2.6), (0, 1.0, 8), (0, 1, 0)))
    _check64(rtc, orc, renderer, w, cam)


def test_aliased_world_ids(rtc, orc, renderer):
    """More than 255 shapes: world_id wraps (src/shape.rs:661-667, u8) and compute_refractive
    matches containers by id, so shapes alias.  The device path reproduces that."""
    rnd = random.Random(5)
    w = rtc.World(rtc.Light(rtc.Color(1, 1, 1), (-5, 8, -6)))
    w.add_shape(rtc.Plane(rtc.Matrix.identity(), rtc.Material(reflectiveness=0.3, transparency=0.5, refractive_index=1.3)))
    for i in range(300):
        r = rnd.uniform(0.2, 0.9)
        glass = i % 3 == 0
        mat = rtc.Material(color=rtc.Color(rnd.random(), rnd.random(), rnd.random()),
                           transparency=0.9 if glass else 0.0, reflectiveness=0.6 if glass else 0.0,
                           refractive_index=rnd.uniform(1.1, 2.0))
        w.add_shape(rtc.Sphere(rtc.Matrix.identity().scaling(r, r, r).translation(rnd.uniform(-4, 4), rnd.uniform(0.2, 3), rnd.uniform(-1, 6)), mat))
    assert len({s.world_id for s in w.shapes}) < len(w.shapes)
    cam = rtc.Camera.new_with_transform(96, 64, 1.1, rtc.Matrix.make_view_transform((0, 1.5, -3.0), (0, 1, 3), (0, 1, 0)))
    _check64(rtc, orc, renderer, w, cam)


def test_output_formats(rtc, orc, renderer):
    world = rtc.scenes.reflect_refract_scene()
    cam = rtc.scenes.reflect_refract_camera(160, 80)
    renderer.upload(world)
    d = cam.flatten()
    f64, _, _ = renderer.render(d, rtc.Renderer.make_opts(out_format=rtc.ffi.OUT_F64_RGB))
    f32, _, _ = renderer.render(d, rtc.Renderer.make_opts(out_format=rtc.ffi.OUT_F32_RGB))
    u8, _, _ = renderer.render(d, rtc.Renderer.make_opts(out_format=rtc.ffi.OUT_U8_RGB))
    assert np.array_equal(f32, f64.astype(np.float32))
    assert np.array_equal(u8, orc.quantize_u8(f64))  # Color::scale, src/color.rs:100-114


def test_unsupported_inputs_are_errors_not_fallbacks(rtc, renderer):
    w2 = rtc.World.default()
    w2.shapes[0].material.color = None
    with pytest.raises(rtc.ffi.RtcError):
        renderer.upload(w2)
    renderer.upload(rtc.World.default())
    cam = rtc.Camera.new(16, 16, 1.0)
    d = cam.flatten()
    d.samples = 256   # antialiasing_samples is a u8 in the reference (src/camera.rs:26)
    with pytest.raises(rtc.ffi.RtcError) as e:
        renderer.render(d, rtc.Renderer.make_opts())
    assert e.value.code == rtc.ffi.RTC_ERR_UNSUPPORTED
    fresh = rtc.Renderer([0])
    with pytest.raises(rtc.ffi.RtcError) as e:
        fresh.render(rtc.Camera.new(4, 4, 1.0).flatten(), rtc.Renderer.make_opts())
    assert e.value.code == rtc.ffi.RTC_ERR_NO_SCENE


# ---- Cube (src/shape.rs:500-630) -----------------------------------------------------
def _cube_world(rtc, seed=7, n_cubes=9, with_spheres=True):
    return rtc.scenes.cube_scene(seed, n_cubes, with_spheres)


@pytest.mark.parametrize("seed", [7, 8, 9])
def test_cubes_f64_parity(rtc, orc, renderer, seed):
    w, cam = _cube_world(rtc, seed)
    _check64(rtc, orc, renderer, w, cam)


def test_cubes_only_axis_aligned_rays_f64(rtc, orc, renderer):
    """Rays parallel to cube faces (|direction| < EPSILON on an axis -> the +-infinity branch of check_axis,
    src/shape.rs:549-559), seen head-on by a camera looking down -z at unit cubes; depth 0 and 5."""
    I = rtc.Matrix.identity
    w = rtc.World(rtc.Light(rtc.Color(1, 1, 1), (0, 0, 10)))
    w.add_shape(rtc.Cube(I(), rtc.Material(color=rtc.Color(0.8, 0.3, 0.2), reflectiveness=0.5)))
    w.add_shape(rtc.Cube(I().translation(2.5, 0, 0), rtc.Material(color=rtc.Color(0.1, 0.1, 0.1), transparency=0.9, reflectiveness=0.6, refractive_index=1.4)))
    w.add_shape(rtc.Cube(I().scaling(8, 8, 0.5).translation(0, 0, -4), rtc.Material(color=rtc.Color(0.4, 0.5, 0.9))))
    cam = rtc.Camera.new_with_transform(96, 64, 0.9, rtc.Matrix.make_view_transform((0, 0, 8), (0, 0, 0), (0, 1, 0)))
    for depth in (0, 5):
        _check64(rtc, orc, renderer, w, cam, max_depth=depth)


def test_cubes_with_aliased_world_ids_f64_and_f32(rtc, orc, renderer):
    """> 255 shapes (u8 world_ids wrap, src/shape.rs:661-667) with cubes among them: the containers of
    compute_refractive are toggled by id, cubes included."""
    w, cam = _cube_world(rtc, 11, n_cubes=300, with_spheres=True)
    _check64(rtc, orc, renderer, w, cam)
    _check32(rtc, orc, renderer, w, cam, min_ok=0.995)


@pytest.mark.parametrize("seed", [7, 8])
def test_cubes_f32(rtc, orc, renderer, seed):
    w, cam = _cube_world(rtc, seed)
    ok, mm, stats = _check32(rtc, orc, renderer, w, cam, min_ok=0.998)
    assert mm == 0   # closed room


# ---- f32 fast mode ---------------------------------------------------------------------------------
def _check32(rtc, orc, renderer, world, cam, rect=1, max_depth=5, min_ok=0.999, ray_tol=0.002, general=False, mask_tol=1e-3):
    """general=True forces the general f32 kernel (render_f32.cu); otherwise scenes that fit take the
    constant-bank small-scene kernel (render_f32_small.cu).  Hit mask: an all-f32 primary ray can fall on the other side
    of a silhouette than the f64 one, so open scenes get a stated allowance (mask_tol, a fraction of the pixels; the
    callers with closed rooms assert zero); big scenes, whose geometry runs in f64, are held to an exact mask
    (_lattice_check)."""
    renderer.upload(world)
    want, want_hit, cnt = orc.OracleWorld(world).render(cam, rect=rect, max_depth=max_depth, threads=NCPU)
    opts = rtc.Renderer.make_opts(precision=rtc.ffi.PRECISION_F32_FAST, rect=rect, max_depth=max_depth,
                                  out_format=rtc.ffi.OUT_F32_RGB, flags=rtc.ffi.OPT_F32_GENERAL_KERNEL if general else 0)
    got, hit, stats = renderer.render(cam.flatten(), opts)
    assert np.isfinite(got).all()
    q_got = orc.quantize_u8(got.astype(np.float64)).astype(int)   # Color::scale, src/color.rs:100-114
    q_want = orc.quantize_u8(want).astype(int)
    ok = (np.abs(q_got - q_want).max(axis=2) <= 1).mean()
    assert ok >= min_ok, f"only {ok:.5f} of pixels within 1/255"
    mask_mismatch = int((hit != want_hit).sum())
    assert mask_mismatch <= mask_tol * hit.size, f"f32 hit mask differs from the oracle's on {mask_mismatch} of {hit.size} pixels"
    assert abs(int(stats.rays) - int(cnt.rays)) <= ray_tol * cnt.rays, (stats.rays, cnt.rays)
    return ok, mask_mismatch, stats


@pytest.mark.parametrize("general", [False, True])
@pytest.mark.parametrize("size,samples", [((320, 160), 1), ((400, 200), 1), ((160, 80), 0), ((33, 17), 1), ((1, 1), 1), ((9, 5), 0)])
def test_f32_reflect_refract(rtc, orc, renderer, size, samples, general):
    world = rtc.scenes.reflect_refract_scene()
    cam = rtc.scenes.reflect_refract_camera(size[0], size[1], samples=samples)
    ok, mm, stats = _check32(rtc, orc, renderer, world, cam, general=general, min_ok=0.999 if size[0] * size[1] >= 1000 else 0.97)
    assert mm == 0  # closed room: every primary ray hits in both precisions


@pytest.mark.parametrize("general", [False, True])
@pytest.mark.parametrize("depth", [0, 1, 3, 8])
def test_f32_depth_limits(rtc, orc, renderer, depth, general):
    world = rtc.scenes.reflect_refract_scene()
    cam = rtc.scenes.reflect_refract_camera(200, 100)
    _check32(rtc, orc, renderer, world, cam, max_depth=depth, general=general)


def test_f32_c2_full_hd(rtc, orc, renderer):
    """Config C2: the reflect/refract room at 1920x1080, f32 fast mode vs the oracle (all host threads)."""
    world = rtc.scenes.reflect_refract_scene()
    cam = rtc.scenes.reflect_refract_camera(1920, 1080)
    ok, mm, stats = _check32(rtc, orc, renderer, world, cam)
    assert mm == 0


def _glass_world(rtc, n_glass, glass_plane=False):
    """Uniform spheres only (small-scene kernel), nested and overlapping glass, optional glass plane."""
    rnd = random.Random(100 + n_glass)
    w = rtc.World(rtc.Light(rtc.Color(1, 1, 1), (-6, 9, -7)))
    floor = rtc.Material(specular=0.0, reflectiveness=0.2, pattern=rtc.CheckerPattern(rtc.Color(0.3, 0.3, 0.3), rtc.Color(0.7, 0.7, 0.7)))
    w.add_shape(rtc.Plane(rtc.Matrix.identity().translation(0, -1, 0), floor))
    w.add_shape(rtc.Plane(rtc.Matrix.identity().rotation_x(math.pi / 2).translation(0, 0, 14),
                          rtc.Material(color=rtc.Color(0.5, 0.6, 0.7), specular=0.1)))
    if glass_plane:
        w.add_shape(rtc.Plane(rtc.Matrix.identity().translation(0, -0.5, 0),
                              rtc.Material(color=rtc.Color(0.05, 0.1, 0.1), diffuse=0.2, reflectiveness=0.3, transparency=0.8, refractive_index=1.33)))
    for r, ri in ((2.0, 1.5), (1.4, 2.0), (0.8, 1.2)):   # src/shape.rs:1269-1304 style nesting
        w.add_shape(rtc.Sphere(rtc.Matrix.identity().scaling(r, r, r).translation(0, 1.2, 3),
                               rtc.Material(color=rtc.Color(0.1, 0.1, 0.1), diffuse=0.2, reflectiveness=0.5, transparency=0.9, refractive_index=ri)))
    for i in range(n_glass):
        r = rnd.uniform(0.4, 0.9)
        w.add_shape(rtc.Sphere(rtc.Matrix.identity().scaling(r, r, r).translation(rnd.uniform(-4, 4), rnd.uniform(0, 2), rnd.uniform(0, 7)),
                               rtc.Material(color=rtc.Color(rnd.random() * 0.3, rnd.random() * 0.3, rnd.random() * 0.3), diffuse=0.3,
                                            reflectiveness=0.6, transparency=0.8, refractive_index=rnd.uniform(1.1, 2.0))))
    for i in range(8):
        r = rnd.uniform(0.3, 0.8)
        w.add_shape(rtc.Sphere(rtc.Matrix.identity().scaling(r, r, r).translation(rnd.uniform(-5, 5), rnd.uniform(-0.5, 3), rnd.uniform(-1, 9)),
                               rtc.Material(color=rtc.Color(rnd.random(), rnd.random(), rnd.random()), reflectiveness=rnd.choice([0, 0, 0.5]))))
    cam = rtc.Camera.new_with_transform(240, 160, 1.0, rtc.Matrix.make_view_transform((0, 2.0, -7.0), (0, 1, 3), (0, 1, 0)))
    return w, cam


@pytest.mark.parametrize("general", [False, True])
@pytest.mark.parametrize("n_glass,glass_plane", [(2, False), (12, False), (3, True)])
def test_f32_nested_glass_small_scene(rtc, orc, renderer, n_glass, glass_plane, general):
    """compute_refractive in both f32 kernels: fast n1/n2 test (<= 8 glass spheres), its overflow
    (> 8) and a glass plane (fast test unavailable)."""
    w, cam = _glass_world(rtc, n_glass, glass_plane)
    _check32(rtc, orc, renderer, w, cam, min_ok=0.995, ray_tol=0.01, general=general)


def test_f32_camera_inside_glass(rtc, orc, renderer):
    w, _ = _glass_world(rtc, 2)
    cam = rtc.Camera.new_with_transform(160, 120, 1.3, rtc.Matrix.make_view_transform((0.2, 1.3, 2.6), (0, 1.0, 8), (0, 1, 0)))
    _check32(rtc, orc, renderer, w, cam, min_ok=0.99, ray_tol=0.01)


def test_f32_light_inside_sphere_takes_general_kernel(rtc, orc, renderer):
    """A light inside a (glass) sphere disqualifies the small-scene kernel's shadow test; the general
    kernel's light_inside variant must give the same image as the oracle."""
    w, cam = _glass_world(rtc, 2)
    w.light = rtc.Light(rtc.Color(1, 1, 1), (0, 1.2, 3))   # centre of the nested glass spheres
    _check32(rtc, orc, renderer, w, cam, min_ok=0.99, ray_tol=0.01)


@pytest.mark.parametrize("dtype", [np.float64, np.float32])
@pytest.mark.parametrize("gamma", [1.0, 2.2, 0.5])
def test_pack_rgba8_is_canvas_to_imgbuf(rtc, orc, renderer, dtype, gamma):
    """rtc_pack_rgba8_device against Canvas::to_imgbuf (src/canvas.rs:61-79): channel =
    Color::scale(component.powf(1/gamma), 255) with `as i32` truncation / saturation / NaN -> 0 and the clamp
    (src/color.rs:55-65,100-114), alpha 255.  gamma = 1 (the Canvas default) must be exact; with a gamma the
    device pow may differ from the host's by an ulp, which can move a value across a truncation boundary."""
    import torch
    rng = np.random.default_rng(5)
    n = 1_000_003                                   # not a multiple of 4: exercises the tail pixels
    vals = rng.uniform(-0.2, 1.6, size=(n, 3))
    vals[:256, 0] = np.arange(256) / 255.0          # exact level boundaries
    vals[256:512, 1] = np.nextafter(np.arange(256) / 255.0, -1.0)
    vals[600, :] = (np.nan, np.inf, -np.inf)
    vals[601, :] = (1e300, -1e300, 0.0)
    vals[602, :] = (-0.0, 1.0, 0.999999999)
    with np.errstate(over="ignore"):
        src = vals.astype(dtype)               # 1e300 -> inf in f32, on purpose
    fmt = rtc.ffi.OUT_F64_RGB if dtype == np.float64 else rtc.ffi.OUT_F32_RGB
    dev = torch.from_numpy(src).cuda()
    out = torch.zeros(n, dtype=torch.int32, device="cuda")
    renderer.pack_rgba8_device(dev.data_ptr(), fmt, n, gamma, out.data_ptr())
    got = out.cpu().numpy().view(np.uint8).reshape(n, 4)
    x = src.astype(np.float64)
    if gamma != 1.0:
        with np.errstate(all="ignore"):
            x = np.power(x, np.float64(np.float32(1.0) / np.float32(gamma)))
    want = orc.quantize_u8(x)
    assert (got[:, 3] == 255).all()
    diff = np.abs(got[:, :3].astype(int) - want.astype(int))
    if gamma == 1.0:
        assert diff.max() == 0
    else:
        assert diff.max() <= 1 and (diff != 0).mean() < 1e-4, (diff.max(), (diff != 0).mean())


def test_pack_rgba8_of_a_rendered_frame(rtc, orc, renderer):
    import torch
    world = rtc.scenes.reflect_refract_scene()
    cam = rtc.scenes.reflect_refract_camera(160, 90)
    renderer.upload(world)
    f = rtc.ffi
    dev = torch.empty((90, 160, 3), dtype=torch.float64, device="cuda")
    renderer.render_device(cam.flatten(), rtc.Renderer.make_opts(precision=f.PRECISION_F64_PARITY, out_format=f.OUT_F64_RGB), dev.data_ptr())
    out = torch.zeros(90 * 160, dtype=torch.int32, device="cuda")
    renderer.pack_rgba8_device(dev.data_ptr(), f.OUT_F64_RGB, 90 * 160, 1.0, out.data_ptr())
    got = out.cpu().numpy().view(np.uint8).reshape(90, 160, 4)
    u8, _, _ = renderer.render(cam.flatten(), rtc.Renderer.make_opts(precision=f.PRECISION_F64_PARITY, out_format=f.OUT_U8_RGB))
    assert np.array_equal(got[:, :, :3], u8) and np.array_equal(got[:, :, :3], orc.quantize_u8(dev.cpu().numpy()))


def test_download_frame_matches_rtc_render(rtc, orc, renderer):
    """rtc_download_frame (device frame -> host Canvas buffer, f32 -> f64 widening on the host) delivers
    exactly what rtc_render delivers for the same options."""
    import torch
    world = rtc.scenes.reflect_refract_scene()
    cam = rtc.scenes.reflect_refract_camera(1280, 720)      # > 4 MB: several bands
    renderer.upload(world)
    d = cam.flatten()
    f = rtc.ffi
    want, _, _ = renderer.render(d, rtc.Renderer.make_opts(precision=f.PRECISION_F32_FAST, out_format=f.OUT_F64_RGB))
    dev = torch.empty((720, 1280, 3), dtype=torch.float32, device="cuda")
    renderer.render_device(d, rtc.Renderer.make_opts(precision=f.PRECISION_F32_FAST, out_format=f.OUT_F32_RGB), dev.data_ptr())
    got64 = renderer.download_frame(dev.data_ptr(), f.OUT_F32_RGB, 1280, 720, np.empty((720, 1280, 3), np.float64))
    got32 = renderer.download_frame(dev.data_ptr(), f.OUT_F32_RGB, 1280, 720, np.empty((720, 1280, 3), np.float32))
    assert np.array_equal(got64, want)
    assert np.array_equal(got32.astype(np.float64), want)
    with pytest.raises(rtc.ffi.RtcError):
        renderer.download_frame(dev.data_ptr(), f.OUT_F32_RGB, 1280, 720, np.empty((720, 1280, 3), np.uint8))


def test_f32_reference_rect_and_formats(rtc, orc, renderer):
    world = rtc.scenes.reflect_refract_scene()
    cam = rtc.scenes.reflect_refract_camera(96, 48)
    renderer.upload(world)
    d = cam.flatten()
    mk = lambda fmt: rtc.Renderer.make_opts(precision=rtc.ffi.PRECISION_F32_FAST, rect=rtc.ffi.RECT_REFERENCE_RENDER, out_format=fmt)
    _f32_formats(rtc, orc, renderer, d, mk)
    mk = lambda fmt: rtc.Renderer.make_opts(precision=rtc.ffi.PRECISION_F32_FAST, rect=rtc.ffi.RECT_REFERENCE_RENDER, out_format=fmt,
                                            flags=rtc.ffi.OPT_F32_GENERAL_KERNEL)
    _f32_formats(rtc, orc, renderer, d, mk)


def _f32_formats(rtc, orc, renderer, d, mk):
    f32, hit, _ = renderer.render(d, mk(rtc.ffi.OUT_F32_RGB))
    f64, _, _ = renderer.render(d, mk(rtc.ffi.OUT_F64_RGB))
    u8, _, _ = renderer.render(d, mk(rtc.ffi.OUT_U8_RGB))
    assert (f32[47, :] == 0).all() and (f32[:, 95] == 0).all() and (hit[47, :] == 0).all()
    assert np.array_equal(f64, f32.astype(np.float64))
    assert np.abs(u8.astype(int) - orc.quantize_u8(f64).astype(int)).max() <= 1


@pytest.mark.parametrize("general", [False, True])
def test_f32_bench_scene(rtc, orc, renderer, general):
    world, cam = rtc.scenes.bench_scene()
    _check32(rtc, orc, renderer, world, cam, general=general)


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_f32_random_worlds(rtc, orc, renderer, seed):
    """General affine spheres, all six patterns, nested glass: f32 is held to a looser bar here
    (pattern edges and grazing refraction amplify rounding); the 99.9 % bar is for the named scenes."""
    w, cam = _random_world(rtc, seed)
    _check32(rtc, orc, renderer, w, cam, min_ok=0.98, ray_tol=0.02)


# ---- row-tile sharding through rtc_render_device ---------------------------------------------------
@pytest.mark.parametrize("precision", [0, 1])
@pytest.mark.parametrize("count,tile", [(2, 16), (3, 4), (8, 16)])
def test_shards_compact_gather_and_full_frame_destination(rtc, orc, renderer, precision, count, tile):
    """Every shard rendered on this GPU: (a) compact buffers + rtc_assemble_device (the NCCL-gather
    layout), (b) RTC_OPT_DST_FULL_FRAME straight into one frame (the peer-store layout).  Both must
    equal the unsharded render bit for bit."""
    import torch
    world = rtc.scenes.reflect_refract_scene()
    cam = rtc.scenes.reflect_refract_camera(200, 117, samples=1)
    d = cam.flatten()
    renderer.upload(world)
    fmt = rtc.ffi.OUT_F64_RGB if precision == 0 else rtc.ffi.OUT_F32_RGB
    dt = torch.float64 if precision == 0 else torch.float32
    whole = torch.zeros((117, 200, 3), dtype=dt, device="cuda")
    whole_hit = torch.zeros((117, 200), dtype=torch.uint8, device="cuda")
    renderer.render_device(d, rtc.Renderer.make_opts(precision=precision, out_format=fmt), whole.data_ptr(), whole_hit.data_ptr())
    o0 = rtc.Renderer.make_opts(precision=precision, out_format=fmt, shard_index=0, shard_count=count, tile_rows=tile)
    stride = renderer.shard_rows(d, o0)
    gathered = torch.zeros((count, stride, 200, 3), dtype=dt, device="cuda")
    gathered_hit = torch.zeros((count, stride, 200), dtype=torch.uint8, device="cuda")
    full = torch.zeros_like(whole)
    full_hit = torch.zeros_like(whole_hit)
    rows = 0
    for k in range(count):
        ok = rtc.Renderer.make_opts(precision=precision, out_format=fmt, shard_index=k, shard_count=count, tile_rows=tile)
        rows += renderer.shard_rows(d, ok)
        renderer.render_device(d, ok, gathered[k].data_ptr(), gathered_hit[k].data_ptr())
        okf = rtc.Renderer.make_opts(precision=precision, out_format=fmt, shard_index=k, shard_count=count, tile_rows=tile,
                                     flags=rtc.ffi.OPT_DST_FULL_FRAME)
        renderer.render_device(d, okf, full.data_ptr(), full_hit.data_ptr())
    assert rows == 117
    frame = torch.zeros_like(whole)
    frame_hit = torch.zeros_like(whole_hit)
    renderer.assemble_device(d, o0, gathered.data_ptr(), stride, fmt, frame.data_ptr())
    renderer.assemble_device(d, o0, gathered_hit.data_ptr(), stride, -1, frame_hit.data_ptr())
    torch.cuda.synchronize()
    assert torch.equal(frame, whole) and torch.equal(frame_hit, whole_hit)
    assert torch.equal(full, whole) and torch.equal(full_hit, whole_hit)


@pytest.mark.parametrize("precision", [0, 1])
@pytest.mark.parametrize("count,tile,size", [(2, 4, (200, 117)), (3, 16, (200, 117)), (8, 4, (1920, 1080))])
def test_rtc_render_shards_into_one_host_frame(rtc, renderer, precision, count, tile, size):
    """rtc_render with shard_count > 1 (one-GPU context): every call fills in its shard's rows of the caller's full
    host frame and leaves the other rows alone — what N rank processes do with a frame in shared host memory.  All
    shards together equal the unsharded frame bit for bit (f64 Canvas layout, fast mode widened on the host)."""
    world = rtc.scenes.reflect_refract_scene()
    cam = rtc.scenes.reflect_refract_camera(size[0], size[1])
    d = cam.flatten()
    renderer.upload(world)
    whole, whole_hit, st = renderer.render(d, rtc.Renderer.make_opts(precision=precision))
    frame = np.full((size[1], size[0], 3), -7.0)
    mask = np.full((size[1], size[0]), 99, dtype=np.uint8)
    rays = 0
    for k in range(count):
        o = rtc.Renderer.make_opts(precision=precision, shard_index=k, shard_count=count, tile_rows=tile)
        if k == 1:   # the other shards' rows are untouched
            before = frame.copy()
        stats = rtc.ffi.RtcStats()
        import ctypes as C
        rtc.ffi.check(rtc.ffi.lib().rtc_render(renderer.handle, C.byref(d), C.byref(o), frame.ctypes.data_as(C.c_void_p),
                                               mask.ctypes.data_as(C.c_void_p), C.byref(stats)))
        rays += stats.rays
        if k == 1:
            rows = np.arange(size[1])
            mine = (rows // tile) % count == 1
            assert np.array_equal(frame[~mine], before[~mine]) and not np.array_equal(frame[mine], before[mine])
    assert np.array_equal(frame, whole) and np.array_equal(mask, whole_hit) and rays == st.rays


# ---- configs C4 and C5 of BASELINE.json ------------------------------------------------------------
@pytest.fixture(scope="module")
def c4_flat(rtc):
    """C4: 9,992 random spheres + 8 planes, seed 0xB200 (world_ids wrap at 256 like the reference's u8)."""
    world = rtc.scenes.synthetic_scene()
    assert len(world.shapes) == 10000
    return world.flatten()


def _lattice_check(rtc, orc, renderer, flat, cam, step, n_random, f32_min_ok, depth=5, f32_mask_exact=True):
    """Full-size frames the oracle cannot finish in seconds: the GPU renders the whole frame, the
    oracle a pixel lattice (every `step`-th pixel in x and y + random pixels), BASELINE.md §3."""
    renderer.upload_flat(*flat)
    d = cam.flatten()
    idx = orc.lattice(d.hsize, d.vsize, step=step, n_random=n_random)
    want, want_hit, cnt = orc.OracleWorld(flat=flat).render(cam, max_depth=depth, threads=NCPU, pixels=idx)
    got, hit, st = renderer.render(d, rtc.Renderer.make_opts(precision=rtc.ffi.PRECISION_F64_PARITY, max_depth=depth))
    assert np.array_equal(hit.reshape(-1)[idx], want_hit)
    err = np.abs(got.reshape(-1, 3)[idx] - want).max()
    assert err <= TOL64, f"max |delta| = {err}"
    # fast mode twice on big scenes: the library's own choice (scenes whose streams exceed shared memory go to the parity
    # kernel: the fastest kernel that meets the tolerance) and the general f32 kernel forced (its mixed-precision kRefine form)
    big = flat[2] > 4000
    q_want = orc.quantize_u8(want).astype(int)
    for f32_flags in ((0, rtc.ffi.OPT_F32_GENERAL_KERNEL) if big else (0,)):
        got32, hit32, st32 = renderer.render(d, rtc.Renderer.make_opts(precision=rtc.ffi.PRECISION_F32_FAST, max_depth=depth, out_format=rtc.ffi.OUT_F32_RGB,
                                                                       flags=f32_flags))
        q_got = orc.quantize_u8(got32.reshape(-1, 3)[idx].astype(np.float64)).astype(int)
        ok = (np.abs(q_got - q_want).max(axis=1) <= 1).mean()
        mm = int((hit32.reshape(-1)[idx] != want_hit).sum())
        if f32_flags == 0 and big:
            assert ok >= f32_min_ok and mm == 0 and st32.rays == st.rays, (ok, mm)
    assert mm == 0 if f32_mask_exact else mm <= 1e-3 * idx.size, f"f32 hit mask differs from the oracle's on {mm} of {idx.size} lattice pixels"
    if ok < f32_min_ok:   # say where it fails before failing: the fraction per recursion depth limit
        per_depth = {}
        for dd in range(0, depth):
            w_d, _, _ = orc.OracleWorld(flat=flat).render(cam, max_depth=dd, threads=NCPU, pixels=idx)
            g_d, _, _ = renderer.render(d, rtc.Renderer.make_opts(precision=rtc.ffi.PRECISION_F32_FAST, max_depth=dd, out_format=rtc.ffi.OUT_F32_RGB,
                                                                  flags=rtc.ffi.OPT_F32_GENERAL_KERNEL if big else 0))
            qd = orc.quantize_u8(g_d.reshape(-1, 3)[idx].astype(np.float64)).astype(int)
            per_depth[dd] = float((np.abs(qd - orc.quantize_u8(w_d).astype(int)).max(axis=1) <= 1).mean())
        raise AssertionError(f"f32: only {ok:.5f} of lattice pixels within 1/255 at depth {depth}; by depth limit: {per_depth}")
    assert abs(int(st32.rays) - int(st.rays)) <= 0.01 * st.rays
    return st, st32, ok


def test_c4_synthetic_10k_small_frame(rtc, orc, renderer, c4_flat):
    """Every pixel of a 96x54 frame of the 10k-shape scene: f64 parity (aliased world_ids, affine
    spheres, glass) and fast mode — the kRefine kernel (f64 ray geometry behind the f32 hierarchy) at the
    north star's 99.9 % bar, hit mask exact on an open scene with genuine misses."""
    cam = rtc.scenes.synthetic_camera(96, 54)
    st, st32, ok = _lattice_check(rtc, orc, renderer, c4_flat, cam, step=1, n_random=0, f32_min_ok=0.999)
    assert 0 < st.primary_hits < 96 * 54          # open scene: genuine misses in the mask


def test_c4_f32_meets_the_bar_at_every_depth_limit(rtc, orc, renderer, c4_flat):
    """Thousands of r = 0.05..0.35 spheres are strongly diverging mirrors: a hit-point error of one f32
    ulp (4e-6 at |p| ~ 40) is amplified ~10^3 per sphere-to-sphere bounce, so an all-f32 path leaves the
    f64 reference's ray tree after two or three bounces (round 1: 97.5 % at depth 5).  The kRefine kernels carry the
    geometry of the radiance rays in f64: the bar is the north star's at every depth limit."""
    cam = rtc.scenes.synthetic_camera(192, 108)
    for depth in (0, 1, 2, 3):
        _lattice_check(rtc, orc, renderer, c4_flat, cam, step=1, n_random=0, f32_min_ok=0.999, depth=depth)


def test_c4_synthetic_10k_4k_lattice(rtc, orc, renderer, c4_flat):
    cam = rtc.scenes.synthetic_camera(3840, 2160)
    _lattice_check(rtc, orc, renderer, c4_flat, cam, step=32, n_random=1024, f32_min_ok=0.999)


def _big_mixed_world(rtc):
    """6,000 random spheres (every 10th affine) + 300 cubes and 60 needle-like ellipsoids: more than fits the
    shared-memory streams, so the f32 general kernel keeps them in global memory behind the bounding-sphere filter."""
    import math
    from rtc_b200.scene import Cube, Material, Sphere, Color
    from rtc_b200.transform import Matrix
    w = rtc.scenes.synthetic_scene(n_spheres=6000, seed=0xC0BE)
    u = lambda i, k: rtc.scenes._u01(0xC0BE5, i, k)
    for i in range(300):
        m = Matrix.identity().scaling(0.1 + 0.5 * u(i, 0), 0.1 + 0.5 * u(i, 1), 0.1 + 0.5 * u(i, 2))
        m = m.rotation_x(6.283 * u(i, 3)).rotation_y(6.283 * u(i, 4)).translation(-18 + 36 * u(i, 5), 0.5 + 9 * u(i, 6), -18 + 36 * u(i, 7))
        glass = i % 7 == 0
        w.add_shape(Cube(m, Material(color=Color(0.2 + 0.7 * u(i, 8), 0.2 + 0.7 * u(i, 9), 0.3), reflectiveness=0.5 if glass else 0.1,
                                     transparency=0.8 if glass else 0.0, refractive_index=1.5 if glass else 1.0)))
    for i in range(60):
        m = Matrix.identity().scaling(0.02, 1.5 + 2 * u(i, 10), 0.05).rotation_z(6.283 * u(i, 11)).rotation_y(6.283 * u(i, 12))
        m = m.translation(-15 + 30 * u(i, 13), 2 + 6 * u(i, 14), -15 + 30 * u(i, 15))
        w.add_shape(Sphere(m, Material(color=Color(0.9, 0.4, 0.1), reflectiveness=0.3)))
    return w


def test_bounding_sphere_filter_changes_nothing(rtc, renderer, c4_flat):
    """Scenes whose streams stay in global memory run the exact shape tests of the general f32 kernel (forced here: by
    default such scenes go to the parity kernel) behind a box / sphere hierarchy (render_f32.cu warp_query).  The
    hierarchy is conservative: the frame, the hit mask and the ray counters are those of the plain brute-force loops
    (RTC_OPT_NO_BOUND_FILTER), bit for bit — spheres, affine spheres, cubes, needle-shaped ellipsoids, aliased world_ids."""
    import torch
    ffi = rtc.ffi
    for flat, cam in ((c4_flat, rtc.scenes.synthetic_camera(640, 360)),
                      (_big_mixed_world(rtc).flatten(), rtc.scenes.synthetic_camera(480, 270))):
        renderer.upload_flat(*flat)
        d = cam.flatten()
        out = []
        for flags in (ffi.OPT_F32_GENERAL_KERNEL, ffi.OPT_F32_GENERAL_KERNEL | ffi.OPT_NO_BOUND_FILTER):
            buf = torch.empty((d.vsize, d.hsize, 3), dtype=torch.float32, device="cuda")
            hit = torch.empty((d.vsize, d.hsize), dtype=torch.uint8, device="cuda")
            st = ffi.RtcStats()
            renderer.render_device(d, rtc.Renderer.make_opts(precision=ffi.PRECISION_F32_FAST, out_format=ffi.OUT_F32_RGB, flags=flags),
                                   buf.data_ptr(), hit.data_ptr(), None, st)
            out.append((buf, hit, (st.rays, st.primary_hits, st.shaded_hits, st.shadow_rays)))
        assert torch.equal(out[0][1], out[1][1])
        assert torch.equal(out[0][0], out[1][0])
        assert out[0][2] == out[1][2] and out[0][2][0] > d.hsize * d.vsize


def test_f64_line_distance_filter_changes_nothing(rtc, renderer, c4_flat):
    """The parity kernel runs a sphere's / cube's exact f64 test only when an f32 test says the ray's line passes within
    its bounding sphere plus a rounding margin (render_f64.cu intersect_world).  Skipped shapes have no root, so frame,
    hit mask and counters are those of the plain loops (RTC_OPT_NO_BOUND_FILTER) bit for bit: the room (streams in
    shared memory), 10k spheres (global memory), cubes / needles / affine spheres, resample pass, two lights."""
    import torch
    ffi = rtc.ffi
    two = rtc.scenes.reflect_refract_scene()
    two.add_light(rtc.Light(rtc.Color(0.3, 0.5, 0.9), (3.5, 3.0, -3.0)))
    cube_world, cube_cam = rtc.scenes.cube_scene(7, hsize=120, vsize=84)
    for world, flat, cam in ((rtc.scenes.reflect_refract_scene(), None, rtc.scenes.reflect_refract_camera(640, 320)),
                             (rtc.scenes.reflect_refract_scene(), None, rtc.scenes.reflect_refract_camera(160, 80, samples=3)),
                             (two, None, rtc.scenes.reflect_refract_camera(200, 100)),
                             (cube_world, None, cube_cam),
                             (None, c4_flat, rtc.scenes.synthetic_camera(320, 180)),
                             (None, _big_mixed_world(rtc).flatten(), rtc.scenes.synthetic_camera(240, 135))):
        if world is not None:
            renderer.upload(world)
        else:
            renderer.upload_flat(*flat)
        d = cam.flatten()
        out = []
        for flags in (0, ffi.OPT_NO_BOUND_FILTER):
            buf = torch.empty((d.vsize, d.hsize, 3), dtype=torch.float64, device="cuda")
            hit = torch.empty((d.vsize, d.hsize), dtype=torch.uint8, device="cuda")
            st = ffi.RtcStats()
            renderer.render_device(d, rtc.Renderer.make_opts(precision=ffi.PRECISION_F64_PARITY, flags=flags), buf.data_ptr(), hit.data_ptr(), None, st)
            out.append((buf, hit, (st.rays, st.primary_hits, st.shaded_hits, st.shadow_rays), st.kernel_ms))
        assert torch.equal(out[0][1], out[1][1]) and torch.equal(out[0][0], out[1][0])
        assert out[0][2] == out[1][2] and out[0][2][0] >= d.hsize * d.vsize


def test_big_mixed_world_f64_parity_and_f32(rtc, orc, renderer):
    """The same mixed world against the oracle on a small frame: f64 parity; fast mode at the 99.9 % bar with an exact
    hit mask (cubes, needle ellipsoids and affine spheres through the f64 exact tests of the kRefine kernel)."""
    cam = rtc.scenes.synthetic_camera(96, 54)
    _lattice_check(rtc, orc, renderer, _big_mixed_world(rtc).flatten(), cam, step=1, n_random=0, f32_min_ok=0.999)


def test_c3_4k_lattice(rtc, orc, renderer):
    """C3, the frame every bench number is quoted on: the reflect/refract room at 3840x2160, 1 sample per pixel, depth
    5 — both precisions against the oracle on every 16th pixel in x and y plus 4096 random pixels."""
    world = rtc.scenes.reflect_refract_scene()
    cam = rtc.scenes.reflect_refract_camera(3840, 2160)
    st, st32, ok = _lattice_check(rtc, orc, renderer, world.flatten(), cam, step=16, n_random=4096, f32_min_ok=0.999)
    assert st.primary_hits == 3840 * 2160 and st.rays == 103212494


def test_c5_8k_supersampled_lattice(rtc, orc, renderer):
    """C5: the reflect/refract room at 7680x4320 with the deterministic 2x2 supersample (samples = 0)."""
    world = rtc.scenes.reflect_refract_scene()
    cam = rtc.scenes.reflect_refract_camera(7680, 4320, samples=0)
    st, st32, ok = _lattice_check(rtc, orc, renderer, world.flatten(), cam, step=32, n_random=2048, f32_min_ok=0.999)
    assert st.primary_hits == 7680 * 4320 * 4


@pytest.mark.parametrize("n_spheres", [100, 200, 1500])
def test_fast_mode_dispatch_by_scene_size(rtc, renderer, n_spheres):
    """Fast mode promises the tolerance, not a data type: beyond 128 shapes the library runs the parity kernel in its
    hierarchical form (faster there than the general f32 kernel, and exact), so the fast frame IS the parity frame; up to
    128 shapes, and with RTC_OPT_F32_GENERAL_KERNEL at any size, the general f32 kernel renders it — within 1/255 of the
    parity frame on >= 99.9 % of the pixels, hit mask identical."""
    f = rtc.ffi
    renderer.upload_flat(*rtc.scenes.synthetic_scene(n_spheres=n_spheres, seed=0xB200).flatten())
    d = rtc.scenes.synthetic_camera(192, 108).flatten()
    want, want_hit, st = renderer.render(d, rtc.Renderer.make_opts(precision=f.PRECISION_F64_PARITY))
    fast, fast_hit, st1 = renderer.render(d, rtc.Renderer.make_opts(precision=f.PRECISION_F32_FAST))
    gen, gen_hit, st2 = renderer.render(d, rtc.Renderer.make_opts(precision=f.PRECISION_F32_FAST, flags=f.OPT_F32_GENERAL_KERNEL))
    if n_spheres + 8 > 128:
        # (a fast-mode frame crosses PCIe as f32 and is widened on the host: the parity kernel's colours rounded to f32)
        assert np.array_equal(fast, want.astype(np.float32).astype(np.float64)) and np.array_equal(fast_hit, want_hit) and st1.rays == st.rays
    for img, hit in ((fast, fast_hit), (gen, gen_hit)):
        q = np.clip(np.rint(np.clip(img, 0.0, 1.0) * 255.0), 0, 255).astype(int)
        qw = np.clip(np.rint(np.clip(want, 0.0, 1.0) * 255.0), 0, 255).astype(int)
        ok = (np.abs(q - qw).max(axis=2) <= 1).mean()
        assert ok >= 0.999, f"{n_spheres} spheres: only {ok:.5f} of the pixels within 1/255 of the parity frame"
        assert np.array_equal(hit, want_hit)


@pytest.mark.parametrize("n_planes,n_spheres", [(1, 0), (0, 1), (1, 1), (2, 1), (1, 2), (3, 3), (3, 4), (2, 5), (5, 8), (15, 47), (16, 48)])
def test_small_scene_kernel_odd_and_even_shape_counts(rtc, orc, renderer, n_planes, n_spheres):
    """The small-scene kernel tests shapes in pairs and an odd count's last shape on its own, from the last slot of its
    constant-bank array (render_f32_small.cu): every parity of plane / sphere counts, up to the kernel's limits of 16 planes
    and 48 spheres, against the oracle — with glass and mirrors among the last shapes, which are the ones that move."""
    rnd = random.Random(1000 * n_planes + n_spheres)
    w = rtc.World(rtc.Light(rtc.Color(1, 1, 1), (-3.0, 9.0, -8.0)))
    for k in range(n_planes):
        m = rtc.Material(color=rtc.Color(rnd.random(), rnd.random(), rnd.random()), specular=0.3,
                         reflectiveness=0.4 if k == n_planes - 1 else 0.0)
        if k % 3 == 0:
            t = rtc.Matrix.identity().translation(0, -1.0 - 0.3 * k, 0)
        elif k % 3 == 1:
            t = rtc.Matrix.identity().rotation_x(math.pi / 2).translation(0, 0, 14.0 + k)
        else:
            t = rtc.Matrix.identity().rotation_z(math.pi / 2).translation(-9.0 - k, 0, 0)
        w.add_shape(rtc.Plane(t, m))
    for k in range(n_spheres):
        r = rnd.uniform(0.35, 0.9)
        last = k == n_spheres - 1
        m = rtc.Material(color=rtc.Color(rnd.random(), rnd.random(), rnd.random()), diffuse=0.7, specular=0.4,
                         reflectiveness=0.6 if last else 0.0, transparency=0.8 if last else 0.0, refractive_index=1.5 if last else 1.0)
        pos = (0.0, 0.6, 2.0) if last else (rnd.uniform(-6, 6), rnd.uniform(-0.5, 4), rnd.uniform(3, 11))
        w.add_shape(rtc.Sphere(rtc.Matrix.identity().scaling(r, r, r).translation(*pos), m))
    cam = rtc.Camera.new_with_transform(160, 100, 1.0, rtc.Matrix.make_view_transform((0, 1.5, -7.0), (0, 0.8, 3), (0, 1, 0)))
    for general in (False, True):
        _check32(rtc, orc, renderer, w, cam, general=general, min_ok=0.998, mask_tol=2e-3)
