"""The oracle is pinned by the reference's own unit-test known answers."""
import subprocess

import numpy as np


def test_oracle_passes_every_reference_kat(orc):
    p = subprocess.run([orc.KAT_PATH], stdout=subprocess.PIPE, text=True)
    lines = p.stdout.strip().splitlines()
    assert p.returncode == 0, p.stdout
    assert lines[-1].startswith("KAT summary: 77 run, 0 failed"), lines[-1]
    assert not [l for l in lines if l.startswith("FAIL")]


def test_oracle_render1_through_flat_descriptors(orc, rtc):
    """test_render1 (src/camera.rs:216-231) again, but through World.flatten() -> orc_world_create:
    pins the descriptor path the GPU parity tests use."""
    import math
    world = rtc.World.default()
    cam = rtc.Camera.new_with_transform(11, 11, math.pi / 2.,
                                        rtc.Matrix.make_view_transform((0., 0., -5.), (0., 0., 0.), (0., 1., 0.)))
    rgb, hit, cnt = orc.render(world, cam, rect=rtc.ffi.RECT_REFERENCE_RENDER)
    assert np.allclose(rgb[5, 5], [0.38066, 0.47583, 0.2855], atol=1e-4)
    assert hit[5, 5] == 1
    assert (rgb[10, :] == 0).all() and (rgb[:, 10] == 0).all()  # Camera::render's off-by-one
    full, _, _ = orc.render(world, cam, rect=rtc.ffi.RECT_FULL)
    assert np.array_equal(full[:10, :10], rgb[:10, :10])


def test_oracle_c1_ray_count(orc, rtc):
    """C1: the reflect/refract room at 320x160 casts 650,692 rays and every
    primary ray hits (closed room)."""
    import os
    world = rtc.scenes.reflect_refract_scene()
    cam = rtc.scenes.reflect_refract_camera(320, 160)
    rgb, hit, cnt = orc.render(world, cam, threads=os.cpu_count() or 1)
    assert cnt.rays == 650692
    assert cnt.shadow_rays * 2 == cnt.rays
    assert hit.all()
    assert np.isfinite(rgb).all() and rgb.max() > 0.5


def test_oracle_threads_do_not_change_results(orc, rtc):
    world = rtc.scenes.reflect_refract_scene()
    cam = rtc.scenes.reflect_refract_camera(64, 32)
    a, ha, ca = orc.render(world, cam, threads=1)
    b, hb, cb = orc.render(world, cam, threads=4)
    assert np.array_equal(a, b) and np.array_equal(ha, hb) and ca.rays == cb.rays


def test_oracle_pixel_subset_matches_full(orc, rtc):
    world = rtc.scenes.reflect_refract_scene()
    cam = rtc.scenes.reflect_refract_camera(64, 32, samples=0)
    full, hit, _ = orc.render(world, cam)
    idx = orc.lattice(64, 32, step=4, n_random=50)
    sub, shit, _ = orc.render(world, cam, pixels=idx)
    assert np.array_equal(sub, full.reshape(-1, 3)[idx])
    assert np.array_equal(shit, hit.reshape(-1)[idx])
    assert set(np.unique(hit)) <= {15}
