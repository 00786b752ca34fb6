"""Generate tests/golden/*.npz: frames of the ORACLE (the CPU restatement pinned by the reference's
77 unit-test known answers) on fixed scenes.  The reference itself cannot run here (Rust, no
toolchain), so these pin the oracle against regressions and give the GPU tests fixed vectors.
usage: python tools/make_golden.py"""
import math
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import orc  # noqa: E402
import rtc_b200 as rtc  # noqa: E402


def cases():
    yield "c1_room_64x32", rtc.scenes.reflect_refract_scene(), rtc.scenes.reflect_refract_camera(64, 32), rtc.ffi.RECT_FULL, 5
    yield "c1_room_64x32_render_rect", rtc.scenes.reflect_refract_scene(), rtc.scenes.reflect_refract_camera(64, 32), rtc.ffi.RECT_REFERENCE_RENDER, 5
    yield "c1_room_32x16_ss2x2", rtc.scenes.reflect_refract_scene(), rtc.scenes.reflect_refract_camera(32, 16, samples=0), rtc.ffi.RECT_FULL, 5
    cam = rtc.scenes.reflect_refract_camera(40, 20, samples=5)   # render_pixel's resample pass, offsets per include/rtc.h
    cam.aa_seed = 7
    yield "c1_room_40x20_aa5_seed7", rtc.scenes.reflect_refract_scene(), cam, rtc.ffi.RECT_FULL, 5
    yield "c1_room_48x24_depth2", rtc.scenes.reflect_refract_scene(), rtc.scenes.reflect_refract_camera(48, 24), rtc.ffi.RECT_FULL, 2
    w, cam = rtc.scenes.bench_scene()
    cam = rtc.Camera.new_with_transform(40, 30, math.pi / 3.0, rtc.Matrix.make_view_transform((0., 7., 0.), (0., 0., 0.), (1., 0., 0.)))
    yield "criterion_bench_40x30", w, cam, rtc.ffi.RECT_FULL, 5
    cam = rtc.Camera.new_with_transform(11, 11, math.pi / 2., rtc.Matrix.make_view_transform((0., 0., -5.), (0., 0., 0.), (0., 1., 0.)))
    yield "default_world_11x11", rtc.World.default(), cam, rtc.ffi.RECT_REFERENCE_RENDER, 5
    w, cam = rtc.scenes.cube_scene(7, hsize=60, vsize=42)
    yield "cube_room_60x42", w, cam, rtc.ffi.RECT_FULL, 5
    yield "synthetic_200_spheres_48x27", rtc.scenes.synthetic_scene(n_spheres=200), rtc.scenes.synthetic_camera(48, 27), rtc.ffi.RECT_FULL, 5


if __name__ == "__main__":
    out = os.path.join(ROOT, "tests", "golden")
    os.makedirs(out, exist_ok=True)
    only = sys.argv[1:]   # optional: names to (re)generate; default all
    for name, world, cam, rect, depth in cases():
        if only and name not in only:
            continue
        rgb, hit, cnt = orc.render(world, cam, rect=rect, max_depth=depth, threads=os.cpu_count() or 1)
        np.savez_compressed(os.path.join(out, name + ".npz"), rgb=rgb, hit=hit, rect=rect, depth=depth,
                            counters=np.array([cnt.rays, cnt.shadow_rays, cnt.shaded_hits], dtype=np.uint64))
        print(name, rgb.shape, "rays", cnt.rays, "mean", rgb.mean())
