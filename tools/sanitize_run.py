"""One small frame through every kernel and every exchange entry point of the library — the program compute-sanitizer
runs (memcheck / racecheck / initcheck, SURVEY.md §5 "race detection").  Frames are tiny so that a sanitizer pass
takes seconds; each case is also compared with a second render of itself (bit-identical) so that a hazard which
changes a pixel shows up even when the tool does not flag it.
usage: compute-sanitizer --tool memcheck python tools/sanitize_run.py [--multi]
       python tools/sanitize_run.py --checked     (the pool refuses compute-sanitizer: the library built with every index
                                                   asserted, `make -C raytracer-challenge_b200/csrc checked`, and every device
                                                   frame between guard bands that must come back untouched)"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
if "--checked" in sys.argv:
    os.environ["RTC_B200_LIB"] = os.path.join(ROOT, "raytracer-challenge_b200", "csrc", "librtc_b200_checked.so")
import numpy as np
import torch

import rtc_b200 as rtc
from rtc_b200 import ffi

multi = "--multi" in sys.argv
r = rtc.Renderer([0])
room = rtc.scenes.reflect_refract_scene()
two = rtc.scenes.reflect_refract_scene()
two.add_light(rtc.Light(rtc.Color(0.3, 0.5, 0.9), (3.5, 3.0, -3.0)))
cube_world, cube_cam = rtc.scenes.cube_scene(7, hsize=48, vsize=32)
mid = rtc.scenes.synthetic_scene(n_spheres=1500, seed=0xB200)   # > 255 shapes: aliased ids, the shared-memory streams of the general kernels
big = rtc.scenes.synthetic_scene(n_spheres=7000, seed=0xB200)   # streams beyond 96 KB: global-memory streams, the hierarchy queries (kRefine kernels)


def both(world, cam, **kw):
    r.upload(world)
    d = cam.flatten()
    for prec in (ffi.PRECISION_F64_PARITY, ffi.PRECISION_F32_FAST):
        for fmt in (ffi.OUT_F64_RGB, ffi.OUT_F32_RGB, ffi.OUT_U8_RGB):
            o = rtc.Renderer.make_opts(precision=prec, out_format=fmt, **kw)
            a, ma, _ = r.render(d, o)
            b, mb, _ = r.render(d, o)
            assert np.array_equal(a, b) and np.array_equal(ma, mb), ("not reproducible", prec, fmt)


def guarded(world, cam, **kw):
    """rtc_render_device into device frames with a guard band of 4 rows of a sentinel on either side (and behind the hit
    mask): a store outside the frame shows up as a changed guard byte."""
    r.upload(world)
    d = cam.flatten()
    G = 4
    for prec in (ffi.PRECISION_F64_PARITY, ffi.PRECISION_F32_FAST):
        for fmt, dt in ((ffi.OUT_F64_RGB, torch.float64), (ffi.OUT_F32_RGB, torch.float32), (ffi.OUT_U8_RGB, torch.uint8)):
            for shards in (1, 3):
                for k in range(shards):
                    o = rtc.Renderer.make_opts(precision=prec, out_format=fmt, shard_index=k, shard_count=shards, tile_rows=4, **kw)
                    rows = r.shard_rows(d, o)
                    buf = torch.full((rows + 2 * G, d.hsize, 3), 77, dtype=dt, device="cuda")
                    msk = torch.full((rows + 2 * G, d.hsize), 77, dtype=torch.uint8, device="cuda")
                    r.render_device(d, o, buf[G:].data_ptr(), msk[G:].data_ptr())
                    torch.cuda.synchronize()
                    for t in (buf, msk):
                        assert bool((t[:G] == 77).all()) and bool((t[rows + G:] == 77).all()), ("guard band overwritten", prec, fmt, shards, k)
    print("guarded ok:", d.hsize, d.vsize, kw, flush=True)


if "--checked" in sys.argv:
    print("library:", ffi.LIB_PATH, flush=True)
    assert ffi.LIB_PATH.endswith("_checked.so")
guarded(room, rtc.scenes.reflect_refract_camera(64, 36))
guarded(room, rtc.scenes.reflect_refract_camera(33, 17, samples=0))
guarded(room, rtc.scenes.reflect_refract_camera(40, 21, samples=5))
guarded(room, rtc.scenes.reflect_refract_camera(64, 36), flags=ffi.OPT_F32_GENERAL_KERNEL)
guarded(two, rtc.scenes.reflect_refract_camera(48, 24))
guarded(cube_world, cube_cam)
guarded(mid, rtc.scenes.synthetic_camera(48, 27))
guarded(mid, rtc.scenes.synthetic_camera(48, 27), flags=ffi.OPT_F32_GENERAL_KERNEL)   # the general f32 kernel with 1500 shapes in shared memory (default: parity kernel)
guarded(big, rtc.scenes.synthetic_camera(48, 27))
guarded(big, rtc.scenes.synthetic_camera(48, 27), flags=ffi.OPT_F32_GENERAL_KERNEL)   # the kRefine form of the general f32 kernel (default: parity kernel)
guarded(big, rtc.scenes.synthetic_camera(7, 5), flags=ffi.OPT_F32_GENERAL_KERNEL)     # fewer rays than a warp has lanes
guarded(big, rtc.scenes.synthetic_camera(24, 14), flags=ffi.OPT_NO_BOUND_FILTER)
guarded(room, rtc.scenes.reflect_refract_camera(1024, 512))   # tile scheduling pre-pass (>= 4096 tiles)
both(room, rtc.scenes.reflect_refract_camera(64, 36))
both(room, rtc.scenes.reflect_refract_camera(33, 17, samples=0))
both(room, rtc.scenes.reflect_refract_camera(40, 20, samples=5))
both(room, rtc.scenes.reflect_refract_camera(64, 36), flags=ffi.OPT_F32_GENERAL_KERNEL)
both(two, rtc.scenes.reflect_refract_camera(48, 24))
both(cube_world, cube_cam)
both(mid, rtc.scenes.synthetic_camera(48, 27))
both(mid, rtc.scenes.synthetic_camera(48, 27), flags=ffi.OPT_F32_GENERAL_KERNEL)
both(big, rtc.scenes.synthetic_camera(48, 27))
both(big, rtc.scenes.synthetic_camera(48, 27), flags=ffi.OPT_F32_GENERAL_KERNEL)
both(big, rtc.scenes.synthetic_camera(32, 18), flags=ffi.OPT_NO_BOUND_FILTER)
# a 4K-sized tile count takes the tile scheduling pre-pass (classify_tiles_small_kernel): one larger frame
r.upload(room)
d = rtc.scenes.reflect_refract_camera(1024, 512).flatten()
o = rtc.Renderer.make_opts(precision=ffi.PRECISION_F32_FAST, out_format=ffi.OUT_F32_RGB)
a, _, _ = r.render(d, o, want_mask=False, want_stats=False)
b, _, _ = r.render(d, o, want_mask=False, want_stats=False)
assert np.array_equal(a, b)
# shards: compact + assemble, full-frame destination, flags, pack, download
d = rtc.scenes.reflect_refract_camera(96, 50).flatten()
whole = torch.zeros((50, 96, 3), dtype=torch.float32, device="cuda")
r.render_device(d, o, whole.data_ptr())
o0 = rtc.Renderer.make_opts(precision=ffi.PRECISION_F32_FAST, out_format=ffi.OUT_F32_RGB, shard_index=0, shard_count=3, tile_rows=4)
stride = r.shard_rows(d, o0)
gathered = torch.zeros((3, stride, 96, 3), dtype=torch.float32, device="cuda")
full = torch.zeros_like(whole)
for k in range(3):
    r.render_device(d, rtc.Renderer.make_opts(precision=ffi.PRECISION_F32_FAST, out_format=ffi.OUT_F32_RGB, shard_index=k, shard_count=3, tile_rows=4),
                    gathered[k].data_ptr())
    r.render_device(d, rtc.Renderer.make_opts(precision=ffi.PRECISION_F32_FAST, out_format=ffi.OUT_F32_RGB, shard_index=k, shard_count=3, tile_rows=4,
                                              flags=ffi.OPT_DST_FULL_FRAME), full.data_ptr())
frame = torch.zeros_like(whole)
r.assemble_device(d, o0, gathered.data_ptr(), stride, ffi.OUT_F32_RGB, frame.data_ptr())
torch.cuda.synchronize()
assert torch.equal(frame, whole) and torch.equal(full, whole)
flags = torch.zeros(64, dtype=torch.int32, device="cuda")
r.signal_device(flags.data_ptr() + 4, 7)
r.wait_device(flags.data_ptr() + 4, 1, 7)
torch.cuda.synchronize()
out = torch.zeros(50 * 96, dtype=torch.int32, device="cuda")
r.pack_rgba8_device(whole.data_ptr(), ffi.OUT_F32_RGB, 50 * 96, 2.2, out.data_ptr())
got = r.download_frame(whole.data_ptr(), ffi.OUT_F32_RGB, 96, 50, np.empty((50, 96, 3), np.float64))
assert np.array_equal(got, whole.cpu().numpy().astype(np.float64))
if multi and torch.cuda.device_count() >= 2:
    r2 = rtc.Renderer([0, 1])
    r2.upload(room)
    d = rtc.scenes.reflect_refract_camera(128, 70).flatten()
    for prec in (0, 1):
        o = rtc.Renderer.make_opts(precision=prec, tile_rows=4)
        a, ma, _ = r2.render(d, o)
        b, mb, _ = r.render(d, o)
        assert np.array_equal(a, b) and np.array_equal(ma, mb), "2-GPU context frame differs from the 1-GPU frame"
print("sanitize_run: ok")
