"""Classify f32-vs-oracle pixel mismatches of the C4 scene by the primary hit's shape kind/material.
usage (GPU box): python tools/diag_f32.py [w h]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import rtc_b200 as rtc, orc
w, h = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (192, 108)
world = rtc.scenes.synthetic_scene()
flat = world.flatten()
cam = rtc.scenes.synthetic_camera(w, h)
r = rtc.Renderer([0]); r.upload_flat(*flat)
for depth in (0, 1, 5):
    want, wh, cnt = orc.OracleWorld(flat=flat).render(cam, max_depth=depth, threads=os.cpu_count())
    got, gh, st = r.render(cam.flatten(), rtc.Renderer.make_opts(precision=1, out_format=rtc.ffi.OUT_F32_RGB, max_depth=depth))
    g64, _, _ = r.render(cam.flatten(), rtc.Renderer.make_opts(precision=0, max_depth=depth))
    q1 = orc.quantize_u8(got.astype(np.float64)).astype(int); q0 = orc.quantize_u8(want).astype(int)
    bad = np.abs(q1 - q0).max(axis=2) > 1
    print(f"depth {depth}: bad {bad.mean():.4f} ({bad.sum()} px), mask mismatches {(gh != wh).sum()}, rays f32 {st.rays} oracle {cnt.rays}, f64 maxerr {np.abs(g64-want).max():.2e}")
    ys, xs = np.nonzero(bad)
    for y, x in list(zip(ys, xs))[:12]:
        print("   ", x, y, "want", want[y, x].round(4), "got", got[y, x].round(4))
