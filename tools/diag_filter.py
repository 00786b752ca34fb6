"""Bounding-sphere filter on / off in the general f32 kernel: how many pixels differ, and by how much.
usage: python tools/diag_filter.py"""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import rtc_b200 as rtc
import test_gpu_parity as T
ffi = rtc.ffi
r = rtc.Renderer([0])
for name, flat, cam in (("c4", rtc.scenes.synthetic_scene().flatten(), rtc.scenes.synthetic_camera(640, 360)),
                        ("mixed", T._big_mixed_world(rtc).flatten(), rtc.scenes.synthetic_camera(480, 270))):
    r.upload_flat(*flat)
    d = cam.flatten()
    for depth in (0, 1, 5):
        out = []
        for flags in (0, ffi.OPT_NO_BOUND_FILTER, 0):
            buf = torch.empty((d.vsize, d.hsize, 3), dtype=torch.float32, device="cuda")
            st = ffi.RtcStats()
            r.render_device(d, rtc.Renderer.make_opts(precision=ffi.PRECISION_F32_FAST, out_format=ffi.OUT_F32_RGB, flags=flags, max_depth=depth),
                            buf.data_ptr(), None, None, st)
            out.append((buf, (st.rays, st.shaded_hits)))
        diff = (out[0][0] != out[1][0]).any(dim=2)
        rep = (out[0][0] != out[2][0]).any(dim=2)
        idx = diff.nonzero()[:5].tolist()
        print(name, "depth", depth, "pixels differing:", int(diff.sum()), "max", float((out[0][0] - out[1][0]).abs().max()),
              "repeat differs:", int(rep.sum()), "counters", out[0][1], out[1][1], "at", idx, flush=True)
