"""Fixed cost of a render launch: kernel_ms of shard 0 of N of the 4K C3 frame (f32 small kernel) for
N = 1, 2, 4, 8, 16, 32 at depth 5 and depth 0; T = a + b * pixels separates the per-launch cost
(start-up + the tail of the longest per-pixel ray chain) from the throughput term."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import rtc_b200 as rtc
from rtc_b200 import ffi
r = rtc.Renderer([0])
r.upload(rtc.scenes.reflect_refract_scene())
cam = rtc.scenes.reflect_refract_camera(3840, 2160).flatten()
buf = torch.empty((2160, 3840, 3), dtype=torch.float32, device="cuda")
tile = int(sys.argv[1]) if len(sys.argv) > 1 else 16
flags = int(sys.argv[2]) if len(sys.argv) > 2 else 0
for depth in (5, 0):
    for n in (1, 2, 4, 8, 16, 32):
        opts = rtc.Renderer.make_opts(precision=1, out_format=ffi.OUT_F32_RGB, max_depth=depth, shard_index=0, shard_count=n, tile_rows=tile, flags=flags)
        ms = []
        for _ in range(8):
            st = ffi.RtcStats()
            r.render_device(cam, opts, buf.data_ptr(), None, None, st)
            ms.append(st.kernel_ms)
        rows = r.shard_rows(cam, opts)
        print(f"depth {depth} shards {n:2d} rows {rows:4d} kernel_ms min={min(ms):.4f} med={sorted(ms)[4]:.4f}  ideal={min_1 / n if n > 1 else min(ms):.4f}" if n > 1 else
              f"depth {depth} shards {n:2d} rows {rows:4d} kernel_ms min={min(ms):.4f} med={sorted(ms)[4]:.4f}")
        if n == 1:
            min_1 = min(ms)
