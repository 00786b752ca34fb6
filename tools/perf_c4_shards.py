"""C4 (10k shapes, 4K) sharded 8 ways on ONE GPU: kernel time of every shard for several tile heights — which part of the
rank-to-rank spread of the 8-GPU bench (11.3-15.1 ms) is the work each shard carries.  usage: python tools/perf_c4_shards.py [precision]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import rtc_b200 as rtc
from rtc_b200 import ffi
prec = int(sys.argv[1]) if len(sys.argv) > 1 else 1
r = rtc.Renderer([0]); r.upload_flat(*rtc.scenes.synthetic_scene().flatten())
cam = rtc.scenes.synthetic_camera(3840, 2160).flatten()
buf = torch.empty((2160, 3840, 3), dtype=torch.float64, device="cuda")
for tile in (4, 8, 12, 20, 28):
    ms = []
    for k in range(8):
        o = rtc.Renderer.make_opts(precision=prec, out_format=ffi.OUT_F32_RGB, shard_index=k, shard_count=8, tile_rows=tile)
        best = 1e9
        for _ in range(2):
            st = ffi.RtcStats()
            r.render_device(cam, o, buf.data_ptr(), None, None, st)
            best = min(best, st.kernel_ms)
        ms.append(best)
    print(f"tile_rows {tile:2d}: shards " + " ".join(f"{m:6.2f}" for m in ms) + f"   max {max(ms):6.2f}  mean {sum(ms) / 8:6.2f}  max/mean {max(ms) / (sum(ms) / 8):.3f}", flush=True)
