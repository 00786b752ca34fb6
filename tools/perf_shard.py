"""kernel time of shard 0 of N of the 4K C3 frame (what one rank of an N-GPU job renders), plain launch on a
side stream.  usage: python tools/perf_shard.py [shards] [tile_rows] [flags]   (RTC_B200_LIB honoured)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import rtc_b200 as rtc
from rtc_b200 import ffi
n = int(sys.argv[1]) if len(sys.argv) > 1 else 8
tile = int(sys.argv[2]) if len(sys.argv) > 2 else 4
flags = int(sys.argv[3]) if len(sys.argv) > 3 else 0
r = rtc.Renderer([0]); r.upload(rtc.scenes.reflect_refract_scene())
cam = rtc.scenes.reflect_refract_camera(3840, 2160).flatten()
buf = torch.empty((2160, 3840, 3), dtype=torch.float32, device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
side = torch.cuda.Stream()
opts = rtc.Renderer.make_opts(precision=1, out_format=ffi.OUT_F32_RGB, shard_index=0, shard_count=n, tile_rows=tile, flags=flags)
ms = []
with torch.cuda.stream(side):
    for i in range(12):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(side); r.render_device(cam, opts, buf.data_ptr(), None, side.cuda_stream, None); e1.record(side)
        side.synchronize()
        if i >= 2: ms.append(e0.elapsed_time(e1))
print(f"shards={n} tile_rows={tile} flags={flags}: min={min(ms):.4f} med={sorted(ms)[len(ms) // 2]:.4f} ms  checksum={float(buf.double().sum()):.4f}")
