"""HBM roofline of rtc_pack_rgba8_device (Canvas::to_imgbuf on the device): GB/s of algorithmic bytes
(24 or 12 read + 4 written per pixel) against MEASURED_PEAKS.json, 4K and 8K frames, gamma 1 and 2.2.
usage: python tools/pack_bench.py"""
import json, os, sys
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, root)
import torch
import rtc_b200 as rtc
from rtc_b200 import ffi
peak = json.load(open(os.path.join(root, "MEASURED_PEAKS.json"))).get("hbm_gbs") if os.path.exists(os.path.join(root, "MEASURED_PEAKS.json")) else None
r = rtc.Renderer([0])
side = torch.cuda.Stream()
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for (w, h) in ((3840, 2160), (7680, 4320)):
    n = w * h
    for dt, fmt, name in ((torch.float64, ffi.OUT_F64_RGB, "f64"), (torch.float32, ffi.OUT_F32_RGB, "f32")):
        src = torch.rand((n, 3), dtype=dt, device="cuda")
        out = torch.empty(n, dtype=torch.int32, device="cuda")
        for gamma in (1.0, 2.2):
            ms = []
            with torch.cuda.stream(side):
                for i in range(12):
                    flush.zero_()
                    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    a.record(side); r.pack_rgba8_device(src.data_ptr(), fmt, n, gamma, out.data_ptr(), side.cuda_stream); b.record(side)
                    side.synchronize()
                    if i >= 2: ms.append(a.elapsed_time(b))
            byts = n * (3 * src.element_size() + 4)
            t = min(ms)
            print(f"{w}x{h} {name} gamma {gamma}: {t * 1e3:7.1f} us  {byts / 1e6:7.1f} MB  {byts / t / 1e6:7.0f} GB/s" + (f"  = {100 * byts / t / 1e6 / peak:.0f} % of {peak:.0f}" if peak else ""))
