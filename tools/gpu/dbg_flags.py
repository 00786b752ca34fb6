import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch, ctypes as C
import rtc_b200 as rtc
from rtc_b200 import ffi
r = rtc.Renderer([0])
r.upload(rtc.scenes.reflect_refract_scene())
d = rtc.scenes.reflect_refract_camera(960, 540).flatten()
def status(tag):
    try:
        r.wait_result(); print(tag, "status ok", flush=True)
    except ffi.RtcError as e:
        print(tag, "STATUS", e, flush=True)
for flags in (0, 2, 0, 2):
    o = rtc.Renderer.make_opts(precision=ffi.PRECISION_F32_FAST, out_format=ffi.OUT_F32_RGB, flags=flags)
    got = torch.zeros((540, 960, 3), dtype=torch.float32, device="cuda")
    words = torch.zeros(64, dtype=torch.int32, device="cuda")
    one = torch.tensor([5], dtype=torch.int32).pin_memory()
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    torch.cuda.synchronize()
    status(f"flags={flags} before")
    t = time.perf_counter()
    r.render_device_synced(d, o, got.data_ptr(), s1.cuda_stream, wait_flag=words.data_ptr(), wait_value=5, done_flag=words.data_ptr() + 4, done_value=9)
    time.sleep(0.05)
    print("  query while waiting:", s1.query())
    with torch.cuda.stream(s2):
        words[0:1].copy_(one, non_blocking=True)
    s1.synchronize(); s2.synchronize()
    print("  done after %.1f ms, words[:2] = %s" % ((time.perf_counter() - t) * 1e3, words[:2].tolist()))
    status(f"flags={flags} after blocked call")
    r.render_device_synced(d, o, got.data_ptr(), s1.cuda_stream, wait_flag=words.data_ptr(), wait_value=3, done_flag=words.data_ptr() + 4, done_value=10)
    s1.synchronize()
    status(f"flags={flags} after satisfied call")
    r.wait_device(words.data_ptr(), 2, 4)
    status(f"flags={flags} after plain wait_device")
