"""Where does the end of a frame go?  Needs a library built with -DRTC_F32S_TRACE (set RTC_B200_LIB):
every warp of the f32 small-scene kernel records its start time, the time it found the tile counter
exhausted, its end time and its loop iterations (total / after exhaustion).
usage: RTC_B200_LIB=build/variants/lib_trace.so python tools/trace_tail.py [shard_count] [tile_rows] [flags]"""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import rtc_b200 as rtc
from rtc_b200 import ffi

n = int(sys.argv[1]) if len(sys.argv) > 1 else 8
tile = int(sys.argv[2]) if len(sys.argv) > 2 else 4
flags = int(sys.argv[3]) if len(sys.argv) > 3 else 0
r = rtc.Renderer([0])
r.upload(rtc.scenes.reflect_refract_scene())
cam = rtc.scenes.reflect_refract_camera(3840, 2160).flatten()
buf = torch.empty((2160, 3840, 3), dtype=torch.float32, device="cuda")
lib = ctypes.CDLL(ffi.LIB_PATH)
opts = rtc.Renderer.make_opts(precision=1, out_format=ffi.OUT_F32_RGB, shard_index=0, shard_count=n, tile_rows=tile, flags=flags)
for rep in range(3):
    st = ffi.RtcStats()
    r.render_device(cam, opts, buf.data_ptr(), None, None, st)
warps = 148 * 8 * 4   # upper bound of resident warps (7 blocks of 4 warps per SM)
raw = np.zeros(16384 * 6, dtype=np.uint64)
rc = lib.rtc_debug_read_trace(raw.ctypes.data_as(ctypes.c_void_p), ctypes.c_int(raw.size))
assert rc == 0, rc
t = raw[:16384 * 4].reshape(16384, 4)
last = raw[16384 * 4:].reshape(16384, 2)
ran = t[:, 0] > 0
t, last = t[ran], last[ran]
t0, t1, t2 = t[:, 0].astype(np.int64), t[:, 1].astype(np.int64), t[:, 2].astype(np.int64)
iters, tail = (t[:, 3] >> np.uint64(32)).astype(np.int64), (t[:, 3] & np.uint64(0xffffffff)).astype(np.int64)
base = t0.min()
us = lambda x: (x - base) / 1e3
print(f"kernel_ms {st.kernel_ms:.4f}  warps {len(t)}  rays {st.rays}")
print(f"start  : min 0 max {us(t0.max()):.1f} us")
seen = t1 > 0
print(f"exhaust: first {us(t1[seen].min()):.1f} median {np.median(us(t1[seen])):.1f} last {us(t1[seen].max()):.1f} us ({seen.sum()} warps saw it inside the loop)")
print(f"end    : p10 {np.percentile(us(t2), 10):.1f} median {np.median(us(t2)):.1f} p90 {np.percentile(us(t2), 90):.1f} p99 {np.percentile(us(t2), 99):.1f} max {us(t2.max()):.1f} us")
print(f"iters  : mean {iters.mean():.1f} max {iters.max()}   after exhaustion: mean {tail.mean():.1f} p90 {np.percentile(tail, 90):.0f} max {tail.max()}")
tail_us = (t2 - np.where(seen, t1, t2)) / 1e3
print(f"time after exhaustion per warp: mean {tail_us.mean():.1f} p90 {np.percentile(tail_us, 90):.1f} max {tail_us.max():.1f} us;  us/iter in tail {tail_us.sum() / max(tail.sum(), 1):.2f}")
hist, edges = np.histogram(us(t2), bins=12)
print("end-time histogram (us):", " ".join(f"{edges[i]:.0f}:{hist[i]}" for i in range(len(hist))))

# the warps that end last: which tile did they fetch last, when was it handed out, what does its centre pixel see?
tiles_x = (3840 + 7) // 8
rows = [r_ for r_ in range(2160) if (r_ // tile) % n == 0] if n > 1 else list(range(2160))
order_ = np.argsort(-t2)[:16]
print("latest warps: end us | iterations after exhaustion | last tile (tx, ty) -> frame pixel of its corner | hand-out rank of", int(st.kernel_launches), "launch(es)")
for w in order_:
    tid, rk = int(last[w, 0]), int(last[w, 1])
    ty_, tx_ = divmod(tid, tiles_x)
    fy = rows[min(ty_ * 4, len(rows) - 1)]
    print(f"  {us(t2[w]):7.1f} | {int(tail[w]):3d} | tile ({tx_:3d},{ty_:3d}) -> pixel ({tx_ * 8:4d},{fy:4d}) | rank {rk}")
