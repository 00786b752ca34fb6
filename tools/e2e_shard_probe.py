"""Per-rank breakdown of the sharded end-to-end step (rtc_render with shard opts into one frame in shared host memory).
usage: python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 --master-port 29512 tools/e2e_shard_probe.py"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist
from multiprocessing import shared_memory
import rtc_b200 as rtc
from rtc_b200 import ffi
rank, N, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("gloo")
W, H = 3840, 2160
r = rtc.Renderer([lr]); r.upload(rtc.scenes.reflect_refract_scene())
cam = rtc.scenes.reflect_refract_camera(W, H).flatten()
name = [None]
if rank == 0:
    shm = shared_memory.SharedMemory(create=True, size=H * W * 3 * 8); name[0] = shm.name
dist.broadcast_object_list(name, src=0)
if rank != 0:
    shm = shared_memory.SharedMemory(name=name[0])
bufs = {"shm": np.ndarray((H, W, 3), dtype=np.float64, buffer=shm.buf), "private": np.zeros((H, W, 3), dtype=np.float64)}
for label, fmt in (("f64 canvas", ffi.OUT_F64_RGB), ("f32 out", ffi.OUT_F32_RGB)):
    for where in ("shm", "private"):
        host = bufs[where] if fmt == ffi.OUT_F64_RGB else np.zeros((H, W, 3), dtype=np.float32)
        o = rtc.Renderer.make_opts(precision=1, out_format=fmt, shard_index=rank, shard_count=N, tile_rows=4)
        best = None
        for it in range(8):
            dist.barrier()
            t = time.perf_counter()
            _, _, st = r.render(cam, o, want_mask=False, out=host)
            dt = (time.perf_counter() - t) * 1e3
            if it >= 2 and (best is None or dt < best[0]): best = (dt, st.kernel_ms, st.d2h_ms, st.gather_ms)
        print(f"rank {rank}/{N} {label:10s} {where:8s} wall {best[0]:6.3f} kernel {best[1]:6.3f} after-kernel {best[2]:6.3f} host-deliver {best[3]:6.3f}", flush=True)
        dist.barrier()
del bufs
shm.close()
if rank == 0: shm.unlink()
