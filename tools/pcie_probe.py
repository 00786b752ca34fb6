import torch, time
x = torch.empty(99532800, dtype=torch.uint8, device="cuda")
h = torch.empty(99532800, dtype=torch.uint8).pin_memory()
for n in (99532800, 12441600):
    for _ in range(3):
        torch.cuda.synchronize(); t=time.perf_counter(); h[:n].copy_(x[:n], non_blocking=True); torch.cuda.synchronize(); dt=time.perf_counter()-t
    print(f"D2H {n/1e6:.1f} MB: {dt*1e3:.3f} ms = {n/dt/1e9:.1f} GB/s")
for _ in range(3):
    torch.cuda.synchronize(); t=time.perf_counter(); x.copy_(h, non_blocking=True); torch.cuda.synchronize(); dt=time.perf_counter()-t
print(f"H2D: {dt*1e3:.3f} ms = {99532800/dt/1e9:.1f} GB/s")
import numpy as np, os
a = np.empty(99532800//4, dtype=np.float32); b = np.empty(99532800//4, dtype=np.float64)
t=time.perf_counter(); b[:] = a; dt=time.perf_counter()-t
print(f"single-thread widen: {dt*1e3:.1f} ms; cpus={os.cpu_count()}")
