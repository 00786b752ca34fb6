#!/usr/bin/env python
"""bench.py — Mrays/s and ms/frame of the Camera::render hot path on the 4K reflect/refract scene.

  python bench.py --gpus N --steps K --warmup W            (N>1: launched by torchrun, one rank per GPU)
  python bench.py --impl reference ...                     (the reference's CPU path: the oracle port)

A "step" is one frame: config C3 of BASELINE.json — the chapter-11 reflect/refract room
(jamis.yml) at 3840x2160, 1 sample per pixel, recursion depth 5 — rendered by the f32 fast kernel
(the f64 parity kernel is timed beside it and reported under "f64").  A "ray" is one
World::intersect call (primary + reflected + refracted + shadow); the per-frame ray
count is deterministic and taken from the f64 parity kernel, whose counters the GPU tests pin to
the oracle's.  Prints ONE JSON line on rank 0.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOAD = "C3: reflect-refract room (jamis.yml, 13 shapes), 3840x2160, 1 spp, depth 5"
W, H, DEPTH, SAMPLES = 3840, 2160, 5, 1   # the metric's configuration (BASELINE.json configs[2]); --workload c5 switches to configs[4]
N_SPHERES, N_PLANES = 7, 6
FLOP_PER_RAY = 54 * N_SPHERES + 13 * N_PLANES   # algorithmic flops per World::intersect
WORKLOAD_KEY = "c3"


def make_scene(rtc):
    return rtc.scenes.synthetic_scene() if WORKLOAD_KEY == "c4" else rtc.scenes.reflect_refract_scene()


def make_camera(rtc):
    if WORKLOAD_KEY == "c4":
        return rtc.scenes.synthetic_camera(W, H, samples=SAMPLES)
    return rtc.scenes.reflect_refract_camera(W, H, samples=SAMPLES)
FLOP_PER_SHADED_HIT = 150
# Executed-instruction figures of the timed kernel from the kept `ncu --set full` capture (a number measured under a
# profiler is never a bench value: these describe the kernel, the timings above are CUDA events).  Refreshed per round.
HW_PROFILE = {
    "c3": {"source": "profiles/r02_ncu_f32_small_c3.txt", "fma_pipe_pct": 42.1, "alu_pipe_pct": 42.6, "issue_active_pct": 86.9,
           "lanes_per_instruction": 28.53, "registers": 72, "dram_bytes_per_launch": 66.83e6,
           "dram_note": "65.7 MB written + 1.2 MB read per 4K launch; the frame is 99.5 MB, the rest was still in the 126 MB L2 when the kernel ended"},
    "c4": {"source": "profiles/r02_ncu_f64_c4_960x540.txt (the parity kernel, which fast mode takes for this scene; a 960x540 frame)",
           "fp64_pipe_pct": 9.0, "fma_pipe_pct": 23.7, "alu_pipe_pct": 21.7, "issue_active_pct": 54.0, "lanes_per_instruction": 8.02,
           "registers": 64, "dram_bytes_per_launch": None},
}
FP32_FORMULA_PEAK = 148 * 128 * 2 * 1.965e9 / 1e12  # TFLOP/s, non-tensor, at clocks.max.sm


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        self.rows, self.proc, self.gpu = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.gpu)], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [c.strip() for c in line.split(",")]))

    def stop(self, t0, t1):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        rows = [r for t, r in self.rows if t0 - 0.05 <= t <= t1 + 0.15 and len(r) >= 9] or [r for _, r in self.rows if len(r) >= 9]
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        sm = sorted(float(r[1]) for r in rows)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in rows for n, v in zip(names, r[5:9]) if v.lower().startswith("active")})
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(rows[0][2]), "reasons": reasons, "samples": len(rows),
                "power_w_max": max(float(r[3]) for r in rows)}


def measured_peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        return {}


def cpu_info():
    model = "unknown"
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                model = line.split(":", 1)[1].strip()
                break
    except Exception:
        pass
    return model, os.cpu_count() or 1


def oracle_sample(threads, target_seconds=12.0):
    """Times the oracle (the CPU restatement of the reference's Camera::render path) on a bounded
    sample of the SAME workload: a pixel lattice of the 4K frame sized to ~target_seconds of CPU work."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import numpy as np
    import orc
    import rtc_b200 as rtc
    world = make_scene(rtc)
    cam = make_camera(rtc)
    ow = orc.OracleWorld(world)
    # calibrate on a coarse lattice, then size the real sample
    idx = orc.lattice(W, H, step=64, n_random=0)
    _, _, c0 = ow.render(cam, threads=threads, pixels=idx)
    rate = c0.rays / max(c0.seconds, 1e-6)
    rays_per_px = c0.rays / idx.size
    n_px = int(min(W * H, max(idx.size, rate * target_seconds / rays_per_px)))
    step = max(1, int((W * H / n_px) ** 0.5))
    idx = orc.lattice(W, H, step=step, n_random=0)
    _, _, c = ow.render(cam, threads=threads, pixels=idx)
    return {"mrays_per_s": c.rays / c.seconds / 1e6, "rays": int(c.rays), "pixels": int(idx.size), "seconds": c.seconds,
            "ms_per_frame_extrapolated": (W * H / idx.size) * c.seconds * 1e3,
            "sample": (f"the whole {W}x{H} frame" if step == 1 else f"every {step}th pixel in x and y of the {W}x{H} frame") +
                      f" ({idx.size} px, {c.rays} rays, {c.seconds:.1f} s)"}


def arm_config(N, TILE):
    """`config` of the JSON line: the workload, identical in both arms (the driver compares the two dicts); what differs
    between the arms — how each is timed, the frame's work counts — sits beside it (`timing_note`, `work`)."""
    return {"workload": WORKLOAD, "n_gpus": N,
            "parallelism": f"interleaved {TILE}-row tiles over {N} GPU(s)" if N > 1 else "one GPU, whole frame",
            "l2": "inputs are KBs, the output frame is written once; a 256 MB buffer is rewritten between timed GPU iterations (outside the events)"}


def oracle_frame_check(frames, step=32, n_random=2048):
    """The oracle as the CHECKER of the frames the bench just timed (outside every timed region): a pixel lattice of the
    workload's frame, f64 parity frame within 1e-9, f32 fast frame within 1/255 after the 8-bit rule on >= 99.9 %."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import numpy as np
    import orc
    import rtc_b200 as rtc
    world, cam = make_scene(rtc), make_camera(rtc)
    idx = orc.lattice(W, H, step=step, n_random=n_random)
    want, want_hit, _ = orc.OracleWorld(world).render(cam, threads=os.cpu_count() or 1, pixels=idx, max_depth=DEPTH)
    out = {"lattice": f"every {step}th pixel in x and y + {n_random} random pixels ({idx.size} px) against the CPU oracle"}
    if "f64" in frames:
        out["f64_max_abs_err"] = float(np.abs(frames["f64"].reshape(-1, 3)[idx] - want).max())
        out["f64_within_1e-9"] = bool(out["f64_max_abs_err"] <= 1e-9)
    if "f32" in frames:
        q_got = orc.quantize_u8(frames["f32"].reshape(-1, 3)[idx].astype(np.float64)).astype(int)
        q_want = orc.quantize_u8(want).astype(int)
        out["f32_fraction_within_1_of_255"] = float((np.abs(q_got - q_want).max(axis=1) <= 1).mean())
        out["f32_meets_99.9pct"] = bool(out["f32_fraction_within_1_of_255"] >= 0.999)
    return out


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path.  The Rust crate cannot
    be built here (no cargo/rustc), so this is the oracle port — a line-by-line C++
    restatement pinned by the reference's unit-test known answers — on all host threads
    (= Camera::render_async's rayon pool, src/camera.rs:144-160)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    model, cores = cpu_info()
    per_step = max(1.0, min(15.0, 150.0 / max(1, args.steps + args.warmup)))
    for _ in range(args.warmup):
        oracle_sample(cores, per_step)
    runs = [oracle_sample(cores, per_step) for _ in range(args.steps)]
    rays = sum(r["rays"] for r in runs)
    secs = sum(r["seconds"] for r in runs)
    value = rays / secs / 1e6
    line = {
        "impl": "reference", "metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": secs / max(1, args.steps) * 1e3,
        "ms_per_frame_extrapolated": sum(r["ms_per_frame_extrapolated"] for r in runs) / len(runs),
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": arm_config(args.gpus, args.tile_rows),
        "timing_note": "host wall clock around the oracle's Camera::render_pixel loop over the sample named in cpu_baseline.sample, all host threads",
        "cpu_baseline": {"value": value, "unit": "Mrays/s", "cores": cores, "kind": "port", "cpu": model,
                         "sample": runs[-1]["sample"]},
        "e2e": {"value": value, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


def run_gpu(args):
    import numpy as np
    import torch
    import torch.distributed as dist
    import rtc_b200 as rtc
    from rtc_b200 import ffi
    from rtc_b200.sharded import ShardedFrame

    rank = int(os.environ.get("RANK", "0"))
    world_size = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    N = world_size
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the render path has no CPU fallback)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if N > 1:
        # stdout carries exactly one JSON line: whatever NCCL prints while its communicators come up (the version
        # banner under NCCL_DEBUG) is sent to stderr by pointing fd 1 at fd 2 until the first collective is done
        sys.stdout.flush()
        saved_stdout = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=dev)
            warm = torch.zeros(1, device=dev)
            dist.all_reduce(warm)
            dist.barrier()
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved_stdout, 1)
            os.close(saved_stdout)
    r = rtc.Renderer([local_rank])
    world = make_scene(rtc)
    cam = make_camera(rtc)
    flat = world.flatten()
    camd = cam.flatten()
    r.upload_flat(*flat)
    TILE = args.tile_rows
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2
    # Everything timed runs on ONE non-default stream: the render kernels (rtc_render_device gets the raw
    # cudaStream_t), the L2 flush, NCCL's stream dependencies and the CUDA events.  (A NULL stream would
    # mean "the context's own stream + host synchronisation" to rtc_render_device.)
    side = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(side)
    stream = side.cuda_stream
    assert stream != 0
    token = torch.zeros(1, dtype=torch.int32, device=dev)

    def sync_all():
        if N > 1:
            dist.barrier()
        torch.cuda.synchronize()

    class Path:
        """One (precision, output format, exchange mode) way of producing the frame on rank 0."""

        def __init__(self, precision, fmt, mode, signal="flags"):
            self.sf = ShardedFrame(W, H, fmt, rank, N, dev, tile_rows=TILE, mode=mode, signal=signal)
            self.sf.connect(r)
            self.opts = self.sf.shard_opts(rtc.Renderer.make_opts, precision=precision, rect=ffi.RECT_FULL, max_depth=DEPTH)
            self.fmt = fmt

        def dst(self):
            d = self.sf.destination()
            return d if isinstance(d, int) else d.data_ptr()

        def kernel_only(self):
            r.render_device(camd, self.opts, self.dst(), None, stream)

        def frame(self):
            """render this rank's interleaved row tiles, then the exchange step that completes rank 0's frame"""
            if N == 1:
                r.render_device(camd, self.opts, self.dst(), None, stream)
                return
            self.sf.render(r, camd, self.opts, stream)   # p2p: the other ranks' kernels wait for 'previous frame consumed' and store their flag themselves
            if self.sf.mode == "nccl":
                self.sf.exchange(lambda g, f: r.assemble_device(camd, self.opts, g.data_ptr(), self.sf.stride_rows, self.fmt, f.data_ptr(), stream))
            else:
                # completion flags over NVLink: no collective (pixels are already in rank 0's frame); rank 0 has no consumer here, so
                # its wait kernel frees the frame for the next one itself (back-pressure flag)
                self.sf.exchange(stream=stream, release=True)

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        sync_all()
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
        t0 = time.time()
        for a, b in ev:
            flush.zero_()            # L2 flush between timed iterations, outside the event pair
            if N > 1:
                dist.all_reduce(token)   # line the ranks up before the step starts (outside the event pair)
            a.record()
            fn()
            b.record()
        sync_all()
        t1 = time.time()
        ms = torch.tensor([a.elapsed_time(b) for a, b in ev], dtype=torch.float64, device=dev)
        if N > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)   # max over ranks, per step
        return ms.cpu().numpy(), t0, t1

    # deterministic work counts of the frame, from the f64 parity kernel's counters (== the oracle's)
    o64c = rtc.Renderer.make_opts(precision=ffi.PRECISION_F64_PARITY, rect=ffi.RECT_FULL, max_depth=DEPTH, out_format=ffi.OUT_F64_RGB,
                                  shard_index=rank, shard_count=N, tile_rows=TILE)
    o32c = rtc.Renderer.make_opts(precision=ffi.PRECISION_F32_FAST, rect=ffi.RECT_FULL, max_depth=DEPTH, out_format=ffi.OUT_F32_RGB,
                                  shard_index=rank, shard_count=N, tile_rows=TILE)
    my_rows = r.shard_rows(camd, o32c)
    scratch = torch.empty((r.shard_rows(camd, o64c), W, 3), dtype=torch.float64, device=dev)
    st = ffi.RtcStats()
    r.render_device(camd, o64c, scratch.data_ptr(), None, None, st)
    counts = torch.tensor([st.rays, st.shaded_hits, st.shadow_rays, st.primary_hits], dtype=torch.float64, device=dev)
    st32 = ffi.RtcStats()
    r.render_device(camd, o32c, scratch.data_ptr(), None, None, st32)
    counts32 = torch.tensor([st32.rays], dtype=torch.float64, device=dev)
    if N > 1:
        dist.all_reduce(counts)
        dist.all_reduce(counts32)
    rays, shaded = float(counts[0]), float(counts[1])
    flops = rays * FLOP_PER_RAY + shaded * FLOP_PER_SHADED_HIT
    del scratch

    main = Path(ffi.PRECISION_F32_FAST, ffi.OUT_F32_RGB, "p2p")
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ms32, t0, t1 = timed(main.frame, args.steps, max(3, args.warmup))
    clocks = sampler.stop(t0, t1) if rank == 0 else None
    total_ms = float(ms32.sum())
    ms_per_step = total_ms / args.steps
    value = rays * args.steps / (total_ms * 1e-3) / 1e6
    # kernel alone (this rank's shard), for the roofline
    k32, _, _ = timed(main.kernel_only, args.steps, 3)
    kernel_ms = float(k32.mean())
    extra = {}
    frame_check = None
    if N > 1:
        # same frame through the NCCL gather + de-interleave, and a bit-exact check of both assembled
        # frames against rank 0's own unsharded render
        alt = Path(ffi.PRECISION_F32_FAST, ffi.OUT_F32_RGB, "nccl")
        msn, _, _ = timed(alt.frame, args.steps, 3)
        # kernel alone with a LOCAL destination (compact shard buffer) vs the peer-mapped frame, per rank
        def per_rank(fn):
            for _ in range(3):
                fn()
            sync_all()
            evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
            for a, b in evs:
                flush.zero_(); dist.all_reduce(token); a.record(); fn(); b.record()
            sync_all()
            mine = torch.tensor([sum(a.elapsed_time(b) for a, b in evs) / len(evs)], dtype=torch.float64, device=dev)
            allr = [torch.zeros_like(mine) for _ in range(N)]
            dist.all_gather(allr, mine)
            return [round(float(x), 4) for x in allr]
        extra["kernel_ms_per_rank"] = {"peer_frame_destination": per_rank(main.kernel_only), "local_destination": per_rank(alt.kernel_only)}
        extra["gather_nccl"] = {"ms_per_frame": float(msn.mean()), "mrays_per_s": rays / (float(msn.mean()) * 1e-3) / 1e6,
                                "path": "compact shard buffers, dist.gather (NCCL send/recv) to rank 0, rtc_assemble_device"}
        ar = Path(ffi.PRECISION_F32_FAST, ffi.OUT_F32_RGB, "p2p", signal="allreduce")
        msa, _, _ = timed(ar.frame, args.steps, 3)
        extra["p2p_allreduce_signal"] = {"ms_per_frame": float(msa.mean()), "path": "peer stores + 4-byte NCCL all-reduce as completion signal"}
        ar.sf.close()
        main.frame(); alt.frame(); sync_all()
        if rank == 0:
            ref = torch.empty((H, W, 3), dtype=torch.float32, device=dev)
            r.render_device(camd, rtc.Renderer.make_opts(precision=ffi.PRECISION_F32_FAST, rect=ffi.RECT_FULL, max_depth=DEPTH,
                                                         out_format=ffi.OUT_F32_RGB), ref.data_ptr(), None, None, None)
            torch.cuda.synchronize()
            frame_check = {"p2p_equals_1gpu_frame": bool(torch.equal(main.sf.frame, ref)),
                           "nccl_equals_1gpu_frame": bool(torch.equal(alt.sf.frame, ref))}
        alt.sf.close()
        del alt
    n64 = max(1, min(args.steps, 5))
    p64 = Path(ffi.PRECISION_F64_PARITY, ffi.OUT_F64_RGB, "p2p")
    ms64, _, _ = timed(p64.frame, n64, 3)
    k64, _, _ = timed(p64.kernel_only, n64, 3)

    # ---- e2e: the public API with HOST buffers: scene H2D + render + Canvas (f64) D2H every step ----
    oe = rtc.Renderer.make_opts(precision=ffi.PRECISION_F32_FAST, rect=ffi.RECT_FULL, max_depth=DEPTH, out_format=ffi.OUT_F64_RGB,
                                shard_index=rank, shard_count=N, tile_rows=TILE)
    oe64 = rtc.Renderer.make_opts(precision=ffi.PRECISION_F64_PARITY, rect=ffi.RECT_FULL, max_depth=DEPTH, out_format=ffi.OUT_F64_RGB,
                                  shard_index=rank, shard_count=N, tile_rows=TILE)
    h2d = C.sizeof(ffi.RtcShape) * flat[2] + C.sizeof(ffi.RtcLight) + C.sizeof(ffi.RtcCamera)
    e2e_steps = max(1, min(args.steps, 10))
    # One Canvas frame (f64, Canvas.pixels' layout).  N = 1: a numpy array; N > 1: host memory all rank processes map (POSIX
    # shared memory created by rank 0) — every rank's rtc_render fills in the rows of its shard over its own PCIe link; no
    # gather, no collective on the data path.  The frame is registered (rtc_register_host_buffer), so the library may deliver
    # by DMA straight into it; h2d / d2h bytes are per rank.
    from rtc_b200.sharded import SharedHostFrame
    shared = SharedHostFrame(W, H, rank, N, numa_interleave=args.numa_interleave)
    host = shared.array
    shared.register(r)

    def e2e_run(opts, steps):
        def step():
            r.upload_flat(*flat)
            r.render(camd, opts, want_mask=False, out=host, want_stats=False)
        for _ in range(2):
            step()
        sync_all()
        ts = []
        for _ in range(steps):
            sync_all()
            a = time.perf_counter()
            step()
            torch.cuda.synchronize()
            ts.append(time.perf_counter() - a)
        t = torch.tensor(ts, dtype=torch.float64, device=dev)
        if N > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)   # per step: the slowest rank
        sync_all()
        return t

    def with_flags(o, fl):
        o2 = ffi.RtcRenderOpts.from_buffer_copy(bytes(o))
        o2.flags |= fl
        return o2

    # fast mode, the library's own delivery choice: N = 1 mixes — most row bands cross the one PCIe link as f32 and are widened
    # by the host threads, the last few go as f64 by DMA straight into the frame, so that link and host threads finish together;
    # N > 1 DMA-writes f64 rows straight into the registered frame (no host pass)
    e2e_t = e2e_run(oe, e2e_steps)
    e2e_ms = float(e2e_t.mean()) * 1e3
    direct_default = N > 1
    delivery = r.delivery_info()          # bytes the last frame moved device -> host on this rank, bands and bands sent direct
    d2h = delivery["d2h_bytes"]
    e2e_frame_ok = None
    if N > 1 and rank == 0:
        ref = torch.empty((H, W, 3), dtype=torch.float32, device=dev)
        r.render_device(camd, rtc.Renderer.make_opts(precision=ffi.PRECISION_F32_FAST, rect=ffi.RECT_FULL, max_depth=DEPTH,
                                                     out_format=ffi.OUT_F32_RGB), ref.data_ptr(), None, None, None)
        torch.cuda.synchronize()
        ref_host = ref.cpu().numpy().astype(np.float64)
        del ref
        e2e_frame_ok = bool(np.array_equal(host, ref_host))
    # the other delivery for comparison, and parity mode end to end (direct: 24 B/pixel cross PCIe either way)
    alt_t = e2e_run(with_flags(oe, ffi.OPT_DELIVER_STAGED if direct_default else ffi.OPT_DELIVER_DIRECT), max(1, e2e_steps // 2))
    staged_t = e2e_run(with_flags(oe, ffi.OPT_DELIVER_STAGED), max(1, e2e_steps // 2)) if N == 1 else None
    e64_t = e2e_run(oe64, max(1, min(e2e_steps, 3)))
    frames = {}
    if N == 1:
        frames["f64"] = host.copy()                       # the parity frame just delivered end to end
        e2e_run(oe, 1)
        frames["f32"] = host.copy()
    # ---- the drop-in's own multi-GPU path: ONE process, rtc_context_create over the N GPUs, one rtc_render call ----
    single = None
    if N > 1:
        # the other rank processes must leave their GPUs IDLE meanwhile: an NCCL barrier is a kernel spinning on every GPU,
        # which the driver would time-slice against rank 0's kernels there; they wait on the CPU (gloo) instead
        cpu_group = dist.new_group(backend="gloo")
        torch.cuda.synchronize()
        dist.barrier(group=cpu_group)
        if rank == 0:
            try:
                rs = rtc.Renderer(list(range(N)))
                rs.upload_flat(*flat)
                rs.register_host_buffer(host)
                osp = rtc.Renderer.make_opts(precision=ffi.PRECISION_F32_FAST, rect=ffi.RECT_FULL, max_depth=DEPTH, out_format=ffi.OUT_F64_RGB, tile_rows=TILE)
                ts = []
                for i in range(2 + e2e_steps):
                    host[0, 0, 0] = -1.0
                    a = time.perf_counter()
                    rs.upload_flat(*flat)
                    rs.render(camd, osp, want_mask=False, out=host, want_stats=False)
                    ts.append(time.perf_counter() - a)
                ts = ts[2:]
                single = {"ms_per_frame": sum(ts) / len(ts) * 1e3, "ms_per_frame_min": min(ts) * 1e3, "steps": len(ts),
                          "value": rays / (sum(ts) / len(ts)) / 1e6, "unit": "Mrays/s", "frame_equals_1gpu_frame": bool(np.array_equal(host, ref_host)),
                          "path": f"one process: rtc_context_create over {N} GPUs, rtc_scene_upload + one rtc_render into the registered f64 Canvas frame "
                                  "(one host thread per GPU, interleaved row tiles, pitched DMA per GPU over its own PCIe link); the other rank "
                                  "processes idle at a barrier meanwhile"}
                dfr = torch.empty((H, W, 3), dtype=torch.float32, device=dev)
                osd = rtc.Renderer.make_opts(precision=ffi.PRECISION_F32_FAST, rect=ffi.RECT_FULL, max_depth=DEPTH, out_format=ffi.OUT_F32_RGB, tile_rows=TILE)
                kms = []
                for i in range(3 + e2e_steps):
                    st1 = ffi.RtcStats()
                    rs.render_device(camd, osd, dfr.data_ptr(), None, None, st1)
                    kms.append(st1.kernel_ms)
                single["device_frame_ms"] = sum(kms[3:]) / len(kms[3:])
                single["device_frame_equals_1gpu_frame"] = bool(np.array_equal(dfr.cpu().numpy().astype(np.float64), ref_host))
                rs.unregister_host_buffer(host)
                rs.close()
            except Exception as ex:   # reported, never hidden
                single = {"error": repr(ex)}
        dist.barrier(group=cpu_group)
        sync_all()

    if rank == 0:
        peaks = measured_peaks()
        p32 = r.measure_peak_tflops(ffi.PRECISION_F32_FAST)
        pk64 = r.measure_peak_tflops(ffi.PRECISION_F64_PARITY)
        my_share = my_rows / H
        achieved = flops * my_share / (kernel_ms * 1e-3) / 1e12
        hw = HW_PROFILE.get(WORKLOAD_KEY) if N == 1 else None
        c4 = WORKLOAD_KEY == "c4"
        line = {
            "metric": "Mrays/s", "value": value, "unit": "Mrays/s", "n_gpus": N, "steps": args.steps, "warmup": max(3, args.warmup),
            "ms_per_step": ms_per_step, "ms_per_frame": ms_per_step, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": arm_config(N, TILE),
            "timing_note": "CUDA events per step on the launch stream, max over ranks" +
                           ("; the render kernels store pixels straight into rank 0's frame over NVLink (CUDA IPC peer mapping); completion = one peer-stored "
                            "flag per rank written by the render kernel's last block, rank 0 waits on its stream (no collective); back-pressure flag "
                            "from rank 0's consumer" if N > 1 else ""),
            "work": {"rays_per_frame": int(rays), "rays_per_frame_f32_kernel": int(counts32[0]), "shaded_hits_per_frame": int(shaded)},
            # `frac` is the ALGORITHMIC fraction SURVEY.md 8(d) defines (flops of the brute-force loops / time / FFMA peak); `hw`
            # is what the hardware counters of the kept ncu capture say for the same kernel (profiles/, from_profile).  On C4
            # the hierarchy skips ~98 % of the algorithmic shape tests, so an algorithmic fraction says nothing there: null.
            "roofline": {"bound": "fp32_non_tensor", "achieved": None if c4 else achieved, "peak": p32, "unit": "TFLOP/s",
                         "frac": None if c4 else achieved / p32,
                         "frac_note": "C4: the bounding-sphere hierarchy skips most algorithmic tests; see hw (executed-instruction utilisation)" if c4 else
                                      "algorithmic flops (54 / ray x sphere, 13 / ray x plane, 150 / shaded hit) over the kernel time, against the FFMA peak",
                         "hw": hw,
                         "traffic": (hw or {}).get("dram_bytes_per_launch"), "traffic_source": "from_profile" if hw else None,
                         "peak_source": "FFMA micro-benchmark on this GPU in this run (rtc_measure_peak_tflops)",
                         "peak_formula": FP32_FORMULA_PEAK, "frac_of_formula": None if c4 else achieved / FP32_FORMULA_PEAK,
                         "algorithmic_gflop_per_frame": flops / 1e9, "kernel_ms": kernel_ms,
                         "kernel": "render_f64_kernel: fast mode sends scenes whose streams exceed shared memory to the parity kernel, the fastest kernel "
                                   "that meets the fast-mode tolerance there (the mixed-precision form of the general f32 kernel: RTC_OPT_F32_GENERAL_KERNEL)"
                                   if c4 else "render_f32_small_kernel",
                         "hbm_gbs_measured_peak": peaks.get("hbm_gbs"),
                         "hbm_gbs_achieved_frame_write": W * H * 12 * my_share / (kernel_ms * 1e-3) / 1e9},
            "f64": {"ms_per_frame": float(ms64.mean()), "mrays_per_s": rays / (float(ms64.mean()) * 1e-3) / 1e6,
                    "kernel_ms": float(k64.mean()), "tflops_algorithmic": flops * my_share / (float(k64.mean()) * 1e-3) / 1e12,
                    "peak_dfma_tflops_measured": pk64,
                    "e2e": {"ms_per_frame": float(e64_t.mean()) * 1e3, "value": rays / float(e64_t.mean()) / 1e6, "unit": "Mrays/s",
                            "d2h_bytes_per_step": int(my_rows * W * 3 * 8),
                            "path": "rtc_scene_upload + rtc_render (f64 parity) delivered by DMA straight into the registered f64 Canvas frame"}},
            "e2e": {"value": rays / (e2e_ms * 1e-3) / 1e6, "unit": "Mrays/s", "ms_per_frame": e2e_ms, "h2d_bytes_per_step": int(h2d),
                    "d2h_bytes_per_step": int(d2h), "steps": e2e_steps,
                    "ms_per_frame_min": float(e2e_t.min()) * 1e3, "ms_per_frame_median": float(e2e_t.median()) * 1e3,
                    "pcie_gbs_per_link": d2h / (e2e_ms * 1e-3) / 1e9, "frame_pages": shared.numa,
                    "row_bands": delivery["bands"], "row_bands_direct": delivery["bands_direct"],
                    "path": "rtc_scene_upload + rtc_render (f32 fast, F64_RGB Canvas layout) into a registered host buffer, mixed delivery: the first "
                            "row bands cross PCIe as f32 (half the bytes over the one link) and host threads widen them into the Canvas, the last "
                            "row_bands_direct bands are written as f64 by the kernel and go by DMA straight into the Canvas (link and host threads side "
                            "by side; the split follows the previous frame's timing; the frame does not depend on it)" if N == 1 else
                            "every rank: rtc_scene_upload + rtc_render of its shard (interleaved row tiles) into ONE registered f64 Canvas frame in host memory "
                            "shared by the rank processes — pitched DMA of f64 rows over the rank's own PCIe link, no staging, no host pass, no gather; "
                            "bytes are per rank; time = max over ranks per step",
                    "other_delivery": {"ms_per_frame": float(alt_t.mean()) * 1e3,
                                       "path": "forced direct f64 DMA" if N == 1 else "staged: f32 over PCIe + widening by the rank's host threads (round 1's path)"},
                    "staged_delivery": None if staged_t is None else {"ms_per_frame": float(staged_t.mean()) * 1e3,
                                                                      "path": "all bands f32 over PCIe + widening by host threads (round 1's path)"}},
            # kernels of this library inside the timed region, all ranks: rank 0 runs the tile scheduling pre-pass + the render
            # kernel (+ the flag wait and the 'consumed' flag store when N > 1); every other rank: its render kernel, which waits for
            # the back-pressure flag and stores its completion flag itself (up to four GPUs preceded by the tile pre-pass: heavy tiles
            # first; beyond that a linear order keeps the pixel stream into rank 0's NVLink ingress even)
            "gpu_launches": args.steps * (2 if N == 1 else 3 + (N - 1) * (2 if N <= 4 else 1)),
            "gpu_launches_per_step": {"rank0": ["classify_tiles_small_kernel", "render_f32_small_kernel"] + (["wait_kernel"] if N > 1 else []),
                                      "other_ranks": ([] if N == 1 else ["render_f32_small_kernel"] if N > 4 else
                                                      ["classify_tiles_small_kernel", "render_f32_small_kernel"])},
            "clocks": clocks,
        }
        if c4:
            line["gpu_launches"] = args.steps * (1 if N == 1 else 2 + 3 * (N - 1))
            line["gpu_launches_per_step"] = {"rank0": ["render_f64_kernel"] + (["wait_kernel"] if N > 1 else []),
                                             "other_ranks": ["wait_kernel", "render_f64_kernel", "signal_kernel"] if N > 1 else []}
        line.update(extra)
        if single is not None:
            line["e2e_single_process"] = single
        if frame_check is not None:
            if e2e_frame_ok is not None:
                frame_check["e2e_shared_host_frame_equals_1gpu_frame"] = e2e_frame_ok
            line["frame_check"] = frame_check
        if N == 1:
            line["frame_check"] = oracle_frame_check(frames) if WORKLOAD_KEY != "c4" else oracle_frame_check(frames, step=64, n_random=512)
        if N == 1 and not args.no_cpu_baseline:
            model, cores = cpu_info()
            s = oracle_sample(cores, 12.0)
            line["cpu_baseline"] = {"value": s["mrays_per_s"], "unit": "Mrays/s", "cores": cores, "kind": "port", "cpu": model,
                                    "sample": s["sample"], "ms_per_frame_extrapolated": s["ms_per_frame_extrapolated"]}
        print(json.dumps(line), flush=True)
    del host
    torch.cuda.synchronize()
    shared.close()
    main.sf.close()
    p64.sf.close()
    if N > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--tile-rows", type=int, default=4, help="rows per interleaved tile (multi-GPU sharding)")
    ap.add_argument("--numa-interleave", action="store_true",
                    help="spread the pages of the shared host frame over the machine's memory nodes (mbind) instead of first touch by rank 0")
    ap.add_argument("--workload", default="c3", choices=["c3", "c4", "c5"],
                    help="c3 (default): the metric's 4K frame; c4: BASELINE.json configs[3], 10k random spheres + planes at 4K; "
                         "c5: configs[4], 7680x4320 with the deterministic 2x2 supersample")
    args = ap.parse_args()
    global WORKLOAD, W, H, SAMPLES, WORKLOAD_KEY, N_SPHERES, N_PLANES, FLOP_PER_RAY
    WORKLOAD_KEY = args.workload
    if args.workload == "c5":
        WORKLOAD = "C5: reflect-refract room (jamis.yml, 13 shapes), 7680x4320, deterministic 2x2 supersample, depth 5"
        W, H, SAMPLES = 7680, 4320, 0
    if args.workload == "c4":
        WORKLOAD = "C4: 9,992 random spheres + 8 planes (seed 0xB200, u8 world_ids wrap), 3840x2160, 1 spp, depth 5"
        N_SPHERES, N_PLANES = 9992, 8
        FLOP_PER_RAY = 54 * N_SPHERES + 13 * N_PLANES
    if args.impl == "reference":
        return run_reference(args)
    return run_gpu(args)


if __name__ == "__main__":
    sys.exit(main())
