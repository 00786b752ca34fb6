"""Multi-GPU frame assembly for ``Camera::render``: one process per GPU
(``torch.distributed``), the frame cut into interleaved row tiles (tile k -> rank k % world), the
scene replicated on every GPU, and ONE exchange step that puts the finished frame on rank 0.

Two exchange modes:

* ``"nccl"``  — every rank renders into a compact shard buffer, ``dist.gather`` (grouped NCCL
  send/recv over NVLink) collects the buffers on rank 0, ``rtc_assemble_device`` de-interleaves.
* ``"p2p"``   — rank 0 exports its frame buffer through CUDA IPC (``rtc_ipc_export``), the other
  ranks map it (``rtc_ipc_open``) and their render kernels store finished pixels straight into it
  over NVLink (``RTC_OPT_DST_FULL_FRAME``): the transfer overlaps the rendering pixel by pixel.
  No collective is left.  A flag array in rank 0's memory carries the protocol, frame k = 1, 2, ...:
  word r (r >= 1) = "rank r's pixels of frame k are in" (release-stored at system scope after them,
  by the last block of the render kernel itself: ``rtc_render_device_synced``), word 0 = "rank 0's
  consumer is done with frame k" (``release()``).  Rank 0 waits for words 1.. on its stream
  (``rtc_wait_device``); the other ranks' kernels wait for word 0 >= k - 1 before their first store —
  back-pressure: a rank that runs ahead cannot overwrite a frame rank 0 is still reading.

The reference has no multi-device path (rayon threads over one ``Canvas``, src/camera.rs:144-160);
the tile-to-rank arithmetic here is the only "protocol", and it is pure index math so the same class
runs on CPU tensors with the ``gloo`` backend in the tests (the renderer is then a stand-in callable).
"""
import torch
import torch.distributed as dist

from . import ffi


def shard_rows(vsize, tile_rows, index, count):
    """Rows of a ``vsize``-row frame owned by shard ``index`` of ``count`` (same as rtc_shard_rows)."""
    if count <= 1:
        return vsize
    n_tiles = (vsize + tile_rows - 1) // tile_rows
    return sum(min(vsize, (t + 1) * tile_rows) - t * tile_rows for t in range(index, n_tiles, count))


def frame_rows_of_shard(vsize, tile_rows, index, count):
    """Frame row of every compact row of shard ``index``, in compact order (a LongTensor)."""
    rows = []
    n_tiles = (vsize + tile_rows - 1) // tile_rows
    for t in range(index, n_tiles, count):
        rows.extend(range(t * tile_rows, min(vsize, (t + 1) * tile_rows)))
    return torch.tensor(rows, dtype=torch.long)


def deinterleave(gathered, vsize, tile_rows):
    """Host/torch form of rtc_assemble_device: ``gathered`` is [count, stride_rows, ...] compact
    shard buffers; returns the [vsize, ...] frame.  Used for CPU tensors (tests) only."""
    count = gathered.shape[0]
    frame = torch.empty((vsize,) + tuple(gathered.shape[2:]), dtype=gathered.dtype, device=gathered.device)
    for k in range(count):
        rows = frame_rows_of_shard(vsize, tile_rows, k, count).to(gathered.device)
        frame[rows] = gathered[k, : rows.numel()]
    return frame


_TORCH_DTYPE = {ffi.OUT_F64_RGB: torch.float64, ffi.OUT_F32_RGB: torch.float32, ffi.OUT_U8_RGB: torch.uint8}


class ShardedFrame:
    """Buffers and exchange step of one camera's frame across the ranks of the default process
    group.  Per frame a rank (1) renders its tiles into ``destination()`` with the options from
    ``shard_opts()`` — ``Renderer.render_device`` on GPUs, any stand-in on CPU tensors in the
    tests — and (2) calls ``exchange()``."""

    def __init__(self, hsize, vsize, out_format, rank, world, device, tile_rows=16, mode="p2p", channels=3, signal="flags"):
        self.hsize, self.vsize, self.out_format = hsize, vsize, out_format
        self.rank, self.world, self.device, self.tile_rows, self.mode = rank, world, device, tile_rows, mode
        self.dtype = _TORCH_DTYPE[out_format]
        self.stride_rows = shard_rows(vsize, tile_rows, 0, world)     # shard 0 owns the most rows
        self.my_rows = shard_rows(vsize, tile_rows, rank, world)
        self.channels = channels
        shape_tail = (hsize, channels)
        self.frame = None
        self.shard = None
        self.gathered = None
        self.remote_ptr = None
        self._ipc = []
        self._token = None
        self.signal = signal          # p2p completion: "flags" (peer-stored sequence numbers) or "allreduce" (4-byte NCCL all-reduce)
        self.flags = None             # rank 0: int32[world]; other ranks: self.remote_flags (peer pointer)
        self.remote_flags = None
        self.seq = 0
        self._renderer = None
        if world == 1:
            self.frame = torch.zeros((vsize,) + shape_tail, dtype=self.dtype, device=device)
        elif mode == "nccl":
            self.shard = torch.zeros((self.stride_rows,) + shape_tail, dtype=self.dtype, device=device)
            if rank == 0:
                self.gathered = torch.empty((world, self.stride_rows) + shape_tail, dtype=self.dtype, device=device)
                self.frame = torch.empty((vsize,) + shape_tail, dtype=self.dtype, device=device)
        elif mode == "p2p":
            if rank == 0:
                self.frame = torch.zeros((vsize,) + shape_tail, dtype=self.dtype, device=device)
                self.flags = torch.zeros(max(world, 64), dtype=torch.int32, device=device)
        else:
            raise ValueError(f"unknown exchange mode {mode!r}")

    # ---- p2p set-up: rank 0 exports the frame, the others map it ---------------------------------
    def connect(self, renderer):
        self._renderer = renderer
        if self.world == 1 or self.mode != "p2p":
            return
        payload = [None]
        if self.rank == 0:
            payload = [(renderer.ipc_export(self.frame.data_ptr()), renderer.ipc_export(self.flags.data_ptr()))]
        dist.broadcast_object_list(payload, src=0)
        if self.rank != 0:
            (fh, foff), (gh, goff) = payload[0]
            self.remote_ptr = renderer.ipc_open(fh, foff)
            self._ipc.append((self.remote_ptr, foff))
            self.remote_flags = renderer.ipc_open(gh, goff)
            self._ipc.append((self.remote_flags, goff))
        dist.barrier()

    def close(self):
        if self._ipc:
            torch.cuda.synchronize()
            for ptr, offset in self._ipc:
                self._renderer.ipc_close(ptr, offset)
            self._ipc = []

    # ---- one frame ---------------------------------------------------------------------------------
    def shard_opts(self, make_opts, **kw):
        flags = kw.pop("flags", 0)
        if self.world > 1 and self.mode == "p2p":
            flags |= ffi.OPT_DST_FULL_FRAME
            if self.rank != 0 and self.world > 4:
                # peer stores from MORE than four GPUs: keep the pixel stream over NVLink even in time (87 MB per 4K frame into
                # rank 0's 900 GB/s ingress is 97 us of a 150 us kernel at N = 8, and the heavy-tiles-first order ends with a
                # burst of cheap pixels over that limit: +27 us).  Up to four GPUs the link has room and the heavy-first order
                # is the faster kernel (a 1/2 shard 0.625 -> 0.601 ms, a 1/4 shard 0.332 -> 0.316; profiles/r02_tail_trace.txt).
                flags |= ffi.OPT_LINEAR_TILE_ORDER
        return make_opts(out_format=self.out_format, shard_index=self.rank, shard_count=self.world,
                         tile_rows=self.tile_rows, flags=flags, **kw)

    def destination(self):
        """Device pointer (or CPU tensor) this rank's render call writes to."""
        if self.world == 1:
            return self.frame
        if self.mode == "nccl":
            return self.shard
        return self.frame if self.rank == 0 else self.remote_ptr

    def render(self, renderer, cam_desc, opts, stream):
        """Enqueue this rank's render of the next frame into ``destination()``.  In p2p mode with flags the non-zero
        ranks use ``rtc_render_device_synced``: wait for 'previous frame consumed', store the completion flag from the
        kernel's last block."""
        dst = self.destination()
        dst = dst if isinstance(dst, int) else dst.data_ptr()
        if self.world > 1 and self.mode == "p2p" and self.signal == "flags" and self.rank != 0:
            k = self.seq + 1
            renderer.render_device_synced(cam_desc, opts, dst, stream, wait_flag=self.remote_flags, wait_value=k - 1,
                                          done_flag=self.remote_flags + 4 * self.rank, done_value=k)
            self._fused = True
        else:
            renderer.render_device(cam_desc, opts, dst, None, stream)
            self._fused = False

    def release(self, stream=None):
        """Rank 0, after the consumer of the frame ``exchange()`` returned has been enqueued on ``stream``: the other
        ranks may overwrite the frame (word 0 of the flag array = frames consumed).  No-op elsewhere."""
        if self.world > 1 and self.mode == "p2p" and self.signal == "flags" and self.rank == 0:
            self._renderer.signal_device(self.flags.data_ptr(), self.seq, stream)

    def exchange(self, assemble=None, stream=None, release=False):
        """The exchange step after every rank's render call was enqueued.  Returns the frame on
        rank 0 (None elsewhere).  ``assemble(gathered, frame)`` de-interleaves on the device
        (rtc_assemble_device); without it the torch index form is used (CPU tensors).  release=True (p2p with flags): rank 0
        has no consumer to wait for and frees the frame for the next one in the same launch — otherwise call release()
        after the consumer."""
        if self.world == 1:
            return self.frame
        if self.mode == "nccl":
            dist.gather(self.shard, [self.gathered[i] for i in range(self.world)] if self.rank == 0 else None, dst=0)
            if self.rank != 0:
                return None
            if assemble is not None:
                assemble(self.gathered, self.frame)
            else:
                self.frame = deinterleave(self.gathered, self.vsize, self.tile_rows)
            return self.frame
        # p2p: all peer stores are complete once every rank's kernel has finished.
        if self.signal == "flags":
            self.seq += 1
            if self.rank == 0:
                if release:   # no consumer of our own between frames: wait and "frame consumed" in one launch
                    self._renderer.wait_signal_device(self.flags.data_ptr() + 4, self.world - 1, self.seq, self.flags.data_ptr(), self.seq, stream)
                else:
                    self._renderer.wait_device(self.flags.data_ptr() + 4, self.world - 1, self.seq, stream)
                return self.frame
            if not getattr(self, "_fused", False):   # the render call was not render(): store the flag behind it
                self._renderer.signal_device(self.remote_flags + 4 * self.rank, self.seq, stream)
            self._fused = False
            return None
        # fallback: a 4-byte stream-ordered all-reduce as completion signal (no host synchronisation)
        if self._token is None:
            self._token = torch.zeros(1, dtype=torch.int32, device=self.device)
        dist.all_reduce(self._token)
        return self.frame if self.rank == 0 else None


def _mbind_interleave(addr, nbytes):
    """Spread the pages of a not-yet-touched mapping round-robin over the machine's memory nodes (mbind(2),
    MPOL_INTERLEAVE; for a shared-memory segment the policy belongs to the segment, so it holds for every process that
    maps it).  Eight GPUs hang off two sockets: a frame whose pages all sit on the socket of the thread that touched them
    first makes half the GPUs' DMA cross the socket link.  Returns a word for the bench line; never raises."""
    import ctypes
    try:
        nodes = []
        for part in open("/sys/devices/system/node/has_memory").read().strip().split(","):
            lo, _, hi = part.partition("-")
            nodes += list(range(int(lo), int(hi or lo) + 1))
        if len(nodes) < 2:
            return "one memory node: nothing to interleave"
        mask = (ctypes.c_ulong * 16)()
        for n in nodes:
            mask[n // 64] |= 1 << (n % 64)
        libc = ctypes.CDLL(None, use_errno=True)
        page = 4096
        start = addr & ~(page - 1)
        length = (addr + nbytes - start + page - 1) & ~(page - 1)
        rc = libc.syscall(237, ctypes.c_void_p(start), ctypes.c_ulong(length), 3, mask, ctypes.c_ulong(16 * 64), 0)   # SYS_mbind (x86-64), MPOL_INTERLEAVE
        if rc != 0:
            return "mbind refused (errno %d): first touch by rank 0" % ctypes.get_errno()
        return "pages interleaved over %d memory nodes" % len(nodes)
    except Exception as e:   # no /sys node files, not x86-64, ...
        return "interleave unavailable (%s): first touch by rank 0" % type(e).__name__


class SharedHostFrame:
    """One Canvas frame (hsize x vsize x 3, f64 by default) in host memory that all rank processes of a node map:
    POSIX shared memory created by rank 0, its name broadcast over the process group.  Every rank's ``rtc_render`` with
    ``shard_index = rank, shard_count = world`` fills in the rows of its shard (include/rtc.h, rtc_render_opts) — its
    own PCIe link, its own host threads, no gather and no collective on the data path.  ``array`` is the numpy view."""

    def __init__(self, hsize, vsize, rank, world, dtype="float64", numa_interleave=False):
        import numpy as np
        from multiprocessing import shared_memory
        self.rank, self.world = rank, world
        self.numa = "first touch by rank 0"
        nbytes = int(hsize) * int(vsize) * 3 * np.dtype(dtype).itemsize
        name = [None]
        if rank == 0:
            self._shm = shared_memory.SharedMemory(create=True, size=nbytes)
            name[0] = self._shm.name
        if world > 1:
            dist.broadcast_object_list(name, src=0)
        if rank != 0:
            self._shm = shared_memory.SharedMemory(name=name[0])
            try:   # rank 0 unlinks the segment; keep the other ranks' resource trackers out of it
                from multiprocessing import resource_tracker
                resource_tracker.unregister(self._shm._name, "shared_memory")
            except Exception:
                pass
        self.array = np.ndarray((vsize, hsize, 3), dtype=dtype, buffer=self._shm.buf)
        self._registered_with = None
        if numa_interleave and rank == 0:
            self.numa = _mbind_interleave(self.array.ctypes.data, nbytes)

    def register(self, renderer):
        """rtc_register_host_buffer on this process's mapping of the segment: this rank's rtc_render then delivers its
        rows by DMA straight into the shared frame (no staging, no host threads).  Rank 0 touches the pages first."""
        if self.rank == 0:
            self.array[:] = 0.0
        if self.world > 1:
            dist.barrier()
        renderer.register_host_buffer(self.array)
        self._registered_with = renderer

    def close(self):
        """Collective: every rank drops its view, then rank 0 removes the segment."""
        if self._registered_with is not None:
            self._registered_with.unregister_host_buffer(self.array)
            self._registered_with = None
        self.array = None
        if self.world > 1:
            dist.barrier()
        self._shm.close()
        if self.rank == 0:
            self._shm.unlink()
