"""The N>1 path on CPU: world_size-2 (and 3) `gloo` process groups drive the same ShardedFrame class
bench.py uses on GPUs.  Each rank's "render" is the oracle on exactly the rows the shard arithmetic
gives it; the exchange step must reproduce the oracle's unsharded frame on rank 0."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, vsize, hsize, tile, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import rtc_b200 as rtc
        import orc
        from rtc_b200 import ffi
        from rtc_b200.sharded import ShardedFrame, frame_rows_of_shard, shard_rows
        scene = rtc.scenes.reflect_refract_scene()
        cam = rtc.scenes.reflect_refract_camera(hsize, vsize)
        sf = ShardedFrame(hsize, vsize, ffi.OUT_F64_RGB, rank, world, torch.device("cpu"), tile_rows=tile, mode="nccl")
        opts = sf.shard_opts(rtc.Renderer.make_opts)
        assert opts.shard_index == rank and opts.shard_count == world and opts.tile_rows == tile and opts.flags == 0
        # the library's own row count agrees with the host arithmetic (no GPU needed for rtc_shard_rows)
        L = ffi.lib()
        import ctypes as C
        assert L.rtc_shard_rows(C.byref(cam.flatten()), C.byref(opts)) == shard_rows(vsize, tile, rank, world) == sf.my_rows
        rows = frame_rows_of_shard(vsize, tile, rank, world).numpy()
        pix = (rows[:, None] * hsize + np.arange(hsize)[None, :]).reshape(-1)
        rgb, _, _ = orc.render(scene, cam, pixels=pix)              # this rank's tiles only
        sf.destination()[: rows.size] = torch.from_numpy(rgb.reshape(rows.size, hsize, 3))
        frame = sf.exchange()
        if rank == 0:
            want, _, _ = orc.render(scene, cam)
            q.put(bool(np.array_equal(frame.numpy(), want)))
        else:
            assert frame is None
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,vsize,tile", [(2, 40, 16), (2, 33, 4), (3, 50, 8)])
def test_sharded_frame_over_gloo_matches_unsharded(world, vsize, tile, _built):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, vsize, 48, tile, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert q.get(timeout=5) is True


def _worker_shared(rank, world, port, vsize, hsize, tile, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import rtc_b200 as rtc
        import orc
        from rtc_b200.sharded import SharedHostFrame, frame_rows_of_shard
        scene = rtc.scenes.reflect_refract_scene()
        cam = rtc.scenes.reflect_refract_camera(hsize, vsize)
        shared = SharedHostFrame(hsize, vsize, rank, world)
        if rank == 0:
            shared.array[:] = -1.0
        dist.barrier()
        rows = frame_rows_of_shard(vsize, tile, rank, world).numpy()
        pix = (rows[:, None] * hsize + np.arange(hsize)[None, :]).reshape(-1)
        rgb, _, _ = orc.render(scene, cam, pixels=pix)              # what rtc_render(shard opts) delivers: this rank's rows
        shared.array[rows] = rgb.reshape(rows.size, hsize, 3)
        dist.barrier()
        if rank == 0:
            want, _, _ = orc.render(scene, cam)
            q.put(bool(np.array_equal(shared.array, want)))
        shared.close()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,vsize,tile", [(2, 33, 4), (3, 50, 8)])
def test_shared_host_frame_over_gloo(world, vsize, tile, _built):
    """The end-to-end N > 1 path of bench.py on CPU: one frame in shared host memory, every rank fills in the rows of
    its shard (here the oracle stands in for rtc_render's shard mode), no exchange step."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker_shared, args=(r, world, port, vsize, 48, tile, q)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert q.get(timeout=5) is True


def test_deinterleave_index_math():
    from rtc_b200.sharded import deinterleave, frame_rows_of_shard, shard_rows
    for vsize in (1, 15, 16, 17, 100):
        for count in (1, 2, 3, 8):
            for tile in (1, 4, 16):
                stride = shard_rows(vsize, tile, 0, count)
                frame = torch.arange(vsize * 5, dtype=torch.float64).reshape(vsize, 5)
                g = torch.zeros(count, stride, 5, dtype=torch.float64)
                seen = []
                for k in range(count):
                    rows = frame_rows_of_shard(vsize, tile, k, count)
                    assert rows.numel() == shard_rows(vsize, tile, k, count) <= stride
                    g[k, : rows.numel()] = frame[rows]
                    seen += rows.tolist()
                assert sorted(seen) == list(range(vsize))
                assert torch.equal(deinterleave(g, vsize, tile), frame)
