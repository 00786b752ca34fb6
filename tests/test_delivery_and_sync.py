"""rtc_render's delivery paths (staged vs direct DMA into a registered Canvas buffer), several GPUs in one process
(skipped below two devices), and the completion-flag protocol (fused into the small-scene kernel, back-pressure, a wait
that times out without poisoning the CUDA context).  All through the C ABI."""
import ctypes as C
import time

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _room(rtc, w, h, samples=1):
    return rtc.scenes.reflect_refract_scene(), rtc.scenes.reflect_refract_camera(w, h, samples=samples).flatten()


@pytest.mark.parametrize("precision", [0, 1])
@pytest.mark.parametrize("size", [(640, 362), (1920, 1080)])
def test_direct_delivery_into_a_registered_buffer_equals_staged(rtc, renderer, precision, size):
    """The frame DMA-ed straight into a registered caller buffer (no staging, no host pass) is the staged frame bit for
    bit: whole frames in both precisions (fast mode: the library would stage a whole f32 frame on one GPU, so direct is
    forced), F64 / F32 / U8 outputs."""
    f = rtc.ffi
    world, d = _room(rtc, *size)
    renderer.upload(world)
    for fmt, dt in ((f.OUT_F64_RGB, np.float64), (f.OUT_F32_RGB, np.float32), (f.OUT_U8_RGB, np.uint8)):
        want, want_hit, st = renderer.render(d, rtc.Renderer.make_opts(precision=precision, out_format=fmt))
        buf = np.full((size[1], size[0], 3), 7, dtype=dt)
        renderer.register_host_buffer(buf)
        try:
            got, hit, st2 = renderer.render(d, rtc.Renderer.make_opts(precision=precision, out_format=fmt, flags=f.OPT_DELIVER_DIRECT), out=buf)
            assert got is buf and np.array_equal(buf, want) and np.array_equal(hit, want_hit) and st2.rays == st.rays
            buf[:] = 7
            renderer.render(d, rtc.Renderer.make_opts(precision=precision, out_format=fmt), out=buf, want_mask=False, want_stats=False)   # the library's own choice
            assert np.array_equal(buf, want)
            buf[:] = 7
            renderer.render(d, rtc.Renderer.make_opts(precision=precision, out_format=fmt, flags=f.OPT_DELIVER_STAGED), out=buf, want_mask=False)
            assert np.array_equal(buf, want)
        finally:
            renderer.unregister_host_buffer(buf)


@pytest.mark.parametrize("size", [(1024, 516), (3840, 2160), (3840, 2161)])
def test_mixed_delivery_of_a_fast_frame_equals_staged(rtc, renderer, size):
    """A fast-mode frame into a registered F64_RGB Canvas buffer on one GPU: the first row bands cross PCIe as f32 and are
    widened by host threads, the last ones are written as f64 by the kernel and DMA-ed straight into the buffer.  Either
    way every value is the kernel's f32 widened exactly, so the frame is the staged frame bit for bit — with the split
    fixed at half the bands (both delivery flags) and with the library's own choice over several frames in a row (the
    split follows the previous frame's timing)."""
    f = rtc.ffi
    world, d = _room(rtc, *size)
    renderer.upload(world)
    want, want_hit, st = renderer.render(d, rtc.Renderer.make_opts(precision=1))          # unregistered: staged
    buf = np.full((size[1], size[0], 3), 7.0)
    renderer.register_host_buffer(buf)
    try:
        got, hit, st2 = renderer.render(d, rtc.Renderer.make_opts(precision=1, flags=f.OPT_DELIVER_DIRECT | f.OPT_DELIVER_STAGED), out=buf)
        assert np.array_equal(buf, want) and np.array_equal(hit, want_hit) and st2.rays == st.rays
        for _ in range(5):
            buf[:] = 7.0
            got, hit, st2 = renderer.render(d, rtc.Renderer.make_opts(precision=1), out=buf)
            assert np.array_equal(buf, want) and np.array_equal(hit, want_hit) and st2.rays == st.rays
    finally:
        renderer.unregister_host_buffer(buf)


def test_delivery_info_reports_what_the_last_render_moved(rtc, renderer):
    """rtc_delivery_info: bytes copied device -> host, row bands and bands sent by DMA straight into the caller's buffer, for the
    staged, the direct and the mixed delivery of one fast-mode frame into an F64_RGB buffer."""
    f = rtc.ffi
    w, h = 1024, 516
    world, d = _room(rtc, w, h)
    renderer.upload(world)
    buf = np.zeros((h, w, 3))
    renderer.render(d, rtc.Renderer.make_opts(precision=1), out=buf, want_mask=False)              # not registered: staged, f32 over PCIe
    info = renderer.delivery_info()
    assert info["d2h_bytes"] == w * h * 3 * 4 and info["bands_direct"] == 0 and info["bands"] >= 1
    renderer.register_host_buffer(buf)
    try:
        renderer.render(d, rtc.Renderer.make_opts(precision=1, flags=f.OPT_DELIVER_DIRECT), out=buf, want_mask=False)
        info = renderer.delivery_info()
        assert info["d2h_bytes"] == w * h * 3 * 8 and info["bands_direct"] == info["bands"] >= 1
        renderer.render(d, rtc.Renderer.make_opts(precision=1, flags=f.OPT_DELIVER_DIRECT | f.OPT_DELIVER_STAGED), out=buf, want_mask=False)
        info = renderer.delivery_info()
        assert 0 < info["bands_direct"] < info["bands"] and w * h * 3 * 4 < info["d2h_bytes"] < w * h * 3 * 8
        renderer.render(d, rtc.Renderer.make_opts(precision=0), out=buf, want_mask=False)          # parity mode, registered: direct
        info = renderer.delivery_info()
        assert info["d2h_bytes"] == w * h * 3 * 8 and info["bands_direct"] == info["bands"]
    finally:
        renderer.unregister_host_buffer(buf)


@pytest.mark.parametrize("precision", [0, 1])
@pytest.mark.parametrize("count,tile,size", [(2, 4, (200, 117)), (3, 16, (200, 117)), (8, 4, (1920, 1080)), (4, 4, (64, 3)),
                                             (2, 4, (3840, 2160)), (3, 20, (2560, 1441))])   # the last two: several row bands per shard
def test_shards_deliver_directly_into_one_registered_frame(rtc, renderer, precision, count, tile, size):
    """What N rank processes do with a registered frame in shared host memory: every shard's rows arrive by pitched 2-D
    DMA at their frame rows (the frame's last, shorter tile by a 1-D copy); all shards together are the unsharded frame."""
    world, d = _room(rtc, *size)
    renderer.upload(world)
    whole, whole_hit, st = renderer.render(d, rtc.Renderer.make_opts(precision=precision))
    frame = np.full((size[1], size[0], 3), -7.0)
    mask = np.full((size[1], size[0]), 99, dtype=np.uint8)
    renderer.register_host_buffer(frame)
    try:
        rays = 0
        for k in range(count):
            o = rtc.Renderer.make_opts(precision=precision, shard_index=k, shard_count=count, tile_rows=tile)
            before = frame.copy()
            stats = rtc.ffi.RtcStats()
            rtc.ffi.check(rtc.ffi.lib().rtc_render(renderer.handle, C.byref(d), C.byref(o), frame.ctypes.data_as(C.c_void_p),
                                                   mask.ctypes.data_as(C.c_void_p), C.byref(stats)))
            rays += stats.rays
            mine = (np.arange(size[1]) // tile) % count == k
            assert np.array_equal(frame[~mine], before[~mine])
        assert np.array_equal(frame, whole) and np.array_equal(mask, whole_hit) and rays == st.rays
    finally:
        renderer.unregister_host_buffer(frame)


def test_a_buffer_another_context_registered_can_be_registered_again(rtc, renderer):
    """Two contexts of one process deliver into the same long-lived frame (bench.py: a rank's own context and the one-process
    several-GPU context): the second registration finds the range page-locked, uses it, and leaves the registration to its owner."""
    world, d = _room(rtc, 320, 180)
    renderer.upload(world)
    other = rtc.Renderer([0])
    other.upload(world)
    buf = np.full((180, 320, 3), -1.0)
    renderer.register_host_buffer(buf)
    other.register_host_buffer(buf)
    o = rtc.Renderer.make_opts(precision=1, flags=rtc.ffi.OPT_DELIVER_DIRECT)
    want, _, _ = renderer.render(d, rtc.Renderer.make_opts(precision=1))
    other.render(d, o, out=buf, want_mask=False, want_stats=False)
    assert np.array_equal(buf, want)
    other.unregister_host_buffer(buf)          # not the owner: the range stays page-locked
    buf[:] = -1.0
    renderer.render(d, o, out=buf, want_mask=False, want_stats=False)
    assert np.array_equal(buf, want)
    renderer.unregister_host_buffer(buf)
    other.close()


def test_direct_delivery_needs_a_registered_buffer(rtc, renderer):
    world, d = _room(rtc, 64, 32)
    renderer.upload(world)
    with pytest.raises(rtc.ffi.RtcError) as e:
        renderer.render(d, rtc.Renderer.make_opts(flags=rtc.ffi.OPT_DELIVER_DIRECT))
    assert e.value.code == rtc.ffi.RTC_ERR_INVALID_ARG
    buf = np.zeros((32, 64, 3))
    with pytest.raises(rtc.ffi.RtcError):
        renderer.unregister_host_buffer(buf)


# ---- several GPUs in one process: rtc_context_create(ids, N) + rtc_render / rtc_render_device ---------------------------
def _n_devices(rtc):
    return rtc.ffi.lib().rtc_device_count()


@pytest.mark.parametrize("precision", [0, 1])
def test_one_process_two_gpus_equals_one_gpu(rtc, renderer, precision):
    """The drop-in's own multi-GPU path (a Rust camera.render(&world) is one call in one process, src/camera.rs:144-160):
    a context over two GPUs renders interleaved row tiles on one host thread per GPU; staged and direct delivery, the
    ragged last tile, hit mask and counters — bit-identical to the one-GPU frame."""
    if _n_devices(rtc) < 2:
        pytest.skip("needs two GPUs")
    r2 = rtc.Renderer([0, 1])
    for size, tile in (((320, 163), 4), ((1920, 1080), 4), ((200, 50), 16)):
        world, d = _room(rtc, *size)
        renderer.upload(world)
        r2.upload(world)
        o = rtc.Renderer.make_opts(precision=precision, tile_rows=tile)
        want, want_hit, st = renderer.render(d, o)
        got, hit, st2 = r2.render(d, o)
        assert np.array_equal(got, want) and np.array_equal(hit, want_hit)
        assert (st2.rays, st2.shaded_hits, st2.primary_hits) == (st.rays, st.shaded_hits, st.primary_hits) and st2.kernel_launches == 2
        buf = np.full((size[1], size[0], 3), -3.0)
        r2.register_host_buffer(buf)
        try:
            r2.render(d, o, out=buf, want_mask=False, want_stats=False)
            assert np.array_equal(buf, want)
        finally:
            r2.unregister_host_buffer(buf)
    r2.close()


@pytest.mark.parametrize("precision", [0, 1])
def test_one_process_two_gpus_device_frame_by_peer_stores(rtc, renderer, precision):
    """rtc_render_device on a two-GPU context: the second GPU's kernel stores its tiles straight into the frame on the
    first GPU (peer access), events order the streams — equal to the one-GPU frame, also when enqueued on a user stream."""
    if _n_devices(rtc) < 2:
        pytest.skip("needs two GPUs")
    import torch
    f = rtc.ffi
    r2 = rtc.Renderer([0, 1])
    world, d = _room(rtc, 640, 362)
    renderer.upload(world)
    r2.upload(world)
    fmt, dt = (f.OUT_F64_RGB, torch.float64) if precision == 0 else (f.OUT_F32_RGB, torch.float32)
    o = rtc.Renderer.make_opts(precision=precision, out_format=fmt, tile_rows=4)
    want = torch.zeros((362, 640, 3), dtype=dt, device="cuda:0")
    want_hit = torch.zeros((362, 640), dtype=torch.uint8, device="cuda:0")
    renderer.render_device(d, o, want.data_ptr(), want_hit.data_ptr())
    got = torch.zeros_like(want)
    hit = torch.zeros_like(want_hit)
    st = f.RtcStats()
    r2.render_device(d, o, got.data_ptr(), hit.data_ptr(), None, st)
    assert torch.equal(got, want) and torch.equal(hit, want_hit) and st.kernel_launches == 2 and st.rays > 640 * 362
    with torch.cuda.device(0):
        s = torch.cuda.Stream()
        got.zero_()
        torch.cuda.synchronize()
        with torch.cuda.stream(s):
            r2.render_device(d, o, got.data_ptr(), None, s.cuda_stream)
        s.synchronize()
    assert torch.equal(got, want)
    r2.close()


# ---- completion flags ----------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("flags", [0, 2])   # 0: the small-scene kernel (fused flags); 2 = RTC_OPT_F32_GENERAL_KERNEL (flag kernels around the launch)
def test_render_device_synced_waits_for_the_consumer_and_signals(rtc, renderer, flags):
    """rtc_render_device_synced: the render does not start before *wait_flag >= wait_value (back-pressure: the consumer of
    the previous frame releases the destination) and stores done_value when the pixels are out.  The flag is released
    here by a DMA copy from pinned host memory on a second stream (copy engines need no SM; the waiting kernel holds all)."""
    import torch
    f = rtc.ffi
    world, d = _room(rtc, 960, 540)
    renderer.upload(world)
    o = rtc.Renderer.make_opts(precision=f.PRECISION_F32_FAST, out_format=f.OUT_F32_RGB, flags=flags)
    want = torch.zeros((540, 960, 3), dtype=torch.float32, device="cuda")
    renderer.render_device(d, o, want.data_ptr())
    words = torch.zeros(64, dtype=torch.int32, device="cuda")
    one = torch.tensor([5], dtype=torch.int32).pin_memory()
    got = torch.zeros_like(want)
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    torch.cuda.synchronize()
    renderer.render_device_synced(d, o, got.data_ptr(), s1.cuda_stream, wait_flag=words.data_ptr(), wait_value=5,
                                  done_flag=words.data_ptr() + 4, done_value=9)
    time.sleep(0.05)
    assert not s1.query()                       # still waiting for the consumer
    with torch.cuda.stream(s2):
        words[0:1].copy_(one, non_blocking=True)
    s1.synchronize()
    s2.synchronize()
    assert int(words[1]) == 9 and torch.equal(got, want)
    renderer.wait_result()                      # no timeout recorded
    # already satisfied: straight through, sequence numbers compare with wrap-around
    renderer.render_device_synced(d, o, got.data_ptr(), s1.cuda_stream, wait_flag=words.data_ptr(), wait_value=3, done_flag=words.data_ptr() + 4, done_value=10)
    s1.synchronize()
    assert int(words[1]) == 10


def test_wait_signal_device_releases_only_when_every_flag_arrived(rtc):
    import torch
    f = rtc.ffi
    r = rtc.Renderer([0])
    r.set_wait_timeout(0.05)
    words = torch.zeros(64, dtype=torch.int32, device="cuda")
    words[1:4] = 4
    torch.cuda.synchronize()
    r.wait_signal_device(words.data_ptr() + 4, 3, 4, words.data_ptr(), 4)
    r.wait_result()
    assert int(words[0]) == 4
    r.wait_signal_device(words.data_ptr() + 4, 5, 4, words.data_ptr(), 9)    # flags 3 and 4 never arrive: no release
    with pytest.raises(f.RtcError) as e:
        r.wait_result()
    assert e.value.code == f.RTC_ERR_TIMEOUT and int(words[0]) == 4
    r.close()


def test_a_wait_that_times_out_is_reported_and_the_context_survives(rtc, orc):
    """A rank that never signals: rtc_wait_device gives up after the context's timeout, rtc_wait_result names the missing
    flag with RTC_ERR_TIMEOUT, and the context renders on (round 1 trapped: every buffer of the process was lost)."""
    import torch
    f = rtc.ffi
    r = rtc.Renderer([0])
    world, d = _room(rtc, 64, 32)
    r.upload(world)
    r.set_wait_timeout(0.05)
    words = torch.zeros(64, dtype=torch.int32, device="cuda")
    words[1] = 3
    torch.cuda.synchronize()
    r.wait_device(words.data_ptr() + 4, 3, 3)   # flags of "ranks" 1..3: only rank 1 has signalled frame 3
    with pytest.raises(f.RtcError) as e:
        r.wait_result()
    assert e.value.code == f.RTC_ERR_TIMEOUT and "flag 1" in str(e.value)
    r.wait_result()                              # the record was consumed
    want, want_hit, _ = orc.render(world, rtc.scenes.reflect_refract_camera(64, 32))
    got, hit, _ = r.render(d, rtc.Renderer.make_opts())
    assert np.abs(got - want).max() <= 1e-9 and np.array_equal(hit, want_hit)
    with pytest.raises(f.RtcError):
        r.set_wait_timeout(-1.0)
    r.close()


@pytest.mark.parametrize("general", [False, True])
def test_f32_light_lying_in_a_plane(rtc, orc, renderer, general):
    """A light exactly in the floor plane (f(light) == 0).  The reference's shadow ray meets that plane at t == distance up
    to rounding and `h.t < distance` (src/shape.rs:724) is then decided by the last bit; the fast kernels' any-hit
    plane term f(p) * f(light) would be 0 <= 0 — everything shadowed.  They treat such a plane as never shadowing, which
    is the f64 image with the light a hair above the plane."""
    I = rtc.Matrix.identity
    w = rtc.World(rtc.Light(rtc.Color(1, 1, 1), (2.0, 0.0, -3.0)))
    w.add_shape(rtc.Plane(I(), rtc.Material(color=rtc.Color(0.6, 0.6, 0.6), specular=0.0)))
    w.add_shape(rtc.Plane(I().rotation_x(np.pi / 2).translation(0, 0, 6), rtc.Material(color=rtc.Color(0.3, 0.5, 0.7), specular=0.1)))
    for k, (x, y, z, rad) in enumerate(((0, 1.0, 1, 1.0), (-2.2, 0.6, 0.5, 0.6), (1.8, 0.4, -0.5, 0.4))):
        w.add_shape(rtc.Sphere(I().scaling(rad, rad, rad).translation(x, y, z), rtc.Material(color=rtc.Color(0.9 - 0.3 * k, 0.3 + 0.2 * k, 0.2), reflectiveness=0.2 * k)))
    cam = rtc.Camera.new_with_transform(240, 160, 1.0, rtc.Matrix.make_view_transform((0, 2.5, -7.0), (0, 0.8, 1), (0, 1, 0)))
    lifted = rtc.World(rtc.Light(rtc.Color(1, 1, 1), (2.0, 1e-7, -3.0)))
    for s in w.shapes:
        lifted.add_shape(s)
    want, _, _ = orc.OracleWorld(lifted).render(cam, threads=4)
    renderer.upload(w)
    got, _, _ = renderer.render(cam.flatten(), rtc.Renderer.make_opts(precision=rtc.ffi.PRECISION_F32_FAST, out_format=rtc.ffi.OUT_F32_RGB,
                                                                       flags=rtc.ffi.OPT_F32_GENERAL_KERNEL if general else 0))
    q_got, q_want = orc.quantize_u8(got.astype(np.float64)).astype(int), orc.quantize_u8(want).astype(int)
    ok = (np.abs(q_got - q_want).max(axis=2) <= 1).mean()
    assert ok >= 0.995, f"only {ok:.5f} within 1/255"
