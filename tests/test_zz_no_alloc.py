"""rtc_render_device_noalloc — the entry point for callers that cannot call cudaMalloc
(no counterpart in the reference).  The small-scene fast kernel with its scratch (tile counter words,
tile hand-out order) in a static device array: its frames must be the frames of the existing entry point
rtc_render_device — bit for bit, which is the tolerance the same-mode comparisons of this suite use — on two calls in a row
(the array keeps the last call's contents), and sizes beyond the stated maximum must be rejected with a clear error.

This file is named to be collected LAST (a failure here hides nothing behind it), and passes
alone in a fresh process as well as after the other tests."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _frames(rtc, r, cam_desc, opts, stream, with_mask):
    """(existing entry point, no-alloc entry point) -> device tensors of the same frame"""
    import torch
    h, w = cam_desc.vsize, cam_desc.hsize
    rows = r.shard_rows(cam_desc, opts) if not (opts.flags & rtc.ffi.OPT_DST_FULL_FRAME) else h
    want = torch.full((rows, w, 3), -5.0, dtype=torch.float32, device="cuda")
    got = torch.full((rows, w, 3), -6.0, dtype=torch.float32, device="cuda")
    want_hit = torch.full((rows, w), 7, dtype=torch.uint8, device="cuda") if with_mask else None
    got_hit = torch.full((rows, w), 9, dtype=torch.uint8, device="cuda") if with_mask else None
    torch.cuda.synchronize()
    r.render_device(cam_desc, opts, want.data_ptr(), want_hit.data_ptr() if with_mask else None, stream.cuda_stream)
    stream.synchronize()
    r.render_device_noalloc(cam_desc, opts, got.data_ptr(), stream.cuda_stream, got_hit.data_ptr() if with_mask else None)
    stream.synchronize()
    return want, got, want_hit, got_hit


def test_noalloc_entry_point_equals_render_device_on_calls_in_a_row(rtc):
    import torch
    f = rtc.ffi
    r = rtc.Renderer([0])
    r.upload(rtc.scenes.reflect_refract_scene())
    s = torch.cuda.Stream()
    mk = lambda **kw: rtc.Renderer.make_opts(precision=f.PRECISION_F32_FAST, out_format=f.OUT_F32_RGB, **kw)
    # two calls in a row on the frame that takes the tile scheduling pre-pass (16,384 tiles: order words in use), then smaller
    # and differently shaped frames on top of the stale contents, a shard, the hit mask, the 2x2 supersample, depth 0
    cases = [(rtc.scenes.reflect_refract_camera(1024, 512), mk(), False),
             (rtc.scenes.reflect_refract_camera(1024, 512), mk(), True),
             (rtc.scenes.reflect_refract_camera(333, 117), mk(), True),
             (rtc.scenes.reflect_refract_camera(1920, 1080), mk(shard_index=1, shard_count=3, tile_rows=4), False),
             (rtc.scenes.reflect_refract_camera(640, 360, samples=0), mk(), True),
             (rtc.scenes.reflect_refract_camera(1024, 512), mk(max_depth=0), False),
             (rtc.scenes.reflect_refract_camera(1024, 512), mk(), True)]
    for cam, opts, with_mask in cases:
        want, got, want_hit, got_hit = _frames(rtc, r, cam.flatten(), opts, s, with_mask)
        assert torch.equal(got, want), "frame of rtc_render_device_noalloc differs from rtc_render_device's"
        if with_mask:
            assert torch.equal(got_hit, want_hit)
    # the second call of a pair again, back to back on the stream without a synchronise in between
    d = rtc.scenes.reflect_refract_camera(1024, 512).flatten()
    a = torch.zeros((512, 1024, 3), dtype=torch.float32, device="cuda")
    b = torch.zeros_like(a)
    torch.cuda.synchronize()
    r.render_device_noalloc(d, mk(), a.data_ptr(), s.cuda_stream)
    r.render_device_noalloc(d, mk(), b.data_ptr(), s.cuda_stream)
    s.synchronize()
    assert torch.equal(a, b) and torch.equal(a, want)
    r.close()


def test_noalloc_entry_point_allocates_nothing(rtc):
    """Device memory in use is the same before and after the calls (first frames of new sizes included: that is where
    rtc_render_device grows the context's tile-order buffer)."""
    import torch
    f = rtc.ffi
    r = rtc.Renderer([0])
    r.upload(rtc.scenes.reflect_refract_scene())
    s = torch.cuda.Stream()
    o = rtc.Renderer.make_opts(precision=f.PRECISION_F32_FAST, out_format=f.OUT_F32_RGB)
    out = torch.zeros((2160, 3840, 3), dtype=torch.float32, device="cuda")
    r.render_device_noalloc(rtc.scenes.reflect_refract_camera(64, 32).flatten(), o, out.data_ptr(), s.cuda_stream)
    s.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    for size in ((1024, 512), (3840, 2160), (1920, 1080)):
        r.render_device_noalloc(rtc.scenes.reflect_refract_camera(*size).flatten(), o, out.data_ptr(), s.cuda_stream)
    s.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 >= free0, f"device memory in use grew by {free0 - free1} bytes"
    r.close()


def test_noalloc_entry_point_rejects_what_it_cannot_do(rtc):
    import torch
    f = rtc.ffi
    r = rtc.Renderer([0])
    r.upload(rtc.scenes.reflect_refract_scene())
    s = torch.cuda.Stream()
    tiny = torch.zeros(16, dtype=torch.float32, device="cuda")
    fast = rtc.Renderer.make_opts(precision=f.PRECISION_F32_FAST, out_format=f.OUT_F32_RGB)
    # beyond the static scratch: 1024 x 1025 tiles > RTC_NOALLOC_MAX_TILES (rejected before anything is touched)
    big = rtc.scenes.reflect_refract_camera(8192, 4100).flatten()
    assert (8192 // 8) * (4100 // 4) > f.NOALLOC_MAX_TILES
    with pytest.raises(f.RtcError) as e:
        r.render_device_noalloc(big, fast, tiny.data_ptr(), s.cuda_stream)
    assert e.value.code == f.RTC_ERR_UNSUPPORTED and "tiles" in str(e.value) and str(f.NOALLOC_MAX_TILES) in str(e.value)
    # the largest named config fits: 7680 x 4320
    assert (7680 // 8) * (4320 // 4) <= f.NOALLOC_MAX_TILES
    small = rtc.scenes.reflect_refract_camera(32, 16).flatten()
    with pytest.raises(f.RtcError) as e:   # parity mode has no such variant
        r.render_device_noalloc(small, rtc.Renderer.make_opts(precision=f.PRECISION_F64_PARITY), tiny.data_ptr(), s.cuda_stream)
    assert e.value.code == f.RTC_ERR_UNSUPPORTED
    with pytest.raises(f.RtcError) as e:   # nor the general fast kernel
        r.render_device_noalloc(small, rtc.Renderer.make_opts(precision=f.PRECISION_F32_FAST, out_format=f.OUT_F32_RGB, flags=f.OPT_F32_GENERAL_KERNEL),
                                tiny.data_ptr(), s.cuda_stream)
    assert e.value.code == f.RTC_ERR_UNSUPPORTED
    with pytest.raises(f.RtcError) as e:   # a stream is required
        rtc.ffi.check(rtc.ffi.lib().rtc_render_device_noalloc(r.handle, small, fast, tiny.data_ptr(), None, None))
    assert e.value.code == f.RTC_ERR_INVALID_ARG
    w, cam = rtc.scenes.cube_scene(7)      # a scene outside the small-scene form (cubes)
    r.upload(w)
    with pytest.raises(f.RtcError) as e:
        r.render_device_noalloc(small, fast, tiny.data_ptr(), s.cuda_stream)
    assert e.value.code == f.RTC_ERR_UNSUPPORTED
    # the context is still good
    r.upload(rtc.scenes.reflect_refract_scene())
    out = torch.zeros((16, 32, 3), dtype=torch.float32, device="cuda")
    r.render_device_noalloc(small, fast, out.data_ptr(), s.cuda_stream)
    s.synchronize()
    assert float(out.abs().sum()) > 0.0
    r.close()
