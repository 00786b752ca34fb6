"""Host-only self-check of the scene flattening pass (rtc_scene_selfcheck; no device needed): every shape in exactly one
intersect stream per precision, Morton reordering only for big streams, and every bounding sphere of the line-distance
filter enclosed by its group's and its block's sphere — what lets the kernels skip whole groups without losing a hit."""
import ctypes as C
import math

import pytest


def _check(rtc, world, n_lights=1):
    light, shapes, n = world.flatten()
    lights = world.flatten_lights()
    out = rtc.ffi.RtcSelfcheck()
    rtc.ffi.check(rtc.ffi.lib().rtc_scene_selfcheck(lights, len(lights), shapes, n, C.byref(out)))
    return out


def test_room_is_small_and_unsorted(rtc):
    out = _check(rtc, rtc.scenes.reflect_refract_scene())
    assert (out.n_planes, out.n_uniform_spheres, out.n_general_spheres, out.n_cubes) == (6, 7, 0, 0)
    assert out.streams_are_permutations == 1 and out.morton_sorted == 0 and out.small_scene_form == 1
    assert out.n_groups_f64 == 1 and out.n_blocks_f64 == 1
    assert out.worst_enclosure_f32 <= 0.0 and out.worst_enclosure_f64 <= 0.0
    two = rtc.scenes.reflect_refract_scene()
    two.add_light(rtc.Light(rtc.Color(0.3, 0.5, 0.9), (3.5, 3.0, -3.0)))
    assert _check(rtc, two).small_scene_form == 0   # the constant-bank form folds one light into its shadow terms


def test_ten_thousand_shapes_are_sorted_and_enclosed(rtc):
    world = rtc.scenes.synthetic_scene()
    out = _check(rtc, world)
    assert out.n_planes == 8 and out.n_uniform_spheres + out.n_general_spheres == 9992 and out.n_general_spheres > 500
    assert out.streams_are_permutations == 1 and out.morton_sorted == 1 and out.small_scene_form == 0
    assert out.n_groups_f32 == math.ceil(math.ceil(9992 / 8)) and out.n_blocks_f32 == math.ceil(out.n_groups_f32 / 32)
    assert out.n_groups_f64 == 1249 and out.n_blocks_f64 == 40
    assert -1.0 < out.worst_enclosure_f32 <= 0.0 and -1.0 < out.worst_enclosure_f64 <= 0.0


def test_cubes_needles_and_aliased_ids(rtc):
    import test_gpu_parity as T
    out = _check(rtc, T._big_mixed_world(rtc))
    assert out.n_cubes > 0 and out.streams_are_permutations == 1
    assert out.worst_enclosure_f32 <= 0.0 and out.worst_enclosure_f64 <= 0.0


def test_argument_checks(rtc):
    world = rtc.World.default()
    light, shapes, n = world.flatten()
    out = rtc.ffi.RtcSelfcheck()
    L = rtc.ffi.lib()
    assert L.rtc_scene_selfcheck(C.byref(light), 0, shapes, n, C.byref(out)) == rtc.ffi.RTC_ERR_INVALID_ARG
    assert L.rtc_scene_selfcheck(C.byref(light), 1, shapes, n, None) == rtc.ffi.RTC_ERR_INVALID_ARG
    assert L.rtc_scene_selfcheck(C.byref(light), 1, shapes, n, C.byref(out)) == 0 and out.n_uniform_spheres == 2


@pytest.mark.parametrize("n_spheres", [70, 300, 2500])
def test_kd_order_keeps_every_level_enclosed_and_the_streams_permutations(rtc, n_spheres):
    """The k-d order of big scenes (rtc_api.cu, spatial_order: nth_element per node, leaves sorted) through the host-only
    self-check: the streams are permutations of the input shapes, the stream is in spatial order, and every group /
    block bound encloses the level below it (worst enclosure slack <= 0 up to rounding) — for sizes around the sort
    threshold (64), one block (256) and many blocks."""
    import ctypes as C
    f = rtc.ffi
    light, shapes, n = rtc.scenes.synthetic_scene(n_spheres=n_spheres, seed=7).flatten()
    out = f.RtcSelfcheck()
    f.check(f.lib().rtc_scene_selfcheck(C.byref(light), 1, shapes, n, C.byref(out)))
    assert out.streams_are_permutations == 1 and out.morton_sorted == 1
    assert out.n_uniform_spheres + out.n_general_spheres == n_spheres and out.n_planes == 8
    assert out.n_groups_f64 == (n_spheres + 7) // 8 and out.n_blocks_f64 == (out.n_groups_f64 + 31) // 32
    assert out.worst_enclosure_f32 <= 1e-6 and out.worst_enclosure_f64 <= 1e-4


def test_shared_host_frame_single_rank_and_numa_option(rtc):
    """SharedHostFrame on one rank (no process group needed): a POSIX shared-memory Canvas frame; the NUMA option never
    raises — it reports what it did (one memory node here: nothing to interleave)."""
    import numpy as np
    from rtc_b200.sharded import SharedHostFrame, _mbind_interleave
    for opt in (False, True):
        fr = SharedHostFrame(48, 20, 0, 1, numa_interleave=opt)
        assert fr.array.shape == (20, 48, 3) and fr.array.dtype == np.float64 and isinstance(fr.numa, str)
        fr.array[:] = 1.5
        assert float(fr.array.sum()) == 1.5 * 20 * 48 * 3
        fr.close()
    buf = np.zeros(1 << 16)
    assert isinstance(_mbind_interleave(buf.ctypes.data, buf.nbytes), str)
