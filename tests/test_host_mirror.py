"""The product's host-side mirror (transform.py / scene.py / camera.py) must build bit-identical
matrices and descriptors to the oracle's restatement of src/transform.rs and src/camera.rs."""
import ctypes as C
import math
import random

import numpy as np
import pytest


def _chain_both(rtc, orc, ops):
    m = rtc.Matrix.identity()
    names = ["translation", "scaling", "rotation_x", "rotation_y", "rotation_z", "shearing"]
    nargs = [3, 3, 1, 1, 1, 6]
    codes = (C.c_int * len(ops))()
    args = (C.c_double * (6 * len(ops)))()
    for i, (op, a) in enumerate(ops):
        m = getattr(m, names[op])(*a[:nargs[op]])
        codes[i] = op
        for k in range(6):
            args[6 * i + k] = a[k] if k < len(a) else 0.0
    out = (C.c_double * 16)()
    assert orc.lib().orc_matrix_chain(codes, args, len(ops), out) == 0
    return m, list(out)


def test_transform_chains_bit_identical(rtc, orc):
    rnd = random.Random(7)
    for _ in range(200):
        ops = []
        for _ in range(rnd.randint(1, 5)):
            op = rnd.randint(0, 5)
            a = [rnd.uniform(-3, 3) for _ in range(6)]
            if op == 1:
                a = [rnd.choice([-1, 1]) * rnd.uniform(0.2, 3) for _ in range(6)]
            ops.append((op, a))
        m, want = _chain_both(rtc, orc, ops)
        assert m.to_row_major() == want
        inv_want = (C.c_double * 16)()
        det = C.c_double()
        src = (C.c_double * 16)(*want)
        rc = orc.lib().orc_matrix_inverse(src, inv_want, C.byref(det))
        assert (rc == 0) == m.is_invertable()
        if rc == 0:
            assert m.inverse().to_row_major() == list(inv_want)
            assert m.determinant == det.value
            t = (C.c_double * 16)()
            orc.lib().orc_matrix_transpose(inv_want, t)
            assert m.inverse().transpose().to_row_major() == list(t)


def test_not_invertible_raises_like_the_reference_panics(rtc):
    with pytest.raises(ValueError):
        rtc.Matrix.identity().scaling(0.002, 0.002, 0.002).inverse()  # |det| = 8e-9 <= 1e-8 (src/transform.rs:37,177)


def test_view_transform_and_camera_bit_identical(rtc, orc):
    rnd = random.Random(11)
    for _ in range(50):
        frm = [rnd.uniform(-5, 5) for _ in range(3)]
        to = [rnd.uniform(-5, 5) for _ in range(3)]
        up = [rnd.uniform(-1, 1), 1.0, rnd.uniform(-1, 1)]
        out = (C.c_double * 16)()
        orc.lib().orc_view_transform((C.c_double * 3)(*frm), (C.c_double * 3)(*to), (C.c_double * 3)(*up), out)
        view = rtc.Matrix.make_view_transform(frm, to, up)
        assert view.to_row_major() == list(out)
        h, v, fov = rnd.randint(1, 500), rnd.randint(1, 500), rnd.uniform(0.3, 2.5)
        cam = rtc.Camera.new_with_transform(h, v, fov, view)
        want = rtc.ffi.RtcCamera()
        assert orc.lib().orc_camera_new(h, v, fov, out, C.byref(want)) == 0
        got = cam.flatten()
        assert (got.half_width, got.half_height, got.pixel_size) == (want.half_width, want.half_height, want.pixel_size)
        assert list(got.view_inv) == list(want.view_inv)


def test_camera_kats(rtc):
    """src/camera.rs:174-185 (pixel size 0.01 for 200x125 and 125x200 at fov pi/2)."""
    assert abs(rtc.Camera.new(200, 125, math.pi / 2).pixel_size - 0.01) < 1e-4
    assert abs(rtc.Camera.new(125, 200, math.pi / 2).pixel_size - 0.01) < 1e-4


def test_world_ids_wrap_like_release_build(rtc):
    w = rtc.World(rtc.Light())
    for _ in range(300):
        w.add_shape(rtc.Sphere())
    ids = [s.world_id for s in w.shapes]
    assert ids[0] == 1 and ids[254] == 255 and ids[255] == 0 and ids[256] == 1


def test_plane_set_transform_keeps_stale_inverse_transpose(rtc):
    """src/shape.rs:446-449: Plane::set_transform transposes the OLD inverse-transpose."""
    p = rtc.Plane()
    p.set_transform(rtc.Matrix.identity().rotation_x(0.7).translation(0, -1, 0))
    assert p.transform_transpose == rtc.Matrix.identity()
    s = rtc.Sphere()
    s.set_transform(rtc.Matrix.identity().rotation_x(0.7))
    assert s.transform_transpose == s.transform.transpose()


def test_reflect_refract_scene_descriptor(rtc):
    w = rtc.scenes.reflect_refract_scene()
    light, arr, n = w.flatten()
    assert n == 13
    assert [arr[i].kind for i in range(n)] == [1] * 6 + [0] * 7
    assert [arr[i].world_id for i in range(n)] == list(range(1, 14))
    assert arr[0].pattern_kind == rtc.ffi.PATTERN_CHECKER and arr[2].pattern_kind == rtc.ffi.PATTERN_STRIPE
    assert arr[11].transparency == 0.9 and arr[11].reflective == 0.9 and arr[11].refractive_index == 1.5
    assert list(light.position) == [-4.9, 4.9, -1.0]


def test_canvas_u8_rule_matches_oracle(rtc, orc):
    vals = np.array([[[0.999, 1.0, 1.9], [-0.5, 0.5, float("nan")], [1e12, -1e12, 0.00392]]], dtype=np.float64)
    c = rtc.Canvas(3, 1, vals)
    assert np.array_equal(c.to_u8(), orc.quantize_u8(vals))
    assert c.to_u8().reshape(-1).tolist() == [254, 255, 255, 0, 127, 0, 255, 0, 0]
