"""The C-ABI library loads and exports every symbol include/rtc.h declares (no GPU needed), and
fails loudly — never falls back — when there is no device."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "rtc.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(rtc_[a-z_0-9]+)\s*\(", src)))


def test_library_exports_every_declared_symbol(rtc):
    L = rtc.ffi.lib()
    names = _declared()
    assert len(names) >= 11
    assert sorted(n for n, _, _ in rtc.ffi.EXPORTS) == names
    for n in names:
        assert getattr(L, n) is not None
    assert L.rtc_abi_version() == 1


def test_struct_layouts_match_header(rtc):
    f = rtc.ffi
    assert C.sizeof(f.RtcShape) == 8 + 16 * 8 * 2 + 8 + 3 * 8 + 7 * 8 + 16 * 8 + 6 * 8
    assert C.sizeof(f.RtcLight) == 48
    assert C.sizeof(f.RtcCamera) == 8 + 3 * 8 + 16 * 8 + 8
    assert C.sizeof(f.RtcRenderOpts) == 32
    assert C.sizeof(f.RtcStats) == 5 * 8 + 5 * 8


def test_field_offsets_match_a_c_compiler(rtc, tmp_path):
    """Every field of every struct of include/rtc.h at the offset gcc gives it == the ctypes mirror (ffi.py), which is
    what the Rust #[repr(C)] structs of INTEGRATION.md have to match too."""
    import shutil
    import subprocess
    if not shutil.which("gcc"):
        pytest.skip("no gcc")
    f = rtc.ffi
    structs = {"rtc_shape": f.RtcShape, "rtc_light": f.RtcLight, "rtc_camera": f.RtcCamera, "rtc_render_opts": f.RtcRenderOpts,
               "rtc_stats": f.RtcStats, "rtc_selfcheck": f.RtcSelfcheck}
    lines = ['#include <stdio.h>', '#include <stddef.h>', '#include "rtc.h"', 'int main(void) {']
    for cname, ct in structs.items():
        lines.append(f'  printf("{cname} size %zu\\n", sizeof({cname}));')
        for fname, _ in ct._fields_:
            lines.append(f'  printf("{cname} {fname} %zu\\n", offsetof({cname}, {fname}));')
    lines += ['  return 0;', '}']
    src = tmp_path / "offsets.c"
    src.write_text("\n".join(lines))
    exe = tmp_path / "offsets"
    subprocess.run(["gcc", "-std=c99", "-I", os.path.join(ROOT, "include"), "-o", str(exe), str(src)], check=True)
    out = subprocess.run([str(exe)], check=True, stdout=subprocess.PIPE, text=True).stdout
    seen = 0
    for line in out.splitlines():
        cname, field, value = line.split()
        ct = structs[cname]
        want = C.sizeof(ct) if field == "size" else getattr(ct, field).offset
        assert int(value) == want, line
        seen += 1
    assert seen == sum(len(ct._fields_) + 1 for ct in structs.values())


def test_shard_rows_partition_the_frame(rtc):
    L = rtc.ffi.lib()
    cam = rtc.ffi.RtcCamera()
    for vsize in (1, 15, 16, 17, 160, 2160, 4321):
        cam.hsize, cam.vsize = 8, vsize
        for count in (1, 2, 3, 4, 8):
            for tile in (1, 8, 16):
                rows = []
                for idx in range(count):
                    o = rtc.Renderer.make_opts(shard_index=idx, shard_count=count, tile_rows=tile)
                    rows.append(L.rtc_shard_rows(C.byref(cam), C.byref(o)))
                assert sum(rows) == vsize, (vsize, count, tile, rows)


def test_no_device_is_a_loud_error_not_a_fallback(rtc):
    L = rtc.ffi.lib()
    if L.rtc_device_count() > 0:
        pytest.skip("a CUDA device is present")
    ctx = C.c_void_p()
    rc = L.rtc_context_create(None, 1, C.byref(ctx))
    assert rc == rtc.ffi.RTC_ERR_NO_DEVICE
    assert b"no CPU fallback" in L.rtc_last_error()
    with pytest.raises(rtc.ffi.RtcError):
        rtc.Renderer([0])
    with pytest.raises(rtc.ffi.RtcError):
        rtc.scenes.reflect_refract_camera(8, 4).render(rtc.scenes.reflect_refract_scene())


def test_product_package_never_touches_the_oracle():
    pkg = os.path.join(ROOT, "raytracer-challenge_b200")
    for dirpath, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cu", ".h", ".cuh", "Makefile")):
                text = open(os.path.join(dirpath, fn), errors="ignore").read()
                assert "liborc" not in text and "oracle/" not in text and "import orc" not in text, fn
