"""The C-ABI library loads and exports every symbol include/rtb.h declares; host-side entry points work without a
GPU; compute entry points FAIL LOUDLY (RTB_ERR_NO_DEVICE) when no CUDA device is usable -- there is no CPU path."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

import _util

rtb = _util.rtb


def _header_symbols():
    src = open(os.path.join(_util.ROOT, "include", "rtb.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(rtb_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_all_exported():
    syms = _header_symbols()
    assert len(syms) >= 20
    lib = rtb.lib()
    for s in syms:
        assert hasattr(lib, s), "librtb.so does not export %s" % s
    assert sorted(rtb.ABI_SYMBOLS) == syms


def test_library_has_sm100a_code():
    out = subprocess.run(["cuobjdump", "-lelf", rtb.LIB_PATH], capture_output=True, text=True).stdout
    assert "sm_100a" in out, out


def test_struct_sizes_match_header(tmp_path):
    """ctypes mirrors of the ABI structs have the sizes the C compiler gives them."""
    src = tmp_path / "sz.c"
    src.write_text('#include "rtb.h"\n#include <stdio.h>\nint main(){printf("%zu %zu %zu %zu\\n", sizeof(rtb_scene_desc), '
                   'sizeof(rtb_render_params), sizeof(rtb_stats), sizeof(rtb_texture_desc));return 0;}\n')
    exe = tmp_path / "sz"
    subprocess.run(["gcc", "-I", os.path.join(_util.ROOT, "include"), str(src), "-o", str(exe)], check=True)
    sizes = [int(x) for x in subprocess.run([str(exe)], capture_output=True, text=True).stdout.split()]
    assert sizes == [C.sizeof(rtb.SceneDesc), C.sizeof(rtb.RenderParams), C.sizeof(rtb.Stats), C.sizeof(rtb.TextureDesc)]


def test_invalid_arguments_are_reported():
    h = C.c_void_p()
    rc = rtb.lib().rtb_host_scene_load_gltf(b"/nonexistent/file.gltf", 8, 8, 1, C.byref(h))
    assert rc == rtb.RTB_ERR_IO and rtb.lib().rtb_last_error()
    rc = rtb.lib().rtb_host_scene_load_gltf(b"x.gltf", 0, 8, 1, C.byref(h))
    assert rc == rtb.RTB_ERR_INVALID
    with pytest.raises(rtb.RtbError):
        rtb.HostScene.from_triangles(np.zeros((0, 9), np.float32))


def test_render_row_partition_helpers():
    p = rtb.RenderParams(100, 1080, 1, 0, 1080, 1, 0, 0, 4, 1, 8)
    rows = rtb.lib().rtb_render_local_rows(C.byref(p))
    assert rows == 272  # blocks 1,5,9,... of 8 rows: 135 blocks -> 34 blocks for index 1... checked below
    seen = [rtb.lib().rtb_render_global_row(C.byref(p), lr) for lr in range(rows)]
    assert seen[:9] == [8, 9, 10, 11, 12, 13, 14, 15, 40]
    allrows = []
    for idx in range(4):
        q = rtb.RenderParams(100, 1080, 1, 0, 1080, 1, 0, 0, 4, idx, 8)
        n = rtb.lib().rtb_render_local_rows(C.byref(q))
        allrows += [rtb.lib().rtb_render_global_row(C.byref(q), lr) for lr in range(n)]
    assert sorted(allrows) == list(range(1080))
    bad = rtb.RenderParams(100, 1080, 1, 0, 1080, 1, 0, 0, 4, 1, 6)   # interleave_rows not a multiple of 4
    assert rtb.lib().rtb_render_local_rows(C.byref(bad)) == -1


def test_ppm_writer(tmp_path):
    img = (np.arange(5 * 7 * 3) % 256).astype(np.uint8).reshape(5, 7, 3)
    path = tmp_path / "o.ppm"
    rtb.write_ppm(path, img)
    data = open(path, "rb").read()
    assert data == b"P6\n7 5\n255\n" + img.tobytes()   # Canvas::write_to (src/render/canvas.h:76-89)


@pytest.mark.skipif(rtb.device_count() > 0, reason="checks the behaviour WITHOUT a CUDA device")
def test_no_cpu_fallback(host_scene):
    hs = host_scene("s7_1")
    with pytest.raises(rtb.RtbError) as e:
        rtb.Scene(hs, 0)
    assert e.value.code == rtb.RTB_ERR_NO_DEVICE


@pytest.mark.skipif(rtb.device_count() > 0, reason="checks the behaviour WITHOUT a CUDA device")
def test_cli_aborts_without_device(tmp_path):
    """the CLI keeps the reference's abort-on-exception behaviour (uncaught std::runtime_error -> SIGABRT)"""
    exe = os.path.join(_util.ROOT, "build", "solution")
    r = subprocess.run([exe, _util.scene_path("s7_1"), "8", "8", "1", str(tmp_path / "o.ppm")], capture_output=True, text=True)
    assert r.returncode == -6 and "no CUDA device" in r.stderr
    assert r.stdout.startswith("Loading scene.\n19 nodes in BVH.")
    r = subprocess.run([exe, "a", "b"], capture_output=True, text=True)
    assert r.returncode == -6 and "Invalid arguments - 3 (expected: 5)" in r.stderr
