"""Device-side scene assembly on the B200: rtb_scene_create_raw uploads the raw glTF views + the
builder's permutation and assembles on the GPU; every device buffer must equal, byte for byte, what rtb_scene_create uploads
from the host-assembled arrays."""
import ctypes as C
import time

import numpy as np
import pytest

import _util
import test_formats

rtb = _util.rtb
pytestmark = pytest.mark.gpu
BUFS = [rtb.BUF_NODES, rtb.BUF_TRIS, rtb.BUF_SHADE, rtb.BUF_TANGENTS, rtb.BUF_BIG_LEAVES, rtb.BUF_BVH_SCALARS]


def _same_buffers(hs):
    a = rtb.Scene(hs, 0)                       # rtb_scene_create_raw
    b = rtb.Scene(hs, 0, assembled=True)       # rtb_scene_create
    for which in BUFS:
        x, y = a.read_buffer(which), b.read_buffer(which)
        assert x.shape == y.shape and x.tobytes() == y.tobytes(), "buffer %d differs" % which
    assert a.traversal_bytes == b.traversal_bytes
    return a, b


@pytest.mark.parametrize("name", _util.SMALL_SCENES + ["sponza_small"])
def test_device_assembled_buffers_equal_the_host_assembled_upload(name):
    hs = rtb.HostScene.load_gltf(_util.scene_path(name), 64, 48, 4)
    a, b = _same_buffers(hs)
    # and the scene behaves the same: one ray batch, one small frame
    g = _util.load_golden(name)
    rays = _util.box_rays(g["root_box"], 4096, seed=1000)
    ha, hb = a.closest_hit(rays), b.closest_hit(rays)
    for k in ("id", "t", "u", "v", "inside"):
        assert np.array_equal(ha[k], hb[k]), k
    ra, rb = a.render(32, 24, 2), b.render(32, 24, 2)
    assert np.array_equal(ra["accum"], rb["accum"]) and np.array_equal(ra["rgb"], rb["rgb"])


@pytest.mark.parametrize("normals", [False, True])
def test_soup_1m(normals):
    """cfg3's scene: 1 M triangles (flat normals computed on the device / caller's normals verbatim); upload bytes before / after"""
    tri = _util.gen_scenes.soup_triangles(1000000, seed=1234, half_extent=10.0).reshape(-1, 9)
    nrm = _util.gen_scenes.splitmix_u01(5, 0, tri.size).reshape(-1, 9).astype(np.float32) if normals else None
    hs = rtb.HostScene.from_triangles(tri, nrm, emissive_from=999999)
    t0 = time.perf_counter()
    a = rtb.Scene(hs, 0)
    t_raw = time.perf_counter() - t0
    t0 = time.perf_counter()
    b = rtb.Scene(hs, 0, assembled=True)
    t_flat = time.perf_counter() - t0
    for which in BUFS:
        assert a.read_buffer(which).tobytes() == b.read_buffer(which).tobytes(), "buffer %d differs" % which
    print("\n1 M soup: rtb_scene_create_raw %.1f MB in %.3f s, rtb_scene_create %.1f MB in %.3f s" % (
        a.upload_bytes / 1e6, t_raw, b.upload_bytes / 1e6, t_flat))
    if not normals:
        assert a.upload_bytes < b.upload_bytes


def test_big_leaves_interleaved_views_and_glb(tmp_path):
    base = _util.gen_scenes.soup_triangles(64, seed=9, half_extent=2.0).reshape(-1, 9)
    tri = np.concatenate([base, np.repeat(base[:1], 19, axis=0), np.repeat(base[5:6], 9, axis=0)]).astype(np.float32)
    a, _ = _same_buffers(rtb.HostScene.from_triangles(tri))
    assert a.read_buffer(rtb.BUF_BIG_LEAVES).size >= 5 * 8
    src = _util.scene_path("sponza_small")
    for path in (test_formats._interleave(src, str(tmp_path)), test_formats._to_glb(src, str(tmp_path))):
        _same_buffers(rtb.HostScene.load_gltf(path, 64, 48, 4))
    _same_buffers(rtb.HostScene.from_triangles(base[:3]))      # the whole BVH is one leaf


def test_errors():
    hs = rtb.HostScene.load_gltf(_util.scene_path("s7_4"), 64, 48, 4)
    d, g = hs.raw()
    meshes = (rtb.RawMesh * g.n_meshes)(*[g.meshes[i] for i in range(g.n_meshes)])
    meshes[0].n_vertices = 1
    with pytest.raises(rtb.RtbError, match="vertex index out of range"):
        rtb.Scene(d, 0, raw=rtb.RawGeometry(g.n_meshes, meshes, g.tri_source))
    src = hs.tri_source().copy()
    src[3] = d.n_tris + 5
    with pytest.raises(rtb.RtbError, match="tri_source"):
        rtb.Scene(d, 0, raw=rtb.RawGeometry(g.n_meshes, g.meshes, src.ctypes.data_as(C.c_void_p)))
    aabb, links = hs.nodes()
    l2, a2 = links.copy(), aabb.copy()
    k = l2[0, 0]
    l2[[k, k + 1]] = l2[[k + 1, k]]
    a2[[k, k + 1]] = a2[[k + 1, k]]
    d2 = rtb.SceneDesc.from_buffer_copy(d)
    d2.node_links = l2.ctypes.data_as(C.c_void_p).value
    d2.node_aabb = a2.ctypes.data_as(C.c_void_p).value
    with pytest.raises(rtb.RtbError, match="order"):
        rtb.Scene(d2, 0, raw=g)
    meshes = (rtb.RawMesh * g.n_meshes)(*[g.meshes[i] for i in range(g.n_meshes)])
    meshes[0].position_stride = 10
    with pytest.raises(rtb.RtbError, match="aligned"):
        rtb.Scene(d, 0, raw=rtb.RawGeometry(g.n_meshes, meshes, g.tri_source))
    rtb.Scene(hs, 0).close()       # and the scene still loads
