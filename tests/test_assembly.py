"""Device-side scene assembly, checked where no GPU exists: csrc/rtb_assemble.cuh's per-triangle and
per-node functions, compiled for the host by nvcc (tests/devsim), must write byte for byte the buffers that rtb_layout.hpp
builds from the host-assembled arrays (what rtb_scene_create uploads).  The kernels that run the same functions on the B200
are compared with the same host path in tests/test_gpu_assembly.py."""
import ctypes as C
import os

import numpy as np
import pytest

import _util
import test_formats

rtb = _util.rtb
KEYS = ("tris", "shade", "tan", "nodes", "big", "scalars")


def _equal(a, b):
    for k in KEYS:
        assert a[k].shape == b[k].shape, k
        assert a[k].tobytes() == b[k].tobytes(), k


def _check(hs):
    d, g = hs.raw()
    got, err = _util.devsim_assemble(d, g)
    assert err is None, err
    _equal(got, _util.devsim_layout(hs.desc))
    return got


@pytest.mark.parametrize("name", _util.SMALL_SCENES + ["sponza_small"])
def test_assembled_buffers_equal_the_host_layout(name):
    hs = rtb.HostScene.load_gltf(_util.scene_path(name), 64, 48, 4)
    got = _check(hs)
    assert got["scalars"][1] == (hs.desc.n_nodes - 1) // 2
    if name == "sponza_small":
        assert got["tan"].shape[0] == hs.desc.n_tris       # normal-mapped materials: the tangent records are written


def test_raw_description_has_no_host_assembled_attributes():
    hs = rtb.HostScene.load_gltf(_util.scene_path("sponza_small"), 64, 48, 4)
    d, g = hs.raw()
    assert not d.tri_nrm and not d.tri_tan and not d.tri_uv and d.tri_pos
    assert g.n_meshes == d.n_meshes and sum(g.meshes[i].n_tris for i in range(g.n_meshes)) == d.n_tris
    src = hs.tri_source()
    assert np.array_equal(np.sort(src), np.arange(d.n_tris, dtype=np.uint32))     # the build's permutation


@pytest.mark.parametrize("normals", [False, True])
def test_triangle_soup(normals):
    tri = _util.gen_scenes.soup_triangles(20000, seed=5, half_extent=3.0).reshape(-1, 9)
    nrm = None
    if normals:
        nrm = _util.gen_scenes.splitmix_u01(77, 0, tri.size).reshape(-1, 9).astype(np.float32)
    hs = rtb.HostScene.from_triangles(tri, nrm, emissive_from=19990)
    got = _check(hs)
    assert np.array_equal(got["shade"][:, 15].view(np.uint32) == 1, hs.tris()["mesh"] == 1)


def test_big_leaves_and_negative_zero():
    """coincident triangles cannot be split (leaves of more than 7 triangles: the table path); -0.0 coordinates keep their sign"""
    base = _util.gen_scenes.soup_triangles(64, seed=9, half_extent=2.0).reshape(-1, 9)
    same = np.repeat(base[:1], 19, axis=0)
    tri = np.concatenate([base, same, -np.abs(base[:3]) * np.float32(0.0), np.repeat(base[5:6], 9, axis=0)]).astype(np.float32)
    tri[70, :] = np.float32(-0.0)
    tri[70, 3] = 1
    tri[70, 7] = 1
    hs = rtb.HostScene.from_triangles(tri)
    got = _check(hs)
    assert got["scalars"][3] >= 3 + 2        # runs of ceil(20 / 7) and ceil(10 / 7) entries
    assert np.any(got["tris"].view(np.uint32) == 0x80000000)


@pytest.mark.parametrize("name", ["s7_4", "sponza_small"])
def test_interleaved_and_glb_inputs(name, tmp_path):
    """byteStride views (several attributes share one byte range) and the .glb container go through the same raw path"""
    src = _util.scene_path(name)
    for path in (test_formats._interleave(src, str(tmp_path)), test_formats._to_glb(src, str(tmp_path))):
        _check(rtb.HostScene.load_gltf(path, 64, 48, 4))


def test_single_leaf_scene():
    tri = _util.gen_scenes.soup_triangles(3, seed=1, half_extent=1.0).reshape(-1, 9)
    hs = rtb.HostScene.from_triangles(tri)
    got = _check(hs)
    assert got["scalars"][1] == 0 and got["scalars"][0] < 0


def test_errors_are_reported():
    hs = rtb.HostScene.load_gltf(_util.scene_path("s7_4"), 64, 48, 4)
    d, g = hs.raw()
    # a vertex index beyond the views
    meshes = (rtb.RawMesh * g.n_meshes)(*[g.meshes[i] for i in range(g.n_meshes)])
    meshes[0].n_vertices = 1
    bad = rtb.RawGeometry(g.n_meshes, meshes, g.tri_source)
    got, err = _util.devsim_assemble(d, bad)
    assert got is None and "vertex index out of range" in err
    # a source beyond the meshes' triangles
    src = hs.tri_source().copy()
    src[3] = d.n_tris + 5
    bad = rtb.RawGeometry(g.n_meshes, g.meshes, src.ctypes.data_as(C.c_void_p))
    got, err = _util.devsim_assemble(d, bad)
    assert got is None and "tri_source" in err
    # nodes in another order (children swapped in the array): the order property is checked, not assumed
    aabb, links = hs.nodes()
    l2, a2 = links.copy(), aabb.copy()
    root_l = l2[0, 0]
    l2[[root_l, root_l + 1]] = l2[[root_l + 1, root_l]]
    a2[[root_l, root_l + 1]] = a2[[root_l + 1, root_l]]
    d2 = rtb.SceneDesc.from_buffer_copy(d)
    d2.node_links = l2.ctypes.data_as(C.c_void_p).value
    d2.node_aabb = a2.ctypes.data_as(C.c_void_p).value
    got, err = _util.devsim_assemble(d2, g)
    assert got is None and "order" in err
