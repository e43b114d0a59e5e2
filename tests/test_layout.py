"""Host front end (glTF load + BVH build, raytracing-hw_b200/host/) produces the reference's node and triangle
layout: digests of every array equal the ones the reference produced (tests/golden), and -- where oracle/_ref is
present -- the arrays themselves equal the reference's, node for node and triangle for triangle."""
import hashlib

import numpy as np
import pytest

import _util

SCENES = _util.SMALL_SCENES + ["sponza_small"]


def digest(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


@pytest.mark.parametrize("name", SCENES)
def test_layout_digests_match_reference(name, host_scene):
    g = _util.load_golden(name)
    hs = host_scene(name)
    a, l = hs.nodes()
    tr = hs.tris()
    assert a.shape[0] == int(g["n_nodes"]) and tr["pos"].shape[0] == int(g["n_tris"])
    got = [digest(a), digest(l), digest(tr["pos"]), digest(tr["nrm"]), digest(tr["tan"]), digest(tr["uv"]), digest(tr["mesh"])]
    assert got == list(g["layout_sha256"])
    la, ll = hs.light_nodes()
    assert [digest(la), digest(ll), digest(hs.light_pos())] == list(g["light_sha256"])
    cam = hs.camera()
    assert np.array_equal(cam["pos"], g["cam_pos"]) and np.array_equal(cam["axes"], g["cam_axes"])


def test_leaf_and_node_invariants(host_scene):
    """what the device layout conversion relies on: leaf <=> count > 0 <=> no children; internal
    nodes have a split axis in {0,1,2}; every triangle is in exactly one leaf."""
    hs = host_scene("s6_1")
    a, l = hs.nodes()
    leaf = l[:, 4] > 0
    assert np.all((l[leaf, 0] == -1) & (l[leaf, 1] == -1))
    assert np.all((l[~leaf, 0] > 0) & (l[~leaf, 1] > 0) & (l[~leaf, 2] >= 0) & (l[~leaf, 2] <= 2))
    cover = np.zeros(hs.tris()["pos"].shape[0], np.int32)
    for f, c in l[leaf][:, 3:5]:
        cover[f:f + c] += 1
    assert np.all(cover == 1)
    n, counts, mx = hs.bvh_stats()
    assert n == a.shape[0] and counts[0] == int(np.sum(~leaf)) and mx == int(l[:, 4].max())


def test_parallel_build_equals_sequential():
    """the task-parallel builder numbers nodes analytically; its output must not depend on the thread count"""
    import os
    import subprocess
    import sys
    tri = _util.gen_scenes.soup_triangles(60000, seed=5, half_extent=4.0).reshape(-1, 9)
    code = ("import sys, hashlib, numpy as np; sys.path.insert(0, %r); import _util;"
            "t=_util.gen_scenes.soup_triangles(60000, seed=5, half_extent=4.0).reshape(-1,9);"
            "hs=_util.rtb.HostScene.from_triangles(t); a,l=hs.nodes();"
            "print(hashlib.sha256(a.tobytes()+l.tobytes()+hs.tris()['pos'].tobytes()).hexdigest())") % os.path.dirname(__file__)
    outs = []
    for th in ("1", "8"):
        env = dict(os.environ, OMP_NUM_THREADS=th)
        outs.append(subprocess.run([sys.executable, "-c", code], env=env, capture_output=True, text=True, check=True).stdout.strip())
    assert outs[0] == outs[1] and len(outs[0]) == 64
    assert tri.shape == (60000, 9)


@pytest.mark.ref
@pytest.mark.parametrize("n_tris,half", [(36, 1.0), (10000, 2.15), (100000, 4.64)])
def test_soup_layout_equals_reference(n_tris, half, ref):
    tri = _util.gen_scenes.soup_triangles(n_tris, seed=1234, half_extent=half).reshape(-1, 9)
    hs = _util.rtb.HostScene.from_triangles(tri, emissive_from=n_tris - 3)
    rs = ref.RefScene.from_triangles(tri, emissive_from=n_tris - 3)
    a, l = hs.nodes()
    ra, rl = rs.nodes()
    assert np.array_equal(a, ra) and np.array_equal(l.astype(np.int64), rl)
    assert np.array_equal(hs.tris()["pos"], rs.tris()["pos"]) and np.array_equal(hs.tris()["nrm"], rs.tris()["nrm"])
    la, ll = hs.light_nodes()
    rla, rll = rs.light_nodes()
    assert np.array_equal(la, rla) and np.array_equal(ll.astype(np.int64), rll)
    assert np.array_equal(hs.light_pos(), rs.light_tris()["pos"])


@pytest.mark.ref
@pytest.mark.parametrize("wide", ["300", "5000"])
def test_all_threads_path_of_big_nodes_equals_reference(wide, ref, monkeypatch):
    """nodes of >= 128 k triangles are split by all threads together: std::sort replayed as tasks (same partitions, per-block
    insertion sort), chunked box scans, the partition point found by search.  RTB_BUILD_WIDE forces that path onto small
    nodes: the layout must stay the reference's, ties included (the axis-aligned Sponza stand-in is full of equal centroids)"""
    monkeypatch.setenv("RTB_BUILD_WIDE", wide)
    tri = _util.gen_scenes.soup_triangles(100000, seed=1234, half_extent=4.64).reshape(-1, 9)
    # duplicates on purpose: exact ties inside big nodes
    tri[5000:9000] = tri[1000:5000]
    hs = _util.rtb.HostScene.from_triangles(tri, emissive_from=100000 - 3)
    rs = ref.RefScene.from_triangles(tri, emissive_from=100000 - 3)
    a, l = hs.nodes()
    ra, rl = rs.nodes()
    assert np.array_equal(a, ra) and np.array_equal(l.astype(np.int64), rl)
    assert np.array_equal(hs.tris()["pos"], rs.tris()["pos"])
    p = _util.scene_path("sponza_small")
    hs = _util.rtb.HostScene.load_gltf(p, 64, 48, 4)
    rs = ref.RefScene.from_gltf(p, 64, 48, 4)
    a, l = hs.nodes()
    ra, rl = rs.nodes()
    assert np.array_equal(a, ra) and np.array_equal(l.astype(np.int64), rl)
    assert np.array_equal(hs.tris()["pos"], rs.tris()["pos"]) and np.array_equal(hs.tris()["mesh"], rs.tris()["mesh"])


@pytest.mark.ref
def test_gltf_scene_equals_reference_fully(ref):
    """every array of the flat scene, including textures and materials, against the reference's Scene"""
    p = _util.scene_path("sponza_small")
    hs = _util.rtb.HostScene.load_gltf(p, 64, 48, 4)
    rs = ref.RefScene.from_gltf(p, 64, 48, 4)
    tr, rt = hs.tris(), rs.tris()
    for k in tr:
        assert np.array_equal(tr[k], rt[k]), k
    m, nt = hs.meshes()
    rm, rnt = rs.meshes()
    assert np.array_equal(m, rm) and np.array_equal(nt, rnt)
    tx, rtx = hs.textures(), rs.textures()
    assert len(tx) == len(rtx) == 69 and all(np.array_equal(x, y) for x, y in zip(tx, rtx))
    cam, rcam = hs.camera(), rs.camera()
    assert np.array_equal(cam["pos"], rcam["pos"]) and np.array_equal(cam["axes"], rcam["axes"])
    assert cam["max_distance"] == rcam["max_distance"] and cam["ray_depth"] == rcam["ray_depth"] == 6


def test_build_bvh_entry_point_equals_scene_build():
    """rtb_build_bvh (BVH::buildBVH alone, for a caller that keeps its own scene): order, nodes and links equal what the scene
    loaders build from the same triangles -- which the tests above pin against the reference"""
    tri = _util.gen_scenes.soup_triangles(20000, seed=11, half_extent=2.7).reshape(-1, 9)
    hs = _util.rtb.HostScene.from_triangles(tri)
    loader_order = np.argsort(hs.tri_source())          # rows of tris()["pos"] back in the caller's order
    pos_in = np.ascontiguousarray(hs.tris()["pos"])[loader_order]
    order, aabb, links, info = _util.rtb.build_bvh(pos_in)
    a, l = hs.nodes()
    assert np.array_equal(order, hs.tri_source()) and np.array_equal(aabb.view(np.uint32), a.view(np.uint32)) and np.array_equal(links, l)
    assert set(info) >= {"on_gpu", "levels", "rounds"}
    with pytest.raises(_util.rtb.RtbError):
        _util.rtb.lib().rtb_build_bvh  # (symbol present)
        import ctypes as C
        nn = C.c_size_t(0)
        _util.rtb._check(_util.rtb.lib().rtb_build_bvh(_util.rtb._ptr(pos_in), pos_in.shape[0], _util.rtb._ptr(order), _util.rtb._ptr(aabb),
                                                       _util.rtb._ptr(links), 3, C.byref(nn), None))   # arrays too small
