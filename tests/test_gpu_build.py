"""GPU BVH builder (csrc/rtb_build.cu: libstdc++'s introsort replayed breadth first, segmented SAH sweeps) against the host
builder, whose layout is pinned node for node against the compiled reference (tests/test_layout.py, test_gpu_baseline_sizes.py):
node boxes, links, split axes and the triangle order must be EQUAL, ties included."""
import os

import numpy as np
import pytest

import _util

rtb = _util.rtb
pytestmark = pytest.mark.gpu


def build(tri, **env):
    old = {k: os.environ.get(k) for k in env}
    os.environ.update({k: str(v) for k, v in env.items()})
    try:
        hs = rtb.HostScene.from_triangles(tri)
    finally:
        for k, v in old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v
    a, l = hs.nodes()
    return a.copy(), l.copy(), hs.tri_source().copy(), hs.build_info(), hs.build_seconds


def same(x, y):
    assert x[0].shape == y[0].shape, "node counts differ: %s vs %s" % (x[0].shape, y[0].shape)
    assert np.array_equal(x[2], y[2]), "triangle order differs at %d places" % int(np.sum(x[2] != y[2]))
    assert np.array_equal(x[1], y[1]), "links differ"
    assert np.array_equal(x[0].view(np.uint32), y[0].view(np.uint32)), "boxes differ (bits)"


def soup(n, half, seed=1234):
    return _util.gen_scenes.soup_triangles(n, seed=seed, half_extent=half).reshape(-1, 9)


@pytest.mark.parametrize("below", [5, 64, 4096])
def test_soup_200k_equals_host(below):
    tri = soup(200000, 5.85)
    host = build(tri, RTB_BUILD_GPU=0)
    gpu = build(tri, RTB_BUILD_GPU=1, RTB_BUILD_GPU_BELOW=below, RTB_BUILD_GPU_STRICT=1)
    assert not host[3]["on_gpu"] and gpu[3]["on_gpu"] and gpu[3]["levels"] > 3 and gpu[3]["rounds"] > 10
    same(host, gpu)


def test_ties_and_duplicates_equal_host():
    """quantised coordinates (thousands of equal centroids per node) and every triangle present twice: the order among
    equal keys is whatever introsort's partition history leaves, at every level"""
    tri = soup(60000, 4.0, seed=9)
    tri = np.round(tri * 8.0) / 8.0
    tri = np.concatenate([tri, tri[::2]], axis=0).astype(np.float32)
    host = build(tri, RTB_BUILD_GPU=0)
    for below in (5, 1024):
        gpu = build(tri, RTB_BUILD_GPU=1, RTB_BUILD_GPU_BELOW=below, RTB_BUILD_GPU_STRICT=1)
        assert gpu[3]["on_gpu"]
        same(host, gpu)


def test_heap_sorted_ranges_equal_host():
    """introsort's depth limit cut to 3: most ranges run out of depth and are heap-sorted (libstdc++'s __partial_sort),
    on the GPU with the same moves as on the host"""
    tri = soup(50000, 3.7, seed=3)
    host = build(tri, RTB_BUILD_GPU=0, RTB_BUILD_SORT_DEPTH=3)
    gpu = build(tri, RTB_BUILD_GPU=1, RTB_BUILD_GPU_BELOW=5, RTB_BUILD_SORT_DEPTH=3, RTB_BUILD_GPU_STRICT=1)
    assert gpu[3]["on_gpu"] and gpu[3]["heap_ranges"] > 0
    same(host, gpu)


def test_sponza_stand_in_equals_host():
    """axis-aligned, tie-ridden geometry with real materials: the glTF path (load -> build) on the GPU builder"""
    p = _util.scene_path("sponza_small")
    old = os.environ.get("RTB_BUILD_GPU"), os.environ.get("RTB_BUILD_GPU_BELOW")
    try:
        os.environ["RTB_BUILD_GPU"] = "0"
        h = rtb.HostScene.load_gltf(p, 64, 64, 1)
        os.environ["RTB_BUILD_GPU"] = "1"
        os.environ["RTB_BUILD_GPU_BELOW"] = "5"
        g = rtb.HostScene.load_gltf(p, 64, 64, 1)
    finally:
        for k, v in zip(("RTB_BUILD_GPU", "RTB_BUILD_GPU_BELOW"), old):
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v
    assert g.build_info()["on_gpu"]
    ha, hl = h.nodes()
    ga, gl = g.nodes()
    assert np.array_equal(ha.view(np.uint32), ga.view(np.uint32)) and np.array_equal(hl, gl)
    assert np.array_equal(h.tri_source(), g.tri_source())


def test_cfg3_one_million_default_path():
    """1 M-triangle soup (BASELINE cfg3): the default policy sends it to the GPU builder; equal to the host build, which
    test_gpu_baseline_sizes.py pins against the compiled reference at this size"""
    tri = soup(1000000, 10.0)
    gpu = build(tri)
    assert gpu[3]["on_gpu"], "scenes of >= 1 M triangles build on the GPU by default"
    host = build(tri, RTB_BUILD_GPU=0)
    same(host, gpu)
    print("1 M build: host %.3f s, GPU path %.3f s (%s)" % (host[4], gpu[4], gpu[3]))


def test_build_bvh_entry_point_on_gpu():
    """rtb_build_bvh, the drop-in for BVH::buildBVH alone: 600 k triangles go to the GPU builder by default; same answer as the host"""
    tri = soup(600000, 8.4, seed=21)
    pos = tri.copy()
    pos[:, 3:6] = tri[:, 3:6] - tri[:, 0:3]
    pos[:, 6:9] = tri[:, 6:9] - tri[:, 0:3]
    o1, a1, l1, info = rtb.build_bvh(pos)
    assert info["on_gpu"]
    os.environ["RTB_BUILD_GPU"] = "0"
    try:
        o0, a0, l0, info0 = rtb.build_bvh(pos)
    finally:
        os.environ.pop("RTB_BUILD_GPU", None)
    assert not info0["on_gpu"]
    assert np.array_equal(o0, o1) and np.array_equal(l0, l1) and np.array_equal(a0.view(np.uint32), a1.view(np.uint32))


def test_cfg5_scene_equals_host():
    """BASELINE cfg5's scene (10 M-triangle soup + the emissive triangle, through the glTF loader): the GPU build -- which meets
    introsort ranges that run out of depth here and heap-sorts them -- equals the host build, nodes (bits), links, triangle order"""
    p = _util.scene_path("soup_10m")
    os.environ["RTB_BUILD_GPU"] = "1"
    try:
        g = rtb.HostScene.load_gltf(p, 64, 64, 1)
        gi = g.build_info()
        ga, gl = g.nodes()
        gs = g.tri_source().copy()
        import hashlib
        gd = (hashlib.sha256(ga.tobytes()).hexdigest(), hashlib.sha256(gl.tobytes()).hexdigest())
        gsec = g.build_seconds
        del g, ga, gl
        os.environ["RTB_BUILD_GPU"] = "0"
        h = rtb.HostScene.load_gltf(p, 64, 64, 1)
        ha, hl = h.nodes()
        hd = (hashlib.sha256(ha.tobytes()).hexdigest(), hashlib.sha256(hl.tobytes()).hexdigest())
        assert gi["on_gpu"] and not h.build_info()["on_gpu"]
        assert np.array_equal(gs, h.tri_source()), "triangle order differs"
        assert gd == hd, "nodes differ"
        print("cfg5 build: host %.2f s, GPU path %.2f s, %s" % (h.build_seconds, gsec, gi))
    finally:
        os.environ.pop("RTB_BUILD_GPU", None)
