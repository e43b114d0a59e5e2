"""Parity of the CUDA closest-hit path (rtb_closest_hit / rtb_closest_hit_device, replacing BVH::intersect) through
the C ABI, on the B200: bit-exact triangle ids (and t, u, v, inside) against the reference's golden answers, the C
twin and -- when it travelled to the box -- the compiled reference itself."""
import numpy as np
import pytest

import _util

rtb = _util.rtb
pytestmark = pytest.mark.gpu
SCENES = _util.SMALL_SCENES + ["sponza_small"]

_dev = {}


def device_scene(hs):
    if id(hs) not in _dev:
        _dev[id(hs)] = rtb.Scene(hs, 0)
    return _dev[id(hs)]


@pytest.mark.parametrize("name", SCENES)
@pytest.mark.parametrize("cull", [rtb.CULL_REFERENCE, rtb.CULL_GLOBAL])
def test_golden_bit_exact(name, cull, host_scene):
    g = _util.load_golden(name)
    sc = device_scene(host_scene(name))
    rays = _util.box_rays(g["root_box"], g["hit_id"].shape[0], seed=1000)
    h = sc.closest_hit(rays, cull=cull)
    assert np.array_equal(h["id"], g["hit_id"])
    for k in ("t", "u", "v", "inside"):
        assert np.array_equal(h[k], g["hit_" + k]), k
    h = sc.closest_hit(_util.camera_rays(_util.golden_camera(g), g["cam_hit_id"].shape[0], seed=2000), cull=cull)
    assert np.array_equal(h["id"], g["cam_hit_id"]) and np.array_equal(h["t"], g["cam_hit_t"])


@pytest.mark.parametrize("n_tris,half,n_rays", [(100_000, 4.64, 1_000_000), (1_000_000, 10.0, 2_000_000)])
def test_soup_vs_twin(n_tris, half, n_rays):
    """cfg3's generator at sizes the CPU oracle finishes in seconds; 1e-5 relative on t is the stated
    tolerance, the test demands equality"""
    tri = _util.gen_scenes.soup_triangles(n_tris, seed=1234, half_extent=half).reshape(-1, 9)
    rays = _util.gen_scenes.soup_rays(n_rays, seed=42, half_extent=half)
    hs = rtb.HostScene.from_triangles(tri)
    sc = rtb.Scene(hs, 0)
    want = _util.twin_bvh(hs).closest_hit(rays, counters=True)
    assert 0.5 < np.mean(want["id"] >= 0) < 1.0
    ref_cull = sc.closest_hit(rays, cull=rtb.CULL_REFERENCE, counters=True)
    glob = sc.closest_hit(rays, cull=rtb.CULL_GLOBAL, counters=True)
    for got in (ref_cull, glob):
        assert np.array_equal(got["id"], want["id"])
        hit = want["id"] >= 0
        assert np.all(np.abs(got["t"][hit] - want["t"][hit]) <= 1e-5 * np.abs(want["t"][hit]))
        for k in ("t", "u", "v", "inside"):
            assert np.array_equal(got[k], want[k]), k
    # CULL_REFERENCE performs exactly the reference's tests.  CULL_GLOBAL (the ray queue with parked leaves) tests at
    # least what the in-order global-best traversal tests, and not much more
    assert (ref_cull["stats"]["box_tests"], ref_cull["stats"]["tri_tests"]) == (want["box_tests"], want["tri_tests"])
    inorder = _util.devsim_closest_hit(hs, rays, rtb.CULL_GLOBAL)
    assert inorder["box_tests"] < ref_cull["stats"]["box_tests"]
    assert inorder["box_tests"] <= glob["stats"]["box_tests"] <= 1.5 * inorder["box_tests"]
    assert inorder["tri_tests"] <= glob["stats"]["tri_tests"] <= 1.5 * inorder["tri_tests"]


@pytest.mark.ref
def test_soup_vs_compiled_reference(ref):
    tri = _util.gen_scenes.soup_triangles(100_000, seed=77, half_extent=4.64).reshape(-1, 9)
    rays = _util.gen_scenes.soup_rays(500_000, seed=78, half_extent=4.64)
    sc = rtb.Scene(rtb.HostScene.from_triangles(tri), 0)
    want = ref.RefScene.from_triangles(tri).closest_hit(rays)
    got = sc.closest_hit(rays)
    assert np.array_equal(got["id"], want["id"].astype(np.int32))
    for k in ("t", "u", "v", "inside"):
        assert np.array_equal(got[k], want[k]), k


def test_edge_cases():
    tri = np.array([[0, 0, 0, 1, 0, 0, 0, 1, 0], [0, 0, 1, 1, 0, 1, 0, 1, 1], [5, 5, 5, 5, 5, 5, 5, 5, 5]], np.float32)
    hs = rtb.HostScene.from_triangles(tri)
    sc = rtb.Scene(hs, 0)
    rays = np.array([[0.25, 0.25, -1, 0, 0, 1], [0.25, 0.25, 0.5, 0, 0, 1], [0.25, 0.25, 0.5, 0, 0, -1],
                     [0.25, 0.25, 2, 0, 0, 1], [-1, 0.25, 0, 1, 0, 0], [0.25, 0.25, 0, 0, 0, 1], [2, 2, -1, 0, 0, 1],
                     [0.25, 0.25, -1, 0, 0, 1e-30], [0.1, 0.1, -3, 1e-9, 0, 1]], np.float32)
    want = _util.twin_bvh(hs).closest_hit(rays)
    for cull in (0, 1):
        got = sc.closest_hit(rays, cull=cull)
        for k in ("id", "t", "u", "v", "inside"):
            assert np.array_equal(got[k], want[k]), (cull, k)
    # empty batch, single ray, ragged batch sizes around the block size, optional outputs omitted
    out = sc.closest_hit(np.zeros((0, 6), np.float32))
    assert out["id"].shape == (0,)
    for n in (1, 127, 128, 129):
        r = np.repeat(rays[:1], n, axis=0)
        o = dict(id=np.full(n, 99, np.int32))
        got = sc.closest_hit(r, out=o)
        assert np.all(got["id"] == want["id"][0])
    # a one-triangle scene: the root is a leaf
    one = rtb.Scene(rtb.HostScene.from_triangles(tri[:1]), 0)
    assert one.closest_hit(rays[:1])["id"][0] == 0


def test_device_pointer_entry_matches_host_entry(host_scene):
    import torch
    g = _util.load_golden("s6_1")
    sc = device_scene(host_scene("s6_1"))
    rays = _util.box_rays(g["root_box"], 4096, seed=1000)
    d_rays = torch.from_numpy(rays).cuda()
    d_id = torch.empty(4096, dtype=torch.int32, device="cuda")
    d_t = torch.empty(4096, dtype=torch.float32, device="cuda")
    d_cnt = torch.zeros(2, dtype=torch.int64, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    sc.closest_hit_device(d_rays, 4096, rtb.CULL_GLOBAL, d_id, d_t=d_t, d_counters=d_cnt, stream=st)
    torch.cuda.synchronize()
    assert np.array_equal(d_id.cpu().numpy(), g["hit_id"]) and np.array_equal(d_t.cpu().numpy(), g["hit_t"])
    assert int(d_cnt[0]) > 4096


def test_full_size_properties():
    """cfg3 at BASELINE.json's size (1 M triangles, 64 M rays) through size-independent properties: the two culling
    policies agree on every id and t; re-running is idempotent; a 2 M-ray prefix equals the CPU oracle."""
    import torch
    n_tris, chunk, n_chunks = 1_000_000, 4_000_000, 16
    tri = _util.gen_scenes.soup_triangles(n_tris, seed=1234, half_extent=10.0).reshape(-1, 9)
    hs = rtb.HostScene.from_triangles(tri)
    sc = rtb.Scene(hs, 0)
    st = torch.cuda.current_stream().cuda_stream
    total_hits, checksum = 0, 0
    first = None
    for c in range(n_chunks):
        rays = _util.gen_scenes.soup_rays(chunk, seed=42, half_extent=10.0, start=c * chunk)
        d_rays = torch.from_numpy(rays).cuda()
        ids = [torch.empty(chunk, dtype=torch.int32, device="cuda") for _ in range(3)]
        ts = [torch.empty(chunk, dtype=torch.float32, device="cuda") for _ in range(3)]
        for k, cull in enumerate((rtb.CULL_GLOBAL, rtb.CULL_REFERENCE, rtb.CULL_GLOBAL)):
            sc.closest_hit_device(d_rays, chunk, cull, ids[k], d_t=ts[k], stream=st)
        torch.cuda.synchronize()
        assert torch.equal(ids[0], ids[1]) and torch.equal(ts[0], ts[1]), "culling policies disagree in chunk %d" % c
        assert torch.equal(ids[0], ids[2]) and torch.equal(ts[0], ts[2])
        total_hits += int((ids[0] >= 0).sum())
        checksum ^= int(ids[0].to(torch.int64).sum())
        if c == 0:
            first = (rays[:2_000_000], ids[0][:2_000_000].cpu().numpy(), ts[0][:2_000_000].cpu().numpy())
    assert 0.90 < total_hits / (chunk * n_chunks) < 0.93       # ~92 % hit rate
    want = _util.twin_bvh(hs).closest_hit(first[0])
    assert np.array_equal(first[1], want["id"]) and np.array_equal(first[2], want["t"])


# RTB_CULL_GLOBAL against the reference's recursion on rays aimed at shared vertices / edges: {scene: (rays whose id differs,
# of 200 000; largest relative gap of t -- ours is the smaller one)}.  Not exact ties: near-ties.
NEAR_TIES = {"s7_1": (205, 1.4e-5), "s7_4": (904, 5.6e-5), "s6_1": (212, 5.3e-6), "soup_10k": (0, 0.0), "sponza_small": (471, 3.4e-4)}


@pytest.mark.parametrize("name", SCENES)
def test_decision_boundary_rays(name, host_scene):
    """rays ON the box test's decision boundaries (zero direction components, origins on box planes, rays through box
    corners; tests/_util.py: edge_rays): the queue kernel (form 3: box_test_z, one pop / decode site, 256-bit loads)
    against the in-order global-best traversal of the same arithmetic -- every bit -- and against the reference's
    recursion (twin): same hit / miss, equal t,u,v where the ids agree, the rest are the documented near-tie deviation
    (NEAR_TIES: measured count and gap per scene)"""
    g = _util.load_golden(name)
    hs = host_scene(name)
    rays = _util.edge_rays(hs, g["root_box"], 200_000, seed=79)
    got = rtb.Scene(hs, 0).closest_hit(rays, cull=rtb.CULL_GLOBAL)
    first = _util.devsim_closest_hit_lane(hs, rays, 0)
    for k in ("id", "t", "u", "v"):
        assert np.array_equal(got[k], first[k]), k
    want = _util.twin_bvh(hs).closest_hit(rays)
    assert np.array_equal(got["id"] >= 0, want["id"] >= 0)
    same = got["id"] == want["id"]
    hit = want["id"] >= 0
    for k in ("t", "u", "v", "inside"):
        assert np.array_equal(got[k][hit & same], want[k][hit & same]), k
    tie = hit & ~same
    # measured on these seeded batches (200 000 rays each; the CPU mirror of the kernel gives the same numbers): a regression
    # in the culling bound shows as a larger count or a wider gap
    assert np.count_nonzero(tie) <= NEAR_TIES[name][0], (name, np.count_nonzero(tie))
    assert np.all(got["t"][tie] <= want["t"][tie])
    if tie.any():
        assert float(np.max((want["t"][tie] - got["t"][tie]) / want["t"][tie])) <= NEAR_TIES[name][1]
    ref = rtb.Scene(hs, 0).closest_hit(rays, cull=rtb.CULL_REFERENCE)
    for k in ("id", "t", "u", "v", "inside"):
        assert np.array_equal(ref[k], want[k]), k


def test_big_leaves_and_exact_ties():
    """leaves of more than 7 triangles (coincident triangles the builder cannot split) and the exact ties they produce"""
    import test_devsim
    hs = test_devsim._big_leaf_scene()
    rs = np.random.RandomState(8)
    rays = np.empty((200000, 6), np.float32)
    rays[:, 0:3] = rs.rand(200000, 3) * 1.4 - 0.2
    rays[:, 3:6] = rs.rand(200000, 3) * 2 - 1
    rays[:50000, 0:2] = rs.rand(50000, 2) * 0.5 + 0.15
    rays[:50000, 2] = -1
    rays[:50000, 3:6] = [0, 0, 1]
    want = _util.twin_bvh(hs).closest_hit(rays)
    sc = rtb.Scene(hs, 0)
    for cull in (rtb.CULL_REFERENCE, rtb.CULL_GLOBAL):
        got = sc.closest_hit(rays, cull=cull)
        for k in ("id", "t", "u", "v", "inside"):
            assert np.array_equal(got[k], want[k]), (cull, k)
