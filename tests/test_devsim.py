"""CPU-side check of the restated device arithmetic: the device headers (raytracing-hw_b200/csrc/*.cuh) compiled
for the host by nvcc (tests/devsim/) against the reference's golden answers.  This is NOT the parity proof of the
CUDA path (that is tests/test_gpu_*.py, on the B200) -- it catches restatement errors where no GPU exists."""
import ctypes as C

import numpy as np
import pytest

import _util

rtb = _util.rtb
SCENES = _util.SMALL_SCENES + ["sponza_small"]


@pytest.mark.parametrize("name", SCENES)
@pytest.mark.parametrize("cull", [rtb.CULL_REFERENCE, rtb.CULL_GLOBAL])
def test_closest_hit_bit_exact(name, cull, host_scene):
    g = _util.load_golden(name)
    hs = host_scene(name)
    rays = _util.box_rays(g["root_box"], g["hit_id"].shape[0], seed=1000)
    h = _util.devsim_closest_hit(hs, rays, cull)
    assert np.array_equal(h["id"], g["hit_id"])
    for k in ("t", "u", "v", "inside"):
        assert np.array_equal(h[k], g["hit_" + k]), k
    h = _util.devsim_closest_hit(hs, _util.camera_rays(_util.golden_camera(g), g["cam_hit_id"].shape[0], seed=2000), cull)
    assert np.array_equal(h["id"], g["cam_hit_id"]) and np.array_equal(h["t"], g["cam_hit_t"])


@pytest.mark.parametrize("name", SCENES)
def test_queue_lane_state_machine(name, host_scene):
    """the queue kernels' per-lane state machine (rtb_trace.cuh: LaneTraversal, bounded branch-free pops) gives the
    reference's hits and performs exactly the in-order global-best traversal's tests"""
    g = _util.load_golden(name)
    hs = host_scene(name)
    rays = _util.box_rays(g["root_box"], g["hit_id"].shape[0], seed=1000)
    h = _util.devsim_closest_hit_lane(hs, rays, 0)
    assert np.array_equal(h["id"], g["hit_id"])
    for k in ("t", "u", "v"):
        assert np.array_equal(h[k], g["hit_" + k]), k
    inorder = _util.devsim_closest_hit(hs, rays, rtb.CULL_GLOBAL)
    assert (h["box_tests"], h["tri_tests"]) == (inorder["box_tests"], inorder["tri_tests"])
    cam = _util.devsim_closest_hit_lane(hs, _util.camera_rays(_util.golden_camera(g), g["cam_hit_id"].shape[0], seed=2000), 0)
    assert np.array_equal(cam["id"], g["cam_hit_id"]) and np.array_equal(cam["t"], g["cam_hit_t"])


@pytest.mark.parametrize("name", SCENES)
def test_lane_form3_equals_first_form(name, host_scene):
    """form 3 of the lane state machine (one pop / decode site per step, box test without the `inside` flag and the
    zero-direction guards: rtb_traverse.cuh box_test_z) = the reference's hits, with exactly the first form's tests"""
    g = _util.load_golden(name)
    hs = host_scene(name)
    rays = _util.box_rays(g["root_box"], g["hit_id"].shape[0], seed=1000)
    h = _util.devsim_closest_hit_lane(hs, rays, 3)
    assert np.array_equal(h["id"], g["hit_id"])
    for k in ("t", "u", "v"):
        assert np.array_equal(h[k], g["hit_" + k]), k
    first = _util.devsim_closest_hit_lane(hs, rays, 0)
    assert (h["box_tests"], h["tri_tests"]) == (first["box_tests"], first["tri_tests"])
    cam = _util.devsim_closest_hit_lane(hs, _util.camera_rays(_util.golden_camera(g), g["cam_hit_id"].shape[0], seed=2000), 3)
    assert np.array_equal(cam["id"], g["cam_hit_id"]) and np.array_equal(cam["t"], g["cam_hit_t"])


@pytest.mark.parametrize("name", SCENES)
@pytest.mark.parametrize("every", [1, 7, 40])
def test_lane_form3_suspend_resume(name, every, host_scene):
    """a ray suspended every `every` steps and resumed in a fresh lane (LaneTraversal::save3 / load3: what the queue kernel
    does with the tail of a launch) performs the same tests and gives the same answer as the uninterrupted ray"""
    g = _util.load_golden(name)
    hs = host_scene(name)
    rays = np.concatenate([_util.box_rays(g["root_box"], 4096, seed=1000), _util.edge_rays(hs, g["root_box"], 4096, seed=78)])
    whole = _util.devsim_closest_hit_lane(hs, rays, 3)
    cut = _util.devsim_closest_hit_lane(hs, rays, 100 + every)
    assert (cut["box_tests"], cut["tri_tests"]) == (whole["box_tests"], whole["tri_tests"])
    for k in ("id", "t", "u", "v"):
        assert np.array_equal(cut[k], whole[k]), k


@pytest.mark.parametrize("name", SCENES)
@pytest.mark.parametrize("form", [4, 5, 6, 10001, 10007, 20001, 20007])
def test_lane_form3_stack_ring(name, form, host_scene):
    """the top of the form-3 stack in a ring (what the batched queue kernel keeps in shared memory; LaneTraversal::push3 /
    pre3 / entry3) with the array as backing store: a 2-entry ring overflows and refills all the time, an 8-entry one
    rarely; alone and combined with suspend / resume; and the variant with the top entry in registers (6, 2000x).  Same
    tests, same answers as the plain array."""
    g = _util.load_golden(name)
    hs = host_scene(name)
    rays = np.concatenate([_util.box_rays(g["root_box"], 4096, seed=1000), _util.edge_rays(hs, g["root_box"], 4096, seed=78)])
    plain = _util.devsim_closest_hit_lane(hs, rays, 3)
    ring = _util.devsim_closest_hit_lane(hs, rays, form)
    assert (ring["box_tests"], ring["tri_tests"]) == (plain["box_tests"], plain["tri_tests"])
    for k in ("id", "t", "u", "v"):
        assert np.array_equal(ring[k], plain[k]), k


@pytest.mark.parametrize("name", SCENES)
def test_lane_form3_on_decision_boundaries(name, host_scene):
    """rays that sit on the box test's decision boundaries (zero direction components, origins on box planes, rays
    through box corners): form 3 vs the twin (the pinned restatement of the reference's recursion)"""
    g = _util.load_golden(name)
    hs = host_scene(name)
    rays = _util.edge_rays(hs, g["root_box"], 60000, seed=77)
    first = _util.devsim_closest_hit_lane(hs, rays, 0)
    h = _util.devsim_closest_hit_lane(hs, rays, 3)
    # the same visits (box_test_z decides like box_test_t on every box met) and the same answers, bit for bit
    assert (h["box_tests"], h["tri_tests"]) == (first["box_tests"], first["tri_tests"])
    for k in ("id", "t", "u", "v"):
        assert np.array_equal(h[k], first[k]), k
    # against the reference's recursion.  Rays aimed exactly at a shared vertex are the documented near-tie deviation of global-best
    # culling: several triangles answer with parameters a few ulp apart and the reference's
    # sibling-only culling keeps whichever its rounded entry distance lets through; global-best culling returns the
    # smallest.  Everything else must agree exactly.
    want = _util.twin_bvh(hs).closest_hit(rays)
    assert np.array_equal(h["id"] >= 0, want["id"] >= 0)
    hit = want["id"] >= 0
    same = h["id"] == want["id"]
    for k in ("t", "u", "v"):
        assert np.array_equal(h[k][hit & same], want[k][hit & same]), k
    tie = hit & ~same
    assert np.count_nonzero(tie) <= 0.005 * rays.shape[0]
    assert np.all(h["t"][tie] <= want["t"][tie])
    # (a grazing triangle's parameter is ill-conditioned: the two candidates were seen up to 3.4e-4 apart)
    assert np.all(np.abs(h["t"][tie] - want["t"][tie]) <= 1e-3 * want["t"][tie])
    assert 0.05 < np.mean(hit) < 0.999


@pytest.mark.parametrize("name", ["s6_1", "soup_10k"])
def test_reference_cull_counts_equal_twin(name, host_scene):
    """CULL_REFERENCE visits exactly the boxes and triangles the reference's recursion visits"""
    g = _util.load_golden(name)
    hs = host_scene(name)
    rays = _util.box_rays(g["root_box"], 4096, seed=1000)
    want = _util.twin_bvh(hs).closest_hit(rays, counters=True)
    got = _util.devsim_closest_hit(hs, rays, rtb.CULL_REFERENCE)
    assert (got["box_tests"], got["tri_tests"]) == (want["box_tests"], want["tri_tests"])
    glob = _util.devsim_closest_hit(hs, rays, rtb.CULL_GLOBAL)
    assert glob["box_tests"] <= got["box_tests"] and glob["tri_tests"] <= got["tri_tests"]


@pytest.mark.parametrize("name", SCENES)
def test_light_pdf_bit_exact(name, host_scene):
    g = _util.load_golden(name)
    if "light_pdf" not in g.files:
        pytest.skip("scene without emissive triangles")
    assert np.array_equal(_util.devsim_light_pdf(host_scene(name), g["light_rays"]), g["light_pdf"])


@pytest.mark.parametrize("name", SCENES)
def test_paths_bit_exact(name, host_scene):
    """Scene::intersect for explicit rays + engine seeds: same radiance bits, same engine state afterwards"""
    g = _util.load_golden(name)
    hs = host_scene(name)
    rays = _util.camera_rays(_util.golden_camera(g), g["path_rgb"].shape[0], seed=4000)
    bad = 0
    for i in range(rays.shape[0]):
        rgb, after = _util.devsim_trace_path(hs, rays[i, 0:3], rays[i, 3:6], 1 + i * 7919)
        if after != int(g["path_engine_after"][i]) or not np.array_equal(rgb, g["path_rgb"][i]):
            bad += 1
    assert bad == 0


@pytest.mark.parametrize("name", SCENES)
@pytest.mark.parametrize("reference_cull", [False, True])
def test_render_bit_exact(name, reference_cull, host_scene):
    """the whole per-pixel loop (seeding, jitter, camera, bounce loop, accumulation, resolve) vs the reference's
    deterministic render (per-pixel normal-distribution state, pixel 0 seeded with 1)"""
    g = _util.load_golden(name)
    hs = host_scene(name)
    H, W, _ = g["render_accum"].shape
    out = _util.devsim_render(hs, W, H, 4, pixel0_seed=1, reference_cull=reference_cull)
    same = np.all(out["accum"] == g["render_accum"], axis=-1)
    assert same.all(), "%d of %d pixels differ" % (np.count_nonzero(~same), same.size)
    assert np.array_equal(out["rgb"], g["render_rgb"])
    assert out["rays"] >= W * H * 4


@pytest.mark.parametrize("name", SCENES)
def test_wavefront_steps_bit_exact(name, host_scene):
    """the wavefront integrator's per-slot steps (rtb_wavefront.cuh), driven on the host, give the sums and bytes of
    the reference's deterministic render -- and the same ray / light-query / test counts as the megakernel form"""
    g = _util.load_golden(name)
    hs = host_scene(name)
    H, W, _ = g["render_accum"].shape
    out = _util.devsim_render(hs, W, H, 4, pixel0_seed=1, wavefront=True)
    same = np.all(out["accum"] == g["render_accum"], axis=-1)
    assert same.all(), "%d of %d pixels differ" % (np.count_nonzero(~same), same.size)
    assert np.array_equal(out["rgb"], g["render_rgb"])
    mega = _util.devsim_render(hs, W, H, 4, pixel0_seed=1)
    for k in ("rays", "light_queries", "box_tests", "tri_tests"):
        assert out[k] == mega[k], k


def test_wavefront_row_ranges_and_interleave(host_scene):
    """slots map to the same pixels (and seeds) under a row range and under the multi-GPU row-block interleave"""
    hs = host_scene("s7_4")
    W, H, S = 40, 36, 3
    full = _util.devsim_render(hs, W, H, S, wavefront=True)["accum"]
    part = _util.devsim_render(hs, W, H, S, rows=(8, 29), wavefront=True)["accum"]
    assert np.array_equal(part, full[8:29])
    for idx in range(3):
        il = _util.devsim_render(hs, W, H, S, wavefront=True, interleave=(3, idx, 4))["accum"]
        rows = [r for r in range(H) if (r // 4) % 3 == idx]
        assert np.array_equal(il, full[rows])


def test_libm_known_answers():
    """glibc logf / powf / pow(x,5) restatements (csrc/rtb_libm.cuh) against the reference host's libm"""
    import os
    import subprocess
    exe = os.path.join(_util.ROOT, "tests", "devsim", "libm_check")
    if not os.path.exists(exe):
        subprocess.run(["make", "-C", os.path.dirname(exe), "libm_check"], check=True, stdout=subprocess.DEVNULL)
    r = subprocess.run([exe, "1009"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout


def test_gpu_builder_sort_replay_equals_std_sort():
    """the GPU builder's sort (csrc/rtb_build_steps.cuh: introsort replayed breadth first, one partition step of every live
    range per round; heap-sorted ranges with libstdc++'s own moves) driven by plain loops and compared with std::sort /
    std::__partial_sort themselves on distinct, tied, sorted, reversed, all-equal and organ-pipe keys, one and many segments"""
    import os
    import subprocess
    exe = os.path.join(_util.ROOT, "tests", "devsim", "build_check")
    subprocess.run(["make", "-C", os.path.dirname(exe), "build_check"], check=True, stdout=subprocess.DEVNULL)
    r = subprocess.run([exe, "4"], capture_output=True, text=True)
    assert r.returncode == 0 and r.stdout.startswith("ok"), r.stdout


def test_gpu_builder_source_emulated_on_the_host_equals_host_builder():
    """csrc/rtb_build.cu itself -- kernels and level loop, the source the B200 runs -- compiled for the host (tests/devsim/emul:
    a launch runs its threads one after the other over host memory) against the host builder behind rtb_build_bvh: node boxes
    (bits), links, split axes, triangle order, on random / quantised / duplicated soups, a flat axis-aligned grid with zeros of
    both signs, coincident triangles (no cut improves: std::partition on z in the node's thread; handed back above 256), tiny
    scenes and forced depth limits (heap-sorted ranges); then the hand-over between the two builders (host/bvh_builder.cpp:
    adopt_all) with the emulated builder in the GPU's place, hand-over sizes 5 / 64 / 512"""
    import os
    import subprocess
    d = os.path.join(_util.ROOT, "tests", "devsim")
    subprocess.run(["make", "-C", d, "build_emul_check"], check=True, stdout=subprocess.DEVNULL)
    r = subprocess.run([os.path.join(d, "build_emul_check")], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and r.stdout.strip().endswith("ok") and "FAILED" not in r.stdout, r.stdout[-3000:]


def _big_leaf_scene():
    """40 + 9 coincident triangles (their centroids cannot be split: leaves of 40 and 9 triangles, and every ray that hits
    them produces exact ties, which the first triangle in the reference's order must win) among 300 random ones"""
    rs = np.random.RandomState(7)
    a = np.array([[0.2, 0.2, 0.5, 0.6, 0.0, 0.0, 0.0, 0.6, 0.0]], np.float32)
    b = np.array([[0.1, 0.5, 0.3, 0.5, 0.1, 0.1, 0.0, 0.4, 0.2]], np.float32)
    soup = rs.rand(300, 9).astype(np.float32)
    soup[:, 3:] = (soup[:, 3:] - 0.5) * 0.3
    tri = np.concatenate([np.repeat(a, 40, axis=0), soup[:150], np.repeat(b, 9, axis=0), soup[150:]])
    return rtb.HostScene.from_triangles(tri)


def test_big_leaves_and_exact_ties():
    hs = _big_leaf_scene()
    _, links = hs.nodes()
    assert links[:, 4].max() == 40 and np.count_nonzero(links[:, 4] > 7) >= 2
    rs = np.random.RandomState(8)
    rays = np.empty((20000, 6), np.float32)
    rays[:, 0:3] = rs.rand(20000, 3) * 1.4 - 0.2
    rays[:, 3:6] = rs.rand(20000, 3) * 2 - 1
    rays[:5000, 0:2] = rs.rand(5000, 2) * 0.5 + 0.15      # aimed at the coincident stack
    rays[:5000, 2] = -1
    rays[:5000, 3:6] = [0, 0, 1]
    want = _util.twin_bvh(hs).closest_hit(rays, counters=True)
    assert np.count_nonzero(want["id"][:5000] >= 0) > 1000
    for form in (0, 3, 105):
        h = _util.devsim_closest_hit_lane(hs, rays, form)
        for k in ("id", "t", "u", "v"):
            assert np.array_equal(h[k], want[k]), (form, k)
    ordered = _util.devsim_closest_hit(hs, rays, rtb.CULL_REFERENCE)
    assert (ordered["box_tests"], ordered["tri_tests"]) == (want["box_tests"], want["tri_tests"])


def test_wavefront_resume_is_bit_identical(host_scene):
    """rtb_render_resume's contract on the host build of the wavefront integrator: a frame rendered in chunks of samples
    through the 32-byte per-pixel state (engine, cached normal half, sum) = the frame rendered at once, bit for bit --
    and every intermediate image is the one-shot image at that sample count"""
    hs = host_scene("s6_1")
    W, H = 24, 16
    whole = _util.devsim_render_resume(hs, W, H, 7)
    part = _util.devsim_render_resume(hs, W, H, 2)
    assert np.array_equal(part["rgb"], _util.devsim_render(hs, W, H, 2, pixel0_seed=1, wavefront=True)["rgb"])
    assert np.all(part["state"][..., 3] == 2)
    part = _util.devsim_render_resume(hs, W, H, 3, state=part["state"], sample_begin=2)
    assert np.array_equal(part["accum"], _util.devsim_render(hs, W, H, 3, pixel0_seed=1, wavefront=True)["accum"])
    part = _util.devsim_render_resume(hs, W, H, 7, state=part["state"], sample_begin=3)
    assert np.array_equal(part["accum"], whole["accum"]) and np.array_equal(part["rgb"], whole["rgb"])
    assert np.array_equal(part["state"], whole["state"])


def test_wavefront_resume_across_a_different_split(host_scene):
    """the per-pixel state is indexed by the call's local rows: a frame started on a 3-way row-block split (3 GPUs),
    its states merged into global row order (what the CLI's checkpoint file holds), carries on as one range -- and the
    other way round -- with the bits of the one-call frame"""
    hs = host_scene("s6_1")
    W, H, R = 24, 40, 8
    whole = _util.devsim_render_resume(hs, W, H, 5)
    p = rtb.RenderParams(W, H, 2, 0, H, 1, 0, 0)
    full = np.zeros((H, W, 8), np.uint32)
    for idx in range(3):
        p.interleave_count, p.interleave_index, p.interleave_rows = 3, idx, R
        part = _util.devsim_render_resume(hs, W, H, 2, interleave=(3, idx, R))
        for lr in range(part["state"].shape[0]):
            full[rtb.lib().rtb_render_global_row(C.byref(p), lr)] = part["state"][lr]
    assert np.array_equal(full, _util.devsim_render_resume(hs, W, H, 2)["state"])
    rest = _util.devsim_render_resume(hs, W, H, 5, state=full, sample_begin=2)
    assert np.array_equal(rest["accum"], whole["accum"]) and np.array_equal(rest["state"], whole["state"])
    # and back onto a 2-way split
    p2 = rtb.RenderParams(W, H, 5, 0, H, 1, 2, 0)
    for idx in range(2):
        p2.interleave_count, p2.interleave_index, p2.interleave_rows = 2, idx, R
        n = rtb.lib().rtb_render_local_rows(C.byref(p2))
        rows = [rtb.lib().rtb_render_global_row(C.byref(p2), lr) for lr in range(n)]
        part = _util.devsim_render_resume(hs, W, H, 5, state=full[rows], sample_begin=2, interleave=(2, idx, R))
        assert np.array_equal(part["accum"], whole["accum"][rows]) and np.array_equal(part["rgb"], whole["rgb"][rows])
