"""The sample-range split (rtb_render_params::substreams) and the multi-GPU layer of librtb (rtb_comm_*, rtb_render_dist,
rtb_render_multi) on the B200.  On a one-GPU box the collective paths run with one rank; with >= 2 visible devices
(gpurun --gpus N) the NCCL combine itself is checked: tile split == the one-GPU frame bit for bit, sample split ==
the one-GPU K-stream frame up to the rounding of the final sum."""
import os
import subprocess

import numpy as np
import pytest

import _util

rtb = _util.rtb
pytestmark = pytest.mark.gpu
NDEV = rtb.device_count()


def test_substreams_equal_host_state_machine(host_scene):
    """K streams per pixel on the GPU == the host build of the same integrator, bit for bit; streams rendered by
    separate calls add up to the K-stream frame; stream 0 is the reference's own stream"""
    hs = host_scene("s7_4")
    sc = rtb.Scene(hs, 0)
    W, H, S, K = 40, 30, 12, 4
    got = sc.render(W, H, S, substreams=K)
    want = _util.devsim_render(hs, W, H, S, substreams=K)
    assert np.array_equal(got["accum"], want["accum"]) and np.array_equal(got["rgb"], want["rgb"])
    parts = [sc.render(W, H, S, substreams=K, substream_range=(k, k + 1)) for k in range(K)]
    assert all(p["rgb"] is None for p in parts)
    tot = parts[0]["accum"].copy()
    for p in parts[1:]:
        tot = tot + p["accum"]
    assert np.array_equal(tot, got["accum"])
    assert np.array_equal(parts[0]["accum"], sc.render(W, H, S // K)["accum"])
    assert got["stats"]["closest_hit_rays"] == sum(p["stats"]["closest_hit_rays"] for p in parts)
    # K = 1 / 0 is the bit-exact mode; more streams than samples leaves the surplus streams empty
    assert np.array_equal(sc.render(W, H, S, substreams=1)["accum"], sc.render(W, H, S)["accum"])
    many = sc.render(W, H, 3, substreams=8)
    assert np.isfinite(many["accum"]).all() and many["stats"]["closest_hit_rays"] >= W * H * 3
    # misuse is refused
    for bad in (dict(substreams=4, megakernel=True), dict(substreams=4, substream_range=(2, 5)), dict(substreams=4, substream_range=(3, 3))):
        with pytest.raises(rtb.RtbError):
            sc.render(W, H, S, **bad)
    with pytest.raises(rtb.RtbError):   # partial streams cannot be resolved
        p = rtb.frame_params(W, H, S, substreams=4)
        p.substream_begin, p.substream_end = 0, 2
        sc.render(W, H, S, params=p, want_rgb=True)


def test_substreams_statistical_gate(ref, host_scene):
    """the sample split is a different but statistically equivalent stream: same gate as the as-shipped comparison"""
    W, H, S = 96, 96, 32
    p = _util.scene_path("s7_4")
    rs = ref.RefScene.from_gltf(p, W, H, S)
    r1, _ = rs.render()
    r2, _ = rs.render()
    sc = rtb.Scene(rtb.HostScene.load_gltf(p, W, H, S), 0)
    floor = _util.rmse8(r1, r2)
    lin = lambda img: (img.astype(np.float64) / 255.0) ** 2.2
    m_ref = _util.luminance(lin(r1)).mean()
    for K in (2, 8):
        out = sc.render(W, H, S, substreams=K)
        assert _util.rmse8(out["rgb"], r1) <= 1.25 * floor + 0.5, K
        assert abs(_util.luminance(lin(out["rgb"])).mean() - m_ref) <= 0.02 * m_ref, K


def test_render_dist_one_rank(host_scene):
    """rtb_render_dist without a communicator (one rank) is rtb_render: both splits"""
    sc = rtb.Scene(host_scene("s7_4"), 0)
    W, H, S = 64, 40, 8
    rgb = np.empty((H, W, 3), np.uint8)
    st = rtb.render_dist(sc, None, rtb.frame_params(W, H, S), rtb.SPLIT_TILE, rgb)
    want = sc.render(W, H, S)
    assert np.array_equal(rgb, want["rgb"]) and st["closest_hit_rays"] == want["stats"]["closest_hit_rays"]
    rtb.render_dist(sc, None, rtb.frame_params(W, H, S, substreams=4), rtb.SPLIT_SAMPLE, rgb)
    assert np.array_equal(rgb, sc.render(W, H, S, substreams=4)["rgb"])
    img, _ = rtb.render_multi([sc], rtb.frame_params(W, H, S), rtb.SPLIT_TILE)
    assert np.array_equal(img, want["rgb"])
    c = rtb.Comm(None, 1, 0, 0)      # a one-rank communicator needs no NCCL
    rtb.render_dist(sc, c, rtb.frame_params(W, H, S), rtb.SPLIT_SAMPLE, rgb)
    assert np.array_equal(rgb, want["rgb"])
    c.close()


@pytest.mark.skipif(NDEV < 2, reason="needs >= 2 CUDA devices (gpurun --gpus N)")
@pytest.mark.parametrize("name,W,H,S", [("s7_4", 200, 150, 8), ("sponza_small", 322, 181, 6)])
def test_render_multi_all_devices(name, W, H, S, host_scene):
    """one rank per visible device inside ONE process (threads), combined over NCCL: what the CLI runs"""
    hs = host_scene(name)
    scenes = [rtb.Scene(hs, d) for d in range(NDEV)]
    one = scenes[0].render(W, H, S)
    img, st = rtb.render_multi(scenes, rtb.frame_params(W, H, S), rtb.SPLIT_TILE)
    assert np.array_equal(img, one["rgb"]), "tile split over %d devices differs from the one-GPU frame" % NDEV
    assert st["closest_hit_rays"] == one["stats"]["closest_hit_rays"]
    K = NDEV            # rtb_render_dist runs max(substreams, world) streams per pixel; with S < K the surplus streams are empty
    kone = scenes[0].render(W, H, S, substreams=K)
    img2, st2 = rtb.render_multi(scenes, rtb.frame_params(W, H, S, substreams=K), rtb.SPLIT_SAMPLE)
    diff = np.abs(img2.astype(np.int32) - kone["rgb"].astype(np.int32))
    assert diff.max() <= 1 and np.count_nonzero(diff) <= 1e-3 * diff.size, (diff.max(), np.count_nonzero(diff))
    assert st2["closest_hit_rays"] == kone["stats"]["closest_hit_rays"]
    # a second frame reuses the kept communicators
    img3, _ = rtb.render_multi(scenes, rtb.frame_params(W, H, S), rtb.SPLIT_TILE)
    assert np.array_equal(img3, img)
    for s in scenes:
        s.close()


@pytest.mark.skipif(NDEV < 2, reason="needs >= 2 CUDA devices (gpurun --gpus N)")
def test_cli_all_devices_equals_one_device(tmp_path):
    """build/solution on every visible device (NCCL gather inside librtb) writes the PPM of the one-device run:
    1920x1080 x 8 spp on the Sponza stand-in; RT_SPLIT=sample stays within the statistical gate of the tile frame"""
    exe = os.path.join(_util.ROOT, "build", "solution")
    scene = _util.scene_path("sponza")
    outs = {}
    for tag, env in (("one", dict(RT_GPUS="1")), ("all", dict(RT_GPUS=str(NDEV))), ("sample", dict(RT_GPUS=str(NDEV), RT_SPLIT="sample"))):
        out = tmp_path / (tag + ".ppm")
        r = subprocess.run([exe, scene, "1920", "1080", "8", str(out)], capture_output=True, text=True, env=dict(os.environ, **env), timeout=900)
        assert r.returncode == 0, r.stderr[-2000:]
        outs[tag] = open(out, "rb").read()
    assert outs["one"] == outs["all"]
    hdr = len(b"P6\n1920 1080\n255\n")
    a = np.frombuffer(outs["one"][hdr:], np.uint8).astype(np.float64)
    b = np.frombuffer(outs["sample"][hdr:], np.uint8).astype(np.float64)
    assert abs(a.mean() - b.mean()) <= 0.02 * a.mean()
