"""rtb_closest_hit_static: the closest-hit batch for callers that must not allocate device memory.

Some users of librtb.so cannot let cudaMalloc / cudaFree run on their thread: a rank thread next to an NCCL collective
in flight (both calls wait for the whole device -- the 8-GPU hang of round 2), a plugin whose host has given all device
memory to its own pool, a frame loop that cannot afford a device-wide synchronisation.  For them the library has one
entry point that allocates nothing: the queue kernel's ray cursor comes from a 16 KB table compiled into the library
instead of from the scene's cudaMalloc'ed cursors (include/rtb.h, csrc/rtb_kernels.cu).

What is checked here, through the C ABI: the answers are those of rtb_closest_hit_device bit for bit (ids, t, u, v,
inside -- integer / decision-exact work, no tolerance), on two calls in a row with batches of different length (the
table keeps the previous call's cursor; a call that relied on it being zero would lose or repeat rays on the second
call), in both culling modes, with the test counters, on a second stream; and that a batch beyond the cursor's range
is refused with a message instead of being served by another kernel.

This file is named so that pytest collects it last: under `pytest -x` a failure here hides nothing.  It builds its own
scene and buffers, so it passes alone in a fresh process as well as after the other tests."""
import ctypes as C

import numpy as np
import pytest

import _util

rtb = _util.rtb

pytestmark = pytest.mark.gpu


def _scene():
    tri = _util.gen_scenes.soup_triangles(20_000, seed=77, half_extent=4.0).reshape(-1, 9)
    return rtb.Scene(rtb.HostScene.from_triangles(tri), 0)


def _run(call, d_rays, n, cull, stream, counters=False):
    import torch
    out = {
        "id": torch.full((n,), -7, dtype=torch.int32, device="cuda"),
        "t": torch.full((n,), -7.0, dtype=torch.float32, device="cuda"),
        "u": torch.full((n,), -7.0, dtype=torch.float32, device="cuda"),
        "v": torch.full((n,), -7.0, dtype=torch.float32, device="cuda"),
        "inside": torch.full((n,), 9, dtype=torch.uint8, device="cuda"),
    }
    cnt = torch.zeros(2, dtype=torch.int64, device="cuda") if counters else None
    torch.cuda.synchronize()   # the fills above ran on torch's stream; `stream` may be another one
    call(d_rays, n, cull, out["id"], d_t=out["t"], d_u=out["u"], d_v=out["v"], d_inside=out["inside"], d_counters=cnt, stream=stream)
    torch.cuda.synchronize()
    res = {k: v.cpu().numpy() for k, v in out.items()}
    if counters:
        res["counters"] = cnt.cpu().numpy()
    return res


def _same(a, b):
    for k in a:
        # bit patterns, not values: -0.0 vs 0.0 or a NaN would count as a difference
        assert np.array_equal(a[k].view(np.uint8), b[k].view(np.uint8)), k


def test_static_entry_equals_device_entry_twice_in_a_row():
    import torch
    sc = _scene()
    side = torch.cuda.Stream()
    streams = [torch.cuda.current_stream().cuda_stream, side.cuda_stream]
    # two calls in a row per configuration, the second one longer and on other rays than the first
    for cull in (rtb.CULL_GLOBAL, rtb.CULL_REFERENCE):
        for st in streams:
            for n, seed in ((50_000, 5), (131_071, 6)):
                rays = _util.gen_scenes.soup_rays(n, seed=seed, half_extent=4.0)
                d_rays = torch.from_numpy(rays).cuda()
                want = _run(sc.closest_hit_device, d_rays, n, cull, st)
                got = _run(sc.closest_hit_static, d_rays, n, cull, st)
                assert (want["id"] >= 0).sum() > n // 10 and (want["id"] < 0).sum() > 0   # hits and misses both occur
                _same(want, got)
    # more calls than the table has cursors (64), so every cursor is used a second time with another call's value in it
    rays = _util.gen_scenes.soup_rays(4_099, seed=9, half_extent=4.0)
    d_rays = torch.from_numpy(rays).cuda()
    want = _run(sc.closest_hit_device, d_rays, 4_099, rtb.CULL_GLOBAL, streams[0])
    for _ in range(70):
        _same(want, _run(sc.closest_hit_static, d_rays, 4_099, rtb.CULL_GLOBAL, streams[0]))
    # the counting instantiation of the kernel
    want = _run(sc.closest_hit_device, d_rays, 4_099, rtb.CULL_GLOBAL, streams[0], counters=True)
    got = _run(sc.closest_hit_static, d_rays, 4_099, rtb.CULL_GLOBAL, streams[0], counters=True)
    assert want["counters"][0] > 4_099
    _same(want, got)


def test_static_entry_allocates_nothing_and_refuses_oversized_batches():
    import torch
    sc = _scene()
    n = 10_000
    d_rays = torch.from_numpy(_util.gen_scenes.soup_rays(n, seed=3, half_extent=4.0)).cuda()
    d_id = torch.empty(n, dtype=torch.int32, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    sc.closest_hit_static(None, 0, rtb.CULL_GLOBAL, None, stream=st)   # the set-up call: n == 0 is enough
    sc.closest_hit_static(d_rays, n, rtb.CULL_GLOBAL, d_id, stream=st)
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    for _ in range(20):
        sc.closest_hit_static(d_rays, n, rtb.CULL_GLOBAL, d_id, stream=st)
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert free1 == free0, "device memory in use changed across rtb_closest_hit_static calls"
    # one ray more than the 32-bit cursor serves: refused before anything is launched or read
    before = d_id.clone()
    with pytest.raises(rtb.RtbError) as err:
        sc.closest_hit_static(d_rays, rtb.CLOSEST_HIT_STATIC_MAX_RAYS + 1, rtb.CULL_GLOBAL, d_id, stream=st)
    assert "rtb_closest_hit_static" in str(err.value) and str(rtb.CLOSEST_HIT_STATIC_MAX_RAYS) in str(err.value)
    torch.cuda.synchronize()
    assert torch.equal(before, d_id)
    rc = rtb.lib().rtb_closest_hit_static(None, None, 0, rtb.CULL_GLOBAL, None, None, None, None, None, None, None)
    assert rc == rtb.RTB_ERR_INVALID
