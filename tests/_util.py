"""Shared helpers of the test-suite: module loading, stand-in scenes, seeded ray batches, the devsim binding."""
import ctypes as C
import importlib
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
if os.path.join(ROOT, "scenes") not in sys.path:
    sys.path.insert(0, os.path.join(ROOT, "scenes"))

import gen_scenes  # noqa: E402

rtb = importlib.import_module("raytracing-hw_b200")

SMALL_SCENES = ["s7_1", "s7_4", "s6_1", "soup_10k"]


def scene_path(name):
    return gen_scenes.ensure(name)


def box_rays(root_box, n, seed):
    """n rays: origin uniform in the root box grown by 25 %, direction uniform in [-1,1]^3 (un-normalised).
    Counter-based (splitmix64), so the batch is a pure function of (box, n, seed)."""
    u = gen_scenes.splitmix_u01(seed, 0, n * 6).reshape(n, 6)
    lo = np.asarray(root_box[:3], np.float32)
    hi = np.asarray(root_box[3:], np.float32)
    c = (lo + hi) * np.float32(0.5)
    h = (hi - lo) * np.float32(0.625)
    r = np.empty((n, 6), np.float32)
    r[:, 0:3] = c + (u[:, 0:3] * np.float32(2) - np.float32(1)) * h
    r[:, 3:6] = u[:, 3:6] * np.float32(2) - np.float32(1)
    return r


def edge_rays(hs, root_box, n, seed):
    """n rays built to sit ON the decision boundaries of the box test: axis-parallel and in-plane directions (zero
    components), origins at triangle vertices (i.e. exactly on node-box planes), rays aimed exactly at vertices (the
    corners of leaf boxes), origins one exact unit behind a box plane.  Pure function of (scene, n, seed)."""
    u = gen_scenes.splitmix_u01(seed, 0, n * 8).reshape(n, 8)
    pos = hs.tris()["pos"]
    p0 = pos[:, 0:3]
    verts = np.concatenate([p0, p0 + pos[:, 3:6], p0 + pos[:, 6:9]]).astype(np.float32)
    lo = np.asarray(root_box[:3], np.float32)
    hi = np.asarray(root_box[3:], np.float32)
    rnd = ((lo + hi) * np.float32(0.5) + (u[:, 0:3] * np.float32(2) - np.float32(1)) * (hi - lo) * np.float32(0.625)).astype(np.float32)
    va = verts[np.minimum((u[:, 3] * len(verts)).astype(np.int64), len(verts) - 1)]
    vb = verts[np.minimum((u[:, 4] * len(verts)).astype(np.int64), len(verts) - 1)]
    axis = np.minimum((u[:, 5] * 3).astype(np.int64), 2)
    sign = np.where(u[:, 6] < 0.5, np.float32(-1), np.float32(1)).astype(np.float32)
    e = np.zeros((n, 3), np.float32)
    e[np.arange(n), axis] = sign
    kind = np.minimum((u[:, 7] * 6).astype(np.int64), 5)
    r = np.empty((n, 6), np.float32)
    rd = (u[:, 0:3] * np.float32(2) - np.float32(1)).astype(np.float32)
    planar = rd.copy()
    planar[np.arange(n), axis] = 0
    for k, (o, d) in enumerate([(va, e),                      # from a vertex along an axis
                                (rnd, e),                     # anywhere, along an axis
                                (rnd, planar),                # one zero component
                                (rnd, va - rnd),              # aimed exactly at a vertex
                                (va - e, e),                  # exactly one unit behind a vertex, along the axis through it
                                (va, vb - va)]):              # vertex to vertex
        m = kind == k
        r[m, 0:3] = o[m]
        r[m, 3:6] = d[m]
    bad = np.all(r[:, 3:6] == 0, axis=1)
    r[bad, 3:6] = np.float32(1)
    return r


def camera_rays(cam, n, seed):
    """n rays from the camera position into the view cone (what primary rays look like), un-normalised."""
    u = gen_scenes.splitmix_u01(seed, 0, n * 2).reshape(n, 2)
    ax = cam["axes"]
    tx = (u[:, 0:1] * 2 - 1) * cam["tan_half_fov"][0]
    ty = (u[:, 1:2] * 2 - 1) * cam["tan_half_fov"][1]
    d = tx * ax[0][None, :] + ty * ax[1][None, :] + ax[2][None, :]
    r = np.empty((n, 6), np.float32)
    r[:, 0:3] = cam["pos"][None, :]
    r[:, 3:6] = d
    return r


def golden_camera(g):
    """the camera a golden file was generated with (position, axes, tan(fov/2) as make_golden.py evaluated it)"""
    return dict(pos=g["cam_pos"], axes=g["cam_axes"], tan_half_fov=g["cam_tan_half_fov"])


def load_golden(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"))


# ------------------------------------------------------------------------------------------------ devsim
_devsim = None


def devsim():
    """tests/devsim/libdevsim.so: the device headers compiled for the host (CPU-side check of the restated
    arithmetic; test tooling, not a fallback of the product)."""
    global _devsim
    if _devsim is None:
        path = os.path.join(ROOT, "tests", "devsim", "libdevsim.so")
        if not os.path.exists(path):
            subprocess.run(["make", "-C", os.path.dirname(path), "libdevsim.so"], check=True, stdout=subprocess.DEVNULL)
        L = C.CDLL(path)
        vp = C.c_void_p
        L.devsim_last_error.restype = C.c_char_p
        L.devsim_closest_hit.argtypes = [C.POINTER(rtb.SceneDesc), vp, C.c_size_t, C.c_int, vp, vp, vp, vp, vp, vp]
        L.devsim_closest_hit_lane.argtypes = [C.POINTER(rtb.SceneDesc), vp, C.c_size_t, C.c_uint, vp, vp, vp, vp, vp]
        L.devsim_light_pdf.argtypes = [C.POINTER(rtb.SceneDesc), vp, C.c_size_t, vp]
        L.devsim_render.argtypes = [C.POINTER(rtb.SceneDesc), C.POINTER(rtb.RenderParams), vp, vp, vp]
        L.devsim_render_wavefront.argtypes = [C.POINTER(rtb.SceneDesc), C.POINTER(rtb.RenderParams), vp, vp, vp]
        L.devsim_render_wavefront_resume.argtypes = [C.POINTER(rtb.SceneDesc), C.POINTER(rtb.RenderParams), vp, vp, vp, vp]
        L.devsim_trace_path.argtypes = [C.POINTER(rtb.SceneDesc), vp, vp, C.c_uint, vp, C.POINTER(C.c_uint)]
        L.devsim_layout.argtypes = [C.POINTER(rtb.SceneDesc), vp, vp, vp, vp, vp, C.c_int, vp]
        L.devsim_assemble.argtypes = [C.POINTER(rtb.SceneDesc), C.POINTER(rtb.RawGeometry), vp, vp, vp, vp, vp, C.c_int, vp]
        _devsim = L
    return _devsim


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def devsim_closest_hit(hs, rays, cull):
    rays = np.ascontiguousarray(rays, np.float32)
    n = rays.shape[0]
    out = dict(id=np.empty(n, np.int32), t=np.empty(n, np.float32), u=np.empty(n, np.float32), v=np.empty(n, np.float32),
               inside=np.empty(n, np.uint8))
    cnt = np.zeros(2, np.uint64)
    rc = devsim().devsim_closest_hit(C.byref(hs.desc), _p(rays), n, cull, _p(out["id"]), _p(out["t"]), _p(out["u"]), _p(out["v"]),
                                     _p(out["inside"]), _p(cnt))
    assert rc == 0, devsim().devsim_last_error()
    out["box_tests"], out["tri_tests"] = int(cnt[0]), int(cnt[1])
    return out


def devsim_closest_hit_lane(hs, rays, schedule_seed):
    """the queue kernels' per-lane state machine (deferred leaves) on the host, phases drawn from `schedule_seed`"""
    rays = np.ascontiguousarray(rays, np.float32)
    n = rays.shape[0]
    out = dict(id=np.empty(n, np.int32), t=np.empty(n, np.float32), u=np.empty(n, np.float32), v=np.empty(n, np.float32))
    cnt = np.zeros(2, np.uint64)
    rc = devsim().devsim_closest_hit_lane(C.byref(hs.desc), _p(rays), n, schedule_seed, _p(out["id"]), _p(out["t"]), _p(out["u"]),
                                          _p(out["v"]), _p(cnt))
    assert rc == 0, devsim().devsim_last_error()
    out["box_tests"], out["tri_tests"] = int(cnt[0]), int(cnt[1])
    return out


def devsim_light_pdf(hs, rays):
    rays = np.ascontiguousarray(rays, np.float32)
    out = np.empty(rays.shape[0], np.float32)
    assert devsim().devsim_light_pdf(C.byref(hs.desc), _p(rays), rays.shape[0], _p(out)) == 0
    return out


def devsim_render(hs, width, height, samples, rows=None, pixel0_seed=1, reference_cull=False, wavefront=False, interleave=None,
                  substreams=0, substream_range=None):
    r0, r1 = rows if rows is not None else (0, height)
    p = rtb.RenderParams(width, height, samples, r0, r1, pixel0_seed, 0, 2 if reference_cull else 0)
    p.substreams = substreams
    if substream_range is not None:
        p.substream_begin, p.substream_end = substream_range
    wavefront = wavefront or substreams > 1
    n_rows = r1 - r0
    if interleave is not None:
        p.interleave_count, p.interleave_index, p.interleave_rows = interleave
        n_rows = rtb.lib().rtb_render_local_rows(C.byref(p))
    accum = np.empty((n_rows, width, 3), np.float32)
    rgb = np.empty((n_rows, width, 3), np.uint8)
    cnt = np.zeros(4, np.uint64)
    fn = devsim().devsim_render_wavefront if wavefront else devsim().devsim_render
    assert fn(C.byref(hs.desc), C.byref(p), _p(accum), _p(rgb), _p(cnt)) == 0, devsim().devsim_last_error()
    return dict(accum=accum, rgb=rgb, rays=int(cnt[0]), light_queries=int(cnt[1]), box_tests=int(cnt[2]), tri_tests=int(cnt[3]))


def devsim_render_resume(hs, width, height, samples, state=None, sample_begin=0, pixel0_seed=1, interleave=None):
    """the host build of the wavefront integrator with rtb_render_resume's per-pixel state (interleave: (count, index,
    rows) of the multi-GPU row-block split; buffers then hold the call's local rows)"""
    p = rtb.RenderParams(width, height, samples, 0, height, pixel0_seed, sample_begin, 0)
    n_rows = height
    if interleave is not None:
        p.interleave_count, p.interleave_index, p.interleave_rows = interleave
        n_rows = rtb.lib().rtb_render_local_rows(C.byref(p))
    accum = np.empty((n_rows, width, 3), np.float32)
    rgb = np.empty((n_rows, width, 3), np.uint8)
    state = np.zeros((n_rows, width, 8), np.uint32) if state is None else np.ascontiguousarray(state, np.uint32).copy()
    assert devsim().devsim_render_wavefront_resume(C.byref(hs.desc), C.byref(p), _p(state), _p(accum), _p(rgb), None) == 0, devsim().devsim_last_error()
    return dict(accum=accum, rgb=rgb, state=state)


def devsim_trace_path(hs, origin, direction, seed):
    o = np.asarray(origin, np.float32)
    d = np.asarray(direction, np.float32)
    rgb = np.zeros(3, np.float32)
    after = C.c_uint()
    assert devsim().devsim_trace_path(C.byref(hs.desc), _p(o), _p(d), seed, _p(rgb), C.byref(after)) == 0
    return rgb, after.value


def _layout_buffers(n_tris, n_nodes):
    return dict(tris=np.zeros((n_tris, 16), np.float32), shade=np.zeros((n_tris, 16), np.float32), tan=np.zeros((n_tris, 12), np.float32),
                nodes=np.zeros((max(0, (n_nodes - 1) // 2), 64), np.uint8), big=np.zeros((max(4096, n_tris), 2), np.int32), scalars=np.zeros(5, np.int32))


def _trim(b):
    b["nodes"] = b["nodes"][:b["scalars"][1]]
    b["big"] = b["big"][:b["scalars"][3]]
    if not b["scalars"][4]:
        b["tan"] = b["tan"][:0]
    return b


def devsim_layout(desc):
    """the device buffers rtb_scene_create builds on the host from the flat, host-assembled description"""
    b = _layout_buffers(desc.n_tris, desc.n_nodes)
    rc = devsim().devsim_layout(C.byref(desc), _p(b["tris"]), _p(b["shade"]), _p(b["tan"]), _p(b["nodes"]), _p(b["big"]), len(b["big"]), _p(b["scalars"]))
    assert rc == 0, devsim().devsim_last_error()
    return _trim(b)


def devsim_assemble(desc, raw):
    """the same buffers from the raw views: rtb_assemble.cuh's per-triangle / per-node functions run on the host.
    Returns (buffers, None) or (None, error text)"""
    b = _layout_buffers(desc.n_tris, desc.n_nodes)
    rc = devsim().devsim_assemble(C.byref(desc), C.byref(raw), _p(b["tris"]), _p(b["shade"]), _p(b["tan"]), _p(b["nodes"]), _p(b["big"]), len(b["big"]),
                                  _p(b["scalars"]))
    if rc != 0:
        return None, devsim().devsim_last_error().decode()
    return _trim(b), None


def twin_bvh(hs, lights=False):
    from oracle import twin
    if lights:
        a, l = hs.light_nodes()
        return twin.Bvh(a, l, hs.light_pos())
    a, l = hs.nodes()
    return twin.Bvh(a, l, hs.tris()["pos"])


def rmse8(a, b):
    d = a.astype(np.float64) - b.astype(np.float64)
    return float(np.sqrt(np.mean(d * d)))


def luminance(rgb_lin):
    return 0.2126 * rgb_lin[..., 0] + 0.7152 * rgb_lin[..., 1] + 0.0722 * rgb_lin[..., 2]
