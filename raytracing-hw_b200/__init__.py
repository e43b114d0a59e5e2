"""ctypes binding of librtb.so (include/rtb.h) -- host plumbing for the tests and bench.py.

The product is the C ABI + the sm_100a kernels behind it; this module only loads the shared library, builds the
argument structs and hands over numpy / torch buffers.  It never computes anything itself and has NO fallback:
if librtb.so is missing or no CUDA device is usable, the calls raise.

Import (the directory name contains a hyphen):  rtb = importlib.import_module("raytracing-hw_b200")
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# RTB_LIB: another build of the same library (A/B runs of kernel variants, tools/gpu_run.sh variants)
LIB_PATH = os.environ.get("RTB_LIB") or os.path.join(_HERE, "librtb.so")

RTB_OK, RTB_ERR_INVALID, RTB_ERR_NO_DEVICE, RTB_ERR_CUDA, RTB_ERR_IO = 0, 1, 2, 3, 4
CULL_REFERENCE, CULL_GLOBAL = 0, 1
RENDER_COUNTERS, RENDER_REFERENCE_CULL, RENDER_MEGAKERNEL, RENDER_STAGE_TIMES = 1, 2, 4, 8
REQUEST_COUNTERS = 0xFFFFFFFFFFFFFFFF
CLOSEST_HIT_STATIC_MAX_RAYS = 0xFFFFFEFF

# every symbol include/rtb.h declares (tests/test_abi.py checks the library exports all of them)
ABI_SYMBOLS = [
    "rtb_last_error", "rtb_device_count",
    "rtb_host_scene_load_gltf", "rtb_host_scene_from_triangles", "rtb_host_scene_desc", "rtb_host_scene_bvh_stats",
    "rtb_host_scene_build_seconds", "rtb_host_scene_build_info", "rtb_build_bvh", "rtb_host_scene_destroy",
    "rtb_host_scene_raw_geometry", "rtb_scene_create_raw", "rtb_scene_read_buffer",
    "rtb_scene_create", "rtb_scene_destroy", "rtb_scene_device", "rtb_scene_traversal_bytes", "rtb_scene_upload_bytes", "rtb_scene_launches", "rtb_scene_stage_ms",
    "rtb_closest_hit", "rtb_closest_hit_device", "rtb_closest_hit_static", "rtb_light_pdf", "rtb_measure_l2",
    "rtb_render", "rtb_render_device", "rtb_resolve_device", "rtb_render_local_rows", "rtb_render_global_row",
    "rtb_render_resume", "rtb_render_resume_device", "rtb_render_state_bytes", "rtb_scene_set_progress",
    "rtb_release_cached_memory",
    "rtb_comm_unique_id", "rtb_comm_create", "rtb_comm_destroy", "rtb_comm_world", "rtb_comm_rank", "rtb_dist_rank_params",
    "rtb_render_dist_device", "rtb_render_dist", "rtb_render_multi",
    "rtb_write_ppm",
]
SPLIT_TILE, SPLIT_SAMPLE = 0, 1
UNIQUE_ID_BYTES = 128


class TextureDesc(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("rgba", C.c_void_p)]


class SceneDesc(C.Structure):
    _fields_ = [
        ("n_nodes", C.c_uint32), ("node_aabb", C.c_void_p), ("node_links", C.c_void_p),
        ("n_tris", C.c_uint32), ("tri_pos", C.c_void_p), ("tri_nrm", C.c_void_p), ("tri_tan", C.c_void_p),
        ("tri_uv", C.c_void_p), ("tri_mesh", C.c_void_p),
        ("n_meshes", C.c_uint32), ("mesh_material", C.c_void_p), ("mesh_normal_transform", C.c_void_p),
        ("n_textures", C.c_uint32), ("textures", C.POINTER(TextureDesc)),
        ("n_lights", C.c_uint32), ("light_pos", C.c_void_p),
        ("n_light_nodes", C.c_uint32), ("light_node_aabb", C.c_void_p), ("light_node_links", C.c_void_p),
        ("cam_pos", C.c_float * 3), ("cam_axes", C.c_float * 9), ("cam_tan_half_fov", C.c_float * 2),
        ("ray_depth", C.c_int32), ("max_distance", C.c_float), ("bg_color", C.c_float * 3),
    ]


class RawMesh(C.Structure):   # rtb_raw_mesh
    _fields_ = [("n_tris", C.c_uint32), ("n_vertices", C.c_uint32), ("indices", C.c_void_p), ("index_bytes", C.c_int32),
                ("index_stride", C.c_int32), ("position", C.c_void_p), ("normal", C.c_void_p), ("tangent", C.c_void_p),
                ("texcoord", C.c_void_p), ("position_stride", C.c_int32), ("normal_stride", C.c_int32),
                ("tangent_stride", C.c_int32), ("texcoord_stride", C.c_int32), ("normal_mode", C.c_int32),
                ("mesh_id", C.c_int32), ("flags", C.c_int32), ("transform", C.c_double * 16)]


class RawGeometry(C.Structure):   # rtb_raw_geometry
    _fields_ = [("n_meshes", C.c_uint32), ("meshes", C.POINTER(RawMesh)), ("tri_source", C.c_void_p)]


NORMALS_TRANSFORM, NORMALS_VERBATIM, NORMALS_FLAT = 0, 1, 2
RAW_WORLD_POSITIONS = 1
BUF_NODES, BUF_TRIS, BUF_SHADE, BUF_TANGENTS, BUF_BIG_LEAVES, BUF_BVH_SCALARS = range(6)


class RenderParams(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("samples", C.c_int32), ("row_begin", C.c_int32),
                ("row_end", C.c_int32), ("pixel0_seed", C.c_uint32), ("sample_begin", C.c_int32), ("flags", C.c_int32),
                ("interleave_count", C.c_int32), ("interleave_index", C.c_int32), ("interleave_rows", C.c_int32),
                ("substreams", C.c_int32), ("substream_begin", C.c_int32), ("substream_end", C.c_int32)]


class Stats(C.Structure):
    _fields_ = [("closest_hit_rays", C.c_uint64), ("light_pdf_queries", C.c_uint64), ("box_tests", C.c_uint64),
                ("tri_tests", C.c_uint64), ("kernel_ms", C.c_double), ("h2d_ms", C.c_double), ("d2h_ms", C.c_double),
                ("launches", C.c_uint64), ("light_box_tests", C.c_uint64), ("light_tri_tests", C.c_uint64)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


PROGRESS_FN = C.CFUNCTYPE(C.c_int, C.c_int, C.c_void_p)   # rtb_progress_fn
PIXEL_STATE_BYTES = 32


class RtbError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("librtb error %d: %s" % (code, msg))
        self.code = code


_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError("%s is missing: run `python -c 'import __graft_entry__ as g; g.build()'` or ./build.sh "
                               "(there is no fallback path)" % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        L.rtb_last_error.restype = C.c_char_p
        L.rtb_host_scene_build_seconds.restype = C.c_double
        L.rtb_host_scene_build_seconds.argtypes = [C.c_void_p]
        L.rtb_host_scene_build_info.argtypes = [C.c_void_p, C.POINTER(C.c_double * 8)]
        L.rtb_build_bvh.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.POINTER(C.c_size_t), C.POINTER(C.c_double * 8)]
        L.rtb_scene_traversal_bytes.restype = C.c_size_t
        L.rtb_scene_traversal_bytes.argtypes = [C.c_void_p]
        L.rtb_scene_upload_bytes.restype = C.c_size_t
        L.rtb_scene_upload_bytes.argtypes = [C.c_void_p]
        L.rtb_scene_launches.restype = C.c_ulonglong
        L.rtb_scene_launches.argtypes = [C.c_void_p]
        L.rtb_scene_stage_ms.argtypes = [C.c_void_p, C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(C.c_ulonglong)]
        L.rtb_host_scene_load_gltf.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_void_p)]
        L.rtb_host_scene_from_triangles.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, C.c_void_p, C.c_void_p,
                                                    C.c_float, C.c_float, C.c_float, C.POINTER(C.c_void_p)]
        L.rtb_host_scene_desc.argtypes = [C.c_void_p, C.POINTER(SceneDesc)]
        L.rtb_host_scene_bvh_stats.argtypes = [C.c_void_p, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64 * 5), C.POINTER(C.c_uint64)]
        L.rtb_host_scene_destroy.argtypes = [C.c_void_p]
        L.rtb_host_scene_destroy.restype = None
        L.rtb_scene_create.argtypes = [C.POINTER(SceneDesc), C.c_int, C.POINTER(C.c_void_p)]
        L.rtb_host_scene_raw_geometry.argtypes = [C.c_void_p, C.POINTER(SceneDesc), C.POINTER(RawGeometry)]
        L.rtb_scene_create_raw.argtypes = [C.POINTER(SceneDesc), C.POINTER(RawGeometry), C.c_int, C.POINTER(C.c_void_p)]
        L.rtb_scene_read_buffer.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.POINTER(C.c_size_t)]
        L.rtb_scene_destroy.argtypes = [C.c_void_p]
        L.rtb_scene_destroy.restype = None
        L.rtb_scene_device.argtypes = [C.c_void_p]
        L.rtb_closest_hit.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int] + [C.c_void_p] * 5 + [C.POINTER(Stats)]
        L.rtb_closest_hit_device.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int] + [C.c_void_p] * 7
        L.rtb_closest_hit_static.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int] + [C.c_void_p] * 7
        L.rtb_light_pdf.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
        L.rtb_measure_l2.argtypes = [C.c_int, C.c_size_t, C.POINTER(C.c_double), C.POINTER(C.c_double)]
        L.rtb_render.argtypes = [C.c_void_p, C.POINTER(RenderParams), C.c_void_p, C.c_void_p, C.POINTER(Stats)]
        L.rtb_render_device.argtypes = [C.c_void_p, C.POINTER(RenderParams), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.rtb_render_resume.argtypes = [C.c_void_p, C.POINTER(RenderParams), C.c_void_p, C.c_void_p, C.c_void_p, C.POINTER(Stats)]
        L.rtb_render_resume_device.argtypes = [C.c_void_p, C.POINTER(RenderParams), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.rtb_release_cached_memory.argtypes = []
        L.rtb_release_cached_memory.restype = None
        L.rtb_render_state_bytes.argtypes = [C.POINTER(RenderParams)]
        L.rtb_render_state_bytes.restype = C.c_size_t
        L.rtb_scene_set_progress.argtypes = [C.c_void_p, PROGRESS_FN, C.c_void_p]
        L.rtb_resolve_device.argtypes = [C.c_void_p, C.c_size_t, C.c_int, C.c_void_p, C.c_void_p]
        L.rtb_write_ppm.argtypes = [C.c_char_p, C.c_int, C.c_int, C.c_void_p]
        L.rtb_comm_unique_id.argtypes = [C.c_void_p]
        L.rtb_comm_create.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_void_p)]
        L.rtb_comm_destroy.argtypes = [C.c_void_p]
        L.rtb_comm_destroy.restype = None
        L.rtb_comm_world.argtypes = [C.c_void_p]
        L.rtb_comm_rank.argtypes = [C.c_void_p]
        L.rtb_dist_rank_params.argtypes = [C.POINTER(RenderParams), C.c_int, C.c_int, C.c_int, C.POINTER(RenderParams)]
        L.rtb_render_dist_device.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(RenderParams), C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.rtb_render_dist.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(RenderParams), C.c_int, C.c_void_p, C.POINTER(Stats)]
        L.rtb_render_multi.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.POINTER(RenderParams), C.c_int, C.c_void_p, C.POINTER(Stats)]
        _lib = L
    return _lib


def _check(rc):
    if rc != RTB_OK:
        raise RtbError(rc, lib().rtb_last_error().decode(errors="replace"))


def _ptr(a):
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        return a.ctypes.data_as(C.c_void_p)
    if hasattr(a, "data_ptr"):      # torch tensor (host-pinned or device)
        return C.c_void_p(a.data_ptr())
    return C.c_void_p(int(a))


def build_bvh(tri_pos9):
    """BVH::buildBVH alone (rtb_build_bvh): tri_pos9 = (n, 9) float32 rows p0, U, V.  Returns (order, node_aabb, node_links, info)."""
    pos = np.ascontiguousarray(tri_pos9, np.float32).reshape(-1, 9)
    n = pos.shape[0]
    cap = max(1, 2 * n - 1)
    order = np.empty(n, np.uint32)
    aabb = np.empty((cap, 6), np.float32)
    links = np.empty((cap, 5), np.int32)
    nn, o = C.c_size_t(0), (C.c_double * 8)()
    _check(lib().rtb_build_bvh(_ptr(pos), n, _ptr(order), _ptr(aabb), _ptr(links), cap, C.byref(nn), C.byref(o)))
    info = dict(on_gpu=bool(o[0]), levels=int(o[1]), rounds=int(o[2]), heap_ranges=int(o[3]), host_fallback_nodes=int(o[4]), gpu_seconds=o[5], kernel_seconds=o[6])
    return order, aabb[:nn.value].copy(), links[:nn.value].copy(), info


def device_count():
    return lib().rtb_device_count()


def _np_view(ptr, shape, dtype):
    n = int(np.prod(shape))
    if n == 0 or not ptr:
        return np.zeros(shape, dtype)
    buf = (C.c_char * (n * np.dtype(dtype).itemsize)).from_address(ptr)
    return np.frombuffer(buf, dtype=dtype).reshape(shape)


class HostScene:
    """Host-side scene (glTF load + BVH build on the C++ host): mirrors `Scene` after its constructor."""

    def __init__(self, handle):
        self.h = C.c_void_p(handle)
        self._desc = None
        self._raw = None

    @property
    def desc(self):
        """the flat description (rtb_host_scene_desc): normals / tangents / texcoords are assembled on the HOST at the first use"""
        if self._desc is None:
            d = SceneDesc()
            _check(lib().rtb_host_scene_desc(self.h, C.byref(d)))
            self._desc = d
        return self._desc

    def raw(self):
        """(SceneDesc without host-assembled attributes, RawGeometry): what rtb_scene_create_raw takes"""
        if self._raw is None:
            d, g = SceneDesc(), RawGeometry()
            _check(lib().rtb_host_scene_raw_geometry(self.h, C.byref(d), C.byref(g)))
            self._raw = (d, g)
        return self._raw

    def tri_source(self):
        d, g = self.raw()
        return _np_view(g.tri_source, (d.n_tris,), np.uint32)

    @classmethod
    def load_gltf(cls, path, width, height, samples):
        h = C.c_void_p()
        _check(lib().rtb_host_scene_load_gltf(str(path).encode(), width, height, samples, C.byref(h)))
        return cls(h.value)

    @classmethod
    def from_triangles(cls, verts, normals=None, emissive_from=None, cam_pos=(0, 0, 30),
                       cam_axes=((1, 0, 0), (0, 1, 0), (0, 0, -1)), fov=(0.7, 0.7), max_distance=1e9):
        verts = np.ascontiguousarray(verts, np.float32).reshape(-1, 9)
        n = verts.shape[0]
        nrm = None if normals is None else np.ascontiguousarray(normals, np.float32).reshape(-1, 9)
        pos = np.ascontiguousarray(cam_pos, np.float32)
        axes = np.ascontiguousarray(cam_axes, np.float32).reshape(9)
        h = C.c_void_p()
        _check(lib().rtb_host_scene_from_triangles(_ptr(verts), _ptr(nrm), n, n if emissive_from is None else emissive_from,
                                                   _ptr(pos), _ptr(axes), fov[0], fov[1], max_distance, C.byref(h)))
        return cls(h.value)

    def close(self):
        if self.h:
            lib().rtb_host_scene_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def build_seconds(self):
        return lib().rtb_host_scene_build_seconds(self.h)

    def build_info(self):
        """where the BVH was built: dict(on_gpu, levels, rounds, heap_ranges, host_fallback_nodes, gpu_seconds, kernel_seconds)"""
        o = (C.c_double * 8)()
        _check(lib().rtb_host_scene_build_info(self.h, C.byref(o)))
        return dict(on_gpu=bool(o[0]), levels=int(o[1]), rounds=int(o[2]), heap_ranges=int(o[3]), host_fallback_nodes=int(o[4]),
                    gpu_seconds=o[5], kernel_seconds=o[6])

    def bvh_stats(self):
        n, c, m = C.c_uint64(), (C.c_uint64 * 5)(), C.c_uint64()
        _check(lib().rtb_host_scene_bvh_stats(self.h, C.byref(n), C.byref(c), C.byref(m)))
        return n.value, list(c), m.value

    # numpy views into the host scene's arrays (valid while the HostScene lives)
    def nodes(self):
        d = self.desc
        return _np_view(d.node_aabb, (d.n_nodes, 6), np.float32), _np_view(d.node_links, (d.n_nodes, 5), np.int32)

    def light_nodes(self):
        d = self.desc
        return _np_view(d.light_node_aabb, (d.n_light_nodes, 6), np.float32), _np_view(d.light_node_links, (d.n_light_nodes, 5), np.int32)

    def tris(self):
        d = self.desc
        n = d.n_tris
        return dict(pos=_np_view(d.tri_pos, (n, 9), np.float32), nrm=_np_view(d.tri_nrm, (n, 9), np.float32),
                    tan=_np_view(d.tri_tan, (n, 12), np.float32), uv=_np_view(d.tri_uv, (n, 6), np.float32),
                    mesh=_np_view(d.tri_mesh, (n,), np.int32))

    def light_pos(self):
        return _np_view(self.desc.light_pos, (self.desc.n_lights, 9), np.float32)

    def meshes(self):
        d = self.desc
        return _np_view(d.mesh_material, (d.n_meshes, 16), np.float32), _np_view(d.mesh_normal_transform, (d.n_meshes, 16), np.float64)

    def textures(self):
        out = []
        for i in range(self.desc.n_textures):
            t = self.desc.textures[i]
            out.append(_np_view(t.rgba, (t.height, t.width, 4), np.uint8))
        return out

    def camera(self):
        d = self.desc
        return dict(pos=np.array(d.cam_pos[:], np.float32), axes=np.array(d.cam_axes[:], np.float32).reshape(3, 3),
                    tan_half_fov=np.array(d.cam_tan_half_fov[:], np.float32), max_distance=d.max_distance, ray_depth=d.ray_depth)


class Scene:
    """Device-resident scene on one GPU; the drop-in for BVH::intersect and Scene::render."""

    def __init__(self, host_scene_or_desc, device=0, assembled=False, raw=None):
        """A HostScene goes through rtb_scene_create_raw (the triangles are assembled on the GPU from the raw views) unless
        assembled=True; a bare SceneDesc through rtb_scene_create (host-assembled arrays), or with raw= a RawGeometry."""
        self._keep = (host_scene_or_desc, raw)
        h = C.c_void_p()
        if isinstance(host_scene_or_desc, HostScene) and not assembled:
            d, g = host_scene_or_desc.raw()
            _check(lib().rtb_scene_create_raw(C.byref(d), C.byref(g), device, C.byref(h)))
        elif raw is not None:
            _check(lib().rtb_scene_create_raw(C.byref(host_scene_or_desc), C.byref(raw), device, C.byref(h)))
        else:
            desc = host_scene_or_desc.desc if isinstance(host_scene_or_desc, HostScene) else host_scene_or_desc
            _check(lib().rtb_scene_create(C.byref(desc), device, C.byref(h)))
        self.h = h
        self.device = device

    def read_buffer(self, which):
        """device buffer `which` (BUF_*) as bytes (test / inspection aid)"""
        n = C.c_size_t(0)
        _check(lib().rtb_scene_read_buffer(self.h, which, None, C.byref(n)))
        out = np.zeros(n.value, np.uint8)
        if n.value:
            _check(lib().rtb_scene_read_buffer(self.h, which, _ptr(out), C.byref(n)))
        return out

    def close(self):
        if self.h:
            lib().rtb_scene_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def traversal_bytes(self):
        return lib().rtb_scene_traversal_bytes(self.h)

    @property
    def upload_bytes(self):
        return lib().rtb_scene_upload_bytes(self.h)

    @property
    def launches(self):
        return lib().rtb_scene_launches(self.h)

    def stage_ms(self):
        """(queue-kernel ms, everything-else ms, queue launches) of the last render with RENDER_STAGE_TIMES"""
        a, b, n = C.c_double(), C.c_double(), C.c_ulonglong()
        _check(lib().rtb_scene_stage_ms(self.h, C.byref(a), C.byref(b), C.byref(n)))
        return a.value, b.value, n.value

    def closest_hit(self, rays, cull=CULL_GLOBAL, counters=False, out=None):
        """rays: (n,6) float32 host array (numpy or pinned torch).  Returns dict(id,t,u,v,inside,stats)."""
        if isinstance(rays, np.ndarray):
            rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 6)
        else:   # a (pinned) torch tensor: must already be packed float32
            assert str(rays.dtype) == "torch.float32" and rays.is_contiguous() and rays.shape[-1] == 6, "rays: contiguous float32 (n, 6)"
        n = rays.shape[0]
        if out is None:
            out = dict(id=np.empty(n, np.int32), t=np.empty(n, np.float32), u=np.empty(n, np.float32),
                       v=np.empty(n, np.float32), inside=np.empty(n, np.uint8))
        else:
            for key, dt in (("id", np.int32), ("t", np.float32), ("u", np.float32), ("v", np.float32), ("inside", np.uint8)):
                a = out.get(key)
                if isinstance(a, np.ndarray):
                    assert a.dtype == dt and a.flags["C_CONTIGUOUS"] and a.size >= n, "out[%r]: contiguous %s of >= n elements" % (key, np.dtype(dt).name)
        st = Stats()
        if counters:
            st.box_tests = REQUEST_COUNTERS
        _check(lib().rtb_closest_hit(self.h, _ptr(rays), n, cull, _ptr(out["id"]), _ptr(out.get("t")), _ptr(out.get("u")),
                                     _ptr(out.get("v")), _ptr(out.get("inside")), C.byref(st)))
        out = dict(out)
        out["stats"] = st.as_dict()
        return out

    def closest_hit_device(self, d_rays, n, cull, d_id, d_t=None, d_u=None, d_v=None, d_inside=None, d_counters=None, stream=0):
        _check(lib().rtb_closest_hit_device(self.h, _ptr(d_rays), n, cull, _ptr(d_id), _ptr(d_t), _ptr(d_u), _ptr(d_v),
                                            _ptr(d_inside), _ptr(d_counters), C.c_void_p(stream)))

    def closest_hit_static(self, d_rays, n, cull, d_id, d_t=None, d_u=None, d_v=None, d_inside=None, d_counters=None, stream=0):
        """rtb_closest_hit_static: closest_hit_device for callers that must not allocate -- the ray cursor comes from a
        table inside librtb.so; at most CLOSEST_HIT_STATIC_MAX_RAYS rays per call (include/rtb.h)."""
        _check(lib().rtb_closest_hit_static(self.h, _ptr(d_rays), n, cull, _ptr(d_id), _ptr(d_t), _ptr(d_u), _ptr(d_v),
                                            _ptr(d_inside), _ptr(d_counters), C.c_void_p(stream)))

    def light_pdf(self, rays):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 6)
        out = np.empty(rays.shape[0], np.float32)
        _check(lib().rtb_light_pdf(self.h, _ptr(rays), rays.shape[0], _ptr(out)))
        return out

    def render(self, width, height, samples, rows=None, pixel0_seed=1, counters=False, want_accum=True, want_rgb=True,
               accum=None, rgb=None, reference_cull=False, megakernel=False, params=None, substreams=0, substream_range=None):
        """Replaces Scene::render.  HOST output buffers (numpy or pinned torch); copies are inside the call.
        substreams = K > 1: the sample-range split (K engine streams per pixel); substream_range=(k0, k1) renders only
        those streams and returns their partial sums (no rgb)."""
        if params is None:
            r0, r1 = rows if rows is not None else (0, height)
            flags = (1 if counters else 0) | (2 if reference_cull else 0) | (4 if megakernel else 0)
            params = RenderParams(width, height, samples, r0, r1, pixel0_seed, 0, flags)
            params.substreams = substreams
            if substream_range is not None:
                params.substream_begin, params.substream_end = substream_range
                if tuple(substream_range) != (0, max(1, substreams)):
                    want_rgb = False
        n = lib().rtb_render_local_rows(C.byref(params))
        if n < 0:
            raise RtbError(RTB_ERR_INVALID, "bad render parameters")
        if accum is None and want_accum:
            accum = np.empty((n, params.width, 3), np.float32)
        if rgb is None and want_rgb:
            rgb = np.empty((n, params.width, 3), np.uint8)
        st = Stats()
        _check(lib().rtb_render(self.h, C.byref(params), _ptr(accum), _ptr(rgb), C.byref(st)))
        return dict(accum=accum, rgb=rgb, stats=st.as_dict())

    def render_resume(self, width, height, samples, state=None, sample_begin=0, rows=None, pixel0_seed=1, params=None):
        """Progressive rendering (rtb_render_resume): samples [sample_begin, samples) of every pixel.  `state` (uint32
        array [rows, width, 8]) is what an earlier call with samples == sample_begin returned; the returned state is
        the one after `samples` samples.  accum / rgb are those of ONE call with `samples` samples, bit for bit."""
        if params is None:
            r0, r1 = rows if rows is not None else (0, height)
            params = RenderParams(width, height, samples, r0, r1, pixel0_seed, sample_begin, 0)
        n = lib().rtb_render_local_rows(C.byref(params))
        if n < 0:
            raise RtbError(RTB_ERR_INVALID, "bad render parameters")
        if state is None:
            if params.sample_begin > 0:
                raise RtbError(RTB_ERR_INVALID, "sample_begin > 0 needs the pixel state of the earlier call")
            state = np.zeros((n, params.width, PIXEL_STATE_BYTES // 4), np.uint32)
        else:
            state = np.ascontiguousarray(state, np.uint32).copy()
            assert state.nbytes == lib().rtb_render_state_bytes(C.byref(params))
        accum = np.empty((n, params.width, 3), np.float32)
        rgb = np.empty((n, params.width, 3), np.uint8)
        st = Stats()
        _check(lib().rtb_render_resume(self.h, C.byref(params), _ptr(state), _ptr(accum), _ptr(rgb), C.byref(st)))
        return dict(accum=accum, rgb=rgb, state=state, stats=st.as_dict())

    def set_progress(self, fn):
        """fn(percent) is called on the rendering thread while a frame renders (rtb_scene_set_progress); None: off."""
        self._progress = PROGRESS_FN(lambda pct, _user: int(fn(pct) or 0)) if fn is not None else PROGRESS_FN()
        _check(lib().rtb_scene_set_progress(self.h, self._progress, None))

    def render_device(self, params, d_accum=None, d_rgb8=None, d_counters=None, stream=0):
        """DEVICE output pointers (e.g. torch tensors), asynchronous on `stream` (a cudaStream_t handle)."""
        _check(lib().rtb_render_device(self.h, C.byref(params), _ptr(d_accum), _ptr(d_rgb8), _ptr(d_counters), C.c_void_p(stream)))


def frame_params(width, height, samples, pixel0_seed=1, flags=0, substreams=0, interleave_rows=8):
    """rtb_render_params of a WHOLE frame, as rtb_render_dist / rtb_render_multi take it"""
    p = RenderParams(width, height, samples, 0, height, pixel0_seed, 0, flags, 0, 0, interleave_rows)
    p.substreams = substreams
    return p


def rank_params(frame, split, world, rank):
    """what rank `rank` of `world` renders for `frame` (rtb_dist_rank_params)"""
    out = RenderParams()
    _check(lib().rtb_dist_rank_params(C.byref(frame), split, world, rank, C.byref(out)))
    return out


def unique_id():
    """rank 0: the bytes every rank hands to Comm (rtb_comm_unique_id = ncclGetUniqueId)"""
    buf = (C.c_char * UNIQUE_ID_BYTES)()
    _check(lib().rtb_comm_unique_id(buf))
    return bytes(buf)


class Comm:
    """One rank's communicator (rtb_comm): NCCL over NVLink, one rank per GPU.  Collective constructor."""

    def __init__(self, uid, world, rank, device):
        h = C.c_void_p()
        self._uid = C.create_string_buffer(uid, UNIQUE_ID_BYTES) if uid is not None else None
        _check(lib().rtb_comm_create(self._uid, world, rank, device, C.byref(h)))
        self.h, self.world, self.rank = h, world, rank

    def close(self):
        if self.h:
            lib().rtb_comm_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def render_dist_device(scene, comm, frame, split, d_accum_root=None, d_rgb8_root=None, d_counters=None, stream=0):
    """every rank calls; rank 0's device buffers receive the whole frame (rtb_render_dist_device)"""
    _check(lib().rtb_render_dist_device(scene.h, comm.h if comm is not None else None, C.byref(frame), split, _ptr(d_accum_root),
                                        _ptr(d_rgb8_root), _ptr(d_counters), C.c_void_p(stream)))


def render_dist(scene, comm, frame, split, rgb_root=None):
    """same with a HOST image on rank 0; returns this rank's stats"""
    st = Stats()
    _check(lib().rtb_render_dist(scene.h, comm.h if comm is not None else None, C.byref(frame), split, _ptr(rgb_root), C.byref(st)))
    return st.as_dict()


def render_multi(scenes, frame, split):
    """one process, one thread per device (what the CLI runs): returns (H x W x 3 uint8 image, summed stats)"""
    arr = (C.c_void_p * len(scenes))(*[s.h for s in scenes])
    rgb = np.empty((frame.height, frame.width, 3), np.uint8)
    st = Stats()
    _check(lib().rtb_render_multi(arr, len(scenes), C.byref(frame), split, _ptr(rgb), C.byref(st)))
    return rgb, st.as_dict()


def release_cached_memory():
    """hand the device blocks kept from destroyed scenes back to the driver (rtb_release_cached_memory)"""
    lib().rtb_release_cached_memory()


def measure_l2(device=0, nbytes=32 << 20):
    """(streaming GB/s, random 64-byte gather GB/s) of the device's L2 on an nbytes buffer"""
    a, b = C.c_double(0.0), C.c_double(0.0)
    _check(lib().rtb_measure_l2(device, nbytes, C.byref(a), C.byref(b)))
    return a.value, b.value


def resolve_device(d_accum, n_pixels, samples, d_rgb8, stream=0):
    _check(lib().rtb_resolve_device(_ptr(d_accum), n_pixels, samples, _ptr(d_rgb8), C.c_void_p(stream)))


def write_ppm(path, rgb):
    rgb = np.ascontiguousarray(rgb, np.uint8)
    _check(lib().rtb_write_ppm(str(path).encode(), rgb.shape[1], rgb.shape[0], _ptr(rgb)))
