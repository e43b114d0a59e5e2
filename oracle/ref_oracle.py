"""TEST INFRASTRUCTURE ONLY: ctypes binding of oracle/_ref/libref_oracle.so (the UNMODIFIED reference compiled
from /root/reference by oracle/Makefile, plus oracle/ref_harness.cpp).  Only tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs may import this module."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_ref", "libref_oracle.so")
_lib = None


def available():
    return os.path.exists(LIB_PATH)


def lib():
    global _lib
    if _lib is None:
        if not available():
            raise RuntimeError("oracle/_ref/libref_oracle.so is missing: run `make -C oracle ref` where /root/reference exists")
        L = C.CDLL(LIB_PATH)
        vp, sz, i32, f32, f64 = C.c_void_p, C.c_size_t, C.c_int, C.c_float, C.c_double
        L.ref_last_error.restype = C.c_char_p
        L.ref_scene_from_gltf.restype = vp
        L.ref_scene_from_gltf.argtypes = [C.c_char_p, i32, i32, i32, i32]
        L.ref_scene_from_triangles.restype = vp
        L.ref_scene_from_triangles.argtypes = [vp, vp, sz, sz, i32, i32, i32, vp, vp, f32, f32, f32, i32]
        L.ref_scene_destroy.argtypes = [vp]
        for n in ("ref_num_nodes", "ref_num_tris", "ref_num_meshes", "ref_num_textures", "ref_num_lights", "ref_num_light_nodes"):
            getattr(L, n).restype = sz
            getattr(L, n).argtypes = [vp]
        L.ref_get_nodes.argtypes = [vp, vp, vp]
        L.ref_get_light_nodes.argtypes = [vp, vp, vp]
        L.ref_get_tris.argtypes = [vp, vp, vp, vp, vp, vp]
        L.ref_get_light_tris.argtypes = [vp, vp, vp, vp, vp, vp]
        L.ref_get_meshes.argtypes = [vp, vp, vp]
        L.ref_get_texture_info.argtypes = [vp, sz, vp, vp, vp]
        L.ref_get_texture_data.argtypes = [vp, sz, vp]
        L.ref_get_camera.argtypes = [vp] + [vp] * 7
        L.ref_closest_hit.restype = f64
        L.ref_closest_hit.argtypes = [vp, vp, sz, vp, vp, vp, vp, vp, i32]
        L.ref_light_pdf.argtypes = [vp, vp, sz, vp]
        L.ref_render.restype = f64
        L.ref_render.argtypes = [vp, vp, i32, i32]
        L.ref_render_deterministic.restype = f64
        L.ref_render_deterministic.argtypes = [vp, C.c_uint, i32, vp, vp]
        L.ref_render_deterministic_rows.restype = f64
        L.ref_render_deterministic_rows.argtypes = [vp, C.c_uint, i32, i32, i32, vp, vp]
        L.ref_trace_path.argtypes = [vp, vp, vp, C.c_uint, vp, vp]
        L.ref_minstd_first.restype = C.c_uint
        L.ref_minstd_first.argtypes = [C.c_uint]
        L.ref_uniform_first.restype = f32
        L.ref_uniform_first.argtypes = [C.c_uint, f32, f32]
        L.ref_normal_seq.argtypes = [C.c_uint, i32, vp]
        L.ref_libm.argtypes = [i32, vp, sz, vp, f32]
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class RefScene:
    """A reference `Scene` (src/core/scene.h:14-38): BVH built by the reference, queried by the reference."""

    def __init__(self, handle):
        if not handle:
            raise RuntimeError("reference scene creation failed: " + lib().ref_last_error().decode())
        self.h = handle

    @classmethod
    def from_gltf(cls, path, width, height, samples, quiet=True):
        return cls(lib().ref_scene_from_gltf(path.encode(), width, height, samples, int(quiet)))

    @classmethod
    def from_triangles(cls, verts, normals=None, emissive_from=None, width=64, height=64, samples=1,
                       cam_pos=(0, 0, 30), cam_axes=((1, 0, 0), (0, 1, 0), (0, 0, -1)), fov=(0.7, 0.7), max_distance=1e9, quiet=True):
        verts = np.ascontiguousarray(verts, np.float32).reshape(-1, 9)
        n = verts.shape[0]
        nrm = None if normals is None else np.ascontiguousarray(normals, np.float32).reshape(-1, 9)
        pos = np.asarray(cam_pos, np.float32)
        axes = np.asarray(cam_axes, np.float32).reshape(9)
        return cls(lib().ref_scene_from_triangles(_p(verts), _p(nrm), n, n if emissive_from is None else emissive_from,
                                                  width, height, samples, _p(pos), _p(axes), fov[0], fov[1], max_distance, int(quiet)))

    def close(self):
        if self.h:
            lib().ref_scene_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- layout
    def _nodes(self, count_fn, get_fn):
        n = count_fn(self.h)
        aabb = np.zeros((n, 6), np.float32)
        links = np.zeros((n, 5), np.int64)
        get_fn(self.h, _p(aabb), _p(links))
        return aabb, links

    def nodes(self):
        return self._nodes(lib().ref_num_nodes, lib().ref_get_nodes)

    def light_nodes(self):
        return self._nodes(lib().ref_num_light_nodes, lib().ref_get_light_nodes)

    def _tris(self, n, get_fn):
        pos = np.zeros((n, 9), np.float32)
        nrm = np.zeros((n, 9), np.float32)
        tan = np.zeros((n, 12), np.float32)
        uv = np.zeros((n, 6), np.float32)
        mesh = np.zeros(n, np.int32)
        get_fn(self.h, _p(pos), _p(nrm), _p(tan), _p(uv), _p(mesh))
        return dict(pos=pos, nrm=nrm, tan=tan, uv=uv, mesh=mesh)

    def tris(self):
        return self._tris(lib().ref_num_tris(self.h), lib().ref_get_tris)

    def light_tris(self):
        return self._tris(lib().ref_num_lights(self.h), lib().ref_get_light_tris)

    def meshes(self):
        n = lib().ref_num_meshes(self.h)
        mat = np.zeros((n, 16), np.float32)
        nt = np.zeros((n, 16), np.float64)
        lib().ref_get_meshes(self.h, _p(mat), _p(nt))
        return mat, nt

    def textures(self):
        out = []
        for i in range(lib().ref_num_textures(self.h)):
            w, h, c = C.c_int(), C.c_int(), C.c_int()
            lib().ref_get_texture_info(self.h, i, C.byref(w), C.byref(h), C.byref(c))
            data = np.zeros((h.value, w.value, c.value), np.uint8)
            lib().ref_get_texture_data(self.h, i, _p(data))
            out.append(data)
        return out

    def camera(self):
        pos, axes, fov = np.zeros(3, np.float32), np.zeros(9, np.float32), np.zeros(2, np.float32)
        size = np.zeros(2, np.int32)
        md, rd, sp = C.c_float(), C.c_int(), C.c_int()
        lib().ref_get_camera(self.h, _p(pos), _p(axes), _p(fov), _p(size), C.byref(md), C.byref(rd), C.byref(sp))
        return dict(pos=pos, axes=axes.reshape(3, 3), fov=fov, size=size, max_distance=md.value, ray_depth=rd.value, samples=sp.value)

    # ---- queries
    def closest_hit(self, rays, threads=0):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 6)
        n = rays.shape[0]
        idx = np.zeros(n, np.int64)
        t, u, v = np.zeros(n, np.float32), np.zeros(n, np.float32), np.zeros(n, np.float32)
        inside = np.zeros(n, np.uint8)
        sec = lib().ref_closest_hit(self.h, _p(rays), n, _p(idx), _p(t), _p(u), _p(v), _p(inside), threads)
        return dict(id=idx, t=t, u=u, v=v, inside=inside, seconds=sec)

    def light_pdf(self, rays):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 6)
        out = np.zeros(rays.shape[0], np.float32)
        lib().ref_light_pdf(self.h, _p(rays), rays.shape[0], _p(out))
        return out

    def render(self, threads=0, quiet=True):
        cam = self.camera()
        rgb = np.zeros((cam["size"][1], cam["size"][0], 3), np.uint8)
        sec = lib().ref_render(self.h, _p(rgb), threads, int(quiet))
        return rgb, sec

    def render_deterministic(self, pixel0_seed=1, nproc=0, rows=None):
        """rows=(r0, r1): only that band of the frame (the outputs then hold r1 - r0 rows)"""
        cam = self.camera()
        H, W = int(cam["size"][1]), int(cam["size"][0])
        r0, r1 = rows if rows is not None else (0, H)
        rgb = np.zeros((r1 - r0, W, 3), np.uint8)
        acc = np.zeros((r1 - r0, W, 3), np.float32)
        sec = lib().ref_render_deterministic_rows(self.h, pixel0_seed, nproc, r0, r1, _p(acc), _p(rgb))
        if sec < 0:
            raise RuntimeError("ref_render_deterministic failed")
        return rgb, acc, sec

    def trace_path(self, origin, direction, engine_seed):
        o = np.asarray(origin, np.float32)
        d = np.asarray(direction, np.float32)
        rgb = np.zeros(3, np.float32)
        after = C.c_uint()
        lib().ref_trace_path(self.h, _p(o), _p(d), engine_seed, _p(rgb), C.byref(after))
        return rgb, after.value


def libm(which, x, arg=0.0):
    """which: 0 logf, 1 powf(x,arg), 2 (float)pow(1-x,5) [double pow], 3 tanf  -- evaluated by the host's glibc."""
    x = np.ascontiguousarray(x, np.float32)
    out = np.zeros_like(x)
    lib().ref_libm(which, _p(x), x.size, _p(out), arg)
    return out


def normal_seq(seed, n):
    out = np.zeros(n, np.float32)
    lib().ref_normal_seq(seed, n, _p(out))
    return out


def write_ppm(path, rgb):
    h, w, _ = rgb.shape
    with open(path, "wb") as f:
        f.write(b"P6\n%d %d\n255\n" % (w, h))
        f.write(np.ascontiguousarray(rgb, np.uint8).tobytes())
