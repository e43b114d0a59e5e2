"""TEST INFRASTRUCTURE ONLY: ctypes binding of oracle/libtwin.so (oracle/twin.c, the plain-C restatement of the
reference's traversal hot path).  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs may import this module; the product (librtb.so) never does."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libtwin.so")
_lib = None


class TwinBvh(C.Structure):
    _fields_ = [("n_nodes", C.c_uint32), ("aabb", C.c_void_p), ("links", C.c_void_p), ("n_tris", C.c_uint32),
                ("tri_pos", C.c_void_p)]


def build():
    subprocess.run(["make", "-C", _HERE, "twin"], check=True, stdout=subprocess.DEVNULL)


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        L = C.CDLL(LIB_PATH)
        vp = C.c_void_p
        L.twin_closest_hit.restype = C.c_double
        L.twin_closest_hit.argtypes = [C.POINTER(TwinBvh), vp, C.c_size_t, vp, vp, vp, vp, vp, vp, C.c_int]
        L.twin_light_pdf.argtypes = [C.POINTER(TwinBvh), vp, C.c_size_t, vp]
        L.twin_minstd_first.restype = C.c_uint32
        L.twin_minstd_first.argtypes = [C.c_uint32]
        L.twin_uniform_first.restype = C.c_float
        L.twin_uniform_first.argtypes = [C.c_uint32, C.c_float, C.c_float]
        L.twin_normal_seq.argtypes = [C.c_uint32, C.c_int, vp]
        L.twin_resolve.argtypes = [vp, C.c_size_t, C.c_int, vp]
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


class Bvh:
    """Flat reference-order BVH (aabb (n,6) f32, links (n,5) i32, tri_pos (m,9) f32) held for the twin."""

    def __init__(self, aabb, links, tri_pos):
        self.aabb = np.ascontiguousarray(aabb, np.float32)
        self.links = np.ascontiguousarray(links, np.int32)
        self.pos = np.ascontiguousarray(tri_pos, np.float32)
        self.c = TwinBvh(self.aabb.shape[0], _p(self.aabb), _p(self.links), self.pos.shape[0], _p(self.pos))

    def closest_hit(self, rays, counters=False, threads=0):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 6)
        n = rays.shape[0]
        out = dict(id=np.empty(n, np.int32), t=np.empty(n, np.float32), u=np.empty(n, np.float32),
                   v=np.empty(n, np.float32), inside=np.empty(n, np.uint8))
        cnt = np.zeros(2, np.uint64)
        sec = lib().twin_closest_hit(C.byref(self.c), _p(rays), n, _p(out["id"]), _p(out["t"]), _p(out["u"]), _p(out["v"]),
                                     _p(out["inside"]), _p(cnt) if counters else None, threads)
        out["seconds"] = sec
        if counters:
            out["box_tests"], out["tri_tests"] = int(cnt[0]), int(cnt[1])
        return out

    def light_pdf(self, rays):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 6)
        out = np.empty(rays.shape[0], np.float32)
        lib().twin_light_pdf(C.byref(self.c), _p(rays), rays.shape[0], _p(out))
        return out


def minstd_first(seed):
    return lib().twin_minstd_first(seed)


def uniform_first(seed, a, b):
    return lib().twin_uniform_first(seed, a, b)


def normal_seq(seed, n):
    out = np.empty(n, np.float32)
    lib().twin_normal_seq(seed, n, _p(out))
    return out


def resolve(accum, samples):
    accum = np.ascontiguousarray(accum, np.float32)
    out = np.empty(accum.shape, np.uint8)
    lib().twin_resolve(_p(accum), accum.size, samples, _p(out))
    return out
