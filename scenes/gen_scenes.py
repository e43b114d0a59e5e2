#!/usr/bin/env python3
"""Deterministic generators for the stand-in scenes of BASELINE.json's five configs.

Every glTF example of the reference points at a geometry buffer that was never shipped
(SURVEY.md section 0.3, Appendix A), so each real-scene config runs on a STAND-IN whose camera,
node transforms and materials are the values of the reference's example JSON
(examples/practice7_1.gltf, practice7_4.gltf, practice6_1.gltf, sponza/sponza.gltf) and whose
geometry/texture buffers are regenerated procedurally with the same primitive partition and counts.
Nothing here reads /root/reference; the numbers below were transcribed from those JSON files.

Usage:  python scenes/gen_scenes.py <name> [outdir]      name in SCENES (see bottom) or "all-small"
"""
import base64
import json
import os
import struct
import sys
import zlib

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
GEN_DIR = os.path.join(HERE, "_gen")

# ----------------------------------------------------------------------------- counter-based uniforms
_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def splitmix_u01(seed, start, count):
    """u[i] = top 24 bits of splitmix64-finalised (seed + start + i), times 2^-24, as float32.
    Counter based: any sub-range can be produced independently."""
    with np.errstate(over="ignore"):
        x = np.arange(start, start + count, dtype=np.uint64) + np.uint64(seed)
        x = x + np.uint64(0x9E3779B97F4A7C15)
        z = x
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        z = z ^ (z >> np.uint64(31))
    return ((z >> np.uint64(40)).astype(np.float32) * np.float32(2.0 ** -24)).astype(np.float32)


def soup_triangles(n, seed=1234, half_extent=10.0, jitter=0.15, start=0):
    """n x 3 x 3 float32 world-space vertices: centre uniform in [-E,E]^3, each vertex = centre + U[-j,j]^3."""
    u = splitmix_u01(seed, start * 12, n * 12).reshape(n, 12)
    E = np.float32(half_extent)
    c = (u[:, 0:3] * (E * np.float32(2)) - E).astype(np.float32)
    J = np.float32(jitter)
    off = (u[:, 3:12] * (J * np.float32(2)) - J).astype(np.float32).reshape(n, 3, 3)
    return (c[:, None, :] + off).astype(np.float32)


def soup_rays(n, seed=42, half_extent=10.0, start=0):
    """n x 6 float32: origin uniform in [-E,E]^3, direction uniform in [-1,1]^3 (NOT normalised: the
    kernels / the reference normalise exactly as Ray::Ray does, src/geometry/primitive.cpp:12-15)."""
    u = splitmix_u01(seed, start * 6, n * 6).reshape(n, 6)
    E = np.float32(half_extent)
    r = np.empty((n, 6), np.float32)
    r[:, 0:3] = u[:, 0:3] * (E * np.float32(2)) - E
    r[:, 3:6] = u[:, 3:6] * np.float32(2) - np.float32(1)
    return r


# ----------------------------------------------------------------------------- tiny PNG writer
def write_png(path, img):
    """img: HxWx3 or HxWx4 uint8.  Sub filter on every row (compresses smooth procedural content well)."""
    h, w, c = img.shape
    ct = {3: 2, 4: 6}[c]
    raw = img.astype(np.int16)
    sub = raw.copy()
    sub[:, 1:, :] -= raw[:, :-1, :]
    sub = (sub & 0xFF).astype(np.uint8).reshape(h, w * c)
    rows = np.concatenate([np.full((h, 1), 1, np.uint8), sub], axis=1)

    def chunk(tag, data):
        body = tag + data
        return struct.pack(">I", len(data)) + body + struct.pack(">I", zlib.crc32(body) & 0xFFFFFFFF)

    with open(path, "wb") as f:
        f.write(b"\x89PNG\r\n\x1a\n")
        f.write(chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, ct, 0, 0, 0)))
        f.write(chunk(b"IDAT", zlib.compress(rows.tobytes(), 6)))
        f.write(chunk(b"IEND", b""))


# ----------------------------------------------------------------------------- glTF writer
class GltfBuilder:
    def __init__(self):
        self.bin = bytearray()
        self.views, self.accessors = [], []
        self.meshes, self.nodes, self.materials, self.cameras = [], [], [], []
        self.images, self.textures = [], []
        self.ext_used = set()

    def _view(self, data, target):
        while len(self.bin) % 4:
            self.bin.append(0)
        off = len(self.bin)
        self.bin += data
        self.views.append({"buffer": 0, "byteLength": len(data), "byteOffset": off, "target": target})
        return len(self.views) - 1

    def accessor(self, arr, kind):
        arr = np.ascontiguousarray(arr)
        if kind == "SCALAR":
            ct = 5123 if arr.dtype == np.uint16 else 5125
            v = self._view(arr.tobytes(), 34963)
            acc = {"bufferView": v, "componentType": ct, "count": int(arr.size), "type": "SCALAR"}
        else:
            arr = arr.astype(np.float32)
            v = self._view(arr.tobytes(), 34962)
            acc = {"bufferView": v, "componentType": 5126, "count": int(arr.shape[0]), "type": kind}
            if kind == "VEC3":
                acc["max"] = [float(x) for x in arr.max(axis=0)]
                acc["min"] = [float(x) for x in arr.min(axis=0)]
        self.accessors.append(acc)
        return len(self.accessors) - 1

    def add_primitive(self, pos, nrm, idx, material, uv=None, tan=None):
        idx = np.asarray(idx)
        idx = idx.astype(np.uint16) if pos.shape[0] <= 65535 else idx.astype(np.uint32)
        attrs = {"POSITION": self.accessor(pos, "VEC3"), "NORMAL": self.accessor(nrm, "VEC3")}
        if tan is not None:
            attrs["TANGENT"] = self.accessor(tan, "VEC4")
        if uv is not None:
            attrs["TEXCOORD_0"] = self.accessor(uv, "VEC2")
        return {"attributes": attrs, "indices": self.accessor(idx, "SCALAR"), "material": material}

    def add_mesh(self, name, prims):
        self.meshes.append({"name": name, "primitives": prims})
        return len(self.meshes) - 1

    def write(self, path, embed=False):
        g = {"asset": {"generator": "rt-b200 stand-in generator", "version": "2.0"}, "scene": 0,
             "scenes": [{"name": "Scene", "nodes": list(range(len(self.nodes)))}],
             "nodes": self.nodes, "cameras": self.cameras, "materials": self.materials, "meshes": self.meshes,
             "accessors": self.accessors, "bufferViews": self.views}
        if self.ext_used:
            g["extensionsUsed"] = sorted(self.ext_used)
        if self.images:
            g["images"] = self.images
            g["textures"] = self.textures
            g["samplers"] = [{"magFilter": 9729, "minFilter": 9987}]
        binname = os.path.splitext(os.path.basename(path))[0] + ".bin"
        if embed:
            uri = "data:application/octet-stream;base64," + base64.b64encode(bytes(self.bin)).decode()
        else:
            uri = binname
            with open(os.path.join(os.path.dirname(path), binname), "wb") as f:
                f.write(self.bin)
        g["buffers"] = [{"byteLength": len(self.bin), "uri": uri}]
        with open(path, "w") as f:
            json.dump(g, f)
        return path


def material(name, base=None, metallic=None, roughness=None, emissive_strength=None, extra_ext=None):
    pbr = {}
    if base is not None:
        pbr["baseColorFactor"] = base
    if metallic is not None:
        pbr["metallicFactor"] = metallic
    if roughness is not None:
        pbr["roughnessFactor"] = roughness
    m = {"doubleSided": True, "name": name, "pbrMetallicRoughness": pbr}
    if emissive_strength is not None:
        m["emissiveFactor"] = [1, 1, 1]
        m["extensions"] = {"KHR_materials_emissive_strength": {"emissiveStrength": emissive_strength}}
    if extra_ext:
        m.setdefault("extensions", {}).update(extra_ext)
    return m


# ----------------------------------------------------------------------------- mesh helpers
def plane_mesh():
    """Blender's default plane as exported: verts (+-1,0,+-1), indices 0,1,3,0,3,2, normal +Y."""
    pos = np.array([[-1, 0, 1], [1, 0, 1], [-1, 0, -1], [1, 0, -1]], np.float32)
    nrm = np.tile(np.array([[0, 1, 0]], np.float32), (4, 1))
    uv = np.array([[0, 0], [1, 0], [0, 1], [1, 1]], np.float32)
    tan = np.tile(np.array([[1, 0, 0, 1]], np.float32), (4, 1))
    idx = np.array([0, 1, 3, 0, 3, 2])
    return pos, nrm, uv, tan, idx


def cube_mesh(lo=(-1, -1, -1), hi=(1, 1, 1)):
    """24-vertex flat-shaded box, 12 triangles."""
    lo, hi = np.array(lo, np.float32), np.array(hi, np.float32)
    P, N, U, I = [], [], [], []
    for axis in range(3):
        for sgn in (-1, 1):
            a, b = (axis + 1) % 3, (axis + 2) % 3
            n = np.zeros(3, np.float32)
            n[axis] = sgn
            base = len(P)
            for (sa, sb) in ((0, 0), (1, 0), (1, 1), (0, 1)):
                p = np.zeros(3, np.float32)
                p[axis] = hi[axis] if sgn > 0 else lo[axis]
                p[a] = hi[a] if sa else lo[a]
                p[b] = hi[b] if sb else lo[b]
                P.append(p)
                N.append(n)
                U.append([sa, sb])
            quad = [0, 1, 2, 0, 2, 3] if sgn > 0 else [0, 2, 1, 0, 3, 2]
            I += [base + q for q in quad]
    return np.array(P, np.float32), np.array(N, np.float32), np.array(U, np.float32), np.array(I)


def icosphere(subdiv, radius):
    t = (1.0 + 5.0 ** 0.5) / 2.0
    v = [(-1, t, 0), (1, t, 0), (-1, -t, 0), (1, -t, 0), (0, -1, t), (0, 1, t), (0, -1, -t), (0, 1, -t),
         (t, 0, -1), (t, 0, 1), (-t, 0, -1), (-t, 0, 1)]
    f = [(0, 11, 5), (0, 5, 1), (0, 1, 7), (0, 7, 10), (0, 10, 11), (1, 5, 9), (5, 11, 4), (11, 10, 2), (10, 7, 6),
         (7, 1, 8), (3, 9, 4), (3, 4, 2), (3, 2, 6), (3, 6, 8), (3, 8, 9), (4, 9, 5), (2, 4, 11), (6, 2, 10),
         (8, 6, 7), (9, 8, 1)]
    v = [np.array(x, np.float64) / np.linalg.norm(x) for x in v]
    for _ in range(subdiv - 1):
        cache, nf = {}, []

        def mid(a, b):
            k = (min(a, b), max(a, b))
            if k not in cache:
                m = v[a] + v[b]
                v.append(m / np.linalg.norm(m))
                cache[k] = len(v) - 1
            return cache[k]

        for (a, b, c) in f:
            ab, bc, ca = mid(a, b), mid(b, c), mid(c, a)
            nf += [(a, ab, ca), (b, bc, ab), (c, ca, bc), (ab, bc, ca)]
        f = nf
    pos = (np.array(v) * radius).astype(np.float32)
    nrm = np.array(v).astype(np.float32)
    return pos, nrm, np.array(f).reshape(-1)


def grid_surface(fn, nu, nv, flat=False):
    """Tessellate p = fn(u, v), (u,v) in [0,1]^2, into nu x nv quads (2 triangles each).
    Returns pos, nrm, uv, tan(vec4), idx.  flat=True duplicates vertices per quad (4 verts / quad)."""
    u = np.linspace(0, 1, nu + 1)
    v = np.linspace(0, 1, nv + 1)
    U, V = np.meshgrid(u, v, indexing="ij")
    P = fn(U, V).astype(np.float64)                      # (nu+1, nv+1, 3)
    eps = 1e-4
    Pu = (fn(np.clip(U + eps, 0, 1), V) - fn(np.clip(U - eps, 0, 1), V))
    Pv = (fn(U, np.clip(V + eps, 0, 1)) - fn(U, np.clip(V - eps, 0, 1)))
    Nn = np.cross(Pu, Pv)
    ln = np.linalg.norm(Nn, axis=-1, keepdims=True)
    Nn = np.where(ln > 1e-20, Nn / np.maximum(ln, 1e-20), np.array([0, 1, 0.0]))
    Tn = Pu / np.maximum(np.linalg.norm(Pu, axis=-1, keepdims=True), 1e-20)
    ii, jj = np.meshgrid(np.arange(nu), np.arange(nv), indexing="ij")
    if not flat:
        vid = lambda a, b: (a * (nv + 1) + b)
        q = np.stack([vid(ii, jj), vid(ii + 1, jj), vid(ii + 1, jj + 1), vid(ii, jj), vid(ii + 1, jj + 1), vid(ii, jj + 1)], -1)
        pos = P.reshape(-1, 3)
        nrm = Nn.reshape(-1, 3)
        uv = np.stack([U, V], -1).reshape(-1, 2)
        tan = np.concatenate([Tn.reshape(-1, 3), np.ones((pos.shape[0], 1))], 1)
        return pos.astype(np.float32), nrm.astype(np.float32), uv.astype(np.float32), tan.astype(np.float32), q.reshape(-1)
    corners = [(0, 0), (1, 0), (1, 1), (0, 1)]
    pos = np.stack([P[ii + a, jj + b] for a, b in corners], 2).reshape(-1, 3)
    uv = np.stack([np.stack([U[ii + a, jj + b], V[ii + a, jj + b]], -1) for a, b in corners], 2).reshape(-1, 2)
    fnrm = np.cross(P[ii + 1, jj] - P[ii, jj], P[ii, jj + 1] - P[ii, jj])
    fnrm /= np.maximum(np.linalg.norm(fnrm, axis=-1, keepdims=True), 1e-20)
    nrm = np.repeat(fnrm.reshape(-1, 1, 3), 4, 1).reshape(-1, 3)
    tan = np.concatenate([np.stack([Tn[ii + a, jj + b] for a, b in corners], 2).reshape(-1, 3), np.ones((pos.shape[0], 1))], 1)
    base = (np.arange(nu * nv) * 4)[:, None]
    q = base + np.array([0, 1, 2, 0, 2, 3])[None, :]
    return pos.astype(np.float32), nrm.astype(np.float32), uv.astype(np.float32), tan.astype(np.float32), q.reshape(-1)


def torus_fn(R, r):
    def fn(U, V):
        a, b = 2 * np.pi * U, 2 * np.pi * V
        return np.stack([(R + r * np.cos(b)) * np.cos(a), r * np.sin(b), (R + r * np.cos(b)) * np.sin(a)], -1)
    return fn


def camera(yfov, aspect, zfar=100, znear=0.10000000149011612, name="Camera"):
    return {"name": name, "perspective": {"aspectRatio": aspect, "yfov": yfov, "zfar": zfar, "znear": znear}, "type": "perspective"}


# ----------------------------------------------------------------------------- scenes
def gen_s7_1(out):
    """Stand-in for examples/practice7_1.gltf (Cornell box: 6 planes + 2 boxes = 36 triangles, 2 light triangles)."""
    g = GltfBuilder()
    g.ext_used.add("KHR_materials_emissive_strength")
    g.cameras.append(camera(0.9272952079772949, 1))
    g.materials += [
        material("Wall", [0.800000011920929] * 3 + [1], metallic=0),
        material("Blue wall", [0.20000000298023224, 0.20000000298023224, 0.800000011920929, 1], roughness=0),
        material("Red wall", [0.800000011920929, 0.20000000298023224, 0.20000000298023224, 1], roughness=0),
        material("Light", [0, 0, 0, 1], metallic=0, roughness=0, emissive_strength=5),
        material("Cube", [0.800000011920929] * 3 + [1], metallic=0),
    ]
    g.nodes.append({"camera": 0, "name": "Camera", "translation": [0, 0, 6]})
    pl = plane_mesh()
    planes = [
        (0, {"scale": [2, 2, 2], "translation": [0, -2, 0]}),
        (0, {"scale": [2, 2, 2], "translation": [0, 2, 0]}),
        (0, {"rotation": [0.7071068286895752, 0, 0, 0.7071068286895752], "scale": [2, 2, 2], "translation": [0, 0, -2]}),
        (1, {"rotation": [0.5, 0.5, -0.5, 0.5], "scale": [2, 2, 2], "translation": [2, 0, 0]}),
        (2, {"rotation": [0.5, 0.5, -0.5, 0.5], "scale": [2, 2, 2], "translation": [-2, 0, 0]}),
        (3, {"translation": [0, 1.9805891513824463, 0]}),
    ]
    for k, (mat, trs) in enumerate(planes):
        m = g.add_mesh("Plane.%03d" % k, [g.add_primitive(pl[0], pl[1], pl[4], mat, uv=pl[2])])
        g.nodes.append(dict(mesh=m, name="Plane.%03d" % k, **trs))
    c1 = cube_mesh((-1, -1, -1), (1, 0.5, 1))
    c2 = cube_mesh()
    m = g.add_mesh("Cube.002", [g.add_primitive(c1[0], c1[1], c1[3], 4, uv=c1[2])])
    g.nodes.append({"mesh": m, "name": "Cube", "rotation": [0, 0.2583293914794922, 0, 0.9660568833351135],
                    "scale": [0.5, 1, 0.5], "translation": [-0.7352063655853271, -1, -0.6174172163009644]})
    m = g.add_mesh("Cube.003", [g.add_primitive(c2[0], c2[1], c2[3], 4, uv=c2[2])])
    g.nodes.append({"mesh": m, "name": "Cube.001", "rotation": [0, -0.17631350457668304, 0, 0.9843340516090393],
                    "scale": [0.5, 0.5, 0.5], "translation": [0.9308327436447144, -1.5, 0]})
    return g.write(out)


def gen_s7_4(out):
    """Stand-in for examples/practice7_4.gltf (5 walls, 2 single-triangle lights, 80-triangle metal icosphere)."""
    g = GltfBuilder()
    g.ext_used |= {"KHR_materials_emissive_strength", "KHR_materials_specular", "KHR_materials_ior"}
    g.cameras.append(camera(0.6911112070083618, 1, name="Camera.001"))
    g.materials += [
        material("Material.004", metallic=0, roughness=0.5),
        material("Material.006", metallic=0, roughness=0.5),
        material("Material.005", metallic=0, roughness=0.5),
        material("Material.001", [1, 0.25, 0.125, 1], metallic=0, roughness=0.5),
        material("Wall", [0.125, 0.25, 1, 1], metallic=0, roughness=0.5),
        material("Material.002", [0, 0, 0, 1], metallic=0, roughness=0, emissive_strength=10,
                 extra_ext={"KHR_materials_specular": {"specularColorFactor": [0, 0, 0]},
                            "KHR_materials_ior": {"ior": 1.4500000476837158}}),
        material("Material.003", roughness=0),
    ]
    g.nodes.append({"camera": 0, "name": "Camera", "translation": [0, 0, 7]})
    pl = plane_mesh()
    planes = [
        (0, {"scale": [2, 2, 2], "translation": [0, -2, 0]}),
        (1, {"scale": [2, 2, 2], "translation": [0, 2, 0]}),
        (2, {"rotation": [0.7071068286895752, 0, 0, 0.7071067094802856], "scale": [2, 2.000000238418579, 2.000000238418579], "translation": [0, 0, -2]}),
        (3, {"rotation": [0.5, 0.5, -0.5, 0.5], "scale": [2, 2, 2], "translation": [-2, 0, 0]}),
        (4, {"rotation": [0.5, 0.5, -0.5, 0.5], "scale": [2, 2, 2], "translation": [2, 0, 0]}),
    ]
    for k, (mat, trs) in enumerate(planes):
        m = g.add_mesh("Plane.%03d" % k, [g.add_primitive(pl[0], pl[1], pl[4], mat, uv=pl[2], tan=pl[3])])
        g.nodes.append(dict(mesh=m, name="Plane.%03d" % k, **trs))
    # single-triangle corner lights inside accessor bounds [0,-1,0]..[1,0,1]
    lp = np.array([[0, 0, 0], [1, 0, 0], [0, -1, 1]], np.float32)
    ln = np.tile(np.cross(lp[1] - lp[0], lp[2] - lp[0]) / np.linalg.norm(np.cross(lp[1] - lp[0], lp[2] - lp[0])), (3, 1)).astype(np.float32)
    luv = np.array([[0, 0], [1, 0], [0, 1]], np.float32)
    ltan = np.tile(np.array([[1, 0, 0, 1]], np.float32), (3, 1))
    for k, trs in enumerate([{"translation": [-2, 2, -2]},
                             {"rotation": [0, -0.7071068286895752, 0, 0.7071067094802856], "translation": [2, 2, -2]}]):
        m = g.add_mesh("Plane.%03d" % (5 + k), [g.add_primitive(lp, ln, np.array([0, 1, 2]), 5, uv=luv, tan=ltan)])
        g.nodes.append(dict(mesh=m, name="Plane.%03d" % (5 + k), **trs))
    sp, sn, si = icosphere(2, 0.5)
    suv = np.stack([0.5 + np.arctan2(sn[:, 2], sn[:, 0]) / (2 * np.pi), 0.5 - np.arcsin(np.clip(sn[:, 1], -1, 1)) / np.pi], 1).astype(np.float32)
    stan = np.concatenate([np.stack([-sn[:, 2], np.zeros(len(sn)), sn[:, 0]], 1), np.ones((len(sn), 1))], 1).astype(np.float32)
    m = g.add_mesh("Icosphere", [g.add_primitive(sp, sn, si, 6, uv=suv, tan=stan)])
    g.nodes.append({"mesh": m, "name": "Icosphere", "scale": [2, 2, 2], "translation": [0, -1, 0]})
    return g.write(out)


def gen_s6_1(out):
    """Stand-in for examples/practice6_1.gltf: plane (2), emissive torus 48x12 flat quads (1152), cube (12),
    and a 15 744-triangle procedural 'blob' in place of Suzanne (82 x 96 quads), in Suzanne's accessor bounds."""
    g = GltfBuilder()
    g.ext_used |= {"KHR_materials_emissive_strength", "KHR_materials_specular", "KHR_materials_ior"}
    g.cameras.append(camera(0.5274237051253257, 1.3333333333333333))
    spec = {"KHR_materials_specular": {"specularColorFactor": [0, 0, 0]}, "KHR_materials_ior": {"ior": 1.4500000476837158}}
    g.materials += [
        material("Material.001", [0.800000011920929] * 3 + [1], metallic=0, roughness=0, extra_ext=spec),
        material("Material.002", [0, 0, 0, 1], metallic=0, roughness=0.5, emissive_strength=10),
        material("Material.003", [0.08583559095859528, 0.16039040684700012, 0.8000000715255737, 1], metallic=0, roughness=0.5, extra_ext=spec),
        material("Material.006", [0.8000000715255737, 0.12973719835281372, 0.044038381427526474, 1], roughness=0, extra_ext=spec),
    ]
    g.nodes.append({"camera": 0, "name": "Camera",
                    "rotation": [-0.20997299253940582, 0.3857799470424652, 0.09062844514846802, 0.8937962055206299],
                    "translation": [7.358891487121582, 4.958309173583984, 6.925790786743164]})
    pl = plane_mesh()
    m = g.add_mesh("Plane", [g.add_primitive(pl[0], pl[1], pl[4], 0, uv=pl[2])])
    g.nodes.append({"mesh": m, "name": "Plane", "scale": [3, 3, 3]})
    tp, tn, tuv, _, ti = grid_surface(torus_fn(1.0, 0.0625), 48, 12, flat=True)
    m = g.add_mesh("Torus", [g.add_primitive(tp, tn, ti, 1, uv=tuv)])
    g.nodes.append({"mesh": m, "name": "Torus", "translation": [0, 0.2639119327068329, 0]})
    c = cube_mesh()
    m = g.add_mesh("Cube.001", [g.add_primitive(c[0], c[1], c[3], 2, uv=c[2])])
    g.nodes.append({"mesh": m, "name": "Cube", "rotation": [0, -0.1469700187444687, 0, 0.9891409277915955],
                    "scale": [0.7720770239830017, 0.37140536308288574, 0.37140533328056335],
                    "translation": [1.5750142335891724, 0.317305326461792, -1.5621356964111328]})

    def blob(U, V):
        th, ph = np.pi * (0.02 + 0.96 * V), 2 * np.pi * U
        r = 1.0 + 0.18 * np.sin(5 * ph) * np.sin(4 * th) + 0.1 * np.cos(9 * th + 2 * ph)
        return np.stack([1.25 * r * np.sin(th) * np.cos(ph), 0.9 * r * np.cos(th), 0.75 * r * np.sin(th) * np.sin(ph)], -1)

    bp, bn, buv, _, bi = grid_surface(blob, 82, 96, flat=True)
    m = g.add_mesh("Suzanne", [g.add_primitive(bp, bn, bi, 3, uv=buv)])
    g.nodes.append({"mesh": m, "name": "Suzanne",
                    "rotation": [0.09827060997486115, 0.48264047503471375, 0.029198922216892242, 0.8697980046272278],
                    "scale": [0.6423652172088623] * 3,
                    "translation": [-1.8808989524841309, 0.871877908706665, -1.0177081823349]})
    return g.write(out)


def gen_soup(out, n, half_extent=10.0, seed=1234, with_light=True, cam_z=None, yfov=0.7):
    """cfg3 / cfg5 triangle soup as a glTF: flat normals, u32 indices, diffuse default-ish material, one large
    emissive triangle above the cloud, camera on +Z looking down -Z."""
    g = GltfBuilder()
    g.ext_used.add("KHR_materials_emissive_strength")
    g.cameras.append(camera(yfov, 0))   # aspectRatio 0 -> W/H is used (src/io/scene_parser.cpp:122)
    g.materials += [material("Soup", [0.7, 0.7, 0.7, 1], metallic=0, roughness=0.5),
                    material("Light", [0, 0, 0, 1], metallic=0, roughness=0.5, emissive_strength=20)]
    E = float(half_extent)
    g.nodes.append({"camera": 0, "name": "Camera", "translation": [0, 0, cam_z if cam_z is not None else 3.0 * E]})
    tri = soup_triangles(n, seed, half_extent)
    pos = tri.reshape(-1, 3)
    fn = np.cross(tri[:, 1] - tri[:, 0], tri[:, 2] - tri[:, 0])
    fn /= np.maximum(np.linalg.norm(fn, axis=1, keepdims=True), 1e-30)
    nrm = np.repeat(fn.astype(np.float32), 3, axis=0)
    idx = np.arange(3 * n, dtype=np.uint32)
    m = g.add_mesh("Soup", [g.add_primitive(pos, nrm, idx, 0)])
    g.nodes.append({"mesh": m, "name": "Soup"})
    if with_light:
        lp = np.array([[-2 * E, 1.5 * E, -2 * E], [2 * E, 1.5 * E, -2 * E], [0, 1.5 * E, 3 * E]], np.float32)
        ln = np.tile(np.array([[0, -1, 0]], np.float32), (3, 1))
        m = g.add_mesh("Light", [g.add_primitive(lp, ln, np.array([0, 1, 2]), 1)])
        g.nodes.append({"mesh": m, "name": "Light"})
    return g.write(out)


# Sponza stand-in: per-primitive (material index, triangle count) of examples/sponza/sponza.gltf's mesh 0, in
# order, and the model-space bounds of each primitive's POSITION accessor (transcribed from the JSON).
SPONZA_PRIMS = [
    (0, 31436, [-680.26, -28.05, -341.14], [551.23, 225.10, 266.44]),
    (1, 3472, [-992.96, 42.32, -261.84], [861.89, 73.12, 187.85]),
    (2, 17688, [-990.74, -4.03, -251.64], [859.30, 56.20, 179.31]),
    (3, 4086, [-1428.27, 40.82, -163.39], [1299.87, 758.55, 88.63]),
    (4, 796, [-1866.98, -126.44, -1139.03], [1746.69, 1347.20, 1039.71]),
    (5, 10168, [-1052.88, 211.40, -318.11], [921.94, 1275.01, 245.81]),
    (6, 5876, [-1429.45, 215.80, -645.15], [1302.20, 1263.93, 574.90]),
    (7, 2816, [-1057.22, -2.46, -316.87], [925.60, 221.00, 246.61]),
    (8, 21, [-1432.22, -2.51, -645.43], [1302.22, 415.53, 574.19]),
    (9, 7088, [-1430.31, 182.65, -646.97], [1303.27, 695.32, 575.73]),
    (10, 880, [-509.68, -2.22, -651.54], [1315.52, 716.85, 605.01]),
    (11, 23208, [-1037.65, 506.63, -309.71], [907.41, 699.61, 234.77]),
    (12, 18, [-119.41, 256.73, 568.66], [-8.26, 348.20, 575.51]),
    (13, 16496, [-958.70, 274.96, -289.70], [830.29, 886.14, 209.12]),
    (14, 16512, [-783.79, 339.44, -324.53], [651.96, 520.59, 252.29]),
    (15, 16512, [-786.90, 339.44, -324.53], [648.85, 520.59, 252.29]),
    (16, 11008, [-418.29, 339.44, -324.53], [283.35, 520.59, 252.29]),
    (17, 14336, [-577.62, -0.31, -295.25], [820.17, 283.68, 213.48]),
    (18, 18944, [-946.24, -0.31, -295.25], [820.17, 283.68, 213.48]),
    (19, 14336, [-946.24, -0.31, -289.26], [448.03, 283.68, 213.48]),
    (20, 32, [-644.93, 133.72, -221.49], [513.36, 214.35, 145.13]),
    (21, 19828, [-654.18, 98.67, -246.83], [522.61, 216.95, 170.48]),
    (22, 9184, [-1276.67, -0.69, -523.50], [1194.17, 133.02, 478.26]),
    (23, 3042, [-1427.43, 79.90, -114.69], [1299.03, 264.76, 42.85]),
    (24, 14484, [-1920.95, 1280.51, -1182.81], [1799.91, 1429.43, 1105.43]),
]
# (base colour, normal, metallic-roughness) IMAGE index per material 0..24, from the JSON's textures[].source
SPONZA_MAT_IMAGES = [
    (1, 0, 2), (4, 3, 5), (7, 6, 8), (10, 9, 11), (13, 12, 14), (16, 15, 17), (19, 18, 20), (22, 21, 23), (25, 24, 26),
    (28, 27, 29), (31, 30, 32), (34, 33, 35), (36, None, None), (38, 37, 39), (41, 40, 42), (44, 43, 42), (46, 45, 42),
    (48, 47, 49), (51, 50, 49), (53, 52, 49), (55, 54, 56), (58, 57, 59), (61, 60, 62), (64, 63, 65), (67, 66, 68),
]


def _proc_texture(kind, k, size):
    """Low-entropy procedural 8-bit texture (compresses well as PNG).  kind: 'base' | 'normal' | 'mr'."""
    y, x = np.mgrid[0:size, 0:size].astype(np.float32) / size
    f = 4 + (k % 5) * 2
    if kind == "base":
        brick = ((np.floor(y * f) % 2) * 0.5 + x * f) % 1.0
        mortar = ((brick < 0.06) | ((y * f) % 1.0 < 0.08)).astype(np.float32)
        hue = np.array([0.55 + 0.4 * np.sin(k * 1.3), 0.5 + 0.35 * np.sin(k * 2.1 + 1), 0.45 + 0.35 * np.sin(k * 0.7 + 2)], np.float32)
        shade = 0.75 + 0.25 * np.sin(2 * np.pi * (x * 3 + y * 2 + 0.1 * k))
        img = (hue[None, None, :] * shade[..., None]) * (1 - 0.55 * mortar[..., None])
    elif kind == "normal":
        a = 2 * np.pi * f
        nx = 0.25 * np.cos(a * x) * (0.6 + 0.4 * np.sin(a * y))
        ny = 0.25 * np.cos(a * y) * (0.6 + 0.4 * np.sin(a * x))
        nz = np.sqrt(np.maximum(0.0, 1 - nx * nx - ny * ny))
        img = np.stack([nx, ny, nz], -1) * 0.5 + 0.5
    else:  # metallic-roughness: G = roughness, B = metallic
        rough = 0.55 + 0.35 * np.sin(2 * np.pi * (x * f * 0.5 + 0.07 * k)) * np.sin(2 * np.pi * y * f * 0.5)
        metal = ((np.floor(x * f) + np.floor(y * f) + k) % 7 == 0).astype(np.float32) * (0.8 if k % 3 == 0 else 0.0)
        img = np.stack([np.ones_like(rough), rough, metal], -1)
    return np.clip(np.round(img * 255), 0, 255).astype(np.uint8)


def gen_sponza(out, tex_size=1024):
    """Stand-in for examples/sponza/sponza.gltf: same nodes / camera / 26 materials / 69 images / primitive
    partition (25 groups with the JSON's triangle counts, 262 267 + 2 triangles), geometry regenerated as an
    atrium-like procedural mesh in the groups' bounds, textures regenerated procedurally at 1024x1024."""
    outdir = os.path.dirname(out)
    g = GltfBuilder()
    g.ext_used.add("KHR_materials_emissive_strength")
    g.cameras.append(camera(0.769618570804596, 1.7777777777777777, name="Camera.001"))
    # images / textures: one glTF texture per image (the reference only uses textures[].source)
    kinds = {}
    for (b, n, mr) in SPONZA_MAT_IMAGES:
        kinds[b] = "base"
        if n is not None:
            kinds[n] = "normal"
        if mr is not None:
            kinds[mr] = "mr"
    for k in range(69):
        name = "tex_%02d.png" % k
        kind = kinds.get(k, "base")
        size = 4 if k == 42 else tex_size       # the real set has one 4x4 image
        p = os.path.join(outdir, name)
        if not os.path.exists(p):
            write_png(p, _proc_texture(kind, k, size))
        g.images.append({"mimeType": "image/png", "name": name, "uri": name})
        g.textures.append({"sampler": 0, "source": k})
    for k, (b, n, mr) in enumerate(SPONZA_MAT_IMAGES):
        pbr = {"baseColorFactor": [0.5879999995231628] * 3 + [1], "baseColorTexture": {"index": b}}
        m = {"name": "Material_%d" % k, "pbrMetallicRoughness": pbr}
        if mr is not None:
            pbr["metallicRoughnessTexture"] = {"index": mr}
        else:
            pbr["metallicFactor"] = 0
        if n is not None:
            m["normalTexture"] = {"index": n}
        g.materials.append(m)
    g.materials.append(material("Material.001", [0, 0, 0, 1], metallic=0, roughness=0.5, emissive_strength=100.0))

    prims = []
    for gi, (mat, ntri, lo, hi) in enumerate(SPONZA_PRIMS):
        prims.append(_sponza_group(g, gi, mat, ntri, np.array(lo), np.array(hi)))
    m0 = g.add_mesh("Mesh_0", prims)
    g.nodes.append({"mesh": m0, "name": "Mesh_0", "scale": [0.00800000037997961] * 3})
    g.nodes.append({"camera": 0, "name": "Camera",
                    "rotation": [-0.013672934845089912, 0.789801299571991, 0.007879254408180714, 0.6131597757339478],
                    "translation": [9.900933265686035, 0.9098806381225586, -1.275167465209961]})
    pl = plane_mesh()
    m1 = g.add_mesh("Plane", [g.add_primitive(pl[0], pl[1], pl[4], 25, uv=pl[2], tan=pl[3])])
    g.nodes.append({"mesh": m1, "name": "Plane", "scale": [14.854633331298828, 5.98621129989624, 9.352560043334961],
                    "translation": [-0.3724990487098694, 15.377957344055176, 0]})
    return g.write(out)


def _sponza_group(g, gi, mat, ntri, lo, hi):
    """Geometry of one material group: group 4 = outer shell (floor + 4 walls, open top), group 24 = a roof ring
    around the open atrium, the others = rows of bulged columns / wavy drapes filling the group's bounds.
    The triangle count of the JSON is matched exactly (a strip of small quads/triangles pads the remainder)."""
    parts = []          # (pos, nrm, uv, tan, idx)
    size = hi - lo
    nq_total = ntri // 2

    def add(fn, nu, nv):
        parts.append(grid_surface(fn, nu, nv))
        return nu * nv

    used = 0
    if gi == 4:     # shell: floor + four walls
        faces = [
            lambda U, V: np.stack([lo[0] + size[0] * U, np.full_like(U, lo[1]), lo[2] + size[2] * V], -1),
            lambda U, V: np.stack([lo[0] + size[0] * U, lo[1] + size[1] * V, np.full_like(U, lo[2])], -1),
            lambda U, V: np.stack([lo[0] + size[0] * U, lo[1] + size[1] * V, np.full_like(U, hi[2])], -1),
            lambda U, V: np.stack([np.full_like(U, lo[0]), lo[1] + size[1] * V, lo[2] + size[2] * U], -1),
            lambda U, V: np.stack([np.full_like(U, hi[0]), lo[1] + size[1] * V, lo[2] + size[2] * U], -1),
        ]
        per = nq_total // 5
        nu = max(1, int(np.sqrt(per)))
        nv = max(1, per // nu)
        for f in faces:
            used += add(f, nu, nv)
    elif gi == 24:  # roof ring: four slabs around an open rectangle (0.55 x 0.5 of the footprint)
        ox, oz = 0.275 * size[0], 0.25 * size[2]
        cx, cz = (lo[0] + hi[0]) / 2, (lo[2] + hi[2]) / 2
        rects = [(lo[0], cx - ox, lo[2], hi[2]), (cx + ox, hi[0], lo[2], hi[2]),
                 (cx - ox, cx + ox, lo[2], cz - oz), (cx - ox, cx + ox, cz + oz, hi[2])]
        per = nq_total // 4
        nu = max(1, int(np.sqrt(per)))
        nv = max(1, per // nu)
        for (x0, x1, z0, z1) in rects:
            def f(U, V, x0=x0, x1=x1, z0=z0, z1=z1):
                yy = lo[1] + size[1] * (0.5 + 0.5 * np.sin(np.pi * V) * np.sin(np.pi * U))
                return np.stack([x0 + (x1 - x0) * U, yy, z0 + (z1 - z0) * V], -1)
            used += add(f, nu, nv)
    else:
        drape = gi in (14, 15, 16, 17, 18, 19)       # the curtain / fabric groups of the real scene
        ncol = 6 + (gi % 5) * 2
        if nq_total < 60:       # the three tiny groups (21, 18, 32 triangles): one coarse column
            ncol = 1
        per = max(3, (nq_total * 9 // 10) // ncol)
        nu = max(3, int(np.sqrt(per * 1.5)))
        nv = max(1, per // nu)
        for c in range(ncol):
            side = -1 if c % 2 == 0 else 1
            fx = (c // 2 + 0.5) / ((ncol + 1) // 2)
            cx = lo[0] + size[0] * (0.04 + 0.92 * fx)
            cz = (lo[2] + hi[2]) / 2 + side * 0.42 * size[2]
            if drape:
                def f(U, V, cx=cx, cz=cz, c=c):
                    w = 0.42 * size[0] / ((ncol + 1) // 2)
                    xx = cx + w * (U - 0.5) * 2
                    zz = cz + 0.035 * size[2] * np.sin(6 * np.pi * U + c) * (0.3 + V)
                    return np.stack([xx, hi[1] - size[1] * V, zz], -1)
            else:
                rad = max(6.0, 0.03 * min(size[0], size[2]) * (1 + 0.2 * (gi % 3)))
                def f(U, V, cx=cx, cz=cz, rad=rad, c=c):
                    a = 2 * np.pi * U
                    r = rad * (1 + 0.25 * np.sin(np.pi * V * (2 + gi % 4)) + 0.1 * np.cos(4 * a))
                    return np.stack([cx + r * np.cos(a), lo[1] + size[1] * V, cz + r * np.sin(a)], -1)
            used += add(f, nu, nv)
        # two bounding "beams" so that the group really spans its accessor bounds
    rem_tri = ntri - used * 2
    if rem_tri > 0:     # padding strip of small triangles along the group's lower +z edge
        k = rem_tri
        xs = np.linspace(lo[0], hi[0], k + 2)
        pos = np.stack([np.stack([xs[:-2], np.full(k, lo[1]), np.full(k, hi[2])], -1),
                        np.stack([xs[1:-1], np.full(k, lo[1] + 0.02 * size[1] + 1.0), np.full(k, hi[2])], -1),
                        np.stack([xs[2:], np.full(k, lo[1]), np.full(k, hi[2] - 0.01 * size[2])], -1)], 1).reshape(-1, 3)
        tri = pos.reshape(k, 3, 3)
        fn = np.cross(tri[:, 1] - tri[:, 0], tri[:, 2] - tri[:, 0])
        fn /= np.maximum(np.linalg.norm(fn, axis=1, keepdims=True), 1e-20)
        nrm = np.repeat(fn, 3, 0)
        uv = np.tile(np.array([[0, 0], [0.5, 1], [1, 0]]), (k, 1))
        tan = np.tile(np.array([[1, 0, 0, 1.0]]), (3 * k, 1))
        parts.append((pos.astype(np.float32), nrm.astype(np.float32), uv.astype(np.float32), tan.astype(np.float32), np.arange(3 * k)))
    assert rem_tri >= 0, (gi, ntri, used)
    pos, nrm, uv, tan, idx, base = [], [], [], [], [], 0
    for (p, n, u, t, i) in parts:
        pos.append(p); nrm.append(n); uv.append(u * (3 + gi % 4)); tan.append(t); idx.append(i + base)
        base += p.shape[0]
    pos, nrm, uv, tan, idx = map(np.concatenate, (pos, nrm, uv, tan, idx))
    assert idx.size == 3 * ntri, (gi, idx.size // 3, ntri)
    assert pos.shape[0] <= 65535, (gi, pos.shape[0])
    return g.add_primitive(pos, nrm, idx, mat, uv=uv, tan=tan)


# ----------------------------------------------------------------------------- registry
SCENES = {
    "s7_1": lambda d: gen_s7_1(os.path.join(d, "s7_1.gltf")),
    "s7_4": lambda d: gen_s7_4(os.path.join(d, "s7_4.gltf")),
    "s6_1": lambda d: gen_s6_1(os.path.join(d, "s6_1.gltf")),
    "soup_10k": lambda d: gen_soup(os.path.join(d, "soup_10k.gltf"), 10_000, 2.15),
    "soup_100k": lambda d: gen_soup(os.path.join(d, "soup_100k.gltf"), 100_000, 4.64),
    "soup_1m": lambda d: gen_soup(os.path.join(d, "soup_1m.gltf"), 1_000_000, 10.0),
    "soup_10m": lambda d: gen_soup(os.path.join(d, "soup_10m.gltf"), 10_000_000, 21.5, cam_z=65.0),
    "sponza": lambda d: gen_sponza(os.path.join(d, "sponza", "sponza.gltf")),
    "sponza_small": lambda d: gen_sponza(os.path.join(d, "sponza_small", "sponza.gltf"), tex_size=64),
}


def ensure(name, outdir=GEN_DIR):
    """Generate scene `name` if it is not there yet; returns the .gltf path."""
    sub = name if name.startswith("sponza") else ""
    path = os.path.join(outdir, sub, ("sponza" if sub else name) + ".gltf")
    if not os.path.exists(path):
        os.makedirs(os.path.dirname(path), exist_ok=True)
        SCENES[name](outdir)
    return path


if __name__ == "__main__":
    names = sys.argv[1:2] or ["all-small"]
    out = sys.argv[2] if len(sys.argv) > 2 else GEN_DIR
    if names == ["all-small"]:
        names = ["s7_1", "s7_4", "s6_1", "soup_10k"]
    for nm in names:
        print(ensure(nm, out))
