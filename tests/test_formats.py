"""Input formats beyond the reference's: interleaved vertex buffers (bufferView.byteStride)
and the binary .glb container must load to exactly the scene the tightly packed .gltf + .bin gives."""
import json
import os
import struct

import numpy as np
import pytest

import _util

rtb = _util.rtb
COMPS = {"SCALAR": 1, "VEC2": 2, "VEC3": 3, "VEC4": 4}


def _interleave(src_gltf, dst_dir):
    """rewrites every primitive's vertex attributes into one interleaved bufferView (with byteStride)"""
    g = json.load(open(src_gltf))
    base = os.path.dirname(src_gltf)
    blob = open(os.path.join(base, g["buffers"][0]["uri"]), "rb").read()
    out = bytearray(blob)
    done = {}
    for mesh in g["meshes"]:
        for prim in mesh["primitives"]:
            names = [a for a in ("POSITION", "NORMAL", "TEXCOORD_0", "TANGENT") if a in prim["attributes"]]
            key = tuple(prim["attributes"][a] for a in names)
            if key in done:
                continue
            done[key] = True
            cols = []
            for a in names:
                acc = g["accessors"][prim["attributes"][a]]
                bv = g["bufferViews"][acc["bufferView"]]
                start = bv.get("byteOffset", 0) + acc.get("byteOffset", 0)
                n = COMPS[acc["type"]]
                cols.append(np.frombuffer(blob, np.float32, acc["count"] * n, start).reshape(acc["count"], n))
            inter = np.concatenate(cols, axis=1).astype(np.float32)
            stride = inter.shape[1] * 4
            while len(out) % 4:
                out.append(0)
            view = {"buffer": 0, "byteOffset": len(out), "byteLength": inter.nbytes, "byteStride": stride}
            out += inter.tobytes()
            g["bufferViews"].append(view)
            off = 0
            for a, c in zip(names, cols):
                acc = g["accessors"][prim["attributes"][a]]
                acc["bufferView"] = len(g["bufferViews"]) - 1
                acc["byteOffset"] = off
                off += c.shape[1] * 4
    g["buffers"][0] = {"uri": "interleaved.bin", "byteLength": len(out)}
    for im in g.get("images", []):     # textures stay where they are
        if "uri" in im and not im["uri"].startswith("data:"):
            im["uri"] = os.path.relpath(os.path.join(base, im["uri"]), dst_dir)
    open(os.path.join(dst_dir, "interleaved.bin"), "wb").write(bytes(out))
    path = os.path.join(dst_dir, "interleaved.gltf")
    json.dump(g, open(path, "w"))
    return path


def _to_glb(src_gltf, dst_dir):
    g = json.load(open(src_gltf))
    base = os.path.dirname(src_gltf)
    blob = open(os.path.join(base, g["buffers"][0]["uri"]), "rb").read()
    g["buffers"][0] = {"byteLength": len(blob)}
    for im in g.get("images", []):
        if "uri" in im and not im["uri"].startswith("data:"):
            im["uri"] = os.path.relpath(os.path.join(base, im["uri"]), dst_dir)
    js = json.dumps(g).encode()
    js += b" " * (-len(js) % 4)
    binc = blob + b"\0" * (-len(blob) % 4)
    total = 12 + 8 + len(js) + 8 + len(binc)
    path = os.path.join(dst_dir, "scene.glb")
    with open(path, "wb") as f:
        f.write(b"glTF" + struct.pack("<II", 2, total))
        f.write(struct.pack("<II", len(js), 0x4E4F534A) + js)
        f.write(struct.pack("<II", len(binc), 0x004E4942) + binc)
    return path


def _same_scene(a, b):
    ta, tb = a.tris(), b.tris()
    for k in ta:
        assert np.array_equal(ta[k], tb[k]), k
    for x, y in zip(a.nodes(), b.nodes()):
        assert np.array_equal(x, y)
    assert a.desc.n_lights == b.desc.n_lights


@pytest.mark.parametrize("name", ["s7_4", "s6_1", "sponza_small"])
def test_interleaved_byte_stride(name, tmp_path):
    src = _util.scene_path(name)
    want = rtb.HostScene.load_gltf(src, 64, 48, 4)
    got = rtb.HostScene.load_gltf(_interleave(src, str(tmp_path)), 64, 48, 4)
    _same_scene(want, got)


@pytest.mark.parametrize("name", ["s7_4", "sponza_small"])
def test_glb_container(name, tmp_path):
    src = _util.scene_path(name)
    want = rtb.HostScene.load_gltf(src, 64, 48, 4)
    got = rtb.HostScene.load_gltf(_to_glb(src, str(tmp_path)), 64, 48, 4)
    _same_scene(want, got)


def test_truncated_glb_is_an_error(tmp_path):
    p = _to_glb(_util.scene_path("s7_4"), str(tmp_path))
    raw = open(p, "rb").read()
    open(p, "wb").write(raw[:40])
    with pytest.raises(Exception):
        rtb.HostScene.load_gltf(p, 64, 48, 4)
