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


# ------------------------------------------------------------------------------------------------ image decoders
def _image_gltf(dst_dir, image_uris):
    """the cfg2 stand-in with `image_uris` added as images (the loaders decode every image of a file, used or not)"""
    src = _util.scene_path("s7_4")
    g = json.load(open(src))
    base = os.path.dirname(src)
    for b in g["buffers"]:
        if "uri" in b and not b["uri"].startswith("data:"):
            b["uri"] = os.path.relpath(os.path.join(base, b["uri"]), dst_dir)
    g["images"] = [{"uri": u} for u in image_uris]
    path = os.path.join(dst_dir, "images.gltf")
    json.dump(g, open(path, "w"))
    return path


def _decoders_agree(path, ref, n):
    hs = rtb.HostScene.load_gltf(path, 16, 16, 1)
    rs = ref.RefScene.from_gltf(path, 16, 16, 1)
    tx, rtx = hs.textures(), rs.textures()
    assert len(tx) == len(rtx) == n
    for i, (a, b) in enumerate(zip(tx, rtx)):
        assert a.shape == b.shape, (i, a.shape, b.shape)
        assert np.array_equal(a, b), "image %d: %d bytes differ (max %d)" % (i, np.count_nonzero(a != b), np.abs(a.astype(int) - b.astype(int)).max())


@pytest.mark.ref
def test_jpeg_and_png_decoders_equal_the_reference_on_generated_files(ref, tmp_path):
    """own baseline-JPEG and PNG decoders against the reference's (tinygltf -> stb_image) on generated files: every
    chroma layout (4:4:4, 4:2:2, 4:2:0, 4:4:0 via a rotated 4:2:2 is not writable by PIL, so h2v1 / h2v2 / grey), odd
    sizes (partial MCUs), restart intervals, optimised Huffman tables, several qualities; PNG grey / RGB / RGBA / palette
    (+ tRNS) / 1-2-4 bit"""
    PIL = pytest.importorskip("PIL.Image")
    rs = np.random.RandomState(5)
    yy, xx = np.mgrid[0:67, 0:93]
    smooth = np.stack([(xx * 2.7) % 256, (yy * 3.1 + xx) % 256, (xx * yy * 0.05) % 256], -1).astype(np.uint8)
    noisy = rs.randint(0, 256, (40, 24, 3)).astype(np.uint8)
    uris = []

    def add(img, name, **kw):
        PIL.fromarray(img).save(os.path.join(str(tmp_path), name), **kw)
        uris.append(name)

    for k, img in enumerate((smooth, noisy, smooth[:8, :8], smooth[:1, :1], smooth[:17, :1])):
        for sub in (0, 1, 2):                       # PIL: 0 = 4:4:4, 1 = 4:2:2, 2 = 4:2:0
            add(img, "j%d_%d.jpg" % (k, sub), quality=85, subsampling=sub)
    add(smooth, "q30.jpg", quality=30, subsampling=2, optimize=True)
    add(smooth, "q98.jpg", quality=98, subsampling=0, optimize=True)
    add(smooth, "rst.jpg", quality=80, subsampling=2, restart_marker_blocks=3)
    add(smooth[..., 0], "grey.jpg", quality=75)
    add(smooth, "rgb.png")
    add(np.dstack([smooth, smooth[..., :1]]), "rgba.png")
    add(smooth[..., 1], "grey.png")
    PIL.fromarray(smooth).quantize(16).save(os.path.join(str(tmp_path), "pal.png")); uris.append("pal.png")
    PIL.fromarray(smooth).quantize(200).save(os.path.join(str(tmp_path), "pal8.png"), transparency=3); uris.append("pal8.png")
    PIL.fromarray((smooth[..., 0] > 100)).save(os.path.join(str(tmp_path), "bit1.png")); uris.append("bit1.png")
    _decoders_agree(_image_gltf(str(tmp_path), uris), ref, len(uris))


SPONZA_DIR = "/root/reference/examples/sponza"


@pytest.mark.ref
@pytest.mark.skipif(not os.path.isdir(SPONZA_DIR), reason="the reference checkout (with the real Sponza textures) is not here")
def test_real_sponza_textures_equal_the_reference(ref, tmp_path):
    """the REAL 69 Sponza textures (65 baseline JPEGs + 4 PNGs, the files the reference ships) through the product's
    decoders and through the reference's: every RGBA8 byte equal"""
    files = sorted(f for f in os.listdir(SPONZA_DIR) if f.lower().endswith((".jpg", ".png")))
    assert len(files) == 69
    _decoders_agree(_image_gltf(str(tmp_path), [os.path.relpath(os.path.join(SPONZA_DIR, f), str(tmp_path)) for f in files]), ref, 69)


def test_bad_images_are_errors(tmp_path):
    jpg = os.path.join(str(tmp_path), "x.jpg")
    open(jpg, "wb").write(b"\xff\xd8\xff\xc2\x00\x0b\x08\x00\x10\x00\x10\x01\x01\x11\x00\xff\xd9")     # a progressive frame header
    with pytest.raises(Exception) as e:
        rtb.HostScene.load_gltf(_image_gltf(str(tmp_path), ["x.jpg"]), 16, 16, 1)
    assert "progressive" in str(e.value)
    open(jpg, "wb").write(b"\xff\xd8\xff\xdb\x00")
    with pytest.raises(Exception):
        rtb.HostScene.load_gltf(_image_gltf(str(tmp_path), ["x.jpg"]), 16, 16, 1)


def test_oversized_and_truncated_jpegs(tmp_path):
    """a header that announces 32768 x 32768 pixels is refused before anything is allocated; a file cut in the middle of
    its entropy data (the EOI marker met while bits are still wanted) decodes to something or is refused -- it does not
    run the bit reader's counter below zero (checked under -fsanitize in development; here: no crash, a full-size image)"""
    jpg = os.path.join(str(tmp_path), "x.jpg")
    sof0 = b"\xff\xc0\x00\x0b\x08\x80\x00\x80\x00\x01\x01\x11\x00"     # baseline, 8 bit, 32768 x 32768, one component
    open(jpg, "wb").write(b"\xff\xd8" + sof0 + b"\xff\xd9")
    with pytest.raises(Exception) as e:
        rtb.HostScene.load_gltf(_image_gltf(str(tmp_path), ["x.jpg"]), 16, 16, 1)
    assert "too large" in str(e.value)
    PIL = pytest.importorskip("PIL.Image")
    yy, xx = np.mgrid[0:64, 0:64]
    img = np.stack([(xx * 3) % 256, (yy * 5) % 256, (xx + yy) % 256], -1).astype(np.uint8)
    PIL.fromarray(img).save(jpg, quality=90, subsampling=2)
    data = open(jpg, "rb").read()
    for cut in (len(data) // 2, len(data) - 20, len(data) - 3):
        open(jpg, "wb").write(data[:cut] + b"\xff\xd9")
        try:
            hs = rtb.HostScene.load_gltf(_image_gltf(str(tmp_path), ["x.jpg"]), 16, 16, 1)
        except Exception:
            continue
        assert hs.textures()[0].shape == (64, 64, 4)


def test_vertex_index_out_of_range_is_an_error(tmp_path):
    """a malformed file whose index accessor points past its vertex accessors is refused, not read out of bounds"""
    src = _util.scene_path("s7_4")
    g = json.load(open(src))
    base = os.path.dirname(src)
    blob = bytearray(open(os.path.join(base, g["buffers"][0]["uri"]), "rb").read())
    prim = g["meshes"][0]["primitives"][0]
    acc = g["accessors"][prim["indices"]]
    bv = g["bufferViews"][acc["bufferView"]]
    off = bv.get("byteOffset", 0) + acc.get("byteOffset", 0)
    width = 2 if acc["componentType"] == 5123 else 4
    blob[off:off + width] = b"\xff" * width
    g["buffers"][0]["uri"] = "bad.bin"
    open(os.path.join(str(tmp_path), "bad.bin"), "wb").write(bytes(blob))
    path = os.path.join(str(tmp_path), "bad.gltf")
    json.dump(g, open(path, "w"))
    with pytest.raises(Exception) as e:
        rtb.HostScene.load_gltf(path, 16, 16, 1)
    assert "vertex index out of range" in str(e.value)
    # an accessor whose count would wrap the range arithmetic
    g2 = json.load(open(src))
    g2["buffers"][0]["uri"] = os.path.relpath(os.path.join(base, g2["buffers"][0]["uri"]), str(tmp_path))
    g2["accessors"][prim["attributes"]["POSITION"]]["count"] = 2 ** 62
    path2 = os.path.join(str(tmp_path), "huge.gltf")
    json.dump(g2, open(path2, "w"))
    with pytest.raises(Exception) as e:
        rtb.HostScene.load_gltf(path2, 16, 16, 1)
    assert "out of range" in str(e.value)
