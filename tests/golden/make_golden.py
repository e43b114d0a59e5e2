#!/usr/bin/env python3
"""Generates tests/golden/*.npz from the UNMODIFIED reference (oracle/_ref/libref_oracle.so, compiled from
/root/reference by oracle/Makefile).  Run in the build container (where /root/reference exists):

    python tests/golden/make_golden.py

The reference ships no tests or golden vectors, so these fixtures are the reference's own
answers on seeded inputs: BVH layout digests, closest-hit batches (BVH::intersect), light pdfs
(ManyLightsDistribution::pdf), single paths (Scene::intersect with a fresh per-path normal-distribution state) and a
deterministic per-pixel render (ref_harness.cpp: ref_render_deterministic), plus <random> and libm known answers.
The scenes and ray batches are regenerated from seeds by scenes/gen_scenes.py / tests/_util.py."""
import hashlib
import os
import sys

import numpy as np

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import _util  # noqa: E402
from oracle import ref_oracle as ro  # noqa: E402

N_RAYS, N_LIGHT, N_PATHS = 4096, 1024, 384
RENDER = (32, 24, 4)
SCENES = _util.SMALL_SCENES + ["sponza_small"]


def digest(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def light_rays(light_pos, root_box, n, seed):
    """rays aimed at random points of random emissive triangles (so that the all-hits query has hits)."""
    r = _util.box_rays(root_box, n, seed)
    u = _util.gen_scenes.splitmix_u01(seed + 1, 0, n * 3).reshape(n, 3)
    k = np.minimum((u[:, 0] * light_pos.shape[0]).astype(np.int64), light_pos.shape[0] - 1)
    a, b = u[:, 1] * 0.5, u[:, 2] * 0.5
    tgt = light_pos[k, 0:3] + a[:, None] * light_pos[k, 3:6] + b[:, None] * light_pos[k, 6:9]
    r[:, 3:6] = (tgt - r[:, 0:3]).astype(np.float32)
    return r


def main():
    out = {}
    # <random> / libm known answers
    seeds = np.array([1, 5, 7, 12345, 2147483646, 2147483647, 4000000000], np.uint32)
    out["kat_seeds"] = seeds
    out["kat_minstd_first"] = np.array([ro.lib().ref_minstd_first(int(s)) for s in seeds], np.uint32)
    out["kat_uniform_first"] = np.array([ro.lib().ref_uniform_first(int(s), -0.5, 0.5) for s in seeds], np.float32)
    out["kat_normal_seq"] = np.stack([ro.normal_seq(int(s), 33) for s in seeds])
    x = _util.gen_scenes.splitmix_u01(99, 0, 4096)
    out["kat_x"] = x
    out["kat_logf"] = ro.libm(0, x)
    out["kat_powf_inv22"] = ro.libm(1, x, 1.0 / 2.2)
    out["kat_pow5"] = ro.libm(2, x)
    np.savez_compressed(os.path.join(_util.GOLDEN, "kat.npz"), **out)
    print("kat.npz")

    for name in SCENES:
        path = _util.scene_path(name)
        W, H, S = RENDER
        rs = ro.RefScene.from_gltf(path, W, H, S)
        g = {}
        aabb, links = rs.nodes()
        tr = rs.tris()
        g["n_nodes"] = np.int64(aabb.shape[0])
        g["n_tris"] = np.int64(tr["pos"].shape[0])
        g["layout_sha256"] = np.array([digest(aabb), digest(links.astype(np.int32)), digest(tr["pos"]), digest(tr["nrm"]),
                                       digest(tr["tan"]), digest(tr["uv"]), digest(tr["mesh"])])
        la, ll = rs.light_nodes()
        lt = rs.light_tris()
        g["light_sha256"] = np.array([digest(la), digest(ll.astype(np.int32)), digest(lt["pos"])])
        g["root_box"] = aabb[0].copy()
        rays = _util.box_rays(aabb[0], N_RAYS, seed=1000)
        h = rs.closest_hit(rays)
        g["hit_id"] = h["id"].astype(np.int32)
        for k in ("t", "u", "v", "inside"):
            g["hit_" + k] = h[k]
        cam = rs.camera()
        cam["tan_half_fov"] = np.tan(cam["fov"] / 2).astype(np.float32)
        g["cam_pos"], g["cam_axes"], g["cam_tan_half_fov"] = cam["pos"], cam["axes"], cam["tan_half_fov"]
        crays = _util.camera_rays(cam, N_RAYS, seed=2000)
        h = rs.closest_hit(crays)
        g["cam_hit_id"] = h["id"].astype(np.int32)
        g["cam_hit_t"] = h["t"]
        if lt["pos"].shape[0] > 0:
            lr = light_rays(lt["pos"], aabb[0], N_LIGHT, seed=3000)
            g["light_rays"] = lr
            g["light_pdf"] = rs.light_pdf(lr)
        prays = _util.camera_rays(cam, N_PATHS, seed=4000)
        rgb = np.zeros((N_PATHS, 3), np.float32)
        after = np.zeros(N_PATHS, np.uint32)
        for i in range(N_PATHS):
            rgb[i], after[i] = rs.trace_path(prays[i, 0:3], prays[i, 3:6], 1 + i * 7919)
        g["path_rgb"], g["path_engine_after"] = rgb, after
        img, acc, _ = rs.render_deterministic(pixel0_seed=1)
        g["render_accum"], g["render_rgb"] = acc, img
        np.savez_compressed(os.path.join(_util.GOLDEN, name + ".npz"), **g)
        print(name, int(g["n_nodes"]), int(g["n_tris"]), "hit rate %.3f / cam %.3f" % (np.mean(g["hit_id"] >= 0), np.mean(g["cam_hit_id"] >= 0)))


if __name__ == "__main__":
    main()
