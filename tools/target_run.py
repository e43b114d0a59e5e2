#!/usr/bin/env python3
"""The target sentence of BASELINE.json's north_star, run through the product's CLI: `./run.sh sponza.gltf 1920 1080 1024
out.ppm` on every visible GPU (one process, a thread per GPU: rtb_render_multi), timed by wall clock like a user would,
and the image compared with the one-GPU frame of the same command -- which tests/test_gpu_baseline_sizes.py proves
bit-identical to the reference's own per-pixel stream:
  tile split (the CLI default): the PPM must be the one-GPU PPM byte for byte;
  RT_SPLIT=sample             : other engine streams, so only statistics can agree: the mean level must be within
                                0.1 % of the bit-exact frame's (an energy bias of 1 % would fail by 10 x); the RMSE over
                                the RGB8 bytes is the Monte-Carlo noise of two independent estimates and is reported,
                                not gated (the reference's per-pixel seeds are fixed, so the CLI cannot produce a second
                                independent bit-exact frame to measure a floor with; tests/test_gpu_dist.py gates the
                                RMSE against the as-shipped reference's run-to-run floor at small sizes).
usage: tools/target_run.py [--spp 1024] [--out gpurun_out/x.json]"""
import argparse
import hashlib
import json
import os
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "scenes"))
import gen_scenes  # noqa: E402

# tolerance of the sample-split frame's mean level against the bit-exact frame's (see the docstring)
MEAN_REL_MAX = 0.001


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--spp", type=int, default=1024)
    ap.add_argument("--gpus", type=int, default=0, help="0: all visible")
    ap.add_argument("--out", default="")
    a = ap.parse_args()
    import torch
    ndev = a.gpus or torch.cuda.device_count()
    scene = gen_scenes.ensure("sponza")
    exe = os.path.join(ROOT, "run.sh")
    W, H = 1920, 1080
    res = {"command": "./run.sh sponza.gltf %d %d %d out.ppm" % (W, H, a.spp), "gpus": ndev, "runs": {}}
    frames = {}
    for tag, env in (("one_gpu", {"RT_GPUS": "1"}), ("tile", {"RT_GPUS": str(ndev)}), ("sample", {"RT_GPUS": str(ndev), "RT_SPLIT": "sample"})):
        out = "/tmp/target_%s.ppm" % tag
        t0 = time.perf_counter()
        r = subprocess.run([exe, scene, str(W), str(H), str(a.spp), out], capture_output=True, text=True,
                           env=dict(os.environ, RT_STATS="1", **env), timeout=1500)
        wall = time.perf_counter() - t0
        if r.returncode != 0:
            res["runs"][tag] = {"rc": r.returncode, "stderr": r.stderr[-1500:]}
            continue
        data = open(out, "rb").read()
        hdr = len(b"P6\n%d %d\n255\n" % (W, H))
        frames[tag] = np.frombuffer(data[hdr:], np.uint8)
        res["runs"][tag] = {"rc": 0, "wall_s": round(wall, 3), "mpix_spp_per_s_wall": round(W * H * a.spp / wall * 1e-6, 1),
                            "sha256": hashlib.sha256(data).hexdigest(), "stdout": r.stdout.strip().splitlines()[-6:]}
        for ln in r.stderr.strip().splitlines():
            if ln.startswith("{"):   # RT_STATS: render seconds without the load (first frame: includes creating the communicators)
                res["runs"][tag]["stats"] = json.loads(ln)
    ok = all(v.get("rc") == 0 for v in res["runs"].values())
    if ok:
        res["tile_equals_one_gpu_bytes"] = bool(np.array_equal(frames["tile"], frames["one_gpu"]))
        x, y = frames["one_gpu"].astype(np.float64), frames["sample"].astype(np.float64)
        res["sample_vs_bit_exact"] = {"rmse_rgb8": float(np.sqrt(np.mean((x - y) ** 2))), "mean_rel_diff": float(abs(x.mean() - y.mean()) / x.mean()),
                                      "max_abs_diff": int(np.abs(x - y).max()), "mean_rel_max": MEAN_REL_MAX}
        ok = res["tile_equals_one_gpu_bytes"] and res["sample_vs_bit_exact"]["mean_rel_diff"] <= MEAN_REL_MAX
    res["ok"] = bool(ok)
    line = json.dumps(res)
    if a.out:
        open(a.out, "w").write(line + "\n")
    print(line)
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
