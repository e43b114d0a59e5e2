#!/usr/bin/env python3
"""BVH build times (BVH::buildBVH, src/core/bvh.cpp:5-175, same layout on every path): the host builder on all threads
and the GPU builder (csrc/rtb_build.cu) with several hand-over sizes.  One JSON line per case."""
import importlib
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "scenes"))
import gen_scenes  # noqa: E402

rtb = importlib.import_module("raytracing-hw_b200")


def main():
    sizes = [(1_000_000, 10.0), (10_000_000, 21.5)]
    if len(sys.argv) > 1:
        sizes = [(int(a), 10.0 * (int(a) / 1e6) ** (1 / 3)) for a in sys.argv[1:]]
    for n, h in sizes:
        tri = gen_scenes.soup_triangles(n, seed=1234, half_extent=h).reshape(-1, 9)
        digest0 = None
        for label, env in (("host", {"RTB_BUILD_GPU": "0"}), ("gpu below 4096", {"RTB_BUILD_GPU": "1", "RTB_BUILD_GPU_BELOW": "4096"}),
                           ("gpu below 256", {"RTB_BUILD_GPU": "1", "RTB_BUILD_GPU_BELOW": "256"}),
                           ("gpu below 64", {"RTB_BUILD_GPU": "1", "RTB_BUILD_GPU_BELOW": "64"}),
                           ("gpu below 5 (whole tree)", {"RTB_BUILD_GPU": "1", "RTB_BUILD_GPU_BELOW": "5"})):
            os.environ.update(env)
            for rep in range(2):
                t = time.time()
                hs = rtb.HostScene.from_triangles(tri)
                total = time.time() - t
                a, l = hs.nodes()
                import hashlib
                dg = hashlib.sha256(a.tobytes() + l.tobytes() + hs.tri_source().tobytes()).hexdigest()[:16]
                if digest0 is None:
                    digest0 = dg
                info = hs.build_info()
                print(json.dumps({"tris": n, "path": label, "rep": rep, "build_seconds": round(hs.build_seconds, 3), "from_triangles_seconds": round(total, 3),
                                  "layout_equals_host": dg == digest0, "cpus": os.cpu_count(), **{k: (round(v, 3) if isinstance(v, float) else v) for k, v in info.items()}}), flush=True)
                del hs


if __name__ == "__main__":
    main()
