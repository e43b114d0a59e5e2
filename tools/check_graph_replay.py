#!/usr/bin/env python3
"""The CUDA-graph replay of the slot groups' iterations and the suspension of a launch's tail rays (rtb_wavefront.cu,
rtb_trace.cuh) against direct launches without suspension: same sums, same bytes.
Each variant runs in its own process (the switches are read once).  usage: check_graph_replay.py [scene W H S]"""
import hashlib
import importlib
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def child(scene, W, H, S):
    import numpy as np
    import _util
    rtb = _util.rtb
    hs = rtb.HostScene.load_gltf(_util.scene_path(scene), W, H, S)
    sc = rtb.Scene(hs, 0)
    digs = []
    for _ in range(2):   # the second frame replays the cached graph
        out = sc.render(W, H, S)
        digs.append(hashlib.sha256(np.ascontiguousarray(out["accum"]).tobytes() + np.ascontiguousarray(out["rgb"]).tobytes()).hexdigest())
    st = out["stats"]
    print("RESULT", digs[0], digs[1], int(st["closest_hit_rays"]), int(st["launches"]))


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "--child":
        child(sys.argv[2], int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5]))
        sys.exit(0)
    args = sys.argv[1:] or ["sponza", "640", "360", "24"]
    res = {}
    for name, env in (("direct_g1", {"RTB_WF_GROUPS": "1", "RTB_TRACE_TAIL": "0"}), ("tail96_g1", {"RTB_WF_GROUPS": "1"}), ("tail8_g4", {"RTB_WF_GROUPS": "4", "RTB_TRACE_TAIL": "8"}), ("direct_g8", {"RTB_WF_GROUPS": "8", "RTB_WF_GRAPH": "0"}),
                      ("graph_g8", {"RTB_WF_GROUPS": "8", "RTB_WF_GRAPH": "1"}), ("graph_g3", {"RTB_WF_GROUPS": "3", "RTB_WF_GRAPH": "1"})):
        e = dict(os.environ)
        e.update(env)
        r = subprocess.run([sys.executable, os.path.abspath(__file__), "--child"] + args, env=e, capture_output=True, text=True, timeout=600)
        line = [l for l in r.stdout.splitlines() if l.startswith("RESULT")]
        if not line:
            print(name, "FAILED", r.stderr[-2000:])
            sys.exit(1)
        res[name] = line[0].split()[1:]
        print(name, res[name])
    ref = res["direct_g1"][0]
    ok = all(v[0] == ref and v[1] == ref for v in res.values())
    print("graph replay == direct launches:", ok)
    sys.exit(0 if ok else 1)
