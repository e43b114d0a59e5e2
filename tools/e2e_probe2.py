"""Where the end-to-end overhead of a frame goes (scene create / first render on a fresh scene / second render /
destroy), for a light and a heavy scene.  usage: python tools/e2e_probe2.py"""
import sys, time
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests"); sys.path.insert(0, "/root/repo/scenes")
import _util, torch
rtb = _util.rtb
import gen_scenes
for name, W, H, S in (("s7_4", 1280, 720, 256), ("sponza", 1920, 1080, 32)):
    gen_scenes.ensure(name)
    hs = rtb.HostScene.load_gltf(_util.scene_path(name), W, H, S)
    for k in range(3):
        torch.cuda.synchronize()
        t0 = time.perf_counter(); sc = rtb.Scene(hs, 0); torch.cuda.synchronize(); t1 = time.perf_counter()
        o1 = sc.render(W, H, S, want_accum=False); t2 = time.perf_counter()
        o2 = sc.render(W, H, S, want_accum=False); t3 = time.perf_counter()
        sc.close(); torch.cuda.synchronize(); t4 = time.perf_counter()
        print("%s: create %.1f ms, first render %.1f ms (kernels %.1f), second %.1f ms (kernels %.1f), destroy %.1f ms" % (
            name, (t1 - t0) * 1e3, (t2 - t1) * 1e3, o1["stats"]["kernel_ms"], (t3 - t2) * 1e3, o2["stats"]["kernel_ms"], (t4 - t3) * 1e3), flush=True)
