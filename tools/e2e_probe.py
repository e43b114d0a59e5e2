import sys, time, importlib
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/tests"); sys.path.insert(0, "/root/repo/scenes")
import _util, torch
rtb = _util.rtb
import gen_scenes
gen_scenes.ensure("sponza")
hs = rtb.HostScene.load_gltf(_util.scene_path("sponza"), 1920, 1080, 8)
for k in range(3):
    t0 = time.perf_counter(); sc = rtb.Scene(hs, 0); torch.cuda.synchronize(); t1 = time.perf_counter()
    out = sc.render(1920, 1080, 2, want_accum=False); t2 = time.perf_counter()
    out = sc.render(1920, 1080, 2, want_accum=False); t3 = time.perf_counter()
    sc.close(); torch.cuda.synchronize(); t4 = time.perf_counter()
    print("create %.1f ms, first render(2spp) %.1f ms, second %.1f ms, destroy %.1f ms" % ((t1-t0)*1e3, (t2-t1)*1e3, (t3-t2)*1e3, (t4-t3)*1e3))
