import importlib,sys,time,os
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "scenes"))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gen_scenes
rtb=importlib.import_module("raytracing-hw_b200")
print("cpus", os.cpu_count(), flush=True)
for n,h in ((1_000_000,10.0),(10_000_000,21.5)):
    tri=gen_scenes.soup_triangles(n, seed=1234, half_extent=h).reshape(-1,9)
    for rep in range(2):
        t=time.time(); hs=rtb.HostScene.from_triangles(tri); print(n, 'build_seconds', round(hs.build_seconds,3), 'total', round(time.time()-t,3), flush=True)
        del hs
