"""Device BVH build probe: build time per scene (best of 5 uploads), with RT_BUILD_TRACE=1 the per-level times."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT]
import cuda_raytracer_b200 as rt
SCENES = os.path.join(ROOT, "scenes")
r = rt.Renderer(0)
for name in sys.argv[1:] or ["teapot", "lamp"]:
    dev = rt.load_scene(name + ".scene", use_bvh="device", cwd=SCENES)
    best, wall = 1e30, 1e30
    for rep in range(5):
        t = time.perf_counter()
        r.upload(dev)
        wall = min(wall, (time.perf_counter() - t) * 1e3)
        best = min(best, r.bvh_device_ms)
    print(f"{name}: device build {best:.2f} ms (events), upload call {wall:.2f} ms wall", flush=True)
r.close()
