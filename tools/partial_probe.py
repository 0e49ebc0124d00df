import json, os, sys
sys.path[:0] = [os.path.dirname(os.path.dirname(os.path.abspath(__file__)))]
import cuda_raytracer_b200 as rt
scene = rt.load_scene("teapot.scene", cwd=os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "scenes")).set_image(1000, 1000, 1000, 10)
r = rt.Renderer(0); r.upload(scene)
def best(**kw):
    b = None
    for _ in range(5):
        r.render_device(sort=False, **kw)
        if b is None or r.last_stats.loop_seconds < b.loop_seconds: b = r.last_stats
    return b
for first, count in ((0, 120), (0, 125), (0, 140), (0, 5), (0, 20), (0, 100), (5, 120)):
    s = best(sample_first=first, sample_count=count)
    print(json.dumps({"samples": [first, first + count], "ms": round(s.loop_seconds * 1e3, 3), "ms_per_20_samples": round(s.loop_seconds * 1e3 * 20 / count, 3), "launches": s.kernel_launches}))
