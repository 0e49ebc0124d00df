"""Reorder pass on / off: per-kernel-class CUDA-event times of the same frame with sort = 0 and sort = 1 at
two key widths. What the reorder could buy at best is the drop in trace time; what it costs is the sort itself plus the gathers
the permutation adds to trace and shade.   usage: reorder_probe.py <scene> <w> <h> <spp> <bounces>"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT]
import cuda_raytracer_b200 as rt
name = sys.argv[1]
w, h, spp, bounces = (int(v) for v in sys.argv[2:6])
scene = rt.load_scene(name + ".scene", cwd=os.path.join(ROOT, "scenes")).set_image(w, h, spp, bounces)
r = rt.Renderer(0)
r.upload(scene)
for label, kw in (("sort0", dict(sort=False)), ("sort15", dict(sort=True, sort_bits=15)), ("sort24", dict(sort=True, sort_bits=24))):
    best = None
    for _ in range(3):
        r.render_device(timing=True, pass_group=4, **kw)
        if best is None or r.last_stats.loop_seconds < best.loop_seconds:
            best = r.last_stats
    s = best
    print(json.dumps({"scene": name, "image": [w, h, spp, bounces], "mode": label, "loop_ms": round(s.loop_seconds * 1e3, 2), "trace_ms": round(s.trace_ms, 2),
                      "shade_ms": round(s.shade_ms, 2), "sort_ms": round(s.sort_ms, 2), "mrays_per_s": round(s.live_ray_bounces / s.loop_seconds * 1e-6),
                      "trace_ms_per_bounce": [round(v, 2) for v in s.trace_ms_per_bounce[:bounces]],
                      "shade_ms_per_bounce": [round(v, 2) for v in s.shade_ms_per_bounce[:bounces]]}), flush=True)
r.close()
