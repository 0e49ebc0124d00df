"""Per-bounce kernel times (event pair per launch, one wave) of two shares of a frame: where one rank's share is slower than
another's with the same number of rays.  usage: share_bounces.py <scene> <w> <h> <spp> <bounces> <N> <rankA> <rankB>"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT]
import cuda_raytracer_b200 as rt
from cuda_raytracer_b200.multi_gpu import sample_range
name = sys.argv[1]
w, h, spp, bounces, N = (int(v) for v in sys.argv[2:7])
scene = rt.load_scene(name + ".scene", cwd=os.path.join(ROOT, "scenes")).set_image(w, h, spp, bounces)
r = rt.Renderer(0)
r.upload(scene)
for rank in (int(v) for v in sys.argv[7:]):
    f, c = sample_range(spp, rank, N)
    best = None
    for _ in range(4):
        r.render_device(sort=False, sample_first=f, sample_count=c, timing=True)
        if best is None or r.last_stats.loop_seconds < best.loop_seconds:
            best = r.last_stats
    print(json.dumps({"rank": rank, "samples": [f, f + c], "timed_ms": round(best.loop_seconds * 1e3, 3), "live": [int(v) for v in best.live_per_bounce[:bounces]],
                      "trace_us": [round(v * 1e3, 1) for v in best.trace_ms_per_bounce[:bounces]],
                      "shade_us": [round(v * 1e3, 1) for v in best.shade_ms_per_bounce[:bounces]]}), flush=True)
    # every pass of the share on its own: the slow one, if there is one
    per_pass = []
    for p in range(f // 20, (f + c + 19) // 20):
        lo, hi = max(f, p * 20), min(f + c, p * 20 + 20)
        r.render_device(sort=False, sample_first=lo, sample_count=hi - lo, timing=True)
        r.render_device(sort=False, sample_first=lo, sample_count=hi - lo, timing=True)
        s = r.last_stats
        per_pass.append({"pass": p, "samples": hi - lo, "ms": round(s.loop_seconds * 1e3, 3), "trace_us_bounce_2_on": [round(v * 1e3) for v in s.trace_ms_per_bounce[2:bounces]]})
    print(json.dumps({"rank": rank, "per_pass": per_pass}), flush=True)
r.close()
