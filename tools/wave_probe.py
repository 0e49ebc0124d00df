"""Where the fixed cost of a wave goes (teapot 1000x1000, 1000 spp, 10 bounces; rank 0's share of 8 = samples [0, 125)):
per-bounce kernel times of small sample ranges (event timing: one pipe), then the share under wave-size / pipe-count variants."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT]
import cuda_raytracer_b200 as rt
scene = rt.load_scene("teapot.scene", cwd=os.path.join(ROOT, "scenes")).set_image(1000, 1000, 1000, 10)


def best(r, **kw):
    b = None
    for _ in range(5):
        r.render_device(sort=False, **kw)
        if b is None or r.last_stats.loop_seconds < b.loop_seconds:
            b = r.last_stats
    return b


r = rt.Renderer(0)
r.upload(scene)
for first, count in ((0, 5), (0, 20), (0, 60), (0, 125)):
    s = best(r, sample_first=first, sample_count=count, timing=True)
    print(json.dumps({"samples": [first, first + count], "timed_ms": round(s.loop_seconds * 1e3, 3),
                      "live": [int(v) for v in s.live_per_bounce[:10]],
                      "trace_us": [round(v * 1e3, 1) for v in s.trace_ms_per_bounce[:10]],
                      "shade_us": [round(v * 1e3, 1) for v in s.shade_ms_per_bounce[:10]]}), flush=True)
r.close()
for pipes in (1, 2, 3, 4):
    os.environ["RT_PIPES"] = str(pipes)
    r = rt.Renderer(0)
    r.upload(scene)
    for group in (0, 1, 2, 3, 4, 7):
        s = best(r, sample_first=0, sample_count=125, pass_group=group)
        print(json.dumps({"pipes": pipes, "pass_group": group, "ms": round(s.loop_seconds * 1e3, 3), "launches": s.kernel_launches}), flush=True)
    r.close()
