"""What one GPU's share of a frame costs against frame / N (the strong-scaling loss that is NOT the collective):
usage: share_probe.py <scene> <w> <h> <spp> <bounces> <N> [ranks]  — renders the whole frame, then rank 0's sample share of N, with
RT_PIPES / pass_group variants."""
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
def best(**kw):
    b = None
    for _ in range(5):
        r.render_device(sort=False, **kw)
        if b is None or r.last_stats.loop_seconds < b.loop_seconds:
            b = r.last_stats
    return b
full = best()
print(json.dumps({"what": "whole frame", "ms": round(full.loop_seconds * 1e3, 3), "ideal_share_ms": round(full.loop_seconds * 1e3 / N, 3), "launches": full.kernel_launches}))
ranks = [int(v) for v in sys.argv[7].split(",")] if len(sys.argv) > 7 else [0, N - 1]
for rank in ranks:
    f, c = sample_range(spp, rank, N)
    for label, kw in (("auto", {}), ("single_pipe", dict(single_pipe=True)), ("group1", dict(pass_group=1)), ("group2", dict(pass_group=2)), ("group3", dict(pass_group=3)), ("group8", dict(pass_group=8))):
        s = best(sample_first=f, sample_count=c, **kw)
        print(json.dumps({"what": f"rank {rank} share [{f}, {f + c})", "variant": label, "ms": round(s.loop_seconds * 1e3, 3), "live": s.live_ray_bounces, "launches": s.kernel_launches,
                          "vs_ideal": round(full.loop_seconds / N / s.loop_seconds * (s.live_ray_bounces * N / full.live_ray_bounces), 4)}), flush=True)
r.close()
