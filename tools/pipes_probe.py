"""Two-pipe pass loop A/B: whole-loop time (ev_start..ev_stop, no per-kernel events) per wave size, under RT_PIPES=1|2."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT]
import numpy as np
import cuda_raytracer_b200 as rt

name = sys.argv[1]
w, h, spp, bounces = (int(v) for v in sys.argv[2:6])
groups = [int(g) for g in sys.argv[6].split(",")]
scene = rt.load_scene(name + ".scene", cwd=os.path.join(ROOT, "scenes")).set_image(w, h, spp, bounces)
r = rt.Renderer(0)
r.upload(scene)
r.render_device(sort=False, pass_count=1)
print(f"RT_PIPES={os.environ.get('RT_PIPES', 'default')} {name} {w}x{h}x{spp}x{bounces}")
for g in groups:
    best = None
    for rep in range(3):
        r.render_device(sort=False, pass_group=g)
        s = r.last_stats
        if best is None or s.loop_seconds < best["loop_seconds"]:
            best = s.as_dict()
    img = r.render(sort=False, pass_group=g)
    print(f" group {g:2d}: loop {best['loop_seconds']*1e3:9.2f} ms {best['live_ray_bounces']/best['loop_seconds']*1e-6:9.1f} Mrays/s "
          f"live {best['live_ray_bounces']} launches {best['kernel_launches']} sum {float(img.astype(np.float64).sum()):.6e}")
r.close()
