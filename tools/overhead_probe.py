"""Host-side overhead of one rt_render_device call: wall time of the call against the CUDA-event time of its pass loop."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT]
import cuda_raytracer_b200 as rt
SCENES = os.path.join(ROOT, "scenes")
r = rt.Renderer(0)
for name, w, h, spp in (("spheres", 1000, 1000, 1000), ("cornell", 1000, 1000, 100), ("cornell", 1000, 1000, 20), ("lamp", 1000, 1000, 1000)):
    scene = rt.load_scene(name + ".scene", cwd=SCENES).set_image(w, h, spp, 10)
    r.upload(scene)
    for group in (0, 16, 1):
        r.render_device(sort=False, pass_group=group)
        best = (1e9, 0, 0)
        for rep in range(3):
            t = time.perf_counter()
            r.render_device(sort=False, pass_group=group)
            wall = (time.perf_counter() - t) * 1e3
            st = r.last_stats
            if wall < best[0]: best = (wall, st.loop_seconds * 1e3, st.total_seconds * 1e3, st.kernel_launches)
        print(f"{name} {w}x{h}x{spp} group {group:2d}: wall {best[0]:8.2f} ms  loop (events) {best[1]:8.2f} ms  total_seconds {best[2]:8.2f} ms  launches {best[3]}")
r.close()

# the e2e shape of bench.py: upload + render into a pinned host buffer
import numpy as np, torch
r = rt.Renderer(0)
for name, w, h, spp in (("spheres", 1000, 1000, 1000), ("lamp", 1000, 1000, 200)):
    scene = rt.load_scene(name + ".scene", cwd=SCENES).set_image(w, h, spp, 10)
    r.upload(scene)
    pinned = torch.empty(w * h * 3, dtype=torch.float32).pin_memory().numpy().reshape(h, w, 3)
    pageable = np.empty((h, w, 3), np.float32)
    for label, fn in (("render_device", lambda: r.render_device(sort=False)),
                      ("render -> pinned", lambda: r.render(sort=False, out=pinned)),
                      ("render -> pageable", lambda: r.render(sort=False, out=pageable)),
                      ("upload + render -> pinned", lambda: (r.upload(scene), r.render(sort=False, out=pinned))),
                      ("render_device again", lambda: r.render_device(sort=False))):
        fn()
        ts = []
        for rep in range(4):
            t = time.perf_counter(); fn(); ts.append((time.perf_counter() - t) * 1e3)
        print(f"{name}: {label:28s} wall {min(ts):8.2f} ms (max {max(ts):8.2f})  loop {r.last_stats.loop_seconds*1e3:8.2f} ms")
r.close()
