"""Short single-GPU program for ncu: one pass of lamp.scene (10 bounce_kernel launches, optional reorder)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import cuda_raytracer_b200 as rt
name = sys.argv[1] if len(sys.argv) > 1 else "lamp"
w, h = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (1000, 1000)
sort = len(sys.argv) > 4 and sys.argv[4] == "sort"
scene = rt.load_scene(name + ".scene", cwd=os.path.join(ROOT, "scenes")).set_image(w, h, 20, 10)
r = rt.Renderer(0)
r.upload(scene)
r.render_device(sort=sort)
st = r.last_stats
print(f"{name} {w}x{h}: {st.live_ray_bounces / st.loop_seconds * 1e-6:.1f} Mrays/s, {st.kernel_launches} launches")
r.close()
