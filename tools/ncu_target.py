"""Short single-GPU program for ncu: ONE pass (20 spp) of a scene at the given size, 10 bounces (one trace + one shade
launch per bounce). Run it plain first (writes gpurun_out/ncu_target[_<nodes>].json: live rays per bounce, build id, the
command), then under ncu with the same arguments.   usage: ncu_target.py [scene] [w h] [sort|nosort] [f32|q16|w4]"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import cuda_raytracer_b200 as rt
from cuda_raytracer_b200.core import build_id, kernel_id
name = sys.argv[1] if len(sys.argv) > 1 else "lamp"
w, h = (int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (1000, 1000)
sort = len(sys.argv) > 4 and sys.argv[4] == "sort"
nodes = sys.argv[5] if len(sys.argv) > 5 else None
scene = rt.load_scene(name + ".scene", cwd=os.path.join(ROOT, "scenes")).set_image(w, h, 20, 10)
r = rt.Renderer(0, nodes=nodes)
r.upload(scene)
r.render_device(sort=sort)
st = r.last_stats
print(f"{name} {w}x{h} nodes={r.node_mode}: {st.live_ray_bounces / st.loop_seconds * 1e-6:.1f} Mrays/s, {st.kernel_launches} launches")
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump({"build_id": build_id(), "kernel_id": kernel_id(), "workload": f"{name}.scene {w}x{h}, one pass of 20 spp, 10 bounces, nodes={r.node_mode}, sort={int(sort)}",
           "command": "python " + " ".join(["tools/ncu_target.py"] + sys.argv[1:]), "live_per_bounce": [int(v) for v in st.live_per_bounce[:10]],
           "mrays_per_s_cold_single_pass": st.live_ray_bounces / st.loop_seconds * 1e-6},
          open(os.path.join(ROOT, "gpurun_out", f"ncu_target{'_' + nodes if nodes else ''}.json"), "w"))
r.close()
