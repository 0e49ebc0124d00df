"""Node-format A/B: the same frame with the tracer walking f32 / q16 / w4 nodes.
usage: nodes_probe.py <scene> <w> <h> <spp> <bounces> [modes=f32,q16,w4] [reps=3]
Prints one JSON line per mode: Mrays/s of the product path (two pipes), per-bounce kernel times (one wave at a time, CUDA
events), N_int / N_tri of the counting variant, live-ray counters (must be identical across modes)."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT]
import cuda_raytracer_b200 as rt

SCENES = os.path.join(ROOT, "scenes")
name = sys.argv[1] if len(sys.argv) > 1 else "lamp"
w, h, spp, bounces = (int(v) for v in (sys.argv[2:6] if len(sys.argv) > 5 else (1000, 1000, 100, 10)))
modes = sys.argv[6].split(",") if len(sys.argv) > 6 else ["f32", "q16", "w4"]
reps = int(sys.argv[7]) if len(sys.argv) > 7 else 3
use_bvh = os.environ.get("PROBE_BVH", "host")
scene = rt.load_scene(name + ".scene", use_bvh={"host": True, "device": "device", "sbvh": "sbvh"}[use_bvh], cwd=SCENES).set_image(w, h, spp, bounces)
for mode in modes:
    r = rt.Renderer(0, nodes=mode)
    r.upload(scene)
    r.render_device(sort=False, pass_count=1)  # warm-up
    r.render_device(sort=False, pass_count=1, collect_stats=True)
    st = r.last_stats
    n_int, n_tri = st.internal_visits / st.live_ray_bounces, st.triangle_tests / st.live_ray_bounces
    best = None
    for _ in range(reps):
        r.render_device(sort=False)
        if best is None or r.last_stats.loop_seconds < best.loop_seconds:
            best = r.last_stats
    tb = None
    for _ in range(max(1, reps - 1)):
        r.render_device(sort=False, timing=True)
        if tb is None or r.last_stats.loop_seconds < tb.loop_seconds:
            tb = r.last_stats
    print(json.dumps({"scene": name, "image": [w, h, spp, bounces], "bvh": use_bvh, "nodes": r.node_mode, "mrays_per_s": best.live_ray_bounces / best.loop_seconds * 1e-6,
                      "ms_per_frame": best.loop_seconds * 1e3, "live_ray_bounces": best.live_ray_bounces, "n_int": round(n_int, 3), "n_tri": round(n_tri, 3),
                      "timing_mode_ms": round(tb.loop_seconds * 1e3, 2), "trace_ms": round(tb.trace_ms, 2), "shade_ms": round(tb.shade_ms, 2),
                      "trace_ms_per_bounce": [round(v, 2) for v in tb.trace_ms_per_bounce[:bounces]],
                      "shade_ms_per_bounce": [round(v, 2) for v in tb.shade_ms_per_bounce[:bounces]]}), flush=True)
    r.close()
