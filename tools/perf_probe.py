"""Perf probe: Mrays/s of the core (sort on/off) and the reference GPU binary on one scene/config."""
import os, re, subprocess, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")]
import numpy as np
import cuda_raytracer_b200 as rt

SCENES = os.path.join(ROOT, "scenes")
name = sys.argv[1] if len(sys.argv) > 1 else "lamp"
w, h, spp, bounces = (int(v) for v in (sys.argv[2:6] if len(sys.argv) > 5 else (1000, 1000, 100, 10)))
variants = sys.argv[6].split(",") if len(sys.argv) > 6 else ["nosort", "sort15", "sort24"]
scene = rt.load_scene(name + ".scene", use_bvh="device" if "devbvh" in sys.argv else True, cwd=SCENES).set_image(w, h, spp, bounces)
r = rt.Renderer(0)
r.upload(scene)
r.render_device(sort=False, pass_count=1)  # warm-up
r.render_device(sort=False, pass_count=1, collect_stats=True)
st = r.last_stats
n_int = st.internal_visits / st.live_ray_bounces; n_tri = st.triangle_tests / st.live_ray_bounces
print(f"{name} {w}x{h}x{spp}x{bounces}: N_int {n_int:.2f} N_tri {n_tri:.2f} bytes/ray-bounce {104+64*n_int+48*n_tri:.0f} live/primary {st.live_ray_bounces/st.primary_rays:.3f}")
print(" live per bounce (1 pass):", [int(v) for v in st.live_per_bounce[:bounces]])
for v in variants:
    kw = dict(sort=v.startswith("sort"), sort_bits=int(v[4:]) if v.startswith("sort") else 0,
              pass_group=int(v[5:]) if v.startswith("group") else 1, root_start="rootstart" in sys.argv)
    best = None
    for rep in range(3):
        r.render_device(timing=True, **kw)
        s = r.last_stats
        if best is None or s.loop_seconds < best.loop_seconds: best = s
    s = best
    mr = s.live_ray_bounces / s.loop_seconds * 1e-6
    print(f" {v:8s}: loop {s.loop_seconds*1e3:8.2f} ms  {mr:9.1f} Mrays/s  trace {s.trace_ms:8.2f} ms  shade {s.shade_ms:8.2f} ms  reorder {s.sort_ms:7.2f} ms  launches {s.kernel_launches}  "
          f"roofline {mr*1e6*(104+64*n_int+48*n_tri)/6543.7e9:.3f}")
    print("   trace ms/bounce:", " ".join(f"{v:.2f}" for v in s.trace_ms_per_bounce[:bounces]))
    print("   shade ms/bounce:", " ".join(f"{v:.2f}" for v in s.shade_ms_per_bounce[:bounces]))
r.close()
ref = os.path.join(ROOT, "oracle", "_ref", "raytracing_ref")
if os.path.exists(ref) and "noref" not in sys.argv:
    tmp = f"/tmp/probe_{name}.scene"
    src = open(os.path.join(SCENES, name + ".scene")).read()
    open(tmp, "w").write(re.sub(r"(?m)^image .*$", f"image {w} {h} {spp} {bounces} 1", src))
    for flags in ([], ["no_sort"]):
        out = subprocess.run([ref, tmp] + flags, cwd=SCENES, capture_output=True, text=True).stdout
        m = re.search(r"GPU Took ([0-9.eE+-]+)s", out)
        if m:
            t = float(m.group(1))
            print(f" reference GPU binary {' '.join(flags) or 'sort':8s}: GPU Took {t:.3f} s  -> {st.live_ray_bounces/st.primary_rays*w*h*spp/t*1e-6:9.1f} Mrays/s (incl. alloc/upload)")
        else:
            print(" reference GPU binary failed:", out[-300:])
