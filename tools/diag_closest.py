"""Diagnostic: print closest-hit disagreements between the CUDA core and the CPU oracle / reference kernel."""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")]
import numpy as np
import pyoracle as po
import cuda_raytracer_b200 as rt
from util import load, oracle_arrays, oracle_closest, scene_rays, closest_hit_mismatches, SCENES

r = rt.Renderer(0)
for name in sys.argv[1:] or ["cornell"]:
    scene = load(name).set_image(64, 48, 25, 10, 1.0)
    r.upload(scene)
    A = oracle_arrays(scene)
    rays = scene_rays(A, 200_000, seed=7)
    prim, t = r.trace_closest(rays)
    wp, wt, _ = oracle_closest(A, rays, 1)
    bad, ties = closest_hit_mismatches(prim, t, wp, wt, 1e-5)
    print(name, "vs oracle: bad", len(bad), "ties", ties)
    for i in bad[:12]:
        tri = A.triangles[wp[i] - len(A.spheres)] if wp[i] >= 0 else None
        print(f"  ray {i}: gpu prim {prim[i]} t {t[i]!r}  oracle prim {wp[i]} t {wt[i]!r} rel {abs(t[i]-wt[i])/max(abs(wt[i]),1e-30):.2e} d {rays[i,3:]}")
    L = po.shim()
    if L is not None:
        ref = po.RefScene(name + ".scene", cwd=SCENES)
        tr = np.zeros(len(rays), np.float32); ir = np.zeros(len(rays), np.int32)
        rc = L.ref_closest_hit_gpu(ref.handle, rays, len(rays), tr, ir)
        bad2, ties2 = closest_hit_mismatches(prim, t, ir, tr, 1e-5)
        bad3, ties3 = closest_hit_mismatches(ir, tr, wp, wt, 1e-5)
        print(name, "vs reference kernel: rc", rc, "bad", len(bad2), "ties", ties2, "| reference kernel vs oracle: bad", len(bad3), "ties", ties3)
        for i in bad2[:8]:
            print(f"  ray {i}: gpu prim {prim[i]} t {t[i]!r}  refkernel prim {ir[i]} t {tr[i]!r}")
