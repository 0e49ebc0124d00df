"""Host-packed vs device-packed tree: closest-hit differences on the same rays (diagnostic)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")]
import numpy as np
import cuda_raytracer_b200 as rt
from util import oracle_arrays, scene_rays
SCENES = os.path.join(ROOT, "scenes")
for name in sys.argv[1:]:
    host = rt.load_scene(name + ".scene", cwd=SCENES).set_image(64, 48, 20, 6, 1.0)
    dev = rt.load_scene(name + ".scene", use_bvh="device", cwd=SCENES).set_image(64, 48, 20, 6, 1.0)
    rays = scene_rays(oracle_arrays(host), 200_000, seed=31)
    for order in ("hd", "dh", "hd"):
        r = rt.Renderer(0)
        res = {}
        for which in order:
            r.upload(host if which == "h" else dev)
            p, t = r.trace_closest(rays, with_spheres=True, stats=True)
            res[which] = (p.copy(), t.copy(), r.last_stats.internal_visits)
        r.close()
        bad = np.nonzero(res["h"][1] != res["d"][1])[0]
        print(name, order, "diff t:", len(bad), "visits", res["h"][2], res["d"][2], "first", bad[:5],
              [(float(res["h"][1][i]), float(res["d"][1][i]), rays[i].tolist()) for i in bad[:2]])
