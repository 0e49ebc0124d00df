"""Diagnostic on the CPU (test tooling: drives the oracle, not the product): which rays of one pass of a scene need the most BVH
records per closest-hit query, bounce by bounce — the rays that set the duration of a late-bounce trace launch on the GPU
(profiles/r2g_waves.md: one launch of lamp 1000x1000's pass 23 lasts 2.1 ms instead of 0.33).

usage: slow_ray_study.py <scene> <w> <h> <spp> <bounces> <pass> [first_bounce_to_examine [walk_form 1|2]]
The pass is stepped with the oracle under the GPU path's seeds (ray index x 4137874753 + 279220567 x (remaining x 20 + bounce),
raytracing.cu:89,229,236); before every step from `first_bounce` on, the live rays are walked once more in chunks of 256 with
the visit counters on (near-child-first order, as the CUDA core walks), the heaviest chunks are resolved ray by ray."""
import ctypes as C
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "oracle")]
import pyoracle as po  # noqa: E402
from util import load, oracle_arrays  # noqa: E402

name = sys.argv[1]
w, h, spp, bounces, pass_index = (int(v) for v in sys.argv[2:7])
first_bounce = int(sys.argv[7]) if len(sys.argv) > 7 else 2
walk_form = int(sys.argv[8]) if len(sys.argv) > 8 else 1  # 1 = the reference's slab test (near child first), 2 = the CUDA core's form
A = oracle_arrays(load(name).set_image(w, h, spp, bounces, 1.0))
S = A.oracle_struct()
L = po.port()
rays_to_cast = min(spp - pass_index * 20, 20)
remaining = spp - pass_index * 20 - rays_to_cast
n = w * h * rays_to_cast
data = np.zeros((n, 12), np.float32)
L.oracle_generate(C.byref(S), data.reshape(-1), rays_to_cast, 0, n, remaining)
index = np.arange(n, dtype=np.uint64)


def visits(rays6):
    c = po.Counters()
    m = len(rays6)
    t, idx = np.zeros(m, np.float32), np.zeros(m, np.int32)
    L.oracle_closest_hit_ex(C.byref(S), np.ascontiguousarray(rays6), m, walk_form, 1, t, idx, C.byref(c))
    return c.internal_visits, c.triangle_tests


for b in range(bounces):
    live = np.flatnonzero(~(data[:, 6:9] == 0).all(axis=1)) if b > 0 else np.arange(n)
    line = {"bounce": b, "live": int(len(live))}
    if b >= first_bounce and len(live):
        rays = data[live, :6]
        chunk = 256
        sums = np.array([visits(rays[k:k + chunk])[0] for k in range(0, len(rays), chunk)], np.int64)
        line["visits_per_ray_mean"] = round(float(sums.sum()) / len(rays), 2)
        heavy = np.argsort(sums)[-3:][::-1]
        worst = []
        for ck in heavy:
            per_ray = [visits(rays[j:j + 1]) for j in range(ck * chunk, min(len(rays), (ck + 1) * chunk))]
            j = int(np.argmax([v[0] for v in per_ray]))
            rid = int(live[ck * chunk + j])
            worst.append({"ray_index_in_pass": rid, "pixel": [rid // rays_to_cast % w, rid // rays_to_cast // w], "internal_visits": int(per_ray[j][0]),
                          "triangle_tests": int(per_ray[j][1]), "origin": [round(float(v), 4) for v in data[rid, :3]],
                          "direction": [round(float(v), 5) for v in data[rid, 3:6]]})
        worst.sort(key=lambda r: -r["internal_visits"])
        line["heaviest_rays"] = worst
        line["heaviest_over_mean"] = round(worst[0]["internal_visits"] / max(1e-9, line["visits_per_ray_mean"]), 1)
    print(json.dumps(line), flush=True)
    seeds = ((index * 4137874753 + 279220567 * (remaining * 20 + b)) & 0xFFFFFFFF).astype(np.uint32)
    L.oracle_process(C.byref(S), data.reshape(-1), seeds, n)
