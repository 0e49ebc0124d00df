#!/usr/bin/env python3
"""Generate tests/golden/ref_vectors.npz from the REFERENCE's own code (oracle/_ref/libref_shim.so).

Run in the build container (needs /root/reference at shim build time):  python tests/golden/make_golden.py
The reference has no test-suite or golden files of its own, so these vectors are outputs of the
reference's functions (precise host build, see oracle/Makefile) on seeded inputs. They pin the C restatement
(oracle/rt_oracle.c) and the product's host scene layer on machines where /root/reference does not exist.
"""
import ctypes as C
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path[:0] = [ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "tests")]

import pyoracle as po  # noqa: E402
from util import ALL_SCENES, SCENES, scene_rays  # noqa: E402


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def main():
    R = po.shim()
    assert R is not None, "build oracle/_ref first (make -C oracle ref)"
    out = {}

    seeds = np.array([0, 1, 2, 12345, 279220567, 0x7FFFFFFF, 0x80000000, 0xFFFFFFFF], np.uint32)
    seq = np.zeros((len(seeds), 32), np.uint32)
    flt = np.zeros((len(seeds), 6), np.float32)
    for i, s in enumerate(seeds):
        R.ref_xor_sequence(int(s), 32, seq[i])
        R.ref_random_floats(int(s), flt[i])
    out["rng_seeds"], out["rng_sequences"], out["rng_floats"] = seeds, seq, flt

    idx = np.array([0, 1, 5, 31, 19_999_999, 165_887_999], np.int32)
    sd = np.array([0, 3, 20, 980, 19_999], np.int32)
    out["seed_ray_index"], out["seed_seed"] = idx, sd
    out["seed_gen"] = np.array([[R.ref_gpu_gen_seed(int(i), int(s)) for s in sd] for i in idx], np.uint32)
    out["seed_proc"] = np.array([[R.ref_gpu_proc_seed(int(i), int(s)) for s in sd] for i in idx], np.uint32)

    rs = np.random.RandomState(1)
    d = rs.randn(4096, 3).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    d[:10] = [[0, 0, 1], [0, 0, -1], [1, 0, 0], [0, 1, 0], [-1, 0, 0], [0, -1, 0], [0.70710677, 0.70710677, 0], [0, 0, 0],
              [1e-30, 0, 1], [0.5, 0.5, -0.70710677]]
    uv = np.zeros((len(d), 2), np.float32)
    R.ref_project(d, len(d), uv)
    out["project_dirs"], out["project_uv"] = d, uv

    # slab test known answers, incl. boundary cases (ray in a face plane, zero direction components)
    boxes = np.array([[0, 0, 0, 1, 1, 1]] * 8 + [[-1, 2, -3, 4, 2, 5]] * 2, np.float32)  # the last two are flat in y
    rays = np.array([[-1, .5, .5, 1, 0, 0], [2, .5, .5, 1, 0, 0], [.5, .5, .5, 0, 0, 1], [-1, 0, .5, 1, 0, 0], [-1, 1, 1, 1, 0, 0],
                     [-1, -1, -1, 1, 1, 1], [-1, 1.0001, .5, 1, 0, 0], [0, 0, 0, 0, 1, 0], [0, 0, 0, 0, 1, 0], [0, 2, 0, 1, 0, 0]], np.float32)
    tmax = np.array([1e30, 1e30, 1e30, 1e30, 1e30, 1e30, 1e30, 0.5, 1e30, 1e30], np.float32)
    hit = np.zeros(len(rays), np.int32)
    tmin = np.zeros(len(rays), np.float32)
    for i in range(len(rays)):
        t = C.c_float()
        hit[i] = R.ref_ray_aabb(boxes[i], rays[i], float(tmax[i]), C.byref(t))
        tmin[i] = t.value
    out["aabb_boxes"], out["aabb_rays"], out["aabb_tmax"], out["aabb_hit"], out["aabb_tmin"] = boxes, rays, tmax, hit, tmin

    for name in ALL_SCENES:
        for use_bvh in (True, False):
            tag = f"{name}_{'bvh' if use_bvh else 'nobvh'}"
            ref = po.RefScene(name + ".scene", use_bvh=use_bvh, cwd=SCENES)
            A = ref.arrays()
            out[f"scene_{tag}_counts"] = np.array([len(A.spheres), len(A.triangles), len(A.materials), len(A.bvh)], np.int64)
            out[f"scene_{tag}_sha"] = np.array([sha(A.spheres), sha(A.triangles), sha(A.material_indices), sha(A.materials), sha(A.bvh)])
            if use_bvh:
                out[f"scene_{tag}_floats"] = np.concatenate([np.atleast_1d(A.f[k]) for k, _ in po.FLOAT_NAMES]).astype(np.float32)
                out[f"scene_{tag}_image"] = np.array([A.width, A.height, A.ray_count, A.bounces], np.int64)
        # closest hit on seeded rays (the product loader's arrays are byte-identical, so the same rays are reproducible)
        ref = po.RefScene(name + ".scene", cwd=SCENES).set_image(32, 24, 25, 6, 1.0)
        A = ref.arrays()
        n = 4000
        r6 = scene_rays(A, n, seed=99)
        t = np.zeros(n, np.float32)
        i = np.zeros(n, np.int32)
        R.ref_closest_hit(ref.handle, r6, n, t, i)
        out[f"hit_{name}_t"], out[f"hit_{name}_idx"] = t, i
        # generate + 3 process steps on the first 2048 rays of a pass, GPU seeds
        m = 2048
        g = np.zeros((m, 12), np.float32)
        R.ref_generate(ref.handle, g, 20, 0, m, None, 4)
        out[f"gen_{name}"] = g.copy()
        for b in range(3):
            s = np.array([R.ref_gpu_proc_seed(k, 4 * 20 + b) for k in range(m)], np.uint32)
            R.ref_process(ref.handle, g, s, m)
            out[f"proc_{name}_b{b}"] = g.copy()
        # whole pass loop, both seed modes (25 spp = one full + one ragged pass)
        for mode in (0, 1):
            img = np.zeros((24 * 32, 3), np.float32)
            live = C.c_longlong()
            R.ref_render(ref.handle, mode, img, C.byref(live))
            out[f"render_{name}_mode{mode}"] = img
            out[f"render_{name}_mode{mode}_live"] = np.array([live.value], np.int64)

    path = os.path.join(HERE, "ref_vectors.npz")
    np.savez_compressed(path, **out)
    print(f"wrote {path}: {len(out)} arrays, {os.path.getsize(path) / 1024:.0f} KiB")


if __name__ == "__main__":
    main()
