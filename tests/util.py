"""Shared helpers of the test-suite: scene loading through the product's host layer, oracle plumbing, comparisons."""
import ctypes as C
import os

import numpy as np

import pyoracle as po
import cuda_raytracer_b200 as rt

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SCENES = os.path.join(ROOT, "scenes")
GOLDEN = os.path.join(ROOT, "tests", "golden")
ALL_SCENES = ["cornell", "cornell_plus", "spheres", "teapot", "glass_teapot", "lamp"]

_scene_cache = {}


def load(name, use_bvh=True):
    """Product host loader (cached per (name, use_bvh)); the image block is reset by the caller with set_image."""
    key = (name, use_bvh)
    if key not in _scene_cache:
        _scene_cache[key] = rt.load_scene(name + ".scene", use_bvh=use_bvh, cwd=SCENES)
    return _scene_cache[key]


def oracle_arrays(scene):
    """pyoracle.SceneArrays built from the product loader's arrays (which tests/test_host_scene.py pins to the reference)."""
    a = scene.arrays()
    floats = {k: a[k] for k in ("camera_position", "forward", "up", "vertical_fov", "exposure", "min_coord", "inv_dimensions",
                                "scaled_right", "scaled_up", "near_plane_top_left", "inv_width", "inv_height")}
    return po.SceneArrays(a["spheres"], a["triangles"], a["material_indices"], a["materials"], a["bvh"], a["env"],
                          a["env_w"], a["env_h"], (a["width"], a["height"], a["ray_count"], a["bounces"]), floats)


def scene_rays(arrays, n, seed):
    """n seeded rays: origins uniform in the (slightly inflated) scene box, directions uniform on the sphere."""
    rs = np.random.RandomState(seed)
    if len(arrays.triangles):
        lo, hi = arrays.bvh[0]["min"].astype(np.float64), arrays.bvh[0]["max"].astype(np.float64)
    else:
        lo, hi = np.full(3, -8.0), np.full(3, 8.0)
    ext = hi - lo
    o = lo - 0.1 * ext + 1.2 * ext * rs.rand(n, 3)
    d = rs.randn(n, 3)
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    return np.ascontiguousarray(np.concatenate([o, d], 1), np.float32)


def oracle_closest(arrays, rays, near_first, with_spheres=False):
    """Closest hit per ray from the oracle: the BVH alone (Scene::bvh_closest_hit_distance), or preceded by process_ray's
    sphere loop (with_spheres)."""
    S = arrays.oracle_struct()
    n = len(rays)
    t = np.zeros(n, np.float32)
    idx = np.zeros(n, np.int32)
    c = po.Counters()
    po.port().oracle_closest_hit_ex(C.byref(S), rays, n, 1 if near_first else 0, 1 if with_spheres else 0, t, idx, C.byref(c))
    return idx, t, c


def oracle_render(arrays, seed_mode=1, pass_first=0, pass_count=-1):
    S = arrays.oracle_struct()
    img = np.zeros((arrays.height, arrays.width, 3), np.float32)
    c = po.Counters()
    po.port().oracle_render(C.byref(S), seed_mode, pass_first, pass_count, img.reshape(-1), C.byref(c))
    return img, c


def oracle_render_select(arrays, seed_mode=1, pass_first=0, pass_count=-1, ray_first=0, ray_count=-1, window=None):
    """The oracle's pass loop over a subset of the frame's rays (pass range, ray range of the frame's sequence, pixel window
    (x0, y0, w, h)); full-size image out."""
    S = arrays.oracle_struct()
    img = np.zeros((arrays.height, arrays.width, 3), np.float32)
    c = po.Counters()
    win = np.array(window, np.int32) if window is not None else None
    po.port().oracle_render_select(C.byref(S), seed_mode, pass_first, pass_count, int(ray_first), int(ray_count),
                                   win.ctypes.data if win is not None else None, img.reshape(-1), C.byref(c))
    return img, c


def oracle_render_samples(arrays, sample_first, sample_count, seed_mode=1):
    """The oracle's pass loop over samples [sample_first, sample_first + sample_count) of every pixel."""
    S = arrays.oracle_struct()
    img = np.zeros((arrays.height, arrays.width, 3), np.float32)
    c = po.Counters()
    po.port().oracle_render_samples(C.byref(S), seed_mode, int(sample_first), int(sample_count), img.reshape(-1), C.byref(c))
    return img, c


def scene_diag(arrays):
    """Diagonal of the triangle root box (1.0 when there are no triangles)."""
    if not len(arrays.triangles):
        return 1.0
    return float(np.linalg.norm(arrays.bvh[0]["max"].astype(np.float64) - arrays.bvh[0]["min"].astype(np.float64)))


def closest_hit_mismatches(prim_a, t_a, prim_b, t_b, rel=1e-5, abs_tol=0.0):
    """Indices where two closest-hit answers disagree beyond the stated tolerance: primitive ids must be equal unless
    the two hit distances are within tolerance (a tie between adjacent primitives); t within `rel` relative
    (+ `abs_tol`, used only against the un-contracted CPU arithmetic: Moller-Trumbore subtracts scene-sized coordinates,
    so an FMA-contracting build moves a tiny t by ~ulp(scene size), not ulp(t))."""
    diff = np.abs(t_a.astype(np.float64) - t_b.astype(np.float64))
    t_close = diff <= rel * np.maximum(np.abs(t_a), np.abs(t_b)).astype(np.float64) + abs_tol
    both_miss = (prim_a < 0) & (prim_b < 0)
    ok = both_miss | t_close
    return np.nonzero(~ok)[0], int(((prim_a != prim_b) & ok & ~both_miss).sum())


def tonemap8(img_sum, exposure, ray_count):
    out = np.zeros(img_sum.shape, np.uint8)
    po.port().oracle_tonemap(np.ascontiguousarray(img_sum, np.float32).reshape(-1), img_sum.shape[0] * img_sum.shape[1],
                             float(exposure), int(ray_count), out.reshape(-1))
    return out


def read_png_rgb8(path):
    """Minimal PNG reader (8-bit RGB, non-interlaced; all five filter types) -> uint8[H, W, 3]. Enough for the files
    stb_image_write and rt_write_png produce; keeps the tests free of imaging dependencies."""
    import struct
    import zlib
    data = open(path, "rb").read()
    assert data[:8] == b"\x89PNG\r\n\x1a\n"
    pos, idat, w = 8, b"", None
    while pos < len(data):
        n, kind = struct.unpack(">I4s", data[pos:pos + 8])
        body = data[pos + 8:pos + 8 + n]
        if kind == b"IHDR":
            w, h, depth, ctype, _, _, interlace = struct.unpack(">IIBBBBB", body)
            assert depth == 8 and ctype == 2 and interlace == 0
        elif kind == b"IDAT":
            idat += body
        pos += 12 + n
    raw = np.frombuffer(zlib.decompress(idat), np.uint8).reshape(h, 1 + 3 * w)
    out = np.zeros((h, 3 * w), np.uint8)
    prev = np.zeros(3 * w, np.int32)
    for y in range(h):
        f, line = int(raw[y, 0]), raw[y, 1:].astype(np.int32)
        if f == 0:
            cur = line
        elif f == 2:
            cur = (line + prev) & 255
        else:  # sub / average / paeth depend on the pixel to the left: sequential over x, vectorised over the 3 channels
            cur = np.zeros(3 * w, np.int32)
            for x in range(0, 3 * w, 3):
                a = cur[x - 3:x] if x else np.zeros(3, np.int32)
                b = prev[x:x + 3]
                c = prev[x - 3:x] if x else np.zeros(3, np.int32)
                if f == 1:
                    pred = a
                elif f == 3:
                    pred = (a + b) >> 1
                else:
                    p = a + b - c
                    pa, pb, pc = np.abs(p - a), np.abs(p - b), np.abs(p - c)
                    pred = np.where((pa <= pb) & (pa <= pc), a, np.where(pb <= pc, b, c))
                cur[x:x + 3] = (line[x:x + 3] + pred) & 255
        out[y] = cur
        prev = cur
    return out.reshape(h, w, 3)
