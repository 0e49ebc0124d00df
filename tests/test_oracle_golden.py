"""CPU tests: the C restatement (oracle/rt_oracle.c) against golden vectors produced by the REFERENCE's own code.

tests/golden/ref_vectors.npz is written by tests/golden/make_golden.py from oracle/_ref/libref_shim.so, i.e. the
reference's scene.cu / random.cuh functions compiled in place (precise host math). The reference ships no tests or
golden files of its own, so these vectors are the pin. Everything here is bit-exact: the restatement
performs the same IEEE operations in the same order and both sides use the same libm.

The scene arrays fed to the oracle come from the PRODUCT's host loader, which tests/test_host_scene.py pins
byte-for-byte to the reference loader, so this file needs neither /root/reference nor a GPU.
"""
import ctypes as C

import numpy as np
import pytest

import pyoracle as po
from util import ALL_SCENES, GOLDEN, load, oracle_arrays, oracle_closest, oracle_render, scene_rays

G = np.load(GOLDEN + "/ref_vectors.npz")
P = po.port()


def test_pcg_sequences_bit_exact():
    """xor_srand / xor_rand (random.cuh:13-30)."""
    for seed, want in zip(G["rng_seeds"], G["rng_sequences"]):
        got = np.zeros(32, np.uint32)
        P.oracle_xor_sequence(int(seed), 32, got)
        assert (got == want).all()


def test_random_floats_bit_exact():
    """random01 / random02 / random_radians / random_on_sphere (random.cuh:32-75), including the double-precision
    angle scale (Q6) and the u32 -> float rounding that can return exactly 1.0f (Q8)."""
    for seed, want in zip(G["rng_seeds"], G["rng_floats"]):
        got = np.zeros(6, np.float32)
        P.oracle_random_floats(int(seed), got)
        assert got.tobytes() == want.tobytes()
    got = np.zeros(6, np.float32)
    P.oracle_random_floats(0xFFFFFFFF, got)
    assert np.isfinite(got).all()


def test_gpu_seed_truncation():
    """64-bit seed products truncated to u32 (Q4): raytracing.cu:89,229,236, scene.cu:81."""
    for a, i in enumerate(G["seed_ray_index"]):
        for b, s in enumerate(G["seed_seed"]):
            assert P.oracle_gpu_gen_seed(int(i), int(s)) == G["seed_gen"][a, b]
            assert P.oracle_gpu_proc_seed(int(i), int(s)) == G["seed_proc"][a, b]


def test_equal_area_projection_bit_exact():
    """equal_area_project_sphere_to_square (scene.cu:284-318) incl. axis directions and the zero vector."""
    d = np.ascontiguousarray(G["project_dirs"])
    got = np.zeros((len(d), 2), np.float32)
    P.oracle_project(d, len(d), got)
    want = G["project_uv"]
    both_nan = np.isnan(got) & np.isnan(want)
    assert ((got == want) | both_nan).all()


def test_slab_test_known_answers():
    """ray_aabb_intersection (scene.cu:109-132): boundary rays, zero direction components, flat boxes, tmax cut-off."""
    for box, ray, tmax, hit, tmin in zip(G["aabb_boxes"], G["aabb_rays"], G["aabb_tmax"], G["aabb_hit"], G["aabb_tmin"]):
        t = C.c_float()
        got = P.oracle_ray_aabb(np.ascontiguousarray(box), np.ascontiguousarray(ray), float(tmax), C.byref(t))
        assert got == hit
        if hit:
            assert np.float32(t.value) == tmin


@pytest.mark.parametrize("name", ALL_SCENES)
def test_closest_hit_golden(name):
    """Scene::bvh_closest_hit_distance (scene.cu:134-241) in the reference's far-first order: bit-exact t and index.
    Near-first (the CUDA core's order) may only differ on exact-t ties (Q5)."""
    scene = load(name).set_image(32, 24, 25, 6, 1.0)
    A = oracle_arrays(scene)
    rays = scene_rays(A, 4000, seed=99)
    idx, t, ctr = oracle_closest(A, rays, near_first=False)
    assert t.tobytes() == G[f"hit_{name}_t"].tobytes()
    assert (idx == G[f"hit_{name}_idx"]).all()
    idx_n, t_n, ctr_n = oracle_closest(A, rays, near_first=True)
    assert t_n.tobytes() == t.tobytes()
    assert (idx_n != idx).sum() <= 4
    if len(A.bvh) > 1:
        assert ctr_n.internal_visits <= ctr.internal_visits  # near-first never visits more (Q5)


@pytest.mark.parametrize("name", ALL_SCENES)
def test_generate_and_process_golden(name):
    """generate_initial_rays (scene.cu:78-105) then three process_ray steps (scene.cu:320-487) with the GPU seed
    schedule, all four material classes + env-map lookup: bit-exact RayData records."""
    scene = load(name).set_image(32, 24, 25, 6, 1.0)
    A = oracle_arrays(scene)  # keeps the arrays alive: the struct holds raw pointers
    S = A.oracle_struct()
    m = 2048
    g = np.zeros((m, 12), np.float32)
    P.oracle_generate(C.byref(S), g.reshape(-1), 20, 0, m, 4)
    assert g.tobytes() == G[f"gen_{name}"].tobytes()
    for b in range(3):
        seeds = np.array([P.oracle_gpu_proc_seed(k, 4 * 20 + b) for k in range(m)], np.uint32)
        P.oracle_process(C.byref(S), g.reshape(-1), seeds, m)
        want = G[f"proc_{name}_b{b}"]
        same = (g.view(np.uint32) == want.view(np.uint32)) | (np.isnan(g) & np.isnan(want))
        assert same.all(), f"{name} bounce {b}: {np.count_nonzero(~same.all(axis=1))} records differ"


@pytest.mark.parametrize("name", ["spheres", "cornell_plus", "cornell"])
def test_sphere_leg_against_golden_process_records(name):
    """oracle_closest_hit_ex(with_spheres=1) = process_ray's sphere loop (scene.cu:338-372) + the BVH. The reference has no
    separate function for it, so it is pinned through the reference's own process_ray records in tests/golden: for every
    ray of bounce 0, 1 and 2 the golden next origin is origin + t * dir (scene.cu:397) bit for bit when the oracle reports
    a hit, and the golden ray is dead (transmitted = 0, origin unchanged) exactly when it reports a miss. spheres.scene
    holds the radius-40000 spheres at distance 4e5 (FP32 cancellation), cornell_plus rays that start ON its spheres."""
    scene = load(name).set_image(32, 24, 25, 6, 1.0)
    A = oracle_arrays(scene)
    m = 2048
    state = G[f"gen_{name}"]
    sphere_hits = 0
    for b in range(3):
        nxt = G[f"proc_{name}_b{b}"]
        live = ~(state[:, 6:9] == 0).all(axis=1)
        rays = np.ascontiguousarray(state[:, :6])
        idx, t, _ = oracle_closest(A, rays, near_first=False, with_spheres=True)
        idx_n, t_n, _ = oracle_closest(A, rays, near_first=True, with_spheres=True)
        assert t_n.tobytes() == t.tobytes() and (idx_n != idx).sum() <= 2
        hit = live & (idx >= 0)
        miss = live & (idx < 0)
        with np.errstate(over="ignore", invalid="ignore"):
            want_origin = (state[:, 0:3] + t[:, None] * state[:, 3:6]).astype(np.float32)
        assert (want_origin[hit].view(np.uint32) == nxt[hit, 0:3].view(np.uint32)).all(), f"{name} bounce {b}"
        assert (nxt[miss, 6:9] == 0).all() and (nxt[miss, 0:3].view(np.uint32) == state[miss, 0:3].view(np.uint32)).all()
        if len(A.spheres) == 0:
            idx0, t0, _ = oracle_closest(A, rays, near_first=False)
            assert t0.tobytes() == t.tobytes() and (idx0 == idx).all()
        sphere_hits += int((hit & (idx < len(A.spheres))).sum())
        state = nxt
    if len(A.spheres):
        assert sphere_hits > 50  # the golden rays are the first 2048 of a pass (top rows): few reach the spheres of cornell_plus


@pytest.mark.parametrize("name", ALL_SCENES)
@pytest.mark.parametrize("mode", [0, 1])
def test_render_loop_golden(name, mode):
    """The whole pass loop: cpu_raytrace (raytracing.cu:122-163, mode 0) and gpu_raytrace's no_sort seed schedule
    (raytracing.cu:222-254, mode 1); 25 spp = one full and one ragged pass. Sums and live-ray counts are exact."""
    scene = load(name).set_image(32, 24, 25, 6, 1.0)
    img, ctr = oracle_render(oracle_arrays(scene), seed_mode=mode)
    want = G[f"render_{name}_mode{mode}"].reshape(24, 32, 3)
    assert ctr.live_ray_bounces == int(G[f"render_{name}_mode{mode}_live"][0])
    same = (img.view(np.uint32) == want.view(np.uint32)) | (np.isnan(img) & np.isnan(want))
    assert same.all()


def test_render_pass_split_adds_up():
    """Passes are independent units (raytracing.cu:222-254): disjoint pass ranges sum to the whole frame (the multi-GPU
    spp split)."""
    scene = load("cornell_plus").set_image(24, 24, 70, 5, 1.0)  # passes of 20,20,20,10
    A = oracle_arrays(scene)
    whole, c = oracle_render(A, 1)
    parts = [oracle_render(A, 1, f, n) for f, n in ((0, 1), (1, 2), (3, 1))]
    total = sum(p[0].astype(np.float64) for p in parts)
    assert np.allclose(total, whole, rtol=1e-5, atol=1e-5 * whole.mean())
    assert sum(p[1].live_ray_bounces for p in parts) == c.live_ray_bounces


def test_render_select_subsets_add_up():
    """oracle_render_select: ray ranges of the frame's sequence and pixel windows keep every ray's own seeds, so disjoint
    subsets sum to the golden-pinned oracle_render frame, counters included (the checker of the product's ray-range split
    and of the 4K crop test)."""
    from util import oracle_render_select
    scene = load("cornell_plus").set_image(32, 24, 50, 5, 1.0)  # 3 passes: 20, 20, 10
    A = oracle_arrays(scene)
    whole, cw = oracle_render(A, 1)
    total = 32 * 24 * 50
    cuts = [0, 7, 32 * 24 * 20 - 1, 32 * 24 * 20 + 13, 32 * 24 * 45, total]
    acc, live = np.zeros_like(whole, dtype=np.float64), 0
    for a, b in zip(cuts[:-1], cuts[1:]):
        img, c = oracle_render_select(A, 1, ray_first=a, ray_count=b - a)
        acc += img
        live += c.live_ray_bounces
    assert live == cw.live_ray_bounces
    assert np.allclose(acc, whole, rtol=1e-5, atol=1e-6)
    win, c1 = oracle_render_select(A, 1, window=(5, 3, 16, 9))
    assert win[3:12, 5:21].tobytes() == whole[3:12, 5:21].tobytes()  # same rays, same per-pixel summation order
    outside = win.copy()
    outside[3:12, 5:21] = 0
    assert not outside.any() and 0 < c1.live_ray_bounces < cw.live_ray_bounces
    one, _ = oracle_render_select(A, 1, pass_first=1, pass_count=1)
    ref, _ = oracle_render(A, 1, 1, 1)
    assert one.tobytes() == ref.tobytes()
    # sample ranges of every pixel (the multi-GPU split): 50 samples cut at 0 | 7 | 20 | 33 | 50 add up, counters included;
    # samples [20, 40) are exactly pass 1
    from util import oracle_render_samples
    acc, live = np.zeros_like(whole, dtype=np.float64), 0
    for a, b in ((0, 7), (7, 20), (20, 33), (33, 50)):
        img, c = oracle_render_samples(A, a, b - a)
        acc += img
        live += c.live_ray_bounces
    assert live == cw.live_ray_bounces and np.allclose(acc, whole, rtol=1e-5, atol=1e-6)
    assert oracle_render_samples(A, 20, 20)[0].tobytes() == ref.tobytes()


def test_sort_key_degenerate_as_in_reference():
    """The reference key (scene.cu:44-60,480-485) collapses to at most 64 live values + 0xFFFFFFFF (Q1)."""
    scene = load("teapot").set_image(32, 24, 20, 4, 1.0)
    A = oracle_arrays(scene)
    S = A.oracle_struct()
    rays = scene_rays(A, 3000, seed=3)
    one = np.ones(3, np.float32)
    keys = {P.oracle_sort_key(C.byref(S), np.ascontiguousarray(r[:3]), np.ascontiguousarray(r[3:]), one) for r in rays}
    assert len(keys) <= 64
    dead = P.oracle_sort_key(C.byref(S), np.zeros(3, np.float32), np.array([0, 0, 1], np.float32), np.zeros(3, np.float32))
    assert dead == 0xFFFFFFFF


def test_bloom_and_tonemap_known_answers():
    """raytracing.cu:21-74 and :286-303 on a hand-checkable image."""
    w, h, rc = 16, 12, 10
    img = np.zeros((h, w, 3), np.float32)
    img[6, 8] = 100.0  # luminance 100 > 0.7 * 10
    img[0, 0] = 5.0    # below the threshold: contributes nothing to the bloom
    out = img.copy()
    P.oracle_bloom(out.reshape(-1), w, h, rc)
    # 11-tap H then V box blur; the divisor is the number of taps INSIDE the image (raytracing.cu:40-47,58-65)
    assert np.isclose(out[6, 8, 0], 100.0 + 100.0 / 121.0, rtol=1e-6)  # rows 1..11, columns 3..13: all 11 taps inside
    assert np.isclose(out[1, 3, 0], 100.0 / 9.0 / 7.0, rtol=1e-6)     # columns 0..8 (9 taps) and rows 0..6 (7 taps) inside
    assert out[0, 0, 0] == np.float32(5.0) or np.isclose(out[0, 0, 0], 5.0)
    assert out[6, 14, 0] == 0.0  # 6 columns away: outside the radius-5 box
    rgb8 = np.zeros((h, w, 3), np.uint8)
    P.oracle_tonemap(img.reshape(-1), w * h, 1.0, rc, rgb8.reshape(-1))
    c = np.float32(100.0 / rc)
    assert rgb8[6, 8, 0] == int(np.sqrt(c / (c + 1)) * np.float32(255.999))
    assert rgb8[3, 3, 0] == 0
