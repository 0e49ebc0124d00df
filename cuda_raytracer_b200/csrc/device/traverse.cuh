// Closest-hit query: sphere loop + stack traversal of the packed two-child BVH with slab and Moller-Trumbore tests.
// Semantics follow Scene::process_ray's sphere loop (scene.cu:338-372) and Scene::bvh_closest_hit_distance
// (scene.cu:134-241) except for the visiting ORDER, which is near-child-first here (the reference pops the far
// child first); results differ only on exact-t ties.
#pragma once

#include "device_scene.cuh"
#include "rng_math.cuh"

namespace rtd {

// Per-thread traversal stack in shared memory, entry-major so that the 32 lanes of a warp hit 32 different banks.
template <int BLOCK>
struct SharedStack {
    uint32_t ref[kStackDepth][BLOCK];
    float dist[kStackDepth][BLOCK];
};

struct TraceCounters {
    uint32_t internal_visits = 0;
    uint32_t triangle_tests = 0;
};

// Slab test with the reference's boundary-safe semantics (Tavian Barnes' form, scene.cu:109-132: tmin starts at 0,
// tmax at `limit`, a NaN from 0 * inf leaves the running interval untouched) restated so that it costs 11 ALU-pipe
// instructions per box instead of 19: the reference evaluates min(max(t1,tmin), max(t2,tmin)) per axis; for non-NaN
// inputs that is max(min(t1,t2), tmin), and min(t1,t2) is t of the bound the ray meets first, which the SIGN of the
// direction picks ahead of time (rounding is monotonic, so (lo-o)*inv <= (hi-o)*inv whenever inv >= 0). fmaxf/fminf
// drop a NaN operand exactly as the reference's chain does, and the three-operand forms compile to FMNMX3.
// Arithmetic per bound is (bound - o) * inv as the reference writes it; there is no a*b+c shape to contract.
// RT_SIGN_MASK (default 1): the signs are kept as all-ones / all-zeros words and the near / far bound of an axis is a
// bitwise select (one LOP3 each, same count as the FSELs) — predicates would have to be rebuilt from registers at every
// visited record (3 LOP3 per visit) because they do not survive the loop.
#ifndef RT_SIGN_MASK
#define RT_SIGN_MASK 1
#endif
#if RT_BOX_CH
struct RaySigns {
    float x, y, z;  // |1 / d|: the centre / half-extent form needs no signs
};
__device__ __forceinline__ RaySigns ray_signs(float3 inv) { return RaySigns{fabsf(inv.x), fabsf(inv.y), fabsf(inv.z)}; }
#elif RT_SIGN_MASK
struct RaySigns {
    uint32_t x, y, z;  // 0xFFFFFFFF when the direction component is negative (including -0: 1/-0 = -inf), else 0
};
__device__ __forceinline__ RaySigns ray_signs(float3 inv)
{
    return RaySigns{(uint32_t)(__float_as_int(inv.x) >> 31), (uint32_t)(__float_as_int(inv.y) >> 31), (uint32_t)(__float_as_int(inv.z) >> 31)};
}
__device__ __forceinline__ float pick(uint32_t m, float a, float b)  // m ? a : b
{
    return __uint_as_float((__float_as_uint(a) & m) | (__float_as_uint(b) & ~m));
}
#else
struct RaySigns {
    bool x, y, z;  // direction component negative (including -0: 1/-0 = -inf)
};
__device__ __forceinline__ RaySigns ray_signs(float3 inv) { return RaySigns{signbit(inv.x) != 0, signbit(inv.y) != 0, signbit(inv.z) != 0}; }
__device__ __forceinline__ float pick(bool m, float a, float b) { return m ? a : b; }
#endif

// RT_SLAB_FMA (default 1): t of a bound is taken as fma(bound, inv, -(o * inv)) instead of (bound - o) * inv: one FFMA
// instead of FADD + FMUL per bound (12 instructions fewer per visited record). The two forms differ by rounding only
// (absolute error ~ulp(o * inv) instead of a relative one), near and far t of a flat box are still computed identically,
// and for an axis-parallel ray (inv = +-inf) the FMA form yields NaN = "no constraint on this axis" where the exact
// form yields +-inf: the walk then visits a superset of the records, and hits come from the exact triangle test.
// `so` is what slab_origin() returned for the ray.
#ifndef RT_SLAB_FMA
#define RT_SLAB_FMA 1
#endif
#if RT_BOX_CH && !RT_SLAB_FMA
#error "RT_BOX_CH needs RT_SLAB_FMA"
#endif
__device__ __forceinline__ float3 slab_origin(float3 o, float3 inv)
{
#if RT_SLAB_FMA
    return f3(-(o.x * inv.x), -(o.y * inv.y), -(o.z * inv.z));
#else
    (void)inv;
    return o;
#endif
}
__device__ __forceinline__ float slab_t(float bound, float so, float inv)
{
#if RT_SLAB_FMA
    return __fmaf_rn(bound, inv, so);
#else
    return (bound - so) * inv;
#endif
}
#if RT_BOX_CH
// the six floats are {centre.xyz, half-extent.xyz}; a flat box (h = 0) gives t_near == t_far exactly, an axis-parallel ray
// (inv = +-inf) gives NaN on that axis = no constraint (a superset of the exact form's visits, as with RT_SLAB_FMA)
__device__ __forceinline__ bool slab(float cx, float cy, float cz, float hx, float hy, float hz, float3 so, float3 inv,
                                     RaySigns ainv, float limit, float &t_enter)
{
    const float tx = __fmaf_rn(cx, inv.x, so.x), ty = __fmaf_rn(cy, inv.y, so.y), tz = __fmaf_rn(cz, inv.z, so.z);
    const float tn = fmaxf(fmaxf(__fmaf_rn(-hx, ainv.x, tx), __fmaf_rn(-hy, ainv.y, ty)), fmaxf(__fmaf_rn(-hz, ainv.z, tz), 0.0f));
    const float tf = fminf(fminf(__fmaf_rn(hx, ainv.x, tx), __fmaf_rn(hy, ainv.y, ty)), fminf(__fmaf_rn(hz, ainv.z, tz), limit));
    t_enter = tn;
    return tn <= tf;
}
#else
__device__ __forceinline__ bool slab(float lox, float loy, float loz, float hix, float hiy, float hiz, float3 so, float3 inv,
                                     RaySigns sg, float limit, float &t_enter)
{
    const float nx = pick(sg.x, hix, lox), fx = pick(sg.x, lox, hix);
    const float ny = pick(sg.y, hiy, loy), fy = pick(sg.y, loy, hiy);
    const float nz = pick(sg.z, hiz, loz), fz = pick(sg.z, loz, hiz);
    const float tn = fmaxf(fmaxf(slab_t(nx, so.x, inv.x), slab_t(ny, so.y, inv.y)), fmaxf(slab_t(nz, so.z, inv.z), 0.0f));
    const float tf = fminf(fminf(slab_t(fx, so.x, inv.x), slab_t(fy, so.y, inv.y)), fminf(slab_t(fz, so.z, inv.z), limit));
    t_enter = tn;
    return tn <= tf;
}
#endif

// scene.cu:338-372. Returns through closest/prim; runs before the BVH and seeds `closest`.
// Contraction is forbidden (explicit _rn operations in the host's evaluation order, IEEE square root): the discriminant
// b^2 - (|c - o|^2 - r^2) cancels catastrophically for the far, huge spheres of spheres.scene (radius 4e4 at 4e5: both terms
// ~1.6e9 with an ulp of 128) and for every ray that STARTS on such a sphere, so a fused multiply-add moves a grazing root by
// whole units. Written this way the sphere distance is bit-identical to the reference's host code (the oracle's sphere leg,
// tests/test_gpu_parity.py::test_closest_hit_with_spheres); the loop runs over 0-4 spheres and costs nothing measurable.
__device__ __forceinline__ void intersect_spheres(const DeviceScene &sc, float3 o, float3 d, float &closest, int &prim)
{
    for (int i = 0; i < sc.sphere_count; i++) {
        const float4 s = __ldg(sc.spheres + i);
        const float3 offset = f3(__fsub_rn(s.x, o.x), __fsub_rn(s.y, o.y), __fsub_rn(s.z, o.z));
        const float minus_half_b = dot_nofma(offset, d);
        const float quarter_c = __fsub_rn(dot_nofma(offset, offset), __fmul_rn(s.w, s.w));
        const float quarter_disc = __fsub_rn(__fmul_rn(minus_half_b, minus_half_b), quarter_c);
        if (quarter_disc < 0.0f) continue;
        const float half_root = __fsqrt_rn(quarter_disc);
        float t = __fsub_rn(minus_half_b, half_root);
        if (t < closest && t > kTMin) { closest = t; prim = i; continue; }
        t = __fadd_rn(minus_half_b, half_root);
        if (t < closest && t > kTMin) { closest = t; prim = i; continue; }
    }
}

// ---- quantised boxes (kNodesQ16 / kNodesW4, layout in device_scene.cuh) ------------------------------------------------
// Per ray: t(bound) = f * a + b with f = 1 + q / 32768 the float a byte permute builds from the stored 16 bits,
// a = q_scale / d and b = (q_bias - o) / d; `near` / `far` are the permute selectors that take the low or the high half
// of a {lo | hi << 16} word according to the sign of the direction component (0x7104: bytes {K0, w0, w1, K3} of
// {w, K = 0x3F000000}; 0x7324: {K0, w2, w3, K3}). Rounding: a and b carry a few ulp of (extent / d), i.e. ~1e-7 of the scene
// extent in coordinates = 0.004 grid cells; the packer adds a whole cell of margin on every side, so the test is
// conservative for origins up to ~100 scene extents away from the tree. An axis-parallel ray (1 / d = +-inf) yields
// NaN = "no constraint on this axis", as with RT_SLAB_FMA. Hits always come from the exact triangle test.
struct QuantRay {
    float ax, ay, az, bx, by, bz;
    uint32_t nx, ny, nz, fx, fy, fz;
};
__device__ __forceinline__ QuantRay quant_ray(const DeviceScene &sc, float3 o, float3 inv)
{
    QuantRay q;
    q.ax = sc.q_scale[0] * inv.x; q.ay = sc.q_scale[1] * inv.y; q.az = sc.q_scale[2] * inv.z;
    q.bx = (sc.q_bias[0] - o.x) * inv.x; q.by = (sc.q_bias[1] - o.y) * inv.y; q.bz = (sc.q_bias[2] - o.z) * inv.z;
    q.nx = __float_as_int(inv.x) < 0 ? 0x7324u : 0x7104u; q.fx = q.nx ^ 0x0220u;
    q.ny = __float_as_int(inv.y) < 0 ? 0x7324u : 0x7104u; q.fy = q.ny ^ 0x0220u;
    q.nz = __float_as_int(inv.z) < 0 ? 0x7324u : 0x7104u; q.fz = q.nz ^ 0x0220u;
    return q;
}
// (prmt.b32 directly: __byte_perm masks its selector with 0x7777 first, one LOP3 per selector and visit)
__device__ __forceinline__ float quant_f(uint32_t w, uint32_t sel)
{
    float f;
    asm("prmt.b32 %0, %1, 0x3F000000, %2;" : "=f"(f) : "r"(w), "r"(sel));
    return f;
}
// one child: wx / wy / wz = its {lo | hi << 16} words
__device__ __forceinline__ bool slab_q(uint32_t wx, uint32_t wy, uint32_t wz, const QuantRay &q, float limit, float &t_enter)
{
    const float tn = fmaxf(fmaxf(__fmaf_rn(quant_f(wx, q.nx), q.ax, q.bx), __fmaf_rn(quant_f(wy, q.ny), q.ay, q.by)),
                           fmaxf(__fmaf_rn(quant_f(wz, q.nz), q.az, q.bz), 0.0f));
    const float tf = fminf(fminf(__fmaf_rn(quant_f(wx, q.fx), q.ax, q.bx), __fmaf_rn(quant_f(wy, q.fy), q.ay, q.by)),
                           fminf(__fmaf_rn(quant_f(wz, q.fz), q.az, q.bz), limit));
    t_enter = tn;
    return tn <= tf;
}
struct __align__(32) Uint8 {
    uint4 lo, hi;
};
__device__ __forceinline__ Uint8 ldg256u(const void *p)
{
    Uint8 r;
    asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r.lo.x), "=r"(r.lo.y), "=r"(r.lo.z), "=r"(r.lo.w), "=r"(r.hi.x), "=r"(r.hi.y), "=r"(r.hi.z), "=r"(r.hi.w)
                 : "l"(p));
    return r;
}

// RT_FAST_RCP (default 1): 1 / a of the triangle test through MUFU.RCP alone (rcp.approx, <= 1 ulp, subnormal inputs
// handled) instead of the IEEE division sequence (range check + 3 fix-up instructions + a slow-path call): u, v and t
// move by at most an ulp, far inside the 1e-5 relative bar on t; the reference's own GPU build divides with
// --use_fast_math.
#ifndef RT_FAST_RCP
#define RT_FAST_RCP 1
#endif
// RT_TRI_FTZ (default 1): rcp.approx.ftz is ONE MUFU.RCP; without .ftz ptxas wraps it in a range scaling (2 FSETP + 2 FSEL + 2
// FMUL) for subnormal and huge operands. A subnormal determinant (|a| < 1.2e-38: a triangle ~1e-19 across) is then treated as
// the reference treats a == 0 (no hit); NaN still falls through the compares as in the reference.
#ifndef RT_TRI_FTZ
#define RT_TRI_FTZ 1
#endif
__device__ __forceinline__ float tri_rcp(float a)
{
#if RT_FAST_RCP && RT_TRI_FTZ
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(a));
    return r;
#elif RT_FAST_RCP
    float r;
    asm("rcp.approx.f32 %0, %1;" : "=f"(r) : "f"(a));
    return r;
#else
    return 1.0f / a;
#endif
}
__device__ __forceinline__ bool tri_degenerate(float a)
{
#if RT_FAST_RCP && RT_TRI_FTZ
    return fabsf(a) < 1.17549435e-38f;
#else
    return a == 0.0f;
#endif
}

// Moller-Trumbore on the precomputed-edge record (scene.cu:164-194). Updates closest/prim on an accepted hit.
__device__ __forceinline__ void intersect_triangle(const float4 *__restrict__ tris, uint32_t tri, int prim_base, float3 o, float3 d,
                                                   float &closest, int &prim)
{
    const float4 *t4 = tris + 4 * (size_t)tri;
    const Float8 t8 = ldg256(t4);
    const float e2z = __ldg(reinterpret_cast<const float *>(t4 + 2));
    float3 p1 = f3(t8.lo.x, t8.lo.y, t8.lo.z), e1 = f3(t8.lo.w, t8.hi.x, t8.hi.y), e2 = f3(t8.hi.z, t8.hi.w, e2z);
    float3 h = cross(d, e2);
    float a = dot(h, e1);
    if (tri_degenerate(a)) return;
    float inv_a = tri_rcp(a);
    float3 s = o - p1;
    float u = dot(s, h) * inv_a;
    if (u < 0.0f || u > 1.0f) return;
    float3 q = cross(s, e1);
    float v = dot(d, q) * inv_a;
    if (v < 0.0f || u + v > 1.0f) return;
    float t = dot(e2, q) * inv_a;
    if (t <= kTMin || t >= closest) return;  // NaN passes, exactly as the reference's two compares do
    closest = t;
    prim = prim_base + (int)tri;
}

// The same test as a pure function: t of an accepted hit (kTMin < t < limit), else kNoTriHit.
constexpr float kNoTriHit = 3e38f;
__device__ __forceinline__ float triangle_t(const float4 *__restrict__ tris, uint32_t tri, float3 o, float3 d, float limit)
{
    const float4 *t4 = tris + 4 * (size_t)tri;
    const Float8 t8 = ldg256(t4);
    const float e2z = __ldg(reinterpret_cast<const float *>(t4 + 2));
    float3 p1 = f3(t8.lo.x, t8.lo.y, t8.lo.z), e1 = f3(t8.lo.w, t8.hi.x, t8.hi.y), e2 = f3(t8.hi.z, t8.hi.w, e2z);
    float3 h = cross(d, e2);
    float a = dot(h, e1);
    float inv_a = tri_rcp(a);
    float3 s = o - p1;
    float u = dot(s, h) * inv_a;
    float3 q = cross(s, e1);
    float v = dot(d, q) * inv_a;
    float t = dot(e2, q) * inv_a;
    const bool hit = !tri_degenerate(a) && u >= 0.0f && u <= 1.0f && v >= 0.0f && u + v <= 1.0f && t > kTMin && t < limit;
    return hit ? t : kNoTriHit;
}

// RT_LEAF_SHARE (A/B, default 0 — measured 20 % SLOWER on lamp 4K, 15 % on teapot, 28 % on cornell_plus, profiles/r2b_tracer.md:
// the sequential test leaves early at the first failed barycentric check, ~3.9 warp instructions per test and ray as profiled,
// while the dealt-out test runs to the end and pays ~70 instructions of scan / shuffle / segmented-minimum overhead per round):
// the leaf phase of a warp as ONE collective step. After the descent loop ~14 of the 32 lanes hold a
// leaf of 1-4 triangles; tested lane by lane (loop over each lane's count) that is 3-4 rounds of a ~50-instruction test at
// 7-14 active lanes, each round waiting for its own triangle loads: 30 % of the tracer's warp instructions at 12 lanes and its
// largest block of stall samples on the incoherent bounces (profiles/r2a). Here the warp's (ray, triangle) pairs are numbered
// with a prefix sum over the lanes' counts and dealt out to all 32 lanes, 32 pairs per round: a lane fetches the owner's ray
// with shuffles, tests ONE triangle, and a segmented minimum over the pair's neighbours hands the nearest accepted hit back to
// the owner. Ties go to the lower triangle index and every test sees the owner's distance limit from before the leaf, so the
// result equals the sequential loop's (which accepts t < closest in index order). A NaN distance is dropped here where the
// sequential form (and the reference) would accept it: only inputs with NaN / infinite coordinates produce one.
#ifndef RT_LEAF_SHARE
#define RT_LEAF_SHARE 0
#endif
template <bool STATS>
__device__ __forceinline__ void leaf_phase_shared(const DeviceScene &sc, bool has_leaf, uint32_t cur, float3 o, float3 d, float &closest,
                                                  int &prim, uint32_t *table, TraceCounters &ctr)
{
    constexpr uint32_t FULL = 0xffffffffu;
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t c = has_leaf ? ((cur >> kLeafCountShift) & 15u) : 0u;
    const uint32_t first = cur & kLeafStartMask;
    if (STATS) ctr.triangle_tests += c;
    uint32_t incl = c;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
        const uint32_t t = __shfl_up_sync(FULL, incl, off);
        if (lane >= (uint32_t)off) incl += t;
    }
    const uint32_t total = __shfl_sync(FULL, incl, 31);
    const uint32_t s = incl - c;  // this owner's first pair
    float best_t = closest;
    uint32_t best_k = 0xffffffffu;
    for (uint32_t base = 0; base < total; base += 32u) {
        const uint32_t lo = max(s, base), hi = min(s + c, base + 32u);
        for (uint32_t x = lo; x < hi; x++) table[x - base] = lane | ((x - s) << 8);  // pair -> (owner lane, triangle of its leaf)
        __syncwarp();
        const bool valid = base + lane < total;
        const uint32_t e = valid ? table[lane] : lane;
        const uint32_t own = e & 31u;
        uint32_t kk = e >> 8;
        const float3 ro = f3(__shfl_sync(FULL, o.x, own), __shfl_sync(FULL, o.y, own), __shfl_sync(FULL, o.z, own));
        const float3 rd = f3(__shfl_sync(FULL, d.x, own), __shfl_sync(FULL, d.y, own), __shfl_sync(FULL, d.z, own));
        const float limit = __shfl_sync(FULL, closest, own);
        const uint32_t tri = __shfl_sync(FULL, first, own) + kk;
        float t = kNoTriHit;
        if (valid) t = triangle_t(sc.tris, tri, ro, rd, limit);
        // segmented minimum: the pairs of one owner sit in consecutive lanes (at most 15), lower lane = lower triangle index
#pragma unroll
        for (int off = 1; off < 16; off <<= 1) {
            const float t2 = __shfl_down_sync(FULL, t, off);
            const uint32_t k2 = __shfl_down_sync(FULL, kk, off), o2 = __shfl_down_sync(FULL, own, off);
            if (lane + (uint32_t)off < 32u && o2 == own && t2 < t) { t = t2; kk = k2; }
        }
        const uint32_t head = hi > lo ? lo - base : 0u;  // the owner's first pair of this round holds its minimum
        const float rt = __shfl_sync(FULL, t, head);
        const uint32_t rk = __shfl_sync(FULL, kk, head);
        if (hi > lo && rt < best_t) { best_t = rt; best_k = rk; }
        __syncwarp();  // the table is rewritten by the next round
    }
    if (best_k != 0xffffffffu) {
        closest = best_t;
        prim = sc.sphere_count + (int)(first + best_k);
    }
}

template <int BLOCK, bool STATS>
__device__ __forceinline__ void traverse_bvh(const DeviceScene &sc, float3 o, float3 d, float &closest, int &prim,
                                             SharedStack<BLOCK> &stack, TraceCounters &ctr)
{
    const float3 inv = f3(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);
    const RaySigns sg = ray_signs(inv);
    const float3 so = slab_origin(o, inv);
    const float4 *__restrict__ nodes = sc.nodes;
    const int lane = threadIdx.x;
    uint32_t cur = 0;  // record 0 is the root
    int sp = 0;
    for (;;) {
        if (!(cur & kLeafBit)) {
            const float4 *n = nodes + 4 * (size_t)cur;
            const Float8 na = ldg256(n), nb = ldg256(n + 2);
            const uint2 q3 = make_uint2(__float_as_uint(na.hi.z), __float_as_uint(nb.hi.z));
            if (STATS) ctr.internal_visits++;
            float d0, d1;
            bool h0 = slab(na.lo.x, na.lo.y, na.lo.z, na.lo.w, na.hi.x, na.hi.y, so, inv, sg, closest, d0);
            bool h1 = slab(nb.lo.x, nb.lo.y, nb.lo.z, nb.lo.w, nb.hi.x, nb.hi.y, so, inv, sg, closest, d1);
            if (h0 && h1) {
                bool zero_near = d0 < d1;
                uint32_t far_ref = zero_near ? q3.y : q3.x;
                stack.ref[sp][lane] = far_ref;
                stack.dist[sp][lane] = zero_near ? d1 : d0;
                sp++;
                cur = zero_near ? q3.x : q3.y;
                continue;
            }
            if (h0) { cur = q3.x; continue; }
            if (h1) { cur = q3.y; continue; }
        } else {
            uint32_t first = cur & kLeafStartMask;
            uint32_t count = (cur >> kLeafCountShift) & 15u;
            if (STATS) ctr.triangle_tests += count;
            for (uint32_t k = 0; k < count; k++) intersect_triangle(sc.tris, first + k, sc.sphere_count, o, d, closest, prim);
        }
        // pop, culling entries that can no longer beat the current hit (scene.cu:150-153)
        float entry;
        do {
            if (sp == 0) return;
            sp--;
            cur = stack.ref[sp][lane];
            entry = stack.dist[sp][lane];
        } while (entry >= closest);
    }
}

// ---- production tracer: persistent warps with per-lane refill ------------------------------------------------------
//
// The per-thread loop above keeps a whole warp busy until its LONGEST ray ends; on incoherent bounces ncu showed
// 5-9 of 32 lanes active per issued instruction (profiles/r1_bounce_fused_lamp1k.md). Here every lane owns one ray
// at a time and the warp keeps walking while enough lanes are busy; when fewer than kRefillBelow lanes still have
// a ray, the warp converges, finished lanes take the next rays from the warp's batch (one global atomic per
// kBatch rays) and the walk resumes ("persistent while-while with dynamic fetch", Aila & Laine 2009, restated for
// independent thread scheduling: every vote is an explicit *_sync at a point where all 32 lanes arrive).
//
// IO policy:  bool load(uint32_t i, float3 &o, float3 &d, float &closest, int &prim)   false = nothing to trace
//             void store(uint32_t i, float closest, int prim)
//             void begin(uint32_t i, Stack &stack, int &sp, uint32_t &cur)           where the walk of ray i starts
constexpr int kTraceBlock = 128;            // threads per CTA of the tracer kernels
constexpr uint32_t kDoneRef = 0xFFFFFFFFu;  // never a valid leaf ref: start would exceed the 2^27 triangle limit
constexpr int kBatch = 512;                 // default: rays a warp claims per global atomic (re-swept in r1m: 192 -> 512 = +0.6-1 %)
constexpr int kRefillBelow = 16;           // default: refill when fewer lanes than this still hold a ray
constexpr int kNodeMin = 12;                // default for TraceTuning::node_min
#ifndef RT_TUNE_RUNTIME
#define RT_TUNE_RUNTIME 0
#endif
struct TraceTuning {
    uint32_t batch;         // rays per claim; the kernel shrinks it when the bounce has fewer rays than batch x warps
    uint32_t refill_below;
    uint32_t node_min;      // leave the descent loop when fewer lanes than this are still descending
};

// RT_BRANCHFREE (A/B, default 0): child selection after the two box tests as one straight-line sequence instead of four branch
// bodies. Measured -0.7 % (lamp 4K), -1.1 % (teapot), -0.7 % (cornell_plus): the bodies are short and ptxas already predicates
// most of them; the straight-line form makes every lane pay for the push (profiles/r2b_tracer.md).
#ifndef RT_BRANCHFREE
#define RT_BRANCHFREE 0
#endif
#ifndef RT_STACK_TOP
#define RT_STACK_TOP 8
#endif
constexpr int kStackTop = RT_STACK_TOP;  // entries per thread kept in shared memory; deeper ones go to local memory

// Traversal stack: the first kStackTop entries of every thread live in shared memory ([entry][thread]: the 32 lanes of
// a warp hit 32 banks whatever their individual depths are, one data-pipe wavefront per access), the rest in a
// thread-local array. A purely local stack scatters: lanes at different depths touch different 128-byte lines, and
// ncu showed those accesses as 28 % of the L1 data-pipe load of the tracer (profiles/r1c). 8 entries x 128 threads x
// 8 B = 8 KB per CTA leaves the L1 to the BVH, unlike a full-depth shared stack (measured in round 1: -5 %, removed).
template <int BLOCK, int SLOTS = kStackSlots>
struct HybridStack {
    // shared-window address of smem[0][threadIdx.x], taken once: through a generic pointer the compiler re-derives the
    // window base (S2R SR_CgaCtaId + 2 integer ops) at every push and pop
    uint32_t top;
    uint2 deep[SLOTS - kStackTop > 0 ? SLOTS - kStackTop : 1];
    __device__ __forceinline__ explicit HybridStack(uint2 *smem) : top((uint32_t)__cvta_generic_to_shared(smem + threadIdx.x))
    {
        asm volatile("" : "+r"(top));  // opaque: otherwise ptxas may rematerialise the address (S2UR + S2R + 3 ALU ops) at every visit
    }
    __device__ __forceinline__ void push(int sp, uint32_t r, float d)
    {
        if (sp < kStackTop)
            asm volatile("st.shared.v2.u32 [%0], {%1, %2};" ::"r"(top + (uint32_t)sp * (BLOCK * 8u)), "r"(r), "r"(__float_as_uint(d)));
        else
            deep[sp - kStackTop] = make_uint2(r, __float_as_uint(d));
    }
    __device__ __forceinline__ void get(int sp, uint32_t &r, float &d) const
    {
        uint2 v;
        if (sp < kStackTop)
            asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(top + (uint32_t)sp * (BLOCK * 8u)));
        else
            v = deep[sp - kStackTop];
        r = v.x;
        d = __uint_as_float(v.y);
    }
};

template <bool STATS, int NODES, class Stack, class IO>
__device__ __forceinline__ void trace_persistent(const DeviceScene &sc, uint32_t n, uint32_t *cursor, Stack &stack, IO &io,
                                                 TraceCounters &ctr, TraceTuning tune)
{
    constexpr uint32_t FULL = 0xffffffffu;
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t lt = (1u << lane) - 1u;
    const float4 *__restrict__ nodes = sc.nodes;

    // Small bounces (late bounces hold ~0.2 % of the pass) are latency-bound: spread them over every resident warp
    // instead of queueing `batch` rays behind a few warps.
    const uint32_t warps = gridDim.x * (blockDim.x >> 5);
    const uint32_t batch = max(4u, min(tune.batch, (n + warps - 1) / warps));
#if RT_TUNE_RUNTIME  // sweeps: thresholds from RT_REFILL / RT_NODEMIN (one LDCU per visited record)
    const uint32_t refill_below = tune.refill_below;
    const uint32_t node_min = tune.node_min;
#else
    constexpr uint32_t refill_below = kRefillBelow, node_min = kNodeMin;
#endif
#if RT_LEAF_SHARE
    __shared__ uint32_t leaf_table[kTraceBlock];  // 32 entries per warp: (owner lane, triangle) of the pairs of one leaf round
#endif
    uint32_t pool_next = 0, pool_end = 0;  // warp-uniform: unclaimed part of the current batch
    bool exhausted = false;                // warp-uniform: the global cursor ran past n
    bool busy = false;
    uint32_t slot = 0, cur = kDoneRef;
    int sp = 0, prim = -1;
    float closest = kMissT;
    float3 o = f3(0, 0, 0), d = f3(0, 0, 1), inv = f3(0, 0, 1), so = f3(0, 0, 0);
    RaySigns sg = ray_signs(inv);
    QuantRay qr = quant_ray(sc, o, inv);
    const uint4 *__restrict__ qnodes = sc.qnodes;

    for (;;) {
        // ---- refill: all 32 lanes are here ----
        uint32_t need = __ballot_sync(FULL, !busy);
        while (need) {
            if (pool_next == pool_end) {
                if (exhausted) break;
                uint32_t b = 0;
                if (lane == 0) b = atomicAdd(cursor, batch);
                b = __shfl_sync(FULL, b, 0);
                if (b >= n) { exhausted = true; break; }
                pool_next = b;
                pool_end = min(b + batch, n);
            }
            const uint32_t avail = pool_end - pool_next;
            const uint32_t rank = __popc(need & lt);
            if (!busy && rank < avail) {
                slot = pool_next + rank;
                closest = kMissT;
                prim = -1;
                if (io.load(slot, o, d, closest, prim)) {
                    inv = f3(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);
                    if (NODES == kNodesF32) {
                        sg = ray_signs(inv);
                        so = slab_origin(o, inv);
                    } else {
                        qr = quant_ray(sc, o, inv);
                    }
                    io.begin(slot, stack, sp, cur);  // record 0 (the root), or the tile's entry records for primary rays
                    busy = true;
                }
            }
            pool_next += min((uint32_t)__popc(need), avail);
            need = __ballot_sync(FULL, !busy);
            if (!need) break;
            // lanes whose load() declined stay idle this round; lanes beyond `avail` retry on the next batch
            if (pool_next != pool_end) break;
        }
        if (!__any_sync(FULL, busy)) {
            if (exhausted && pool_next == pool_end) break;
            continue;
        }

#if RT_LEAF_SHARE
        // ---- walk in warp-uniform rounds: descent (lane by lane) -> collective leaf phase -> pop / retire ----
        for (;;) {
            if (busy) {
            // descend through internal records until this lane holds a leaf (or runs dry)
            while (!(cur & kLeafBit)) {
                if (STATS) ctr.internal_visits++;
                if (NODES == kNodesW4) {
                    // 4-wide node: two 32-byte halves {children 0,1} {children 2,3}; hit children sorted by entry distance
                    // (5 compare-exchanges), the nearest becomes the current reference, the others go on the stack far first
                    const uint4 *nd = qnodes + 4 * (size_t)cur;
                    const Uint8 na = ldg256u(nd), nb = ldg256u(nd + 2);
                    constexpr float kNoHit = 3e38f;
                    float e0, e1, e2, e3;
                    const bool h0 = slab_q(na.lo.x, na.lo.y, na.lo.z, qr, closest, e0), h1 = slab_q(na.lo.w, na.hi.x, na.hi.y, qr, closest, e1);
                    const bool h2 = slab_q(nb.lo.x, nb.lo.y, nb.lo.z, qr, closest, e2), h3 = slab_q(nb.lo.w, nb.hi.x, nb.hi.y, qr, closest, e3);
                    float k0 = h0 ? e0 : kNoHit, k1 = h1 ? e1 : kNoHit, k2 = h2 ? e2 : kNoHit, k3 = h3 ? e3 : kNoHit;
                    uint32_t r0 = na.hi.z, r1 = na.hi.w, r2 = nb.hi.z, r3 = nb.hi.w;
                    const int hits = (int)h0 + (int)h1 + (int)h2 + (int)h3;
#define RT_CSWAP(ka, ra, kb, rb) { const bool sw = kb < ka; const float tk = sw ? ka : kb; ka = sw ? kb : ka; kb = tk; const uint32_t tr = sw ? ra : rb; ra = sw ? rb : ra; rb = tr; }
                    RT_CSWAP(k0, r0, k1, r1) RT_CSWAP(k2, r2, k3, r3) RT_CSWAP(k0, r0, k2, r2) RT_CSWAP(k1, r1, k3, r3) RT_CSWAP(k1, r1, k2, r2)
#undef RT_CSWAP
                    if (hits > 3) { stack.push(sp, r3, k3); sp++; }
                    if (hits > 2) { stack.push(sp, r2, k2); sp++; }
                    if (hits > 1) { stack.push(sp, r1, k1); sp++; }
                    if (hits > 0) {
                        cur = r0;
                    } else {
                        float entry;
                        do {
                            if (sp == 0) { cur = kDoneRef; break; }
                            sp--;
                            stack.get(sp, cur, entry);
                        } while (entry >= closest);
                    }
                } else {
                    float d0, d1;
                    bool h0, h1;
                    uint2 q3;
                    if (NODES == kNodesQ16) {  // one 32-byte record: 12 bounds on the 15-bit grid + the two child refs
                        const Uint8 nq = ldg256u(qnodes + 2 * (size_t)cur);
                        q3 = make_uint2(nq.hi.z, nq.hi.w);
                        h0 = slab_q(nq.lo.x, nq.lo.y, nq.lo.z, qr, closest, d0);
                        h1 = slab_q(nq.lo.w, nq.hi.x, nq.hi.y, qr, closest, d1);
                    } else {
                        const float4 *nd = nodes + 4 * (size_t)cur;
                        // One 32-byte half per child (box + ref). Fetching the halves in opposite order on odd and even lanes
                        // (to spread one LDG.256 over all four sector positions of the L1 lines) was measured: no gain, and
                        // it makes tie-breaking lane-dependent, so every lane reads child 0 first.
                        const Float8 na = ldg256(nd), nb = ldg256(nd + 2);
                        q3 = make_uint2(__float_as_uint(na.hi.z), __float_as_uint(nb.hi.z));
                        h0 = slab(na.lo.x, na.lo.y, na.lo.z, na.lo.w, na.hi.x, na.hi.y, so, inv, sg, closest, d0);
                        h1 = slab(nb.lo.x, nb.lo.y, nb.lo.z, nb.lo.w, nb.hi.x, nb.hi.y, so, inv, sg, closest, d1);
                    }
#if RT_BRANCHFREE
                    // one straight-line selection instead of four branch bodies (both / only 0 / only 1 / none): on incoherent
                    // bounces each body ran with 4-15 of the warp's ~22 descending lanes (profiles/r2a: 12 % of the warp instructions)
                    const bool zero_near = h0 && (!h1 || d0 < d1);  // child 0 is the one to enter
                    const uint32_t near_ref = zero_near ? q3.x : q3.y;
                    if (h0 && h1) {
                        stack.push(sp, zero_near ? q3.y : q3.x, zero_near ? d1 : d0);
                        sp++;
                    }
                    if (h0 || h1) {
                        cur = near_ref;
                    } else {
                        float entry;
                        do {
                            if (sp == 0) { cur = kDoneRef; break; }
                            sp--;
                            stack.get(sp, cur, entry);
                        } while (entry >= closest);
                    }
#else
                    if (h0 && h1) {
                        const bool zero_near = d0 < d1;
                        stack.push(sp, zero_near ? q3.y : q3.x, zero_near ? d1 : d0);
                        sp++;
                        cur = zero_near ? q3.x : q3.y;
                    } else if (h0) {
                        cur = q3.x;
                    } else if (h1) {
                        cur = q3.y;
                    } else {
                        float entry;
                        do {
                            if (sp == 0) { cur = kDoneRef; break; }
                            sp--;
                            stack.get(sp, cur, entry);
                        } while (entry >= closest);
                    }
#endif
                }
                // lanes that already hold a leaf wait at the loop exit: stop descending once only a few are left
                if (__popc(__activemask()) < node_min) break;
            }
            }
            __syncwarp();
            const bool has_leaf = busy && (cur & kLeafBit) && cur != kDoneRef;
            if (__any_sync(FULL, has_leaf)) leaf_phase_shared<STATS>(sc, has_leaf, cur, o, d, closest, prim, leaf_table + (threadIdx.x & ~31u), ctr);
            if (has_leaf) {
                float entry;  // pop, culling entries that can no longer beat the current hit (scene.cu:150-153)
                do {
                    if (sp == 0) { cur = kDoneRef; break; }
                    sp--;
                    stack.get(sp, cur, entry);
                } while (entry >= closest);
            }
            if (busy && cur == kDoneRef) {
                io.store(slot, closest, prim);
                busy = false;
            }
            if ((uint32_t)__popc(__ballot_sync(FULL, busy)) < refill_below) break;
        }
#else
        // ---- walk: lanes leave individually, the warp reconverges at the __syncwarp below ----
        while (busy) {
            // descend through internal records until this lane holds a leaf (or runs dry)
            while (!(cur & kLeafBit)) {
                if (STATS) ctr.internal_visits++;
                if (NODES == kNodesW4) {
                    // 4-wide node: two 32-byte halves {children 0,1} {children 2,3}; hit children sorted by entry distance
                    // (5 compare-exchanges), the nearest becomes the current reference, the others go on the stack far first
                    const uint4 *nd = qnodes + 4 * (size_t)cur;
                    const Uint8 na = ldg256u(nd), nb = ldg256u(nd + 2);
                    constexpr float kNoHit = 3e38f;
                    float e0, e1, e2, e3;
                    const bool h0 = slab_q(na.lo.x, na.lo.y, na.lo.z, qr, closest, e0), h1 = slab_q(na.lo.w, na.hi.x, na.hi.y, qr, closest, e1);
                    const bool h2 = slab_q(nb.lo.x, nb.lo.y, nb.lo.z, qr, closest, e2), h3 = slab_q(nb.lo.w, nb.hi.x, nb.hi.y, qr, closest, e3);
                    float k0 = h0 ? e0 : kNoHit, k1 = h1 ? e1 : kNoHit, k2 = h2 ? e2 : kNoHit, k3 = h3 ? e3 : kNoHit;
                    uint32_t r0 = na.hi.z, r1 = na.hi.w, r2 = nb.hi.z, r3 = nb.hi.w;
                    const int hits = (int)h0 + (int)h1 + (int)h2 + (int)h3;
#define RT_CSWAP(ka, ra, kb, rb) { const bool sw = kb < ka; const float tk = sw ? ka : kb; ka = sw ? kb : ka; kb = tk; const uint32_t tr = sw ? ra : rb; ra = sw ? rb : ra; rb = tr; }
                    RT_CSWAP(k0, r0, k1, r1) RT_CSWAP(k2, r2, k3, r3) RT_CSWAP(k0, r0, k2, r2) RT_CSWAP(k1, r1, k3, r3) RT_CSWAP(k1, r1, k2, r2)
#undef RT_CSWAP
                    if (hits > 3) { stack.push(sp, r3, k3); sp++; }
                    if (hits > 2) { stack.push(sp, r2, k2); sp++; }
                    if (hits > 1) { stack.push(sp, r1, k1); sp++; }
                    if (hits > 0) {
                        cur = r0;
                    } else {
                        float entry;
                        do {
                            if (sp == 0) { cur = kDoneRef; break; }
                            sp--;
                            stack.get(sp, cur, entry);
                        } while (entry >= closest);
                    }
                } else {
                    float d0, d1;
                    bool h0, h1;
                    uint2 q3;
                    if (NODES == kNodesQ16) {  // one 32-byte record: 12 bounds on the 15-bit grid + the two child refs
                        const Uint8 nq = ldg256u(qnodes + 2 * (size_t)cur);
                        q3 = make_uint2(nq.hi.z, nq.hi.w);
                        h0 = slab_q(nq.lo.x, nq.lo.y, nq.lo.z, qr, closest, d0);
                        h1 = slab_q(nq.lo.w, nq.hi.x, nq.hi.y, qr, closest, d1);
                    } else {
                        const float4 *nd = nodes + 4 * (size_t)cur;
                        // One 32-byte half per child (box + ref). Fetching the halves in opposite order on odd and even lanes
                        // (to spread one LDG.256 over all four sector positions of the L1 lines) was measured: no gain, and
                        // it makes tie-breaking lane-dependent, so every lane reads child 0 first.
                        const Float8 na = ldg256(nd), nb = ldg256(nd + 2);
                        q3 = make_uint2(__float_as_uint(na.hi.z), __float_as_uint(nb.hi.z));
                        h0 = slab(na.lo.x, na.lo.y, na.lo.z, na.lo.w, na.hi.x, na.hi.y, so, inv, sg, closest, d0);
                        h1 = slab(nb.lo.x, nb.lo.y, nb.lo.z, nb.lo.w, nb.hi.x, nb.hi.y, so, inv, sg, closest, d1);
                    }
#if RT_BRANCHFREE
                    // one straight-line selection instead of four branch bodies (both / only 0 / only 1 / none): on incoherent
                    // bounces each body ran with 4-15 of the warp's ~22 descending lanes (profiles/r2a: 12 % of the warp instructions)
                    const bool zero_near = h0 && (!h1 || d0 < d1);  // child 0 is the one to enter
                    const uint32_t near_ref = zero_near ? q3.x : q3.y;
                    if (h0 && h1) {
                        stack.push(sp, zero_near ? q3.y : q3.x, zero_near ? d1 : d0);
                        sp++;
                    }
                    if (h0 || h1) {
                        cur = near_ref;
                    } else {
                        float entry;
                        do {
                            if (sp == 0) { cur = kDoneRef; break; }
                            sp--;
                            stack.get(sp, cur, entry);
                        } while (entry >= closest);
                    }
#else
                    if (h0 && h1) {
                        const bool zero_near = d0 < d1;
                        stack.push(sp, zero_near ? q3.y : q3.x, zero_near ? d1 : d0);
                        sp++;
                        cur = zero_near ? q3.x : q3.y;
                    } else if (h0) {
                        cur = q3.x;
                    } else if (h1) {
                        cur = q3.y;
                    } else {
                        float entry;
                        do {
                            if (sp == 0) { cur = kDoneRef; break; }
                            sp--;
                            stack.get(sp, cur, entry);
                        } while (entry >= closest);
                    }
#endif
                }
                // lanes that already hold a leaf wait at the loop exit: stop descending once only a few are left
                if (__popc(__activemask()) < node_min) break;
            }
            if ((cur & kLeafBit) && cur != kDoneRef) {
                const uint32_t first = cur & kLeafStartMask;
                const uint32_t count = (cur >> kLeafCountShift) & 15u;
                if (STATS) ctr.triangle_tests += count;
                for (uint32_t k = 0; k < count; k++) intersect_triangle(sc.tris, first + k, sc.sphere_count, o, d, closest, prim);
                float entry;  // pop, culling entries that can no longer beat the current hit (scene.cu:150-153)
                do {
                    if (sp == 0) { cur = kDoneRef; break; }
                    sp--;
                    stack.get(sp, cur, entry);
                } while (entry >= closest);
            }
            if (cur == kDoneRef) {
                io.store(slot, closest, prim);
                busy = false;
                break;
            }
            if (__popc(__activemask()) < refill_below) break;  // heuristic only: correctness never depends on this mask
        }
#endif
        __syncwarp();
    }
}

}  // namespace rtd
