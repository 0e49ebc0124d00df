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

// Boundary-safe slab test (Tavian Barnes' min/max form, scene.cu:109-132): tmin starts at 0, tmax at `limit`.
// Arithmetic is (bound - o) * inv exactly as the reference writes it; there is no a*b+c shape to contract.
__device__ __forceinline__ bool slab(float lox, float loy, float loz, float hix, float hiy, float hiz, float3 o, float3 inv,
                                     float limit, float &t_enter)
{
    float tmin = 0.0f, tmax = limit;
    float t1 = (lox - o.x) * inv.x, t2 = (hix - o.x) * inv.x;
    tmin = fminf(fmaxf(t1, tmin), fmaxf(t2, tmin));
    tmax = fmaxf(fminf(t1, tmax), fminf(t2, tmax));
    t1 = (loy - o.y) * inv.y; t2 = (hiy - o.y) * inv.y;
    tmin = fminf(fmaxf(t1, tmin), fmaxf(t2, tmin));
    tmax = fmaxf(fminf(t1, tmax), fminf(t2, tmax));
    t1 = (loz - o.z) * inv.z; t2 = (hiz - o.z) * inv.z;
    tmin = fminf(fmaxf(t1, tmin), fmaxf(t2, tmin));
    tmax = fmaxf(fminf(t1, tmax), fminf(t2, tmax));
    t_enter = tmin;
    return tmin <= tmax;
}

// scene.cu:338-372. Returns through closest/prim; runs before the BVH and seeds `closest`.
__device__ __forceinline__ void intersect_spheres(const DeviceScene &sc, float3 o, float3 d, float &closest, int &prim)
{
    for (int i = 0; i < sc.sphere_count; i++) {
        float4 s = __ldg(sc.spheres + i);
        float3 offset = f3(s.x, s.y, s.z) - o;
        float minus_half_b = dot(offset, d);
        float quarter_c = length2(offset) - s.w * s.w;
        float quarter_disc = minus_half_b * minus_half_b - quarter_c;
        if (quarter_disc < 0.0f) continue;
        float half_root = sqrtf(quarter_disc);
        float t = minus_half_b - half_root;
        if (t < closest && t > kTMin) { closest = t; prim = i; continue; }
        t = minus_half_b + half_root;
        if (t < closest && t > kTMin) { closest = t; prim = i; continue; }
    }
}

// Moller-Trumbore on the precomputed-edge record (scene.cu:164-194). Updates closest/prim on an accepted hit.
__device__ __forceinline__ void intersect_triangle(const float4 *__restrict__ tris, uint32_t tri, int prim_base, float3 o, float3 d,
                                                   float &closest, int &prim)
{
    const float4 *t4 = tris + 3 * (size_t)tri;
    float4 a4 = __ldg(t4), b4 = __ldg(t4 + 1), c4 = __ldg(t4 + 2);
    float3 p1 = f3(a4.x, a4.y, a4.z), e1 = f3(b4.x, b4.y, b4.z), e2 = f3(c4.x, c4.y, c4.z);
    float3 h = cross(d, e2);
    float a = dot(h, e1);
    if (a == 0.0f) return;
    float inv_a = 1.0f / a;
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

template <int BLOCK, bool STATS>
__device__ __forceinline__ void traverse_bvh(const DeviceScene &sc, float3 o, float3 d, float &closest, int &prim,
                                             SharedStack<BLOCK> &stack, TraceCounters &ctr)
{
    const float3 inv = f3(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);
    const float4 *__restrict__ nodes = sc.nodes;
    const int lane = threadIdx.x;
    uint32_t cur = 0;  // record 0 is the root
    int sp = 0;
    for (;;) {
        if (!(cur & kLeafBit)) {
            const float4 *n = nodes + 4 * (size_t)cur;
            float4 q0 = __ldg(n), q1 = __ldg(n + 1), q2 = __ldg(n + 2);
            uint4 q3 = __ldg(reinterpret_cast<const uint4 *>(n + 3));
            if (STATS) ctr.internal_visits++;
            float d0, d1;
            bool h0 = slab(q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, o, inv, closest, d0);
            bool h1 = slab(q1.z, q1.w, q2.x, q2.y, q2.z, q2.w, o, inv, closest, d1);
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

}  // namespace rtd
