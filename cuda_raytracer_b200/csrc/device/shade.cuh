// Ray generation and per-bounce shading of the render path.
//   generate_ray  <- Scene::generate_initial_rays                    (scene.cu:78-105)
//   shade_miss    <- env rotation + equal-area projection + lookup  (scene.cu:376-395, 284-318)
//   shade_hit     <- material scatter: diffuse / metal / dielectric (scene.cu:396-477)
#pragma once

#include "device_scene.cuh"
#include "rng_math.cuh"

namespace rtd {

// Primary ray `ray_index` of a pass of `rays_per_pixel` samples. Contraction is forbidden so that the direction is
// bit-identical to the host oracle (the whole bounce-0 hit comparison rests on identical rays).
__device__ __forceinline__ bool generate_ray(const DeviceScene &sc, uint32_t ray_index, const FastDiv &rays_per_pixel, uint32_t gen_seed_term,
                                             unsigned long long inc, float3 &o, float3 &d)
{
    Pcg rng = pcg_seed(ray_index * 2239826922u + gen_seed_term, inc);  // 298592570346 mod 2^32 = 2239826922 (scene.cu:81)
    const uint32_t pixel = fast_div(ray_index, rays_per_pixel);
    const uint32_t y = fast_div(pixel, sc.div_width), x = pixel - y * (uint32_t)sc.width;
    if (y >= (uint32_t)sc.height) return false;
    float xs = __fmul_rn(__fadd_rn((float)x, rnd01(rng)), sc.inv_width);
    float ys = __fmul_rn(__fadd_rn((float)y, rnd01(rng)), sc.inv_height);
    float3 v;
    v.x = __fsub_rn(__fadd_rn(sc.top_left[0], __fmul_rn(xs, sc.right[0])), __fmul_rn(ys, sc.up[0]));
    v.y = __fsub_rn(__fadd_rn(sc.top_left[1], __fmul_rn(xs, sc.right[1])), __fmul_rn(ys, sc.up[1]));
    v.z = __fsub_rn(__fadd_rn(sc.top_left[2], __fmul_rn(xs, sc.right[2])), __fmul_rn(ys, sc.up[2]));
    float inv_len = 1.0f / sqrtf(dot_nofma(v, v));
    d = f3(__fmul_rn(inv_len, v.x), __fmul_rn(inv_len, v.y), __fmul_rn(inv_len, v.z));
    o = f3(sc.cam[0], sc.cam[1], sc.cam[2]);
    return true;
}

// pbrt's equal-area octahedral sphere -> square map (scene.cu:284-318)
__device__ __forceinline__ float2 sphere_to_square(float3 dir)
{
    float x = fabsf(dir.x), y = fabsf(dir.y), z = fabsf(dir.z);
    float r = sqrtf(1.0f - fminf(z, 1.0f));
    float a = fmaxf(x, y), b = fminf(x, y);
    b = (a == 0.0f) ? 0.0f : b / a;
    float phi = (float)(0.6366197723675814 * (double)atanf(b));  // (2/pi) * atan(b) in double, as the reference (Q6)
    if (x < y) phi = 1.0f - phi;
    float v = phi * r;
    float u = r - v;
    if (dir.z < 0.0f) {
        float old_v = v;
        v = 1.0f - u;
        u = 1.0f - old_v;
    }
    u = copysignf(u, dir.x);
    v = copysignf(v, dir.y);
    return make_float2((u + 1.0f) * 0.5f, (v + 1.0f) * 0.5f);
}

// Sky radiance seen along d (nearest texel; the row stride is the map HEIGHT as in the reference, Q7).
__device__ __forceinline__ float3 shade_miss(const DeviceScene &sc, float3 d)
{
    // hard-coded rotation + y/z swap of the test data's map (scene.cu:378-382); double products narrowed to float
    float dx = (float)((double)d.x * -0.386527 + (double)d.z * 0.922278);
    float dy = (float)((double)d.x * -0.922278 + (double)d.z * -0.386527);
    float2 uv = sphere_to_square(f3(dx, dy, d.y));
    float cu = fmaxf(fminf(uv.x, 1.0f), 0.0f), cv = fmaxf(fminf(uv.y, 1.0f), 0.0f);
    int tx = (int)((double)(cu * (float)(sc.env_w - 1)) + 0.5);
    int ty = (int)((double)(cv * (float)(sc.env_h - 1)) + 0.5);
    float4 texel = __ldg(sc.env + (size_t)ty * sc.env_h + tx);
    return f3(texel.x, texel.y, texel.z);
}

// One surface interaction. In: ray (o,d), hit distance t, primitive id, throughput T. Out: next ray, new
// throughput, `emitted` = radiance to add for this hit (material.emitted * T_before).
__device__ __forceinline__ void shade_hit(const DeviceScene &sc, float3 &o, float3 &d, float t, int prim, float3 &T, Pcg &rng,
                                          float3 &emitted)
{
    float3 hit = o + t * d;
    float3 n;
    uint32_t material;
    if (prim < sc.sphere_count) {
        float4 s = __ldg(sc.spheres + prim);
        n = (1.0f / s.w) * (hit - f3(s.x, s.y, s.z));
        material = __ldg(sc.material_indices + prim);
    } else {  // words 8-15 of the triangle record: e2.z, n.xyz, material index
        const Float8 t8 = ldg256(sc.tris + 4 * (size_t)(prim - sc.sphere_count) + 2);
        n = f3(t8.lo.y, t8.lo.z, t8.lo.w);
        material = __float_as_uint(t8.hi.x);
    }
    const float4 *m = sc.materials + 3 * (size_t)material;
    float4 m0 = __ldg(m), m1 = __ldg(m + 1), m2 = __ldg(m + 2);
    float3 diffuse = f3(m0.x, m0.y, m0.z), specular = f3(m1.x, m1.y, m1.z);
    float metallicity = m0.w, roughness = m1.w, ior = m2.w;

    emitted = f3(m2.x, m2.y, m2.z) * T;
    o = hit;

    bool front = dot(n, d) < 0.0f;
    if (!front) n = -n;
    float3 nr = unit_fast(n + roughness * rnd_on_sphere(rng));
    float cos_theta = dot(nr, d);

    if (ior == 0.0f) {
        if (rnd01(rng) <= metallicity) {  // inclusive compare (Q8)
            T = T * specular;
            d = d - (2.0f * cos_theta) * nr;  // not renormalised (Q11)
        } else {
            T = T * diffuse;
            d = unit_fast(n + rnd_on_sphere(rng));
        }
    } else {
        float inv_ior = 1.0f / ior;
        if (front) { float tmp = inv_ior; inv_ior = ior; ior = tmp; }
        float sin2 = 1.0f - cos_theta * cos_theta;
        float r0 = (1.0f - ior) / (1.0f + ior);
        r0 *= r0;
        float c = 1.0f + cos_theta;
        float reflectance = r0 + (1.0f - r0) * c * c * c * c * c;
        if (sin2 > inv_ior * inv_ior || rnd01(rng) < reflectance) {
            T = T * specular;
            d = d - (2.0f * cos_theta) * nr;
        } else {
            T = T * diffuse;
            float3 perp = ior * (d - cos_theta * nr);
            float3 par = (-sqrtf(1.0f - length2(perp))) * nr;
            d = unit_fast(par + perp);
        }
    }
}

// Reorder key of a live ray: Morton code of the origin cell (grid over the scene bounds) in the high bits, direction
// octant in the low 3 bits. Replaces the reference's key (scene.cu:44-60,480-485), whose interleave collapses to 6
// useful bits and whose normalisation divides by max instead of extent (Q2).
__device__ __forceinline__ uint32_t spread3(uint32_t v)  // 10 bits -> every third bit
{
    v = (v | (v << 16)) & 0x030000FFu;
    v = (v | (v << 8)) & 0x0300F00Fu;
    v = (v | (v << 4)) & 0x030C30C3u;
    v = (v | (v << 2)) & 0x09249249u;
    return v;
}
__device__ __forceinline__ uint32_t reorder_key(const float *min_coord, const float *inv_extent, float3 o, float3 d, int key_bits)
{
    // cell coordinates with `g` bits per axis, g = (key_bits - 3) / 3
    int g = (key_bits - 3) / 3;
    float scale = (float)(1 << g);
    uint32_t hi = (1u << g) - 1u;
    uint32_t cx = min((uint32_t)fmaxf((o.x - min_coord[0]) * inv_extent[0] * scale, 0.0f), hi);
    uint32_t cy = min((uint32_t)fmaxf((o.y - min_coord[1]) * inv_extent[1] * scale, 0.0f), hi);
    uint32_t cz = min((uint32_t)fmaxf((o.z - min_coord[2]) * inv_extent[2] * scale, 0.0f), hi);
    uint32_t morton = spread3(cx) | (spread3(cy) << 1) | (spread3(cz) << 2);
    uint32_t octant = (d.x < 0.0f ? 1u : 0u) | (d.y < 0.0f ? 2u : 0u) | (d.z < 0.0f ? 4u : 0u);
    return (morton << 3) | octant;
}

}  // namespace rtd
