// float3 helpers and the PCG-XSH-RR 64/32 generator of the render path (reference: random.cuh:5-75, math.cuh).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace rtd {

__device__ __forceinline__ float3 f3(float x, float y, float z) { return make_float3(x, y, z); }
__device__ __forceinline__ float3 operator+(float3 a, float3 b) { return f3(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ float3 operator-(float3 a, float3 b) { return f3(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ float3 operator*(float3 a, float3 b) { return f3(a.x * b.x, a.y * b.y, a.z * b.z); }
__device__ __forceinline__ float3 operator*(float s, float3 a) { return f3(s * a.x, s * a.y, s * a.z); }
__device__ __forceinline__ float3 operator-(float3 a) { return f3(-a.x, -a.y, -a.z); }
__device__ __forceinline__ float dot(float3 a, float3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ float3 cross(float3 a, float3 b)
{
    return f3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
__device__ __forceinline__ float length2(float3 a) { return a.x * a.x + a.y * a.y + a.z * a.z; }
// (1/|v|) * v with IEEE sqrt and division, as math.cuh:110-113 on the host
__device__ __forceinline__ float3 unit(float3 a) { return (1.0f / sqrtf(length2(a))) * a; }
// Shading-only variants on the SFU (MUFU.RSQ / MUFU.SIN / MUFU.COS): ~2 ulp and |err| < 5e-7 for phi in [0, 2 pi],
// far inside the 2e-4 shading tolerance of the parity tests; the reference's own GPU build uses the same fast-math
// instructions (--use_fast_math, build.sh:2). Ray generation keeps the IEEE forms (it is checked bit for bit).
__device__ __forceinline__ float3 unit_fast(float3 a) { return rsqrtf(length2(a)) * a; }
__device__ __forceinline__ bool is_zero(float3 a) { return a.x == 0.0f && a.y == 0.0f && a.z == 0.0f; }

// Same sums with contraction forbidden: used where the result must equal the host oracle bit for bit
// (primary-ray generation, bloom threshold).
__device__ __forceinline__ float dot_nofma(float3 a, float3 b)
{
    return __fadd_rn(__fadd_rn(__fmul_rn(a.x, b.x), __fmul_rn(a.y, b.y)), __fmul_rn(a.z, b.z));
}

// 256-bit read-only load (sm_100a: LDG.E.256): one L1 request for half a BVH record / a triangle's edge block where the
// 128-bit form needs two. p must be 32-byte aligned.
struct __align__(32) Float8 {
    float4 lo, hi;
};
__device__ __forceinline__ Float8 ldg256(const void *p)
{
    Float8 r;
    asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(r.lo.x), "=f"(r.lo.y), "=f"(r.lo.z), "=f"(r.lo.w), "=f"(r.hi.x), "=f"(r.hi.y), "=f"(r.hi.z), "=f"(r.hi.w)
                 : "l"(p));
    return r;
}

struct Pcg {
    unsigned long long state;
    unsigned long long inc;  // already OR-ed with 1
};

__device__ __forceinline__ uint32_t pcg_next(Pcg &r)  // random.cuh:13-23
{
    unsigned long long old = r.state;
    r.state = old * 6364136223846793005ULL + r.inc;
    uint32_t xorshifted = (uint32_t)(((old >> 18) ^ old) >> 27);
    uint32_t rot = (uint32_t)(old >> 59);
    return __funnelshift_r(xorshifted, xorshifted, rot);  // rotate right
}

__device__ __forceinline__ Pcg pcg_seed(uint32_t seed, unsigned long long inc_or_1)  // random.cuh:25-30
{
    Pcg r;
    r.state = (unsigned long long)seed * 6839056345687307ULL;
    r.inc = inc_or_1;
    pcg_next(r);
    return r;
}

// u32 -> float rounds to nearest, then * 2^-32: may return exactly 1.0f, random.cuh:32-35
__device__ __forceinline__ float rnd01(Pcg &r) { return (float)pcg_next(r) * 2.3283064365386963e-10f; }
__device__ __forceinline__ float rnd02(Pcg &r) { return (float)pcg_next(r) * 4.656612873077393e-10f; }
// random.cuh:42-45 multiplies in double by 2*pi/UINT_MAX and narrows; kept so angles match the oracle exactly
__device__ __forceinline__ float rnd_radians(Pcg &r) { return (float)((double)pcg_next(r) * (6.283185307179586 / 4294967295.0)); }

__device__ __forceinline__ float3 rnd_on_sphere(Pcg &r)  // random.cuh:63-75
{
    float phi = rnd_radians(r);
    float r2 = rnd02(r);
    float x = sqrtf(r2 * (2.0f - r2));
    float s, c;
    __sincosf(phi, &s, &c);
    return f3(c * x, s * x, 1.0f - r2);
}

}  // namespace rtd
