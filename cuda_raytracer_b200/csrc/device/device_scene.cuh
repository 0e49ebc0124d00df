// Device-side data layout of the render core (sm_100a).
//
//  * BVH: only INTERNAL nodes are stored, one 64-byte record (4 x float4, 64-byte aligned) per node holding BOTH
//    children's boxes and both child references, so one visit = one cache-line-half fetch instead of the
//    reference's node + two child loads (scene.cu:145-201). Valid because children are always adjacent.
//      half 0 = {c0.centre.xyz, c0.half_extent.xyz, ref0, pad}   half 1 = the same for child 1        (32 bytes each)
//    (RT_BOX_CH=0 builds keep {min.xyz, max.xyz}); the half-extent is rounded up, so the box contains the builder's.
//    ref: bit31 clear -> index of an internal record; bit31 set -> leaf: bits[30:27] = triangle count (1..15),
//    bits[26:0] = first triangle. Larger leaves (no_bvh, SAH refusals) are split into chains of same-box records
//    by the packer. An absent child has a far-away degenerate box (an inverted box would pass the
//    order-agnostic slab test) and ref = kEmptyRef.
//  * Triangles: one 64-byte record (32-byte aligned for LDG.256), BVH order, precomputed-edge form (scene.cuh:15-33):
//      words 0-7  = p1.xyz, e1.xyz, e2.xy      <- the intersection test reads these 32 bytes + word 8
//      words 8-12 = e2.z, n.xyz, material index (as int bits)   <- shading reads words 9-12 with one load
//      words 13-15 unused.
//  * Materials: 3 x float4, the reference's 48-byte record. Spheres: float4. Environment map: float4 texels.
//  * Ray queues: SoA, 40 bytes per live ray:
//      qa = {o.x, o.y, o.z, d.x}   qb = {d.y, d.z, T.r, T.g}   qc = {T.b, ray id (bits)}
//    Radiance is not carried: it is added to the framebuffer when it is found (miss / emissive hit).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace rtd {

// RT_BOX_CH (default 1): a child's box is stored as {centre.xyz, half-extent.xyz} instead of {min.xyz, max.xyz}. The slab test
// then needs no near / far selection by the ray's signs: t_centre = fma(c, 1/d, -o/d), t_near = t_centre - h |1/d|,
// t_far = t_centre + h |1/d| (9 FFMA per box instead of 6 LOP3 + 6 FFMA). The half-extent is rounded UP, so the stored box
// contains the builder's box. Measured (profiles/r1k_box_ch.md): node visit 59 -> 53 instructions, lamp 4K +1.9 %.
#ifndef RT_BOX_CH
#define RT_BOX_CH 1
#endif

constexpr uint32_t kLeafBit = 0x80000000u;
constexpr uint32_t kEmptyRef = 0x80000000u;  // leaf with zero triangles; its box is parked far away so the slab test fails
constexpr int kLeafCountShift = 27;
constexpr uint32_t kLeafStartMask = (1u << kLeafCountShift) - 1u;
constexpr int kMaxLeafTris = 15;
constexpr int kMaxPassGroup = 16;  // passes traced together as one wave (rt_render_opts.pass_group)
constexpr int kStackDepth = 32;  // deepest packed tree the tracer accepts
constexpr int kStackSlots = kStackDepth + 8;  // stack entries per thread: tree depth + the entry records of a primary ray

// accept t iff t > kTMin: the reference rejects `t < 0.005` in DOUBLE (scene.cu:190,357,366); 0.005f is the largest
// float below 0.005, so for float t:  (double)t < 0.005  <=>  t <= 0.005f.
constexpr float kTMin = 0.005f;
constexpr float kMissT = 1e30f;

// Division of a 32-bit index by a launch-invariant divisor (rays per pass, samples per pixel, image width) without the
// ~20-instruction software division: Granlund & Montgomery's multiply-and-shift, exact for every 32-bit numerator.
// m = floor(2^32 (2^L - d) / d) + 1 with L = ceil(log2 d);  n / d = (((n - hi(m n)) >> min(L, 1)) + hi(m n)) >> max(L - 1, 0).
struct FastDiv {
    uint32_t m, sh1, sh2, d;
};
inline FastDiv make_fast_div(uint32_t d)
{
    if (d == 0) d = 1;
    uint32_t L = 0;
    while (L < 32 && (1ull << L) < d) L++;
    const uint64_t m = ((1ull << 32) * ((1ull << L) - d)) / d + 1;
    return FastDiv{(uint32_t)m, L < 1 ? L : 1u, L > 1 ? L - 1 : 0u, d};
}
#ifdef __CUDACC__
__host__ __device__ __forceinline__ uint32_t fast_div(uint32_t n, const FastDiv &f)
{
#ifdef __CUDA_ARCH__
    const uint32_t q = __umulhi(f.m, n);
#else
    const uint32_t q = (uint32_t)(((uint64_t)f.m * n) >> 32);  // host mirror, reachable through rt_index_divide (tests)
#endif
    return (((n - q) >> f.sh1) + q) >> f.sh2;
}
#endif

// Node formats the tracer can walk (rt_ctx::node_mode, environment RT_NODES; SURVEY §8f N3 "wider / compressed BVH"):
//   kNodesF32  the 64-byte two-child record above: two LDG.256 per visit
//   kNodesQ16  the SAME binary tree, each record compressed to 32 bytes = ONE LDG.256 per visit: the twelve bounds of the two
//              children as 15-bit coordinates on a grid over the root box (rounded outwards + one cell of margin), two per
//              word {lo | hi << 16}, words 0-2 = child 0 x/y/z, 3-5 = child 1, words 6-7 = the two child refs. Every stored
//              16-bit value has its top bit set, so that ONE byte permute builds the float 1 + q / 32768 from it
//              (0x3F000000 | v << 8) and the ray's sign picks the near / far bound in the same permute; t of a bound is one
//              FFMA with per-ray constants (traverse.cuh, QuantRay).
//   kNodesW4   a 4-wide tree collapsed from the binary one, 64 bytes per node = two kNodesQ16-style halves (children 0,1 and
//              2,3 with their refs): two LDG.256 per visit, about half the visits.
enum NodeMode : int { kNodesF32 = 0, kNodesQ16 = 1, kNodesW4 = 2 };

struct DeviceScene {
    const float4 *nodes;      // 4 per internal record
    const uint4 *qnodes;      // kNodesQ16: 2 per internal record (same numbering as `nodes`); kNodesW4: 4 per wide node
    float q_scale[3];         // 32768 x cell size of the quantisation grid, per axis
    float q_bias[3];          // grid origin - q_scale: coordinate = q_bias + (1 + q / 32768) * q_scale
    const float4 *tris;       // 4 per triangle (64-byte records)
    const float4 *materials;  // 3 per material
    const float4 *spheres;
    const float4 *env;
    const uint16_t *material_indices;
    int32_t sphere_count, tri_count, env_w, env_h;
    int32_t width, height;
    FastDiv div_width;
    float cam[3], top_left[3], right[3], up[3];
    float inv_width, inv_height;
    float min_coord[3], inv_dim[3];          // the reference's key bounds (scene.cu:822-830), kept for completeness
    float key_min[3], key_inv_extent[3];     // bounds used by reorder_key: triangle root box (or sphere bounds), 1/(max-min)
};

struct RayQueue {
    float4 *qa;
    float4 *qb;
    float2 *qc;
};

// Per-pass device counters (zeroed by one memset per pass).
constexpr int kMaxBounces = 256;  // the reference has no limit; these per-bounce counters need one
struct PassCounters {
    uint32_t cursor[kMaxBounces + 1];  // work-fetch cursor of bounce b
    uint32_t count[kMaxBounces + 1];   // live rays entering bounce b (count[0] unused: known on the host)
    uint32_t tail_bounce;              // 0 = the bounce-by-bounce launches are still running this wave; b >= 2 = tail_kernel took
                                       // over at bounce b (every later trace / shade / tail launch of the wave returns at once)
};

struct RenderTotals {
    unsigned long long live_ray_bounces;
    unsigned long long internal_visits;
    unsigned long long triangle_tests;
    unsigned long long live_per_bounce[32];  // bounces past the 32nd are added to the last entry
};

struct BounceArgs {
    RayQueue in, out;
    float2 *hits;           // hit record {t, prim as int bits} per position of the input queue (trace -> shade)
    const uint32_t *order;  // optional permutation of the input queue (reorder pass), else nullptr
    uint32_t *keys_out;     // optional: reorder key of every ray written to `out`
    PassCounters *pc;
    RenderTotals *totals;
    float *framebuffer;     // 3 floats per pixel, un-normalised sums
    uint32_t n_first;       // rays of the wave = group x pass_rays (bounce 0 only)
    uint32_t pass_rays;     // rays of ONE pass (pixels x rays_per_pixel); ray id = pass-in-wave x pass_rays + index in pass
    uint32_t partial_g;     // member of the wave that is a PARTIAL pass (always the last one; 0xFFFFFFFF = none): the three fields below are its
    uint32_t ray_offset;    // a wave that holds PART of one pass (rt_render_opts.ray_first / ray_count): index in pass = id + ray_offset
    uint32_t sub_count;     // > 0: the wave holds samples [sub_first, sub_first + sub_count) of EVERY pixel of one pass
    uint32_t sub_first;     //      (rt_render_opts.sample_first / sample_count): index in pass = (id / sub_count) * rays_per_pixel + sub_first + id % sub_count
    FastDiv div_sub_count;
    int32_t bounce;
    int32_t rays_per_pixel; // of every pass of this wave
    FastDiv div_pass_rays, div_rays_per_pixel;
    uint32_t gen_seed[kMaxPassGroup];   // per pass of the wave: 709579 * remaining_rays           (scene.cu:81)
    uint32_t proc_seed[kMaxPassGroup];  // per pass of the wave: 279220567 * (remaining*20+bounce)  (raytracing.cu:89,236)
    unsigned long long pcg_inc;  // 820957824423429 | 1 for stream 0 (random.cuh:28)
    int32_t key_bits;
    uint32_t trace_batch, trace_refill_below, trace_node_min;  // TraceTuning
    const uint4 *entry_lists;  // per-tile entry records for primary rays (entry_lists.cuh), nullptr = start at the root
    int32_t entry_tiles_x;
    int32_t bounces_total;  // the frame's bounce limit (tail_kernel runs a ray from a.bounce to its end)
    uint32_t tail_below;    // tail_kernel takes the wave over when at most this many rays enter its bounce (0 = never)
};

}  // namespace rtd
