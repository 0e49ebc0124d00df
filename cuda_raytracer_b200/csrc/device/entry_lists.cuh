// Entry points for primary rays: per screen tile, the few BVH records a ray of that tile can reach.
//
// Every primary ray starts at the camera, and the 20 samples of a pixel and its neighbours walk the same top of the tree
// (Scene::bvh_closest_hit_distance, scene.cu:134-241, starts each of them at the root). Here one thread per tile of
// kEntryTile x kEntryTile pixels cuts the tree once per frame with the tile's (widened) frustum and keeps at most
// kEntryMax records, nearest first; the bounce-0 tracer seeds its stack with them instead of the root.
//
// Why results are unchanged: a record enters the list only if the frustum test cannot exclude it, so every record a ray of
// the tile can hit lies below (or is) an entry. Skipping the box tests of the ancestors is harmless because boxes nest
// exactly (a parent's bounds are min / max of its children's) and the slab test is monotone in the bounds under
// round-to-nearest: a ray that passes a child's test passes every ancestor's. Entries are internal records only (the
// children's boxes, including a leaf child's, are still tested by the ray itself). Each entry carries a lower bound of
// the distance at which any ray from the camera can enter it (distance from the camera to its box, shrunk by 1e-4), used
// by the ordinary `entry >= closest` pop test.
#pragma once

#include "device_scene.cuh"
#include "rng_math.cuh"

namespace rtd {

#ifndef RT_ENTRY_TILE
#define RT_ENTRY_TILE 8
#endif
constexpr int kEntryTile = RT_ENTRY_TILE;  // pixels per tile side
constexpr int kEntryMax = 7;   // entries per tile: {count, 7 refs, 7 distances, pad} = one 64-byte record

struct EntryFrustum {
    float3 o;
    float3 n[4];  // inward normals of the four side planes through the camera
};

// conservative: true unless the box lies wholly outside one side plane (with a relative margin far above float rounding)
__device__ __forceinline__ bool frustum_may_touch(const EntryFrustum &f, const float *lo, const float *hi)
{
    const float a[3] = {lo[0] - f.o.x, lo[1] - f.o.y, lo[2] - f.o.z}, b[3] = {hi[0] - f.o.x, hi[1] - f.o.y, hi[2] - f.o.z};
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const float n[3] = {f.n[i].x, f.n[i].y, f.n[i].z};
        float most = 0.0f, scale = 0.0f;
#pragma unroll
        for (int k = 0; k < 3; k++) {
            most += fmaxf(n[k] * a[k], n[k] * b[k]);
            scale += fabsf(n[k]) * fmaxf(fabsf(a[k]), fabsf(b[k]));
        }
        if (most < -1e-4f * scale) return false;
    }
    return true;
}

__device__ __forceinline__ float box_distance(const float3 o, const float *lo, const float *hi)
{
    const float dx = fmaxf(fmaxf(lo[0] - o.x, o.x - hi[0]), 0.0f), dy = fmaxf(fmaxf(lo[1] - o.y, o.y - hi[1]), 0.0f),
                dz = fmaxf(fmaxf(lo[2] - o.z, o.z - hi[2]), 0.0f);
    return sqrtf(dx * dx + dy * dy + dz * dz) * (1.0f - 1e-4f);
}

// lists: one uint4[4] record per tile. max_entries <= kEntryMax (tuning), rounds bounds the depth of the cut.
__global__ void build_entry_lists_kernel(const DeviceScene sc, int tiles_x, int tiles_y, int max_entries, uint4 *__restrict__ lists)
{
    const int tile = blockIdx.x * blockDim.x + threadIdx.x;
    if (tile >= tiles_x * tiles_y) return;
    const int tx = tile % tiles_x, ty = tile / tiles_x;
    // pixel x covers xs in [x, x + 1] * inv_width (scene.cu:86-87; the jitter can be exactly 1, Q8); half a pixel of margin
    const float xs0 = ((float)(tx * kEntryTile) - 0.5f) * sc.inv_width, xs1 = ((float)min(sc.width, (tx + 1) * kEntryTile) + 0.5f) * sc.inv_width;
    const float ys0 = ((float)(ty * kEntryTile) - 0.5f) * sc.inv_height, ys1 = ((float)min(sc.height, (ty + 1) * kEntryTile) + 0.5f) * sc.inv_height;
    const float3 tl = f3(sc.top_left[0], sc.top_left[1], sc.top_left[2]), r = f3(sc.right[0], sc.right[1], sc.right[2]),
                 u = f3(sc.up[0], sc.up[1], sc.up[2]);
    const float3 c00 = tl + xs0 * r - ys0 * u, c10 = tl + xs1 * r - ys0 * u, c01 = tl + xs0 * r - ys1 * u, c11 = tl + xs1 * r - ys1 * u;
    const float3 mid = 0.25f * (c00 + c10 + c01 + c11);
    EntryFrustum f;
    f.o = f3(sc.cam[0], sc.cam[1], sc.cam[2]);
    f.n[0] = cross(c00, c10);  // top edge
    f.n[1] = cross(c10, c11);  // right edge
    f.n[2] = cross(c11, c01);  // bottom edge
    f.n[3] = cross(c01, c00);  // left edge
#pragma unroll
    for (int i = 0; i < 4; i++)
        if (dot(f.n[i], mid) < 0.0f) f.n[i] = -1.0f * f.n[i];

    uint32_t ref[kEntryMax], next_ref[kEntryMax];
    float dist[kEntryMax], next_dist[kEntryMax];
    int count = 1;
    ref[0] = 0;  // record 0 = the root
    dist[0] = 0.0f;
    for (int round = 0; round < 28; round++) {
        int n_next = 0;
        bool expanded = false;
        for (int e = 0; e < count; e++) {
            const float4 *nd = sc.nodes + 4 * (size_t)ref[e];
            const float4 a0 = __ldg(nd), a1 = __ldg(nd + 1), b0 = __ldg(nd + 2), b1 = __ldg(nd + 3);
#if RT_BOX_CH  // {centre, half-extent} -> bounds, rounded outwards
            const float lo0[3] = {__fsub_rd(a0.x, a0.w), __fsub_rd(a0.y, a1.x), __fsub_rd(a0.z, a1.y)}, hi0[3] = {__fadd_ru(a0.x, a0.w), __fadd_ru(a0.y, a1.x), __fadd_ru(a0.z, a1.y)};
            const float lo1[3] = {__fsub_rd(b0.x, b0.w), __fsub_rd(b0.y, b1.x), __fsub_rd(b0.z, b1.y)}, hi1[3] = {__fadd_ru(b0.x, b0.w), __fadd_ru(b0.y, b1.x), __fadd_ru(b0.z, b1.y)};
#else
            const float lo0[3] = {a0.x, a0.y, a0.z}, hi0[3] = {a0.w, a1.x, a1.y}, lo1[3] = {b0.x, b0.y, b0.z}, hi1[3] = {b0.w, b1.x, b1.y};
#endif
            const uint32_t r0 = __float_as_uint(a1.z), r1 = __float_as_uint(b1.z);
            // an empty child has its box parked far away; treat it as absent
            const bool t0 = r0 != kEmptyRef && frustum_may_touch(f, lo0, hi0), t1 = r1 != kEmptyRef && frustum_may_touch(f, lo1, hi1);
            if (!t0 && !t1) { expanded = true; continue; }  // nothing of this tile below: the entry disappears
            const bool leaf0 = (r0 & kLeafBit) != 0, leaf1 = (r1 & kLeafBit) != 0;
            // a leaf child keeps its parent as the entry (the ray tests the leaf's box itself); so does a full list:
            // the entries still to come need one slot each
            const int later = count - e - 1, children = (t0 ? 1 : 0) + (t1 ? 1 : 0);
            if ((t0 && leaf0) || (t1 && leaf1) || n_next + children + later > max_entries) {
                next_ref[n_next] = ref[e]; next_dist[n_next] = dist[e]; n_next++;
                continue;
            }
            if (t0) { next_ref[n_next] = r0; next_dist[n_next] = box_distance(f.o, lo0, hi0); n_next++; }
            if (t1) { next_ref[n_next] = r1; next_dist[n_next] = box_distance(f.o, lo1, hi1); n_next++; }
            expanded = true;
        }
        count = n_next;
        for (int e = 0; e < count; e++) { ref[e] = next_ref[e]; dist[e] = next_dist[e]; }
        if (!expanded || count == 0) break;
    }
    // nearest first (insertion sort, <= 7 entries)
    for (int i = 1; i < count; i++) {
        const uint32_t rr = ref[i];
        const float dd = dist[i];
        int j = i - 1;
        while (j >= 0 && dist[j] > dd) { ref[j + 1] = ref[j]; dist[j + 1] = dist[j]; j--; }
        ref[j + 1] = rr;
        dist[j + 1] = dd;
    }
    uint32_t w[16];
    w[0] = (uint32_t)count;
    for (int e = 0; e < kEntryMax; e++) {
        w[1 + e] = e < count ? ref[e] : 0u;
        w[8 + e] = __float_as_uint(e < count ? dist[e] : 0.0f);
    }
    w[15] = 0;
    uint4 *out = lists + 4 * (size_t)tile;
#pragma unroll
    for (int q = 0; q < 4; q++) out[q] = make_uint4(w[4 * q], w[4 * q + 1], w[4 * q + 2], w[4 * q + 3]);
}

}  // namespace rtd
