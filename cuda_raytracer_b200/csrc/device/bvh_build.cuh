// Binned-SAH BVH build on the device: the decisions of the reference builder
// (BvhNode::maybe_split, /root/reference/scene.cu:866-1000; Scene::generate_bvh :1002-1036) taken level by level, one
// CTA per node of the level.
//
// Same rules as the host builder (csrc/host/bvh_builder.cpp, which is pinned byte for byte to the reference):
//   * leaf if count <= 4 or the depth limit is reached;
//   * per axis 8 bins over the centroid range, bin = min(7, (int)((c - c_min) * (8 / (c_max - c_min))));
//   * cost of a plane = n_left * half_area(left) + n_right * half_area(right), planes scanned axis-major in ascending
//     position, the first strictly cheaper one wins; an empty side yields 0 * inf = NaN and loses every compare (Q14);
//   * leaf if no plane beats count * half_area(node);
//   * partition by `centroid[axis] < c_min + bin_width * (k + 1)` (NOT by bin index), leaf if one side is empty.
// Every float expression is written with explicit round-to-nearest intrinsics in the host's evaluation order, so the
// tree is the same SET of nodes (boxes, triangle sets, left/right roles) as the host's; node numbering (breadth-first,
// atomically allocated pairs) and the triangle order inside a leaf (file order, the partition is stable) differ, which
// closest-hit results cannot see. tests/test_gpu_bvh_build.py checks both statements.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace rtd {

constexpr int kBuildBlock = 256;
constexpr int kBuildWarps = kBuildBlock / 32;
constexpr int kBins = 8;

struct BuildNode {  // the reference's 32-byte node (scene.cuh:82-100)
    float lo[3], hi[3];
    int32_t child1;  // internal: left child; leaf: END of its triangle range
    int32_t child2;  // internal: right child (= child1 + 1); leaf: START of its triangle range
};
static_assert(sizeof(BuildNode) == 32, "reference node layout");

struct BuildTri {  // build form: three vertices + centroid (scene.cu:520-531)
    float p[4][3];
};
static_assert(sizeof(BuildTri) == 48, "build-form triangle");

struct BuildArgs {
    const BuildTri *tris;
    BuildNode *nodes;
    const int32_t *order_in;   // triangle ids of the ranges of this level's nodes
    int32_t *order_out;        // partitions of the nodes that split
    int32_t *order_final;      // ranges of the nodes that became leaves
    const int32_t *level_nodes;  // node ids of this level
    int32_t *next_nodes;         // node ids of the next level
    int32_t *counters;           // [0] nodes allocated, [1] size of the next level
    int32_t level;
    int32_t max_depth;
};

// float min / max through integer atomics in shared memory (monotone mapping of the IEEE bit pattern)
__device__ __forceinline__ int32_t float_to_ordered(float f)
{
    int32_t i = __float_as_int(f);
    return i >= 0 ? i : i ^ 0x7FFFFFFF;
}
__device__ __forceinline__ float ordered_to_float(int32_t i) { return __int_as_float(i >= 0 ? i : i ^ 0x7FFFFFFF); }

__device__ __forceinline__ float half_area_rn(const float *lo, const float *hi)  // Box::half_area, host evaluation order
{
    const float dx = __fsub_rn(hi[0], lo[0]), dy = __fsub_rn(hi[1], lo[1]), dz = __fsub_rn(hi[2], lo[2]);
    return __fadd_rn(__fadd_rn(__fmul_rn(dx, dy), __fmul_rn(dx, dz)), __fmul_rn(dy, dz));
}

__device__ __forceinline__ float warp_min(float v)
{
    for (int off = 16; off > 0; off >>= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, off));
    return v;
}
__device__ __forceinline__ float warp_max(float v)
{
    for (int off = 16; off > 0; off >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, off));
    return v;
}

__global__ void __launch_bounds__(kBuildBlock) bvh_level_kernel(const BuildArgs a)
{
    __shared__ float red[kBuildWarps][12];                   // per-warp partial bounds
    __shared__ float node_lo[3], node_hi[3], c_min[3], c_max[3];
    __shared__ int32_t bin_lo[kBuildWarps][3][kBins][3], bin_hi[kBuildWarps][3][kBins][3];  // ordered-int boxes per warp
    __shared__ int32_t bin_count[kBuildWarps][3][kBins];
    __shared__ int32_t split_axis, left_total;
    __shared__ float split_position;
    __shared__ int32_t warp_left[kBuildWarps];

    const int node_id = a.level_nodes[blockIdx.x];
    BuildNode &node = a.nodes[node_id];
    const int begin = node.child2, end = node.child1, count = end - begin;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;

    // ---- A: node box over the vertices, centroid range per axis ----
    float lo[3] = {1e30f, 1e30f, 1e30f}, hi[3] = {-1e30f, -1e30f, -1e30f};
    float cl[3] = {1e30f, 1e30f, 1e30f}, ch[3] = {-1e30f, -1e30f, -1e30f};
    for (int i = begin + threadIdx.x; i < end; i += kBuildBlock) {
        const BuildTri &t = a.tris[a.order_in[i]];
#pragma unroll
        for (int k = 0; k < 3; k++) {
            lo[k] = fminf(lo[k], fminf(t.p[0][k], fminf(t.p[1][k], t.p[2][k])));
            hi[k] = fmaxf(hi[k], fmaxf(t.p[0][k], fmaxf(t.p[1][k], t.p[2][k])));
            cl[k] = fminf(cl[k], t.p[3][k]);
            ch[k] = fmaxf(ch[k], t.p[3][k]);
        }
    }
#pragma unroll
    for (int k = 0; k < 3; k++) {
        lo[k] = warp_min(lo[k]); hi[k] = warp_max(hi[k]);
        cl[k] = warp_min(cl[k]); ch[k] = warp_max(ch[k]);
    }
    if (lane == 0) {
#pragma unroll
        for (int k = 0; k < 3; k++) { red[warp][k] = lo[k]; red[warp][3 + k] = hi[k]; red[warp][6 + k] = cl[k]; red[warp][9 + k] = ch[k]; }
    }
    __syncthreads();
    if (threadIdx.x < 3) {
        const int k = threadIdx.x;
        float l = red[0][k], h = red[0][3 + k], c0 = red[0][6 + k], c1 = red[0][9 + k];
        for (int w = 1; w < kBuildWarps; w++) {
            l = fminf(l, red[w][k]); h = fmaxf(h, red[w][3 + k]);
            c0 = fminf(c0, red[w][6 + k]); c1 = fmaxf(c1, red[w][9 + k]);
        }
        node_lo[k] = l; node_hi[k] = h; c_min[k] = c0; c_max[k] = c1;
        node.lo[k] = l; node.hi[k] = h;
    }
    __syncthreads();

    bool leaf = count <= 4 || a.level >= a.max_depth;

    // ---- B: per-warp bins, merged by the plane scan ----
    if (!leaf) {
        for (int i = threadIdx.x; i < kBuildWarps * 3 * kBins; i += kBuildBlock) {
            int32_t *bl = &bin_lo[0][0][0][0] + 3 * i, *bh = &bin_hi[0][0][0][0] + 3 * i;
            bl[0] = bl[1] = bl[2] = float_to_ordered(1e30f);
            bh[0] = bh[1] = bh[2] = float_to_ordered(-1e30f);
            (&bin_count[0][0][0])[i] = 0;
        }
        __syncthreads();
        float to_bin[3];
#pragma unroll
        for (int k = 0; k < 3; k++) to_bin[k] = __fdiv_rn((float)kBins, __fsub_rn(c_max[k], c_min[k]));
        for (int i = begin + threadIdx.x; i < end; i += kBuildBlock) {
            const BuildTri &t = a.tris[a.order_in[i]];
            int32_t vlo[3], vhi[3];
#pragma unroll
            for (int k = 0; k < 3; k++) {
                vlo[k] = float_to_ordered(fminf(t.p[0][k], fminf(t.p[1][k], t.p[2][k])));
                vhi[k] = float_to_ordered(fmaxf(t.p[0][k], fmaxf(t.p[1][k], t.p[2][k])));
            }
#pragma unroll
            for (int axis = 0; axis < 3; axis++) {
                if (c_min[axis] == c_max[axis]) continue;
                const int b = min(kBins - 1, (int)__fmul_rn(__fsub_rn(t.p[3][axis], c_min[axis]), to_bin[axis]));
                atomicAdd(&bin_count[warp][axis][b], 1);
#pragma unroll
                for (int k = 0; k < 3; k++) {
                    atomicMin(&bin_lo[warp][axis][b][k], vlo[k]);
                    atomicMax(&bin_hi[warp][axis][b][k], vhi[k]);
                }
            }
        }
        __syncthreads();

        // ---- C: plane scan, sequential and in the host's order (scene.cu:905-957) ----
        if (threadIdx.x == 0) {
            const float leaf_cost = __fmul_rn(half_area_rn(node_lo, node_hi), (float)count);
            float best_cost = leaf_cost, best_position = 0.0f;
            int best_axis = -1;
            for (int axis = 0; axis < 3; axis++) {
                if (c_min[axis] == c_max[axis]) continue;
                float blo[kBins][3], bhi[kBins][3];
                int bcount[kBins];
                for (int b = 0; b < kBins; b++) {
                    int c = 0;
                    int32_t l[3] = {bin_lo[0][axis][b][0], bin_lo[0][axis][b][1], bin_lo[0][axis][b][2]};
                    int32_t h[3] = {bin_hi[0][axis][b][0], bin_hi[0][axis][b][1], bin_hi[0][axis][b][2]};
                    for (int w = 0; w < kBuildWarps; w++) {
                        c += bin_count[w][axis][b];
                        for (int k = 0; k < 3; k++) { l[k] = min(l[k], bin_lo[w][axis][b][k]); h[k] = max(h[k], bin_hi[w][axis][b][k]); }
                    }
                    bcount[b] = c;
                    for (int k = 0; k < 3; k++) { blo[b][k] = ordered_to_float(l[k]); bhi[b][k] = ordered_to_float(h[k]); }
                }
                float area_left[kBins - 1], area_right[kBins - 1];
                int count_left[kBins - 1];
                float sl_lo[3] = {1e30f, 1e30f, 1e30f}, sl_hi[3] = {-1e30f, -1e30f, -1e30f};
                float sr_lo[3] = {1e30f, 1e30f, 1e30f}, sr_hi[3] = {-1e30f, -1e30f, -1e30f};
                int running = 0;
                for (int k = 0; k + 1 < kBins; k++) {
                    running += bcount[k];
                    count_left[k] = running;
                    for (int c = 0; c < 3; c++) {
                        sl_lo[c] = fminf(sl_lo[c], blo[k][c]); sl_hi[c] = fmaxf(sl_hi[c], bhi[k][c]);
                        sr_lo[c] = fminf(sr_lo[c], blo[kBins - 1 - k][c]); sr_hi[c] = fmaxf(sr_hi[c], bhi[kBins - 1 - k][c]);
                    }
                    area_left[k] = half_area_rn(sl_lo, sl_hi);
                    area_right[kBins - 2 - k] = half_area_rn(sr_lo, sr_hi);
                }
                const float bin_width = __fdiv_rn(__fsub_rn(c_max[axis], c_min[axis]), (float)kBins);
                for (int k = 0; k + 1 < kBins; k++) {
                    const float cost = __fadd_rn(__fmul_rn((float)count_left[k], area_left[k]),
                                                 __fmul_rn((float)(count - count_left[k]), area_right[k]));
                    if (cost != 0.0f && cost < best_cost) {
                        best_axis = axis;
                        best_position = __fadd_rn(c_min[axis], __fmul_rn(bin_width, (float)(k + 1)));
                        best_cost = cost;
                    }
                }
            }
            split_axis = (best_axis < 0 || best_cost >= leaf_cost) ? -1 : best_axis;
            split_position = best_position;
        }
        __syncthreads();
        leaf = split_axis < 0;
    }

    // ---- D: stable partition by the plane; an empty side makes the node a leaf after all ----
    if (!leaf) {
        const int axis = split_axis;
        const float position = split_position;
        int mine = 0;
        for (int i = begin + threadIdx.x; i < end; i += kBuildBlock) mine += a.tris[a.order_in[i]].p[3][axis] < position ? 1 : 0;
        for (int off = 16; off > 0; off >>= 1) mine += __shfl_xor_sync(0xffffffffu, mine, off);
        if (lane == 0) warp_left[warp] = mine;
        __syncthreads();
        if (threadIdx.x == 0) {
            int total = 0;
            for (int w = 0; w < kBuildWarps; w++) total += warp_left[w];
            left_total = total;
        }
        __syncthreads();
        const int n_left = left_total;
        leaf = n_left == 0 || n_left == count;
        if (!leaf) {
            // tiles of kBuildBlock elements in order; running offsets keep both sides in input order
            int left_done = 0, right_done = 0;
            for (int base = begin; base < end; base += kBuildBlock) {
                const int i = base + threadIdx.x;
                const bool valid = i < end;
                const int id = valid ? a.order_in[i] : 0;
                const bool go_left = valid && a.tris[id].p[3][axis] < position;
                const uint32_t ml = __ballot_sync(0xffffffffu, go_left), mv = __ballot_sync(0xffffffffu, valid);
                __syncthreads();  // warp_left of the previous tile has been read
                if (lane == 0) warp_left[warp] = __popc(ml) | (__popc(mv) << 16);
                __syncthreads();
                int l_before = 0, v_before = 0, l_tile = 0, v_tile = 0;
                for (int w = 0; w < kBuildWarps; w++) {
                    const int l = warp_left[w] & 0xFFFF, v = warp_left[w] >> 16;
                    if (w < warp) { l_before += l; v_before += v; }
                    l_tile += l; v_tile += v;
                }
                if (valid) {
                    const uint32_t lt = (1u << lane) - 1u;
                    const int l_rank = l_before + __popc(ml & lt);
                    const int r_rank = (v_before - l_before) + __popc(mv & ~ml & lt);
                    a.order_out[go_left ? begin + left_done + l_rank : begin + n_left + right_done + r_rank] = id;
                }
                left_done += l_tile;
                right_done += v_tile - l_tile;
            }
            if (threadIdx.x == 0) {
                const int pair = atomicAdd(&a.counters[0], 2);
                BuildNode &l = a.nodes[pair], &r = a.nodes[pair + 1];
                l.child2 = begin; l.child1 = begin + n_left;       // pending nodes carry their range, leaf convention
                r.child2 = begin + n_left; r.child1 = end;
                node.child1 = pair;
                node.child2 = pair + 1;
                const int slot = atomicAdd(&a.counters[1], 2);
                a.next_nodes[slot] = pair;
                a.next_nodes[slot + 1] = pair + 1;
            }
            return;
        }
    }
    // leaf: its range is final
    for (int i = begin + threadIdx.x; i < end; i += kBuildBlock) a.order_final[i] = a.order_in[i];
}

__global__ void iota_kernel(int32_t *out, int32_t n)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = i;
}

// Build form, file order -> 64-byte render-form records in BVH order (Scene::generate_bvh's tail, scene.cu:1029-1035:
// e1 = p2 - p1, e2 = p3 - p1, n = normalise(cross(e2, e1))), arithmetic in the host's order without contraction.
__global__ void pack_built_triangles_kernel(const BuildTri *__restrict__ tris, const int32_t *__restrict__ order,
                                            const uint16_t *__restrict__ tri_materials, float4 *__restrict__ out, uint32_t n)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int id = order[i];
    const BuildTri &t = tris[id];
    float e1[3], e2[3], c[3];
#pragma unroll
    for (int k = 0; k < 3; k++) { e1[k] = __fsub_rn(t.p[1][k], t.p[0][k]); e2[k] = __fsub_rn(t.p[2][k], t.p[0][k]); }
    c[0] = __fsub_rn(__fmul_rn(e2[1], e1[2]), __fmul_rn(e2[2], e1[1]));
    c[1] = __fsub_rn(__fmul_rn(e2[2], e1[0]), __fmul_rn(e2[0], e1[2]));
    c[2] = __fsub_rn(__fmul_rn(e2[0], e1[1]), __fmul_rn(e2[1], e1[0]));
    const float len = __fsqrt_rn(__fadd_rn(__fadd_rn(__fmul_rn(c[0], c[0]), __fmul_rn(c[1], c[1])), __fmul_rn(c[2], c[2])));
    const float inv = __fdiv_rn(1.0f, len);
    float4 *o = out + 4 * (size_t)i;
    o[0] = make_float4(t.p[0][0], t.p[0][1], t.p[0][2], e1[0]);
    o[1] = make_float4(e1[1], e1[2], e2[0], e2[1]);
    o[2] = make_float4(e2[2], __fmul_rn(inv, c[0]), __fmul_rn(inv, c[1]), __fmul_rn(inv, c[2]));
    o[3] = make_float4(__uint_as_float((uint32_t)tri_materials[id]), __int_as_float(id), 0.0f, 0.0f);
}

}  // namespace rtd
