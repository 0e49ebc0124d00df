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

#include <cooperative_groups.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "device_scene.cuh"  // RT_BOX_CH: the record layout pack_built_nodes_kernel writes

namespace rtd {

constexpr int kBuildBlock = 256;     // one CTA per node: the many small nodes of the lower levels
constexpr int kBuildBigBlock = 512;  // CTAs of a cluster working on one large node
constexpr int kBuildCluster = 8;     // portable cluster size
constexpr int kBuildBigNode = 2048;  // nodes with at least this many triangles go to a cluster
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
    const int32_t *level_small, *level_big;  // node ids of this level: one CTA each / one cluster each
    int32_t *next_small, *next_big;          // node ids of the next level, by size class
    int32_t *counters;                       // [0] nodes allocated, [1] next level's small nodes, [2] its big nodes
    int32_t big_node_min;
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

// One node per CTA (CLUSTER == 1) or per thread-block cluster (CLUSTER CTAs, each owning a contiguous chunk of the node's
// range): the top levels hold a handful of nodes with 10^4-10^5 triangles each, and one CTA walking such a range was 70 % of
// the whole build. Partial bounds, bins and left-counts of the CTAs are combined through distributed shared memory; every
// CTA of the cluster then takes the same decision from the same merged data (min / max / integer sums: order-free), so
// nothing has to be broadcast back.
template <int CLUSTER>
__device__ __forceinline__ void group_sync()
{
    if (CLUSTER > 1) cooperative_groups::this_cluster().sync();
    else __syncthreads();
}
template <int CLUSTER, class T>
__device__ __forceinline__ const T *peer(const T *p, int rank)
{
    if (CLUSTER > 1) return cooperative_groups::this_cluster().map_shared_rank(const_cast<T *>(p), rank);
    return p;
}

struct TriData {  // one build-form triangle, three 16-byte loads
    float4 a, b, c;  // a = p1.xyz p2.x, b = p2.yz p3.xy, c = p3.z centroid.xyz
    __device__ __forceinline__ float lo(int k) const
    {
        return k == 0 ? fminf(a.x, fminf(a.w, b.z)) : k == 1 ? fminf(a.y, fminf(b.x, b.w)) : fminf(a.z, fminf(b.y, c.x));
    }
    __device__ __forceinline__ float hi(int k) const
    {
        return k == 0 ? fmaxf(a.x, fmaxf(a.w, b.z)) : k == 1 ? fmaxf(a.y, fmaxf(b.x, b.w)) : fmaxf(a.z, fmaxf(b.y, c.x));
    }
    __device__ __forceinline__ float centroid(int k) const { return k == 0 ? c.y : k == 1 ? c.z : c.w; }
};
__device__ __forceinline__ TriData load_tri(const BuildTri *tris, int id)
{
    const float4 *p = reinterpret_cast<const float4 *>(tris + id);
    return TriData{__ldg(p), __ldg(p + 1), __ldg(p + 2)};
}
__device__ __forceinline__ float load_centroid(const BuildTri *tris, int id, int axis) { return __ldg(&tris[id].p[3][axis]); }

template <int BLOCK, int CLUSTER>
__global__ void __launch_bounds__(BLOCK) bvh_level_kernel(const BuildArgs a)
{
    constexpr int WARPS = BLOCK / 32;
    __shared__ float red[WARPS][12];  // per-warp partial bounds
    __shared__ float part[12];        // this CTA's partial bounds: lo, hi, centroid lo, centroid hi (read by the peers)
    __shared__ float node_lo[3], node_hi[3], c_min[3], c_max[3];
    __shared__ int32_t bin_lo[WARPS][3][kBins][3], bin_hi[WARPS][3][kBins][3];  // ordered-int boxes per warp
    __shared__ int32_t bin_count[WARPS][3][kBins];
    __shared__ int32_t cta_lo[3][kBins][3], cta_hi[3][kBins][3], cta_count[3][kBins];  // merged over the warps (read by the peers)
    __shared__ float fin_lo[3][kBins][3], fin_hi[3][kBins][3];                          // merged over the cluster
    __shared__ int32_t fin_count[3][kBins];
    __shared__ int32_t split_axis, cta_left, left_total, left_before;
    __shared__ float split_position;
    __shared__ int32_t warp_left[WARPS];

    const int rank = CLUSTER > 1 ? (int)cooperative_groups::this_cluster().block_rank() : 0;
    const int node_id = (CLUSTER > 1 ? a.level_big : a.level_small)[blockIdx.x / CLUSTER];
    BuildNode &node = a.nodes[node_id];
    const int begin = node.child2, end = node.child1, count = end - begin;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // this CTA's chunk of the range: whole tiles of BLOCK triangles, in range order
    const int per = ((count + CLUSTER - 1) / CLUSTER + BLOCK - 1) / BLOCK * BLOCK;
    const int cb = min(end, begin + rank * per), ce = min(end, cb + per);

    // ---- A: node box over the vertices, centroid range per axis ----
    float lo[3] = {1e30f, 1e30f, 1e30f}, hi[3] = {-1e30f, -1e30f, -1e30f};
    float cl[3] = {1e30f, 1e30f, 1e30f}, ch[3] = {-1e30f, -1e30f, -1e30f};
    for (int i = cb + threadIdx.x; i < ce; i += BLOCK) {
        const TriData t = load_tri(a.tris, a.order_in[i]);
#pragma unroll
        for (int k = 0; k < 3; k++) {
            lo[k] = fminf(lo[k], t.lo(k));
            hi[k] = fmaxf(hi[k], t.hi(k));
            cl[k] = fminf(cl[k], t.centroid(k));
            ch[k] = fmaxf(ch[k], t.centroid(k));
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
    if (threadIdx.x < 12) {
        const int k = threadIdx.x;
        const bool is_min = k < 3 || (k >= 6 && k < 9);
        float v = red[0][k];
        for (int w = 1; w < WARPS; w++) v = is_min ? fminf(v, red[w][k]) : fmaxf(v, red[w][k]);
        part[k] = v;
    }
    group_sync<CLUSTER>();
    if (threadIdx.x < 12) {
        const int k = threadIdx.x;
        const bool is_min = k < 3 || (k >= 6 && k < 9);
        float v = part[k];
        for (int r = 0; r < CLUSTER; r++) {
            const float o = peer<CLUSTER>(part, r)[k];
            v = is_min ? fminf(v, o) : fmaxf(v, o);
        }
        if (k < 3) { node_lo[k] = v; if (rank == 0) node.lo[k] = v; }
        else if (k < 6) { node_hi[k - 3] = v; if (rank == 0) node.hi[k - 3] = v; }
        else if (k < 9) c_min[k - 6] = v;
        else c_max[k - 9] = v;
    }
    __syncthreads();

    bool leaf = count <= 4 || a.level >= a.max_depth;

    // ---- B: per-warp bins -> per-CTA bins -> bins of the whole node ----
    if (!leaf) {
        for (int i = threadIdx.x; i < WARPS * 3 * kBins; i += BLOCK) {
            int32_t *bl = &bin_lo[0][0][0][0] + 3 * i, *bh = &bin_hi[0][0][0][0] + 3 * i;
            bl[0] = bl[1] = bl[2] = float_to_ordered(1e30f);
            bh[0] = bh[1] = bh[2] = float_to_ordered(-1e30f);
            (&bin_count[0][0][0])[i] = 0;
        }
        __syncthreads();
        float to_bin[3];
#pragma unroll
        for (int k = 0; k < 3; k++) to_bin[k] = __fdiv_rn((float)kBins, __fsub_rn(c_max[k], c_min[k]));
        for (int i = cb + threadIdx.x; i < ce; i += BLOCK) {
            const TriData t = load_tri(a.tris, a.order_in[i]);
            int32_t vlo[3], vhi[3];
#pragma unroll
            for (int k = 0; k < 3; k++) {
                vlo[k] = float_to_ordered(t.lo(k));
                vhi[k] = float_to_ordered(t.hi(k));
            }
#pragma unroll
            for (int axis = 0; axis < 3; axis++) {
                if (c_min[axis] == c_max[axis]) continue;
                const int b = min(kBins - 1, (int)__fmul_rn(__fsub_rn(t.centroid(axis), c_min[axis]), to_bin[axis]));
                atomicAdd(&bin_count[warp][axis][b], 1);
#pragma unroll
                for (int k = 0; k < 3; k++) {
                    atomicMin(&bin_lo[warp][axis][b][k], vlo[k]);
                    atomicMax(&bin_hi[warp][axis][b][k], vhi[k]);
                }
            }
        }
        __syncthreads();
        if (threadIdx.x < 3 * kBins * 3) {  // (axis, bin, component): merge the warps of this CTA
            const int k = threadIdx.x % 3, ab = threadIdx.x / 3, axis = ab / kBins, b = ab % kBins;
            int32_t l = bin_lo[0][axis][b][k], h = bin_hi[0][axis][b][k], c = bin_count[0][axis][b];
            for (int w = 1; w < WARPS; w++) {
                l = min(l, bin_lo[w][axis][b][k]);
                h = max(h, bin_hi[w][axis][b][k]);
                c += bin_count[w][axis][b];
            }
            cta_lo[axis][b][k] = l;
            cta_hi[axis][b][k] = h;
            if (k == 0) cta_count[axis][b] = c;
        }
        group_sync<CLUSTER>();
        if (threadIdx.x < 3 * kBins * 3) {  // merge the CTAs of the cluster
            const int k = threadIdx.x % 3, ab = threadIdx.x / 3, axis = ab / kBins, b = ab % kBins;
            int32_t l = cta_lo[axis][b][k], h = cta_hi[axis][b][k], c = 0;
            for (int r = 0; r < CLUSTER; r++) {
                l = min(l, peer<CLUSTER>(&cta_lo[0][0][0], r)[threadIdx.x]);
                h = max(h, peer<CLUSTER>(&cta_hi[0][0][0], r)[threadIdx.x]);
                c += peer<CLUSTER>(&cta_count[0][0], r)[ab];
            }
            fin_lo[axis][b][k] = ordered_to_float(l);
            fin_hi[axis][b][k] = ordered_to_float(h);
            if (k == 0) fin_count[axis][b] = c;
        }
        __syncthreads();

        // ---- C: plane scan, sequential and in the host's order (scene.cu:905-957) ----
        if (threadIdx.x == 0) {
            const float leaf_cost = __fmul_rn(half_area_rn(node_lo, node_hi), (float)count);
            float best_cost = leaf_cost, best_position = 0.0f;
            int best_axis = -1;
            for (int axis = 0; axis < 3; axis++) {
                if (c_min[axis] == c_max[axis]) continue;
                float area_left[kBins - 1], area_right[kBins - 1];
                int count_left[kBins - 1];
                float sl_lo[3] = {1e30f, 1e30f, 1e30f}, sl_hi[3] = {-1e30f, -1e30f, -1e30f};
                float sr_lo[3] = {1e30f, 1e30f, 1e30f}, sr_hi[3] = {-1e30f, -1e30f, -1e30f};
                int running = 0;
                for (int k = 0; k + 1 < kBins; k++) {
                    running += fin_count[axis][k];
                    count_left[k] = running;
                    for (int c = 0; c < 3; c++) {
                        sl_lo[c] = fminf(sl_lo[c], fin_lo[axis][k][c]); sl_hi[c] = fmaxf(sl_hi[c], fin_hi[axis][k][c]);
                        sr_lo[c] = fminf(sr_lo[c], fin_lo[axis][kBins - 1 - k][c]); sr_hi[c] = fmaxf(sr_hi[c], fin_hi[axis][kBins - 1 - k][c]);
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
        for (int i = cb + threadIdx.x; i < ce; i += BLOCK) mine += load_centroid(a.tris, a.order_in[i], axis) < position ? 1 : 0;
        for (int off = 16; off > 0; off >>= 1) mine += __shfl_xor_sync(0xffffffffu, mine, off);
        if (lane == 0) warp_left[warp] = mine;
        __syncthreads();
        if (threadIdx.x == 0) {
            int total = 0;
            for (int w = 0; w < WARPS; w++) total += warp_left[w];
            cta_left = total;
        }
        group_sync<CLUSTER>();
        if (threadIdx.x == 0) {
            int total = 0, before = 0;
            for (int r = 0; r < CLUSTER; r++) {
                const int c = *peer<CLUSTER>(&cta_left, r);
                if (r < rank) before += c;
                total += c;
            }
            left_total = total;
            left_before = before;
        }
        __syncthreads();
        const int n_left = left_total;
        leaf = n_left == 0 || n_left == count;
        if (!leaf) {
            // tiles of BLOCK elements in range order; running offsets keep both sides in input order
            int left_done = left_before, right_done = (cb - begin) - left_before;
            for (int base = cb; base < ce; base += BLOCK) {
                const int i = base + threadIdx.x;
                const bool valid = i < ce;
                const int id = valid ? a.order_in[i] : 0;
                const bool go_left = valid && load_centroid(a.tris, id, axis) < position;
                const uint32_t ml = __ballot_sync(0xffffffffu, go_left), mv = __ballot_sync(0xffffffffu, valid);
                __syncthreads();  // warp_left of the previous tile has been read
                if (lane == 0) warp_left[warp] = __popc(ml) | (__popc(mv) << 16);
                __syncthreads();
                int l_before = 0, v_before = 0, l_tile = 0, v_tile = 0;
                for (int w = 0; w < WARPS; w++) {
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
            if (rank == 0 && threadIdx.x == 0) {
                const int pair = atomicAdd(&a.counters[0], 2);
                BuildNode &l = a.nodes[pair], &r = a.nodes[pair + 1];
                l.child2 = begin; l.child1 = begin + n_left;       // pending nodes carry their range, leaf convention
                r.child2 = begin + n_left; r.child1 = end;
                node.child1 = pair;
                node.child2 = pair + 1;
                const int counts[2] = {n_left, count - n_left};
                for (int c = 0; c < 2; c++) {
                    if (counts[c] >= a.big_node_min) a.next_big[atomicAdd(&a.counters[2], 1)] = pair + c;
                    else a.next_small[atomicAdd(&a.counters[1], 1)] = pair + c;
                }
            }
        }
    }
    // leaf: its range is final
    if (leaf)
        for (int i = cb + threadIdx.x; i < ce; i += BLOCK) a.order_final[i] = a.order_in[i];
    if (CLUSTER > 1) cooperative_groups::this_cluster().sync();  // a CTA's shared memory must outlive its peers' reads
}

__global__ void iota_kernel(int32_t *out, int32_t n)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = i;
}

// Reference-layout nodes -> the tracer's 64-byte two-child records (device_scene.cuh), one thread per internal node.
// Children are allocated in adjacent pairs (2k+1, 2k+2), and every pair belongs to exactly one internal node, so the pair
// number k is a dense record index: record(node) = (node.child1 - 1) / 2, the root's record is 0. A leaf of more than
// `max_leaf` triangles (depth limit or no useful plane) needs a chain of records: *needs_host is raised and the host
// packer takes over.
__global__ void pack_built_nodes_kernel(const BuildNode *__restrict__ nodes, int node_count, float4 *__restrict__ records,
                                        uint32_t leaf_bit, int count_shift, int max_leaf, int32_t *needs_host)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= node_count) return;
    const BuildNode n = nodes[i];
    if (n.child2 <= n.child1) return;  // a leaf lives in its parent's record
    float4 *rec = records + 4 * (size_t)((n.child1 - 1) >> 1);
#pragma unroll
    for (int c = 0; c < 2; c++) {
        const BuildNode ch = nodes[n.child1 + c];
        uint32_t ref;
        if (ch.child2 > ch.child1) {
            ref = (uint32_t)((ch.child1 - 1) >> 1);
        } else {
            const int count = ch.child1 - ch.child2;
            if (count > max_leaf || count <= 0) { *needs_host = 1; return; }
            ref = leaf_bit | ((uint32_t)count << count_shift) | (uint32_t)ch.child2;
        }
#if RT_BOX_CH  // centre + half-extent rounded up (same arithmetic as the host packer's set_box, rt_core.cu)
        float cc[3], hh[3];
#pragma unroll
        for (int k = 0; k < 3; k++) {
            cc[k] = __fadd_rn(__fmul_rn(0.5f, ch.lo[k]), __fmul_rn(0.5f, ch.hi[k]));
            hh[k] = fmaxf(__fsub_ru(ch.hi[k], cc[k]), __fsub_ru(cc[k], ch.lo[k]));
        }
        rec[2 * c] = make_float4(cc[0], cc[1], cc[2], hh[0]);
        rec[2 * c + 1] = make_float4(hh[1], hh[2], __uint_as_float(ref), 0.0f);
#else
        rec[2 * c] = make_float4(ch.lo[0], ch.lo[1], ch.lo[2], ch.hi[0]);
        rec[2 * c + 1] = make_float4(ch.hi[1], ch.hi[2], __uint_as_float(ref), 0.0f);
#endif
    }
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
