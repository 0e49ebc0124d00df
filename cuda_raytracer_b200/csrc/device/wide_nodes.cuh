// Compressed node formats of the tracer (SURVEY §8f N3 "wider / compressed BVH"; layouts in device_scene.cuh):
//   * quantize_records_kernel: the 64-byte two-child records -> 32-byte records on a 15-bit grid (kNodesQ16), on the device;
//   * WideCollapse: the same records collapsed into 4-wide nodes with quantised boxes (kNodesW4), on the host.
// Both keep the binary tree the records came from (the reference's, scene.cu:866-1036, or the device build's): boxes are
// only ever enlarged and the children of a wide node are nodes of that tree, so the set of triangles a ray can be tested
// against is a superset of the exact walk's and closest hits come from the same exact triangle test.
#pragma once

#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>

#include "device_scene.cuh"

namespace rtd {

// Grid: coordinate(q) = q_bias + (1 + q / 32768) * q_scale for q in [0, 32767], in EXACT arithmetic on the two float
// constants (the tracer evaluates it as one FFMA with per-ray constants, traverse.cuh). kQuantSlack cells are left free on
// both sides of the root box so that the outward rounding + one cell of margin never clamps.
constexpr int kQuantSlack = 8;
struct QuantGrid {
    float scale[3], bias[3];
    bool ok;  // false: the root box does not fit the grid with its margin (coordinates far larger than the extent)
};

__host__ __device__ inline double quant_coord(float bias, float scale, int q) { return (double)bias + (1.0 + (double)q / 32768.0) * (double)scale; }

inline QuantGrid make_quant_grid(const float *root_lo, const float *root_hi)
{
    QuantGrid g{};
    g.ok = true;
    for (int k = 0; k < 3; k++) {
        double lo = root_lo[k], hi = root_hi[k];
        if (!(hi >= lo) || !std::isfinite(lo) || !std::isfinite(hi)) { lo = 0.0; hi = 1.0; }  // no triangles: any grid will do
        const double extent = std::max({hi - lo, 1e-5 * std::max(std::fabs(lo), std::fabs(hi)), 1e-20});
        const double cell = extent / (32767.0 - 2.0 * kQuantSlack);
        g.scale[k] = (float)(32768.0 * cell);
        g.bias[k] = (float)(lo - kQuantSlack * cell - (double)g.scale[k]);
        const double c = (double)g.scale[k] / 32768.0;
        if (!(quant_coord(g.bias[k], g.scale[k], 0) <= lo - 2.0 * c) || !(quant_coord(g.bias[k], g.scale[k], 32767) >= hi + 2.0 * c)) g.ok = false;
    }
    return g;
}

// {lo | hi << 16} word of one axis: lo rounded down and hi rounded up on the grid, one more cell outwards, top bits set.
__host__ __device__ inline uint32_t quant_word(float lo, float hi, float bias, float scale)
{
    const double inv = 32768.0 / (double)scale;
    double ql = floor(((double)lo - (double)bias) * inv - 32768.0) - 1.0;
    double qh = ceil(((double)hi - (double)bias) * inv - 32768.0) + 1.0;
    ql = ql < 0.0 ? 0.0 : (ql > 32767.0 ? 32767.0 : ql);
    qh = qh < 0.0 ? 0.0 : (qh > 32767.0 ? 32767.0 : qh);
    return (0x8000u | (uint32_t)ql) | ((0x8000u | (uint32_t)qh) << 16);
}
constexpr uint32_t kQuantEmptyWord = 0x8000FFFFu;  // lo = 32767, hi = 0: fails the slab test for every ray

// bounds of child c of a 64-byte record (4 x float4), rounded outwards
__host__ __device__ inline void record_child_box(const float *rec16, int c, float *lo, float *hi)
{
    const float *b = rec16 + 8 * c;
#if RT_BOX_CH
    for (int k = 0; k < 3; k++) {
        const double l = (double)b[k] - (double)b[3 + k], h = (double)b[k] + (double)b[3 + k];
        float fl = (float)l, fh = (float)h;
#ifdef __CUDA_ARCH__
        fl = __double2float_rd(l);
        fh = __double2float_ru(h);
#else
        if ((double)fl > l) fl = std::nextafterf(fl, -INFINITY);
        if ((double)fh < h) fh = std::nextafterf(fh, INFINITY);
#endif
        lo[k] = fl;
        hi[k] = fh;
    }
#else
    for (int k = 0; k < 3; k++) { lo[k] = b[k]; hi[k] = b[3 + k]; }
#endif
}

#ifdef __CUDACC__
__global__ void quantize_records_kernel(const float4 *__restrict__ records, int record_count, uint4 *__restrict__ out, QuantGrid g)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= record_count) return;
    float rec[16];
    for (int q = 0; q < 4; q++) {
        const float4 v = records[4 * (size_t)i + q];
        rec[4 * q] = v.x; rec[4 * q + 1] = v.y; rec[4 * q + 2] = v.z; rec[4 * q + 3] = v.w;
    }
    uint32_t w[8];
    for (int c = 0; c < 2; c++) {
        const uint32_t ref = __float_as_uint(rec[8 * c + 6]);
        w[6 + c] = ref;
        if (ref == kEmptyRef) {
            w[3 * c] = w[3 * c + 1] = w[3 * c + 2] = kQuantEmptyWord;
            continue;
        }
        float lo[3], hi[3];
        record_child_box(rec, c, lo, hi);
        for (int k = 0; k < 3; k++) w[3 * c + k] = quant_word(lo[k], hi[k], g.bias[k], g.scale[k]);
    }
    out[2 * (size_t)i] = make_uint4(w[0], w[1], w[2], w[3]);
    out[2 * (size_t)i + 1] = make_uint4(w[4], w[5], w[6], w[7]);
}
#endif

// Host: binary records -> 4-wide nodes. A node starts with the two children of a record; while it has fewer than four
// children, the internal child with the largest surface area is replaced by its own two children. Nodes are numbered in
// depth-first order (node 0 = the root). 16 words per node: children 0,1 {x,y,z words} + refs 0,1 | children 2,3 + refs 2,3.
struct WideCollapse {
    const float *records;  // 16 floats per binary record
    QuantGrid grid;
    std::vector<uint32_t> out;
    int stack_need = 0;    // most stack entries any root-to-leaf walk can hold
    int depth = 0;

    struct Child {
        float lo[3], hi[3];
        uint32_t ref;
    };
    static double area(const Child &c)
    {
        const double dx = (double)c.hi[0] - c.lo[0], dy = (double)c.hi[1] - c.lo[1], dz = (double)c.hi[2] - c.lo[2];
        return dx * dy + dy * dz + dz * dx;
    }
    void children_of(uint32_t record, std::vector<Child> &into) const
    {
        const float *rec = records + 16 * (size_t)record;
        for (int c = 0; c < 2; c++) {
            Child ch;
            std::memcpy(&ch.ref, rec + 8 * c + 6, 4);
            if (ch.ref == kEmptyRef) continue;
            record_child_box(rec, c, ch.lo, ch.hi);
            into.push_back(ch);
        }
    }
    // returns the stack need below (and including) this node
    int emit(uint32_t record, int level)
    {
        depth = std::max(depth, level + 1);
        std::vector<Child> ch;
        children_of(record, ch);
        while (ch.size() < 4) {
            int best = -1;
            double best_area = -1.0;
            for (size_t i = 0; i < ch.size(); i++)
                if (!(ch[i].ref & kLeafBit) && area(ch[i]) > best_area) { best = (int)i; best_area = area(ch[i]); }
            if (best < 0) break;
            const uint32_t r = ch[best].ref;
            ch.erase(ch.begin() + best);
            children_of(r, ch);
        }
        const size_t at = out.size();
        out.resize(at + 16);
        int need = (int)ch.size() - 1, below = 0;
        for (int i = 0; i < 4; i++) {
            uint32_t w[3] = {kQuantEmptyWord, kQuantEmptyWord, kQuantEmptyWord}, ref = kEmptyRef;
            if (i < (int)ch.size()) {
                for (int k = 0; k < 3; k++) w[k] = quant_word(ch[i].lo[k], ch[i].hi[k], grid.bias[k], grid.scale[k]);
                ref = ch[i].ref;
                if (!(ref & kLeafBit)) {
                    ref = (uint32_t)(out.size() / 16);
                    below = std::max(below, emit(ch[i].ref, level + 1));
                }
            }
            uint32_t *node = out.data() + at;  // emit() may have reallocated
            const int half = i >> 1, slot = i & 1;
            node[8 * half + 3 * slot] = w[0];
            node[8 * half + 3 * slot + 1] = w[1];
            node[8 * half + 3 * slot + 2] = w[2];
            node[8 * half + 6 + slot] = ref;
        }
        need = std::max(0, need) + below;
        stack_need = std::max(stack_need, need);
        return need;
    }
    void run()
    {
        out.clear();
        stack_need = depth = 0;
        emit(0, 0);
    }
};

}  // namespace rtd
