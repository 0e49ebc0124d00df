// librtcore device side: context, scene packing/upload, the pass loop and the C ABI of include/rt_core.h.
// Replaces gpu_raytrace and its helpers (/root/reference/raytracing.cu:170-284, scene.cu:243-278).
#include <cuda_runtime.h>

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <vector>

#include "../../../include/rt_core.h"
#include "bvh_build.cuh"
#include "kernels.cuh"
#include "radix_sort.cuh"
#include "wave_plan.cuh"
#include "wide_nodes.cuh"

namespace rt {
void set_global_error(const std::string &msg);  // scene_api.cpp
const char *global_error();
}  // namespace rt

using namespace rtd;

struct rt_ctx {
    int device = 0;
    int sm_count = 0;
    cudaStream_t stream = nullptr;
    static constexpr int kMaxPipes = 4;
    cudaStream_t side[kMaxPipes - 1] = {};  // further pipes of the pass loop: wave w runs on pipe w % pipes (render_impl)
    cudaEvent_t ev_fork = nullptr, ev_join[kMaxPipes - 1] = {};
    int pipes = 0;                   // 0 = automatic (render_impl); RT_PIPES=1: one wave at a time, 2..4: that many
    int trace_per_sm = 0;            // RT_TRACE_PER_SM: CTAs per SM of the persistent tracer grid (0 = full occupancy)
    int shade_per_sm = 0;            // RT_SHADE_PER_SM: CTAs per SM of the shade grid (0 = 2048 threads' worth)
    std::string error;

    // scene
    bool has_scene = false;
    DeviceScene ds{};
    void *d_nodes = nullptr, *d_tris = nullptr, *d_materials = nullptr, *d_spheres = nullptr, *d_env = nullptr, *d_matidx = nullptr;
    size_t cap_nodes = 0, cap_tris = 0, cap_materials = 0, cap_spheres = 0, cap_env = 0, cap_matidx = 0, cap_raw = 0;
    void *d_raw = nullptr;    // upload scratch: the caller's raw triangles + texels before the device-side repack
    void *d_build = nullptr;  // device BVH build scratch: nodes, order buffers, level lists, counters
    size_t cap_build = 0;
    int32_t *d_order = nullptr;  // device build: BVH position -> file-order triangle id (inside d_build)
    std::vector<uint8_t> built_nodes;  // device build: the reference-layout nodes it produced (32 B each), fetched on demand
    int built_node_count = 0, built_levels = 0;
    bool device_packed = false;  // the records in d_nodes were written by pack_built_nodes_kernel
    float built_root[8] = {};    // the root node of a device-packed tree (box for the reorder key)
    double bvh_device_ms = 0;
    void *d_entry = nullptr;   // per-tile entry records of the primary rays (entry_lists.cuh)
    size_t cap_entry = 0;
    bool entry_dirty = true;   // scene or camera changed since the records were cut
    int entry_max = kEntryMax; // RT_ENTRY_K: entries per tile, 0 = primary rays start at the root
    double sort_device_ms = 0;
    int ray_count = 0, bounces = 0;
    float exposure = 0;
    int packed_nodes = 0, packed_depth = 0;

    // render state
    float *d_fb = nullptr, *d_fb_tmp = nullptr, *d_fb_tmp2 = nullptr;
    uint8_t *d_rgb8 = nullptr;
    size_t fb_pixels = 0;
    RayQueue queue[2] = {};
    float2 *d_hits = nullptr;  // {t, prim} per queue position of the bounce in flight
    size_t queue_capacity = 0;
    int node_mode = kNodesF32; // node format the tracer walks (rt_set_node_mode, environment RT_NODES = f32 | q16 | w4)
    int active_nodes = kNodesF32;  // format of the uploaded scene (falls back to f32 when the grid cannot hold the scene)
    void *d_qnodes = nullptr;  // compressed nodes (kNodesQ16: 32 B per record; kNodesW4: 64 B per wide node)
    size_t cap_qnodes = 0;
    int wide_nodes = 0, wide_depth = 0, wide_stack_need = 0;
    uint32_t trace_batch = kBatch, trace_refill_below = kRefillBelow, trace_node_min = kNodeMin;  // RT_BATCH / RT_REFILL override
    uint32_t *d_keys[2] = {nullptr, nullptr}, *d_vals[2] = {nullptr, nullptr};
    size_t sort_capacity = 0;
    uint32_t tail_below = 0;   // RT_TAIL: tail_kernel takes a wave over when at most this many rays enter a bounce (0 = off)
    int sort_rank = 2, sort_items = kSortItems;  // RT_SORT_RANK / RT_SORT_ITEMS: ranking variant and keys per thread of the digit pass
    uint32_t *d_sort_scratch = nullptr;  // hist + tile counters + tile status
    size_t sort_scratch_words = 0;
    PassCounters *d_pc = nullptr;
    RenderTotals *d_totals = nullptr;
    cudaEvent_t ev_start = nullptr, ev_stop = nullptr;
    std::vector<cudaEvent_t> ev_pool;  // per-launch timing (rt_render_opts.kernel_timing)
};

namespace {

int fail(rt_ctx *ctx, int status, const std::string &msg)
{
    if (ctx) ctx->error = msg;
    rt::set_global_error(msg);
    return status;
}

#define RT_CUDA(ctx, call)                                                                                   \
    do {                                                                                                     \
        cudaError_t e_ = (call);                                                                             \
        if (e_ != cudaSuccess)                                                                               \
            return fail(ctx, e_ == cudaErrorMemoryAllocation ? RT_ERR_NOMEM : RT_ERR_CUDA,                   \
                        std::string(#call) + ": " + cudaGetErrorString(e_));                                 \
    } while (0)

template <typename T>
void release(T *&p)
{
    if (p) cudaFree(p);
    p = nullptr;
}

// ---- BVH packing: reference 32-byte nodes -> 64-byte two-child records (layout in device_scene.cuh) --------------
struct RefNode {  // scene.cuh:82-100
    float lo[3], hi[3];
    int32_t child1, child2;
};
static_assert(sizeof(RefNode) == 32, "reference node layout");

struct PackedHalf {  // one child: box + reference, 32 bytes = one LDG.256
    float box[6];
    uint32_t ref;
    uint32_t pad;
};
struct PackedRecord {
    PackedHalf child[2];
};
static_assert(sizeof(PackedRecord) == 64, "packed record");

struct Packer {
    const RefNode *nodes;
    std::vector<PackedRecord> out;
    int max_depth = 0;  // most stack entries a walk can hold = levels of tree records (+ 1 below an oversized leaf's list)

    static void set_box(PackedRecord &r, int child, const float *lo, const float *hi)
    {
        float *b = r.child[child].box;
#if RT_BOX_CH  // centre + half-extent rounded up: differences of floats are exact in double, so this is __fsub_ru
        for (int k = 0; k < 3; k++) {
            const float c = 0.5f * lo[k] + 0.5f * hi[k];
            const double hd = std::max((double)hi[k] - (double)c, (double)c - (double)lo[k]);
            float h = (float)hd;
            if ((double)h < hd) h = std::nextafterf(h, INFINITY);
            b[k] = c;
            b[3 + k] = h;
        }
#else
        b[0] = lo[0]; b[1] = lo[1]; b[2] = lo[2];
        b[3] = hi[0]; b[4] = hi[1]; b[5] = hi[2];
#endif
    }
    static void set_empty(PackedRecord &r, int child)
    {
        const float far_away[3] = {3e38f, 3e38f, 3e38f};
        set_box(r, child, far_away, far_away);
        r.child[child].ref = kEmptyRef;
    }
    // triangles [begin, end) under box lo/hi -> a leaf ref, or, when there are > 15, a LIST of same-box records: child 1 = a
    // leaf of 15 triangles, child 0 = the rest of the list. Both boxes are equal, so both are hit at the same entry distance and
    // the tracer (ties go to child 1) tests the leaf first with the rest as the ONE pending stack entry, which it pops before it
    // goes any deeper: a list of any length needs one stack entry more than its parent (a balanced chain needed log2(n / 15)).
    uint32_t range_ref(int begin, int end, const float *lo, const float *hi, int depth)
    {
        int count = end - begin;
        if (count <= 0) return kEmptyRef;
        if (count <= kMaxLeafTris) return kLeafBit | ((uint32_t)count << kLeafCountShift) | (uint32_t)begin;
        max_depth = std::max(max_depth, depth + 1);
        uint32_t head = 0;
        size_t prev = 0;
        bool first = true;
        while (count > kMaxLeafTris) {  // iterative: no_bvh on a large mesh makes lists of 10^4 records
            const size_t idx = out.size();
            out.emplace_back();
            set_box(out[idx], 0, lo, hi);
            set_box(out[idx], 1, lo, hi);
            out[idx].child[1].ref = kLeafBit | ((uint32_t)kMaxLeafTris << kLeafCountShift) | (uint32_t)begin;
            if (first) head = (uint32_t)idx;
            else out[prev].child[0].ref = (uint32_t)idx;
            first = false;
            prev = idx;
            begin += kMaxLeafTris;
            count -= kMaxLeafTris;
        }
        out[prev].child[0].ref = kLeafBit | ((uint32_t)count << kLeafCountShift) | (uint32_t)begin;
        return head;
    }
    uint32_t node_ref(int i, int depth)
    {
        const RefNode &n = nodes[i];
        if (n.child2 <= n.child1) return range_ref(n.child2, n.child1, n.lo, n.hi, depth);  // leaf: [child2, child1)
        max_depth = std::max(max_depth, depth + 1);
        size_t idx = out.size();
        out.emplace_back();
        uint32_t a = node_ref(n.child1, depth + 1);
        uint32_t b = node_ref(n.child2, depth + 1);
        PackedRecord &r = out[idx];
        set_box(r, 0, nodes[n.child1].lo, nodes[n.child1].hi);
        set_box(r, 1, nodes[n.child2].lo, nodes[n.child2].hi);
        r.child[0].ref = a;
        r.child[1].ref = b;
        if (a == kEmptyRef) set_empty(r, 0);
        if (b == kEmptyRef) set_empty(r, 1);
        return (uint32_t)idx;
    }
    void run(int node_count)
    {
        out.clear();
        out.reserve((size_t)node_count / 2 + 2);
        bool root_internal = node_count > 0 && nodes[0].child2 > nodes[0].child1;
        if (root_internal) {
            node_ref(0, 0);
        } else {  // root is a leaf (<= 4 triangles, no_bvh, or no triangles at all): synthesise record 0
            out.emplace_back();
            max_depth = 1;
            uint32_t a = node_count > 0 ? range_ref(nodes[0].child2, nodes[0].child1, nodes[0].lo, nodes[0].hi, 1) : kEmptyRef;
            if (a == kEmptyRef) set_empty(out[0], 0);
            else { set_box(out[0], 0, nodes[0].lo, nodes[0].hi); out[0].child[0].ref = a; }
            set_empty(out[0], 1);
        }
    }
};

// True when pack_built_nodes_kernel can pack this host tree: root internal; every internal node's children are the adjacent
// pair (child1 odd, child2 = child1 + 1) allocated AFTER it; every leaf holds 1..15 triangles. *depth = levels of records.
bool host_tree_packs_on_device(const RefNode *nodes, int count, int *depth)
{
    if (count < 3 || !(nodes[0].child2 > nodes[0].child1)) return false;
    std::vector<uint8_t> level((size_t)count, 0);
    int deepest = 0;
    for (int i = 0; i < count; i++) {
        const RefNode &n = nodes[i];
        if (n.child2 > n.child1) {
            if (n.child1 <= i || (n.child1 & 1) == 0 || n.child2 != n.child1 + 1 || n.child2 >= count || level[i] >= 250) return false;
            level[n.child1] = level[n.child2] = (uint8_t)(level[i] + 1);
            deepest = std::max(deepest, level[i] + 1);
        } else {
            const int tris = n.child1 - n.child2;
            if (tris < 1 || tris > kMaxLeafTris) return false;
        }
    }
    *depth = deepest;
    return true;
}

// Grow-only device buffer: re-uploading a scene of the same size allocates nothing.
int reserve(rt_ctx *ctx, void **dst, size_t *cap, size_t bytes)
{
    bytes = std::max<size_t>(bytes, 16);
    if (*dst && *cap >= bytes) return RT_OK;
    release(*dst);
    *cap = 0;
    RT_CUDA(ctx, cudaMalloc(dst, bytes));
    *cap = bytes;
    return RT_OK;
}

int upload(rt_ctx *ctx, void **dst, size_t *cap, const void *src, size_t bytes)
{
    int rc = reserve(ctx, dst, cap, bytes);
    if (rc != RT_OK) return rc;
    if (bytes) RT_CUDA(ctx, cudaMemcpyAsync(*dst, src, bytes, cudaMemcpyHostToDevice, ctx->stream));
    return RT_OK;
}

void copy3(float *dst, const float *src) { dst[0] = src[0]; dst[1] = src[1]; dst[2] = src[2]; }

int set_image_block(rt_ctx *ctx, const rt_scene_desc *s)
{
    if (s->width <= 0 || s->height <= 0 || s->ray_count < 0 || s->bounces < 0 || s->bounces > kMaxBounces)
        return fail(ctx, RT_ERR_INVALID, "image block: need width,height > 0, ray_count >= 0, 0 <= bounces <= 256");
    DeviceScene &d = ctx->ds;
    ctx->entry_dirty = true;
    d.width = s->width;
    d.div_width = make_fast_div((uint32_t)std::max(1, s->width));
    d.height = s->height;
    copy3(d.cam, s->camera_position);
    copy3(d.top_left, s->near_plane_top_left);
    copy3(d.right, s->scaled_right);
    copy3(d.up, s->scaled_up);
    d.inv_width = s->inv_width;
    d.inv_height = s->inv_height;
    ctx->ray_count = s->ray_count;
    ctx->bounces = s->bounces;
    ctx->exposure = s->exposure;
    return RT_OK;
}

}  // namespace

// ---- context ----------------------------------------------------------------------------------------------------
extern "C" int rt_api_version(void) { return RT_API_VERSION; }

extern "C" int rt_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

extern "C" const char *rt_last_error(const rt_ctx *ctx) { return ctx ? ctx->error.c_str() : rt::global_error(); }

static int init_context(rt_ctx *ctx, int device);
extern "C" void rt_destroy(rt_ctx *ctx);

extern "C" int rt_create(rt_ctx **out, int device)
{
    if (!out) return fail(nullptr, RT_ERR_INVALID, "rt_create: null out pointer");
    *out = nullptr;
    int n = rt_device_count();
    if (n <= 0) return fail(nullptr, RT_ERR_NO_DEVICE, "rt_create: no CUDA device available (this library has no CPU fallback)");
    if (device < 0 || device >= n) return fail(nullptr, RT_ERR_NO_DEVICE, "rt_create: device index out of range");
    rt_ctx *ctx = new rt_ctx();
    ctx->device = device;
    const int rc = init_context(ctx, device);
    if (rc != RT_OK) {  // the message is already in the thread-local slot (rt_last_error(NULL))
        rt_destroy(ctx);
        return rc;
    }
    *out = ctx;
    return RT_OK;
}

static int init_context(rt_ctx *ctx, int device)
{
    RT_CUDA(ctx, cudaSetDevice(device));
    cudaDeviceProp prop;
    RT_CUDA(ctx, cudaGetDeviceProperties(&prop, device));
    ctx->sm_count = prop.multiProcessorCount;
    RT_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
    RT_CUDA(ctx, cudaEventCreateWithFlags(&ctx->ev_fork, cudaEventDisableTiming));
    for (int i = 0; i < rt_ctx::kMaxPipes - 1; i++) {
        RT_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->side[i], cudaStreamNonBlocking));
        RT_CUDA(ctx, cudaEventCreateWithFlags(&ctx->ev_join[i], cudaEventDisableTiming));
    }
    if (const char *e = std::getenv("RT_PIPES")) ctx->pipes = std::min((int)rt_ctx::kMaxPipes, std::max(1, std::atoi(e)));
    if (const char *e = std::getenv("RT_TRACE_PER_SM")) ctx->trace_per_sm = std::max(0, std::atoi(e));
    if (const char *e = std::getenv("RT_SHADE_PER_SM")) ctx->shade_per_sm = std::max(0, std::atoi(e));
    if (const char *e = std::getenv("RT_NODES")) ctx->node_mode = std::strcmp(e, "q16") == 0 ? kNodesQ16 : std::strcmp(e, "w4") == 0 ? kNodesW4 : kNodesF32;
    if (const char *e = std::getenv("RT_SORT_RANK")) ctx->sort_rank = std::min(2, std::max(0, std::atoi(e)));
    if (const char *e = std::getenv("RT_SORT_ITEMS")) { const int v = std::atoi(e); ctx->sort_items = v <= 12 ? 12 : v <= 16 ? 16 : v <= 18 ? 18 : v <= 22 ? 22 : 26; }
    if (const char *e = std::getenv("RT_TAIL")) ctx->tail_below = (uint32_t)std::max(0L, std::atol(e));
    if (const char *e = std::getenv("RT_ENTRY_K")) ctx->entry_max = std::min((int)kEntryMax, std::max(0, std::atoi(e)));
    if (const char *e = std::getenv("RT_BATCH")) ctx->trace_batch = (uint32_t)std::min(4096, std::max(4, std::atoi(e)));
    if (const char *e = std::getenv("RT_NODEMIN")) ctx->trace_node_min = (uint32_t)std::min(33, std::max(0, std::atoi(e)));
    if (const char *e = std::getenv("RT_REFILL")) ctx->trace_refill_below = (uint32_t)std::min(33, std::max(1, std::atoi(e)));
    RT_CUDA(ctx, cudaEventCreate(&ctx->ev_start));
    RT_CUDA(ctx, cudaEventCreate(&ctx->ev_stop));
    RT_CUDA(ctx, cudaMalloc(&ctx->d_pc, rt_ctx::kMaxPipes * sizeof(PassCounters)));  // one per pipe
    RT_CUDA(ctx, cudaMalloc(&ctx->d_totals, sizeof(RenderTotals)));
    return RT_OK;
}

// Page-lock caller-owned host arrays so that uploads / downloads are true asynchronous DMA at PCIe speed (a pageable copy is
// staged through the runtime's bounce buffer at ~7 GB/s and blocks the host). The host scene layer pins its own arrays.
extern "C" int rt_host_pin(const void *ptr, size_t bytes)
{
    if (!ptr || bytes == 0) return RT_OK;
    cudaError_t e = cudaHostRegister(const_cast<void *>(ptr), bytes, cudaHostRegisterPortable | cudaHostRegisterReadOnly);
    if (e == cudaErrorNotSupported || e == cudaErrorInvalidValue) {  // read-only registration is optional
        cudaGetLastError();
        e = cudaHostRegister(const_cast<void *>(ptr), bytes, cudaHostRegisterPortable);
    }
    if (e == cudaSuccess || e == cudaErrorHostMemoryAlreadyRegistered) { cudaGetLastError(); return RT_OK; }
    cudaGetLastError();
    return (e == cudaErrorNoDevice || e == cudaErrorInsufficientDriver) ? RT_ERR_NO_DEVICE : RT_ERR_CUDA;
}
extern "C" void rt_host_unpin(const void *ptr)
{
    if (!ptr) return;
    cudaHostUnregister(const_cast<void *>(ptr));
    cudaGetLastError();
}

extern "C" int rt_set_node_mode(rt_ctx *ctx, int mode)
{
    if (!ctx || mode < kNodesF32 || mode > kNodesW4) return fail(ctx, RT_ERR_INVALID, "rt_set_node_mode: mode must be 0 (f32), 1 (q16) or 2 (w4)");
    ctx->node_mode = mode;
    return RT_OK;
}
extern "C" int rt_get_node_mode(const rt_ctx *ctx) { return ctx ? (ctx->has_scene ? ctx->active_nodes : ctx->node_mode) : RT_ERR_INVALID; }

extern "C" void rt_destroy(rt_ctx *ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    for (cudaStream_t s : ctx->side) if (s) cudaStreamSynchronize(s);
    release(ctx->d_nodes); release(ctx->d_tris); release(ctx->d_materials); release(ctx->d_spheres); release(ctx->d_env);
    release(ctx->d_matidx); release(ctx->d_qnodes); release(ctx->d_raw); release(ctx->d_build);
    release(ctx->d_fb); release(ctx->d_fb_tmp); release(ctx->d_fb_tmp2); release(ctx->d_rgb8);
    for (int i = 0; i < 2; i++) {
        release(ctx->queue[i].qa); release(ctx->queue[i].qb); release(ctx->queue[i].qc);
        release(ctx->d_keys[i]); release(ctx->d_vals[i]);
    }
    release(ctx->d_hits);
    release(ctx->d_sort_scratch); release(ctx->d_pc); release(ctx->d_totals); release(ctx->d_entry);
    for (cudaEvent_t e : ctx->ev_pool) cudaEventDestroy(e);
    if (ctx->ev_start) cudaEventDestroy(ctx->ev_start);
    if (ctx->ev_stop) cudaEventDestroy(ctx->ev_stop);
    if (ctx->ev_fork) cudaEventDestroy(ctx->ev_fork);
    for (cudaEvent_t e : ctx->ev_join) if (e) cudaEventDestroy(e);
    for (cudaStream_t s : ctx->side) if (s) cudaStreamDestroy(s);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

namespace {

// Device BVH build (bvh_build.cuh): level-synchronous, one CTA per node; the host only reads the size of the next
// level. Leaves the build-form triangles in ctx->d_raw, the BVH-order permutation in ctx->d_order and the
// reference-layout nodes in ctx->built_nodes.
int build_bvh_on_device(rt_ctx *ctx, const rt_scene_desc *s, int max_depth)
{
    const int n = s->triangle_count;
    const size_t max_nodes = (size_t)2 * n + 2;
    const size_t bytes = max_nodes * sizeof(BuildNode) + (size_t)3 * n * 4 + 4 * max_nodes * 4 + 64;
    int rc = reserve(ctx, &ctx->d_build, &ctx->cap_build, bytes);
    if (rc != RT_OK) return rc;
    if ((rc = reserve(ctx, &ctx->d_raw, &ctx->cap_raw, std::max<size_t>((size_t)n * 48, (size_t)s->env_width * s->env_height * 12))) != RT_OK) return rc;
    uint8_t *base = static_cast<uint8_t *>(ctx->d_build);
    BuildNode *nodes = reinterpret_cast<BuildNode *>(base);
    int32_t *order_a = reinterpret_cast<int32_t *>(base + max_nodes * sizeof(BuildNode));
    int32_t *order_b = order_a + n, *order_final = order_b + n;
    int32_t *small_list[2] = {order_final + n, order_final + n + max_nodes};  // node ids per level, ping-pong, by size class
    int32_t *big_list[2] = {small_list[1] + max_nodes, small_list[1] + 2 * max_nodes};
    int32_t *counters = big_list[1] + max_nodes;
    const char *big_env = std::getenv("RT_BUILD_BIG");
    const int big_node_min = big_env ? std::max(1, std::atoi(big_env)) : kBuildBigNode;
    ctx->d_order = order_final;

    RT_CUDA(ctx, cudaEventRecord(ctx->ev_start, ctx->stream));
    RT_CUDA(ctx, cudaMemcpyAsync(ctx->d_raw, s->triangles, (size_t)n * 48, cudaMemcpyHostToDevice, ctx->stream));
    BuildNode root{};
    for (int k = 0; k < 3; k++) { root.lo[k] = 1e30f; root.hi[k] = -1e30f; }
    root.child2 = 0;
    root.child1 = n;
    const int32_t init[3] = {1, 0, 0};  // nodes allocated, next level's small / big nodes
    const bool root_big = n >= big_node_min;
    RT_CUDA(ctx, cudaMemcpyAsync(nodes, &root, sizeof(root), cudaMemcpyHostToDevice, ctx->stream));
    RT_CUDA(ctx, cudaMemcpyAsync(counters, init, sizeof(init), cudaMemcpyHostToDevice, ctx->stream));
    RT_CUDA(ctx, cudaMemsetAsync(root_big ? big_list[0] : small_list[0], 0, sizeof(int32_t), ctx->stream));  // the root's id
    if (n > 0) iota_kernel<<<(n + 255) / 256, 256, 0, ctx->stream>>>(order_a, n);

    int small_count = root_big ? 0 : 1, big_count = root_big ? 1 : 0, node_count = 1;
    const bool trace_levels = std::getenv("RT_BUILD_TRACE") != nullptr;  // per-level times on stderr (tuning aid)
    std::vector<cudaEvent_t> marks;
    std::vector<int> sizes;
    auto mark = [&]() {
        if (!trace_levels) return;
        cudaEvent_t e;
        cudaEventCreate(&e);
        cudaEventRecord(e, ctx->stream);
        marks.push_back(e);
    };
    mark();
    for (int level = 0; small_count + big_count > 0 && level <= max_depth; level++) {
        sizes.push_back(small_count + big_count);
        BuildArgs a{};
        a.tris = static_cast<const BuildTri *>(ctx->d_raw);
        a.nodes = nodes;
        a.order_in = (level & 1) ? order_b : order_a;
        a.order_out = (level & 1) ? order_a : order_b;
        a.order_final = order_final;
        a.level_small = small_list[level & 1];
        a.level_big = big_list[level & 1];
        a.next_small = small_list[(level & 1) ^ 1];
        a.next_big = big_list[(level & 1) ^ 1];
        a.counters = counters;
        a.level = level;
        a.max_depth = max_depth;
        a.big_node_min = big_node_min;
        if (big_count > 0) {  // one thread-block cluster per node
            cudaLaunchConfig_t cfg{};
            cfg.gridDim = dim3((unsigned)big_count * kBuildCluster);
            cfg.blockDim = dim3(kBuildBigBlock);
            cfg.stream = ctx->stream;
            cudaLaunchAttribute attr{};
            attr.id = cudaLaunchAttributeClusterDimension;
            attr.val.clusterDim.x = kBuildCluster;
            attr.val.clusterDim.y = 1;
            attr.val.clusterDim.z = 1;
            cfg.attrs = &attr;
            cfg.numAttrs = 1;
            RT_CUDA(ctx, cudaLaunchKernelEx(&cfg, bvh_level_kernel<kBuildBigBlock, kBuildCluster>, a));
        }
        if (small_count > 0) bvh_level_kernel<kBuildBlock, 1><<<small_count, kBuildBlock, 0, ctx->stream>>>(a);
        int32_t c[3];
        RT_CUDA(ctx, cudaMemcpyAsync(c, counters, sizeof(c), cudaMemcpyDeviceToHost, ctx->stream));
        RT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        node_count = c[0];
        small_count = c[1];
        big_count = c[2];
        RT_CUDA(ctx, cudaMemsetAsync(counters + 1, 0, 2 * sizeof(int32_t), ctx->stream));
        mark();
    }
    RT_CUDA(ctx, cudaGetLastError());
    ctx->built_node_count = node_count;
    ctx->built_levels = (int)sizes.size();
    ctx->device_packed = false;
    if (node_count >= 3) {  // pack on the device; the nodes stay there unless rt_scene_download_bvh asks for them
        const int records = (node_count - 1) / 2;
        if ((rc = reserve(ctx, &ctx->d_nodes, &ctx->cap_nodes, (size_t)records * 64)) != RT_OK) return rc;
        pack_built_nodes_kernel<<<(node_count + 255) / 256, 256, 0, ctx->stream>>>(nodes, node_count, (float4 *)ctx->d_nodes, kLeafBit,
                                                                                   kLeafCountShift, kMaxLeafTris, counters + 1);
        int32_t needs_host = 0;
        RT_CUDA(ctx, cudaMemcpyAsync(&needs_host, counters + 1, sizeof(int32_t), cudaMemcpyDeviceToHost, ctx->stream));
        RT_CUDA(ctx, cudaMemcpyAsync(ctx->built_root, nodes, sizeof(BuildNode), cudaMemcpyDeviceToHost, ctx->stream));
        mark();
        RT_CUDA(ctx, cudaEventRecord(ctx->ev_stop, ctx->stream));
        RT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        ctx->device_packed = needs_host == 0;
        ctx->packed_nodes = records;
        ctx->packed_depth = ctx->built_levels;
    }
    if (!ctx->device_packed) {  // tiny tree or a leaf that needs a chain: the host packer works from the nodes
        ctx->built_nodes.resize((size_t)node_count * sizeof(BuildNode));
        RT_CUDA(ctx, cudaMemcpyAsync(ctx->built_nodes.data(), nodes, ctx->built_nodes.size(), cudaMemcpyDeviceToHost, ctx->stream));
        mark();
        RT_CUDA(ctx, cudaEventRecord(ctx->ev_stop, ctx->stream));
        RT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    }
    if (trace_levels) {
        float t0 = 0;
        cudaEventElapsedTime(&t0, ctx->ev_start, marks[0]);
        std::fprintf(stderr, "bvh build: %d triangles, copy-in %.3f ms\n", n, t0);
        for (size_t i = 0; i + 1 < marks.size(); i++) {
            float ms = 0;
            cudaEventElapsedTime(&ms, marks[i], marks[i + 1]);
            if (i < sizes.size()) std::fprintf(stderr, "  level %2zu: %7d nodes %8.3f ms\n", i, sizes[i], ms);
            else std::fprintf(stderr, "  %s: %.3f ms\n", ctx->device_packed ? "pack records" : "nodes to host", ms);
        }
        for (cudaEvent_t e : marks) cudaEventDestroy(e);
    }
    float ms = 0;
    RT_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->ev_start, ctx->ev_stop));
    ctx->bvh_device_ms = ms;
    return RT_OK;
}

}  // namespace

extern "C" int rt_scene_set_image(rt_ctx *ctx, const rt_scene_desc *s)
{
    if (!ctx || !s) return fail(ctx, RT_ERR_INVALID, "rt_scene_set_image: null argument");
    if (!ctx->has_scene) return fail(ctx, RT_ERR_INVALID, "rt_scene_set_image: no scene uploaded");
    return set_image_block(ctx, s);
}

extern "C" int rt_scene_upload(rt_ctx *ctx, const rt_scene_desc *s)
{
    if (!ctx || !s) return fail(ctx, RT_ERR_INVALID, "rt_scene_upload: null argument");
    const bool device_build = s->triangles_form == 1;
    if (s->sphere_count < 0 || s->triangle_count < 0 || s->material_count <= 0 || (!device_build && (s->bvh_node_count <= 0 || !s->bvh_nodes)) ||
        s->triangles_form < 0 || s->triangles_form > 1 || !s->materials || !s->environment_map || s->env_width <= 0 || s->env_height <= 0 ||
        (s->triangle_count > 0 && !s->triangles) || (s->sphere_count > 0 && !s->spheres) ||
        (s->triangle_count + s->sphere_count > 0 && !s->material_indices))
        return fail(ctx, RT_ERR_INVALID, "rt_scene_upload: malformed scene descriptor");
    if ((uint32_t)s->triangle_count > kLeafStartMask) return fail(ctx, RT_ERR_INVALID, "rt_scene_upload: more than 2^27 triangles");
    // the lookup keeps the reference's row stride = map HEIGHT (scene.cu:393): texel index <= (h-1)*h + w-1, inside the
    // map only when width >= height (the reference's own maps are square; it would read out of bounds otherwise)
    if (s->env_width < s->env_height) return fail(ctx, RT_ERR_INVALID, "rt_scene_upload: environment map must be at least as wide as high (the lookup strides rows by the height, as the reference does)");
    RT_CUDA(ctx, cudaSetDevice(ctx->device));
    ctx->has_scene = false;
    int rc = set_image_block(ctx, s);
    if (rc != RT_OK) return rc;

    ctx->built_nodes.clear();
    ctx->built_node_count = 0;
    ctx->device_packed = false;
    ctx->d_order = nullptr;
    ctx->bvh_device_ms = 0;
    if (device_build && (rc = build_bvh_on_device(ctx, s, 30)) != RT_OK) return rc;  // MAX_BVH_DEPTH, scene.cu:10
    Packer packer;
    packer.nodes = reinterpret_cast<const RefNode *>(ctx->built_root);
    if (!device_build) {
        // A host tree in the reference's own shape (children allocated in adjacent pairs after their parent, leaves of
        // 1..15 triangles: what generate_bvh produces) is packed into records ON THE DEVICE from its raw 32-byte
        // nodes: one forward pass on the host validates the shape and finds the depth, nothing else touches the 6 MB.
        // Anything else (no_bvh's root leaf, oversized leaves, a tree in another order) goes through the host packer.
        const RefNode *nodes = static_cast<const RefNode *>(s->bvh_nodes);
        const int count = s->bvh_node_count;
        int depth = 0;
        if (host_tree_packs_on_device(nodes, count, &depth)) {
            if (depth > kStackDepth) return fail(ctx, RT_ERR_INVALID, "rt_scene_upload: BVH deeper than the traversal stack (32)");
            const int records = (count - 1) / 2;
            if ((rc = reserve(ctx, &ctx->d_build, &ctx->cap_build, (size_t)count * sizeof(RefNode) + 16)) != RT_OK) return rc;
            if ((rc = reserve(ctx, &ctx->d_nodes, &ctx->cap_nodes, (size_t)records * 64)) != RT_OK) return rc;
            int32_t *flag = reinterpret_cast<int32_t *>(static_cast<uint8_t *>(ctx->d_build) + (size_t)count * sizeof(RefNode));
            RT_CUDA(ctx, cudaMemcpyAsync(ctx->d_build, nodes, (size_t)count * sizeof(RefNode), cudaMemcpyHostToDevice, ctx->stream));
            pack_built_nodes_kernel<<<(count + 255) / 256, 256, 0, ctx->stream>>>(static_cast<const BuildNode *>(ctx->d_build), count, (float4 *)ctx->d_nodes,
                                                                                  kLeafBit, kLeafCountShift, kMaxLeafTris, flag);
            RT_CUDA(ctx, cudaGetLastError());
            ctx->device_packed = true;
            ctx->packed_nodes = records;
            ctx->packed_depth = depth;
            packer.nodes = nodes;
        }
    }
    if (!ctx->device_packed) {
        packer.nodes = device_build ? reinterpret_cast<const RefNode *>(ctx->built_nodes.data()) : static_cast<const RefNode *>(s->bvh_nodes);
        packer.run(device_build ? (int)(ctx->built_nodes.size() / sizeof(RefNode)) : s->bvh_node_count);
        if (packer.max_depth > kStackDepth) return fail(ctx, RT_ERR_INVALID, "rt_scene_upload: BVH deeper than the traversal stack (32)");
        ctx->packed_nodes = (int)packer.out.size();
        ctx->packed_depth = packer.max_depth;
    }

    const int prims = s->sphere_count + s->triangle_count;
    const size_t texels = (size_t)s->env_width * s->env_height;
    const size_t node_bytes = packer.out.size() * sizeof(PackedRecord);
    const size_t raw_tri = (size_t)s->triangle_count * 48, raw_env = texels * 12;
    if (!ctx->device_packed) {  // packer.out is pageable: the runtime stages the copy before cudaMemcpyAsync returns
        if ((rc = upload(ctx, &ctx->d_nodes, &ctx->cap_nodes, packer.out.data(), node_bytes)) != RT_OK) return rc;
    }
    if ((rc = upload(ctx, &ctx->d_materials, &ctx->cap_materials, s->materials, (size_t)s->material_count * 48)) != RT_OK) return rc;
    if ((rc = upload(ctx, &ctx->d_spheres, &ctx->cap_spheres, s->spheres, (size_t)s->sphere_count * 16)) != RT_OK) return rc;
    if ((rc = upload(ctx, &ctx->d_matidx, &ctx->cap_matidx, s->material_indices, (size_t)prims * sizeof(uint16_t))) != RT_OK) return rc;
    // triangles and texels: raw copy, then repack on the device (pack_triangles_kernel / pack_env_kernel)
    if ((rc = reserve(ctx, &ctx->d_raw, &ctx->cap_raw, std::max(raw_tri, raw_env))) != RT_OK) return rc;
    if ((rc = reserve(ctx, &ctx->d_tris, &ctx->cap_tris, (size_t)s->triangle_count * 64)) != RT_OK) return rc;
    if ((rc = reserve(ctx, &ctx->d_env, &ctx->cap_env, texels * 16)) != RT_OK) return rc;
    if (s->triangle_count > 0 && device_build) {  // build-form triangles are already in d_raw; order = the build's permutation
        pack_built_triangles_kernel<<<(s->triangle_count + 255) / 256, 256, 0, ctx->stream>>>(
            (const BuildTri *)ctx->d_raw, ctx->d_order, (const uint16_t *)ctx->d_matidx + s->sphere_count, (float4 *)ctx->d_tris,
            (uint32_t)s->triangle_count);
    } else if (s->triangle_count > 0) {
        RT_CUDA(ctx, cudaMemcpyAsync(ctx->d_raw, s->triangles, raw_tri, cudaMemcpyHostToDevice, ctx->stream));
        pack_triangles_kernel<<<(s->triangle_count + 255) / 256, 256, 0, ctx->stream>>>(
            (const float *)ctx->d_raw, (const uint16_t *)ctx->d_matidx + s->sphere_count, (float4 *)ctx->d_tris, (uint32_t)s->triangle_count);
    }
    RT_CUDA(ctx, cudaMemcpyAsync(ctx->d_raw, s->environment_map, raw_env, cudaMemcpyHostToDevice, ctx->stream));
    pack_env_kernel<<<(unsigned)((texels + 255) / 256), 256, 0, ctx->stream>>>((const float *)ctx->d_raw, (float4 *)ctx->d_env, texels);
    RT_CUDA(ctx, cudaGetLastError());
    // no synchronisation: copies from the caller's pageable arrays are staged by the runtime before the call returns,
    // and everything else is ordered on the context's stream ahead of the first render kernel

    // compressed node formats: derived from the 64-byte records that are now in d_nodes
    ctx->active_nodes = kNodesF32;
    QuantGrid grid{};
    if (ctx->node_mode != kNodesF32) {
        const RefNode &root_box = packer.nodes[0];
        grid = make_quant_grid(root_box.lo, root_box.hi);
        if (grid.ok || s->triangle_count == 0) ctx->active_nodes = ctx->node_mode;
    }
    if (ctx->active_nodes == kNodesQ16) {
        if ((rc = reserve(ctx, &ctx->d_qnodes, &ctx->cap_qnodes, (size_t)ctx->packed_nodes * 32)) != RT_OK) return rc;
        quantize_records_kernel<<<(ctx->packed_nodes + 127) / 128, 128, 0, ctx->stream>>>((const float4 *)ctx->d_nodes, ctx->packed_nodes,
                                                                                        (uint4 *)ctx->d_qnodes, grid);
        RT_CUDA(ctx, cudaGetLastError());
    } else if (ctx->active_nodes == kNodesW4) {
        std::vector<float> fetched;
        WideCollapse wide;
        wide.grid = grid;
        if (ctx->device_packed) {  // the records exist only on the device
            fetched.resize((size_t)ctx->packed_nodes * 16);
            RT_CUDA(ctx, cudaMemcpyAsync(fetched.data(), ctx->d_nodes, fetched.size() * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
            RT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
            wide.records = fetched.data();
        } else {
            wide.records = reinterpret_cast<const float *>(packer.out.data());
        }
        wide.run();
        ctx->wide_nodes = (int)(wide.out.size() / 16);
        ctx->wide_depth = wide.depth;
        ctx->wide_stack_need = wide.stack_need;
        if (wide.stack_need > kWideStackSlots) return fail(ctx, RT_ERR_INVALID, "rt_scene_upload: 4-wide tree needs more stack entries than the tracer has (72)");
        RT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));  // wide.out is pageable and dies with this call
        if ((rc = upload(ctx, &ctx->d_qnodes, &ctx->cap_qnodes, wide.out.data(), wide.out.size() * sizeof(uint32_t))) != RT_OK) return rc;
        RT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    }

    DeviceScene &d = ctx->ds;
    d.nodes = static_cast<const float4 *>(ctx->d_nodes);
    d.qnodes = static_cast<const uint4 *>(ctx->d_qnodes);
    for (int k = 0; k < 3; k++) { d.q_scale[k] = grid.scale[k]; d.q_bias[k] = grid.bias[k]; }
    d.tris = static_cast<const float4 *>(ctx->d_tris);
    d.materials = static_cast<const float4 *>(ctx->d_materials);
    d.spheres = static_cast<const float4 *>(ctx->d_spheres);
    d.env = static_cast<const float4 *>(ctx->d_env);
    d.material_indices = static_cast<const uint16_t *>(ctx->d_matidx);
    d.sphere_count = s->sphere_count;
    d.tri_count = s->triangle_count;
    d.env_w = s->env_width;
    d.env_h = s->env_height;
    copy3(d.min_coord, s->min_coord);
    copy3(d.inv_dim, s->inv_dimensions);

    // bounds for the reorder key: the triangle root box, else the spheres' bounds
    const RefNode &root = packer.nodes[0];
    float lo[3], hi[3];
    if (s->triangle_count > 0) { copy3(lo, root.lo); copy3(hi, root.hi); }
    else {
        for (int k = 0; k < 3; k++) { lo[k] = 1e30f; hi[k] = -1e30f; }
        for (int i = 0; i < s->sphere_count; i++)
            for (int k = 0; k < 3; k++) {
                float c = s->spheres[4 * i + k], r = std::min(s->spheres[4 * i + 3], 1e3f);
                lo[k] = std::min(lo[k], c - r); hi[k] = std::max(hi[k], c + r);
            }
    }
    for (int k = 0; k < 3; k++) {
        float extent = hi[k] - lo[k];
        d.key_min[k] = lo[k];
        d.key_inv_extent[k] = (extent > 0.0f && extent < 1e30f) ? 1.0f / extent : 0.0f;
    }
    ctx->has_scene = true;
    return RT_OK;
}

namespace {

int ensure_framebuffer(rt_ctx *ctx)
{
    size_t px = (size_t)ctx->ds.width * ctx->ds.height;
    if (px == ctx->fb_pixels && ctx->d_fb) return RT_OK;
    release(ctx->d_fb); release(ctx->d_fb_tmp); release(ctx->d_fb_tmp2); release(ctx->d_rgb8);
    ctx->fb_pixels = 0;
    RT_CUDA(ctx, cudaMalloc(&ctx->d_fb, px * 3 * sizeof(float)));
    RT_CUDA(ctx, cudaMalloc(&ctx->d_fb_tmp, px * 3 * sizeof(float)));
    RT_CUDA(ctx, cudaMalloc(&ctx->d_fb_tmp2, px * 3 * sizeof(float)));
    RT_CUDA(ctx, cudaMalloc(&ctx->d_rgb8, px * 3));
    RT_CUDA(ctx, cudaMemsetAsync(ctx->d_fb, 0, px * 3 * sizeof(float), ctx->stream));
    ctx->fb_pixels = px;
    return RT_OK;
}

int ensure_queues(rt_ctx *ctx, size_t rays, bool sort)
{
    if (rays > ctx->queue_capacity) {
        for (int i = 0; i < 2; i++) { release(ctx->queue[i].qa); release(ctx->queue[i].qb); release(ctx->queue[i].qc); }
        release(ctx->d_hits);
        ctx->queue_capacity = 0;
        RT_CUDA(ctx, cudaMalloc(&ctx->d_hits, rays * sizeof(float2)));
        for (int i = 0; i < 2; i++) {
            RT_CUDA(ctx, cudaMalloc(&ctx->queue[i].qa, rays * sizeof(float4)));
            RT_CUDA(ctx, cudaMalloc(&ctx->queue[i].qb, rays * sizeof(float4)));
            RT_CUDA(ctx, cudaMalloc(&ctx->queue[i].qc, rays * sizeof(float2)));
        }
        ctx->queue_capacity = rays;
    }
    if (sort && rays > ctx->sort_capacity) {
        for (int i = 0; i < 2; i++) { release(ctx->d_keys[i]); release(ctx->d_vals[i]); }
        release(ctx->d_sort_scratch);
        ctx->sort_capacity = 0;
        for (int i = 0; i < 2; i++) {
            RT_CUDA(ctx, cudaMalloc(&ctx->d_keys[i], rays * sizeof(uint32_t)));
            RT_CUDA(ctx, cudaMalloc(&ctx->d_vals[i], rays * sizeof(uint32_t)));
        }
        size_t tiles = (rays + kSortTile - 1) / kSortTile;
        ctx->sort_scratch_words = (size_t)kSortMaxPasses * 256 + 16 + (size_t)kSortMaxPasses * tiles * 256;
        RT_CUDA(ctx, cudaMalloc(&ctx->d_sort_scratch, ctx->sort_scratch_words * sizeof(uint32_t)));
        ctx->sort_capacity = rays;
    }
    return RT_OK;
}

// Enqueue a stable sort of (keys[0], index) on the low `bits` bits; n is read from *count_ptr on the device, `bound`
// is the host-side upper bound used for grid sizing. On return *sorted_vals points at the permutation.
int enqueue_sort(rt_ctx *ctx, const uint32_t *count_ptr, size_t bound, int bits, bool iota, uint32_t **sorted_keys,
                 uint32_t **sorted_vals, int *launches)
{
    int passes = std::min(kSortMaxPasses, std::max(1, (bits + 7) / 8));
    const int items = ctx->sort_items, tile_keys = kSortBlock * items;
    const size_t tiles = (ctx->sort_capacity + tile_keys - 1) / tile_keys;  // stride of the per-pass status arrays (the scratch is sized for the smallest tile)
    uint32_t launch_tiles = (uint32_t)((bound + tile_keys - 1) / tile_keys);
    SortArgs s;
    s.count_ptr = count_ptr;
    s.hist = ctx->d_sort_scratch;
    s.tile_counter = ctx->d_sort_scratch + kSortMaxPasses * 256;
    s.tile_status = ctx->d_sort_scratch + kSortMaxPasses * 256 + 16;
    s.passes = passes;
    s.max_tiles = (uint32_t)tiles;
    // zero histograms, tile counters and the status words of the tiles this launch can touch: ONE memset when the status
    // arrays of the passes are adjacent (the launch covers the whole capacity, or one pass), else one per pass
    const size_t head_words = kSortMaxPasses * 256 + 16;
    if (launch_tiles == tiles || passes == 1) {
        RT_CUDA(ctx, cudaMemsetAsync(ctx->d_sort_scratch, 0, (head_words + ((size_t)(passes - 1) * tiles + launch_tiles) * 256) * sizeof(uint32_t), ctx->stream));
    } else {
        RT_CUDA(ctx, cudaMemsetAsync(ctx->d_sort_scratch, 0, head_words * sizeof(uint32_t), ctx->stream));
        for (int p = 0; p < passes; p++)
            RT_CUDA(ctx, cudaMemsetAsync(s.tile_status + (size_t)p * tiles * 256, 0, (size_t)launch_tiles * 256 * sizeof(uint32_t), ctx->stream));
    }
    int hist_blocks = (int)std::min<size_t>((size_t)ctx->sm_count * 8, std::max<size_t>(1, (bound + kSortBlock - 1) / kSortBlock));
    sort_histogram_kernel<<<hist_blocks, kSortBlock, 0, ctx->stream>>>(ctx->d_keys[0], s);  // its last CTA scans the histograms
    int src = 0;
    for (int p = 0; p < passes; p++) {
        const uint32_t *vin = (p == 0 && iota) ? nullptr : ctx->d_vals[src];
#define RT_SCATTER(RANK, ITEMS)                                                                                                   \
    do {                                                                                                                          \
        constexpr int kStage = kSortBlock * ITEMS * 8;                                                                            \
        if (kStage > 40 * 1024) cudaFuncSetAttribute(sort_scatter_kernel<RANK, ITEMS>, cudaFuncAttributeMaxDynamicSharedMemorySize, kStage); \
        sort_scatter_kernel<RANK, ITEMS><<<launch_tiles, kSortBlock, kStage, ctx->stream>>>(ctx->d_keys[src], vin, ctx->d_keys[src ^ 1], ctx->d_vals[src ^ 1], s, p); \
    } while (0)
        switch (ctx->sort_rank * 100 + items) {
        case 12: RT_SCATTER(0, 12); break;
        case 16: RT_SCATTER(0, 16); break;
        case 18: RT_SCATTER(0, 18); break;
        case 112: RT_SCATTER(1, 12); break;
        case 116: RT_SCATTER(1, 16); break;
        case 118: RT_SCATTER(1, 18); break;
        case 212: RT_SCATTER(2, 12); break;
        case 216: RT_SCATTER(2, 16); break;
        case 222: RT_SCATTER(2, 22); break;
        case 226: RT_SCATTER(2, 26); break;
        default: RT_SCATTER(2, 18); break;
        }
#undef RT_SCATTER
        src ^= 1;
    }
    RT_CUDA(ctx, cudaGetLastError());
    *sorted_keys = ctx->d_keys[src];
    *sorted_vals = ctx->d_vals[src];
    if (launches) *launches += 1 + passes;
    return RT_OK;
}

}  // namespace

// ---- the pass loop (gpu_raytrace, raytracing.cu:170-284) ------------------------------------------------------
namespace {

template <bool FIRST, bool STATS>
void launch_trace_nodes(const rt_ctx *ctx, cudaStream_t stream, int grid, const BounceArgs &a)
{
    switch (ctx->active_nodes) {
    case kNodesQ16: trace_rays_kernel<FIRST, STATS, kNodesQ16><<<grid, kBlock, 0, stream>>>(ctx->ds, a); break;
    case kNodesW4: trace_rays_kernel<FIRST, STATS, kNodesW4><<<grid, kBlock, 0, stream>>>(ctx->ds, a); break;
    default: trace_rays_kernel<FIRST, STATS, kNodesF32><<<grid, kBlock, 0, stream>>>(ctx->ds, a); break;
    }
}
template <bool FIRST>
void launch_trace(const rt_ctx *ctx, cudaStream_t stream, bool stats, int grid, const BounceArgs &a)
{
    if (stats) launch_trace_nodes<FIRST, true>(ctx, stream, grid, a);
    else launch_trace_nodes<FIRST, false>(ctx, stream, grid, a);
}

// Persistent grid of the tracer: every SM filled to the occupancy the kernel variant allows.
int trace_grid_size(rt_ctx *ctx, int *out)
{
    int per_sm = 0;
    switch (ctx->active_nodes) {
    case kNodesQ16: RT_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, trace_rays_kernel<false, false, kNodesQ16>, kBlock, 0)); break;
    case kNodesW4: RT_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, trace_rays_kernel<false, false, kNodesW4>, kBlock, 0)); break;
    default: RT_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, trace_rays_kernel<false, false, kNodesF32>, kBlock, 0)); break;
    }
    if (ctx->trace_per_sm > 0) per_sm = std::min(per_sm, ctx->trace_per_sm);
    *out = ctx->sm_count * std::max(1, per_sm);
    return RT_OK;
}

int render_impl(rt_ctx *ctx, const rt_render_opts *opts_in, rt_stats *stats)
{
    if (!ctx) return fail(nullptr, RT_ERR_INVALID, "rt_render: null context");
    if (!ctx->has_scene) return fail(ctx, RT_ERR_INVALID, "rt_render: no scene uploaded");
    auto wall0 = std::chrono::high_resolution_clock::now();
    rt_render_opts opts{};
    if (opts_in) opts = *opts_in;
    opts.sort = opts_in ? opts_in->sort : 1;
    const int pass_spp = opts.pass_spp > 0 ? opts.pass_spp : 20;  // MAX_RAYS_PER_PIXEL_PER_PASS, raytracing.cu:109
    const size_t pixels = (size_t)ctx->ds.width * ctx->ds.height;
    if (pixels * (size_t)pass_spp >= (1ull << 31)) return fail(ctx, RT_ERR_INVALID, "rt_render: width*height*pass_spp must be < 2^31");
    // The work of this call: SEGMENTS = rays [offset, offset + count) of a pass (wave_plan.cuh): whole passes by default or by pass
    // range; with sample_count / ray_count > 0 a range of every pixel's samples / of the frame's ray sequence, whose first and
    // last pass may be partial.
    using Segment = PassSegment;
    const FrameShape shape{pixels, ctx->ray_count, pass_spp};
    std::vector<Segment> segs;
    auto spp_of = [&](int p) { return shape.spp_of(p); };
    if (const char *bad = call_segments(shape, opts.pass_first, opts.pass_count, opts.ray_first, opts.ray_count, opts.sample_first, opts.sample_count, segs))
        return fail(ctx, RT_ERR_INVALID, bad);
    const int n_segs = (int)segs.size();
    const bool collect = opts.collect_stats != 0;
    const bool timing = opts.kernel_timing != 0;  // CUDA-event pair around every bounce kernel and every reorder pass
    std::vector<int> ev_kind;                       // 0 = trace kernel, 1 = reorder, 2 = shade kernel
    std::vector<int> ev_bounce;
    int cur_bounce = 0;
    cudaStream_t cur_stream = ctx->stream;  // the pipe the wave in hand runs on
    auto mark = [&](int kind) -> int {
        if (!timing) return RT_OK;
        size_t k = ev_kind.size();
        while (ctx->ev_pool.size() <= k) {
            cudaEvent_t e;
            if (cudaEventCreate(&e) != cudaSuccess) return RT_ERR_CUDA;
            ctx->ev_pool.push_back(e);
        }
        ev_kind.push_back(kind);
        ev_bounce.push_back(cur_bounce);
        return cudaEventRecord(ctx->ev_pool[k], cur_stream) == cudaSuccess ? RT_OK : RT_ERR_CUDA;
    };
    const bool sort = opts.sort != 0 && ctx->bounces > 1;
    int key_bits = opts.sort_bits > 0 ? opts.sort_bits : 15;  // 4 bits per axis + octant = two 8-bit digit passes
    key_bits = std::min(30, std::max(6, key_bits));
    key_bits = 3 + 3 * ((key_bits - 3) / 3);
    // random.cuh:28: inc = 820957824423429, used as (inc | 1); other streams pick other odd increments
    const unsigned long long pcg_inc = (820957824423429ULL + 2ULL * 0x9E3779B97F4A7C15ULL * (unsigned long long)std::max(0, opts.stream_id)) | 1ULL;

    RT_CUDA(ctx, cudaSetDevice(ctx->device));
    int rc = ensure_framebuffer(ctx);
    if (rc != RT_OK) return rc;
    const size_t pass_rays_max = pixels * (size_t)std::min(pass_spp, std::max(1, ctx->ray_count));
    int max_group = opts.pass_group;
    if (max_group <= 0) {  // auto: as many passes per wave as fit in 70 % of the free device memory (ids stay < 2^31)
        size_t free_b = 0, total_b = 0;
        RT_CUDA(ctx, cudaMemGetInfo(&free_b, &total_b));
        const size_t have = ctx->queue_capacity * (size_t)(88 + (sort ? 16 : 0));  // already ours, so also "free"
        const size_t per_pass = pass_rays_max * (size_t)(88 + (sort ? 16 : 0));    // 2 x 40 B queues + 8 B hit (+ keys, permutation)
        max_group = (int)std::max<size_t>(1, (free_b + have) / 10 * 7 / std::max<size_t>(1, per_pass));
    }
    const int mem_group = std::max(1, max_group);  // passes whose queues fit in memory (over all pipes), before any other limit
    max_group = std::max(1, std::min({max_group, (int)kMaxPassGroup, std::max(1, n_segs)}));
    // ray ids are 31-bit; the reorder pass keeps tile counts in 30 bits next to its two status flags (radix_sort.cuh)
    const size_t wave_limit = sort ? (1ull << 30) : (1ull << 31);
    if (pass_rays_max >= wave_limit) return fail(ctx, RT_ERR_INVALID, "rt_render: with sort = 1 width*height*pass_spp must be < 2^30");
    while (max_group > 1 && pass_rays_max * (size_t)max_group >= wave_limit) max_group--;
    // PIPES (2 by default, RT_PIPES=1..4): consecutive waves rotate over that many streams, each with its own share of the queues and its own
    // counters. A wave is a chain of dependent launches whose late bounces hold a few 10^4 rays and whose persistent
    // tracer drains warp by warp at the end of every launch; with a second, independent wave in flight those SMs pick up
    // its CTAs instead of idling, and the HBM-latency-bound shade kernel of one wave overlaps the issue-bound tracer of the
    // other. The automatic wave size is divided by the pipes so that the memory in use stays the same. Not with the reorder pass (its
    // scratch is single), not with per-kernel event timing or counting (their numbers are per kernel, alone on the GPU).
    // Automatic: 2; 3 for a call of at most 8 passes (one GPU's share of a small frame). Such a call is two or three waves in all, the
    // fixed chain of late bounces is a tenth of it, and three smaller waves leave less of that chain without company than two
    // (teapot 1000x1000, samples [0, 125): 17.67 ms on 2 pipes, 17.11 on 3, 17.25 on 4; profiles/r2f_partial.md). Whole frames
    // measured no gain from a third pipe (profiles/r1k_box_ch.md).
    // Only where the whole call then fits ONE round of three waves in memory: lamp 4K's share of 4 GPUs is 8 passes of 14.6 GB with
    // room for 8 at a time, which is two pipes of 4 + 4, not three pipes of 2 + 2 + 1 + 1 + 1 + 1.
    const bool small_call = n_segs <= 8 && mem_group / 3 >= (n_segs + 2) / 3;
    const int want_pipes = ctx->pipes > 0 ? ctx->pipes : (small_call ? 3 : 2);
    int pipes = (sort || timing || collect || opts.single_pipe || n_segs < 2) ? 1 : want_pipes;
    pipes = std::max(1, std::min(pipes, n_segs));
    // automatic wave size with several pipes: the work split evenly over the pipes (rounded UP: 5 passes on 2 pipes are waves of
    // 3 + 2, not 2 + 2 + 1 — every wave pays a fixed chain of latency-bound late bounces), within each pipe's share of the memory
    if (pipes > 1 && opts.pass_group <= 0) max_group = std::max(1, std::min({max_group, mem_group / pipes, (n_segs + pipes - 1) / pipes}));
    // the automatic wave size is a guess at what the allocator will grant: on an out-of-memory answer try smaller waves
    for (;;) {
        rc = ensure_queues(ctx, pass_rays_max * (size_t)max_group * (size_t)pipes, sort);
        if (rc != RT_ERR_NOMEM) break;
        cudaGetLastError();
        if (opts.pass_group <= 0 && max_group > 1) max_group = (max_group + 1) / 2;
        else if (pipes > 1) pipes--;  // an explicit wave size that fits once but not `pipes` times
        else break;
    }
    if (rc != RT_OK) return rc;
    const size_t pipe_stride = pass_rays_max * (size_t)max_group;  // queue positions per pipe

    int grid = 0;
    if ((rc = trace_grid_size(ctx, &grid)) != RT_OK) return rc;
    const int shade_grid_max = ctx->sm_count * (ctx->shade_per_sm > 0 ? ctx->shade_per_sm : 2048 / kShadeBlock);

    // entry records of the primary rays: cut once per scene + camera, reused by every pass and frame
    const int tiles_x = (ctx->ds.width + kEntryTile - 1) / kEntryTile, tiles_y = (ctx->ds.height + kEntryTile - 1) / kEntryTile;
    const bool use_entries = ctx->entry_max > 0 && opts.root_start == 0;
    // tail_kernel (kernels.cuh) walks the float records from the queue in its own order: not with the reorder pass (permuted
    // queue), the counting variant (it does not count visits) or the compressed node formats
    const bool use_tail = ctx->tail_below > 0 && !sort && !collect && ctx->active_nodes == kNodesF32;
    if (use_entries && ctx->entry_dirty) {
        if ((rc = reserve(ctx, &ctx->d_entry, &ctx->cap_entry, (size_t)tiles_x * tiles_y * 64)) != RT_OK) return rc;
        build_entry_lists_kernel<<<(tiles_x * tiles_y + 127) / 128, 128, 0, ctx->stream>>>(ctx->ds, tiles_x, tiles_y, ctx->entry_max, (uint4 *)ctx->d_entry);
        ctx->entry_dirty = false;
    }

    RT_CUDA(ctx, cudaMemsetAsync(ctx->d_fb, 0, pixels * 3 * sizeof(float), ctx->stream));
    RT_CUDA(ctx, cudaMemsetAsync(ctx->d_totals, 0, sizeof(RenderTotals), ctx->stream));
    RT_CUDA(ctx, cudaEventRecord(ctx->ev_start, ctx->stream));
    if (pipes > 1) {
        RT_CUDA(ctx, cudaEventRecord(ctx->ev_fork, ctx->stream));
        for (int i = 0; i < pipes - 1; i++) RT_CUDA(ctx, cudaStreamWaitEvent(ctx->side[i], ctx->ev_fork, 0));
    }

    // an error return from inside the loop must not leave launches in flight on the side streams (they write the queues and
    // counters the next call reuses): drain every pipe on the way out
    struct PipeGuard {
        rt_ctx *c;
        bool armed;
        ~PipeGuard()
        {
            if (!armed) return;
            for (cudaStream_t side : c->side) if (side) cudaStreamSynchronize(side);
            cudaStreamSynchronize(c->stream);
        }
    } pipe_guard{ctx, true};
    int launches = 0, passes_done = 0, wave = 0;
    long long primaries = 0;
    // A WAVE is `group` consecutive passes of equal size traced together: every pass keeps the reference's seeds
    // (they depend on the pass, not on the launch), while the late bounces - a few 10^4 rays per pass, latency-bound -
    // and the launch gaps are paid once per wave instead of once per pass.
    // Which passes go into which wave: plan_waves (wave_plan.cuh; partial passes ride as the last member of a wave of whole passes,
    // large runs are dealt out in equal waves, one-round calls first fit).
    const std::vector<WavePlan> plan = plan_waves(segs, shape, max_group, pipes);
    for (const WavePlan &wp : plan) {
        const Segment *member = wp.member;
        const int group = wp.group, partial_member = wp.partial_member;
        const int rays_to_cast = spp_of(member[0].pass);
        const bool partial = partial_member >= 0;
        const Segment &pseg = member[partial ? partial_member : 0];
        const uint32_t pass_rays = partial_member == 0 ? pseg.count : (uint32_t)(pixels * (size_t)rays_to_cast);
        int remaining[kMaxPassGroup];  // the reference's `remaining_rays` after the decrement (:224-225), per pass
        for (int g = 0; g < group; g++) remaining[g] = ctx->ray_count - member[g].pass * pass_spp - rays_to_cast;
        const uint32_t total = partial_member > 0 ? pass_rays * (uint32_t)(group - 1) + pseg.count : pass_rays * (uint32_t)group;
        primaries += total;
        const int pipe = wave++ % pipes;
        cur_stream = pipe ? ctx->side[pipe - 1] : ctx->stream;
        PassCounters *pc = ctx->d_pc + pipe;
        const size_t qoff = (size_t)pipe * pipe_stride;
        auto pipe_queue = [&](int i) { return RayQueue{ctx->queue[i].qa + qoff, ctx->queue[i].qb + qoff, ctx->queue[i].qc + qoff}; };
        RT_CUDA(ctx, cudaMemsetAsync(pc, 0, sizeof(PassCounters), cur_stream));
        const uint32_t *order = nullptr;
        for (int b = 0; b < ctx->bounces; b++) {
            BounceArgs a{};
            cur_bounce = b;
            a.in = pipe_queue((b + 1) & 1);
            a.out = (b + 1 < ctx->bounces) ? pipe_queue(b & 1) : RayQueue{nullptr, nullptr, nullptr};
            a.order = order;
            a.keys_out = (sort && b + 1 < ctx->bounces) ? ctx->d_keys[0] : nullptr;
            a.pc = pc;
            a.totals = ctx->d_totals;
            a.framebuffer = ctx->d_fb;
            a.n_first = total;
            a.pass_rays = pass_rays;
            a.partial_g = partial ? (uint32_t)partial_member : 0xFFFFFFFFu;
            a.ray_offset = partial ? pseg.offset : 0u;
            a.sub_count = partial ? pseg.sub_count : 0u;
            a.sub_first = partial ? pseg.sub_first : 0u;
            a.div_sub_count = make_fast_div(std::max(1u, a.sub_count));
            a.bounce = b;
            a.rays_per_pixel = rays_to_cast;
            a.div_pass_rays = make_fast_div(pass_rays);
            a.div_rays_per_pixel = make_fast_div((uint32_t)rays_to_cast);
            for (int g = 0; g < group; g++) {
                a.gen_seed[g] = 709579u * (uint32_t)remaining[g];                      // scene.cu:81
                a.proc_seed[g] = 279220567u * (uint32_t)(remaining[g] * 20 + b);       // raytracing.cu:89,236
            }
            a.pcg_inc = pcg_inc;
            a.key_bits = key_bits;
            a.hits = ctx->d_hits + qoff;
            a.trace_batch = ctx->trace_batch;
            a.trace_refill_below = ctx->trace_refill_below;
            a.trace_node_min = ctx->trace_node_min;
            a.entry_lists = use_entries ? (const uint4 *)ctx->d_entry : nullptr;
            a.entry_tiles_x = tiles_x;
            a.bounces_total = ctx->bounces;
            a.tail_below = use_tail ? ctx->tail_below : 0u;
            // the live count of bounce b > 0 exists only on the device: size the grids for the pass's upper bound
            const int shade_grid = (int)std::min<size_t>((size_t)shade_grid_max, ((size_t)total + kShadeBlock - 1) / kShadeBlock);
            const int t_grid = (int)std::min<size_t>((size_t)grid, ((size_t)total + kBlock - 1) / kBlock);
            if (use_tail && b >= 2) {  // the wave's stragglers in one launch (a no-op until few enough rays are left)
                const int tail_grid = (int)std::min<size_t>((size_t)ctx->sm_count * kTailCtasPerSM, ((size_t)std::min<uint32_t>(ctx->tail_below, total) + kBlock - 1) / kBlock);
                if ((rc = mark(0)) != RT_OK) return fail(ctx, rc, "event record failed");
                tail_kernel<<<std::max(1, tail_grid), kBlock, 0, cur_stream>>>(ctx->ds, a);
                if ((rc = mark(0)) != RT_OK) return fail(ctx, rc, "event record failed");
                launches++;
            }
            if ((rc = mark(0)) != RT_OK) return fail(ctx, rc, "event record failed");
            if (b == 0) launch_trace<true>(ctx, cur_stream, collect, t_grid, a);
            else launch_trace<false>(ctx, cur_stream, collect, t_grid, a);
            if ((rc = mark(0)) != RT_OK) return fail(ctx, rc, "event record failed");
            if ((rc = mark(2)) != RT_OK) return fail(ctx, rc, "event record failed");
            if (b == 0) shade_kernel<true><<<shade_grid, kShadeBlock, 0, cur_stream>>>(ctx->ds, a);
            else shade_kernel<false><<<shade_grid, kShadeBlock, 0, cur_stream>>>(ctx->ds, a);
            if ((rc = mark(2)) != RT_OK) return fail(ctx, rc, "event record failed");
            launches += 2;
            order = nullptr;
            if (a.keys_out) {
                uint32_t *sk, *sv;
                if ((rc = mark(1)) != RT_OK) return fail(ctx, rc, "event record failed");
                if ((rc = enqueue_sort(ctx, &ctx->d_pc->count[b + 1], total, key_bits, true, &sk, &sv, &launches)) != RT_OK) return rc;
                if ((rc = mark(1)) != RT_OK) return fail(ctx, rc, "event record failed");
                order = sv;
            }
        }
        passes_done += group;
    }
    if (pipes > 1) {
        for (int i = 0; i < pipes - 1; i++) {
            RT_CUDA(ctx, cudaEventRecord(ctx->ev_join[i], ctx->side[i]));
            RT_CUDA(ctx, cudaStreamWaitEvent(ctx->stream, ctx->ev_join[i], 0));
        }
    }
    RT_CUDA(ctx, cudaEventRecord(ctx->ev_stop, ctx->stream));
    RT_CUDA(ctx, cudaGetLastError());
    RT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    pipe_guard.armed = false;  // the main stream waited for every pipe (ev_join)

    if (stats) {
        std::memset(stats, 0, sizeof(*stats));
        float ms = 0;
        RT_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->ev_start, ctx->ev_stop));
        RenderTotals t;
        RT_CUDA(ctx, cudaMemcpy(&t, ctx->d_totals, sizeof(t), cudaMemcpyDeviceToHost));
        stats->loop_seconds = ms * 1e-3;
        stats->primary_rays = primaries;
        stats->live_ray_bounces = (int64_t)t.live_ray_bounces;
        stats->internal_visits = (int64_t)t.internal_visits;
        stats->triangle_tests = (int64_t)t.triangle_tests;
        stats->kernel_launches = launches;
        stats->passes = passes_done;
        for (int b = 0; b < 32; b++) {
            stats->live_per_bounce[b] = (int64_t)t.live_per_bounce[b];
            if (sort && b >= 1) stats->sorted_pairs += (int64_t)t.live_per_bounce[b];
        }
        for (size_t k = 0; k + 1 < ev_kind.size(); k += 2) {
            float e_ms = 0;
            RT_CUDA(ctx, cudaEventElapsedTime(&e_ms, ctx->ev_pool[k], ctx->ev_pool[k + 1]));
            (ev_kind[k] == 0 ? stats->trace_ms : ev_kind[k] == 2 ? stats->shade_ms : stats->sort_ms) += e_ms;
            if (ev_kind[k] == 0) stats->trace_ms_per_bounce[std::min(ev_bounce[k], 31)] += e_ms;
            if (ev_kind[k] == 2) stats->shade_ms_per_bounce[std::min(ev_bounce[k], 31)] += e_ms;
        }
        stats->total_seconds = std::chrono::duration<double>(std::chrono::high_resolution_clock::now() - wall0).count();
    }
    return RT_OK;
}

}  // namespace

extern "C" int rt_render_device(rt_ctx *ctx, const rt_render_opts *opts, float **out_dev, rt_stats *stats)
{
    int rc = render_impl(ctx, opts, stats);
    if (rc == RT_OK && out_dev) *out_dev = ctx->d_fb;
    return rc;
}

extern "C" int rt_render(rt_ctx *ctx, const rt_render_opts *opts, float *out_rgb_sum, rt_stats *stats)
{
    auto wall0 = std::chrono::high_resolution_clock::now();
    if (!out_rgb_sum) return fail(ctx, RT_ERR_INVALID, "rt_render: null output buffer");
    int rc = render_impl(ctx, opts, stats);
    if (rc != RT_OK) return rc;
    RT_CUDA(ctx, cudaMemcpyAsync(out_rgb_sum, ctx->d_fb, ctx->fb_pixels * 3 * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    RT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (stats) stats->total_seconds = std::chrono::duration<double>(std::chrono::high_resolution_clock::now() - wall0).count();
    return RT_OK;
}

extern "C" int rt_framebuffer_upload(rt_ctx *ctx, const float *rgb_sum)
{
    if (!ctx || !rgb_sum) return fail(ctx, RT_ERR_INVALID, "rt_framebuffer_upload: null argument");
    if (!ctx->has_scene) return fail(ctx, RT_ERR_INVALID, "rt_framebuffer_upload: no scene uploaded");
    RT_CUDA(ctx, cudaSetDevice(ctx->device));
    int rc = ensure_framebuffer(ctx);
    if (rc != RT_OK) return rc;
    RT_CUDA(ctx, cudaMemcpyAsync(ctx->d_fb, rgb_sum, ctx->fb_pixels * 3 * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
    RT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return RT_OK;
}

extern "C" int rt_postprocess(rt_ctx *ctx, int bloom, float *out_rgb_sum, uint8_t *out_rgb8)
{
    if (!ctx) return fail(nullptr, RT_ERR_INVALID, "rt_postprocess: null context");
    if (!ctx->d_fb || !ctx->fb_pixels) return fail(ctx, RT_ERR_INVALID, "rt_postprocess: nothing rendered");
    RT_CUDA(ctx, cudaSetDevice(ctx->device));
    const int w = ctx->ds.width, h = ctx->ds.height;
    const size_t px = ctx->fb_pixels;
    const int grid = (int)((px + 127) / 128);
    if (bloom) {
        const float threshold = (float)(0.7 * ctx->ray_count);  // raytracing.cu:376
        bloom_h_kernel<<<grid, 128, 0, ctx->stream>>>(ctx->d_fb, ctx->d_fb_tmp, threshold, w, h);
        bloom_v_add_kernel<<<grid, 128, 0, ctx->stream>>>(ctx->d_fb_tmp, ctx->d_fb, ctx->d_fb_tmp2, w, h);
        std::swap(ctx->d_fb, ctx->d_fb_tmp2);
    }
    if (out_rgb_sum) RT_CUDA(ctx, cudaMemcpyAsync(out_rgb_sum, ctx->d_fb, px * 3 * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    if (out_rgb8) {
        const float scale = ctx->exposure / (float)ctx->ray_count;  // raytracing.cu:290
        tonemap_kernel<<<(int)((px * 3 + 255) / 256), 256, 0, ctx->stream>>>(ctx->d_fb, ctx->d_rgb8, scale, px * 3);
        RT_CUDA(ctx, cudaMemcpyAsync(out_rgb8, ctx->d_rgb8, px * 3, cudaMemcpyDeviceToHost, ctx->stream));
    }
    RT_CUDA(ctx, cudaGetLastError());
    RT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return RT_OK;
}

// ---- kernel-level entry points (parity hooks) -------------------------------------------------------------------
namespace {
struct DeviceBuffer {  // scoped scratch allocation
    void *p = nullptr;
    ~DeviceBuffer() { if (p) cudaFree(p); }
};
int hook_grid(const rt_ctx *ctx, size_t n)
{
    return (int)std::min<size_t>((size_t)ctx->sm_count * kBlocksPerSM, std::max<size_t>(1, (n + kBlock - 1) / kBlock));
}
}  // namespace

extern "C" int rt_trace_closest(rt_ctx *ctx, const float *rays6, int32_t n, int32_t with_spheres, int32_t *out_prim, float *out_t,
                                rt_stats *stats)
{
    if (!ctx || !rays6 || !out_prim || !out_t || n < 0) return fail(ctx, RT_ERR_INVALID, "rt_trace_closest: bad argument");
    if (!ctx->has_scene) return fail(ctx, RT_ERR_INVALID, "rt_trace_closest: no scene uploaded");
    if (n == 0) return RT_OK;
    RT_CUDA(ctx, cudaSetDevice(ctx->device));
    DeviceBuffer rays, prim, t;
    RT_CUDA(ctx, cudaMalloc(&rays.p, (size_t)n * 6 * sizeof(float)));
    RT_CUDA(ctx, cudaMalloc(&prim.p, (size_t)n * sizeof(int)));
    RT_CUDA(ctx, cudaMalloc(&t.p, (size_t)n * sizeof(float)));
    RT_CUDA(ctx, cudaMemcpyAsync(rays.p, rays6, (size_t)n * 6 * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
    RT_CUDA(ctx, cudaMemsetAsync(ctx->d_totals, 0, sizeof(RenderTotals), ctx->stream));
    RT_CUDA(ctx, cudaMemsetAsync(ctx->d_pc, 0, sizeof(PassCounters), ctx->stream));
    RT_CUDA(ctx, cudaEventRecord(ctx->ev_start, ctx->stream));
    {
        int grid = 0, rc = trace_grid_size(ctx, &grid);
        if (rc != RT_OK) return rc;
        grid = (int)std::min<size_t>((size_t)grid, ((size_t)n + kBlock - 1) / kBlock);
        uint32_t *cursor = &ctx->d_pc->cursor[0];
        const TraceTuning tune{ctx->trace_batch, ctx->trace_refill_below, ctx->trace_node_min};
        const float *r6 = (const float *)rays.p;
#define RT_TRACE_HOOK(STATS, NODES) trace_kernel<STATS, NODES><<<grid, kBlock, 0, ctx->stream>>>(ctx->ds, r6, (uint32_t)n, with_spheres, (int *)prim.p, (float *)t.p, cursor, tune, ctx->d_totals)
        switch (ctx->active_nodes) {
        case kNodesQ16: if (stats) RT_TRACE_HOOK(true, kNodesQ16); else RT_TRACE_HOOK(false, kNodesQ16); break;
        case kNodesW4: if (stats) RT_TRACE_HOOK(true, kNodesW4); else RT_TRACE_HOOK(false, kNodesW4); break;
        default: if (stats) RT_TRACE_HOOK(true, kNodesF32); else RT_TRACE_HOOK(false, kNodesF32); break;
        }
#undef RT_TRACE_HOOK
    }
    RT_CUDA(ctx, cudaEventRecord(ctx->ev_stop, ctx->stream));
    RT_CUDA(ctx, cudaGetLastError());
    RT_CUDA(ctx, cudaMemcpyAsync(out_prim, prim.p, (size_t)n * sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    RT_CUDA(ctx, cudaMemcpyAsync(out_t, t.p, (size_t)n * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    RT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (stats) {
        std::memset(stats, 0, sizeof(*stats));
        float ms = 0;
        RT_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->ev_start, ctx->ev_stop));
        RenderTotals tot;
        RT_CUDA(ctx, cudaMemcpy(&tot, ctx->d_totals, sizeof(tot), cudaMemcpyDeviceToHost));
        stats->loop_seconds = ms * 1e-3;
        stats->live_ray_bounces = n;
        stats->internal_visits = (int64_t)tot.internal_visits;
        stats->triangle_tests = (int64_t)tot.triangle_tests;
        stats->kernel_launches = 1;
    }
    return RT_OK;
}

extern "C" int rt_generate_rays(rt_ctx *ctx, int32_t rays_per_pixel, int32_t first, int32_t n, int32_t seed, float *out12)
{
    if (!ctx || !out12 || n < 0 || rays_per_pixel <= 0 || first < 0) return fail(ctx, RT_ERR_INVALID, "rt_generate_rays: bad argument");
    if (!ctx->has_scene) return fail(ctx, RT_ERR_INVALID, "rt_generate_rays: no scene uploaded");
    if (n == 0) return RT_OK;
    RT_CUDA(ctx, cudaSetDevice(ctx->device));
    DeviceBuffer out;
    RT_CUDA(ctx, cudaMalloc(&out.p, (size_t)n * 12 * sizeof(float)));
    RT_CUDA(ctx, cudaMemsetAsync(out.p, 0, (size_t)n * 12 * sizeof(float), ctx->stream));
    generate_kernel<<<(n + 127) / 128, 128, 0, ctx->stream>>>(ctx->ds, make_fast_div((uint32_t)rays_per_pixel), (uint32_t)first, (uint32_t)n, 709579u * (uint32_t)seed,
                                                              820957824423429ULL | 1ULL, (float *)out.p);
    RT_CUDA(ctx, cudaGetLastError());
    RT_CUDA(ctx, cudaMemcpyAsync(out12, out.p, (size_t)n * 12 * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    RT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return RT_OK;
}

extern "C" int rt_process_rays(rt_ctx *ctx, float *ray_data12, const uint32_t *seeds, int32_t n)
{
    if (!ctx || !ray_data12 || !seeds || n < 0) return fail(ctx, RT_ERR_INVALID, "rt_process_rays: bad argument");
    if (!ctx->has_scene) return fail(ctx, RT_ERR_INVALID, "rt_process_rays: no scene uploaded");
    if (n == 0) return RT_OK;
    RT_CUDA(ctx, cudaSetDevice(ctx->device));
    DeviceBuffer data, sd;
    RT_CUDA(ctx, cudaMalloc(&data.p, (size_t)n * 12 * sizeof(float)));
    RT_CUDA(ctx, cudaMalloc(&sd.p, (size_t)n * sizeof(uint32_t)));
    RT_CUDA(ctx, cudaMemcpyAsync(data.p, ray_data12, (size_t)n * 12 * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
    RT_CUDA(ctx, cudaMemcpyAsync(sd.p, seeds, (size_t)n * sizeof(uint32_t), cudaMemcpyHostToDevice, ctx->stream));
    process_kernel<<<hook_grid(ctx, n), kBlock, 0, ctx->stream>>>(ctx->ds, (float *)data.p, (const uint32_t *)sd.p, (uint32_t)n,
                                                                  820957824423429ULL | 1ULL);
    RT_CUDA(ctx, cudaGetLastError());
    RT_CUDA(ctx, cudaMemcpyAsync(ray_data12, data.p, (size_t)n * 12 * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    RT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return RT_OK;
}

extern "C" int rt_sort_pairs(rt_ctx *ctx, uint32_t *keys, uint32_t *values, int32_t n, int32_t bits)
{
    if (!ctx || !keys || !values || n < 0 || n >= (1 << 30) || bits <= 0 || bits > 32) return fail(ctx, RT_ERR_INVALID, "rt_sort_pairs: bad argument (n must be < 2^30)");
    if (n == 0) return RT_OK;
    RT_CUDA(ctx, cudaSetDevice(ctx->device));
    int rc = RT_OK;
    if ((size_t)n > ctx->sort_capacity) {  // allocate the sort buffers only (queues untouched)
        for (int i = 0; i < 2; i++) { release(ctx->d_keys[i]); release(ctx->d_vals[i]); }
        release(ctx->d_sort_scratch);
        ctx->sort_capacity = 0;
        for (int i = 0; i < 2; i++) {
            RT_CUDA(ctx, cudaMalloc(&ctx->d_keys[i], (size_t)n * sizeof(uint32_t)));
            RT_CUDA(ctx, cudaMalloc(&ctx->d_vals[i], (size_t)n * sizeof(uint32_t)));
        }
        size_t tiles = ((size_t)n + kSortTile - 1) / kSortTile;
        ctx->sort_scratch_words = (size_t)kSortMaxPasses * 256 + 16 + (size_t)kSortMaxPasses * tiles * 256;
        RT_CUDA(ctx, cudaMalloc(&ctx->d_sort_scratch, ctx->sort_scratch_words * sizeof(uint32_t)));
        ctx->sort_capacity = (size_t)n;
    }
    DeviceBuffer count;
    uint32_t n32 = (uint32_t)n;
    RT_CUDA(ctx, cudaMalloc(&count.p, sizeof(uint32_t)));
    RT_CUDA(ctx, cudaMemcpyAsync(count.p, &n32, sizeof(uint32_t), cudaMemcpyHostToDevice, ctx->stream));
    RT_CUDA(ctx, cudaMemcpyAsync(ctx->d_keys[0], keys, (size_t)n * sizeof(uint32_t), cudaMemcpyHostToDevice, ctx->stream));
    RT_CUDA(ctx, cudaMemcpyAsync(ctx->d_vals[0], values, (size_t)n * sizeof(uint32_t), cudaMemcpyHostToDevice, ctx->stream));
    uint32_t *sk = nullptr, *sv = nullptr;
    cudaEvent_t e0, e1;
    RT_CUDA(ctx, cudaEventCreate(&e0));
    RT_CUDA(ctx, cudaEventCreate(&e1));
    RT_CUDA(ctx, cudaEventRecord(e0, ctx->stream));
    rc = enqueue_sort(ctx, (const uint32_t *)count.p, (size_t)n, bits, false, &sk, &sv, nullptr);
    cudaEventRecord(e1, ctx->stream);
    if (rc != RT_OK) { cudaEventDestroy(e0); cudaEventDestroy(e1); return rc; }
    RT_CUDA(ctx, cudaMemcpyAsync(keys, sk, (size_t)n * sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
    RT_CUDA(ctx, cudaMemcpyAsync(values, sv, (size_t)n * sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
    RT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    ctx->sort_device_ms = ms;
    return RT_OK;
}
extern "C" double rt_sort_device_ms(const rt_ctx *ctx) { return ctx ? ctx->sort_device_ms : 0.0; }

// ---- bandwidth probe (bench.py's roofline denominators for the cache levels) ------------------------------------------
namespace {
__global__ void __launch_bounds__(256) probe_copy_kernel(const float4 *__restrict__ src, float4 *__restrict__ dst, size_t n, int reps)
{
    for (int r = 0; r < reps; r++)  // many sweeps inside ONE launch: a 32 MB sweep lasts a few microseconds, less than a launch gap
        for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (size_t)gridDim.x * 256) {
            float4 v;
            asm volatile("ld.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(src + i));  // volatile: every sweep loads
            dst[i] = v;
        }
}
}  // namespace

extern "C" int rt_plan_waves(const rt_render_opts *opts_in, int width, int height, int ray_count, int max_group, int pipes,
                             int64_t *out, int64_t out_capacity)
{
    if (width <= 0 || height <= 0 || ray_count < 0 || max_group < 1 || max_group > kMaxPassGroup || pipes < 1 || (!out && out_capacity > 0))
        return RT_ERR_INVALID;
    rt_render_opts opts{};
    if (opts_in) opts = *opts_in;
    const int pass_spp = opts.pass_spp > 0 ? opts.pass_spp : 20;
    const FrameShape shape{(size_t)width * (size_t)height, ray_count, pass_spp};
    if (shape.pixels * (size_t)pass_spp >= (1ull << 31)) return RT_ERR_INVALID;
    std::vector<PassSegment> segs;
    if (call_segments(shape, opts.pass_first, opts.pass_count, opts.ray_first, opts.ray_count, opts.sample_first, opts.sample_count, segs))
        return RT_ERR_INVALID;
    const std::vector<WavePlan> plan = plan_waves(segs, shape, max_group, pipes);
    int64_t at = 0;
    for (const WavePlan &wp : plan) {
        if (at + 2 + 5 * (int64_t)wp.group > out_capacity) return RT_ERR_INVALID;
        out[at++] = wp.group;
        out[at++] = wp.partial_member;
        for (int m = 0; m < wp.group; m++) {
            const PassSegment &g = wp.member[m];
            out[at++] = g.pass; out[at++] = g.offset; out[at++] = g.count; out[at++] = g.sub_first; out[at++] = g.sub_count;
        }
    }
    return (int)plan.size();
}

extern "C" int rt_probe_copy_gbs(rt_ctx *ctx, int64_t bytes, int32_t iters, double *out_gbs)
{
    if (!ctx || !out_gbs || bytes < 4096 || iters <= 0) return fail(ctx, RT_ERR_INVALID, "rt_probe_copy_gbs: bad argument");
    RT_CUDA(ctx, cudaSetDevice(ctx->device));
    DeviceBuffer a, b;
    const size_t n = (size_t)bytes / 16;
    RT_CUDA(ctx, cudaMalloc(&a.p, n * 16));
    RT_CUDA(ctx, cudaMalloc(&b.p, n * 16));
    RT_CUDA(ctx, cudaMemsetAsync(a.p, 1, n * 16, ctx->stream));
    const int grid = (int)std::min<size_t>((size_t)ctx->sm_count * 8, (n + 255) / 256);
    probe_copy_kernel<<<grid, 256, 0, ctx->stream>>>((const float4 *)a.p, (float4 *)b.p, n, 3);  // warm (and fill L2)
    RT_CUDA(ctx, cudaEventRecord(ctx->ev_start, ctx->stream));
    probe_copy_kernel<<<grid, 256, 0, ctx->stream>>>((const float4 *)a.p, (float4 *)b.p, n, iters);
    RT_CUDA(ctx, cudaEventRecord(ctx->ev_stop, ctx->stream));
    RT_CUDA(ctx, cudaGetLastError());
    RT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    float ms = 0;
    RT_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->ev_start, ctx->ev_stop));
    *out_gbs = 2.0 * (double)(n * 16) * iters / (ms * 1e-3) * 1e-9;
    return RT_OK;
}

extern "C" uint32_t rt_index_divide(uint32_t n, uint32_t d) { return rtd::fast_div(n, rtd::make_fast_div(d)); }

// ---- device BVH build: inspection hooks ---------------------------------------------------------------------------
extern "C" double rt_scene_bvh_device_ms(const rt_ctx *ctx) { return ctx ? ctx->bvh_device_ms : 0.0; }

extern "C" int rt_scene_download_bvh(rt_ctx *ctx, void *nodes32, int32_t node_capacity, int32_t *node_count, int32_t *order,
                                     int32_t order_capacity)
{
    if (!ctx || !node_count) return fail(ctx, RT_ERR_INVALID, "rt_scene_download_bvh: null argument");
    if (!ctx->has_scene || ctx->built_node_count <= 0) return fail(ctx, RT_ERR_INVALID, "rt_scene_download_bvh: the BVH was not built on the device");
    const int n = ctx->built_node_count;
    *node_count = n;
    if (nodes32 && ctx->built_nodes.empty()) {  // packed on the device: the nodes are still in the build buffer
        RT_CUDA(ctx, cudaSetDevice(ctx->device));
        ctx->built_nodes.resize((size_t)n * sizeof(BuildNode));
        RT_CUDA(ctx, cudaMemcpyAsync(ctx->built_nodes.data(), ctx->d_build, ctx->built_nodes.size(), cudaMemcpyDeviceToHost, ctx->stream));
        RT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    }
    if (nodes32) {
        if (node_capacity < n) return fail(ctx, RT_ERR_INVALID, "rt_scene_download_bvh: node buffer too small");
        std::memcpy(nodes32, ctx->built_nodes.data(), ctx->built_nodes.size());
    }
    if (order) {
        if (order_capacity < ctx->ds.tri_count) return fail(ctx, RT_ERR_INVALID, "rt_scene_download_bvh: order buffer too small");
        RT_CUDA(ctx, cudaSetDevice(ctx->device));
        RT_CUDA(ctx, cudaMemcpyAsync(order, ctx->d_order, (size_t)ctx->ds.tri_count * sizeof(int32_t), cudaMemcpyDeviceToHost, ctx->stream));
        RT_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    }
    return RT_OK;
}

// ---- rt_sort_pairs_static: the reorder pass for callers that cannot allocate ----------------------------------------------
// No counterpart in the reference. Same kernels as rt_sort_pairs
// (sort_histogram_kernel + one sort_scatter_kernel per digit), but every byte of scratch - digit histograms, tile counters,
// tile status words, the ping-pong (key, value) buffers of the intermediate digit passes - lives in ONE file-scope device
// array instead of cudaMalloc'ed buffers. Every GPU that loads the library holds the array (9.3 MiB) until the process exits.
// All calls on a device share it: it keeps the last call's contents (what must start at zero is cleared per call), and two
// calls that overlap on the GPU (two streams, two host threads, two contexts on one device) race.
namespace {
constexpr int kStaticSortMaxPairs = 1 << 20;  // RT_SORT_STATIC_MAX_PAIRS in rt_core.h
constexpr size_t kStaticSortHeadWords = (size_t)kSortMaxPasses * 256 + 16;  // histograms + tile counters (word 8 of them: the pair count)
constexpr size_t kStaticSortMaxTiles = ((size_t)kStaticSortMaxPairs + kSortTile - 1) / kSortTile;  // for the smallest tile
constexpr size_t kStaticSortStatusWords = (size_t)kSortMaxPasses * kStaticSortMaxTiles * 256;
constexpr size_t kStaticSortWords = kStaticSortHeadWords + kStaticSortStatusWords + 2 * (size_t)kStaticSortMaxPairs;
static_assert(RT_SORT_STATIC_MAX_PAIRS == kStaticSortMaxPairs, "rt_core.h and rt_core.cu disagree on the limit");
}  // namespace

// In HOST code this name is a stand-in for the symbol, not a device address: the address comes from cudaGetSymbolAddress.
static __device__ __align__(256) uint32_t g_static_sort_scratch[kStaticSortWords];

namespace {
__global__ void static_sort_set_count_kernel(uint32_t *count, uint32_t n) { *count = n; }

// Device address of the array on `device` (which must be current): looked up once per device, then cached.
int static_sort_scratch_address(rt_ctx *ctx, int device, uint32_t **out)
{
    constexpr int kMaxDevices = 64;
    static uint32_t *cached[kMaxDevices] = {};
    static std::mutex lock;
    if (device < 0 || device >= kMaxDevices) return fail(ctx, RT_ERR_INVALID, "rt_sort_pairs_static: device index beyond 63");
    std::lock_guard<std::mutex> guard(lock);
    if (!cached[device]) {
        void *ptr = nullptr;
        RT_CUDA(ctx, cudaGetSymbolAddress(&ptr, g_static_sort_scratch));
        if (!ptr) return fail(ctx, RT_ERR_CUDA, "rt_sort_pairs_static: cudaGetSymbolAddress returned a null address");
        cached[device] = static_cast<uint32_t *>(ptr);
    }
    *out = cached[device];
    return RT_OK;
}
}  // namespace

extern "C" int rt_sort_pairs_static(rt_ctx *ctx, const uint32_t *d_keys_in, const uint32_t *d_values_in, uint32_t *d_keys_out,
                                    uint32_t *d_values_out, int32_t n, int32_t bits)
{
    if (!ctx || n < 0 || bits <= 0 || bits > 32) return fail(ctx, RT_ERR_INVALID, "rt_sort_pairs_static: bad argument");
    if (n == 0) return RT_OK;  // an empty batch needs no buffers
    if (!d_keys_in || !d_values_in || !d_keys_out || !d_values_out) return fail(ctx, RT_ERR_INVALID, "rt_sort_pairs_static: null device pointer");
    if (n > kStaticSortMaxPairs)
        return fail(ctx, RT_ERR_INVALID, "rt_sort_pairs_static: n = " + std::to_string(n) + " pairs exceeds the static scratch (RT_SORT_STATIC_MAX_PAIRS = " +
                                             std::to_string(kStaticSortMaxPairs) + "); use rt_sort_pairs, which allocates");
    if (d_keys_in == d_keys_out || d_values_in == d_values_out) return fail(ctx, RT_ERR_INVALID, "rt_sort_pairs_static: inputs and outputs must not alias");
    RT_CUDA(ctx, cudaSetDevice(ctx->device));
    uint32_t *scratch = nullptr;
    int rc = static_sort_scratch_address(ctx, ctx->device, &scratch);
    if (rc != RT_OK) return rc;

    const int passes = std::min(kSortMaxPasses, std::max(1, (bits + 7) / 8));
    const int items = 22, tile_keys = kSortBlock * items;  // the digit pass of the default configuration (ballot ranking)
    const uint32_t tiles = (uint32_t)(((size_t)n + tile_keys - 1) / tile_keys);
    SortArgs s;
    s.hist = scratch;
    s.tile_counter = scratch + kSortMaxPasses * 256;
    s.count_ptr = s.tile_counter + 8;
    s.tile_status = scratch + kStaticSortHeadWords;
    s.passes = passes;
    s.max_tiles = tiles;  // stride of the per-pass status arrays
    uint32_t *tmp_keys = scratch + kStaticSortHeadWords + kStaticSortStatusWords, *tmp_vals = tmp_keys + kStaticSortMaxPairs;
    cudaStream_t stream = ctx->stream;
    // the array holds the previous call's histograms, counters and status flags: clear them (the ping-pong part is overwritten
    // before it is read)
    RT_CUDA(ctx, cudaMemsetAsync(scratch, 0, (kStaticSortHeadWords + (size_t)passes * tiles * 256) * sizeof(uint32_t), stream));
    static_sort_set_count_kernel<<<1, 1, 0, stream>>>(const_cast<uint32_t *>(s.count_ptr), (uint32_t)n);
    const int hist_blocks = (int)std::min<size_t>((size_t)ctx->sm_count * 8, std::max<size_t>(1, ((size_t)n + kSortBlock - 1) / kSortBlock));
    sort_histogram_kernel<<<hist_blocks, kSortBlock, 0, stream>>>(d_keys_in, s);
    constexpr int kStage = kSortBlock * 22 * 8;
    RT_CUDA(ctx, cudaFuncSetAttribute(sort_scatter_kernel<2, 22>, cudaFuncAttributeMaxDynamicSharedMemorySize, kStage));
    // in -> ... -> out through the one temporary pair: the LAST pass must land in the caller's output
    const uint32_t *src_k = d_keys_in, *src_v = d_values_in;
    for (int p = 0; p < passes; p++) {
        const bool to_out = ((passes - 1 - p) & 1) == 0;
        uint32_t *dst_k = to_out ? d_keys_out : tmp_keys, *dst_v = to_out ? d_values_out : tmp_vals;
        sort_scatter_kernel<2, 22><<<tiles, kSortBlock, kStage, stream>>>(src_k, src_v, dst_k, dst_v, s, p);
        src_k = dst_k;
        src_v = dst_v;
    }
    RT_CUDA(ctx, cudaGetLastError());
    RT_CUDA(ctx, cudaStreamSynchronize(stream));
    return RT_OK;
}

// ---- several GPUs of one box (rt_multi_*) ----------------------------------------------------------------------------
#include "multi_device.cuh"
