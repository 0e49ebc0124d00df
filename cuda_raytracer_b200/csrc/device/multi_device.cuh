// Several GPUs of one box behind the C ABI (rt_multi_*, include/rt_core.h): the scene is replicated, every pixel's samples
// are cut into equal shares (one per device, rt_render_opts.sample_first / sample_count), every device renders
// its share from its own host thread, and the float accumulation buffers are summed into device 0's with ONE ncclReduce
// over NVLink (the reference's gpu_raytrace, raytracing.cu:170-284, is single-GPU). Included by rt_core.cu.
//
// NCCL is loaded at run time (dlopen "libnccl.so.2"): the single-GPU library has no NCCL dependency, and inside a process
// that already holds a copy (torch bundles one) the same copy is used. When the device list names one device more than
// once (a box with fewer GPUs than shares: the round-end test box has one) NCCL cannot form the communicator; the sum is
// then a plain add kernel on device 0, which reads the other contexts' buffers directly (same device). The same kernels over
// peer access replace NCCL when RT_MULTI_REDUCE=p2p is set (front-end word reduce=p2p): no communicator start-up.
#pragma once

#include <dlfcn.h>

#include <thread>

namespace {

struct NcclApi {
    // minimal prototypes of the NCCL 2.x entry points used (enum values are part of its ABI: ncclFloat32 = 7, ncclSum = 0)
    typedef int (*CommInitAll)(void **comms, int ndev, const int *devlist);
    typedef int (*CommDestroy)(void *comm);
    typedef int (*Reduce)(const void *send, void *recv, size_t count, int datatype, int op, int root, void *comm, cudaStream_t stream);
    typedef int (*Group)(void);
    typedef const char *(*ErrorString)(int);
    void *handle = nullptr;
    CommInitAll comm_init_all = nullptr;
    CommDestroy comm_destroy = nullptr;
    Reduce reduce = nullptr;
    Group group_start = nullptr, group_end = nullptr;
    ErrorString error_string = nullptr;
    bool load()
    {
        if (handle) return true;
        for (const char *name : {"libnccl.so.2", "libnccl.so"}) {
            handle = dlopen(name, RTLD_NOW | RTLD_LOCAL);
            if (handle) break;
        }
        if (!handle) return false;
        comm_init_all = (CommInitAll)dlsym(handle, "ncclCommInitAll");
        comm_destroy = (CommDestroy)dlsym(handle, "ncclCommDestroy");
        reduce = (Reduce)dlsym(handle, "ncclReduce");
        group_start = (Group)dlsym(handle, "ncclGroupStart");
        group_end = (Group)dlsym(handle, "ncclGroupEnd");
        error_string = (ErrorString)dlsym(handle, "ncclGetErrorString");
        return comm_init_all && comm_destroy && reduce && group_start && group_end;
    }
};

__global__ void __launch_bounds__(256) add_framebuffer_kernel(float *__restrict__ dst, const float *__restrict__ src, size_t n)
{
    for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (size_t)gridDim.x * 256) dst[i] += src[i];
}

}  // namespace

struct rt_multi {
    std::vector<rt_ctx *> ctx;
    std::vector<int> devices;
    std::vector<void *> comms;  // one NCCL communicator per context; empty = add-kernel path
    NcclApi nccl;
    bool use_nccl = false;
    std::string error;
    cudaEvent_t ev_r0 = nullptr, ev_r1 = nullptr;
    std::vector<cudaEvent_t> ev_done;  // per context: its share is rendered (add-kernel path)
    double reduce_ms = 0;
};

namespace {
int mfail(rt_multi *m, int status, const std::string &msg)
{
    if (m) m->error = msg;
    rt::set_global_error(msg);
    return status;
}

// run f(i) for every context on its own host thread (the calling thread takes context 0); returns the first failure
template <class F>
int for_each_device(rt_multi *m, F f)
{
    const int n = (int)m->ctx.size();
    std::vector<int> rc((size_t)n, RT_OK);
    std::vector<std::thread> workers;
    for (int i = 1; i < n; i++) workers.emplace_back([&, i] { rc[i] = f(i); });
    rc[0] = f(0);
    for (std::thread &t : workers) t.join();
    for (int i = 0; i < n; i++)
        if (rc[i] != RT_OK) return mfail(m, rc[i], std::string("device ") + std::to_string(m->devices[i]) + ": " + m->ctx[i]->error);
    return RT_OK;
}
}  // namespace

extern "C" void rt_multi_destroy(rt_multi *m)
{
    if (!m) return;
    for (size_t i = 0; i < m->comms.size(); i++)
        if (m->comms[i]) m->nccl.comm_destroy(m->comms[i]);
    if (!m->ctx.empty() && m->ctx[0]) cudaSetDevice(m->ctx[0]->device);
    if (m->ev_r0) cudaEventDestroy(m->ev_r0);
    if (m->ev_r1) cudaEventDestroy(m->ev_r1);
    for (size_t i = 0; i < m->ev_done.size(); i++)
        if (m->ev_done[i]) { cudaSetDevice(m->ctx[i]->device); cudaEventDestroy(m->ev_done[i]); }
    for (rt_ctx *c : m->ctx) rt_destroy(c);
    delete m;
}

extern "C" int rt_multi_create(rt_multi **out, const int *devices, int n)
{
    if (!out || !devices || n <= 0 || n > 64) return mfail(nullptr, RT_ERR_INVALID, "rt_multi_create: need 1..64 devices");
    *out = nullptr;
    rt_multi *m = new rt_multi();
    bool distinct = true;
    for (int i = 0; i < n; i++) {
        for (int j = 0; j < i; j++) distinct = distinct && devices[i] != devices[j];
        rt_ctx *c = nullptr;
        int rc = rt_create(&c, devices[i]);
        if (rc != RT_OK) { const std::string msg = rt::global_error(); rt_multi_destroy(m); return mfail(nullptr, rc, msg); }
        m->ctx.push_back(c);
        m->devices.push_back(devices[i]);
    }
    cudaSetDevice(devices[0]);
    if (cudaEventCreate(&m->ev_r0) != cudaSuccess || cudaEventCreate(&m->ev_r1) != cudaSuccess) {
        rt_multi_destroy(m);
        return mfail(nullptr, RT_ERR_CUDA, "rt_multi_create: event creation failed");
    }
    // RT_MULTI_REDUCE=p2p: sum the buffers with add kernels on device 0 that read the peers' memory over NVLink (peer access)
    // instead of NCCL. For one-shot use (the command line): ncclCommInitAll + the first collective took ~3.3 s on an 8-GPU box
    // (GPU Took 4.27 s for a 0.48 s loop), more than a whole 4K frame. NCCL stays the default (the reduce itself is 0.8 ms).
    const char *reduce_env = std::getenv("RT_MULTI_REDUCE");
    const bool p2p = reduce_env && std::strcmp(reduce_env, "p2p") == 0;
    if (n > 1 && distinct && !p2p) {
        if (!m->nccl.load()) { rt_multi_destroy(m); return mfail(nullptr, RT_ERR_CUDA, "rt_multi_create: libnccl.so.2 not found (needed to combine the devices' framebuffers)"); }
        m->comms.assign((size_t)n, nullptr);
        const int r = m->nccl.comm_init_all(m->comms.data(), n, devices);
        if (r != 0) {
            const std::string msg = std::string("ncclCommInitAll: ") + (m->nccl.error_string ? m->nccl.error_string(r) : "error");
            m->comms.clear();
            rt_multi_destroy(m);
            return mfail(nullptr, RT_ERR_CUDA, msg);
        }
        m->use_nccl = true;
    } else if (n > 1) {  // contexts 1.. hand their buffers to context 0's add kernel: one device named several times, or peers
        m->ev_done.assign((size_t)n, nullptr);
        for (int i = 0; i < n; i++) {
            if (!distinct && devices[i] != devices[0]) { rt_multi_destroy(m); return mfail(nullptr, RT_ERR_INVALID, "rt_multi_create: a device list with repeats must name one device only"); }
            if (distinct && i > 0) {
                int can = 0;
                cudaSetDevice(devices[0]);
                cudaDeviceCanAccessPeer(&can, devices[0], devices[i]);
                const cudaError_t e = can ? cudaDeviceEnablePeerAccess(devices[i], 0) : cudaErrorPeerAccessUnsupported;
                if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) { rt_multi_destroy(m); return mfail(nullptr, RT_ERR_CUDA, "rt_multi_create: RT_MULTI_REDUCE=p2p needs peer access from the first device to every other"); }
                cudaGetLastError();
            }
            cudaSetDevice(devices[i]);
            cudaEventCreateWithFlags(&m->ev_done[i], cudaEventDisableTiming);
        }
    }
    *out = m;
    return RT_OK;
}

extern "C" int rt_multi_device_count(const rt_multi *m) { return m ? (int)m->ctx.size() : 0; }
extern "C" const char *rt_multi_last_error(const rt_multi *m) { return m ? m->error.c_str() : rt::global_error(); }
extern "C" rt_ctx *rt_multi_context(rt_multi *m, int index) { return (m && index >= 0 && index < (int)m->ctx.size()) ? m->ctx[index] : nullptr; }

extern "C" int rt_multi_set_node_mode(rt_multi *m, int mode)
{
    if (!m) return RT_ERR_INVALID;
    for (rt_ctx *c : m->ctx) {
        int rc = rt_set_node_mode(c, mode);
        if (rc != RT_OK) return mfail(m, rc, c->error);
    }
    return RT_OK;
}

extern "C" int rt_multi_scene_upload(rt_multi *m, const rt_scene_desc *scene)
{
    if (!m || !scene) return mfail(m, RT_ERR_INVALID, "rt_multi_scene_upload: null argument");
    return for_each_device(m, [&](int i) { return rt_scene_upload(m->ctx[i], scene); });
}

// The frame in device 0's framebuffer (un-normalised sums); out_rgb_sum may be NULL (rt_multi_postprocess follows).
extern "C" int rt_multi_render(rt_multi *m, const rt_render_opts *opts_in, float *out_rgb_sum, rt_stats *stats)
{
    if (!m) return mfail(nullptr, RT_ERR_INVALID, "rt_multi_render: null handle");
    const int n = (int)m->ctx.size();
    rt_ctx *c0 = m->ctx[0];
    if (!c0->has_scene) return mfail(m, RT_ERR_INVALID, "rt_multi_render: no scene uploaded");
    auto wall0 = std::chrono::high_resolution_clock::now();
    rt_render_opts base{};
    if (opts_in) base = *opts_in;
    base.sort = opts_in ? opts_in->sort : 1;
    // What this call covers, cut into n equal shares. Default (and with a sample range of the caller's): SAMPLES of every pixel,
    // so that each share spans the whole image and equal shares cost the same. A pass range or ray range of the caller's is
    // honoured as a range of the ray sequence.
    const int pass_spp = base.pass_spp > 0 ? base.pass_spp : 20;
    const int64_t pixels = (int64_t)c0->ds.width * c0->ds.height;
    const int total_passes = (c0->ray_count + pass_spp - 1) / pass_spp;
    const bool by_rays = base.sample_count <= 0 && (base.ray_count > 0 || base.pass_first > 0 || base.pass_count > 0);
    int64_t first = 0, count = c0->ray_count, past_end = c0->ray_count;  // in samples per pixel, or in rays
    if (by_rays) {
        count = past_end = pixels * c0->ray_count;
        if (base.ray_count > 0) {
            first = std::max<int64_t>(0, base.ray_first);
            count = std::max<int64_t>(0, std::min(first + base.ray_count, past_end) - first);
        } else {
            const int p0 = std::min(std::max(0, base.pass_first), total_passes);
            const int p1 = base.pass_count > 0 ? std::min(total_passes, p0 + base.pass_count) : total_passes;
            first = pixels * std::min<int64_t>((int64_t)p0 * pass_spp, c0->ray_count);
            count = pixels * std::min<int64_t>((int64_t)p1 * pass_spp, c0->ray_count) - first;
        }
    } else if (base.sample_count > 0) {
        first = std::max(0, base.sample_first);
        count = std::max<int64_t>(0, std::min<int64_t>(first + base.sample_count, c0->ray_count) - first);
    }
    std::vector<rt_stats> st((size_t)n);
    int rc = for_each_device(m, [&](int i) {
        rt_render_opts o = base;
        const int64_t lo = first + count * i / n, hi = first + count * (i + 1) / n;
        o.pass_first = o.pass_count = 0;
        o.ray_first = o.ray_count = 0;
        o.sample_first = o.sample_count = 0;
        // an empty share: a range past the end renders nothing and leaves a cleared framebuffer
        if (by_rays) { o.ray_first = hi > lo ? lo : past_end; o.ray_count = std::max<int64_t>(1, hi - lo); }
        else { o.sample_first = (int32_t)(hi > lo ? lo : past_end); o.sample_count = (int32_t)std::max<int64_t>(1, hi - lo); }
        int r = render_impl(m->ctx[i], &o, &st[i]);
        if (r == RT_OK && !m->ev_done.empty()) {
            cudaSetDevice(m->ctx[i]->device);
            r = cudaEventRecord(m->ev_done[i], m->ctx[i]->stream) == cudaSuccess ? RT_OK : RT_ERR_CUDA;
        }
        return r;
    });
    if (rc != RT_OK) return rc;

    // combine: every context's stream has finished its share (render_impl synchronises)
    const size_t n_floats = c0->fb_pixels * 3;
    cudaSetDevice(c0->device);
    cudaEventRecord(m->ev_r0, c0->stream);
    if (m->use_nccl) {
        int r = m->nccl.group_start();
        for (int i = 0; i < n && r == 0; i++)
            r = m->nccl.reduce(m->ctx[i]->d_fb, m->ctx[i]->d_fb, n_floats, /*ncclFloat32*/ 7, /*ncclSum*/ 0, 0, m->comms[i], m->ctx[i]->stream);
        const int e = m->nccl.group_end();
        if (r == 0) r = e;
        if (r != 0) return mfail(m, RT_ERR_CUDA, std::string("ncclReduce: ") + (m->nccl.error_string ? m->nccl.error_string(r) : "error"));
    } else {
        const int grid = (int)std::min<size_t>((size_t)c0->sm_count * 8, (n_floats + 255) / 256);
        for (int i = 1; i < n; i++) {
            cudaStreamWaitEvent(c0->stream, m->ev_done[i], 0);
            add_framebuffer_kernel<<<grid, 256, 0, c0->stream>>>(c0->d_fb, m->ctx[i]->d_fb, n_floats);
        }
    }
    cudaSetDevice(c0->device);
    cudaEventRecord(m->ev_r1, c0->stream);
    for (int i = 0; i < n; i++) {
        cudaSetDevice(m->ctx[i]->device);
        if (cudaStreamSynchronize(m->ctx[i]->stream) != cudaSuccess) return mfail(m, RT_ERR_CUDA, std::string("rt_multi_render: ") + cudaGetErrorString(cudaGetLastError()));
    }
    cudaSetDevice(c0->device);
    float red_ms = 0;
    cudaEventElapsedTime(&red_ms, m->ev_r0, m->ev_r1);
    m->reduce_ms = red_ms;
    if (out_rgb_sum) {
        if (cudaMemcpyAsync(out_rgb_sum, c0->d_fb, n_floats * sizeof(float), cudaMemcpyDeviceToHost, c0->stream) != cudaSuccess ||
            cudaStreamSynchronize(c0->stream) != cudaSuccess)
            return mfail(m, RT_ERR_CUDA, std::string("rt_multi_render: download: ") + cudaGetErrorString(cudaGetLastError()));
    }
    if (stats) {
        *stats = st[0];
        for (int i = 1; i < n; i++) {
            stats->loop_seconds = std::max(stats->loop_seconds, st[i].loop_seconds);  // the devices run side by side
            stats->primary_rays += st[i].primary_rays;
            stats->live_ray_bounces += st[i].live_ray_bounces;
            stats->internal_visits += st[i].internal_visits;
            stats->triangle_tests += st[i].triangle_tests;
            stats->sorted_pairs += st[i].sorted_pairs;
            stats->kernel_launches += st[i].kernel_launches;
            stats->passes = std::max(stats->passes, st[i].passes);
            stats->trace_ms = std::max(stats->trace_ms, st[i].trace_ms);
            stats->shade_ms = std::max(stats->shade_ms, st[i].shade_ms);
            stats->sort_ms = std::max(stats->sort_ms, st[i].sort_ms);
            for (int b = 0; b < 32; b++) stats->live_per_bounce[b] += st[i].live_per_bounce[b];
        }
        stats->loop_seconds += red_ms * 1e-3;
        stats->total_seconds = std::chrono::duration<double>(std::chrono::high_resolution_clock::now() - wall0).count();
    }
    return RT_OK;
}

extern "C" double rt_multi_reduce_ms(const rt_multi *m) { return m ? m->reduce_ms : 0.0; }

extern "C" int rt_multi_postprocess(rt_multi *m, int bloom, float *out_rgb_sum, uint8_t *out_rgb8)
{
    if (!m) return mfail(nullptr, RT_ERR_INVALID, "rt_multi_postprocess: null handle");
    int rc = rt_postprocess(m->ctx[0], bloom, out_rgb_sum, out_rgb8);
    if (rc != RT_OK) m->error = m->ctx[0]->error;
    return rc;
}
