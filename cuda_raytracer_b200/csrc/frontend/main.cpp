// `raytracing <scene> [no_sort|cpu|no_gpu|no_bvh]` — drop-in command line of the reference program
// (/root/reference/raytracing.cu:305-398): same argument grammar, exit codes (1 usage / CUDA error, 2 no hardware),
// stdout lines and ./raytracing.png output. The render itself goes through the C ABI (rt_core.h); there is no CPU
// renderer in this product, so `cpu` without a GPU path is refused with exit code 2 like "no hardware".
// Extra words the reference would ignore: out=<path>, stats, no_bloom, device=<n>, gpus=<n> (split the frame's rays over the first
// n GPUs, one NCCL reduce; reduce=p2p sums over peer access instead), host_bvh (build the BVH on the host as the reference does; the default is the same binned-SAH
// build on the device, ~70x faster), gpu_bvh (the default, kept for old command lines), sbvh (host tree with spatial splits).
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "../../../include/rt_scene.h"

int main(int argc, char **argv)
{
    if (argc < 2) {
        std::printf("Usage: %s <scene>\n", argv[0]);
        return 1;
    }
    bool sort = true, cpu = false, gpu = true, bvh = true, bloom = true, stats_on = false, gpu_bvh = true, sbvh = false;
    int device = 0, gpus = 1;
    std::string out_path = "raytracing.png";
    for (int i = 2; i < argc; i++) {
        const char *w = argv[i];
        if (!std::strcmp(w, "no_sort")) sort = false;
        else if (!std::strcmp(w, "cpu")) cpu = true;
        else if (!std::strcmp(w, "no_bvh")) bvh = false;
        else if (!std::strcmp(w, "no_gpu")) gpu = false;
        else if (!std::strcmp(w, "no_bloom")) bloom = false;
        else if (!std::strcmp(w, "stats")) stats_on = true;
        else if (!std::strcmp(w, "gpu_bvh")) gpu_bvh = true;
        else if (!std::strcmp(w, "host_bvh")) gpu_bvh = false;
        else if (!std::strcmp(w, "sbvh")) { sbvh = true; gpu_bvh = false; }
        else if (!std::strncmp(w, "gpus=", 5)) gpus = std::max(1, std::atoi(w + 5));
        else if (!std::strcmp(w, "reduce=p2p")) setenv("RT_MULTI_REDUCE", "p2p", 1);  // peer-access add kernels instead of NCCL (no ~3 s start-up)
        else if (!std::strncmp(w, "out=", 4)) out_path = w + 4;
        else if (!std::strncmp(w, "device=", 7)) device = std::atoi(w + 7);
    }
    if (!gpu) {  // covers `cpu no_gpu` and plain `no_gpu`: nothing this build can render with
        std::printf(cpu ? "This build has no CPU renderer (use the reference's cpu path)\n" : "No raytracing hardware specified\n");
        return 2;
    }
    if (cpu) std::printf("note: `cpu` ignored, this build renders on the GPU only\n");

    rt_host_scene *scene = nullptr;
    if (rt_scene_load(argv[1], bvh ? (gpu_bvh ? 2 : (sbvh ? 3 : 1)) : 0, 1, &scene) != RT_OK) {
        std::printf("Error loading scene: %s\n", rt_last_error(nullptr));
        return 1;
    }
    rt_scene_desc desc;
    rt_scene_describe(scene, &desc);

    auto t0 = std::chrono::high_resolution_clock::now();
    rt_ctx *ctx = nullptr;
    rt_multi *multi = nullptr;
    int rc;
    if (gpus > 1) {  // devices device, device + 1, ...
        std::vector<int> devs;
        for (int i = 0; i < gpus; i++) devs.push_back(device + i);
        rc = rt_multi_create(&multi, devs.data(), gpus);
        if (rc == RT_OK) ctx = rt_multi_context(multi, 0);
        if (rc == RT_OK) rc = rt_multi_scene_upload(multi, &desc);
    } else {
        rc = rt_create(&ctx, device);
        if (rc == RT_OK) rc = rt_scene_upload(ctx, &desc);
    }
    if (rc == RT_OK && desc.triangles_form == 1) {  // the reference's two build lines (scene.cu:1023-1027), from the device build
        int32_t node_count = 0;
        rt_scene_download_bvh(ctx, nullptr, 0, &node_count, nullptr, 0);
        std::printf("BVH Took %gms\nNode count: %d\n", rt_scene_bvh_device_ms(ctx), node_count);
    }
    rt_render_opts opts{};
    opts.sort = sort ? 1 : 0;
    opts.collect_stats = stats_on ? 1 : 0;
    rt_stats st{};
    if (rc == RT_OK) rc = multi ? rt_multi_render(multi, &opts, nullptr, &st) : rt_render_device(ctx, &opts, nullptr, &st);
    std::vector<uint8_t> rgb8((size_t)desc.width * desc.height * 3);
    if (rc == RT_OK) {
        double gpu_s = std::chrono::duration<double>(std::chrono::high_resolution_clock::now() - t0).count();
        std::printf("GPU Took %gs\n", gpu_s);  // same span as the reference's timer: allocation, upload, passes
        rc = multi ? rt_multi_postprocess(multi, bloom ? 1 : 0, nullptr, rgb8.data()) : rt_postprocess(ctx, bloom ? 1 : 0, nullptr, rgb8.data());
    }
    auto teardown = [&]() {
        if (multi) rt_multi_destroy(multi);
        else rt_destroy(ctx);
        rt_scene_free(scene);
    };
    if (rc != RT_OK) {
        std::printf("Error %s\n", multi ? rt_multi_last_error(multi) : rt_last_error(ctx));
        teardown();
        return 1;
    }
    if (stats_on) {
        double mrays = st.live_ray_bounces / st.loop_seconds * 1e-6;
        std::printf("{\"loop_s\": %.6f, \"live_ray_bounces\": %lld, \"mrays_per_s\": %.1f, \"internal_visits\": %lld, "
                    "\"triangle_tests\": %lld, \"launches\": %d, \"passes\": %d, \"gpus\": %d, \"reduce_ms\": %.3f}\n",
                    st.loop_seconds, (long long)st.live_ray_bounces, mrays, (long long)st.internal_visits,
                    (long long)st.triangle_tests, st.kernel_launches, st.passes, gpus, multi ? rt_multi_reduce_ms(multi) : 0.0);
    }
    rc = rt_write_png(out_path.c_str(), desc.width, desc.height, rgb8.data());
    teardown();
    return rc == RT_OK ? 0 : 1;
}
