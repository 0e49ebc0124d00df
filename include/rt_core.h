/* include/rt_core.h — C ABI of the B200-native render core (librtcore.so).
 *
 * This is the drop-in boundary for the reference's GPU render path. The reference has no FFI layer;
 * its de-facto operator boundary is
 *
 *     Vec3 *gpu_raytrace(const Scene *scene, bool sort)      /root/reference/raytracing.cu:170-284
 *
 * followed by the inline bloom block of main()               /root/reference/raytracing.cu:360-392
 * and the host tone map                                       /root/reference/raytracing.cu:286-303.
 *
 * Everything here is `extern "C"`, plain pointers and sizes. Caller owns every host buffer; the
 * library owns all device memory (freed by rt_destroy). Functions return RT_OK (0) or a negative
 * rt_status and never call exit(); rt_last_error() gives the message. One rt_ctx is bound to one
 * CUDA device and must be used by one host thread at a time. There is NO CPU fallback: without a
 * usable CUDA device rt_create fails with RT_ERR_NO_DEVICE.
 *
 * Array layouts are the reference's own (scene.cuh), so a maintainer can pass the fields of a
 * loaded `Scene` straight through (INTEGRATION.md shows the stub).
 */
#ifndef RT_CORE_H
#define RT_CORE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RT_API_VERSION 7

typedef enum rt_status {
    RT_OK = 0,
    RT_ERR_INVALID = -1,   /* bad argument / scene not uploaded */
    RT_ERR_NO_DEVICE = -2, /* no CUDA device, or device index out of range */
    RT_ERR_CUDA = -3,      /* a CUDA runtime call or kernel failed (message in rt_last_error) */
    RT_ERR_NOMEM = -4,
    RT_ERR_IO = -5,        /* host scene layer: file missing / malformed */
    RT_ERR_PARSE = -6      /* host scene layer: e.g. material used before definition */
} rt_status;

/* Replaces the `Scene` POD the reference hands to gpu_raytrace (scene.cuh:102-166). */
typedef struct rt_scene_desc {
    const float *spheres;              /* sphere_count x {cx, cy, cz, radius}            scene.cuh:9-13   */
    int32_t sphere_count;
    const float *triangles;            /* triangle_count x {p1, p2-p1, p3-p1, normal}    scene.cuh:27-33, render form, BVH order */
    int32_t triangle_count;
    const uint16_t *material_indices;  /* sphere_count + triangle_count entries          scene.cuh:114    */
    const float *materials;            /* material_count x {diffuse[3], metallicity, specular[3], roughness, emitted[3], ior}  scene.cuh:42-59 */
    int32_t material_count;
    const void *bvh_nodes;             /* bvh_node_count x 32 B {min[3], max[3], int child1, int child2}  scene.cuh:82-100 */
    int32_t bvh_node_count;
    const float *environment_map;      /* env_width*env_height x rgb                     scene.cuh:132    */
    int32_t env_width, env_height;
    int32_t width, height;             /* output image                                   scene.cuh:129-130 */
    int32_t ray_count;                 /* rays (samples) per pixel                       scene.cuh:157    */
    int32_t bounces;                   /*                                                scene.cuh:156    */
    float exposure;                    /*                                                scene.cuh:139    */
    float camera_position[3];          /*                                                scene.cuh:134    */
    float near_plane_top_left[3];      /* precompute_camera_data, scene.cu:62-76 */
    float scaled_right[3];
    float scaled_up[3];
    float inv_width, inv_height;
    float min_coord[3];                /* scene bounds for the reorder key               scene.cu:822-830 */
    float inv_dimensions[3];
    int32_t triangles_form;            /* 0 = as above (what the reference's load_scene leaves behind). 1 = BUILD form, file order:
                                          {p1, p2, p3, centroid} per triangle (scene.cu:520-531), material_indices in the same
                                          order, bvh_nodes ignored: the core builds the BVH on the device with the reference's
                                          binned-SAH rules (generate_bvh / maybe_split, scene.cu:866-1036). */
} rt_scene_desc;

typedef struct rt_render_opts {
    int32_t sort;          /* 1 = reorder live rays between bounces (reference default), 0 = `no_sort` (raytracing.cu:320-323) */
    int32_t pass_spp;      /* samples per pixel per pass; 0 = 20 = MAX_RAYS_PER_PIXEL_PER_PASS (raytracing.cu:109). Seeds are
                              defined per pass, so changing it changes the random streams (not the statistics). */
    int32_t pass_first;    /* render only passes [pass_first, pass_first+pass_count) of ceil(ray_count/pass_spp); */
    int32_t pass_count;    /* pass_count <= 0 = all remaining. Used for the per-GPU spp split. */
    int32_t collect_stats; /* 1 = count node visits / triangle tests (slower kernel variant) */
    int32_t stream_id;     /* 0 = reference-compatible PCG streams (raytracing.cu:89,229,236 with slot == ray id);
                              k > 0 = disjoint PCG streams: stream k's `inc` differs (random.cuh:28), for multi-GPU runs
                              that do NOT want pass-identical images */
    int32_t sort_bits;     /* key width for the reorder pass, 0 = default */
    int32_t pass_group;    /* passes traced together as one wave: 0 = auto (memory budget, at most 16), 1 = one pass per wave as the
                              reference does. Results do not depend on it beyond float summation order. With sort = 0 two
                              waves are in flight at a time on two streams of the context (environment RT_PIPES=1: one). */
    int32_t kernel_timing; /* 1 = CUDA-event pair around every bounce kernel / reorder pass (rt_stats.trace_ms, shade_ms, sort_ms); one
                              wave at a time, so that every interval is one kernel's own */
    int32_t root_start;    /* 1 = primary rays walk from the BVH root instead of their screen tile's entry records (same results) */
    int32_t single_pipe;   /* 1 = one wave at a time on one stream (as RT_PIPES=1) */
    int32_t reserved;      /* must be 0 */
    int64_t ray_first;     /* with ray_count > 0: render only rays [ray_first, ray_first + ray_count) of the frame's ray sequence */
    int64_t ray_count;     /* (pass 0's rays, then pass 1's, ...; a pass holds width*height*spp_of_that_pass rays, pixel-major) and
                              ignore pass_first / pass_count. A ray's seeds depend on its pass and its index inside the pass
                              (raytracing.cu:89,229,236, scene.cu:81), so disjoint ranges add up to the whole frame exactly as
                              disjoint pass ranges do. */
    int32_t sample_first;  /* with sample_count > 0: render only samples [sample_first, sample_first + sample_count) of EVERY pixel */
    int32_t sample_count;  /* (sample s of a pixel is ray s % pass_spp of that pixel in pass s / pass_spp) and ignore the pass and ray
                              ranges. This is the per-GPU split: every share covers the whole image, so shares of equal size cost
                              the same whatever the picture shows (a ray range of a pass covers a band of rows: sky or geometry). */
} rt_render_opts;

typedef struct rt_stats {
    double loop_seconds;        /* CUDA-event time of the pass loop on the context's stream */
    double total_seconds;       /* host wall time of the call ("GPU Took" equivalent, raytracing.cu:172-281) */
    int64_t primary_rays;
    int64_t live_ray_bounces;   /* non-early-out process_ray equivalents (SURVEY §8d metric unit) */
    int64_t internal_visits;    /* valid when collect_stats */
    int64_t triangle_tests;     /* valid when collect_stats */
    int64_t sorted_pairs;       /* (key,index) pairs pushed through the reorder pass */
    int32_t kernel_launches;    /* kernels of THIS library launched inside the loop */
    int32_t passes;
    double trace_ms;            /* CUDA-event sums per kernel class (0 unless rt_render_opts.reserved bit 0): closest-hit kernels, */
    double shade_ms;            /* shading + compaction kernels, */
    double sort_ms;             /* reorder passes */
    int64_t live_per_bounce[32];
    double trace_ms_per_bounce[32];  /* with reserved bit 0: per-bounce sums over the passes */
    double shade_ms_per_bounce[32];
} rt_stats;

typedef struct rt_ctx rt_ctx;

int rt_api_version(void);
int rt_device_count(void);
/* device: CUDA ordinal (honours CUDA_VISIBLE_DEVICES). */
int rt_create(rt_ctx **out, int device);
/* Node format the closest-hit kernel walks, takes effect at the next rt_scene_upload:
 *   0 = 64-byte two-child records with float boxes (two 256-bit loads per visit);
 *   1 = the same binary tree, 32-byte records with 15-bit quantised boxes (one 256-bit load per visit);
 *   2 = 4-wide nodes collapsed from that tree, 64 bytes, quantised boxes.
 * Same hits in every mode (boxes only grow; hits come from the exact triangle test). Default: environment RT_NODES
 * (f32 | q16 | w4), else the library's default. rt_get_node_mode returns the format of the uploaded scene. */
int rt_set_node_mode(rt_ctx *ctx, int mode);
int rt_get_node_mode(const rt_ctx *ctx);
void rt_destroy(rt_ctx *ctx);
/* ctx may be NULL for errors of rt_create / the host scene layer (thread-local). */
const char *rt_last_error(const rt_ctx *ctx);

/* Optional: page-lock a caller-owned host array (scene arrays handed to rt_scene_upload, the output of rt_render) so that the
 * copies are asynchronous DMA at PCIe speed instead of staged pageable copies (scene.cu:243-265 copies pageable memory). The
 * caller must rt_host_unpin before freeing the array. rt_scene_load pins the arrays of the scenes it returns; without a CUDA
 * device rt_host_pin returns RT_ERR_NO_DEVICE and nothing else changes. */
int rt_host_pin(const void *ptr, size_t bytes);
void rt_host_unpin(const void *ptr);

/* Replaces Scene::copy_from_cpu_async + the __constant__ symbol (scene.cu:243-265, raytracing.cu:19):
 * packs the BVH into 64-byte two-child records and the triangles into 3 x float4, uploads all. */
int rt_scene_upload(rt_ctx *ctx, const rt_scene_desc *scene);
/* After an upload with triangles_form == 1: the tree the device built, in the reference's 32-byte node layout (numbered
 * breadth-first), and order[i] = file-order id of the triangle at BVH position i. Either output may be NULL. */
int rt_scene_download_bvh(rt_ctx *ctx, void *nodes32, int32_t node_capacity, int32_t *node_count, int32_t *order, int32_t order_capacity);
/* CUDA-event time of that build (H2D of the triangles, all levels, packing into the tracer's records), 0 if the host
 * built the tree. */
double rt_scene_bvh_device_ms(const rt_ctx *ctx);
/* Change the image block without re-uploading geometry (width/height/ray_count/bounces/exposure + camera basis). */
int rt_scene_set_image(rt_ctx *ctx, const rt_scene_desc *scene);

/* Replaces gpu_raytrace (raytracing.cu:170-284): out_rgb_sum = 3*width*height floats, row-major, top row first,
 * UN-NORMALISED radiance sums, caller-owned host memory (pageable or pinned). opts/stats may be NULL. */
int rt_render(rt_ctx *ctx, const rt_render_opts *opts, float *out_rgb_sum, rt_stats *stats);
/* Same, but the sums stay in the context's device framebuffer; *out_dev (may be NULL) receives the device pointer
 * (valid until the next render / rt_destroy). Used when a collective (NCCL reduce) follows. The stream is synchronised. */
int rt_render_device(rt_ctx *ctx, const rt_render_opts *opts, float **out_dev, rt_stats *stats);

/* Replaces the bloom block (raytracing.cu:21-74,360-392) and write_framebuffer_to_output_image (:286-303), on device.
 * Operates in place on the context's device framebuffer. Any output pointer may be NULL.
 *   bloom != 0: high-pass at 0.7*ray_count, 11-tap H+V box blur, add.
 *   out_rgb_sum: 3*W*H floats after bloom; out_rgb8: 3*W*H bytes, tone-mapped. */
int rt_postprocess(rt_ctx *ctx, int bloom, float *out_rgb_sum, uint8_t *out_rgb8);
/* Load sums into the device framebuffer (e.g. after a host-side combine); n = 3*W*H floats. */
int rt_framebuffer_upload(rt_ctx *ctx, const float *rgb_sum);

/* ---- several GPUs of one box --------------------------------------------------------------------------------------
 * The reference's gpu_raytrace (raytracing.cu:170-284) drives one GPU. rt_multi_* is the same call over a list of
 * devices: the scene is replicated, the frame's ray sequence is cut into equal contiguous shares (ray_first / ray_count
 * above; every ray keeps the seeds of its pass and index, so the image equals the one-GPU image up to float summation
 * order), each device renders its share from its own host thread, and ONE ncclReduce over NVLink sums the float
 * accumulation buffers into devices[0]'s, where bloom and tone map run. The library owns the NCCL communicators
 * (ncclCommInitAll at rt_multi_create; libnccl.so.2 is loaded at run time, the one-GPU entry points do not need it).
 * A device list that names ONE device several times renders the shares on that device and sums them with a kernel
 * (no NCCL): the same host path on a box with fewer GPUs. One host thread drives an rt_multi at a time. */
typedef struct rt_multi rt_multi;
int rt_multi_create(rt_multi **out, const int *devices, int n);
void rt_multi_destroy(rt_multi *m);
int rt_multi_device_count(const rt_multi *m);
const char *rt_multi_last_error(const rt_multi *m);
/* The per-device context (index 0 holds the combined frame after rt_multi_render); owned by the rt_multi. */
rt_ctx *rt_multi_context(rt_multi *m, int index);
int rt_multi_set_node_mode(rt_multi *m, int mode);
int rt_multi_scene_upload(rt_multi *m, const rt_scene_desc *scene);
/* As rt_render: out_rgb_sum (3*W*H floats, may be NULL when rt_multi_postprocess follows) receives the combined sums. A pass
 * range or ray range in opts selects the rays to split. stats: counters summed over the devices, loop_seconds = the slowest
 * device's pass loop + the reduce (CUDA events). */
int rt_multi_render(rt_multi *m, const rt_render_opts *opts, float *out_rgb_sum, rt_stats *stats);
/* CUDA-event time of the last combine step (ncclReduce, or the add kernels), milliseconds. */
double rt_multi_reduce_ms(const rt_multi *m);
int rt_multi_postprocess(rt_multi *m, int bloom, float *out_rgb_sum, uint8_t *out_rgb8);

/* ---- kernel-level entry points (parity hooks; each runs the production device code) ---- */

/* Closest hit of n rays {ox,oy,oz,dx,dy,dz}: Scene::bvh_closest_hit_distance (scene.cu:134-241), preceded by the
 * sphere loop (scene.cu:338-372) when with_spheres != 0. Miss: prim = -1, t = 1e30. prim is the reference's primitive
 * index (sphere_count + triangle index in BVH order). */
int rt_trace_closest(rt_ctx *ctx, const float *rays6, int32_t n, int32_t with_spheres, int32_t *out_prim, float *out_t,
                     rt_stats *stats);
/* Scene::generate_initial_rays (scene.cu:78-105) for ray indices [first, first+n) of a pass: 12 floats per ray
 * {origin, direction, transmitted, collected}. seed = remaining_rays of the pass. */
int rt_generate_rays(rt_ctx *ctx, int32_t rays_per_pixel, int32_t first, int32_t n, int32_t seed, float *out_ray_data12);
/* One Scene::process_ray step (scene.cu:320-487) per record, rng = xor_srand(seeds[i]) (random.cuh:25-30); in place. */
int rt_process_rays(rt_ctx *ctx, float *ray_data12, const uint32_t *seeds, int32_t n);
/* The reorder pass on its own: stable LSD radix sort of (key, value) pairs on the low `bits` bits. Replaces
 * cub::DeviceRadixSort::SortPairs (raytracing.cu:194-195,240-246). In place on host arrays. */
int rt_sort_pairs(rt_ctx *ctx, uint32_t *keys, uint32_t *values, int32_t n, int32_t bits);
/* CUDA-event time of the sort kernels of the last rt_sort_pairs call (copies excluded), in milliseconds. */
double rt_sort_device_ms(const rt_ctx *ctx);
/* The same stable sort for callers that cannot allocate (CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc). NO COUNTERPART IN THE REFERENCE. DEVICE pointers in and out
 * (keys / values of n pairs each; inputs are not modified and must not alias the outputs); the call makes no cudaMalloc,
 * cudaFree or other allocating call. All scratch (digit histograms, tile status, the buffers of the intermediate digit
 * passes) is one static device array of the library: 9.3 MiB that every GPU which loads the library holds until the process
 * exits, sized for n <= RT_SORT_STATIC_MAX_PAIRS; a larger n is rejected with RT_ERR_INVALID and a message. Every call on a
 * device shares that array: calls that overlap on the GPU (two host threads, two contexts on one device) race, so serialise
 * them. Runs on the context's stream and returns after it has drained. */
#define RT_SORT_STATIC_MAX_PAIRS (1 << 20)
int rt_sort_pairs_static(rt_ctx *ctx, const uint32_t *d_keys_in, const uint32_t *d_values_in, uint32_t *d_keys_out,
                         uint32_t *d_values_out, int32_t n, int32_t bits);
/* Measurement aid: device-to-device copy of `bytes` (read + write counted) repeated `iters` times after a warm-up, GB/s by
 * CUDA events. With a buffer pair that fits the L2 it gives the L2 bandwidth bench.py's roofline.l2_frac is quoted against;
 * with a large one, the HBM copy bandwidth (the figure MEASURED_PEAKS.json holds). */
int rt_probe_copy_gbs(rt_ctx *ctx, int64_t bytes, int32_t iters, double *out_gbs);
/* Test hook (host only: no context, no device): the planning half of rt_render — which rays of which passes a call with these
 * options traces (its segments: whole passes, or the pieces a sample_* / ray_* / pass_* range cuts out of them) and how they
 * are grouped into waves of at most max_group (1..16) passes on `pipes` streams — for a frame of width x height pixels with
 * ray_count samples per pixel. No counterpart in the reference (its loop is one pass at a time, raytracing.cu:222-247).
 * Writes per wave {members, partial_member (-1: none; else the index of the wave's partial pass, always its last member or
 * its only one), then per member: pass, offset, count, sub_first, sub_count} as int64 values; returns the number of waves
 * (>= 0) or RT_ERR_INVALID (option out of range, out_capacity too small). */
int rt_plan_waves(const rt_render_opts *opts, int width, int height, int ray_count, int max_group, int pipes, int64_t *out,
                  int64_t out_capacity);
/* Host mirror of the kernels' index arithmetic (ray id -> pass / pixel / row: n / d by multiply-and-shift with the constants
 * the render call derives from d). Needs no device; d = 0 is treated as 1. Test hook. */
uint32_t rt_index_divide(uint32_t n, uint32_t d);

#ifdef __cplusplus
}
#endif
#endif /* RT_CORE_H */
