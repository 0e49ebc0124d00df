// oracle/ref_shim.cu — TEST INFRASTRUCTURE ONLY (never linked into the product).
//
// Compiles the UNMODIFIED reference sources where they lie (`#include "scene.cu"` resolved through
// `-I$(REF)`, nothing is copied into this repo) and exposes the reference's own functions over a
// plain C interface so that `oracle/rt_oracle.c` (the restatement that travels to the GPU box) can
// be pinned against them, and so golden fixtures can be generated (tests/golden/make_golden.py).
//
// Built by oracle/Makefile into oracle/_ref/libref_shim.so, only where /root/reference exists.
// Flags: precise host math (no -ffast-math, -ffp-contract=off, -fwrapv for SURVEY quirk Q3).
//
// Everything below is a thin caller; all arithmetic is the reference's:
//   load_scene                      /root/reference/scene.cu:569-831
//   Scene::generate_initial_rays    scene.cu:78-105
//   Scene::bvh_closest_hit_distance scene.cu:134-241
//   Scene::process_ray              scene.cu:320-487
//   equal_area_project_sphere_to_square scene.cu:284-318
//   xor_srand / xor_rand / random*  random.cuh:13-75
#include "scene.cu"

#include <cstring>
#include <cstdio>
#include <cub/device/device_radix_sort.cuh>

static_assert(sizeof(RayData) == 48, "RayData layout");
static_assert(sizeof(Triangle) == 48, "Triangle layout");
static_assert(sizeof(BvhNode) == 32, "BvhNode layout");
static_assert(sizeof(Material) == 48, "Material layout");
static_assert(sizeof(Sphere) == 16, "Sphere layout");

// GPU pass-loop seeds, raytracing.cu:89,229,236 and scene.cu:81: 64-bit products
// truncated to unsigned int by xor_srand's parameter.
static inline unsigned gpu_gen_seed(int ray_index, int seed) { return (unsigned)(ray_index * 298592570346 + 709579 * seed); }
static inline unsigned gpu_proc_seed(int ray_index, int seed) { return (unsigned)(ray_index * 4137874753 + 279220567 * seed); }

__global__ void shim_closest_hit_kernel(Scene scene, const float *rays6, int n, int with_spheres, float *out_t, int *out_idx)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Ray ray = {{rays6[6 * i], rays6[6 * i + 1], rays6[6 * i + 2]}, {rays6[6 * i + 3], rays6[6 * i + 4], rays6[6 * i + 5]}};
    float t = 1e30;
    int idx = -1;
    scene.bvh_closest_hit_distance(ray, t, idx);
    out_t[i] = t;
    out_idx[i] = idx;
}

extern "C" {

void *ref_load_scene(const char *path, int use_bvh)
{
    Scene *scene = new Scene{};
    load_scene(scene, path, use_bvh != 0);
    return scene;
}

// ints: sphere_count, triangle_count, material_count, bvh_node_count, width, height, ray_count, bounces, env_w, env_h
void ref_scene_ints(void *s, int *out)
{
    Scene *scene = (Scene *)s;
    int v[10] = {scene->sphere_count, scene->triangle_count, scene->material_count, scene->bvh_node_count, scene->width,
                 scene->height, scene->ray_count, scene->bounces, scene->environment_map_width, scene->environment_map_height};
    memcpy(out, v, sizeof(v));
}

// which: 0 spheres 1 triangles 2 material_indices 3 materials 4 bvh 5 environment_map
const void *ref_scene_array(void *s, int which)
{
    Scene *scene = (Scene *)s;
    switch (which) {
    case 0: return scene->spheres;
    case 1: return scene->triangles;
    case 2: return scene->material_indices;
    case 3: return scene->materials;
    case 4: return scene->bvh;
    case 5: return scene->environment_map;
    }
    return nullptr;
}

// floats: camera_position[3] forward[3] up[3] vertical_fov exposure min_coord[3] inv_dimensions[3]
//         scaled_right[3] scaled_up[3] near_plane_top_left[3] inv_width inv_height   (= 28 floats)
void ref_scene_floats(void *s, float *out)
{
    Scene *scene = (Scene *)s;
    const Vec3 *v3[] = {&scene->camera_position, &scene->forward, &scene->up};
    int k = 0;
    for (auto v : v3) { out[k++] = v->x; out[k++] = v->y; out[k++] = v->z; }
    out[k++] = scene->vertical_fov;
    out[k++] = scene->exposure;
    const Vec3 *w3[] = {&scene->min_coord, &scene->inv_dimensions, &scene->scaled_right, &scene->scaled_up, &scene->near_plane_top_left};
    for (auto v : w3) { out[k++] = v->x; out[k++] = v->y; out[k++] = v->z; }
    out[k++] = scene->inv_width;
    out[k++] = scene->inv_height;
}

void ref_set_image(void *s, int width, int height, int ray_count, int bounces, float exposure)
{
    Scene *scene = (Scene *)s;
    scene->width = width; scene->height = height; scene->ray_count = ray_count; scene->bounces = bounces; scene->exposure = exposure;
    scene->precompute_camera_data();
}

void ref_xor_sequence(unsigned seed, int n, unsigned *out)
{
    xor_random rng;
    xor_srand(&rng, seed);
    for (int i = 0; i < n; i++) out[i] = xor_rand(&rng);
}

// out: random01, random02, random_radians, then random_on_sphere xyz, each from a fresh xor_srand(seed)
void ref_random_floats(unsigned seed, float *out)
{
    xor_random rng;
    xor_srand(&rng, seed); out[0] = random01(&rng);
    xor_srand(&rng, seed); out[1] = random02(&rng);
    xor_srand(&rng, seed); out[2] = random_radians(&rng);
    xor_srand(&rng, seed);
    Vec3 v = random_on_sphere(&rng);
    out[3] = v.x; out[4] = v.y; out[5] = v.z;
}

void ref_project(const float *dirs3, int n, float *out_uv)
{
    for (int i = 0; i < n; i++) {
        Vec3 c = equal_area_project_sphere_to_square({dirs3[3 * i], dirs3[3 * i + 1], dirs3[3 * i + 2]});
        out_uv[2 * i] = c.x;
        out_uv[2 * i + 1] = c.y;
    }
}

int ref_ray_aabb(const float *aabb6, const float *ray6, float tmax, float *tmin_out)
{
    Aabb box;
    box.min_bound = {aabb6[0], aabb6[1], aabb6[2]};
    box.max_bound = {aabb6[3], aabb6[4], aabb6[5]};
    Ray ray = {{ray6[0], ray6[1], ray6[2]}, {ray6[3], ray6[4], ray6[5]}};
    Vec3 n_inv = {1 / ray.direction.x, 1 / ray.direction.y, 1 / ray.direction.z};
    return ray_aabb_intersection(box, ray, n_inv, *tmin_out, tmax) ? 1 : 0;
}

// BVH-only closest hit (the function north_star's dumped-ray check names), host execution.
void ref_closest_hit(void *s, const float *rays6, int n, float *out_t, int *out_idx)
{
    Scene *scene = (Scene *)s;
#pragma omp parallel for schedule(dynamic, 1024)
    for (int i = 0; i < n; i++) {
        Ray ray = {{rays6[6 * i], rays6[6 * i + 1], rays6[6 * i + 2]}, {rays6[6 * i + 3], rays6[6 * i + 4], rays6[6 * i + 5]}};
        float t = 1e30;
        int idx = -1;
        scene->bvh_closest_hit_distance(ray, t, idx);
        out_t[i] = t;
        out_idx[i] = idx;
    }
}

// The same on the GPU, running the reference's device code ("the reference kernel"). Returns 0 or a cudaError.
int ref_closest_hit_gpu(void *s, const float *rays6, int n, float *out_t, int *out_idx)
{
    Scene *scene = (Scene *)s;
    Scene dev = *scene;
    float *d_rays = nullptr, *d_t = nullptr;
    int *d_idx = nullptr;
    cudaError_t e;
#define SHIM_TRY(x) do { e = (x); if (e != cudaSuccess) return (int)e; } while (0)
    SHIM_TRY(cudaMalloc(&dev.triangles, sizeof(Triangle) * (scene->triangle_count + 1)));
    SHIM_TRY(cudaMalloc(&dev.bvh, sizeof(BvhNode) * scene->bvh_node_count));
    SHIM_TRY(cudaMemcpy(dev.triangles, scene->triangles, sizeof(Triangle) * scene->triangle_count, cudaMemcpyHostToDevice));
    SHIM_TRY(cudaMemcpy(dev.bvh, scene->bvh, sizeof(BvhNode) * scene->bvh_node_count, cudaMemcpyHostToDevice));
    SHIM_TRY(cudaMalloc(&d_rays, sizeof(float) * 6 * n));
    SHIM_TRY(cudaMalloc(&d_t, sizeof(float) * n));
    SHIM_TRY(cudaMalloc(&d_idx, sizeof(int) * n));
    SHIM_TRY(cudaMemcpy(d_rays, rays6, sizeof(float) * 6 * n, cudaMemcpyHostToDevice));
    shim_closest_hit_kernel<<<(n + 127) / 128, 128>>>(dev, d_rays, n, 0, d_t, d_idx);
    SHIM_TRY(cudaGetLastError());
    SHIM_TRY(cudaDeviceSynchronize());
    SHIM_TRY(cudaMemcpy(out_t, d_t, sizeof(float) * n, cudaMemcpyDeviceToHost));
    SHIM_TRY(cudaMemcpy(out_idx, d_idx, sizeof(int) * n, cudaMemcpyDeviceToHost));
    cudaFree(dev.triangles); cudaFree(dev.bvh); cudaFree(d_rays); cudaFree(d_t); cudaFree(d_idx);
#undef SHIM_TRY
    return 0;
}

// ray_data: n x 12 floats in RayData order (origin, direction, transmitted, collected).
void ref_generate(void *s, float *ray_data, int rays_per_pixel, int first, int n, const unsigned *seeds_unused, int seed)
{
    Scene *scene = (Scene *)s;
    (void)seeds_unused;
    // generate_initial_rays seeds itself from (ray_index, seed): scene.cu:81
    for (int i = 0; i < n; i++) {
        RayData tmp[1];
        // write into a 1-element window addressed by ray_index
        scene->generate_initial_rays(tmp - (first + i), nullptr, nullptr, rays_per_pixel, first + i, seed);
        memcpy(ray_data + 12 * i, tmp, sizeof(RayData));
    }
}

// One process_ray step per record, rng = xor_srand(seeds[i]).
void ref_process(void *s, float *ray_data, const unsigned *seeds, int n)
{
    Scene *scene = (Scene *)s;
#pragma omp parallel for schedule(dynamic, 1024)
    for (int i = 0; i < n; i++) {
        RayData rd;
        memcpy(&rd, ray_data + 12 * i, sizeof(RayData));
        xor_random rng;
        xor_srand(&rng, seeds[i]);
        scene->process_ray(&rd, nullptr, rng);
        memcpy(ray_data + 12 * i, &rd, sizeof(RayData));
    }
}

// Whole pass loop with the reference's functions. seed_mode 0 = cpu_raytrace's seeds
// (raytracing.cu:142-148), 1 = gpu_raytrace's seeds with slot == ray index, i.e. the `no_sort`
// GPU stream (raytracing.cu:89,229,236). out_rgb: width*height*3 un-normalised sums.
// Returns the number of live process_ray calls (ray-bounces) in *out_live if non-null.
void ref_render(void *s, int seed_mode, float *out_rgb, long long *out_live)
{
    Scene *scene = (Scene *)s;
    const int per_pass = 20;  // MAX_RAYS_PER_PIXEL_PER_PASS, raytracing.cu:109
    int pixel_count = scene->width * scene->height;
    RayData *ray_data = new RayData[(size_t)pixel_count * per_pass];
    memset(out_rgb, 0, sizeof(float) * 3 * pixel_count);
    long long live = 0;
    int remaining_rays = scene->ray_count;
    while (remaining_rays) {
        int rays_to_cast = remaining_rays < per_pass ? remaining_rays : per_pass;
        remaining_rays -= rays_to_cast;
        int total_rays = rays_to_cast * pixel_count;
#pragma omp parallel for schedule(dynamic, 1000)
        for (int i = 0; i < total_rays; i++)
            scene->generate_initial_rays(ray_data, nullptr, nullptr, rays_to_cast, i, remaining_rays);
        for (int b = 0; b < scene->bounces; b++) {
#pragma omp parallel for schedule(dynamic, 1000) reduction(+ : live)
            for (int i = 0; i < total_rays; i++) {
                const Vec3 &tc = ray_data[i].transmitted_color;
                if (!(tc.x == 0 && tc.y == 0 && tc.z == 0)) live++;
                xor_random rng;
                if (seed_mode == 0)
                    xor_srand(&rng, 1905678123 * i + 345903 * (remaining_rays * per_pass + i));
                else
                    xor_srand(&rng, gpu_proc_seed(i, remaining_rays * per_pass + b));
                scene->process_ray(&ray_data[i], nullptr, rng);
            }
        }
        for (int i = 0; i < total_rays; i++) {
            out_rgb[3 * (i / rays_to_cast) + 0] += ray_data[i].collected_color.x;
            out_rgb[3 * (i / rays_to_cast) + 1] += ray_data[i].collected_color.y;
            out_rgb[3 * (i / rays_to_cast) + 2] += ray_data[i].collected_color.z;
        }
    }
    delete[] ray_data;
    if (out_live) *out_live = live;
}

unsigned ref_gpu_gen_seed(int ray_index, int seed) { return gpu_gen_seed(ray_index, seed); }
unsigned ref_gpu_proc_seed(int ray_index, int seed) { return gpu_proc_seed(ray_index, seed); }

// The sort call the reference makes between bounces (raytracing.cu:194-195, 240-243): cub::DeviceRadixSort::SortPairs over
// all 32 key bits of `n` (key, index) pairs. Returns the sorted pairs and the CUDA-event time of the best of `reps` calls
// in milliseconds (device buffers, copies excluded). Used as the timing baseline of the product's own radix sort.
int ref_cub_sort_pairs(unsigned *keys, unsigned *values, int n, int reps, float *best_ms)
{
    unsigned *k[2], *v[2];
    for (int i = 0; i < 2; i++) {
        if (cudaMalloc(&k[i], (size_t)n * 4) != cudaSuccess || cudaMalloc(&v[i], (size_t)n * 4) != cudaSuccess) return -1;
    }
    size_t temp_size = 0;
    void *temp = nullptr;
    cub::DeviceRadixSort::SortPairs(nullptr, temp_size, k[0], k[1], v[0], v[1], n);
    if (cudaMalloc(&temp, temp_size) != cudaSuccess) return -1;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    float best = 1e30f;
    for (int r = 0; r < reps; r++) {
        cudaMemcpy(k[0], keys, (size_t)n * 4, cudaMemcpyHostToDevice);
        cudaMemcpy(v[0], values, (size_t)n * 4, cudaMemcpyHostToDevice);
        cudaEventRecord(e0);
        cub::DeviceRadixSort::SortPairs(temp, temp_size, k[0], k[1], v[0], v[1], n);
        cudaEventRecord(e1);
        if (cudaEventSynchronize(e1) != cudaSuccess) return -2;
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    cudaMemcpy(keys, k[1], (size_t)n * 4, cudaMemcpyDeviceToHost);
    cudaMemcpy(values, v[1], (size_t)n * 4, cudaMemcpyDeviceToHost);
    for (int i = 0; i < 2; i++) { cudaFree(k[i]); cudaFree(v[i]); }
    cudaFree(temp);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    if (best_ms) *best_ms = best;
    return cudaGetLastError() == cudaSuccess ? 0 : -3;
}

}  // extern "C"
