/* oracle/rt_oracle.h — TEST INFRASTRUCTURE ONLY.
 *
 * Plain-C CPU restatement of the reference's render hot path. Only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it.
 * The product (librtcore.so) never links or calls anything in oracle/.
 *
 * Parity status: PINNED. tests/test_oracle_vs_ref.py checks every function here bit-for-bit against
 * the reference's own code compiled in place (oracle/_ref/libref_shim.so, built by oracle/Makefile
 * `ref`), and tests/golden/ holds vectors generated from that shim (tests/golden/make_golden.py) so
 * the pin also holds on machines without /root/reference.
 *
 * All array layouts are the reference's (scene.cuh): Sphere 4 floats, Triangle 12 floats in
 * render form (p1, p2-p1, p3-p1, normal), Material 12 floats, BvhNode 32 bytes
 * (min[3], max[3], int child1, int child2), RayData 12 floats (origin, direction, transmitted,
 * collected), environment map 3 floats per texel.
 */
#ifndef RT_ORACLE_H
#define RT_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct oracle_node {
    float min_bound[3];
    float max_bound[3];
    int32_t child1;
    int32_t child2;
} oracle_node; /* scene.cuh:82-100 */

typedef struct oracle_scene {
    const float *spheres;            int32_t sphere_count;
    const float *triangles;          int32_t triangle_count;
    const uint16_t *material_indices;
    const float *materials;          int32_t material_count;
    const oracle_node *bvh;          int32_t bvh_node_count;
    const float *environment_map;    int32_t environment_map_width, environment_map_height;
    int32_t width, height, ray_count, bounces;
    float exposure;
    float camera_position[3];
    float near_plane_top_left[3];
    float scaled_right[3];
    float scaled_up[3];
    float inv_width, inv_height;
    float min_coord[3];
    float inv_dimensions[3];
} oracle_scene; /* scene.cuh:102-166, the fields the hot path reads */

typedef struct oracle_counters {
    int64_t live_ray_bounces;   /* non-early-out process_ray calls */
    int64_t internal_visits;    /* internal nodes whose children were tested */
    int64_t leaf_visits;
    int64_t triangle_tests;
} oracle_counters;

/* random.cuh:13-30 */
void oracle_xor_sequence(uint32_t seed, int n, uint32_t *out);
/* random.cuh:32-75: out = random01, random02, random_radians, random_on_sphere.xyz (fresh seed each) */
void oracle_random_floats(uint32_t seed, float *out6);
/* scene.cu:284-318 */
void oracle_project(const float *dirs3, int n, float *out_uv2);
/* scene.cu:109-132 */
int oracle_ray_aabb(const float *aabb6, const float *ray6, float tmax, float *tmin_out);
/* raytracing.cu:89,229,236 / scene.cu:81 seeds */
uint32_t oracle_gpu_gen_seed(int32_t ray_index, int32_t seed);
uint32_t oracle_gpu_proc_seed(int32_t ray_index, int32_t seed);

/* scene.cu:134-241. near_first = 0 reproduces the reference's pop order (far child first, Q5);
 * near_first = 1 is the order the CUDA core uses. Starts from t = 1e30, idx = -1. counters may be NULL. */
void oracle_closest_hit(const oracle_scene *s, const float *rays6, int n, int near_first,
                        float *out_t, int32_t *out_idx, oracle_counters *counters);

/* scene.cu:78-105: rays first..first+n-1 of a pass, 12 floats each */
void oracle_generate(const oracle_scene *s, float *ray_data12, int rays_per_pixel, int first, int n, int seed);

/* scene.cu:320-487 (host branch): one step per record, rng = xor_srand(seeds[i]) */
void oracle_process(const oracle_scene *s, float *ray_data12, const uint32_t *seeds, int n);

/* raytracing.cu:122-163 (seed_mode 0) or the gpu_raytrace `no_sort` stream raytracing.cu:222-254
 * (seed_mode 1: slot == ray index). pass_first/pass_count select passes for multi-GPU style splits
 * (pass_count < 0 = all). out_rgb: 3*width*height un-normalised sums, overwritten. */
void oracle_render(const oracle_scene *s, int seed_mode, int pass_first, int pass_count,
                   float *out_rgb, oracle_counters *counters);

/* scene.cu:44-60,480-485 with the device's saturating float->u16 conversion */
uint32_t oracle_sort_key(const oracle_scene *s, const float *origin3, const float *dir3, const float *transmitted3);

/* raytracing.cu:21-74,360-392: in-place bloom on un-normalised sums */
void oracle_bloom(float *rgb, int width, int height, int ray_count);
/* raytracing.cu:286-303 */
void oracle_tonemap(const float *rgb, int pixel_count, float exposure, int ray_count, uint8_t *out_rgb8);

int oracle_max_threads(void);

#ifdef __cplusplus
}
#endif
#endif
