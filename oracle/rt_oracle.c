/* oracle/rt_oracle.c — TEST INFRASTRUCTURE ONLY. See rt_oracle.h for scope and parity status.
 *
 * CPU restatement of the reference's per-ray arithmetic, written from the behaviour described in
 * SURVEY.md §8a and the cited reference lines. Operation ORDER and the float/double mix follow
 * the reference expression by expression so the result is bit-identical to the reference's host
 * build with precise math (tests/test_oracle_golden.py against tests/golden/ref_vectors.npz). Build: -ffp-contract=off, no fast-math.
 */
#include "rt_oracle.h"

#include <limits.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

typedef struct { float x, y, z; } v3;

/* ---- math.cuh:66-141 ---------------------------------------------------------------------- */
static inline v3 V(float x, float y, float z) { v3 r = {x, y, z}; return r; }
static inline v3 ld3(const float *p) { return V(p[0], p[1], p[2]); }
static inline void st3(float *p, v3 a) { p[0] = a.x; p[1] = a.y; p[2] = a.z; }
static inline float dot3(v3 a, v3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }                /* math.cuh:66 */
static inline v3 cross3(v3 a, v3 b) { return V(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); } /* :71 */
static inline float magsq3(v3 a) { return a.x * a.x + a.y * a.y + a.z * a.z; }                    /* :80 */
static inline v3 add3(v3 a, v3 b) { return V(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline v3 sub3(v3 a, v3 b) { return V(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline v3 mul3(v3 a, v3 b) { return V(a.x * b.x, a.y * b.y, a.z * b.z); }
static inline v3 scale3(float s, v3 a) { return V(s * a.x, s * a.y, s * a.z); }                   /* :105 */
static inline v3 neg3(v3 a) { return V(-a.x, -a.y, -a.z); }
static inline v3 normalise3(v3 a) { return scale3(1.0f / sqrtf(magsq3(a)), a); }                  /* :110 */
static inline float clamp01f(float x) { return fmaxf(fminf(x, 1.0f), 0.0f); }                     /* :115-123 */

/* ---- random.cuh --------------------------------------------------------------------------- */
typedef struct { uint64_t state, inc; } pcg;

static inline uint32_t pcg_next(pcg *rng) /* random.cuh:13-23, PCG-XSH-RR 64/32 */
{
    uint64_t old_state = rng->state;
    rng->state = old_state * 6364136223846793005ULL + (rng->inc | 1);
    uint32_t xor_shifted = (uint32_t)(((old_state >> 18) ^ old_state) >> 27);
    uint32_t rot = (uint32_t)(old_state >> 59);
    return (xor_shifted >> rot) | (xor_shifted << ((-rot) & 31));
}

static inline void pcg_seed(pcg *rng, uint32_t seed) /* random.cuh:25-30 */
{
    rng->state = seed * 6839056345687307ULL;
    rng->inc = 820957824423429ULL;
    pcg_next(rng);
}

/* UINT_MAX converts to 2^32 in float, so these scale by exactly 2^-32 / 2^-31 (Q8: can reach 1.0f) */
static inline float rnd01(pcg *rng) { return pcg_next(rng) * (1.0f / UINT_MAX); }                 /* :32 */
static inline float rnd02(pcg *rng) { return pcg_next(rng) * (2.0f / UINT_MAX); }                 /* :37 */
static inline float rnd_radians(pcg *rng) { return (float)(pcg_next(rng) * (M_PI * 2 / UINT_MAX)); } /* :42, double (Q6) */

static inline v3 rnd_on_sphere(pcg *rng) /* random.cuh:63-75 */
{
    float r1 = rnd_radians(rng);
    float r2 = rnd02(rng);
    float x = sqrtf(r2 * (2 - r2));
    return V(cosf(r1) * x, sinf(r1) * x, 1 - r2);
}

void oracle_xor_sequence(uint32_t seed, int n, uint32_t *out)
{
    pcg rng;
    pcg_seed(&rng, seed);
    for (int i = 0; i < n; i++) out[i] = pcg_next(&rng);
}

void oracle_random_floats(uint32_t seed, float *out)
{
    pcg rng;
    pcg_seed(&rng, seed); out[0] = rnd01(&rng);
    pcg_seed(&rng, seed); out[1] = rnd02(&rng);
    pcg_seed(&rng, seed); out[2] = rnd_radians(&rng);
    pcg_seed(&rng, seed);
    v3 v = rnd_on_sphere(&rng);
    out[3] = v.x; out[4] = v.y; out[5] = v.z;
}

/* 64-bit products truncated to 32 bits == 32-bit wrap-around arithmetic */
uint32_t oracle_gpu_gen_seed(int32_t ray_index, int32_t seed)
{
    return (uint32_t)((int64_t)ray_index * 298592570346LL + (int64_t)(709579 * seed)); /* scene.cu:81 */
}
uint32_t oracle_gpu_proc_seed(int32_t ray_index, int32_t seed)
{
    return (uint32_t)((int64_t)ray_index * 4137874753LL + (int64_t)(279220567 * seed)); /* raytracing.cu:89 */
}

/* ---- scene.cu:284-318: pbrt equal-area sphere -> square --------------------------------------- */
static inline void project_sphere_to_square(v3 d, float *out_u, float *out_v)
{
    float x = fabsf(d.x), y = fabsf(d.y), z = fabsf(d.z);
    float r = sqrtf(1 - fminf(z, 1.0f));
    float a = fmaxf(x, y);
    float b = fminf(x, y);
    b = a == 0 ? 0 : b / a;
    float phi = (float)((2 / M_PI) * atanf(b));
    if (x < y) phi = 1 - phi;
    float v = phi * r;
    float u = r - v;
    if (d.z < 0) {
        float old_v = v;
        v = 1 - u;
        u = 1 - old_v;
    }
    u = copysignf(u, d.x);
    v = copysignf(v, d.y);
    *out_u = (u + 1) * 0.5f;
    *out_v = (v + 1) * 0.5f;
}

void oracle_project(const float *dirs3, int n, float *out_uv)
{
    for (int i = 0; i < n; i++) project_sphere_to_square(ld3(dirs3 + 3 * i), &out_uv[2 * i], &out_uv[2 * i + 1]);
}

/* ---- scene.cu:109-132: boundary-safe slab test, tmin starts at 0 -------------------------------- */
static inline int ray_aabb(const float *mn, const float *mx, v3 o, v3 n_inv, float *tmin_out, float tmax)
{
    float tmin = 0.0f;
    float t1 = (mn[0] - o.x) * n_inv.x;
    float t2 = (mx[0] - o.x) * n_inv.x;
    tmin = fminf(fmaxf(t1, tmin), fmaxf(t2, tmin));
    tmax = fmaxf(fminf(t1, tmax), fminf(t2, tmax));
    t1 = (mn[1] - o.y) * n_inv.y;
    t2 = (mx[1] - o.y) * n_inv.y;
    tmin = fminf(fmaxf(t1, tmin), fmaxf(t2, tmin));
    tmax = fmaxf(fminf(t1, tmax), fminf(t2, tmax));
    t1 = (mn[2] - o.z) * n_inv.z;
    t2 = (mx[2] - o.z) * n_inv.z;
    tmin = fminf(fmaxf(t1, tmin), fmaxf(t2, tmin));
    tmax = fmaxf(fminf(t1, tmax), fminf(t2, tmax));
    *tmin_out = tmin;
    return tmin <= tmax;
}

/* DIAGNOSTIC ONLY (near_first == 2 in the closest-hit entry points; tools/slow_ray_study.py): the slab test in the form the CUDA
 * core evaluates it (traverse.cuh: box as centre + half-extent rounded up, t_centre = fma(c, 1/d, -(o * 1/d)), t_near / t_far =
 * fma(-+h, |1/d|, t_centre), NaN = no constraint) — to count the records THAT form visits for a given ray. Not pinned to any
 * reference vector: it restates the product, not the reference. */
static inline int ray_aabb_device_form(const float *mn, const float *mx, v3 o, v3 n_inv, float *tmin_out, float tmax)
{
    float c[3], h[3];
    for (int k = 0; k < 3; k++) {
        c[k] = 0.5f * (mn[k] + mx[k]);
        float hk = fmaxf(mx[k] - c[k], c[k] - mn[k]);
        h[k] = nextafterf(hk, INFINITY);
    }
    const float sox = -(o.x * n_inv.x), soy = -(o.y * n_inv.y), soz = -(o.z * n_inv.z);
    const float ax = fabsf(n_inv.x), ay = fabsf(n_inv.y), az = fabsf(n_inv.z);
    const float tx = fmaf(c[0], n_inv.x, sox), ty = fmaf(c[1], n_inv.y, soy), tz = fmaf(c[2], n_inv.z, soz);
    const float tn = fmaxf(fmaxf(fmaf(-h[0], ax, tx), fmaf(-h[1], ay, ty)), fmaxf(fmaf(-h[2], az, tz), 0.0f));
    const float tf = fminf(fminf(fmaf(h[0], ax, tx), fmaf(h[1], ay, ty)), fminf(fmaf(h[2], az, tz), tmax));
    *tmin_out = tn;
    return tn <= tf;
}

int oracle_ray_aabb(const float *aabb6, const float *ray6, float tmax, float *tmin_out)
{
    v3 d = ld3(ray6 + 3);
    v3 n_inv = V(1 / d.x, 1 / d.y, 1 / d.z);
    return ray_aabb(aabb6, aabb6 + 3, ld3(ray6), n_inv, tmin_out, tmax);
}

/* ---- scene.cu:134-241: stack DFS closest hit --------------------------------------------------- */
#define ORACLE_STACK 31 /* MAX_BVH_DEPTH + 1, scene.cu:10,138 */

static void bvh_closest_hit(const oracle_scene *s, v3 o, v3 d, int near_first, float *closest, int32_t *closest_idx,
                            oracle_counters *c)
{
    v3 n_inv = V(1 / d.x, 1 / d.y, 1 / d.z);
    uint32_t index_stack[ORACLE_STACK];
    float distance_stack[ORACLE_STACK];
    int stack_count = 1;
    index_stack[0] = 0;
    distance_stack[0] = 0;

    while (stack_count) {
        stack_count--;
        float distance = distance_stack[stack_count];
        if (distance >= *closest) continue;

        const oracle_node *node = &s->bvh[index_stack[stack_count]];
        if (node->child2 <= node->child1) { /* is_leaf, scene.cu:859-862: [child2, child1) */
            if (c) c->leaf_visits++;
            for (int i = node->child2; i < node->child1; i++) {
                const float *tri = s->triangles + 12 * (size_t)i;
                v3 p1 = ld3(tri), e1 = ld3(tri + 3), e2 = ld3(tri + 6);
                if (c) c->triangle_tests++;
                /* Moller-Trumbore as written at scene.cu:164-194 */
                v3 h = cross3(d, e2);
                float a = dot3(h, e1);
                if (a == 0) continue;
                float inv_a = 1 / a;
                v3 offset = sub3(o, p1);
                float u = dot3(offset, h) * inv_a;
                if (u < 0 || u > 1) continue;
                v3 q = cross3(offset, e1);
                float v = dot3(d, q) * inv_a;
                if (v < 0 || u + v > 1) continue;
                float t = dot3(e2, q) * inv_a;
                if ((double)t < 0.005 || t >= *closest) continue; /* double compare, Q6 */
                *closest = t;
                *closest_idx = s->sphere_count + i;
            }
        } else {
            if (c) c->internal_visits++;
            float d1, d2;
            const oracle_node *n1 = &s->bvh[node->child1], *n2 = &s->bvh[node->child2];
            int hit1, hit2;
            if (near_first == 2) {
                hit1 = ray_aabb_device_form(n1->min_bound, n1->max_bound, o, n_inv, &d1, *closest);
                hit2 = ray_aabb_device_form(n2->min_bound, n2->max_bound, o, n_inv, &d2, *closest);
            } else {
                hit1 = ray_aabb(n1->min_bound, n1->max_bound, o, n_inv, &d1, *closest);
                hit2 = ray_aabb(n2->min_bound, n2->max_bound, o, n_inv, &d2, *closest);
            }
            if (hit1 && hit2) {
                /* reference: nearer child pushed first => popped last (Q5) */
                int first_is_1 = d1 < d2;
                if (near_first) first_is_1 = !first_is_1;
                if (first_is_1) {
                    index_stack[stack_count] = node->child1; distance_stack[stack_count++] = d1;
                    index_stack[stack_count] = node->child2; distance_stack[stack_count++] = d2;
                } else {
                    index_stack[stack_count] = node->child2; distance_stack[stack_count++] = d2;
                    index_stack[stack_count] = node->child1; distance_stack[stack_count++] = d1;
                }
            } else if (hit1) {
                index_stack[stack_count] = node->child1; distance_stack[stack_count++] = d1;
            } else if (hit2) {
                index_stack[stack_count] = node->child2; distance_stack[stack_count++] = d2;
            }
        }
    }
}

/* ---- scene.cu:338-372: the sphere loop of process_ray; runs BEFORE the BVH and seeds closest / closest_idx ---------- */
static inline void sphere_closest_hit(const oracle_scene *s, v3 origin, v3 dir, float *closest_io, int32_t *idx_io)
{
    float closest = *closest_io;
    int32_t closest_idx = *idx_io;
    for (int i = 0; i < s->sphere_count; i++) {
        v3 center = ld3(s->spheres + 4 * i);
        float radius = s->spheres[4 * i + 3];
        v3 offset = sub3(center, origin);
        float minus_half_b = dot3(offset, dir);
        float quarter_c = magsq3(offset) - radius * radius;
        float quarter_disc = minus_half_b * minus_half_b - quarter_c;
        if (quarter_disc < 0) continue;
        float half_root = sqrtf(quarter_disc);
        float t = minus_half_b - half_root;
        if (t < closest && (double)t >= 0.005) { closest = t; closest_idx = i; continue; }
        t = minus_half_b + half_root;
        if (t < closest && (double)t >= 0.005) { closest = t; closest_idx = i; continue; }
    }
    *closest_io = closest;
    *idx_io = closest_idx;
}

void oracle_closest_hit(const oracle_scene *s, const float *rays6, int n, int near_first, float *out_t, int32_t *out_idx,
                        oracle_counters *counters)
{
    oracle_closest_hit_ex(s, rays6, n, near_first, 0, out_t, out_idx, counters);
}

/* The closest-hit query of process_ray as a whole: sphere loop (scene.cu:338-372) when with_spheres, then the BVH
 * (scene.cu:134-241) continuing from the sphere result. */
void oracle_closest_hit_ex(const oracle_scene *s, const float *rays6, int n, int near_first, int with_spheres, float *out_t,
                           int32_t *out_idx, oracle_counters *counters)
{
    int64_t iv = 0, lv = 0, tt = 0;
#pragma omp parallel for schedule(dynamic, 1024) reduction(+ : iv, lv, tt)
    for (int i = 0; i < n; i++) {
        oracle_counters c = {0, 0, 0, 0};
        float t = 1e30f;
        int32_t idx = -1;
        if (with_spheres) sphere_closest_hit(s, ld3(rays6 + 6 * (size_t)i), ld3(rays6 + 6 * (size_t)i + 3), &t, &idx);
        bvh_closest_hit(s, ld3(rays6 + 6 * (size_t)i), ld3(rays6 + 6 * (size_t)i + 3), near_first, &t, &idx, counters ? &c : NULL);
        out_t[i] = t;
        out_idx[i] = idx;
        iv += c.internal_visits; lv += c.leaf_visits; tt += c.triangle_tests;
    }
    if (counters) {
        counters->internal_visits += iv; counters->leaf_visits += lv; counters->triangle_tests += tt;
    }
}

/* ---- scene.cu:78-105 ---------------------------------------------------------------------------- */
static inline int generate_one(const oracle_scene *s, float *rd, int rays_per_pixel, int ray_index, int seed)
{
    pcg rng;
    pcg_seed(&rng, oracle_gpu_gen_seed(ray_index, seed));
    int framebuffer_index = ray_index / rays_per_pixel;
    int x = framebuffer_index % s->width;
    int y = framebuffer_index / s->width;
    if (y >= s->height) return 0;
    float x_clamped = (x + rnd01(&rng)) * s->inv_width;
    float y_clamped = (y + rnd01(&rng)) * s->inv_height;
    v3 dir = normalise3(sub3(add3(ld3(s->near_plane_top_left), scale3(x_clamped, ld3(s->scaled_right))),
                             scale3(y_clamped, ld3(s->scaled_up))));
    st3(rd, ld3(s->camera_position));
    st3(rd + 3, dir);
    st3(rd + 6, V(1, 1, 1));
    st3(rd + 9, V(0, 0, 0));
    return 1;
}

void oracle_generate(const oracle_scene *s, float *ray_data, int rays_per_pixel, int first, int n, int seed)
{
#pragma omp parallel for schedule(static)
    for (int i = 0; i < n; i++) generate_one(s, ray_data + 12 * (size_t)i, rays_per_pixel, first + i, seed);
}

/* ---- scene.cu:320-487: one bounce of one ray (host branch) --------------------------------------- */
/* returns 1 if the ray was live (not an early-out) */
static int process_one(const oracle_scene *s, float *rd, pcg rng, oracle_counters *c)
{
    v3 origin = ld3(rd), dir = ld3(rd + 3), transmitted = ld3(rd + 6), collected = ld3(rd + 9);
    if (transmitted.x == 0 && transmitted.y == 0 && transmitted.z == 0) return 0; /* scene.cu:326 */

    float closest = 1e30f;
    int32_t closest_idx = -1;

    sphere_closest_hit(s, origin, dir, &closest, &closest_idx);

    bvh_closest_hit(s, origin, dir, 0, &closest, &closest_idx, c);

    if (closest_idx == -1) { /* miss: scene.cu:376-395 */
        float dir_x = (float)(dir.x * -0.386527 + dir.z * 0.922278);
        float dir_y = (float)(dir.x * -0.922278 + dir.z * -0.386527);
        float dir_z = dir.y;
        float u, v;
        project_sphere_to_square(V(dir_x, dir_y, dir_z), &u, &v);
        int texel_x = (int)(clamp01f(u) * (s->environment_map_width - 1) + 0.5);
        int texel_y = (int)(clamp01f(v) * (s->environment_map_height - 1) + 0.5);
        v3 sky = ld3(s->environment_map + 3 * ((size_t)texel_y * s->environment_map_height + texel_x)); /* Q7 */
        collected = add3(collected, mul3(sky, transmitted));
        transmitted = V(0, 0, 0);
    } else { /* hit: scene.cu:396-477 */
        v3 hit_point = add3(origin, scale3(closest, dir));
        origin = hit_point;
        v3 normal;
        if (closest_idx < s->sphere_count) {
            v3 center = ld3(s->spheres + 4 * closest_idx);
            float radius = s->spheres[4 * closest_idx + 3];
            normal = scale3(1 / radius, sub3(hit_point, center));
        } else {
            normal = ld3(s->triangles + 12 * (size_t)(closest_idx - s->sphere_count) + 9);
        }
        const float *m = s->materials + 12 * (size_t)s->material_indices[closest_idx];
        v3 diffuse = ld3(m), specular = ld3(m + 4), emitted = ld3(m + 8);
        float metallicity = m[3], roughness = m[7], index_of_refraction = m[11];

        collected = add3(collected, mul3(emitted, transmitted));
        int front_face = dot3(normal, dir) < 0;
        if (!front_face) normal = neg3(normal);

        v3 rough_normal = normalise3(add3(normal, scale3(roughness, rnd_on_sphere(&rng))));
        float cos_theta = dot3(rough_normal, dir);

        if (index_of_refraction == 0) {
            if (rnd01(&rng) <= metallicity) {
                transmitted = mul3(transmitted, specular);
                dir = sub3(dir, scale3(2 * cos_theta, rough_normal));
            } else {
                transmitted = mul3(transmitted, diffuse);
                dir = normalise3(add3(normal, rnd_on_sphere(&rng)));
            }
        } else {
            float ior = index_of_refraction;
            float inv_ior = 1 / ior;
            if (front_face) { float temp = inv_ior; inv_ior = ior; ior = temp; }
            float sin_theta_squared = 1 - cos_theta * cos_theta;
            float r0 = (1 - ior) / (1 + ior);
            r0 *= r0;
            float cosine = 1 + cos_theta;
            float reflectance = r0 + (1 - r0) * cosine * cosine * cosine * cosine * cosine;
            if (sin_theta_squared > inv_ior * inv_ior || rnd01(&rng) < reflectance) {
                transmitted = mul3(transmitted, specular);
                dir = sub3(dir, scale3(2 * cos_theta, rough_normal));
            } else {
                transmitted = mul3(transmitted, diffuse);
                v3 perp = scale3(ior, sub3(dir, scale3(cos_theta, rough_normal)));
                v3 parallel = scale3(-sqrtf(1 - magsq3(perp)), rough_normal);
                dir = normalise3(add3(parallel, perp));
            }
        }
    }
    st3(rd, origin); st3(rd + 3, dir); st3(rd + 6, transmitted); st3(rd + 9, collected);
    return 1;
}

void oracle_process(const oracle_scene *s, float *ray_data, const uint32_t *seeds, int n)
{
#pragma omp parallel for schedule(dynamic, 1024)
    for (int i = 0; i < n; i++) {
        pcg rng;
        pcg_seed(&rng, seeds[i]);
        process_one(s, ray_data + 12 * (size_t)i, rng, NULL);
    }
}

/* ---- pass loop: raytracing.cu:122-163 (cpu) / :222-254 (gpu, no_sort stream) ----------------------- */
#define ORACLE_PER_PASS 20 /* MAX_RAYS_PER_PIXEL_PER_PASS, raytracing.cu:109 */

void oracle_render(const oracle_scene *s, int seed_mode, int pass_first, int pass_count, float *out_rgb, oracle_counters *counters)
{
    oracle_render_select(s, seed_mode, pass_first, pass_count, 0, -1, NULL, out_rgb, counters);
}

/* The pass loop restricted to a SUBSET of the frame's rays; every selected ray keeps the seeds of its own pass and of its
 * index inside that pass (raytracing.cu:89,229,236, scene.cu:81), so the result is exactly that subset's contribution to
 * the whole frame. Selection = passes [pass_first, pass_first + pass_count) (pass_count < 0: all), AND rays
 * [ray_first, ray_first + ray_count) of the frame's ray sequence (pass 0's rays, then pass 1's, ...; ray_count < 0: all),
 * AND pixels inside window = {x0, y0, w, h} (NULL: the whole image). out_rgb is the full-size image. */
static void render_subset(const oracle_scene *s, int seed_mode, int pass_first, int pass_count, int64_t ray_first, int64_t ray_count,
                          const int32_t *window, int sample_first, int sample_count, float *out_rgb, oracle_counters *counters);

void oracle_render_select(const oracle_scene *s, int seed_mode, int pass_first, int pass_count, int64_t ray_first, int64_t ray_count,
                          const int32_t *window, float *out_rgb, oracle_counters *counters)
{
    render_subset(s, seed_mode, pass_first, pass_count, ray_first, ray_count, window, 0, -1, out_rgb, counters);
}

/* Samples [sample_first, sample_first + sample_count) of EVERY pixel (sample i of a pixel = ray i % 20 of that pixel in pass
 * i / 20): the subset one GPU renders in the multi-GPU split. */
void oracle_render_samples(const oracle_scene *s, int seed_mode, int sample_first, int sample_count, float *out_rgb, oracle_counters *counters)
{
    render_subset(s, seed_mode, 0, -1, 0, -1, NULL, sample_first, sample_count, out_rgb, counters);
}

static void render_subset(const oracle_scene *s, int seed_mode, int pass_first, int pass_count, int64_t ray_first, int64_t ray_count,
                          const int32_t *window, int sample_first, int sample_count, float *out_rgb, oracle_counters *counters)
{
    int pixel_count = s->width * s->height;
    memset(out_rgb, 0, sizeof(float) * 3 * (size_t)pixel_count);
    int64_t live = 0, iv = 0, lv = 0, tt = 0;
    int remaining_rays = s->ray_count;
    int pass = 0;
    int64_t sequence_start = 0;
    int32_t *ids = NULL;
    float *ray_data = NULL;
    size_t cap = 0;
    while (remaining_rays) {
        int rays_to_cast = remaining_rays < ORACLE_PER_PASS ? remaining_rays : ORACLE_PER_PASS;
        remaining_rays -= rays_to_cast;
        int this_pass = pass++;
        int64_t total_rays = (int64_t)rays_to_cast * pixel_count;
        int64_t pass_start = sequence_start;
        sequence_start += total_rays;
        if (this_pass < pass_first || (pass_count >= 0 && this_pass >= pass_first + pass_count)) continue;
        int64_t lo = 0, hi = total_rays;
        if (ray_count >= 0) {
            lo = ray_first - pass_start > 0 ? ray_first - pass_start : 0;
            hi = ray_first + ray_count - pass_start < total_rays ? ray_first + ray_count - pass_start : total_rays;
        }
        if (hi <= lo) continue;
        /* samples of this pass inside the sample range: k in [k0, k1) */
        int k0 = 0, k1 = rays_to_cast;
        if (sample_count >= 0) {
            int pass_s0 = this_pass * ORACLE_PER_PASS;
            k0 = sample_first - pass_s0 > 0 ? sample_first - pass_s0 : 0;
            k1 = sample_first + sample_count - pass_s0 < rays_to_cast ? sample_first + sample_count - pass_s0 : rays_to_cast;
            if (k1 <= k0) continue;
        }
        const int32_t whole_image[4] = {0, 0, s->width, s->height};
        const int32_t *win = window;  /* per pass: a partial sample range is enumerated pixel by pixel over the whole image */
        if (!win && (k0 > 0 || k1 < rays_to_cast)) win = whole_image;
        /* the selected ray indices of this pass */
        size_t n_sel = 0;
        if (win) {
            size_t bound = (size_t)win[2] * win[3] * rays_to_cast;
            if (bound > cap) { cap = bound; ids = (int32_t *)realloc(ids, sizeof(int32_t) * cap); ray_data = (float *)realloc(ray_data, sizeof(float) * 12 * cap); }
            for (int y = win[1]; y < win[1] + win[3]; y++)
                for (int x = win[0]; x < win[0] + win[2]; x++)
                    for (int k = k0; k < k1; k++) {
                        int64_t i = ((int64_t)y * s->width + x) * rays_to_cast + k;
                        if (i >= lo && i < hi) ids[n_sel++] = (int32_t)i;
                    }
        } else {
            n_sel = (size_t)(hi - lo);
            if (n_sel > cap) { cap = n_sel; ids = (int32_t *)realloc(ids, sizeof(int32_t) * cap); ray_data = (float *)realloc(ray_data, sizeof(float) * 12 * cap); }
            for (size_t j = 0; j < n_sel; j++) ids[j] = (int32_t)(lo + (int64_t)j);
        }
#pragma omp parallel for schedule(dynamic, 1000)
        for (size_t j = 0; j < n_sel; j++) generate_one(s, ray_data + 12 * j, rays_to_cast, ids[j], remaining_rays);
        for (int b = 0; b < s->bounces; b++) {
#pragma omp parallel for schedule(dynamic, 1000) reduction(+ : live, iv, lv, tt)
            for (size_t j = 0; j < n_sel; j++) {
                int i = ids[j];
                pcg rng;
                if (seed_mode == 0) /* raytracing.cu:148, shadowed loop variable => no bounce term (Q3) */
                    pcg_seed(&rng, (uint32_t)(1905678123 * i + 345903 * (remaining_rays * ORACLE_PER_PASS + i)));
                else
                    pcg_seed(&rng, oracle_gpu_proc_seed(i, remaining_rays * ORACLE_PER_PASS + b));
                oracle_counters c = {0, 0, 0, 0};
                live += process_one(s, ray_data + 12 * j, rng, counters ? &c : NULL);
                iv += c.internal_visits; lv += c.leaf_visits; tt += c.triangle_tests;
            }
        }
        for (size_t j = 0; j < n_sel; j++) { /* raytracing.cu:114-120, serial order */
            float *px = out_rgb + 3 * (size_t)(ids[j] / rays_to_cast);
            const float *col = ray_data + 12 * j + 9;
            px[0] += col[0]; px[1] += col[1]; px[2] += col[2];
        }
    }
    free(ray_data);
    free(ids);
    if (counters) {
        counters->live_ray_bounces += live; counters->internal_visits += iv;
        counters->leaf_visits += lv; counters->triangle_tests += tt;
    }
}

/* ---- scene.cu:44-60,480-485: the reference's (degenerate, Q1) reorder key ------------------------ */
static inline uint16_t sat_u16(double v) /* device float->u16 conversion saturates; NaN -> 0 */
{
    if (!(v > 0)) return 0;
    if (v >= 65535.0) return 65535;
    return (uint16_t)v;
}
static inline uint16_t interleave_5(uint16_t x)
{
    x = (uint16_t)((x | (x << 8)) & 0x100F);          /* 0b1000000001111 */
    x = (uint16_t)((x | (x << 4)) & 0x1000010100011); /* hex where binary was meant; low 16 bits = 0x0011 */
    x = (uint16_t)((x | (x << 2)) & 0x1249);          /* 0b1001001001001 */
    return x;
}
static inline uint16_t morton_code(v3 v)
{
    uint16_t x = sat_u16(v.x * 31.99), y = sat_u16(v.y * 31.99), z = sat_u16(v.z * 31.99);
    return (uint16_t)(interleave_5(x) | (interleave_5(y) << 1) | (interleave_5(z) << 2));
}
uint32_t oracle_sort_key(const oracle_scene *s, const float *origin3, const float *dir3, const float *transmitted3)
{
    if (transmitted3[0] == 0 && transmitted3[1] == 0 && transmitted3[2] == 0) return 0xFFFFFFFFu;
    v3 on = mul3(sub3(ld3(origin3), ld3(s->min_coord)), ld3(s->inv_dimensions));
    v3 dn = scale3(0.5f, add3(ld3(dir3), V(1, 1, 1)));
    return ((uint32_t)morton_code(on) << 16) | (uint32_t)morton_code(dn);
}

/* ---- raytracing.cu:21-74,360-392: bloom on the un-normalised sums --------------------------------- */
void oracle_bloom(float *rgb, int width, int height, int ray_count)
{
    size_t n = (size_t)width * height;
    float *bright = (float *)malloc(sizeof(float) * 3 * n);
    float *blur = (float *)malloc(sizeof(float) * 3 * n);
    float threshold = (float)(0.7 * ray_count); /* double product narrowed at the call, raytracing.cu:376 */
    const int radius = 5;
    for (size_t i = 0; i < n; i++) { /* high_pass_filter */
        v3 p = ld3(rgb + 3 * i);
        float brightness = dot3(p, V(0.2126f, 0.7152f, 0.0722f));
        st3(bright + 3 * i, brightness > threshold ? p : V(0, 0, 0));
    }
    for (int y = 0; y < height; y++) /* box_blur_horizontal */
        for (int x = 0; x < width; x++) {
            v3 sum = V(0, 0, 0);
            int count = 0;
            for (int dx = -radius; dx <= radius; dx++) {
                int nx = x + dx;
                if (nx >= 0 && nx < width) { sum = add3(sum, ld3(bright + 3 * ((size_t)y * width + nx))); count++; }
            }
            st3(blur + 3 * ((size_t)y * width + x), scale3(1.0f / count, sum));
        }
    for (int y = 0; y < height; y++) /* box_blur_vertical */
        for (int x = 0; x < width; x++) {
            v3 sum = V(0, 0, 0);
            int count = 0;
            for (int dy = -radius; dy <= radius; dy++) {
                int ny = y + dy;
                if (ny >= 0 && ny < height) { sum = add3(sum, ld3(blur + 3 * ((size_t)ny * width + x))); count++; }
            }
            st3(bright + 3 * ((size_t)y * width + x), scale3(1.0f / count, sum));
        }
    for (size_t i = 0; i < 3 * n; i++) rgb[i] += bright[i]; /* add_images */
    free(bright);
    free(blur);
}

/* ---- raytracing.cu:286-303 ----------------------------------------------------------------------- */
void oracle_tonemap(const float *rgb, int pixel_count, float exposure, int ray_count, uint8_t *out)
{
    float scale = exposure / ray_count;
    for (size_t i = 0; i < (size_t)pixel_count * 3; i++) {
        float c = scale * rgb[i];
        out[i] = (uint8_t)(sqrtf(c / (c + 1)) * 255.999f);
    }
}

int oracle_max_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
