/* include/rt_scene.h — C ABI of the host scene layer (part of librtcore.so; host code: it works without a CUDA device, and with
 * one its only CUDA call is rt_host_pin on the arrays of a loaded scene, so that uploads are asynchronous DMA).
 *
 * Replaces, on the host side of the render path:
 *     void load_scene(Scene *scene, const char *filename, bool use_bvh)   /root/reference/scene.cuh:168, scene.cu:569-831
 *       load_ply  scene.cu:491-546        load_pfm  scene.cu:548-567
 *       Scene::precompute_camera_data  scene.cu:62-76
 *       Scene::generate_bvh / BvhNode::maybe_split  scene.cu:866-1036
 * and the PNG write at raytracing.cu:395.
 *
 * The scene-file grammar is the reference's. Differences, all hardening: a trailing '\r' is
 * stripped from every line, missing files and undefined materials return RT_ERR_IO / RT_ERR_PARSE with
 * a message instead of throwing.
 */
#ifndef RT_SCENE_H
#define RT_SCENE_H

#include "rt_core.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct rt_host_scene rt_host_scene;

/* Paths inside the scene file are relative to the current working directory (reference behaviour).
 * use_bvh = 0 reproduces the `no_bvh` flag: one root leaf holding every triangle (scene.cu:820, Q12).
 * use_bvh = 2 defers the build to the device: triangles stay in build form and file order, the descriptor carries
 * triangles_form = 1 and a single root node (its box is still computed, the scene bounds need it).
 * use_bvh = 3 builds the tree on the host WITH spatial splits (not the reference's tree: a triangle that straddles a split
 * plane is referenced from both sides, so the triangle array grows by a few per cent and primitive indices are positions in
 * that array; hits and images are the same, the render core visits 4-6 % fewer records on the bundled meshes).
 * verbose != 0 prints the reference's stdout lines (Loaded environment map..., Triangle count, BVH Took, Node count). */
int rt_scene_load(const char *path, int use_bvh, int verbose, rt_host_scene **out);
void rt_scene_free(rt_host_scene *scene);
/* Fills a descriptor whose pointers alias the host scene's arrays (valid until rt_scene_free / rt_scene_set_image_host). */
int rt_scene_describe(const rt_host_scene *scene, rt_scene_desc *out);
/* Override the `image` line and redo the camera precompute (scene.cu:62-76). exposure < 0 keeps the current value. */
int rt_scene_set_image_host(rt_host_scene *scene, int width, int height, int ray_count, int bounces, float exposure);
/* forward[3], up[3], vertical_fov (radians): the raw camera fields, for tests. */
int rt_scene_camera_raw(const rt_host_scene *scene, float *out7);
double rt_scene_bvh_build_ms(const rt_host_scene *scene);

/* 8-bit RGB PNG, rows top to bottom (what stbi_write_png does at raytracing.cu:395). */
int rt_write_png(const char *path, int width, int height, const uint8_t *rgb8);

#ifdef __cplusplus
}
#endif
#endif /* RT_SCENE_H */
