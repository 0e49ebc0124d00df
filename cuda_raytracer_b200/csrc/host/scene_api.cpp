// C ABI glue of the host scene layer (include/rt_scene.h).
#include <cstring>
#include <new>
#include <string>

#include "../../../include/rt_scene.h"
#include "host_scene.hpp"

struct rt_host_scene {
    rt::HostScene scene;
    bool pinned = false;
};

namespace {
// The big arrays a render context uploads (triangles, nodes, texels): page-locked for the scene's lifetime so that
// rt_scene_upload's copies are asynchronous DMA. A no-op without a CUDA device (rt_host_pin reports it, nothing depends on it).
template <class F>
void for_each_big_array(rt_host_scene *h, F f)
{
    rt::HostScene &s = h->scene;
    f(s.triangles.data(), s.triangles.size() * sizeof(rt::TriRecord));
    f(s.nodes.data(), s.nodes.size() * sizeof(rt::TreeNode));
    f(s.env.data(), s.env.size() * sizeof(float));
    f(s.material_indices.data(), s.material_indices.size() * sizeof(uint16_t));
}
}  // namespace

namespace rt {
thread_local std::string g_last_error;  // shared with the device side (rt_last_error(NULL))
void set_global_error(const std::string &msg) { g_last_error = msg; }
const char *global_error() { return g_last_error.c_str(); }
}  // namespace rt

extern "C" {

int rt_scene_load(const char *path, int use_bvh, int verbose, rt_host_scene **out)
{
    if (!path || !out) { rt::set_global_error("rt_scene_load: null argument"); return RT_ERR_INVALID; }
    *out = nullptr;
    rt_host_scene *h = new (std::nothrow) rt_host_scene();
    if (!h) return RT_ERR_NOMEM;
    try {
        rt::load_scene_file(h->scene, path, use_bvh < 0 ? 0 : (use_bvh > 3 ? 1 : use_bvh), verbose != 0);
    } catch (const rt::LoadError &e) {
        rt::set_global_error(e.message);
        delete h;
        return e.status;
    } catch (const std::exception &e) {
        rt::set_global_error(std::string("rt_scene_load: ") + e.what());
        delete h;
        return RT_ERR_IO;
    }
    if (rt_device_count() > 0) {
        h->pinned = true;
        for_each_big_array(h, [](const void *p, size_t bytes) { if (bytes >= 65536) rt_host_pin(p, bytes); });
    }
    *out = h;
    return RT_OK;
}

void rt_scene_free(rt_host_scene *scene)
{
    if (scene && scene->pinned) for_each_big_array(scene, [](const void *p, size_t bytes) { if (bytes >= 65536) rt_host_unpin(p); });
    delete scene;
}

int rt_scene_describe(const rt_host_scene *h, rt_scene_desc *d)
{
    if (!h || !d) return RT_ERR_INVALID;
    const rt::HostScene &s = h->scene;
    std::memset(d, 0, sizeof(*d));
    d->spheres = reinterpret_cast<const float *>(s.spheres.data());
    d->sphere_count = (int32_t)s.spheres.size();
    d->triangles = reinterpret_cast<const float *>(s.triangles.data());
    d->triangle_count = (int32_t)s.triangles.size();
    d->material_indices = s.material_indices.data();
    d->materials = reinterpret_cast<const float *>(s.materials.data());
    d->material_count = (int32_t)s.materials.size();
    d->bvh_nodes = s.nodes.data();
    d->bvh_node_count = (int32_t)s.nodes.size();
    d->environment_map = s.env.data();
    d->env_width = s.env_w;
    d->env_height = s.env_h;
    d->width = s.width;
    d->height = s.height;
    d->ray_count = s.ray_count;
    d->bounces = s.bounces;
    d->exposure = s.exposure;
    auto put = [](float *dst, const rt::F3 &v) { dst[0] = v.x; dst[1] = v.y; dst[2] = v.z; };
    put(d->camera_position, s.camera_position);
    put(d->near_plane_top_left, s.near_plane_top_left);
    put(d->scaled_right, s.scaled_right);
    put(d->scaled_up, s.scaled_up);
    d->inv_width = s.inv_width;
    d->inv_height = s.inv_height;
    put(d->min_coord, s.min_coord);
    put(d->inv_dimensions, s.inv_dimensions);
    d->triangles_form = s.triangles_form;
    return RT_OK;
}

int rt_scene_set_image_host(rt_host_scene *h, int width, int height, int ray_count, int bounces, float exposure)
{
    if (!h || width <= 0 || height <= 0 || ray_count < 0 || bounces < 0) return RT_ERR_INVALID;
    rt::HostScene &s = h->scene;
    s.width = width;
    s.height = height;
    s.ray_count = ray_count;
    s.bounces = bounces;
    if (exposure >= 0) s.exposure = exposure;
    s.precompute_camera();
    return RT_OK;
}

int rt_scene_camera_raw(const rt_host_scene *h, float *out7)
{
    if (!h || !out7) return RT_ERR_INVALID;
    const rt::HostScene &s = h->scene;
    out7[0] = s.forward.x; out7[1] = s.forward.y; out7[2] = s.forward.z;
    out7[3] = s.up.x; out7[4] = s.up.y; out7[5] = s.up.z;
    out7[6] = s.vertical_fov;
    return RT_OK;
}

double rt_scene_bvh_build_ms(const rt_host_scene *h) { return h ? h->scene.bvh_ms : 0.0; }

}  // extern "C"
