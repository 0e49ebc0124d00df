// Host-side scene representation shared by the loader, the BVH builder and the C ABI glue.
// Layouts of the flat arrays are the reference's (scene.cuh) because they cross the C ABI unchanged.
#pragma once

#include <cmath>
#include <cstdint>
#include <string>
#include <vector>

namespace rt {

struct F3 {
    float x = 0, y = 0, z = 0;
    float operator[](int axis) const { return (&x)[axis]; }
};

// fminf/fmaxf with exactly glibc's x86-64 result (MINSS/MAXSS: ties and +-0 return the second operand) but inlinable;
// NaN operands take the libm path. The reference's host `min`/`max` are fminf/fmaxf (CUDA math headers), and the
// sign of a zero bound matters downstream (1/max_coord), so the tie rule is part of the contract.
inline float min_f(float x, float y) { return (x != x || y != y) ? std::fmin(x, y) : (x < y ? x : y); }
inline float max_f(float x, float y) { return (x != x || y != y) ? std::fmax(x, y) : (x > y ? x : y); }

inline F3 operator+(F3 a, F3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline F3 operator-(F3 a, F3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline F3 operator*(float s, F3 a) { return {s * a.x, s * a.y, s * a.z}; }
inline F3 cross(F3 a, F3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
inline float length(F3 a) { return std::sqrt(a.x * a.x + a.y * a.y + a.z * a.z); }
inline F3 unit(F3 a) { return (1.0f / length(a)) * a; }

// 48-byte triangle record. While the tree is being built slot[1], slot[2] hold the 2nd/3rd vertex and slot[3]
// the centroid; finalise() turns it into the render form {p1, p2-p1, p3-p1, unit normal} (scene.cuh:15-33).
struct TriRecord {
    F3 slot[4];
};
static_assert(sizeof(TriRecord) == 48, "triangle record must match the 48-byte ABI layout");

struct SphereRecord {
    F3 center;
    float radius;
};
static_assert(sizeof(SphereRecord) == 16, "sphere record");

struct MaterialRecord {  // scene.cuh:42-59
    F3 diffuse{1, 1, 1};
    float metallicity = 0;
    F3 specular{1, 1, 1};
    float roughness = 0;
    F3 emitted{0, 0, 0};
    float ior = 0;
};
static_assert(sizeof(MaterialRecord) == 48, "material record");

struct Box {
    F3 lo{1e30f, 1e30f, 1e30f};  // large finite sentinels, as the reference (scene.cuh:70-74)
    F3 hi{-1e30f, -1e30f, -1e30f};
    void grow(F3 p);
    void grow(const TriRecord &t);  // build form: three vertices
    void grow(const Box &b);
    float half_area() const;
};

struct alignas(32) TreeNode {  // scene.cuh:82-100
    Box box;
    int32_t child1 = 0;  // internal: left child index; leaf: END of triangle range
    int32_t child2 = 0;  // internal: right child index (= child1 + 1); leaf: START of triangle range
    bool leaf() const { return child2 <= child1; }
};
static_assert(sizeof(TreeNode) == 32, "tree node");

struct HostScene {
    std::vector<SphereRecord> spheres;
    std::vector<TriRecord> triangles;
    std::vector<uint16_t> material_indices;  // spheres first, then triangles (BVH order)
    std::vector<MaterialRecord> materials;
    std::vector<TreeNode> nodes;
    std::vector<float> env;  // rgb texels
    int env_w = 0, env_h = 0;

    int width = 1920, height = 1080, ray_count = 1, bounces = 3;  // scene.cu:571-574
    float exposure = 0;                                           // zero-initialised Scene, raytracing.cu:344
    F3 camera_position, forward, up;
    float vertical_fov = 0;
    F3 scaled_right, scaled_up, near_plane_top_left;
    float inv_width = 0, inv_height = 0;
    F3 min_coord, inv_dimensions;
    double bvh_ms = 0;
    int triangles_form = 0;  // 1 = build form, file order: the BVH is built by the render core on the device

    void precompute_camera();
    void build_tree(int max_depth, bool verbose, bool spatial_splits = false);  // bvh_builder.cpp
    void defer_tree(bool verbose);                 // bvh_builder.cpp: root node only, triangles stay in build form
    void compute_bounds();
};

struct LoadError {
    int status;
    std::string message;
};

// Throws LoadError. scene_loader.cpp
void load_scene_file(HostScene &scene, const char *path, int use_bvh, bool verbose);  // 0 no_bvh, 1 host build, 2 deferred, 3 host build with spatial splits

}  // namespace rt
