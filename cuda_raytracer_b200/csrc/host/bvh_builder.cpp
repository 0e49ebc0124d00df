// Binned-SAH BVH build on the host (8 bins, leaves of <= 4 triangles, depth <= 30), followed by the conversion
// of the triangles to render form. Reproduces the split decisions of the reference builder
// (/root/reference/scene.cu:866-1036) so that node order and triangle order are identical and primitive indices
// are directly comparable; tests/test_host_scene.py checks the arrays byte for byte against the reference's own
// load_scene. The code is an explicit work-list formulation rather than the reference's recursive member function.
#include "host_scene.hpp"

#include <algorithm>
#include <chrono>
#include <cstdio>

namespace rt {

void Box::grow(F3 p)
{
    lo = {min_f(lo.x, p.x), min_f(lo.y, p.y), min_f(lo.z, p.z)};
    hi = {max_f(hi.x, p.x), max_f(hi.y, p.y), max_f(hi.z, p.z)};
}
void Box::grow(const TriRecord &t)
{
    grow(t.slot[0]);
    grow(t.slot[1]);
    grow(t.slot[2]);
}
void Box::grow(const Box &b)
{
    lo = {min_f(lo.x, b.lo.x), min_f(lo.y, b.lo.y), min_f(lo.z, b.lo.z)};
    hi = {max_f(hi.x, b.hi.x), max_f(hi.y, b.hi.y), max_f(hi.z, b.hi.z)};
}
float Box::half_area() const
{
    F3 d = hi - lo;
    return d.x * d.y + d.x * d.z + d.y * d.z;
}

namespace {

#ifndef RT_HOST_BINS
#define RT_HOST_BINS 8  // the reference's bin count (scene.cu:866-1000); other values only for tree-quality experiments
#endif
constexpr int kBins = RT_HOST_BINS;
#ifndef RT_HOST_LEAF
#define RT_HOST_LEAF 4  // the reference stops splitting at <= 4 triangles (scene.cu:880)
#endif

struct SplitPlane {
    int axis = -1;
    float position = 0;
    float cost = 0;
};

// Cheapest of the 7 bin boundaries on each axis, or axis == -1 if none beats `parent_cost`.
// An empty side has an inverted sentinel box whose half_area overflows to +inf; 0 * inf = NaN fails the `<`
// compare, which is how such planes are rejected — this file must not be built with -ffast-math.
SplitPlane best_plane(const std::vector<TriRecord> &tris, int begin, int end, float parent_cost)
{
    SplitPlane best;
    best.cost = parent_cost;
    const int count = end - begin;
    for (int axis = 0; axis < 3; axis++) {
        float c_min = 1e30f, c_max = -1e30f;
        for (int i = begin; i < end; i++) {
            float c = tris[i].slot[3][axis];
            c_min = min_f(c_min, c);
            c_max = max_f(c_max, c);
        }
        if (c_min == c_max) continue;

        float to_bin = kBins / (c_max - c_min);
        Box bin_box[kBins];
        int bin_count[kBins] = {};
        for (int i = begin; i < end; i++) {
            int b = std::min(kBins - 1, (int)((tris[i].slot[3][axis] - c_min) * to_bin));
            bin_count[b]++;
            bin_box[b].grow(tris[i]);
        }

        float area_left[kBins - 1], area_right[kBins - 1];
        int count_left[kBins - 1];
        Box sweep_l, sweep_r;
        int running = 0;
        for (int k = 0; k + 1 < kBins; k++) {
            running += bin_count[k];
            count_left[k] = running;
            sweep_l.grow(bin_box[k]);
            area_left[k] = sweep_l.half_area();
            sweep_r.grow(bin_box[kBins - 1 - k]);
            area_right[kBins - 2 - k] = sweep_r.half_area();
        }

        float bin_width = (c_max - c_min) / kBins;
        for (int k = 0; k + 1 < kBins; k++) {
            float cost = count_left[k] * area_left[k] + (count - count_left[k]) * area_right[k];
            if (cost != 0 && cost < best.cost) {
                best.axis = axis;
                best.position = c_min + bin_width * (k + 1);
                best.cost = cost;
            }
        }
    }
    return best;
}

struct Builder {
    HostScene &scene;
    std::vector<TreeNode> &nodes;
    uint16_t *tri_material;  // material_indices + sphere_count

    // Depth-first, left subtree completed before the right one starts; children are appended as an adjacent pair
    // before either is refined, which gives child2 == child1 + 1 for every internal node (Q13).
    void refine(int node_index, int depth_left)
    {
        int begin = nodes[node_index].child2, end = nodes[node_index].child1;
        Box box;
        for (int i = begin; i < end; i++) box.grow(scene.triangles[i]);
        nodes[node_index].box = box;

        int count = end - begin;
        if (count <= RT_HOST_LEAF || depth_left == 0) return;

        float leaf_cost = box.half_area() * count;
        SplitPlane plane = best_plane(scene.triangles, begin, end, leaf_cost);
        if (plane.axis < 0 || plane.cost >= leaf_cost) return;

        // in-place two-pointer partition on the centroid; material indices move with their triangles
        int i = begin, j = end - 1;
        while (i <= j) {
            if (scene.triangles[i].slot[3][plane.axis] < plane.position) {
                i++;
            } else {
                std::swap(scene.triangles[i], scene.triangles[j]);
                std::swap(tri_material[i], tri_material[j]);
                j--;
            }
        }
        if (i == begin || i == end) return;

        int left = (int)nodes.size();
        nodes.emplace_back();
        nodes.emplace_back();
        nodes[left].child2 = begin;
        nodes[left].child1 = i;
        nodes[left + 1].child2 = i;
        nodes[left + 1].child1 = end;
        refine(left, depth_left - 1);
        refine(left + 1, depth_left - 1);
        nodes[node_index].child1 = left;
        nodes[node_index].child2 = left + 1;
    }
};

}  // namespace

void HostScene::build_tree(int max_depth, bool verbose)
{
    auto t0 = std::chrono::high_resolution_clock::now();
    const int tri_count = (int)triangles.size();
    nodes.clear();
    nodes.reserve((size_t)tri_count * 2 + 1);
    nodes.emplace_back();
    nodes[0].child2 = 0;
    nodes[0].child1 = tri_count;
    Builder builder{*this, nodes, material_indices.data() + spheres.size()};
    builder.refine(0, max_depth);
    auto t1 = std::chrono::high_resolution_clock::now();
    bvh_ms = std::chrono::duration<double, std::milli>(t1 - t0).count();
    if (verbose) {  // same three stdout lines as the reference (scene.cu:1023-1026)
        std::printf("Triangle count: %d\n", tri_count);
        std::printf("BVH Took %gms\n", bvh_ms);
        std::printf("Node count: %d\n", (int)nodes.size());
    }

    // build form -> render form {p1, e1 = p2-p1, e2 = p3-p1, n = unit(e2 x e1)} (scene.cu:1029-1035, Q9)
    for (TriRecord &t : triangles) {
        t.slot[1] = t.slot[1] - t.slot[0];
        t.slot[2] = t.slot[2] - t.slot[0];
        t.slot[3] = unit(cross(t.slot[2], t.slot[1]));
    }
}

// Deferred build (rt_scene_load with use_bvh = 2): only the root node and its box exist on the host (the scene bounds of
// scene.cu:822-830 need it); the render core builds the tree on the device from the build-form triangles.
void HostScene::defer_tree(bool verbose)
{
    nodes.clear();
    nodes.emplace_back();
    nodes[0].child2 = 0;
    nodes[0].child1 = (int)triangles.size();
    Box box;
    for (const TriRecord &t : triangles) box.grow(t);
    nodes[0].box = box;
    triangles_form = 1;
    bvh_ms = 0;
    if (verbose) std::printf("Triangle count: %d\n", (int)triangles.size());
}

}  // namespace rt
