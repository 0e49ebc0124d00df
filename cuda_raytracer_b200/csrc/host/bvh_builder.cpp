// Binned-SAH BVH build on the host (8 bins, leaves of <= 4 triangles, depth <= 30), followed by the conversion
// of the triangles to render form. Reproduces the split decisions of the reference builder
// (/root/reference/scene.cu:866-1036) so that node order and triangle order are identical and primitive indices
// are directly comparable; tests/test_host_scene.py checks the arrays byte for byte against the reference's own
// load_scene. The code is an explicit work-list formulation rather than the reference's recursive member function.
#include "host_scene.hpp"

#include <algorithm>
#include <chrono>
#include <cstdio>
#include <cstdlib>

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

// ---- optional: spatial splits (Stich, Friedrich, Dietrich: "Spatial Splits in Bounding Volume Hierarchies", 2009) ----
// NOT the reference's builder: selected with rt_scene_load(use_bvh = 3) (front-end word `sbvh`), or RT_SBVH=1 in the
// environment of the loader. Same hits and images, fewer visited records (profiles/r1m_sbvh.md). A node is split either by object (binned SAH on the centroids of the reference boxes) or by space (a plane
// through the node; triangles that straddle it are referenced from both children with their boxes clipped to each side),
// whichever is cheaper under the same cost as above (count x half-area). The output uses the reference's array layouts:
// a triangle referenced from several leaves appears several times in the triangle array.
struct Ref {
    int tri;
    Box box;
};

#ifndef RT_SBVH_OBJ_BINS
#define RT_SBVH_OBJ_BINS 16
#endif
#ifndef RT_SBVH_SPACE_BINS
#define RT_SBVH_SPACE_BINS 32
#endif
constexpr int kObjBins = RT_SBVH_OBJ_BINS, kSpaceBins = RT_SBVH_SPACE_BINS;

inline Box intersect(const Box &a, const Box &b)
{
    Box r;
    r.lo = {max_f(a.lo.x, b.lo.x), max_f(a.lo.y, b.lo.y), max_f(a.lo.z, b.lo.z)};
    r.hi = {min_f(a.hi.x, b.hi.x), min_f(a.hi.y, b.hi.y), min_f(a.hi.z, b.hi.z)};
    return r;
}
inline bool valid(const Box &b) { return b.lo.x <= b.hi.x && b.lo.y <= b.hi.y && b.lo.z <= b.hi.z; }
inline float &comp(F3 &v, int axis) { return (&v.x)[axis]; }

// keep the part of a convex polygon with sign * (p[axis] - pos) >= 0 (Sutherland-Hodgman, one plane)
int clip_polygon(const F3 *in, int n, int axis, float pos, float sign, F3 *out)
{
    int m = 0;
    for (int i = 0; i < n; i++) {
        const F3 a = in[i], b = in[(i + 1) % n];
        const float da = sign * (a[axis] - pos), db = sign * (b[axis] - pos);
        if (da >= 0) out[m++] = a;
        if ((da >= 0) != (db >= 0)) {
            const float t = da / (da - db);
            F3 p = a + t * (b - a);
            comp(p, axis) = pos;
            out[m++] = p;
        }
    }
    return m;
}

// bounding box of (triangle clipped to lo <= p[axis] <= hi), padded by a few ulps on the other axes (the clip vertices are
// rounded) and cut to `within`
Box clipped_box(const TriRecord &t, int axis, float lo, float hi, const Box &within)
{
    F3 a[10], b[10];
    a[0] = t.slot[0]; a[1] = t.slot[1]; a[2] = t.slot[2];
    int n = clip_polygon(a, 3, axis, lo, 1.0f, b);
    n = clip_polygon(b, n, axis, hi, -1.0f, a);
    Box r;
    for (int i = 0; i < n; i++) r.grow(a[i]);
    if (n == 0) return r;
    for (int k = 0; k < 3; k++) {
        if (k == axis) continue;
        const float pad = 4e-7f * max_f(std::fabs(r.lo[k]), std::fabs(r.hi[k])) + 1e-30f;
        comp(r.lo, k) -= pad;
        comp(r.hi, k) += pad;
    }
    return intersect(r, within);
}


struct SpatialBuilder {
    const std::vector<TriRecord> &tris;  // build form
    std::vector<TreeNode> &nodes;
    std::vector<int> leaf_order;         // triangle ids in leaf order, with repetitions
    float root_area = 0;
    size_t ref_budget = 0;               // stop splitting in space once this many references exist
    size_t ref_total = 0;
    float alpha = 1e-5f;                 // spatial splits are tried where the object split's children overlap by more than alpha x root area
    long spatial_splits = 0, object_splits = 0;

    struct Choice {
        bool spatial = false;
        int axis = -1;
        float position = 0, cost = 0;
    };

    Choice object_split(const std::vector<Ref> &refs, float limit, Box *left_box, Box *right_box) const
    {
        Choice best;
        best.cost = limit;
        const int count = (int)refs.size();
        for (int axis = 0; axis < 3; axis++) {
            float c_min = 1e30f, c_max = -1e30f;
            for (const Ref &r : refs) {
                const float c = 0.5f * (r.box.lo[axis] + r.box.hi[axis]);
                c_min = min_f(c_min, c);
                c_max = max_f(c_max, c);
            }
            if (!(c_min < c_max)) continue;
            const float to_bin = kObjBins / (c_max - c_min);
            Box bin_box[kObjBins];
            int bin_count[kObjBins] = {};
            for (const Ref &r : refs) {
                const int b = std::min(kObjBins - 1, (int)((0.5f * (r.box.lo[axis] + r.box.hi[axis]) - c_min) * to_bin));
                bin_count[b]++;
                bin_box[b].grow(r.box);
            }
            Box right_sweep[kObjBins];
            Box acc;
            for (int k = kObjBins - 1; k >= 1; k--) { acc.grow(bin_box[k]); right_sweep[k] = acc; }
            Box left;
            int n_left = 0;
            for (int k = 0; k + 1 < kObjBins; k++) {
                left.grow(bin_box[k]);
                n_left += bin_count[k];
                if (n_left == 0 || n_left == count) continue;
                const float cost = n_left * left.half_area() + (count - n_left) * right_sweep[k + 1].half_area();
                if (cost < best.cost) {
                    best.axis = axis;
                    best.position = c_min + (c_max - c_min) * (float)(k + 1) / kObjBins;
                    best.cost = cost;
                    *left_box = left;
                    *right_box = right_sweep[k + 1];
                }
            }
        }
        return best;
    }

    Choice spatial_split(const std::vector<Ref> &refs, const Box &box, float limit) const
    {
        Choice best;
        best.spatial = true;
        best.cost = limit;
        const int count = (int)refs.size();
        for (int axis = 0; axis < 3; axis++) {
            const float lo = box.lo[axis], hi = box.hi[axis];
            if (!(lo < hi)) continue;
            const float width = (hi - lo) / kSpaceBins, to_bin = kSpaceBins / (hi - lo);
            Box bin_box[kSpaceBins];
            int enter[kSpaceBins] = {}, leave[kSpaceBins] = {};
            auto edge = [&](int k) { return k >= kSpaceBins ? hi : lo + width * (float)k; };
            for (const Ref &r : refs) {
                const int b0 = std::max(0, std::min(kSpaceBins - 1, (int)((r.box.lo[axis] - lo) * to_bin)));
                const int b1 = std::max(b0, std::min(kSpaceBins - 1, (int)((r.box.hi[axis] - lo) * to_bin)));
                enter[b0]++;
                leave[b1]++;
                if (b0 == b1) { bin_box[b0].grow(r.box); continue; }
                for (int b = b0; b <= b1; b++) {
                    const Box part = clipped_box(tris[r.tri], axis, edge(b), edge(b + 1), r.box);
                    if (valid(part)) bin_box[b].grow(part);
                }
            }
            Box right_sweep[kSpaceBins];
            int right_count[kSpaceBins + 1] = {};
            Box acc;
            for (int k = kSpaceBins - 1; k >= 1; k--) {
                acc.grow(bin_box[k]);
                right_sweep[k] = acc;
                right_count[k] = right_count[k + 1] + leave[k];
            }
            Box left;
            int n_left = 0;
            for (int k = 0; k + 1 < kSpaceBins; k++) {
                left.grow(bin_box[k]);
                n_left += enter[k];
                const int n_right = right_count[k + 1];
                if (n_left == 0 || n_right == 0 || (n_left == count && n_right == count)) continue;
                const float cost = n_left * left.half_area() + n_right * right_sweep[k + 1].half_area();
                if (cost < best.cost) {
                    best.axis = axis;
                    best.position = edge(k + 1);
                    best.cost = cost;
                }
            }
        }
        return best;
    }

    void make_leaf(int node_index, const std::vector<Ref> &refs)
    {
        nodes[node_index].child2 = (int)leaf_order.size();
        for (const Ref &r : refs) leaf_order.push_back(r.tri);
        nodes[node_index].child1 = (int)leaf_order.size();
    }

    void refine(int node_index, std::vector<Ref> &refs, int depth_left)
    {
        Box box;
        for (const Ref &r : refs) box.grow(r.box);
        nodes[node_index].box = box;
        const int count = (int)refs.size();
        if (count <= RT_HOST_LEAF || depth_left == 0) { make_leaf(node_index, refs); return; }

        const float leaf_cost = box.half_area() * count;
        Box ol, orr;
        Choice pick = object_split(refs, leaf_cost, &ol, &orr);
        bool try_space = ref_total < ref_budget;
        if (try_space && pick.axis >= 0) {  // only where the object split's children overlap noticeably
            const Box ov = intersect(ol, orr);
            try_space = valid(ov) && ov.half_area() > alpha * root_area;
        }
        if (try_space) {
            const Choice sp = spatial_split(refs, box, pick.cost);
            if (sp.axis >= 0) pick = sp;
        }
        if (pick.axis < 0) { make_leaf(node_index, refs); return; }

        std::vector<Ref> left, right;
        left.reserve(refs.size());
        right.reserve(refs.size() / 2);
        const int axis = pick.axis;
        if (!pick.spatial) {
            for (const Ref &r : refs) (0.5f * (r.box.lo[axis] + r.box.hi[axis]) < pick.position ? left : right).push_back(r);
            object_splits++;
        } else {
            for (const Ref &r : refs) {
                if (r.box.hi[axis] <= pick.position) { left.push_back(r); continue; }
                if (r.box.lo[axis] >= pick.position) { right.push_back(r); continue; }
                const Box lb = clipped_box(tris[r.tri], axis, r.box.lo[axis], pick.position, r.box);
                const Box rb = clipped_box(tris[r.tri], axis, pick.position, r.box.hi[axis], r.box);
                if (valid(lb)) left.push_back(Ref{r.tri, lb});
                if (valid(rb)) right.push_back(Ref{r.tri, rb});
                if (valid(lb) && valid(rb)) ref_total++;
            }
            spatial_splits++;
        }
        if (left.empty() || right.empty() || (left.size() == refs.size() && right.size() == refs.size())) { make_leaf(node_index, refs); return; }
        std::vector<Ref>().swap(refs);  // release before recursing

        const int first = (int)nodes.size();
        nodes.emplace_back();
        nodes.emplace_back();
        refine(first, left, depth_left - 1);
        refine(first + 1, right, depth_left - 1);
        nodes[node_index].child1 = first;
        nodes[node_index].child2 = first + 1;
    }
};

}  // namespace

void HostScene::build_tree(int max_depth, bool verbose, bool spatial_splits)
{
    auto t0 = std::chrono::high_resolution_clock::now();
    const int tri_count = (int)triangles.size();
    nodes.clear();
    nodes.reserve((size_t)tri_count * 2 + 1);
    nodes.emplace_back();
    nodes[0].child2 = 0;
    nodes[0].child1 = tri_count;
    const char *sbvh = std::getenv("RT_SBVH");
    if (((sbvh && std::atoi(sbvh) > 0) || spatial_splits) && tri_count > 0 && max_depth > 0) {  // spatial-split builder (not the reference's tree)
        SpatialBuilder sb{triangles, nodes, {}};
        std::vector<Ref> refs((size_t)tri_count);
        Box root;
        for (int i = 0; i < tri_count; i++) {
            refs[i].tri = i;
            refs[i].box.grow(triangles[i]);
            root.grow(refs[i].box);
        }
        sb.root_area = root.half_area();
        const char *budget = std::getenv("RT_SBVH_BUDGET");  // references allowed, as a multiple of the triangle count
        sb.ref_budget = (size_t)((budget ? std::atof(budget) : 1.5) * tri_count);
        sb.ref_total = (size_t)tri_count;
        if (const char *al = std::getenv("RT_SBVH_ALPHA")) sb.alpha = (float)std::atof(al);
        sb.refine(0, refs, max_depth);
        std::vector<TriRecord> ordered(sb.leaf_order.size());
        std::vector<uint16_t> mats(spheres.size() + sb.leaf_order.size());
        for (size_t i = 0; i < spheres.size(); i++) mats[i] = material_indices[i];
        for (size_t i = 0; i < sb.leaf_order.size(); i++) {
            ordered[i] = triangles[sb.leaf_order[i]];
            mats[spheres.size() + i] = material_indices[spheres.size() + sb.leaf_order[i]];
        }
        triangles.swap(ordered);
        material_indices.swap(mats);
        if (verbose) std::printf("Spatial splits: %ld, object splits: %ld, references: %zu\n", sb.spatial_splits, sb.object_splits, triangles.size());
    } else {
        Builder builder{*this, nodes, material_indices.data() + spheres.size()};
        builder.refine(0, max_depth);
    }
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
