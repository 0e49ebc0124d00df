// Scene-file parser, PLY and PFM readers: the host feed of the render path.
// Behaviour follows /root/reference/scene.cu:489-831 (grammar in SURVEY.md §9); the implementation is a plain
// whitespace tokenizer with explicit error reporting instead of the reference's istream juggling.
#include "host_scene.hpp"

#include <algorithm>
#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>
#include <thread>
#include <unordered_map>

#include "../../../include/rt_core.h"

namespace rt {
namespace {

[[noreturn]] void fail(int status, const std::string &msg) { throw LoadError{status, msg}; }

// One line split at single spaces. The reference alternates getline(' ') and operator>>, which makes every keyword
// after a number arrive behind an empty token; splitting on runs of blanks gives the same token sequence for every
// well-formed file and is indifferent to a trailing CR.
struct Tokens {
    std::vector<std::string> words;
    size_t next = 0;

    explicit Tokens(const std::string &line)
    {
        size_t i = 0, n = line.size();
        while (i < n) {
            while (i < n && (line[i] == ' ' || line[i] == '\t' || line[i] == '\r')) i++;
            size_t start = i;
            while (i < n && line[i] != ' ' && line[i] != '\t' && line[i] != '\r') i++;
            if (i > start) words.emplace_back(line, start, i - start);
        }
    }
    bool more() const { return next < words.size(); }
    const std::string &word() { static const std::string empty; return more() ? words[next++] : empty; }
    // operator>> on a float leaves the target unchanged-or-zero at end of input; a missing number reads as 0.
    float number()
    {
        if (!more()) return 0.0f;
        return std::strtof(words[next++].c_str(), nullptr);
    }
    int integer()
    {
        if (!more()) return 0;
        return (int)std::strtol(words[next++].c_str(), nullptr, 10);
    }
    F3 vec() { F3 v; v.x = number(); v.y = number(); v.z = number(); return v; }
};

TriRecord build_form(F3 a, F3 b, F3 c)
{
    TriRecord t;
    t.slot[0] = a;
    t.slot[1] = b;
    t.slot[2] = c;
    t.slot[3] = (1.0f / 3.0f) * (a + b + c);  // centroid, used by the SAH binning (scene.cu:529)
    return t;
}

// Whole file into memory with one read (the meshes are a few hundred KB each; per-face istream calls were most of the
// load time of lamp.scene).
std::vector<char> read_whole_file(const std::string &path, const char *what)
{
    std::FILE *f = std::fopen(path.c_str(), "rb");
    if (!f) fail(RT_ERR_IO, std::string("cannot open ") + what + " file '" + path + "'");
    std::vector<char> data;
    if (std::fseek(f, 0, SEEK_END) == 0) {
        const long size = std::ftell(f);
        std::rewind(f);
        if (size > 0) {
            data.resize((size_t)size);
            if (std::fread(data.data(), 1, data.size(), f) != data.size()) data.clear();
        }
    }
    std::fclose(f);
    return data;
}

// Binary little-endian PLY as exported for the bundled meshes (scene.cu:491-546): float vertex properties (the
// first three are the position), faces as uint8 count + int32 indices, fan-triangulated.
void read_ply(std::vector<TriRecord> &out, const std::string &path)
{
    const std::vector<char> data = read_whole_file(path, "PLY");
    const char *p = data.data(), *end = p + data.size();
    long vertex_count = -1, face_count = -1;
    int vertex_floats = 0;
    bool in_vertex = false, header_done = false;
    for (int guard = 0; guard < 256 && p < end; guard++) {
        const char *nl = static_cast<const char *>(std::memchr(p, '\n', (size_t)(end - p)));
        if (!nl) break;
        std::string line(p, nl);
        p = nl + 1;
        if (!line.empty() && line.back() == '\r') line.pop_back();
        if (line.rfind("element vertex", 0) == 0) { vertex_count = std::atol(line.c_str() + 14); in_vertex = true; }
        else if (line.rfind("element face", 0) == 0) { face_count = std::atol(line.c_str() + 12); in_vertex = false; }
        else if (line.rfind("element", 0) == 0) in_vertex = false;
        else if (in_vertex && line.rfind("property float", 0) == 0) vertex_floats++;
        else if (line == "end_header") { header_done = true; break; }
    }
    if (!header_done || vertex_count < 0 || face_count < 0 || vertex_floats < 3)
        fail(RT_ERR_IO, "unsupported PLY header in '" + path + "'");

    const size_t vertex_bytes = (size_t)vertex_count * vertex_floats * sizeof(float);
    if ((size_t)(end - p) < vertex_bytes) fail(RT_ERR_IO, "truncated PLY vertex block in '" + path + "'");
    const char *vertices = p;
    p += vertex_bytes;
    auto position = [&](int32_t i) -> F3 {
        if (i < 0 || i >= vertex_count) fail(RT_ERR_IO, "PLY face index out of range in '" + path + "'");
        float v[3];
        std::memcpy(v, vertices + (size_t)i * vertex_floats * sizeof(float), sizeof(v));  // the block is not 4-byte aligned
        return {v[0], v[1], v[2]};
    };
    out.reserve(out.size() + (size_t)face_count);
    int32_t corner[256];
    for (long f = 0; f < face_count; f++) {
        if (p >= end) fail(RT_ERR_IO, "truncated PLY face block in '" + path + "'");
        const int n = (unsigned char)*p++;
        if ((size_t)(end - p) < sizeof(int32_t) * (size_t)n) fail(RT_ERR_IO, "truncated PLY face block in '" + path + "'");
        std::memcpy(corner, p, sizeof(int32_t) * (size_t)n);
        p += sizeof(int32_t) * (size_t)n;
        for (int k = 2; k < n; k++) out.push_back(build_form(position(corner[0]), position(corner[k - 1]), position(corner[k])));
    }
}

// PFM as the reference reads it (scene.cu:548-567): "PF", "W H", scale line, then raw float rgb in file order;
// scale sign / endianness are not interpreted.
void read_pfm(HostScene &scene, const std::string &path)
{
    std::ifstream file(path, std::ios::binary);
    if (!file) fail(RT_ERR_IO, "cannot open PFM file '" + path + "'");
    std::string magic, dims, scale;
    std::getline(file, magic);
    std::getline(file, dims);
    std::getline(file, scale);
    int w = 0, h = 0;
    if (std::sscanf(dims.c_str(), "%d %d", &w, &h) != 2 || w <= 0 || h <= 0) fail(RT_ERR_IO, "bad PFM header in '" + path + "'");
    scene.env.assign((size_t)w * h * 3, 0.0f);
    file.read(reinterpret_cast<char *>(scene.env.data()), (std::streamsize)(scene.env.size() * sizeof(float)));
    if (!file) fail(RT_ERR_IO, "truncated PFM payload in '" + path + "'");
    scene.env_w = w;
    scene.env_h = h;
}

}  // namespace

void HostScene::precompute_camera()  // scene.cu:62-76
{
    F3 right = cross(up, forward);
    float plane_h = 2.0f * std::tan(vertical_fov * 0.5f);
    float plane_w = plane_h * width / height;
    scaled_right = plane_w * right;
    scaled_up = plane_h * up;
    near_plane_top_left = forward - 0.5f * scaled_right + 0.5f * scaled_up;
    inv_width = 1.0f / (width - 1);
    inv_height = 1.0f / (height - 1);
}

void HostScene::compute_bounds()  // scene.cu:822-830; note inv = 1/max, not 1/(max-min) (Q2)
{
    min_coord = nodes[0].box.lo;
    F3 top = nodes[0].box.hi;
    for (const SphereRecord &s : spheres) {
        F3 r{s.radius, s.radius, s.radius};
        F3 hi = s.center + r, lo = s.center - r;
        top = {max_f(top.x, hi.x), max_f(top.y, hi.y), max_f(top.z, hi.z)};
        min_coord = {min_f(min_coord.x, lo.x), min_f(min_coord.y, lo.y), min_f(min_coord.z, lo.z)};
    }
    inv_dimensions = {1 / top.x, 1 / top.y, 1 / top.z};
}

void load_scene_file(HostScene &scene, const char *path, int use_bvh, bool verbose)
{
    std::ifstream file(path);
    if (!file) fail(RT_ERR_IO, std::string("cannot open scene file '") + path + "'");

    std::unordered_map<std::string, uint16_t> material_ids;
    std::vector<uint16_t> sphere_materials, triangle_materials;
    // `ply` lines are decoded after the parse, several files at a time; `at` is where the mesh goes in the list of the
    // triangles written inline (triangle / quad), so the final order is the file's order, as in the reference
    struct PlyJob {
        std::string path;
        uint16_t material;
        size_t at;
        std::vector<TriRecord> triangles;
        int status = RT_OK;
        std::string error;
    };
    std::vector<PlyJob> jobs;
    auto material_of = [&](const std::string &name, int line_no) -> uint16_t {
        auto it = material_ids.find(name);
        if (it == material_ids.end())
            fail(RT_ERR_PARSE, std::string(path) + ":" + std::to_string(line_no) + ": material '" + name + "' used before definition");
        return it->second;
    };

    std::string line;
    for (int line_no = 1; std::getline(file, line); line_no++) {
        Tokens t(line);
        if (!t.more()) continue;
        const std::string cmd = t.word();
        if (cmd == "sky") {  // constant sky = 1x1 map
            F3 c = t.vec();
            scene.env = {c.x, c.y, c.z};
            scene.env_w = scene.env_h = 1;
        } else if (cmd == "sky_map") {
            read_pfm(scene, t.word());
            if (verbose) std::printf("Loaded environment map with size %d,%d\n", scene.env_w, scene.env_h);
        } else if (cmd == "camera") {  // keywords are positional, not matched (scene.cu:617-646)
            t.word(); scene.camera_position = t.vec();
            t.word(); scene.forward = unit(t.vec());
            t.word(); scene.up = unit(t.vec());
            t.word(); scene.vertical_fov = (float)(t.number() * (M_PI / 180));
        } else if (cmd == "material") {
            if (scene.materials.size() >= 65535) fail(RT_ERR_PARSE, "too many materials (uint16 index)");
            material_ids[t.word()] = (uint16_t)scene.materials.size();
            MaterialRecord m;
            while (t.more()) {
                const std::string key = t.word();
                if (key == "diffuse") m.diffuse = t.vec();
                else if (key == "specular") m.specular = t.vec();
                else if (key == "emit") m.emitted = t.vec();
                else if (key == "metallicity") m.metallicity = t.number();
                else if (key == "roughness") m.roughness = t.number();
                else if (key == "ior") m.ior = t.number();
            }
            scene.materials.push_back(m);
        } else if (cmd == "sphere") {
            sphere_materials.push_back(material_of(t.word(), line_no));
            SphereRecord s;
            s.center = t.vec();
            s.radius = t.number();
            scene.spheres.push_back(s);
        } else if (cmd == "triangle") {
            triangle_materials.push_back(material_of(t.word(), line_no));
            F3 a = t.vec(), b = t.vec(), c = t.vec();
            scene.triangles.push_back(build_form(a, b, c));
        } else if (cmd == "quad") {  // (p1,p2,p3) and (p1,p3,p4), Q15
            uint16_t m = material_of(t.word(), line_no);
            F3 a = t.vec(), b = t.vec(), c = t.vec(), d = t.vec();
            scene.triangles.push_back(build_form(a, b, c));
            scene.triangles.push_back(build_form(a, c, d));
            triangle_materials.push_back(m);
            triangle_materials.push_back(m);
        } else if (cmd == "ply") {
            PlyJob job;
            job.material = material_of(t.word(), line_no);
            job.path = t.word();
            job.at = scene.triangles.size();
            jobs.push_back(std::move(job));
        } else if (cmd == "image") {
            scene.width = t.integer();
            scene.height = t.integer();
            scene.ray_count = t.integer();
            scene.bounces = t.integer();
            scene.exposure = t.number();
        }  // unknown commands are ignored, as in the reference
    }

    if (!jobs.empty()) {
        std::atomic<size_t> next_job{0};
        auto worker = [&]() {
            for (size_t j; (j = next_job.fetch_add(1)) < jobs.size();) {
                try { read_ply(jobs[j].triangles, jobs[j].path); }
                catch (const LoadError &e) { jobs[j].status = e.status; jobs[j].error = e.message; }
                catch (const std::exception &e) { jobs[j].status = RT_ERR_IO; jobs[j].error = "reading '" + jobs[j].path + "': " + e.what(); }
            }
        };
        const size_t n_threads = std::min<size_t>({jobs.size(), (size_t)std::max(1u, std::thread::hardware_concurrency()), (size_t)16});
        std::vector<std::thread> pool;
        for (size_t i = 1; i < n_threads; i++) pool.emplace_back(worker);
        worker();
        for (std::thread &th : pool) th.join();
        for (const PlyJob &job : jobs)  // the first failure in file order, as a sequential load would report it
            if (job.status != RT_OK) fail(job.status, job.error);
        std::vector<TriRecord> merged;
        std::vector<uint16_t> merged_materials;
        size_t total = scene.triangles.size();
        for (const PlyJob &job : jobs) total += job.triangles.size();
        merged.reserve(total);
        merged_materials.reserve(total);
        size_t inline_done = 0;
        auto take_inline = [&](size_t upto) {
            merged.insert(merged.end(), scene.triangles.begin() + inline_done, scene.triangles.begin() + upto);
            merged_materials.insert(merged_materials.end(), triangle_materials.begin() + inline_done, triangle_materials.begin() + upto);
            inline_done = upto;
        };
        for (const PlyJob &job : jobs) {
            take_inline(job.at);
            merged.insert(merged.end(), job.triangles.begin(), job.triangles.end());
            merged_materials.insert(merged_materials.end(), job.triangles.size(), job.material);
        }
        take_inline(scene.triangles.size());
        scene.triangles.swap(merged);
        triangle_materials.swap(merged_materials);
    }

    scene.material_indices = sphere_materials;
    scene.material_indices.insert(scene.material_indices.end(), triangle_materials.begin(), triangle_materials.end());
    if (scene.env.empty()) {  // the reference would dereference a null map; give misses a black sky instead
        scene.env = {0, 0, 0};
        scene.env_w = scene.env_h = 1;
    }
    scene.precompute_camera();
    if (use_bvh == 2) scene.defer_tree(verbose);
    else scene.build_tree(use_bvh ? 30 : 0, verbose, use_bvh == 3);  // MAX_BVH_DEPTH, scene.cu:10,820
    scene.compute_bounds();
}

}  // namespace rt
