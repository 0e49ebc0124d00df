// The host-side planning of a render call (no CUDA in here): which rays of which passes the call traces (SEGMENTS) and how
// they are grouped into WAVES. Pure functions of the call's options, so that the CPU test-suite can drive them through
// rt_plan_waves (tests/test_wave_plan.py) with thousands of option sets; render_impl (rt_core.cu) launches what they return.
#pragma once

#include <stdint.h>

#include <algorithm>
#include <vector>

#include "device_scene.cuh"

namespace rtd {

// Rays [offset, offset + count) of a pass; with sub_count > 0: samples [sub_first, sub_first + sub_count) of every pixel
// (count = pixels x sub_count). Seeds are functions of the pass and the ray's index inside it (raytracing.cu:89,229,236,
// scene.cu:81), so any such piece of a pass renders the rays the whole pass would.
struct PassSegment {
    int pass;
    uint32_t offset, count;
    uint32_t sub_first, sub_count;
};

struct FrameShape {
    size_t pixels;
    int ray_count;  // samples per pixel of the frame
    int pass_spp;   // MAX_RAYS_PER_PIXEL_PER_PASS, raytracing.cu:109
    int total_passes() const { return (ray_count + pass_spp - 1) / pass_spp; }
    int spp_of(int p) const { return std::min(ray_count - p * pass_spp, pass_spp); }  // the last pass may be shorter (:224-225)
    bool whole(const PassSegment &g) const { return g.sub_count == 0 && g.offset == 0 && (size_t)g.count == pixels * (size_t)spp_of(g.pass); }
};

// The work of a call: whole passes by default or by pass range; with sample_count > 0 samples [sample_first, +sample_count)
// of every pixel; with ray_count > 0 a range of the frame's ray sequence. The first and last pass of a range may be partial.
// Returns nullptr, or the message of the option that is out of range.
inline const char *call_segments(const FrameShape &f, int pass_first_opt, int pass_count_opt, int64_t ray_first, int64_t ray_count,
                                 int sample_first, int sample_count, std::vector<PassSegment> &segs)
{
    segs.clear();
    const int total_passes = f.total_passes();
    if (sample_count > 0) {  // pass p holds samples [p * pass_spp, p * pass_spp + spp_of(p))
        if (sample_first < 0) return "rt_render: sample_first must be >= 0";
        const int64_t s0 = sample_first, s1 = std::min<int64_t>((int64_t)sample_first + sample_count, f.ray_count);
        for (int p = 0; p < total_passes; p++) {
            const int64_t lo = std::max<int64_t>(s0, (int64_t)p * f.pass_spp), hi = std::min<int64_t>(s1, (int64_t)p * f.pass_spp + f.spp_of(p));
            if (hi <= lo) continue;
            const uint32_t k0 = (uint32_t)(lo - (int64_t)p * f.pass_spp), m = (uint32_t)(hi - lo);
            if ((int)m == f.spp_of(p)) segs.push_back(PassSegment{p, 0u, (uint32_t)(f.pixels * (size_t)m), 0u, 0u});
            else segs.push_back(PassSegment{p, 0u, (uint32_t)(f.pixels * (size_t)m), k0, m});
        }
    } else if (ray_count > 0) {
        if (ray_first < 0) return "rt_render: ray_first must be >= 0";
        int64_t start = 0;
        const int64_t u0 = ray_first, u1 = ray_first + ray_count;
        for (int p = 0; p < total_passes && start < u1; p++) {
            const int64_t n = (int64_t)f.pixels * f.spp_of(p);
            const int64_t lo = std::max(u0, start), hi = std::min(u1, start + n);
            if (hi > lo) segs.push_back(PassSegment{p, (uint32_t)(lo - start), (uint32_t)(hi - lo), 0u, 0u});
            start += n;
        }
    } else {
        const int pass_first = std::max(0, pass_first_opt);
        const int pass_end = pass_count_opt > 0 ? std::min(total_passes, pass_first + pass_count_opt) : total_passes;
        for (int p = pass_first; p < pass_end; p++) segs.push_back(PassSegment{p, 0u, (uint32_t)(f.pixels * (size_t)f.spp_of(p)), 0u, 0u});
    }
    return nullptr;
}

// A WAVE is up to max_group passes of equal size traced together: every pass keeps the reference's seeds (they depend on
// the pass, not on the launch), while the late bounces - a few 10^4 rays per pass, latency-bound - and the launch gaps are
// paid once per wave instead of once per pass. member[partial_member] (if >= 0) is a partial pass; it is the LAST member
// (the member index of a ray is id / pass_rays, which needs full-size members in front) or the only one.
struct WavePlan {
    PassSegment member[kMaxPassGroup];
    int group, partial_member;
};

// Whole passes first, partial passes (the ends of a ray / sample range) after them: a partial pass rides along as the last
// member of a wave of whole passes with the same samples per pixel. On its own it would be a whole chain of latency-bound
// launches for a fraction of a pass: a quarter pass as its own wave cost teapot's 8-GPU share 2 ms of 19
// (profiles/r2f_partial.md).
// A run of equal passes that needs MORE waves than there are pipes is dealt out evenly (50 passes with room for 16: 13 + 13 +
// 12 + 12, not 16 + 16 + 16 + 2; wave count a multiple of the pipes where that fits), so that the pipes finish together: teapot
// 1000x1000 whole frame 130.3 -> 128.7 ms, lamp 1000x1000 240.4 -> 238.2. A run that fits ONE round of waves keeps the
// first-fit plan (6 passes + a partial one on three pipes: (2 + partial) + 3 + 1): there the unequal waves help, because the
// short waves' chains of late bounces run while the long wave still has bulk work (teapot's share of 8: 17.09 ms first-fit,
// 17.44 ms as 2 + 2 + (2 + partial); profiles/r2g_waves.md).
// Reorders segs (whole passes first, stable); max_group in [1, kMaxPassGroup], pipes >= 1.
inline std::vector<WavePlan> plan_waves(std::vector<PassSegment> &segs, const FrameShape &f, int max_group, int pipes)
{
    const size_t pixels = f.pixels;
    const int n_segs = (int)segs.size();
    auto whole = [&](const PassSegment &g) { return f.whole(g); };
    auto spp_of = [&](int p) { return f.spp_of(p); };
    std::stable_partition(segs.begin(), segs.end(), whole);
    const int n_whole = (int)(std::find_if(segs.begin(), segs.end(), [&](const PassSegment &g) { return !whole(g); }) - segs.begin());
    std::vector<WavePlan> plan;
    std::vector<char> rides(n_segs, 0);  // partial passes that found a wave
    for (int wi = 0; wi < n_whole;) {
        const int spp0 = spp_of(segs[wi].pass);
        int run = 0;
        while (wi + run < n_whole && spp_of(segs[wi + run].pass) == spp0) run++;
        std::vector<int> riders;  // partial passes that can ride with this run, largest first
        for (int j = n_whole; j < n_segs; j++)
            if (!rides[j] && spp_of(segs[j].pass) == spp0 && (size_t)segs[j].count < pixels * (size_t)spp0) riders.push_back(j);
        std::stable_sort(riders.begin(), riders.end(), [&](int x, int y) { return segs[x].count > segs[y].count; });
        const int P0 = max_group >= 2 ? std::min<int>((int)riders.size(), run) : 0;
        if ((run + P0 + max_group - 1) / max_group <= pipes) {
            // one round: first fit, in queue order; a slot is kept for a rider while one is waiting
            std::stable_sort(riders.begin(), riders.end());
            size_t next_rider = 0;
            for (int left = run; left > 0;) {
                WavePlan wp{};
                const int room = std::max(1, max_group - (next_rider < riders.size() ? 1 : 0));
                while (wp.group < room && left > 0) { wp.member[wp.group++] = segs[wi++]; left--; }
                wp.partial_member = -1;
                if (next_rider < riders.size() && wp.group < max_group) {
                    const int j = riders[next_rider++];
                    rides[j] = 1;
                    wp.partial_member = wp.group;
                    wp.member[wp.group++] = segs[j];
                }
                plan.push_back(wp);
            }
            continue;
        }
        // W waves: wave k holds run / W whole passes (+ 1 for the first run % W waves); the last P waves carry one rider each
        auto fits = [&](int W, int P) {
            if (W < 1 || W > run || P > W) return false;
            for (int k = 0; k < W; k++)
                if (run / W + (k < run % W ? 1 : 0) + (k >= W - P ? 1 : 0) > max_group) return false;
            return true;
        };
        int W = run, P = 0;
        for (int p = P0; p >= 0; p--) {  // drop riders that do not fit
            int first = 0;
            for (int w = 1; w <= run && !first; w++) if (fits(w, p)) first = w;
            if (!first) continue;
            W = first;
            P = p;
            const int up = (W + pipes - 1) / pipes * pipes;
            if (W > 1 && up != W && fits(up, p)) W = up;
            break;
        }
        for (int k = 0; k < W; k++) {
            WavePlan wp{};
            const int n = run / W + (k < run % W ? 1 : 0);
            for (int m = 0; m < n; m++) wp.member[wp.group++] = segs[wi++];
            wp.partial_member = -1;
            if (k >= W - P) {  // the last wave (the smallest) takes the largest rider
                const int j = riders[W - 1 - k];
                rides[j] = 1;
                wp.partial_member = wp.group;
                wp.member[wp.group++] = segs[j];
            }
            plan.push_back(wp);
        }
    }
    for (int j = n_whole; j < n_segs; j++)  // partial passes without company: each is a wave of its own
        if (!rides[j]) {
            WavePlan wp{};
            wp.member[wp.group++] = segs[j];
            wp.partial_member = 0;
            plan.push_back(wp);
        }
    return plan;
}

}  // namespace rtd
