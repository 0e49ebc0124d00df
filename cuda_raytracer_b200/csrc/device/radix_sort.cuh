// Hand-written stable LSD radix sort of (key u32, value u32) pairs, 8 bits per digit pass. Replaces the reference's
// cub::DeviceRadixSort::SortPairs (raytracing.cu:194-195,240-246) as the ray-reordering pass.
//
// Structure ("one sweep per digit"):
//   1. sort_histogram_kernel: ONE read of the keys builds the global histograms of every digit position.
//   2. its last CTA scans each 256-bin histogram into exclusive offsets (no separate launch).
//   3. sort_scatter_kernel, once per digit: each CTA takes the next tile (dynamic tile id), ranks its keys stably
//      (warp match + per-warp counters in shared memory), obtains its global offset per digit by decoupled look-back
//      over the tile-status array, and scatters keys and values. 8 B read + 8 B written per pair per digit.
//
// The element count lives in DEVICE memory (it is the live-ray count produced by the previous bounce kernel), so the
// host launches for an upper bound and surplus CTAs retire immediately: no host synchronisation inside a pass.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace rtd {

constexpr int kSortBlock = 256;
constexpr int kSortItems = 22;                          // keys per thread (RT_SORT_ITEMS selects among the instantiated 12 / 16 / 18 / 22 / 26)
constexpr int kSortTile = kSortBlock * 12;              // SMALLEST tile among the variants: sizes the tile-status scratch
constexpr int kSortWarps = kSortBlock / 32;
constexpr int kSortMaxPasses = 4;
constexpr uint32_t kFlagAggregate = 1u << 30;           // tile status: local count published
constexpr uint32_t kFlagPrefix = 2u << 30;              // tile status: inclusive prefix published
constexpr uint32_t kFlagMask = 3u << 30;
constexpr uint32_t kValueMask = ~kFlagMask;

struct SortArgs {
    const uint32_t *count_ptr;   // device: number of pairs
    uint32_t *hist;              // [kSortMaxPasses][256] global digit histograms -> exclusive offsets after the scan
    uint32_t *tile_status;       // [passes][max_tiles][256]
    uint32_t *tile_counter;      // [passes] dynamic tile ids
    int passes;
    uint32_t max_tiles;
};

// One read of the keys -> global histograms of every digit position; the LAST CTA to finish turns them into exclusive
// offsets (one launch less per sort than a separate scan kernel). s.tile_counter[kSortMaxPasses] counts the finished CTAs.
__global__ void __launch_bounds__(kSortBlock) sort_histogram_kernel(const uint32_t *__restrict__ keys, SortArgs s)
{
    __shared__ uint32_t h[kSortMaxPasses][256];
    __shared__ uint32_t warp_sum[8];
    __shared__ bool last;
    for (int p = 0; p < s.passes; p++) h[p][threadIdx.x] = 0;
    __syncthreads();
    const uint32_t n = *s.count_ptr;
    for (uint32_t i = blockIdx.x * kSortBlock + threadIdx.x; i < n; i += gridDim.x * kSortBlock) {
        uint32_t k = keys[i];
        for (int p = 0; p < s.passes; p++) atomicAdd(&h[p][(k >> (8 * p)) & 255u], 1u);
    }
    __syncthreads();
    for (int p = 0; p < s.passes; p++) {
        uint32_t v = h[p][threadIdx.x];
        if (v) atomicAdd(&s.hist[p * 256 + threadIdx.x], v);
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) last = atomicAdd(&s.tile_counter[kSortMaxPasses], 1u) == gridDim.x - 1;
    __syncthreads();
    if (!last) return;
    __threadfence();
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    for (int p = 0; p < s.passes; p++) {  // exclusive scan of each 256-bin histogram, in place
        volatile uint32_t *hp = s.hist + p * 256;
        const uint32_t v = hp[threadIdx.x];
        uint32_t incl = v;
        for (int off = 1; off < 32; off <<= 1) {
            uint32_t t = __shfl_up_sync(0xffffffffu, incl, off);
            if (lane >= (uint32_t)off) incl += t;
        }
        if (lane == 31) warp_sum[warp] = incl;
        __syncthreads();
        uint32_t base = 0;
        for (uint32_t w = 0; w < warp; w++) base += warp_sum[w];
        hp[threadIdx.x] = base + incl - v;
        __syncthreads();
    }
}

// One digit pass. values_in == nullptr means "value = element index" (first pass of a fresh permutation).
// The tile is ranked, re-ordered by digit in shared memory and only then written out, so that consecutive threads
// store consecutive addresses of a digit's run (a direct scatter writes 32 different sectors per warp instruction).
//
// RANK selects how a warp ranks its keys inside a round of 32 (A/B in tools/sort_probe.py, RT_SORT_RANK):
//   0  match.any; the leader lane reads, bumps and writes back the warp's digit counter (round 1's kernel)
//   1  match.any; the leader lane bumps the counter with ONE returning shared-memory atomic: no load -> store chain, the 16
//      rounds only meet in memory (shared atomics of one warp apply in program order; __syncwarp orders the lanes)
//   2  no match.any: the peer mask of a lane is built from 8 ballots (one per digit bit), counters as in 1. DEFAULT: MATCH.ANY
//      is the slow instruction of variants 0 / 1 on sm_100a (64 M pairs x 32 bits: 2.58 / 2.55 / 2.32 ms for 0 / 1 / 2 at 18 keys
//      per thread; CUB's onesweep 2.33 ms; profiles/r2_sort_probe.md)
// ITEMS = keys per thread (tile = 256 x ITEMS).
template <int RANK, int ITEMS>
__global__ void __launch_bounds__(kSortBlock, 3) sort_scatter_kernel(const uint32_t *__restrict__ keys_in, const uint32_t *__restrict__ values_in,
                                                                   uint32_t *__restrict__ keys_out, uint32_t *__restrict__ values_out,
                                                                   SortArgs s, int pass)
{
    constexpr int TILE = kSortBlock * ITEMS;
    __shared__ uint32_t warp_hist[kSortWarps][256];  // per-warp digit counts -> exclusive offsets of the warp inside the tile's digit run
    __shared__ uint32_t digit_base[256];             // global position of the tile's first key of each digit MINUS its position in the tile
    __shared__ uint32_t digit_start[256];            // position of that key inside the digit-ordered tile
    __shared__ uint32_t warp_total[kSortWarps];
    extern __shared__ uint32_t sort_stage[];         // dynamic (tiles above 18 keys per thread pass the 48 KB static limit): keys, then values
    uint32_t *stage_key = sort_stage, *stage_val = sort_stage + TILE;
    __shared__ uint32_t tile_id_sh;

    const uint32_t n = *s.count_ptr;
    const uint32_t num_tiles = (n + TILE - 1) / TILE;
    if (threadIdx.x == 0) tile_id_sh = atomicAdd(&s.tile_counter[pass], 1u);
    for (int w = 0; w < kSortWarps; w++) warp_hist[w][threadIdx.x] = 0;
    __syncthreads();
    const uint32_t tile = tile_id_sh;
    if (tile >= num_tiles) return;

    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const uint32_t shift = 8u * pass;
    const uint32_t tile_base = tile * TILE;
    const uint32_t warp_base = tile_base + warp * (ITEMS * 32u);
    const uint32_t tile_n = min((uint32_t)TILE, n - tile_base);
    const uint32_t lt = (1u << lane) - 1u;
    const bool full = tile_n == (uint32_t)TILE;

    uint32_t key[ITEMS], val[ITEMS];
    uint16_t rank[ITEMS];
#pragma unroll
    for (int r = 0; r < ITEMS; r++) {
        uint32_t i = warp_base + r * 32u + lane;
        bool valid = full || i < n;
        key[r] = valid ? __ldcs(keys_in + i) : 0xFFFFFFFFu;
        val[r] = valid ? (values_in ? __ldcs(values_in + i) : i) : 0u;
    }
    // stable ranking inside the warp's segment: rounds in memory order, lanes in order within a round
    uint32_t *my_hist = &warp_hist[warp][0];
#pragma unroll
    for (int r = 0; r < ITEMS; r++) {
        const uint32_t i = warp_base + r * 32u + lane;
        const bool valid = full || i < n;
        const uint32_t digit = (key[r] >> shift) & 255u;
        uint32_t peers;
        if (RANK == 2) {
            peers = full ? 0xffffffffu : __ballot_sync(0xffffffffu, valid);
#pragma unroll
            for (int b = 0; b < 8; b++) {
                const uint32_t m = __ballot_sync(0xffffffffu, (digit >> b) & 1u);
                peers &= ((digit >> b) & 1u) ? m : ~m;
            }
            if (!valid) peers = 1u << lane;
        } else {
            peers = __match_any_sync(0xffffffffu, valid ? digit : (256u + lane));
        }
        const uint32_t leader = __ffs(peers) - 1;
        uint32_t before = 0;
        if (RANK == 0) {
            if (valid && lane == leader) {
                before = my_hist[digit];
                my_hist[digit] = before + __popc(peers);
            }
        } else {
            if (valid && lane == leader) before = atomicAdd(&my_hist[digit], (uint32_t)__popc(peers));
        }
        before = __shfl_sync(0xffffffffu, before, leader);
        rank[r] = (uint16_t)(before + __popc(peers & lt));
        __syncwarp();
    }
    __syncthreads();

    // thread d owns digit d: exclusive scan over the 8 warps, tile total, then decoupled look-back
    const uint32_t dgt = threadIdx.x;
    uint32_t running = 0;
#pragma unroll
    for (int w = 0; w < kSortWarps; w++) {
        uint32_t c = warp_hist[w][dgt];
        warp_hist[w][dgt] = running;
        running += c;
    }
    volatile uint32_t *status = s.tile_status + ((size_t)pass * s.max_tiles) * 256u;
    uint32_t exclusive = 0;
    if (tile == 0) {
        status[dgt] = running | kFlagPrefix;
    } else {
        status[(size_t)tile * 256u + dgt] = running | kFlagAggregate;
    }
    // exclusive scan of the tile's digit totals over the 256 digits -> start of each digit's run inside the tile
    // (before the look-back: its latency hides behind the wait for the predecessors)
    uint32_t incl = running;
    for (int off = 1; off < 32; off <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, incl, off);
        if (lane >= (uint32_t)off) incl += t;
    }
    if (lane == 31) warp_total[warp] = incl;
    __syncthreads();
    uint32_t before_warps = 0;
    for (uint32_t w = 0; w < warp; w++) before_warps += warp_total[w];
    const uint32_t start = before_warps + incl - running;
    digit_start[dgt] = start;
    if (tile != 0) {
        for (int t = (int)tile - 1; t >= 0; t--) {
            uint32_t st;
            do { st = status[(size_t)t * 256u + dgt]; } while ((st & kFlagMask) == 0u);
            exclusive += st & kValueMask;
            if (st & kFlagPrefix) break;
        }
        status[(size_t)tile * 256u + dgt] = (exclusive + running) | kFlagPrefix;
    }
    digit_base[dgt] = s.hist[pass * 256 + dgt] + exclusive - start;
    __syncthreads();

    // re-order the tile by digit in shared memory (stable), then stream it out run by run
#pragma unroll
    for (int r = 0; r < ITEMS; r++) {
        uint32_t i = warp_base + r * 32u + lane;
        if (full || i < n) {
            uint32_t digit = (key[r] >> shift) & 255u;
            uint32_t local = digit_start[digit] + my_hist[digit] + rank[r];
            stage_key[local] = key[r];
            stage_val[local] = val[r];
        }
    }
    __syncthreads();
    for (uint32_t j = threadIdx.x; j < tile_n; j += kSortBlock) {
        const uint32_t k = stage_key[j];
        const uint32_t pos = digit_base[(k >> shift) & 255u] + j;
        keys_out[pos] = k;
        values_out[pos] = stage_val[j];
    }
}

}  // namespace rtd
