// radix_sort.cuh — onesweep-style LSD radix sort of (cell key, particle index) pairs.
//
// Replaces the reference's bitonic network over whole 32-byte records
// (sort.wgsl:27-51 dispatched S(S+1)/2 times, src/simulation.rs:323-347,518-524).
// The sorted KEY array is identical to the reference's (any correct sort gives sort(keys));
// the payload order within equal keys is the stable one (previous order), which is
// deterministic run to run (SURVEY.md §7 hard part 2).
//
// One histogram pass (fused into the hash kernel, kernels_hash.cuh) + one `onesweep_pass`
// launch per 8-bit digit. Each pass: a CTA takes the next tile (atomic ticket), ranks its keys with
// warp-level match_any, publishes its per-digit counts, resolves its global offsets by decoupled
// look-back over earlier tiles' status words, and scatters through shared memory so that global
// stores are contiguous per digit.
#pragma once
#include "common.cuh"

namespace fluid {

constexpr uint32_t STATUS_FLAG_AGG = 1u << 30;   // tile-local count available
constexpr uint32_t STATUS_FLAG_INCL = 2u << 30;  // inclusive prefix over tiles [0..t] available
constexpr uint32_t STATUS_VALUE_MASK = (1u << 30) - 1u;

// exclusive scan of the P x 256 histograms (one CTA of 256 threads per pass)
__global__ void __launch_bounds__(RADIX) k_scan_histograms(uint32_t* __restrict__ hist) {
    __shared__ uint32_t warp_sums[RADIX / 32];
    uint32_t* hp = hist + blockIdx.x * RADIX;
    uint32_t v = hp[threadIdx.x];
    uint32_t incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane_id() >= (uint32_t)o) incl += t;
    }
    if (lane_id() == 31) warp_sums[threadIdx.x >> 5] = incl;
    __syncthreads();
    uint32_t base = 0;
    for (uint32_t w = 0; w < (threadIdx.x >> 5); w++) base += warp_sums[w];
    hp[threadIdx.x] = base + incl - v;
}

// tuning (scripts/sweep_force.sh, 64M-U, ms for the 3 passes): batch 1/4/8/16 at 4 CTAs per SM = 1.15 / 1.01 / 1.04 / 1.11;
// unbounded registers (1 CTA per SM) = 2.2; 5 or 6 CTAs per SM or 8 / 12 / 24 keys per thread are all slower
#ifndef FLUID_LOOKBACK_BATCH
#define FLUID_LOOKBACK_BATCH 4
#endif
#ifndef FLUID_SORT_MIN_CTAS
#define FLUID_SORT_MIN_CTAS 4
#endif
constexpr int LOOKBACK_BATCH = FLUID_LOOKBACK_BATCH;

template <int ITEMS, bool IDENTITY_VALUES>
__global__ void __launch_bounds__(SORT_THREADS, FLUID_SORT_MIN_CTAS)
k_onesweep_pass(const uint32_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                uint32_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out,
                const uint32_t* __restrict__ digit_base,  // [256] exclusive-scanned global histogram of this digit
                uint32_t* __restrict__ tile_counter, volatile uint32_t* __restrict__ tile_status,  // [tiles][256]
                uint32_t n, int shift) {
    constexpr int TILE = SORT_THREADS * ITEMS;
    extern __shared__ __align__(16) uint32_t s_dyn[];  // 2*TILE words
    uint32_t* s_keys = s_dyn;
    uint32_t* s_vals = s_dyn + TILE;
    __shared__ uint32_t s_whist[SORT_WARPS][RADIX];  // per-warp digit counts -> exclusive prefix over warps
    __shared__ uint32_t s_tile_base[RADIX];          // exclusive scan over digits of the tile counts
    __shared__ uint32_t s_goff[RADIX];               // global offset for digit d minus its tile base
    __shared__ uint32_t s_scan[SORT_WARPS];
    __shared__ uint32_t s_tile;

    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    if (tid == 0) s_tile = atomicAdd(tile_counter, 1u);
#pragma unroll
    for (int w = 0; w < SORT_WARPS; w++) s_whist[w][tid] = 0;
    __syncthreads();
    const uint32_t tile = s_tile;
    const uint32_t tile_start = tile * (uint32_t)TILE;
    const uint32_t n_valid = min((uint32_t)TILE, n - tile_start);

    // ---- load (warp-striped: coalesced, and index order == (warp, item, lane) order) ----
    uint32_t key[ITEMS], val[ITEMS];
    uint16_t pos[ITEMS];
    const uint32_t warp_base = tile_start + warp * (32u * ITEMS) + lane;
#pragma unroll
    for (int k = 0; k < ITEMS; k++) {
        uint32_t idx = warp_base + k * 32u;
        bool ok = idx < n;
        key[k] = ok ? __ldg(keys_in + idx) : 0xffffffffu;
        if (IDENTITY_VALUES) val[k] = idx;
        else val[k] = ok ? __ldg(vals_in + idx) : 0u;
    }

    // ---- rank within the warp: match_any on the digit, leader bumps the warp's counter ----
    const uint32_t lt = lanemask_lt();
#pragma unroll
    for (int k = 0; k < ITEMS; k++) {
        uint32_t d = (key[k] >> shift) & (RADIX - 1);
        uint32_t peers = __match_any_sync(0xffffffffu, d);
        int leader = __ffs(peers) - 1;
        uint32_t old = 0;
        if ((int)lane == leader) {
            old = s_whist[warp][d];
            s_whist[warp][d] = old + __popc(peers);
        }
        old = __shfl_sync(0xffffffffu, old, leader);
        pos[k] = (uint16_t)(old + __popc(peers & lt));
        __syncwarp();
    }
    __syncthreads();

    // ---- thread d owns digit d: prefix over warps, tile count, publish aggregate ----
    uint32_t count = 0;
#pragma unroll
    for (int w = 0; w < SORT_WARPS; w++) {
        uint32_t t = s_whist[w][tid];
        s_whist[w][tid] = count;
        count += t;
    }
    if (tid == RADIX - 1) count -= ((uint32_t)TILE - n_valid);  // padding keys carry digit 0xff and rank last
    volatile uint32_t* my_status = tile_status + (size_t)tile * RADIX;
    my_status[tid] = (tile == 0 ? STATUS_FLAG_INCL : STATUS_FLAG_AGG) | count;

    // exclusive scan over digits (tile-local placement); padding re-included so positions stay dense
    uint32_t cnt_padded = count + ((tid == RADIX - 1) ? ((uint32_t)TILE - n_valid) : 0u);
    uint32_t incl = cnt_padded;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= (uint32_t)o) incl += t;
    }
    if (lane == 31) s_scan[warp] = incl;
    __syncthreads();
    uint32_t wbase = 0;
    for (uint32_t w = 0; w < warp; w++) wbase += s_scan[w];
    const uint32_t tile_base = wbase + incl - cnt_padded;
    s_tile_base[tid] = tile_base;
    __syncthreads();

    // ---- scatter into shared memory in tile-sorted order ----
#pragma unroll
    for (int k = 0; k < ITEMS; k++) {
        uint32_t d = (key[k] >> shift) & (RADIX - 1);
        uint32_t p = s_tile_base[d] + s_whist[warp][d] + pos[k];
        s_keys[p] = key[k];
        s_vals[p] = val[k];
    }

    // ---- decoupled look-back for digit `tid` (overlaps the smem scatter of the other warps) ----
    // Predecessor status words are fetched LOOKBACK_BATCH at a time (independent loads, one L2 round trip per batch)
    // and consumed in order: with ~600 resident tiles in near lock step the nearest tile holding an inclusive prefix
    // is typically tens of tiles back, and a one-load-at-a-time walk serialises that many ~1 us L2 round trips.
    uint32_t excl = 0;
    if (tile > 0) {
        int t = (int)tile - 1;
        bool done = false;
        while (!done) {
            uint32_t sw[LOOKBACK_BATCH];
#pragma unroll
            for (int j = 0; j < LOOKBACK_BATCH; j++)
                sw[j] = (t - j >= 0) ? tile_status[(size_t)(t - j) * RADIX + tid] : 0x80000000u;  // before tile 0: inclusive prefix 0
#pragma unroll
            for (int j = 0; j < LOOKBACK_BATCH; j++) {
                if (!done) {
                    uint32_t v = sw[j];
                    while ((v >> 30) == 0u) v = tile_status[(size_t)(t - j) * RADIX + tid];
                    excl += v & STATUS_VALUE_MASK;
                    if ((v >> 30) == 2u) done = true;
                }
            }
            t -= LOOKBACK_BATCH;
        }
        my_status[tid] = STATUS_FLAG_INCL | (excl + count);
    }
    s_goff[tid] = __ldg(digit_base + tid) + excl - tile_base;
    __syncthreads();

    // ---- coalesced-by-digit global stores ----
#pragma unroll
    for (int k = 0; k < ITEMS; k++) {
        uint32_t p = k * SORT_THREADS + tid;
        if (p < n_valid) {
            uint32_t kk = s_keys[p];
            uint32_t d = (kk >> shift) & (RADIX - 1);
            uint32_t dst = s_goff[d] + p;
            keys_out[dst] = kk;
            vals_out[dst] = s_vals[p];
        }
    }
}

}  // namespace fluid
