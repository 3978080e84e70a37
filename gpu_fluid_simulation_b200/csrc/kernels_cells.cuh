// kernels_cells.cuh — gather into key order + cell tables.
//
// Reference: after sort.wgsl has physically permuted the 32-byte records, compute_start_indices
// (compute.wgsl:45-56) writes start_indices[key[i]] = i wherever key[i] != key[i-1], for i >= 1 only,
// into a table that is never cleared (src/simulation.rs:204-209).
//
// Here the sort moved (key, index) pairs only; this pass moves the payload once (R idx 4 + pos 8 +
// vel 8, W pos 8 + pred 8 + vel 8) and, in the same pass, builds two tables:
//   * start_indices  — the reference-compatible table, same write rule, never cleared;
//   * cell_lb        — a dense, fully rewritten lower-bound table over all grid_w*grid_h+1 cells:
//                      cell_lb[c] = first sorted index whose key is >= c, so the particles of the
//                      cell range [a, b] are the contiguous span [cell_lb[a], cell_lb[b+1]).
// The reference quirk that the lowest occupied cell c0 never gets its entry written (i == 0 returns
// early) and so is walked from whatever start_indices[c0] held before, is reproduced by publishing
// quirk = {c0, S} (S = that stale entry): consumers raise the start of any span that contains c0 to
// min(S, end of span) (kernels_sph.cuh thread_spans). cell_lb itself stays clean and monotone.
#pragma once
#include "common.cuh"

namespace fluid {

constexpr int GATHER_THREADS = 256;

// every lane passes (begin, end, value); the warp fills lb[begin..end) = value for each lane in turn
__device__ __forceinline__ void warp_fill_gaps(uint32_t* __restrict__ lb, uint32_t begin, uint32_t end, uint32_t value) {
    // short gaps (the common case: the next occupied cell is 1-2 cells away) are written by the owner
    uint32_t len = end > begin ? end - begin : 0u;
    if (len <= 4u) {
        for (uint32_t c = begin; c < end; c++) lb[c] = value;
        len = 0;
    }
    uint32_t pending = __ballot_sync(0xffffffffu, len > 0u);
    while (pending) {
        int src = __ffs(pending) - 1;
        uint32_t b = __shfl_sync(0xffffffffu, begin, src);
        uint32_t e = __shfl_sync(0xffffffffu, end, src);
        uint32_t v = __shfl_sync(0xffffffffu, value, src);
        for (uint32_t c = b + lane_id(); c < e; c += 32u) lb[c] = v;
        pending &= pending - 1u;
    }
}

__global__ void __launch_bounds__(GATHER_THREADS)
k_gather_cells(const DevParams p, const uint32_t* __restrict__ sorted_keys, const uint32_t* __restrict__ sorted_idx,
               const float2* __restrict__ pos_in, const float2* __restrict__ vel_in,
               float2* __restrict__ pos_s, float2* __restrict__ pred_s, float2* __restrict__ vel_s,
               uint32_t* __restrict__ start_indices, uint32_t* __restrict__ cell_lb, uint32_t* __restrict__ quirk,
               const uint32_t* __restrict__ frame_time) {
    const uint32_t i = blockIdx.x * GATHER_THREADS + threadIdx.x;  // grid is a whole number of warps
    const bool valid = i < p.n;
    uint32_t gap_begin = 0, gap_end = 0, gap_value = 0;
    if (valid) {
        const uint32_t src = __ldg(sorted_idx + i);
        const uint32_t k = __ldg(sorted_keys + i);
        const float2 ps = __ldg(pos_in + src);
        const float2 vs = __ldg(vel_in + src);
        pos_s[i] = ps;
        vel_s[i] = vs;
        const float2 pr = predict_position(p, ps, vs);  // same inputs, same ops as the hash pass: same bits
        pred_s[i] = pr;
        // quirk[2] = wild stamp (common.cuh tame_component): any writer stores the same tick number
        if (!(tame_component(pr.x) && tame_component(pr.y) && tame_component(vs.x) && tame_component(vs.y)))
            quirk[2] = __ldg(frame_time);

        const uint32_t kc = min(k, p.num_cells - 1u);
        if (i == 0) {
            gap_begin = 0; gap_end = kc + 1u; gap_value = 0;
            quirk[0] = kc;
            quirk[1] = start_indices[kc];  // stale entry: nobody writes start_indices[key[0]] this tick
        } else {
            const uint32_t kp = min(__ldg(sorted_keys + i - 1), p.num_cells - 1u);
            if (kc != kp) {
                start_indices[kc] = i;  // compute.wgsl:52-54
                gap_begin = kp + 1u; gap_end = kc + 1u; gap_value = i;
            }
        }
    }
    warp_fill_gaps(cell_lb, gap_begin, gap_end, gap_value);
    // the last particle also closes the table: every cell above its key starts at n
    uint32_t tb = 0, te = 0;
    if (valid && i == p.n - 1u) {
        tb = min(__ldg(sorted_keys + i), p.num_cells - 1u) + 1u;
        te = p.num_cells + 1u;
    }
    if (__any_sync(0xffffffffu, te > tb)) warp_fill_gaps(cell_lb, tb, te, p.n);
}

}  // namespace fluid
