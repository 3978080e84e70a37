// kernels_hash.cuh — K1+K2 fused: predict_next_position + create_spatial_lookup
// (reference compute.wgsl:8-42, funcs.wgsl:206-218), plus the radix-sort digit histograms.
//
// One streaming pass: R position 8 B + velocity 8 B, W key 4 B per particle. The predicted position
// is NOT stored here: it is recomputed bit-identically in the gather pass from the same inputs, which
// saves 16 B/particle of HBM traffic against storing and re-reading it.
#pragma once
#include "common.cuh"
#include "radix_sort.cuh"

namespace fluid {

constexpr int HASH_THREADS = 256;

// warp-aggregated shared-memory histogram update. The array is in last tick's key order, so along the lanes of a warp
// the digits come in runs: the first lane of every run adds the run's length (one shuffle and one vote instead of a
// match_any, which kept the XU pipe at 39 % and led the stall samples in round 1). Any digit order is counted correctly;
// unsorted input just makes the runs short.
__device__ __forceinline__ void hist_add(uint32_t* s_hist, uint32_t digit, bool valid) {
    const uint32_t lane = lane_id();
    const uint32_t d = valid ? digit : 0xffffffffu;
    const uint32_t prev = __shfl_up_sync(0xffffffffu, d, 1);
    const uint32_t heads = __ballot_sync(0xffffffffu, lane == 0u || d != prev);
    if (valid && ((heads >> lane) & 1u)) {
        const uint32_t above = heads & ~((2u << lane) - 1u);  // run heads in the lanes above this one
        const uint32_t next = above ? (uint32_t)__ffs(above) - 1u : 32u;
        atomicAdd(s_hist + digit, next - lane);
    }
}

// the two particles of every lane at once, for the upper digits: the array is in last tick's key order, so a warp's 64
// keys nearly always share them — one vote and one atomic per warp; match_any only when they differ
__device__ __forceinline__ void hist_add_pair_upper(uint32_t* s_hist, uint32_t d0, uint32_t d1, bool v0, bool v1) {
    const uint32_t ref = __shfl_sync(0xffffffffu, d0, 0);  // lane 0's first particle is valid whenever any lane's is
    const bool same = (!v0 || d0 == ref) && (!v1 || d1 == ref);
    if (__all_sync(0xffffffffu, same)) {
        const uint32_t cnt = __popc(__ballot_sync(0xffffffffu, v0)) + __popc(__ballot_sync(0xffffffffu, v1));
        if (lane_id() == 0 && cnt) atomicAdd(s_hist + ref, cnt);
    } else {
        hist_add(s_hist, d0, v0);
        hist_add(s_hist, d1, v1);
    }
}

template <int PASSES>
__global__ void __launch_bounds__(HASH_THREADS)
k_predict_key_hist(const DevParams p, const float2* __restrict__ pos, const float2* __restrict__ vel,
                   uint32_t* __restrict__ keys, uint32_t* __restrict__ g_hist, uint32_t* __restrict__ frame_time,
                   uint32_t bump /* 0 when k_predict_key_count already counted this tick */) {
    __shared__ uint32_t s_hist[PASSES * RADIX];
    for (int i = threadIdx.x; i < PASSES * RADIX; i += HASH_THREADS) s_hist[i] = 0;
    __syncthreads();
    if (blockIdx.x == 0 && threadIdx.x == 0) *frame_time += bump;  // reference: self.tick += 1 (src/simulation.rs:460)

    // two particles per thread per iteration: float4 loads of pos and vel, uint2 store of the keys
    const uint32_t n_pairs = (p.n + 1u) >> 1;
    const float4* pos4 = reinterpret_cast<const float4*>(pos);
    const float4* vel4 = reinterpret_cast<const float4*>(vel);
    const uint32_t stride = gridDim.x * HASH_THREADS;
    const float y_h = rcp_refined(p.h);  // the key's two divisions by h share it (common.cuh: cell_key_shared)
    // loop bound rounded up to a warp multiple so that match_any always sees a full warp
    const uint32_t n_iter_end = (n_pairs + 31u) & ~31u;
    for (uint32_t i = blockIdx.x * HASH_THREADS + threadIdx.x; i < n_iter_end; i += stride) {
        const bool v0 = i < n_pairs;
        const bool v1 = v0 && (2u * i + 1u < p.n);
        uint32_t k0 = 0, k1 = 0;
        if (v0) {
            float4 ps = __ldg(pos4 + i), vs = __ldg(vel4 + i);  // arrays are padded to an even length
            const float2 q0 = predict_position(p, make_float2(ps.x, ps.y), make_float2(vs.x, vs.y));
            const float2 q1 = predict_position(p, make_float2(ps.z, ps.w), make_float2(vs.z, vs.w));
            if (p.key_fast_div) { k0 = cell_key_shared(p, q0, y_h); k1 = cell_key_shared(p, q1, y_h); }
            else { k0 = cell_key(p, q0); k1 = cell_key(p, q1); }
            if (v1) reinterpret_cast<uint2*>(keys)[i] = make_uint2(k0, k1);
            else keys[2u * i] = k0;
        }
#pragma unroll
        for (int q = 0; q < PASSES; q++) {
            const uint32_t d0 = (k0 >> (q * RADIX_BITS)) & (RADIX - 1), d1 = (k1 >> (q * RADIX_BITS)) & (RADIX - 1);
            if (q == 0) {
                hist_add(s_hist, d0, v0);
                hist_add(s_hist, d1, v1);
            } else {
                hist_add_pair_upper(s_hist + q * RADIX, d0, d1, v0, v1);
            }
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < PASSES * RADIX; i += HASH_THREADS) {
        uint32_t c = s_hist[i];
        if (c) atomicAdd(g_hist + i, c);
    }
}

}  // namespace fluid
