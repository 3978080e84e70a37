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

// warp-aggregated shared-memory histogram update: neighbouring particles are in neighbouring cells
// (the array is in last tick's key order), so lanes collide heavily on the upper digits.
__device__ __forceinline__ void hist_add(uint32_t* s_hist, uint32_t digit, bool valid) {
    uint32_t d = valid ? digit : 0xffffffffu;
    uint32_t peers = __match_any_sync(0xffffffffu, d);
    if (valid && (int)lane_id() == __ffs(peers) - 1) atomicAdd(s_hist + digit, (uint32_t)__popc(peers));
}

template <int PASSES>
__global__ void __launch_bounds__(HASH_THREADS)
k_predict_key_hist(const DevParams p, const float2* __restrict__ pos, const float2* __restrict__ vel,
                   uint32_t* __restrict__ keys, uint32_t* __restrict__ g_hist, uint32_t* __restrict__ frame_time) {
    __shared__ uint32_t s_hist[PASSES * RADIX];
    for (int i = threadIdx.x; i < PASSES * RADIX; i += HASH_THREADS) s_hist[i] = 0;
    __syncthreads();
    if (blockIdx.x == 0 && threadIdx.x == 0) *frame_time += 1u;  // reference: self.tick += 1 (src/simulation.rs:460)

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
            hist_add(s_hist + q * RADIX, (k0 >> (q * RADIX_BITS)) & (RADIX - 1), v0);
            hist_add(s_hist + q * RADIX, (k1 >> (q * RADIX_BITS)) & (RADIX - 1), v1);
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < PASSES * RADIX; i += HASH_THREADS) {
        uint32_t c = s_hist[i];
        if (c) atomicAdd(g_hist + i, c);
    }
}

}  // namespace fluid
