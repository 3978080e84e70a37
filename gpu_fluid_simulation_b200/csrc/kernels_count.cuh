// kernels_count.cuh — the key sort as a counting sort over the cells (single-GPU tick, default path).
//
// Replaces, like radix_sort.cuh, the reference's bitonic network (sort.wgsl:27-51 dispatched S(S+1)/2 times,
// src/simulation.rs:323-347,518-524) and feeds the same consumers: sorted keys, the source permutation, the dense
// lower-bound table `cell_lb` and the reference-compatible `start_indices` (compute.wgsl:45-56).
//
// A key IS a cell id (funcs.wgsl:216-218), there are ~4 particles per cell and `cell_lb` — the exclusive prefix sum of
// the per-cell particle counts — has to be built anyway. So instead of three 8-bit radix passes over (key, index):
//   1. k_predict_key_count   predict + key (K1+K2, compute.wgsl:8-42) and count[key] += 1           (one RED per run)
//                            — in steady state not launched at all: the previous tick's force pass has emitted the
//                            keys and the counts (kernels_sph.cuh, "speculation"; fluid_sim.cu enqueue_tick decides)
//   2. k_cell_reduce / k_cell_downsweep   cell_lb = exclusive scan of count; cursor = cell_lb; cell statistics -> host;
//                            clears the count array of the next tick. Grids of <= 32 tiles: k_cell_scan_small, one launch
//   3. k_cell_slots          slot = cursor[key]++ : writes (key, source index) into the cell's block, in ARRIVAL order
//   4. k_rank_gather_cells   puts each cell's block into source-index order (rank = entries of the cell with a
//                            smaller index), which makes the result the STABLE sort — exactly the order the LSD radix
//                            passes produce, so the float sums downstream do not change a bit — moves the payload and
//                            makes the staging plans of the neighbour passes.
// The multi-GPU step (kernels_band.cuh) runs the same steps over a band's cells: k_band_key / k_band_recv count, the scan,
// k_cell_slots, k_band_rank_gather (in-cell order by origin instead of by index).
// Traffic per particle: 20 + 12 + ~56 B against 20 + 44 + 60 B for hash + 3 onesweep passes + gather.
// The atomics are cheap because the array enters a tick in the previous tick's key order: along the lanes of a warp
// the keys come in runs, the head of a run issues one atomic for the whole run, and the lanes of one instruction
// touch a handful of adjacent 32-byte sectors. Arbitrary input order (first tick after an upload) is still sorted
// correctly, only with shorter runs.
// Step 4 costs O(cell population) per particle. The host therefore watches the largest cell population (published
// into pinned host memory by step 2, read without a sync one or more ticks later, or after a sync on the first tick
// after an upload) and takes the radix passes instead while the cells are too crowded (criterion at COUNT_SORT_MAX_*).
#pragma once
#include "common.cuh"
#include "kernels_cells.cuh"
#include "kernels_hash.cuh"
#include "kernels_sph.cuh"

namespace fluid {

constexpr int CELL_SCAN_THREADS = 256;
constexpr int CELL_SCAN_ITEMS = 16;  // consecutive cells per thread (4 x uint4)
constexpr uint32_t CELL_SCAN_TILE = CELL_SCAN_THREADS * CELL_SCAN_ITEMS;
// the counting path is taken while  max population <= clamp(n / 1024, 64, 4096)  (the tail of one crowded cell:
// its threads walk the whole block)  and  sum of population^2 <= 128 n  (the total ranking work); measured on the
// late dam break (cells above 64, 1M particles): scan + slots + rank 0.044 ms against 0.075 ms for the radix passes
constexpr uint32_t COUNT_SORT_MAX_CELL_LO = 64, COUNT_SORT_MAX_CELL_HI = 4096;
constexpr float COUNT_SORT_MAX_MEAN_SQ = 128.0f;

__device__ __forceinline__ uint32_t count_bin(uint32_t key, uint32_t key_lo, uint32_t last_bin) {
    return min(key - min(key, key_lo), last_bin);
}

// count[cell] += 1 for every valid lane; the first lane of every run of equal cells adds the run's length
__device__ __forceinline__ void cell_count_add(uint32_t* __restrict__ count, uint32_t cell, bool valid) {
    atomic_count_run(count, cell, valid);
}

// K1+K2 as in k_predict_key_hist (same arithmetic, same bits), with the per-cell count in place of the digit histograms
__global__ void __launch_bounds__(HASH_THREADS)
k_predict_key_count(const DevParams p, const float2* __restrict__ pos, const float2* __restrict__ vel,
                    uint32_t* __restrict__ keys, uint32_t* __restrict__ count, uint32_t* __restrict__ frame_time) {
    pdl_enter();
    if (blockIdx.x == 0 && threadIdx.x == 0) *frame_time += 1u;  // reference: self.tick += 1 (src/simulation.rs:460)
    const uint32_t n_pairs = (p.n + 1u) >> 1;
    const float4* pos4 = reinterpret_cast<const float4*>(pos);
    const float4* vel4 = reinterpret_cast<const float4*>(vel);
    const uint32_t stride = gridDim.x * HASH_THREADS;
    const float y_h = rcp_refined(p.h);
    const uint32_t last_cell = p.num_cells - 1u;
    const uint32_t n_iter_end = (n_pairs + 31u) & ~31u;  // whole warps take part in the votes
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
        cell_count_add(count, min(k0, last_cell), v0);
        cell_count_add(count, min(k1, last_cell), v1);
    }
}

__device__ __forceinline__ uint32_t block_sum_256(uint32_t v, uint32_t* s_tmp /*8*/) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane_id() == 0) s_tmp[threadIdx.x >> 5] = v;
    __syncthreads();
    uint32_t t = 0;
#pragma unroll
    for (int w = 0; w < CELL_SCAN_THREADS / 32; w++) t += s_tmp[w];
    __syncthreads();
    return t;
}
__device__ __forceinline__ uint32_t block_max_256(uint32_t v, uint32_t* s_tmp /*8*/) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = max(v, __shfl_xor_sync(0xffffffffu, v, o));
    if (lane_id() == 0) s_tmp[threadIdx.x >> 5] = v;
    __syncthreads();
    uint32_t t = 0;
#pragma unroll
    for (int w = 0; w < CELL_SCAN_THREADS / 32; w++) t = max(t, s_tmp[w]);
    __syncthreads();
    return t;
}

// per tile of CELL_SCAN_TILE cells: particle count and largest cell (the count array is zero-padded to whole tiles)
__global__ void __launch_bounds__(CELL_SCAN_THREADS)
k_cell_reduce(const uint32_t* __restrict__ count, uint32_t* __restrict__ tile_sum, uint32_t* __restrict__ tile_max,
              float* __restrict__ tile_sq, uint32_t* __restrict__ frame_time, uint32_t bump) {
    pdl_enter();
    // first kernel of a tick whose keys and counts came from the previous tick's force pass: self.tick += 1 (src/simulation.rs:460)
    if (bump && blockIdx.x == 0 && threadIdx.x == 0) *frame_time += bump;
    __shared__ uint32_t s_tmp[CELL_SCAN_THREADS / 32];
    __shared__ float s_sq[CELL_SCAN_THREADS / 32];
    const uint4* src = reinterpret_cast<const uint4*>(count + (size_t)blockIdx.x * CELL_SCAN_TILE) + threadIdx.x * (CELL_SCAN_ITEMS / 4);
    uint32_t sum = 0, mx = 0;
    float sq = 0.0f;  // a work estimate, not a parity value: any summation order will do
#pragma unroll
    for (int j = 0; j < CELL_SCAN_ITEMS / 4; j++) {
        const uint4 v = __ldg(src + j);
        sum += v.x + v.y + v.z + v.w;
        mx = max(max(mx, max(v.x, v.y)), max(v.z, v.w));
        sq += (float)v.x * (float)v.x + (float)v.y * (float)v.y + (float)v.z * (float)v.z + (float)v.w * (float)v.w;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sq += __shfl_xor_sync(0xffffffffu, sq, o);
    if (lane_id() == 0) s_sq[threadIdx.x >> 5] = sq;
    sum = block_sum_256(sum, s_tmp);  // (its barriers also publish s_sq)
    mx = block_max_256(mx, s_tmp);
    if (threadIdx.x == 0) {
        float t = 0.0f;
#pragma unroll
        for (int w = 0; w < CELL_SCAN_THREADS / 32; w++) t += s_sq[w];
        tile_sum[blockIdx.x] = sum; tile_max[blockIdx.x] = mx; tile_sq[blockIdx.x] = t;
    }
}

// cell_lb[c] = particles in cells below c (exclusive scan), also left in the count array as the cell's write cursor.
// Entries past the last cell of the padded arrays receive n, which is what cell_lb[num_cells] has to hold.
// Tile 0 also publishes the largest cell population of the tick to the host (pinned, mapped memory).
// `zero_next` is the count array of the NEXT tick (the two alternate): clearing it here, tile by tile, replaces a
// per-tick memset, i.e. one stream operation between the force pass and the next tick's first kernel.
__global__ void __launch_bounds__(CELL_SCAN_THREADS)
k_cell_downsweep(uint32_t* __restrict__ count_cursor, uint32_t* __restrict__ cell_lb, const uint32_t* __restrict__ tile_sum,
                 const uint32_t* __restrict__ tile_max, const float* __restrict__ tile_sq, uint32_t ntiles,
                 volatile uint32_t* __restrict__ host_max /* [0] largest population, [1] sum of squares as float bits */,
                 uint32_t* __restrict__ zero_next) {
    pdl_enter();
    __shared__ uint32_t s_tmp[CELL_SCAN_THREADS / 32];
    __shared__ uint32_t s_warp[CELL_SCAN_THREADS / 32];
    uint32_t part = 0;
    for (uint32_t b = threadIdx.x; b < blockIdx.x; b += CELL_SCAN_THREADS) part += __ldg(tile_sum + b);
    const uint32_t tile_base = block_sum_256(part, s_tmp);
    if (blockIdx.x == 0) {
        uint32_t mx = 0;
        for (uint32_t b = threadIdx.x; b < ntiles; b += CELL_SCAN_THREADS) mx = max(mx, __ldg(tile_max + b));
        mx = block_max_256(mx, s_tmp);
        float sq = 0.0f;
        for (uint32_t b = threadIdx.x; b < ntiles; b += CELL_SCAN_THREADS) sq += __ldg(tile_sq + b);
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) sq += __shfl_xor_sync(0xffffffffu, sq, o);
        __shared__ float s_sq[CELL_SCAN_THREADS / 32];
        if (lane_id() == 0) s_sq[threadIdx.x >> 5] = sq;
        __syncthreads();
        if (threadIdx.x == 0) {
            float t = 0.0f;
#pragma unroll
            for (int w = 0; w < CELL_SCAN_THREADS / 32; w++) t += s_sq[w];
            host_max[0] = mx;
            host_max[1] = __float_as_uint(t);
        }
    }
    const size_t first = (size_t)blockIdx.x * CELL_SCAN_TILE + (size_t)threadIdx.x * CELL_SCAN_ITEMS;
    uint4* cc = reinterpret_cast<uint4*>(count_cursor + first);
    uint4* lb = reinterpret_cast<uint4*>(cell_lb + first);
    uint4* zn = reinterpret_cast<uint4*>(zero_next + first);
    uint4 v[CELL_SCAN_ITEMS / 4];
    uint32_t total = 0;
#pragma unroll
    for (int j = 0; j < CELL_SCAN_ITEMS / 4; j++) {
        v[j] = cc[j];
        total += v[j].x + v[j].y + v[j].z + v[j].w;
    }
    uint32_t incl = total;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane_id() >= (uint32_t)o) incl += t;
    }
    if (lane_id() == 31) s_warp[threadIdx.x >> 5] = incl;
    __syncthreads();
    uint32_t run = tile_base + incl - total;
    for (uint32_t w = 0; w < (threadIdx.x >> 5); w++) run += s_warp[w];
#pragma unroll
    for (int j = 0; j < CELL_SCAN_ITEMS / 4; j++) {
        uint4 o;
        o.x = run; run += v[j].x;
        o.y = run; run += v[j].y;
        o.z = run; run += v[j].z;
        o.w = run; run += v[j].w;
        cc[j] = o;
        lb[j] = o;
        zn[j] = make_uint4(0u, 0u, 0u, 0u);
    }
}

// Small grids (the reference's default scene has 71,289 cells = 18 tiles): the scan in ONE launch — at this size a kernel
// boundary (~4 us) costs more than the work. CTA b owns tile b (4096 cells, one uint4 per thread) and gets its base by
// summing the raw counts of the tiles before it itself (at most 31 coalesced uint4 loads per thread, all L2 hits), so no
// CTA waits for another. The counts are therefore read-only here and the cursors go to their own array. The last CTA
// has seen every cell and publishes the statistics. Same cell_lb / cursors / statistics as k_cell_reduce + k_cell_downsweep.
constexpr int CELL_SCAN_SMALL_THREADS = 1024;
constexpr uint32_t CELL_SCAN_SMALL_MAX_TILES = 32;
static_assert(CELL_SCAN_SMALL_THREADS * 4 == CELL_SCAN_TILE, "one uint4 per thread per tile");
__global__ void __launch_bounds__(CELL_SCAN_SMALL_THREADS)
k_cell_scan_small(const uint32_t* __restrict__ count, uint32_t* __restrict__ cursor, uint32_t* __restrict__ cell_lb,
                  volatile uint32_t* __restrict__ host_max, uint32_t* __restrict__ zero_next,
                  uint32_t* __restrict__ frame_time, uint32_t bump) {
    pdl_enter();
    if (bump && blockIdx.x == 0 && threadIdx.x == 0) *frame_time += bump;  // as in k_cell_reduce
    __shared__ uint32_t s_pre[32], s_own[32], s_max[32];
    __shared__ float s_sq[32];
    const uint32_t t = threadIdx.x, lane = lane_id(), warp = t >> 5, b = blockIdx.x;
    const uint4* all = reinterpret_cast<const uint4*>(count);
    uint32_t pre = 0, mx = 0;
    float sq = 0.0f;
#pragma unroll 8
    for (uint32_t k = 0; k < b; k++) {
        const uint4 v = __ldg(all + (size_t)k * CELL_SCAN_SMALL_THREADS + t);
        pre += v.x + v.y + v.z + v.w;
        mx = max(max(mx, max(v.x, v.y)), max(v.z, v.w));
        sq += (float)v.x * (float)v.x + (float)v.y * (float)v.y + (float)v.z * (float)v.z + (float)v.w * (float)v.w;
    }
    const size_t mine = (size_t)b * CELL_SCAN_SMALL_THREADS + t;
    const uint4 v = __ldg(all + mine);
    const uint32_t own = v.x + v.y + v.z + v.w;
    mx = max(max(mx, max(v.x, v.y)), max(v.z, v.w));
    sq += (float)v.x * (float)v.x + (float)v.y * (float)v.y + (float)v.z * (float)v.z + (float)v.w * (float)v.w;
    uint32_t incl = own;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t u = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= (uint32_t)o) incl += u;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        pre += __shfl_xor_sync(0xffffffffu, pre, o);
        mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        sq += __shfl_xor_sync(0xffffffffu, sq, o);
    }
    if (lane == 31) s_own[warp] = incl;
    if (lane == 0) { s_pre[warp] = pre; s_max[warp] = mx; s_sq[warp] = sq; }
    __syncthreads();
    uint32_t run = incl - own;
#pragma unroll
    for (int w = 0; w < CELL_SCAN_SMALL_THREADS / 32; w++) {
        run += s_pre[w];
        if ((uint32_t)w < warp) run += s_own[w];
    }
    uint4 o;
    o.x = run; run += v.x;
    o.y = run; run += v.y;
    o.z = run; run += v.z;
    o.w = run;
    reinterpret_cast<uint4*>(cursor)[mine] = o;
    reinterpret_cast<uint4*>(cell_lb)[mine] = o;
    reinterpret_cast<uint4*>(zero_next)[mine] = make_uint4(0u, 0u, 0u, 0u);
    if (b == gridDim.x - 1u && t == 0) {
        uint32_t m = 0;
        float q = 0.0f;
        for (int w = 0; w < CELL_SCAN_SMALL_THREADS / 32; w++) { m = max(m, s_max[w]); q += s_sq[w]; }
        host_max[0] = m;
        host_max[1] = __float_as_uint(q);
    }
}

// slot = cursor[cell]++, one atomic per run of equal cells along the lanes; the run's lanes take consecutive slots, so a
// cell's block is a concatenation of runs that are each in source order (k_rank_gather_cells orders the runs).
// A warp takes SLOT_ITEMS rows of 32 consecutive particles and issues the rows' loads, then their atomics, back to back:
// with one row per thread the kernel ran at the latency of the load -> atomic -> store chain (ncu: 0.33 ms at 64M for
// 0.84 GB of traffic, full occupancy).
#ifndef FLUID_SLOT_ITEMS
#define FLUID_SLOT_ITEMS 4
#endif
constexpr int SLOT_ITEMS = FLUID_SLOT_ITEMS;
__global__ void __launch_bounds__(GATHER_THREADS)
k_cell_slots(uint32_t n, uint32_t key_lo, uint32_t last_bin, const uint32_t* __restrict__ keys_u, uint32_t* __restrict__ cursor,
             uint32_t* __restrict__ slot_keys, uint32_t* __restrict__ slot_idx, const uint32_t* __restrict__ n_dev) {
    // bin of a key = clamp(key - key_lo, 0, last_bin): (0, num_cells - 1) on one GPU; a band of the multi-GPU step passes the
    // first cell of its table and one bin past the last for the sentinel key of departed particles (count_bin), and
    // the number of entries through `n_dev` (only the device knows it; `n` then bounds the grid)
    pdl_enter();
    if (n_dev) n = __ldg(n_dev);
    const uint32_t lane = lane_id();
    const uint32_t warp = (blockIdx.x * GATHER_THREADS + threadIdx.x) >> 5;
    const uint32_t first = warp * (32u * SLOT_ITEMS) + lane;
    const uint32_t upto = (2u << lane) - 1u;  // lanes 0..lane
    uint32_t k[SLOT_ITEMS], head[SLOT_ITEMS], base[SLOT_ITEMS];
#pragma unroll
    for (int j = 0; j < SLOT_ITEMS; j++) {
        const uint32_t i = first + j * 32u;
        k[j] = i < n ? __ldg(keys_u + i) : 0u;
    }
#pragma unroll
    for (int j = 0; j < SLOT_ITEMS; j++) {
        const bool valid = first + j * 32u < n;
        const uint32_t kc = count_bin(k[j], key_lo, last_bin);
        const uint32_t d = valid ? kc : 0xffffffffu;
        const uint32_t prev = __shfl_up_sync(0xffffffffu, d, 1);
        const uint32_t heads = __ballot_sync(0xffffffffu, lane == 0u || d != prev);
        head[j] = 31u - (uint32_t)__clz(heads & upto);  // lane 0 is always a head
        base[j] = 0;
        if (valid && head[j] == lane) {
            const uint32_t above = heads & ~upto;
            const uint32_t next = above ? (uint32_t)__ffs(above) - 1u : 32u;
            base[j] = atomicAdd(cursor + kc, next - lane);
        }
    }
#pragma unroll
    for (int j = 0; j < SLOT_ITEMS; j++) {
        const uint32_t i = first + j * 32u;
        const uint32_t b = __shfl_sync(0xffffffffu, base[j], head[j]);
        if (i < n) {
            const uint32_t slot = b + (lane - head[j]);
            slot_keys[slot] = k[j];
            slot_idx[slot] = i;
        }
    }
}

// Per slot: final position inside the cell's block = number of entries of the block with a smaller source index (stable
// order); then the payload move and the tables of k_gather_cells (same rules, same bits): start_indices[key] = i at the
// first particle of every cell except sorted index 0 (compute.wgsl:45-56), the quirk words for the lowest occupied
// cell, the wild stamp. cell_lb is already complete (k_cell_downsweep).
// The pass is a chain of three dependent global loads ([key, index] -> [cell bounds, payload] -> [the block's indices])
// and ran at exactly that latency with one slot per thread (0.66 ms at 64M, 4.8 TB/s); a thread therefore takes
// RANK_ITEMS slots and the first RANK_UNROLL entries of a block are fetched as independent loads.
#ifndef FLUID_RANK_ITEMS
#define FLUID_RANK_ITEMS 2
#endif
constexpr int RANK_ITEMS = FLUID_RANK_ITEMS;
constexpr int RANK_UNROLL = 8;
__global__ void __launch_bounds__(GATHER_THREADS)
k_rank_gather_cells(const DevParams p, const uint32_t* __restrict__ slot_keys, const uint32_t* __restrict__ slot_idx,
                    const float2* __restrict__ pos_in, const float2* __restrict__ vel_in,
                    float2* __restrict__ pos_s, float2* __restrict__ pred_s, float2* __restrict__ vel_s,
                    uint32_t* __restrict__ start_indices,
                    const uint32_t* __restrict__ cell_lb, uint32_t* __restrict__ quirk,
                    const uint32_t* __restrict__ frame_time, TilePlan* __restrict__ plans) {
    pdl_enter();
    // the staging plans of the neighbour passes for the tiles this CTA's slots cover (k_tile_plans as a separate launch
    // costs one more kernel boundary per tick): they depend on the slot keys and cell_lb only, both final by now
    {
        constexpr uint32_t TILES_PER_CTA = GATHER_THREADS * RANK_ITEMS / SPH_THREADS;
        static_assert(GATHER_THREADS * RANK_ITEMS % SPH_THREADS == 0, "a CTA covers whole tiles");
        const uint32_t tile = blockIdx.x * TILES_PER_CTA + threadIdx.x;
        if (threadIdx.x < TILES_PER_CTA && tile * SPH_THREADS < p.n) make_tile_plan(p, slot_keys, cell_lb, tile, plans);
    }
    const uint32_t q0 = blockIdx.x * (GATHER_THREADS * RANK_ITEMS) + threadIdx.x;
    uint32_t kc[RANK_ITEMS], src[RANK_ITEMS], lo[RANK_ITEMS], hi[RANK_ITEMS], rank[RANK_ITEMS];
    float2 ps[RANK_ITEMS], vs[RANK_ITEMS];
#pragma unroll
    for (int r = 0; r < RANK_ITEMS; r++) {
        const uint32_t q = q0 + r * GATHER_THREADS;
        const bool ok = q < p.n;  // a slot past the end reads particle 0 and cell 0 (the always-empty border cell) and stores nothing
        kc[r] = ok ? min(__ldg(slot_keys + q), p.num_cells - 1u) : 0u;
        src[r] = ok ? __ldg(slot_idx + q) : 0u;
    }
#pragma unroll
    for (int r = 0; r < RANK_ITEMS; r++) {
        lo[r] = __ldg(cell_lb + kc[r]);
        hi[r] = __ldg(cell_lb + kc[r] + 1u);
        ps[r] = __ldg(pos_in + src[r]);
        vs[r] = __ldg(vel_in + src[r]);
    }
#pragma unroll
    for (int r = 0; r < RANK_ITEMS; r++) {
        uint32_t v[RANK_UNROLL];
#pragma unroll
        for (int t = 0; t < RANK_UNROLL; t++) v[t] = (lo[r] + t < hi[r]) ? __ldg(slot_idx + lo[r] + t) : 0xffffffffu;
        uint32_t c = 0;
#pragma unroll
        for (int t = 0; t < RANK_UNROLL; t++) c += (v[t] < src[r]) ? 1u : 0u;
        rank[r] = c;
    }
#pragma unroll
    for (int r = 0; r < RANK_ITEMS; r++) {
        for (uint32_t j = lo[r] + RANK_UNROLL; j < hi[r]; j++) rank[r] += (__ldg(slot_idx + j) < src[r]) ? 1u : 0u;
        if (q0 + r * GATHER_THREADS >= p.n) continue;
        const uint32_t i = lo[r] + rank[r];
        pos_s[i] = ps[r];
        vel_s[i] = vs[r];
        const float2 pr = predict_position(p, ps[r], vs[r]);  // same inputs, same ops as the key pass: same bits
        pred_s[i] = pr;
        if (!(tame_component(pr.x) && tame_component(pr.y) && tame_component(vs[r].x) && tame_component(vs[r].y)))
            quirk[2] = __ldg(frame_time);
        if (rank[r] == 0u) {
            if (i == 0u) {
                quirk[0] = kc[r];
                quirk[1] = start_indices[kc[r]];  // stale entry: nobody writes start_indices[key[0]] this tick
            } else {
                start_indices[kc[r]] = i;  // compute.wgsl:52-54
            }
        }
    }
}

// the source permutation (sorted index -> index before the tick) is a parity tap only: computed on request from the
// slot arrays, which stay valid until the next tick (fluid_sim_download_permutation)
__global__ void __launch_bounds__(GATHER_THREADS)
k_rank_permutation(uint32_t n, uint32_t num_cells, const uint32_t* __restrict__ slot_keys, const uint32_t* __restrict__ slot_idx,
                   const uint32_t* __restrict__ cell_lb, uint32_t* __restrict__ idx_sorted) {
    const uint32_t q = blockIdx.x * GATHER_THREADS + threadIdx.x;
    if (q >= n) return;
    const uint32_t kc = min(__ldg(slot_keys + q), num_cells - 1u);
    const uint32_t lo = __ldg(cell_lb + kc), hi = __ldg(cell_lb + kc + 1u);
    const uint32_t src = __ldg(slot_idx + q);
    uint32_t rank = 0;
    for (uint32_t j = lo; j < hi; j++) rank += (__ldg(slot_idx + j) < src) ? 1u : 0u;
    idx_sorted[lo + rank] = src;
}

}  // namespace fluid
