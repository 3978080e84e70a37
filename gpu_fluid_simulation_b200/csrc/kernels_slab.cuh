// kernels_slab.cuh — the extra passes of slab (multi-GPU) mode. SURVEY.md §8(e).
//
// The domain is cut into bands of grid rows, one band per rank. Keys stay GLOBAL (funcs.wgsl:216-218), so
// the concatenation of the ranks' sorted key arrays is the single-GPU sorted key array, and
// global_start[c] = local_start[c] + (number of particles on lower ranks).
//
// Per tick and rank:
//   1. k_slab_key_hist    keys of the particles the rank held after the last tick ("natives"); a particle whose
//                         predicted position left the band is an emigrant: it gets the sentinel key (sorts last,
//                         dropped after the sort) and is appended to an emigrant list (destination rank, index)
//   2. [exchange]         emigrants' (pos, vel) go to their new owner. Immigrants from lower ranks are placed
//                         BEFORE the natives and immigrants from higher ranks AFTER them, in source order, so the
//                         stable sort reproduces the single-GPU tie order exactly
//   3. k_slab_key_hist    keys of the immigrants (no emigrant detection); then the ordinary onesweep passes
//   4. k_slab_gather      payload into key order, at an offset that leaves room for the halo row below
//   5. [exchange]         first / last grid row of (pred, vel) to the lower / upper neighbour -> halo rows
//   6. k_slab_halo_keys, k_slab_build_cells   keys of halo particles; cell tables over [halo | owned | halo]
//   7. k_density over the owned range; [exchange] rho of the same rows; k_force_integrate over the owned range
#pragma once
#include "common.cuh"
#include "kernels_cells.cuh"
#include "kernels_hash.cuh"
#include "kernels_count.cuh"
#include "slab_rows.h"

namespace fluid {

// 1 particle per thread (the range may start at an odd index, so no float4 loads here)
template <int PASSES>
__global__ void __launch_bounds__(HASH_THREADS)
k_slab_key_hist(const DevParams p, const float2* __restrict__ pos, const float2* __restrict__ vel, uint32_t count,
                uint32_t* __restrict__ keys, uint32_t* __restrict__ g_hist, const SlabRows rows, int detect_emigrants,
                uint32_t list_index_base, unsigned long long* __restrict__ emig_list, uint32_t* __restrict__ emig_count,
                uint32_t* __restrict__ emig_dest_count, uint32_t emig_cap, uint32_t sentinel_key,
                uint32_t* __restrict__ frame_time) {
    __shared__ uint32_t s_hist[PASSES * RADIX];
    for (int i = threadIdx.x; i < PASSES * RADIX; i += HASH_THREADS) s_hist[i] = 0;
    __syncthreads();
    if (frame_time && blockIdx.x == 0 && threadIdx.x == 0) *frame_time += 1u;
    const uint32_t stride = gridDim.x * HASH_THREADS;
    const uint32_t end = (count + 31u) & ~31u;
    const uint32_t row_lo = rows.bounds[rows.rank], row_hi = rows.bounds[rows.rank + 1];
    const float y_h = rcp_refined(p.h);
    for (uint32_t i = blockIdx.x * HASH_THREADS + threadIdx.x; i < end; i += stride) {
        const bool v = i < count;
        uint32_t k = 0;
        if (v) {
            const float2 q = predict_position(p, __ldg(pos + i), __ldg(vel + i));
            k = p.key_fast_div ? cell_key_shared(p, q, y_h) : cell_key(p, q);
            if (detect_emigrants) {
                const uint32_t row = k / p.grid_w;
                if (row < row_lo || row >= row_hi) {
                    int dest = 0;
#pragma unroll 1
                    for (int r = 0; r < rows.world; r++) if (row >= rows.bounds[r]) dest = r;
                    const uint32_t slot = atomicAdd(emig_count, 1u);
                    atomicAdd(emig_dest_count + dest, 1u);
                    if (slot < emig_cap) emig_list[slot] = ((unsigned long long)dest << 32) | (unsigned long long)(list_index_base + i);
                    k = sentinel_key;
                }
            }
            keys[i] = k;
        }
#pragma unroll
        for (int q = 0; q < PASSES; q++) hist_add(s_hist + q * RADIX, (k >> (q * RADIX_BITS)) & (RADIX - 1), v);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < PASSES * RADIX; i += HASH_THREADS) {
        uint32_t c = s_hist[i];
        if (c) atomicAdd(g_hist + i, c);
    }
}

// payload + keys into key order at `own_begin`
__global__ void __launch_bounds__(GATHER_THREADS)
k_slab_gather(const DevParams p, const uint32_t* __restrict__ sorted_keys, const uint32_t* __restrict__ sorted_idx,
              const float2* __restrict__ pos_in, const float2* __restrict__ vel_in, float2* __restrict__ pos_s,
              float2* __restrict__ pred_s, float2* __restrict__ vel_s, uint32_t* __restrict__ keys_ext,
              uint32_t* __restrict__ wild_stamp, const uint32_t* __restrict__ frame_time) {
    const uint32_t i = blockIdx.x * GATHER_THREADS + threadIdx.x;
    if (i >= p.n) return;
    const uint32_t src = __ldg(sorted_idx + i);
    const float2 ps = __ldg(pos_in + src), vs = __ldg(vel_in + src);
    const uint32_t o = p.own_begin + i;
    const float2 pr = predict_position(p, ps, vs);
    pos_s[o] = ps;
    vel_s[o] = vs;
    pred_s[o] = pr;
    keys_ext[o] = __ldg(sorted_keys + i);
    // wild stamp of this tick (common.cuh tame_component); halo rows are tested where their keys are made
    if (!(tame_component(pr.x) && tame_component(pr.y) && tame_component(vs.x) && tame_component(vs.y))) *wild_stamp = __ldg(frame_time);
}
// counting path (kernels_count.cuh): per slot of [from below | natives | from above] in arrival order: rank inside its
// bin by source index = the stable order of the radix passes; emigrants carry the sentinel key, sit in the last bin
// (positions >= p.n) and are dropped. Same outputs as k_slab_gather.
__global__ void __launch_bounds__(GATHER_THREADS)
k_slab_rank_gather(const DevParams p, uint32_t total, uint32_t key_lo, uint32_t last_bin, const uint32_t* __restrict__ slot_keys,
                   const uint32_t* __restrict__ slot_idx, const uint32_t* __restrict__ bin_lb,
                   const float2* __restrict__ pos_in, const float2* __restrict__ vel_in, float2* __restrict__ pos_s,
                   float2* __restrict__ pred_s, float2* __restrict__ vel_s, uint32_t* __restrict__ keys_ext,
                   uint32_t* __restrict__ wild_stamp, const uint32_t* __restrict__ frame_time) {
    const uint32_t q = blockIdx.x * GATHER_THREADS + threadIdx.x;
    if (q >= total) return;
    const uint32_t k = __ldg(slot_keys + q);
    const uint32_t bin = count_bin(k, key_lo, last_bin);
    if (bin == last_bin) return;  // emigrant
    const uint32_t lo = __ldg(bin_lb + bin), hi = __ldg(bin_lb + bin + 1u);
    const uint32_t src = __ldg(slot_idx + q);
    const float2 ps = __ldg(pos_in + src), vs = __ldg(vel_in + src);
    uint32_t rank = 0;
    for (uint32_t j = lo; j < hi; j++) rank += (__ldg(slot_idx + j) < src) ? 1u : 0u;
    const uint32_t i = lo + rank;
    if (i >= p.n) return;
    const uint32_t o = p.own_begin + i;
    const float2 pr = predict_position(p, ps, vs);
    pos_s[o] = ps;
    vel_s[o] = vs;
    pred_s[o] = pr;
    keys_ext[o] = k;
    if (!(tame_component(pr.x) && tame_component(pr.y) && tame_component(vs.x) && tame_component(vs.y))) *wild_stamp = __ldg(frame_time);
}
__device__ __forceinline__ void stamp_if_wild(float2 pr, float2 vs, uint32_t* wild_stamp, const uint32_t* frame_time) {
    if (!(tame_component(pr.x) && tame_component(pr.y) && tame_component(vs.x) && tame_component(vs.y))) *wild_stamp = __ldg(frame_time);
}

// keys of received halo particles from their predicted positions (same arithmetic as the hash pass)
__global__ void k_slab_halo_keys(const DevParams p, const float2* __restrict__ pred_s, uint32_t* __restrict__ keys_ext,
                                 uint32_t below_begin, uint32_t below_count, uint32_t above_begin, uint32_t above_count,
                                 const float2* __restrict__ vel_s, uint32_t* __restrict__ wild_stamp,
                                 const uint32_t* __restrict__ frame_time) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t i;
    if (t < below_count) i = below_begin + t;
    else if (t < below_count + above_count) i = above_begin + (t - below_count);
    else return;
    const float2 pr = __ldg(pred_s + i);
    keys_ext[i] = cell_key(p, pr);
    stamp_if_wild(pr, __ldg(vel_s + i), wild_stamp, frame_time);
}

// densities of the received halo rows (exchanged after the density pass): the last input of the force pass's stamp
__global__ void k_slab_check_rho_halo(const float* __restrict__ rho, uint32_t below_begin, uint32_t below_count, uint32_t above_begin,
                                      uint32_t above_count, const uint32_t* __restrict__ dyn_hb_ha /* {hb, ha} or null */,
                                      uint32_t own_begin, uint32_t* __restrict__ wild_stamp, const uint32_t* __restrict__ frame_time) {
    if (dyn_hb_ha) {  // device-resolved protocol: the counts live on the device, the windows are pad-sized
        below_count = dyn_hb_ha[0]; above_count = dyn_hb_ha[1];
        below_begin = own_begin - below_count;
    }
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t i;
    if (t < below_count) i = below_begin + t;
    else if (t < below_count + above_count) i = above_begin + (t - below_count);
    else return;
    if (!(__ldg(rho + i) <= DIV_DEN_HI)) *wild_stamp = __ldg(frame_time);
}

// ---- device-resolved protocol: what the host used to read back between the sort and the cell tables --------------
struct SlabDyn {
    uint32_t hb, ha;             // halo particles below / above (the neighbours' last / first row counts)
    uint32_t ext_begin, ext_end; // [halo | owned | halo] in the key-ordered arrays
    uint32_t global_offset;      // particles on lower ranks
    uint32_t quirk_c0, quirk_stale_ext;
    uint32_t error;              // bit 0: a halo row is larger than the pad
};
constexpr uint32_t SLAB_META_WORDS = 8;  // per rank in the gathered array: n, first-row count, last-row count, first key, its table entry
// one thread: every rank's band metadata (all_gather output, device resident) -> this rank's SlabDyn and quirk words
__global__ void k_slab_resolve(const uint32_t* __restrict__ allmeta, int world, int me, uint32_t own_begin, uint32_t n,
                               uint32_t pad, SlabDyn* __restrict__ dyn, uint32_t* __restrict__ quirk) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    uint32_t prefix = 0;
    for (int r = 0; r < me; r++) prefix += allmeta[r * SLAB_META_WORDS];
    uint32_t hb = me > 0 ? allmeta[(me - 1) * SLAB_META_WORDS + 2] : 0u;
    uint32_t ha = me + 1 < world ? allmeta[(me + 1) * SLAB_META_WORDS + 1] : 0u;
    uint32_t err = 0;
    if (hb > pad) { hb = pad; err |= 1u; }
    if (ha > pad) { ha = pad; err |= 1u; }
    uint32_t c0 = 0xffffffffu, stale = 0;
    for (int r = 0; r < world; r++)
        if (allmeta[r * SLAB_META_WORDS] > 0u) { c0 = allmeta[r * SLAB_META_WORDS + 3]; stale = allmeta[r * SLAB_META_WORDS + 4]; break; }
    const uint32_t ext_begin = own_begin - hb;
    uint32_t stale_ext = 0;
    if (c0 != 0xffffffffu) {  // the stale entry is a GLOBAL sorted index: translate into this rank's extended index space
        long long t = (long long)stale - (long long)prefix + (long long)own_begin;
        if (t < (long long)ext_begin) t = ext_begin;
        if (t > 0xffffffffll) t = 0xffffffffll;
        stale_ext = (uint32_t)t;
    }
    dyn->hb = hb; dyn->ha = ha; dyn->ext_begin = ext_begin; dyn->ext_end = own_begin + n + ha;
    dyn->global_offset = prefix; dyn->quirk_c0 = c0; dyn->quirk_stale_ext = stale_ext;
    dyn->error |= err;
    quirk[0] = c0; quirk[1] = stale_ext;
}
// keys of the received halo particles; grid over 2 * pad threads, counts from SlabDyn
__global__ void k_slab_halo_keys_dev(const DevParams p, const float2* __restrict__ pred_s, uint32_t* __restrict__ keys_ext,
                                     const SlabDyn* __restrict__ dyn, uint32_t own_begin, uint32_t n, uint32_t pad,
                                     const float2* __restrict__ vel_s, uint32_t* __restrict__ wild_stamp,
                                     const uint32_t* __restrict__ frame_time) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t i;
    if (t < pad) {
        if (t < pad - dyn->hb) return;  // the window is right-aligned: only its last hb entries are halo particles
        i = own_begin - pad + t;
    } else if (t < 2u * pad) {
        if (t - pad >= dyn->ha) return;
        i = own_begin + n + (t - pad);
    } else return;
    const float2 pr = __ldg(pred_s + i);
    keys_ext[i] = cell_key(p, pr);
    stamp_if_wild(pr, __ldg(vel_s + i), wild_stamp, frame_time);
}
// an empty extended range still needs a valid (all-empty) table
__global__ void k_slab_fill_if_empty(uint32_t* __restrict__ cell_lb, uint32_t first, uint32_t count, const SlabDyn* __restrict__ dyn) {
    if (dyn->ext_end != dyn->ext_begin) return;
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < count) cell_lb[first + i] = dyn->ext_begin;
}

// cell tables over the extended range. start_indices holds GLOBAL sorted indices for owned cells
// (compute.wgsl:45-56: global particle 0 writes nothing); cell_lb holds local (extended) indices.
__global__ void __launch_bounds__(GATHER_THREADS)
k_slab_build_cells(const DevParams p, const uint32_t* __restrict__ keys_ext, uint32_t* __restrict__ start_indices,
                   uint32_t* __restrict__ cell_lb, uint32_t* __restrict__ quirk, uint32_t quirk_c0, uint32_t quirk_stale,
                   uint32_t fill_lo, uint32_t fill_hi, const SlabDyn* __restrict__ dyn, uint32_t grid_first) {
    // [fill_lo, fill_hi): the table entries the neighbour passes of this band can read (rows row_first-1 .. row_end+1);
    // gaps outside it (half the grid for an edge particle of a 2-rank split) are not filled
    // dyn == nullptr: the extended range and the global offset are in `p` (host-resolved protocol) and the grid starts
    // at ext_begin; dyn != nullptr: they were resolved on the device (k_slab_resolve) and the grid covers
    // [grid_first, ...) = the owned range plus a full pad on both sides
    const uint32_t ext_begin = dyn ? dyn->ext_begin : p.ext_begin, ext_end = dyn ? dyn->ext_end : p.ext_end;
    const uint32_t global_offset = dyn ? dyn->global_offset : p.global_offset;
    const uint32_t i = (dyn ? grid_first : ext_begin) + blockIdx.x * GATHER_THREADS + threadIdx.x;
    const bool valid = i >= ext_begin && i < ext_end;
    uint32_t gap_begin = 0, gap_end = 0, gap_value = 0;
    if (!dyn && blockIdx.x == 0 && threadIdx.x == 0) { quirk[0] = quirk_c0; quirk[1] = quirk_stale; }
    if (valid) {
        const uint32_t kc = min(__ldg(keys_ext + i), p.num_cells - 1u);
        const bool owned = i >= p.own_begin && i < p.own_begin + p.n;
        const uint32_t gidx = i - p.own_begin + global_offset;
        if (i == ext_begin) {
            gap_begin = 0; gap_end = kc + 1u; gap_value = ext_begin;
            if (owned && gidx > 0u) start_indices[kc] = gidx;  // previous global particle lives on a lower rank, in another row
        } else {
            const uint32_t kp = min(__ldg(keys_ext + i - 1), p.num_cells - 1u);
            if (kc != kp) {
                if (owned && gidx > 0u) start_indices[kc] = gidx;
                gap_begin = kp + 1u; gap_end = kc + 1u; gap_value = i;
            }
        }
    }
    gap_begin = max(gap_begin, fill_lo);
    gap_end = min(gap_end, fill_hi);
    if (gap_end < gap_begin) gap_end = gap_begin;
    warp_fill_gaps(cell_lb, gap_begin, gap_end, gap_value);
    uint32_t tb = 0, te = 0;
    if (valid && i == ext_end - 1u) {
        tb = max(min(__ldg(keys_ext + i), p.num_cells - 1u) + 1u, fill_lo);
        te = min(p.num_cells + 1u, fill_hi);
        if (te < tb) te = tb;
    }
    if (__any_sync(0xffffffffu, te > tb)) warp_fill_gaps(cell_lb, tb, te, ext_end);
}
// an empty extended range still needs a valid (all-empty) table
__global__ void k_slab_fill_table(uint32_t* __restrict__ cell_lb, uint32_t first, uint32_t count, uint32_t value) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < count) cell_lb[first + i] = value;
}

// meta[0] = owned particles, meta[1] = of them in the band's first grid row, meta[2] = in its last grid row,
// meta[3] = key of the first owned particle, meta[4] = this rank's start_indices entry for that key
__global__ void k_slab_meta(const uint32_t* __restrict__ keys_owned, uint32_t n, uint32_t grid_w, uint32_t row_first,
                            uint32_t row_last, const uint32_t* __restrict__ start_indices, uint32_t num_cells,
                            uint32_t* __restrict__ meta) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    // lower_bound(keys, (row_first + 1) * grid_w) and lower_bound(keys, row_last * grid_w)
    uint32_t targets[2] = {(row_first + 1u) * grid_w, row_last * grid_w};
    uint32_t res[2];
    for (int t = 0; t < 2; t++) {
        uint32_t lo = 0, hi = n;
        while (lo < hi) {
            uint32_t mid = (lo + hi) >> 1;
            if (keys_owned[mid] < targets[t]) lo = mid + 1; else hi = mid;
        }
        res[t] = lo;
    }
    meta[0] = n;
    meta[1] = res[0];
    meta[2] = n - res[1];
    const uint32_t k0 = n ? min(keys_owned[0], num_cells - 1u) : 0xffffffffu;
    meta[3] = k0;
    meta[4] = n ? start_indices[k0] : 0u;
}

// ---- migration payload (phase 2) ------------------------------------------------------------------------
// emigrant list entries are (destination << 32 | state index), appended with an atomic counter: sorting them
// ascending groups them by destination in source order, which is what makes the stable key sort reproduce
// the single-GPU tie order. One CTA sorts up to EMIG_SORT_MAX entries in shared memory (bitonic network).
constexpr int EMIG_SORT_MAX = 4096;
constexpr int EMIG_SORT_THREADS = 1024;
__global__ void __launch_bounds__(EMIG_SORT_THREADS)
k_emig_sort_small(unsigned long long* __restrict__ list, uint32_t m) {
    __shared__ unsigned long long s[EMIG_SORT_MAX];
    uint32_t cap = 2;
    while (cap < m) cap <<= 1;
    for (uint32_t i = threadIdx.x; i < cap; i += EMIG_SORT_THREADS) s[i] = i < m ? list[i] : ~0ull;
    __syncthreads();
    for (uint32_t k = 2; k <= cap; k <<= 1)
        for (uint32_t j = k >> 1; j > 0; j >>= 1) {
            for (uint32_t i = threadIdx.x; i < cap; i += EMIG_SORT_THREADS) {
                const uint32_t l = i ^ j;
                if (l > i) {
                    const unsigned long long a = s[i], b = s[l];
                    const bool up = (i & k) == 0;
                    if ((a > b) == up) { s[i] = b; s[l] = a; }
                }
            }
            __syncthreads();
        }
    for (uint32_t i = threadIdx.x; i < m; i += EMIG_SORT_THREADS) list[i] = s[i];
}
// one compare-exchange step of a global-memory bitonic network (lists longer than EMIG_SORT_MAX; `cap` = m
// rounded up to a power of two, entries past m hold 0xff.. which sorts last)
__global__ void k_emig_sort_step(unsigned long long* __restrict__ list, uint32_t cap, uint32_t k, uint32_t j) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t l = i ^ j;
    if (i >= cap || l <= i) return;
    const unsigned long long a = list[i], b = list[l];
    const bool up = (i & k) == 0;
    if ((a > b) == up) { list[i] = b; list[l] = a; }
}
// payload[t] = (pos, vel) of the t-th emigrant of the sorted list
__global__ void k_emig_pack(const unsigned long long* __restrict__ list, uint32_t m, const float2* __restrict__ pos,
                            const float2* __restrict__ vel, float4* __restrict__ payload) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= m) return;
    const uint32_t idx = (uint32_t)(list[t] & 0xffffffffull);
    const float2 a = pos[idx], b = vel[idx];
    payload[t] = make_float4(a.x, a.y, b.x, b.y);
}
// immigrants: recv[0, nb) came from lower ranks -> state[first_below ...]; recv[nb, nb+na) from higher ranks -> state[first_above ...]
__global__ void k_immig_unpack(const float4* __restrict__ recv, uint32_t nb, uint32_t na, uint32_t first_below,
                               uint32_t first_above, float2* __restrict__ pos, float2* __restrict__ vel) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= nb + na) return;
    const float4 v = recv[t];
    const uint32_t o = t < nb ? first_below + t : first_above + (t - nb);
    pos[o] = make_float2(v.x, v.y);
    vel[o] = make_float2(v.z, v.w);
}

}  // namespace fluid
