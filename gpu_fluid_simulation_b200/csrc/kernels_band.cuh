// kernels_band.cuh — the extra passes of the multi-GPU step: the domain is cut into bands of grid
// rows, one band per rank, and every tick needs ONE exchange between the ranks.
//
// Keys stay GLOBAL (funcs.wgsl:216-218), so the concatenation of the ranks' sorted key arrays is the single-GPU sorted
// key array. A rank holds, in key order, [2 ghost rows below | its owned rows | 2 ghost rows above]: with two ghost rows
// it can compute the density of the inner ghost row itself (bit-identical to the owner's: same neighbours, same order),
// so the force pass of the owned rows never waits for a neighbour's densities.
//
// Per tick and rank (host side: multi.inl):
//   1. k_band_key          K1+K2 (compute.wgsl:8-42) of the particles the rank owned after the last tick ("natives").
//                          The key's row decides everything: still inside the band's extended rows -> counted into the
//                          band's cell histogram, else the particle leaves (sentinel key, dropped by the sort). A COPY of
//                          (pos, vel, origin) goes into the outbox of every other rank whose extended rows contain the
//                          row: as an immigrant where that rank owns the row, as a ghost otherwise. The last CTA writes
//                          the message headers (record count, row of the migration matrix).
//   2. [exchange]          one grouped NCCL send/recv per tick (fixed-size windows, the counts travel in the header), or
//                          device copies when the ranks share a process and a device (tests)
//   3. k_band_recv         appends the received records behind the natives, keys + histogram as in 1; one thread folds the
//                          headers into the per-rank particle counts -> this rank's global offset. NOTHING is read back by
//                          the host: every size lives in BandDyn on the device.
//   4. k_cell_reduce / k_cell_downsweep / k_cell_slots (kernels_count.cuh) over the band's cells
//   5. k_band_rank_gather  the in-cell order. Single GPU: stable = by index before the sort. Here every particle carries its
//                          ORIGIN = its global sorted index after the previous tick (natives: offset + index, implicit;
//                          received records: explicit), and a cell's entries are ranked by origin — which reproduces the
//                          single-GPU order exactly, whatever order the records arrived in. Also start_indices (reference
//                          rule, GLOBAL indices, compute.wgsl:45-56), the stale-entry quirk, the staging plans.
//   6. k_density over owned + 1 ghost row per side, k_force_integrate over owned (kernels_sph.cuh, `band_range`)
#pragma once
#include "common.cuh"
#include "kernels_cells.cuh"
#include "kernels_count.cuh"
#include "kernels_sph.cuh"

namespace fluid {

constexpr int BAND_MAX_RANKS = 8;
constexpr int BAND_GHOST_ROWS = 2;
constexpr uint32_t BAND_SENTINEL = 0xffffffffu;  // key of a particle outside the band's extended rows (dropped by the sort)
constexpr int BAND_HDR_WORDS = 16;               // 64-byte message header
constexpr int BAND_KEY_THREADS = 256;

// error bits latched in BandDyn::error (and forwarded in the headers); fluid_multi_sync reports them
constexpr uint32_t BAND_ERR_WINDOW = 1u;    // more records for one destination than its window holds (records lost)
constexpr uint32_t BAND_ERR_CAPACITY = 2u;  // natives + received exceed the rank's capacity (records lost)
constexpr uint32_t BAND_ERR_COUNT = 4u;     // owned count after the sort differs from the migration matrix's (follows 1 / 2)

struct BandRows {
    uint32_t bounds[BAND_MAX_RANKS + 1];  // rank r owns grid rows [bounds[r], bounds[r+1])
    int world, rank;
    uint32_t table_lo;                    // first grid row the band's cell table covers (extended rows + one empty row per side)
};

// header of one message (outbox -> inbox), BAND_HDR_WORDS words:
//   [0] records pushed (may exceed the window: the receiver clamps and latches BAND_ERR_WINDOW)
//   [1] the sender's latched error bits      [2] the sender's native count
//   [4 + r] natives of the sender whose new owner is rank r (the sender's row of the migration matrix)
// followed by float4 (pos, vel)[cap] and uint32 origin[cap]
struct BandBoxes {
    uint32_t* box[BAND_MAX_RANKS];  // null for the rank itself
    uint32_t cap[BAND_MAX_RANKS];   // records per window (multiple of 4)
};
__host__ __device__ inline size_t band_box_bytes(uint32_t cap) { return (size_t)BAND_HDR_WORDS * 4 + (size_t)cap * 20; }
__device__ __forceinline__ float4* band_box_posvel(uint32_t* box) { return reinterpret_cast<float4*>(box + BAND_HDR_WORDS); }
__device__ __forceinline__ uint32_t* band_box_origin(uint32_t* box, uint32_t cap) { return box + BAND_HDR_WORDS + (size_t)cap * 4; }

// everything about a rank's tick that only the device knows
struct BandDyn {
    uint32_t dens_begin, dens_end;  // key-ordered range the density pass computes (owned + one ghost row per side)
    uint32_t own_begin, own_end;    // owned range of the key-ordered arrays; natives of the next tick = own_end - own_begin
    uint32_t goff;                  // global sorted index of the first owned particle (particles on lower ranks)
    uint32_t goff_prev;             // the same after the previous tick: natives' origin = goff_prev + index
    uint32_t n_nat, n_recv, total;  // this tick: natives, received records (clamped), entries fed to the sort
    uint32_t sorted_n;              // entries inside the extended rows (ghosts + owned)
    uint32_t error;                 // latched BAND_ERR_* bits
    uint32_t done_ctas;             // k_band_key: CTAs finished (self-resetting)
    uint32_t out_count[BAND_MAX_RANKS];  // k_band_key: records pushed per destination (self-resetting)
    uint32_t e_row[BAND_MAX_RANKS];      // k_band_key: natives whose new owner is rank r (self-resetting)
    uint32_t e_sent[BAND_MAX_RANKS];     // the row as sent with this tick's headers
    uint32_t n_all[BAND_MAX_RANKS];      // owned particles of every rank (updated from the migration matrix each tick)
};

__device__ __forceinline__ int band_owner(const BandRows& rows, uint32_t row) {
    int o = 0;
#pragma unroll 1
    for (int r = 1; r < rows.world; r++) if (row >= rows.bounds[r]) o = r;
    return o;
}
// rows a rank keeps particles of: its own and BAND_GHOST_ROWS on each side
__device__ __forceinline__ bool band_in_ext(const BandRows& rows, int r, uint32_t row) {
    return rows.bounds[r + 1] > rows.bounds[r] && row + BAND_GHOST_ROWS >= rows.bounds[r] && row < rows.bounds[r + 1] + BAND_GHOST_ROWS;
}

// shared by k_band_key and k_band_recv: key of a particle of this rank's arrays and its histogram bin
__device__ __forceinline__ uint32_t band_key_and_count(const DevParams& p, const BandRows& rows, float2 q, float y_h, bool valid,
                                                      uint32_t key_lo, uint32_t last_bin, uint32_t* __restrict__ count, uint32_t& row_out) {
    uint32_t k = 0, row = 0;
    bool keep = false;
    if (valid) {
        k = p.key_fast_div ? cell_key_shared(p, q, y_h) : cell_key(p, q);
        row = min(k, p.num_cells - 1u) / p.grid_w;
        keep = band_in_ext(rows, rows.rank, row);
    }
    row_out = row;
    cell_count_add(count, keep ? k - key_lo : 0u, keep);  // (warp-collective)
    (void)last_bin;
    return keep ? k : BAND_SENTINEL;
}

// 1. natives: key, histogram, copies for the other ranks, message headers
__global__ void __launch_bounds__(BAND_KEY_THREADS)
k_band_key(const DevParams p, const BandRows rows, const float2* __restrict__ pos, const float2* __restrict__ vel,
           uint32_t* __restrict__ keys, uint32_t* __restrict__ count, uint32_t key_lo, uint32_t last_bin,
           BandDyn* __restrict__ dyn, const BandBoxes out, const uint32_t* __restrict__ origin_explicit /* first tick after an upload, else null */,
           uint32_t* __restrict__ frame_time, uint32_t* __restrict__ quirk) {
    const uint32_t n_nat = dyn->own_end - dyn->own_begin;
    const uint32_t goff_prev = dyn->goff;
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        *frame_time += 1u;          // reference: self.tick += 1 (src/simulation.rs:460)
        quirk[0] = 0xffffffffu;     // lowest occupied cell of the whole domain: set by k_band_rank_gather if this rank sees it
        quirk[1] = 0u;
    }
    const int me = rows.rank;
    const uint32_t lo = rows.bounds[me], hi = rows.bounds[me + 1];
    const float y_h = rcp_refined(p.h);
    const uint32_t stride = gridDim.x * BAND_KEY_THREADS;
    const uint32_t end = (n_nat + 31u) & ~31u;
    const uint32_t lane = lane_id();
    for (uint32_t i = blockIdx.x * BAND_KEY_THREADS + threadIdx.x; i < end; i += stride) {
        const bool v = i < n_nat;
        float2 ps = make_float2(0.f, 0.f), vs = ps;
        if (v) { ps = __ldg(pos + i); vs = __ldg(vel + i); }
        const float2 q = predict_position(p, ps, vs);
        uint32_t row;
        const uint32_t k = band_key_and_count(p, rows, q, y_h, v, key_lo, last_bin, count, row);
        if (v) keys[i] = k;
        // rows [lo + 2, hi - 2) concern nobody else
        const bool edge = v && !(row >= lo + BAND_GHOST_ROWS && row + BAND_GHOST_ROWS < hi);
        if (!__any_sync(0xffffffffu, edge)) continue;
        const uint32_t org = origin_explicit ? (v ? __ldg(origin_explicit + i) : 0u) : goff_prev + i;
#pragma unroll 1
        for (int d = 0; d < rows.world; d++) {
            if (d == me) continue;
            const bool push = edge && band_in_ext(rows, d, row);
            const uint32_t m = __ballot_sync(0xffffffffu, push);
            if (!m) continue;
            const bool owns = push && row >= rows.bounds[d] && row < rows.bounds[d + 1];
            const uint32_t mo = __ballot_sync(0xffffffffu, owns);
            const int leader = __ffs(m) - 1;
            uint32_t base = 0;
            if ((int)lane == leader) {
                base = atomicAdd(&dyn->out_count[d], (uint32_t)__popc(m));
                if (mo) atomicAdd(&dyn->e_row[d], (uint32_t)__popc(mo));
            }
            base = __shfl_sync(0xffffffffu, base, leader);
            if (push) {
                const uint32_t slot = base + (uint32_t)__popc(m & lanemask_lt());
                if (slot < out.cap[d]) {
                    band_box_posvel(out.box[d])[slot] = make_float4(ps.x, ps.y, vs.x, vs.y);
                    band_box_origin(out.box[d], out.cap[d])[slot] = org;
                }
            }
        }
    }
    // the last CTA to finish writes the headers and resets the counters for the next tick
    __shared__ bool s_last;
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        s_last = atomicAdd(&dyn->done_ctas, 1u) == gridDim.x - 1u;
    }
    __syncthreads();
    if (s_last && threadIdx.x < (uint32_t)rows.world) {
        __threadfence();
        const int d = threadIdx.x;
        const uint32_t cnt = atomicExch(&dyn->out_count[d], 0u);
        uint32_t err = dyn->error;
        if (d != me && cnt > out.cap[d]) { err |= BAND_ERR_WINDOW; atomicOr(&dyn->error, BAND_ERR_WINDOW); }
        const uint32_t e = atomicExch(&dyn->e_row[d], 0u);
        dyn->e_sent[d] = e;
        __syncwarp((1u << rows.world) - 1u);
        if (d != me) {
            uint32_t* h = out.box[d];
            h[0] = cnt; h[1] = err; h[2] = n_nat; h[3] = 0u;
            for (int r = 0; r < BAND_MAX_RANKS; r++) h[4 + r] = r < rows.world ? dyn->e_sent[r] : 0u;
        }
        if (d == 0) dyn->done_ctas = 0u;
    }
}

// 3. received records -> behind the natives; keys + histogram; one thread resolves the tick's sizes
__global__ void __launch_bounds__(BAND_KEY_THREADS)
k_band_recv(const DevParams p, const BandRows rows, const BandBoxes in, float2* __restrict__ pos, float2* __restrict__ vel,
            uint32_t* __restrict__ keys, uint32_t* __restrict__ origin, uint32_t* __restrict__ count, uint32_t key_lo,
            uint32_t last_bin, BandDyn* __restrict__ dyn, uint32_t state_cap) {
    const int me = rows.rank;
    const uint32_t n_nat = dyn->own_end - dyn->own_begin;
    // thread -> (source, slot) over the concatenated windows; off = records of earlier sources
    const uint32_t t = blockIdx.x * BAND_KEY_THREADS + threadIdx.x;
    uint32_t cap_off = 0, off = 0, j = 0;
    int src = -1;
#pragma unroll 1
    for (int s = 0; s < rows.world; s++) {
        if (s == me) continue;
        const uint32_t c = min(__ldg(in.box[s]), in.cap[s]);
        if (src < 0) {
            if (t < cap_off + in.cap[s]) { src = s; j = t - cap_off; }
            else { cap_off += in.cap[s]; off += c; }
        }
    }
    bool v = false;
    uint32_t dst = 0;
    float2 ps = make_float2(0.f, 0.f), vs = ps;
    if (src >= 0) {
        const uint32_t c = min(__ldg(in.box[src]), in.cap[src]);
        dst = n_nat + off + j;
        v = j < c && dst < state_cap;
        if (v) {
            const float4 r = __ldg(band_box_posvel(in.box[src]) + j);
            ps = make_float2(r.x, r.y); vs = make_float2(r.z, r.w);
            pos[dst] = ps; vel[dst] = vs;
            origin[dst] = __ldg(band_box_origin(in.box[src], in.cap[src]) + j);
        }
    }
    const float2 q = predict_position(p, ps, vs);
    uint32_t row;
    const uint32_t k = band_key_and_count(p, rows, q, rcp_refined(p.h), v, key_lo, last_bin, count, row);
    if (v) keys[dst] = k;

    if (blockIdx.x == 0 && threadIdx.x == 0) {
        uint32_t n_recv = 0, err = 0;
        // migration matrix: E[s][r] from the headers, this rank's own row from k_band_key
        uint32_t n_all[BAND_MAX_RANKS];
        for (int r = 0; r < rows.world; r++) n_all[r] = dyn->n_all[r];
        for (int s = 0; s < rows.world; s++) {
            uint32_t e[BAND_MAX_RANKS];
            if (s == me) {
                for (int r = 0; r < rows.world; r++) e[r] = dyn->e_sent[r];
            } else {
                const uint32_t* h = in.box[s];
                const uint32_t c = __ldg(h);
                if (c > in.cap[s]) err |= BAND_ERR_WINDOW;
                n_recv += min(c, in.cap[s]);
                err |= __ldg(h + 1);
                for (int r = 0; r < rows.world; r++) e[r] = __ldg(h + 4 + r);
            }
            for (int r = 0; r < rows.world; r++) { n_all[r] += e[r]; n_all[s] -= e[r]; }
        }
        if (n_nat + n_recv > state_cap) { err |= BAND_ERR_CAPACITY; n_recv = state_cap - n_nat; }
        uint32_t goff = 0;
        for (int r = 0; r < rows.world; r++) { dyn->n_all[r] = n_all[r]; if (r < me) goff += n_all[r]; }
        dyn->goff_prev = dyn->goff;
        dyn->goff = goff;
        dyn->n_nat = n_nat; dyn->n_recv = n_recv; dyn->total = n_nat + n_recv;
        if (err) dyn->error |= err;
    }
}

// 5. in-cell order by origin, payload into key order, tables, plans. Grid over the slots of [natives | received].
__global__ void __launch_bounds__(GATHER_THREADS)
k_band_rank_gather(const DevParams p, const BandRows rows, uint32_t key_lo, uint32_t last_bin,
                   const uint32_t* __restrict__ slot_keys, const uint32_t* __restrict__ slot_idx,
                   const uint32_t* __restrict__ bin_lb, const float2* __restrict__ pos_in, const float2* __restrict__ vel_in,
                   const uint32_t* __restrict__ origin, int explicit_natives, BandDyn* __restrict__ dyn,
                   float2* __restrict__ pos_s, float2* __restrict__ pred_s, float2* __restrict__ vel_s,
                   uint32_t* __restrict__ start_indices, uint32_t* __restrict__ quirk, const uint32_t* __restrict__ frame_time,
                   TilePlan* __restrict__ plans) {
    const uint32_t* lbg = bin_lb - key_lo;  // indexed by global cell id, valid for the table's rows
    const int me = rows.rank;
    const uint32_t gw = p.grid_w;
    const uint32_t own_begin = __ldg(lbg + rows.bounds[me] * gw), own_end = __ldg(lbg + rows.bounds[me + 1] * gw);
    const uint32_t sorted_n = __ldg(bin_lb + last_bin);
    const uint32_t total = dyn->total, n_nat = dyn->n_nat, goff = dyn->goff, goff_prev = dyn->goff_prev;
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        const uint32_t lo = rows.bounds[me], hi = rows.bounds[me + 1];
        dyn->own_begin = own_begin; dyn->own_end = own_end;
        dyn->dens_begin = __ldg(lbg + max(max(lo, 1u) - 1u, rows.table_lo) * gw);
        dyn->dens_end = __ldg(lbg + min(hi + 1u, p.grid_h) * gw);
        dyn->sorted_n = sorted_n;
        if (own_end - own_begin != dyn->n_all[me]) dyn->error |= BAND_ERR_COUNT;
    }
    // staging plan of the tile this CTA's slots cover (tiles count from sorted index 0)
    if (threadIdx.x == 0) {
        static_assert(GATHER_THREADS == SPH_THREADS, "one tile per CTA");
        const uint32_t i0 = blockIdx.x * SPH_THREADS;
        if (i0 < sorted_n) make_tile_plan_range(p, slot_keys, lbg, blockIdx.x, i0, min(i0 + SPH_THREADS, sorted_n), plans);
    }
    const uint32_t q = blockIdx.x * GATHER_THREADS + threadIdx.x;
    if (q >= total) return;
    const uint32_t k = __ldg(slot_keys + q);
    if (k == BAND_SENTINEL) return;
    const uint32_t bin = k - key_lo;
    const uint32_t lo_i = __ldg(bin_lb + bin), hi_i = __ldg(bin_lb + bin + 1u);
    const uint32_t src = __ldg(slot_idx + q);
    const float2 ps = __ldg(pos_in + src), vs = __ldg(vel_in + src);
    const bool implicit = !explicit_natives;
    auto origin_of = [&](uint32_t idx) { return (implicit && idx < n_nat) ? goff_prev + idx : __ldg(origin + idx); };
    const uint32_t mine = origin_of(src);
    uint32_t rank = 0;
    for (uint32_t j = lo_i; j < hi_i; j++) rank += origin_of(__ldg(slot_idx + j)) < mine ? 1u : 0u;
    const uint32_t i = lo_i + rank;
    pos_s[i] = ps;
    vel_s[i] = vs;
    const float2 pr = predict_position(p, ps, vs);  // same inputs, same ops as the key pass: same bits
    pred_s[i] = pr;
    if (!(tame_component(pr.x) && tame_component(pr.y) && tame_component(vs.x) && tame_component(vs.y))) quirk[2] = __ldg(frame_time);
    if (rank == 0u) {
        const uint32_t gidx = goff + i - own_begin;  // global sorted index (ghosts below the band: goff >= their count)
        if (gidx == 0u) {
            // compute.wgsl:50: global particle 0 writes nothing; its cell is walked from the stale entry (a GLOBAL index)
            const long long tl = (long long)start_indices[k] - (long long)goff + (long long)own_begin;
            quirk[0] = k;
            quirk[1] = (uint32_t)min(max(tl, 0ll), 0xffffffffll);
        } else {
            start_indices[k] = gidx;  // compute.wgsl:52-54
        }
    }
}

// LOCAL transport (ranks of one process on one device, or peer-accessible devices): a whole window, header first
__global__ void k_band_copy_box(const uint4* __restrict__ src, uint4* __restrict__ dst, uint32_t cap) {
    const uint32_t cnt = min(reinterpret_cast<const uint32_t*>(src)[0], cap);
    const uint32_t hdr16 = BAND_HDR_WORDS / 4;
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    // header | posvel[cnt] | (gap) | origin[cnt]
    if (i < hdr16) dst[i] = src[i];
    if (i < cnt) dst[hdr16 + i] = src[hdr16 + i];
    const uint32_t org16 = hdr16 + cap;  // origin array starts cap float4 later
    if (i < (cnt + 3u) / 4u) dst[org16 + i] = src[org16 + i];
}

}  // namespace fluid
