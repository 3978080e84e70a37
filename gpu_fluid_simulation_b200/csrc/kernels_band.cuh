// kernels_band.cuh — the extra passes of the multi-GPU step: the domain is cut into bands of grid
// rows, one band per rank, and every tick needs ONE exchange between the ranks.
//
// Keys stay GLOBAL (funcs.wgsl:216-218), so the concatenation of the ranks' sorted key arrays is the single-GPU sorted
// key array. A rank holds, in key order, [2 ghost rows below | its owned rows | 2 ghost rows above]: with two ghost rows
// it can compute the density of the inner ghost row itself (bit-identical to the owner's: same neighbours, same order),
// so the force pass of the owned rows never waits for a neighbour's densities.
//
// Per tick and rank (host side: multi.inl):
//   1. k_band_key          (first tick after an upload / a change of delta; otherwise the previous tick's force pass has already
//                          done this step in its epilogue, band_common.cuh)
//                          K1+K2 (compute.wgsl:8-42) of the particles the rank owned after the last tick ("natives").
//                          The key's row decides everything: still inside the band's extended rows -> counted into the
//                          band's cell histogram, else the particle leaves (sentinel key, dropped by the sort). A COPY of
//                          (pos, vel, origin) goes into the outbox of every other rank whose extended rows contain the
//                          row: as an immigrant where that rank owns the row, as a ghost otherwise. k_band_headers (one
//                          warp) then writes the message headers (record count, row of the migration matrix).
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
#include "band_common.cuh"
#include "kernels_cells.cuh"
#include "kernels_count.cuh"
#include "kernels_sph.cuh"

namespace fluid {

// 1. natives: key, histogram, copies for the other ranks, message headers. Only on the first tick after an upload or a
// change of `delta`: otherwise the previous tick's force pass has done all of this in its epilogue (kernels_sph.cuh).
__global__ void __launch_bounds__(BAND_KEY_THREADS)
k_band_key(const DevParams p, const BandSpec* __restrict__ spec, const float2* __restrict__ pos, const float2* __restrict__ vel,
           uint32_t* __restrict__ keys, uint32_t* __restrict__ count, const uint32_t* __restrict__ origin_explicit /* first tick after an upload, else null */) {
    pdl_enter();
    const BandRows rows = spec->rows;
    BandDyn* dyn = spec->dyn;
    const uint32_t n_nat = dyn->own_end - dyn->own_begin;
    const uint32_t goff_prev = dyn->goff;
    const float y_h = rcp_refined(p.h);
    const uint32_t stride = gridDim.x * BAND_KEY_THREADS;
    const uint32_t end = (n_nat + 31u) & ~31u;
    for (uint32_t i = blockIdx.x * BAND_KEY_THREADS + threadIdx.x; i < end; i += stride) {
        const bool v = i < n_nat;
        float2 ps = make_float2(0.f, 0.f), vs = ps;
        if (v) { ps = __ldg(pos + i); vs = __ldg(vel + i); }
        const float2 q = predict_position(p, ps, vs);
        uint32_t row;
        const uint32_t k = band_key_and_count(p, rows, q, y_h, v, spec->key_lo, count, row);
        if (v) keys[i] = k;
        const uint32_t org = origin_explicit ? (v ? __ldg(origin_explicit + i) : 0u) : goff_prev + i;
        band_push_copies(rows, spec->out, dyn, v, row, ps, vs, org);
    }
}

// 3. received records -> behind the natives; keys + histogram; one thread resolves the tick's sizes
__global__ void __launch_bounds__(BAND_KEY_THREADS)
k_band_recv(const DevParams p, const BandRows rows, const BandBoxes in, float2* __restrict__ pos, float2* __restrict__ vel,
            uint32_t* __restrict__ keys, uint32_t* __restrict__ origin, uint32_t* __restrict__ count, uint32_t key_lo,
            BandDyn* __restrict__ dyn, uint32_t state_cap, uint32_t* __restrict__ frame_time, uint32_t* __restrict__ quirk) {
    pdl_enter();
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
    const uint32_t k = band_key_and_count(p, rows, q, rcp_refined(p.h), v, key_lo, count, row);
    if (v) keys[dst] = k;

    if (blockIdx.x == 0 && threadIdx.x == 0) {
        *frame_time += 1u;          // reference: self.tick += 1 (src/simulation.rs:460); read by the later kernels of the tick
        quirk[0] = 0xffffffffu;     // lowest occupied cell of the whole domain: set by k_band_rank_gather if this rank sees it
        quirk[1] = 0u;
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

// 5. in-cell order by origin, payload into key order, tables, plans. Grid over the slots of [natives | received];
// a thread takes BAND_RANK_ITEMS slots and fetches the first RANK_UNROLL entries of a block as independent loads
// (the pass is a chain of dependent loads, as k_rank_gather_cells).
constexpr int BAND_RANK_ITEMS = 2;
__global__ void __launch_bounds__(GATHER_THREADS)
k_band_rank_gather(const DevParams p, const BandRows rows, uint32_t key_lo, uint32_t last_bin,
                   const uint32_t* __restrict__ slot_keys, const uint32_t* __restrict__ slot_idx,
                   const uint32_t* __restrict__ bin_lb, const float2* __restrict__ pos_in, const float2* __restrict__ vel_in,
                   const uint32_t* __restrict__ origin, int explicit_natives, BandDyn* __restrict__ dyn,
                   float2* __restrict__ pos_s, float2* __restrict__ pred_s, float2* __restrict__ vel_s,
                   uint32_t* __restrict__ start_indices, uint32_t* __restrict__ quirk, const uint32_t* __restrict__ frame_time,
                   TilePlan* __restrict__ plans) {
    pdl_enter();
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
    // staging plans of the tiles this CTA's slots cover (tiles count from sorted index 0)
    {
        constexpr uint32_t TILES_PER_CTA = GATHER_THREADS * BAND_RANK_ITEMS / SPH_THREADS;
        static_assert(GATHER_THREADS * BAND_RANK_ITEMS % SPH_THREADS == 0, "a CTA covers whole tiles");
        const uint32_t tile = blockIdx.x * TILES_PER_CTA + threadIdx.x;
        const uint32_t i0 = tile * SPH_THREADS;
        if (threadIdx.x < TILES_PER_CTA && i0 < sorted_n) make_tile_plan_range(p, slot_keys, lbg, tile, i0, min(i0 + SPH_THREADS, sorted_n), plans);
    }
    const bool implicit = !explicit_natives;
    auto origin_of = [&](uint32_t idx) { return (implicit && idx < n_nat) ? goff_prev + idx : __ldg(origin + idx); };
    const uint32_t q0 = blockIdx.x * (GATHER_THREADS * BAND_RANK_ITEMS) + threadIdx.x;
    uint32_t k[BAND_RANK_ITEMS], src[BAND_RANK_ITEMS], lo_i[BAND_RANK_ITEMS], hi_i[BAND_RANK_ITEMS], rank[BAND_RANK_ITEMS], mine[BAND_RANK_ITEMS];
    bool ok[BAND_RANK_ITEMS];
    float2 ps[BAND_RANK_ITEMS], vs[BAND_RANK_ITEMS];
#pragma unroll
    for (int r = 0; r < BAND_RANK_ITEMS; r++) {
        const uint32_t q = q0 + r * GATHER_THREADS;
        k[r] = q < total ? __ldg(slot_keys + q) : BAND_SENTINEL;
        ok[r] = k[r] != BAND_SENTINEL;  // a slot past the end or a departed particle: reads entry 0 and stores nothing
        src[r] = ok[r] ? __ldg(slot_idx + q) : 0u;
    }
#pragma unroll
    for (int r = 0; r < BAND_RANK_ITEMS; r++) {
        const uint32_t bin = ok[r] ? k[r] - key_lo : 0u;
        lo_i[r] = ok[r] ? __ldg(bin_lb + bin) : 0u;
        hi_i[r] = ok[r] ? __ldg(bin_lb + bin + 1u) : 0u;
        ps[r] = __ldg(pos_in + src[r]);
        vs[r] = __ldg(vel_in + src[r]);
        mine[r] = origin_of(src[r]);
    }
#pragma unroll
    for (int r = 0; r < BAND_RANK_ITEMS; r++) {
        uint32_t v[RANK_UNROLL];
#pragma unroll
        for (int t = 0; t < RANK_UNROLL; t++) v[t] = (lo_i[r] + t < hi_i[r]) ? __ldg(slot_idx + lo_i[r] + t) : 0xffffffffu;
        uint32_t c = 0;
#pragma unroll
        for (int t = 0; t < RANK_UNROLL; t++) c += (v[t] != 0xffffffffu && origin_of(v[t]) < mine[r]) ? 1u : 0u;
        rank[r] = c;
    }
#pragma unroll
    for (int r = 0; r < BAND_RANK_ITEMS; r++) {
        for (uint32_t j = lo_i[r] + RANK_UNROLL; j < hi_i[r]; j++) rank[r] += origin_of(__ldg(slot_idx + j)) < mine[r] ? 1u : 0u;
        if (!ok[r]) continue;
        const uint32_t i = lo_i[r] + rank[r];
        pos_s[i] = ps[r];
        vel_s[i] = vs[r];
        const float2 pr = predict_position(p, ps[r], vs[r]);  // same inputs, same ops as the key pass: same bits
        pred_s[i] = pr;
        if (!(tame_component(pr.x) && tame_component(pr.y) && tame_component(vs[r].x) && tame_component(vs[r].y))) quirk[2] = __ldg(frame_time);
        if (rank[r] == 0u) {
            const uint32_t gidx = goff + i - own_begin;  // global sorted index (ghosts below the band: goff >= their count)
            if (gidx == 0u) {
                // compute.wgsl:50: global particle 0 writes nothing; its cell is walked from the stale entry (a GLOBAL index)
                const long long tl = (long long)start_indices[k[r]] - (long long)goff + (long long)own_begin;
                quirk[0] = k[r];
                quirk[1] = (uint32_t)min(max(tl, 0ll), 0xffffffffll);
            } else {
                start_indices[k[r]] = gidx;  // compute.wgsl:52-54
            }
        }
    }
}

// a key pass that was run ahead by the force pass and is NOT used (the next tick came with another delta): its pushes are
// forgotten before the key pass proper runs (the records it stored are overwritten from slot 0)
__global__ void k_band_forget_pushes(BandDyn* __restrict__ dyn) {
    if (threadIdx.x < BAND_MAX_RANKS) { dyn->out_count[threadIdx.x] = 0u; dyn->e_row[threadIdx.x] = 0u; }
}

// ---- pipelined host transfers of a band (fluid_multi_upload_band_async / _download_band_async) -------------------------
__global__ void k_band_set_word(uint32_t* __restrict__ w, uint32_t v) { *w = v; }
// after an upload whose sizes were all-gathered on the device: the counters band_set_counts writes from the host
__global__ void k_band_counts_from_gathered(BandDyn* __restrict__ dyn, const uint32_t* __restrict__ gathered, int rank, int world,
                                            uint32_t n_total) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    uint32_t goff = 0, total = 0;
    for (int r = 0; r < world; r++) { const uint32_t n = gathered[r]; dyn->n_all[r] = n; total += n; if (r < rank) goff += n; }
    dyn->own_begin = 0; dyn->own_end = gathered[rank];
    dyn->dens_begin = 0; dyn->dens_end = gathered[rank];
    dyn->goff = goff; dyn->goff_prev = goff;
    for (int r = 0; r < BAND_MAX_RANKS; r++) { dyn->out_count[r] = 0; dyn->e_row[r] = 0; dyn->e_sent[r] = 0; }
    if (total != n_total) dyn->error |= BAND_ERR_COUNT;
}
// the band's first min(cap, size) particles in key order -> AoS; info = {size, global offset}
__global__ void k_band_soa_to_aos(FluidParticle* __restrict__ aos, uint32_t cap, const BandDyn* __restrict__ dyn,
                                  const float2* __restrict__ pos, const float2* __restrict__ pred_s, const float2* __restrict__ vel,
                                  const float* __restrict__ rho_s, const uint32_t* __restrict__ keys_s, uint32_t* __restrict__ info) {
    const uint32_t ob = dyn->own_begin, n = dyn->own_end - dyn->own_begin;
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) { info[0] = n; info[1] = dyn->goff; }
    if (i >= min(n, cap)) return;
    const float2 a = pos[i], b = pred_s[ob + i], c = vel[i];
    float4* dst = reinterpret_cast<float4*>(aos + i);
    dst[0] = make_float4(a.x, a.y, b.x, b.y);
    dst[1] = make_float4(c.x, c.y, rho_s[ob + i], __uint_as_float(keys_s[ob + i]));
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

// ---- PEER transport: one rank per process, the other ranks' outboxes mapped through CUDA IPC -----------------------
// The exchange of a tick is a PULL over NVLink: the receiver waits until the sender's header carries this exchange's number
// (k_band_headers stores it last), then copies the header and the records that are there — not the whole window, as the
// fixed-size NCCL messages must. Outboxes are double-buffered by exchange parity, so a sender never waits for a receiver:
// the buffer it fills for exchange x + 1 was last read in exchange x - 1, which every peer finished before it stored
// the header of exchange x that this rank has already waited for.
struct PeerBoxes { const uint32_t* box[BAND_MAX_RANKS]; };  // rank s's outbox for this rank (this exchange's parity), null for itself

__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ uint64_t global_timer_ns() {
    uint64_t t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
// one thread per source rank polls that rank's header; a rank that does not show up within the limit latches BAND_ERR_PEER
// (the pull then copies whatever count the stale header holds, clamped to the window: memory-safe, and the error is reported)
__global__ void k_band_wait_peers(const PeerBoxes src, int world, int me, uint32_t stamp, BandDyn* __restrict__ dyn, uint64_t limit_ns) {
    const int s = threadIdx.x;
    if (s >= world || s == me) return;
    const uint32_t* flag = src.box[s] + 3;
    const uint64_t t0 = global_timer_ns();
    while (ld_acquire_sys(flag) != stamp) {
        if (global_timer_ns() - t0 > limit_ns) { atomicOr(&dyn->error, BAND_ERR_PEER); break; }
        __nanosleep(200);
    }
}
// grid (chunks, world): header, posvel[count], origin[count] of every source into this rank's inboxes
__global__ void k_band_pull(const PeerBoxes src, const BandBoxes in, int me) {
    const int s = blockIdx.y;
    if (s == me) return;
    const uint4* from = reinterpret_cast<const uint4*>(src.box[s]);
    uint4* to = reinterpret_cast<uint4*>(in.box[s]);
    const uint32_t cap = in.cap[s];
    const uint32_t cnt = min(__ldcv(src.box[s]), cap);
    const uint32_t hdr16 = BAND_HDR_WORDS / 4;
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < hdr16) to[i] = __ldcv(from + i);
    if (i < cnt) to[hdr16 + i] = __ldcv(from + hdr16 + i);
    const uint32_t org16 = hdr16 + cap;
    if (i < (cnt + 3u) / 4u) to[org16 + i] = __ldcv(from + org16 + i);
}

}  // namespace fluid
