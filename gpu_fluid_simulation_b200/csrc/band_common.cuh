// band_common.cuh — what the multi-GPU step's key pass needs, shared by kernels_band.cuh (k_band_key, k_band_recv) and the
// force pass of kernels_sph.cuh, whose epilogue runs the NEXT tick's key pass on the position it is about to store
// ("speculation", as on one GPU). See kernels_band.cuh for the protocol.
#pragma once
#include "common.cuh"

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
constexpr uint32_t BAND_ERR_PEER = 8u;      // PEER transport: another rank's records for this tick did not appear within the time limit

struct BandRows {
    uint32_t bounds[BAND_MAX_RANKS + 1];  // rank r owns grid rows [bounds[r], bounds[r+1])
    int world, rank;
    uint32_t table_lo;                    // first grid row the band's cell table covers (extended rows + one empty row per side)
};

// header of one message (outbox -> inbox), BAND_HDR_WORDS words:
//   [0] records pushed (may exceed the window: the receiver clamps and latches BAND_ERR_WINDOW)
//   [1] the sender's latched error bits      [2] the sender's native count
//   [3] the exchange number this header belongs to, stored LAST (PEER transport: the receiver polls it over NVLink)
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
    uint32_t reserved0;
    uint32_t out_count[BAND_MAX_RANKS];  // key pass: records pushed per destination (reset by k_band_headers)
    uint32_t e_row[BAND_MAX_RANKS];      // key pass: natives whose new owner is rank r (reset by k_band_headers)
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

// what a band's key pass needs besides the tick's constants, resident on the device (one per band, written at creation)
struct BandSpec {
    BandRows rows;
    BandBoxes out;
    BandDyn* dyn;
    uint32_t key_lo, last_bin;
};

// key of a particle of this rank's arrays (sentinel if its row left the band's extended rows) + its histogram bin
// (warp-collective: every lane of the warp calls it, `valid` or not)
__device__ __forceinline__ uint32_t band_key_and_count(const DevParams& p, const BandRows& rows, float2 q, float y_h, bool valid,
                                                      uint32_t key_lo, uint32_t* __restrict__ count, uint32_t& row_out) {
    uint32_t k = 0, row = 0;
    bool keep = false;
    if (valid) {
        k = p.key_fast_div ? cell_key_shared(p, q, y_h) : cell_key(p, q);
        row = min(k, p.num_cells - 1u) / p.grid_w;
        keep = band_in_ext(rows, rows.rank, row);
    }
    row_out = row;
    atomic_count_run(count, keep ? k - key_lo : 0u, keep);
    return keep ? k : BAND_SENTINEL;
}

// a COPY of (pos, vel, origin) into the outbox of every other rank whose extended rows contain `row` — as an immigrant
// where that rank owns the row (counted in the migration matrix), as a ghost otherwise. Warp-collective.
__device__ __forceinline__ void band_push_copies(const BandRows& rows, const BandBoxes& out, BandDyn* __restrict__ dyn, bool valid,
                                                 uint32_t row, float2 ps, float2 vs, uint32_t origin) {
    const int me = rows.rank;
    const uint32_t lo = rows.bounds[me], hi = rows.bounds[me + 1];
    // rows [lo + 2, hi - 2) concern nobody else
    const bool edge = valid && !(row >= lo + BAND_GHOST_ROWS && row + BAND_GHOST_ROWS < hi);
    if (!__any_sync(0xffffffffu, edge)) return;
    const uint32_t lane = lane_id();
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
                band_box_origin(out.box[d], out.cap[d])[slot] = origin;
            }
        }
    }
}

// end of a key pass: one warp writes the message headers (record counts, this rank's row of the migration matrix) and resets
// the counters for the next tick. A kernel of its own (one CTA, launched before the exchange): letting the last CTA of the
// key pass do it cost one contended atomic per CTA — 0.4 ms per tick when the key pass is the epilogue of the force
// pass's 125,000 CTAs (32M particles).
__global__ void k_band_headers(const BandSpec* __restrict__ spec, uint32_t stamp) {
    pdl_enter();
    const BandRows rows = spec->rows;
    BandDyn* dyn = spec->dyn;
    const int me = rows.rank;
    const int d = threadIdx.x;
    if (d >= rows.world) return;
    const uint32_t n_nat = dyn->own_end - dyn->own_begin;
    const uint32_t cnt = dyn->out_count[d];
    dyn->out_count[d] = 0u;
    uint32_t err = dyn->error;
    if (d != me && cnt > spec->out.cap[d]) { err |= BAND_ERR_WINDOW; atomicOr(&dyn->error, BAND_ERR_WINDOW); }
    dyn->e_sent[d] = dyn->e_row[d];
    dyn->e_row[d] = 0u;
    __syncwarp((1u << rows.world) - 1u);
    if (d != me) {
        uint32_t* h = spec->out.box[d];
        h[0] = cnt; h[1] = err; h[2] = n_nat;
        for (int r = 0; r < BAND_MAX_RANKS; r++) h[4 + r] = r < rows.world ? dyn->e_sent[r] : 0u;
        // the records were stored by an earlier kernel of this stream (complete, hence in L2, which is where a peer's loads
        // over NVLink are served from); the header words above must be ordered before the stamp for an observer on another GPU
        __threadfence_system();
        *reinterpret_cast<volatile uint32_t*>(h + 3) = stamp;
    }
}

}  // namespace fluid
