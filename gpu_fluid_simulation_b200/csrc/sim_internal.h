// sim_internal.h — the state behind the opaque FluidSim handle.
#pragma once
#include "common.cuh"

namespace fluid {

#ifndef FLUID_SORT_ITEMS
#define FLUID_SORT_ITEMS 16
#endif
constexpr int SORT_ITEMS = FLUID_SORT_ITEMS;                          // keys per thread per onesweep tile
constexpr uint32_t SORT_TILE = SORT_THREADS * SORT_ITEMS;
constexpr int MAX_MARKS = 12;
constexpr int TIMING_RING = 256;

struct Sim {
    int device = 0;
    int num_sms = 148;
    cudaStream_t stream = nullptr;
    FluidSimSettings settings{};
    uint32_t n = 0, capacity = 0;
    uint32_t grid_w = 0, grid_h = 0, num_cells = 0;
    uint32_t sort_passes = 1;
    uint32_t tick = 0;             // reference: pub tick (src/simulation.rs:12)
    FluidSimUniform last_uniform{};

    // particle store, structure of arrays. *_a: state between ticks (input of a tick, output of
    // force+integrate); *_s: this tick's key-ordered snapshot the neighbour passes read.
    float2 *pos_a = nullptr, *vel_a = nullptr;
    float2 *pos_s = nullptr, *vel_s = nullptr, *pred_s = nullptr;
    float* rho = nullptr;
    uint32_t *keys_u = nullptr, *keys_b = nullptr, *keys_c = nullptr;  // unsorted keys, sort ping-pong
    uint32_t *idx_a = nullptr, *idx_b = nullptr;
    uint32_t *keys_sorted = nullptr, *idx_sorted = nullptr;            // aliases into the above
    uint32_t* sort_scratch = nullptr;
    uint4* nb_head = nullptr;           // per sorted particle, density pass -> force pass: starts of its three row spans + flags + first mask
    uint2* nb_mask = nullptr;           // ... and the other two acceptance masks (kernels_sph.cuh NB_*)
    void* tile_plans = nullptr;         // TilePlan per tile of SPH_THREADS sorted particles (kernels_sph.cuh), 64 B each
    size_t sort_scratch_words = 0;
    uint32_t max_tiles = 0;

    uint32_t* start_indices = nullptr;  // reference-compatible table (compute.wgsl:45-56), never cleared
    uint32_t* cell_lb = nullptr;        // dense lower-bound table, num_cells + 1 entries
    uint32_t* quirk = nullptr;          // [0] lowest occupied cell c0, [1] its stale start_indices entry

    // counting sort over the cells (kernels_count.cuh): per-cell counts / write cursors, zero-padded to whole scan tiles
    uint32_t* cell_count = nullptr;              // the array the current tick counts into (one of the two below)
    uint32_t* cell_count_buf[2] = {nullptr, nullptr};  // alternate per counted tick; the scan clears the other one
    int cell_count_cur = 0;
    uint32_t* cell_cursor = nullptr;             // small grids (k_cell_scan_small): the cursors live apart from the counts
    // speculation by the force pass (kernels_sph.cuh): keys (in keys_c) and per-cell counts (in the next count array)
    // of the NEXT tick, valid if that tick uses the same delta and the state has not been replaced in between
    bool spec_valid = false;    // the last enqueued tick's force pass speculated and nothing invalidated it since
    bool count_dirty = false;   // the next count array holds speculative counts (must be cleared if they are not used)
    float spec_delta = 0.0f;
    bool speculate = true;      // FLUID_NO_SPECULATION=1: off
    bool pdl = true;                             // programmatic dependent launch on the tick's kernels (FLUID_NO_PDL=1: off)
    int tile_order = -1;                         // neighbour passes: -1 by size, 0 index order, 1 from both ends (FLUID_TILE_ORDER=0|1)
    bool tiles_both_ends() const;                // which instantiation of the two neighbour kernels the tick launches
    uint32_t *cell_tile_sum = nullptr, *cell_tile_max = nullptr;
    float* cell_tile_sq = nullptr;
    uint32_t cell_tiles = 0;
    volatile uint32_t* cell_max_host = nullptr;  // pinned + mapped, 2 words: largest cell population and sum of squared populations (float bits) of the latest completed scan
    uint32_t* cell_max_dev = nullptr;            // device alias of cell_max_host
    bool cell_max_known = false;                 // false after anything that replaces the state (upload, resize)
    int sort_mode = FLUID_SORT_AUTO;
    int last_sort_path = FLUID_SORT_RADIX;       // what the most recently enqueued tick used
    uint32_t* d_frame_time = nullptr;   // device copy of `tick`, bumped by the first kernel of a tick
    float2* force_field = nullptr;      // reference `texture` (src/simulation.rs:213-218)

    FluidParticle* aos_stage = nullptr;
    // pipelined transfers (fluid_sim_upload_particles_async / _download_particles_async): own copy streams and staging
    // buffers, so that the upload of the next state overlaps the download of the previous result (PCIe is full duplex)
    cudaStream_t h2d_stream = nullptr, d2h_stream = nullptr;
    FluidParticle *aos_in = nullptr, *aos_out = nullptr;
    cudaEvent_t ev_h2d_done = nullptr, ev_in_free = nullptr, ev_out_ready = nullptr, ev_d2h_done = nullptr;
    bool in_used = false, out_used = false;
    float2 *dbg_accel = nullptr, *dbg_abs = nullptr;
    bool debug_force = false;
    bool exact_math = true;   // FLUID_MATH_EXACT / FLUID_MATH_FAST

    // a band of the multi-GPU step (kernels_band.cuh / multi.inl): the handle is owned by a FluidMulti
    bool band = false;
    uint32_t state_off = 0;    // index of the first owned particle in pos_a / vel_a (always 0)
    uint32_t sorted_off = 0;   // index of the first owned particle in pos_s / pred_s / vel_s / rho / sorted keys (band: set by a download)
    uint32_t* lb_view = nullptr;  // the lower-bound table as the neighbour passes index it (by global cell id): cell_lb, or a band's table shifted by its first cell

    // output buffers of the per-frame consumers (render.cu, dormant.cu), grown on demand and kept: a cudaMalloc / cudaFree
    // pair per frame synchronises the whole device
    void* frame_f32 = nullptr; void* frame_u8 = nullptr; size_t frame_px = 0;
    void* diag_out = nullptr; size_t diag_bytes = 0;
    cudaEvent_t ev_frame0 = nullptr, ev_frame1 = nullptr;
    uint32_t *jfa_a = nullptr, *jfa_b = nullptr;  // packed seed ping-pong for fluid_sim_jfa_from_particles
    size_t jfa_pixels = 0;
    uint8_t* mask_dev = nullptr;                  // obstacle mask at texture size (field.inl)
    uint32_t* mask_flag = nullptr;
    void* objects_dev = nullptr;                  // prepared FluidObjects
    uint32_t objects_cap = 0;

    // per-phase device timing: a ring of TIMING_RING ticks x MAX_MARKS events on `stream`
    bool timing = false;
    int n_marks = 0;                 // marks recorded in the tick being enqueued
    int marks_per_tick = 0;          // marks of the last completed enqueue
    uint64_t timed_ticks = 0;        // ticks enqueued since timing was enabled
    cudaEvent_t mark_events[TIMING_RING][MAX_MARKS] = {};
    const char* mark_names[MAX_MARKS] = {};
    uint64_t launches = 0;

    int alloc(uint32_t cap);
    void make_uniform(const FluidTickSettings& t, uint32_t tick_value, FluidSimUniform* out) const;
    DevParams make_params(const FluidSimUniform& u) const;
    void mark(const char* name);
    int enqueue_tick(const FluidTickSettings& t, bool hash_sort_only);
    uint32_t* keys_view() const { return keys_sorted + sorted_off; }
};

}  // namespace fluid
