// fluid_sim.cu — the FluidSim handle and the C ABI of include/fluid_b200.h.
//
// Host-side mirror of the reference's FluidSimulation (src/simulation.rs): owns the particle store
// (structure of arrays in HBM), the reference-compatible start_indices table, the force-field texture
// and one CUDA stream, and enqueues the per-tick kernel sequence in place of the reference's K1, K2, K3 x S(S+1)/2, K4, K5,
// K6 dispatches (src/simulation.rs:512-537). Default (counting sort over the cells, kernels_count.cuh), steady state:
//   cell scan (reduce, downsweep; one launch on small grids) -> slots -> rank+gather+cell tables+tile plans -> density
//   -> force+integrate (which also emits the NEXT tick's keys and per-cell counts) = six launches (five on small grids);
// the first tick after an upload or a change of delta starts with predict+key+count. Crowded cells or FLUID_SORT_RADIX:
//   predict+key+histogram -> scan -> onesweep x P -> gather+cell tables -> tile plans -> density -> force+integrate.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "common.cuh"
#include "kernels_cells.cuh"
#include "kernels_count.cuh"
#include "kernels_hash.cuh"
#include "kernels_sph.cuh"
#include "radix_sort.cuh"
#include "sim_internal.h"

namespace fluid {

static thread_local std::string g_last_error;
void set_error(const std::string& msg) { g_last_error = msg; }
int cuda_fail(cudaError_t e, const char* what, const char* file, int line) {
    char buf[512];
    snprintf(buf, sizeof buf, "CUDA error %d (%s) at %s:%d in %s", (int)e, cudaGetErrorString(e), file, line, what);
    g_last_error = buf;
    cudaGetLastError();
    return FLUID_ERR_CUDA;
}
static int invalid(const char* msg) {
    g_last_error = msg;
    return FLUID_ERR_INVALID_ARGUMENT;
}

// Rust f32::powi (compiler-rt __powisf2), used by the uniform derivation src/simulation.rs:486-490
static float powi_f32(float a, int b) {
    float r = 1.0f;
    for (;;) {
        if (b & 1) r *= a;
        b /= 2;
        if (b == 0) break;
        a *= a;
    }
    return r;
}

// ---- AoS <-> SoA transposition at the boundary (ParticleInstance, src/simulation.rs:126-135) ----
__global__ void k_aos_to_soa(const FluidParticle* __restrict__ aos, uint32_t n, float2* __restrict__ pos,
                             float2* __restrict__ pred, float2* __restrict__ vel, float* __restrict__ rho,
                             uint32_t* __restrict__ keys) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4* src = reinterpret_cast<const float4*>(aos + i);
    float4 lo = __ldg(src), hi = __ldg(src + 1);
    pos[i] = make_float2(lo.x, lo.y);
    pred[i] = make_float2(lo.z, lo.w);
    vel[i] = make_float2(hi.x, hi.y);
    rho[i] = hi.z;
    keys[i] = __float_as_uint(hi.w);
}
__global__ void k_soa_to_aos(FluidParticle* __restrict__ aos, uint32_t n, const float2* __restrict__ pos,
                             const float2* __restrict__ pred, const float2* __restrict__ vel,
                             const float* __restrict__ rho, const uint32_t* __restrict__ keys) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float2 a = pos[i], b = pred[i], c = vel[i];
    float4* dst = reinterpret_cast<float4*>(aos + i);
    dst[0] = make_float4(a.x, a.y, b.x, b.y);
    dst[1] = make_float4(c.x, c.y, rho[i], __uint_as_float(keys[i]));
}

static uint32_t bits_for(uint32_t max_value) {
    uint32_t b = 0;
    while (b < 32 && (max_value >> b) != 0) b++;
    return b == 0 ? 1 : b;
}

int Sim::alloc(uint32_t cap) {
    // free old particle-sized buffers, allocate new; contents are NOT preserved
    void** bufs[] = {(void**)&pos_a, (void**)&vel_a, (void**)&pos_s, (void**)&vel_s, (void**)&pred_s, (void**)&rho,
                     (void**)&keys_u, (void**)&keys_b, (void**)&keys_c, (void**)&idx_a, (void**)&idx_b,
                     (void**)&sort_scratch, (void**)&aos_stage, (void**)&dbg_accel, (void**)&dbg_abs, (void**)&tile_plans,
                     (void**)&nb_head, (void**)&nb_mask};
    for (void** b : bufs) { if (*b) cudaFree(*b); *b = nullptr; }
    capacity = cap;
    const size_t padded = (size_t)cap + 64;
    FLUID_CUDA(cudaMalloc(&pos_a, padded * 8)); FLUID_CUDA(cudaMalloc(&vel_a, padded * 8));
    FLUID_CUDA(cudaMalloc(&pos_s, padded * 8)); FLUID_CUDA(cudaMalloc(&vel_s, padded * 8));
    FLUID_CUDA(cudaMalloc(&pred_s, padded * 8)); FLUID_CUDA(cudaMalloc(&rho, padded * 4));
    FLUID_CUDA(cudaMalloc(&keys_u, padded * 4)); FLUID_CUDA(cudaMalloc(&keys_b, padded * 4));
    FLUID_CUDA(cudaMalloc(&keys_c, padded * 4));
    FLUID_CUDA(cudaMalloc(&idx_a, padded * 4)); FLUID_CUDA(cudaMalloc(&idx_b, padded * 4));
    void* zero_me[] = {pos_a, vel_a, pos_s, vel_s, pred_s};
    for (void* z : zero_me) FLUID_CUDA(cudaMemsetAsync(z, 0, padded * 8, stream));
    void* zero_me4[] = {rho, keys_u, keys_b, keys_c, idx_a, idx_b};
    for (void* z : zero_me4) FLUID_CUDA(cudaMemsetAsync(z, 0, padded * 4, stream));
    max_tiles = (cap + SORT_TILE - 1) / SORT_TILE;
    // [hist MAX_PASSES*256][tile counters MAX_PASSES, padded to 256][status MAX_PASSES*max_tiles*256]
    sort_scratch_words = (size_t)MAX_PASSES * RADIX + RADIX + (size_t)MAX_PASSES * max_tiles * RADIX;
    FLUID_CUDA(cudaMalloc(&sort_scratch, sort_scratch_words * 4));
    FLUID_CUDA(cudaMalloc(&tile_plans, ((size_t)cap / SPH_THREADS + 2) * sizeof(TilePlan)));
    if (padded >= (1ull << 30)) { set_error("capacity must stay below 2^30 particles"); return FLUID_ERR_INVALID_ARGUMENT; }
    FLUID_CUDA(cudaMalloc(&nb_head, padded * sizeof(uint4)));
    FLUID_CUDA(cudaMalloc(&nb_mask, padded * sizeof(uint2)));
    keys_sorted = keys_b;
    return FLUID_OK;
}

void Sim::make_uniform(const FluidTickSettings& t, uint32_t tick_value, FluidSimUniform* o) const {
    // src/simulation.rs:470-497
    const float PI_RS = 3.14159274101257324219f;
    const float h = settings.smoothing_radius;
    memset(o, 0, sizeof *o);
    o->delta = t.delta;
    o->particle_count = n;
    o->sqr_radius = h * h;
    o->frame_time = tick_value;
    o->gravity[0] = t.gravity[0]; o->gravity[1] = t.gravity[1];
    o->bounds[0] = settings.size[0]; o->bounds[1] = settings.size[1];
    o->mouse_pos[0] = t.mouse_pos[0]; o->mouse_pos[1] = t.mouse_pos[1];
    o->smoothing_radius = h;
    o->particle_mass = t.mass;
    o->pressure_constant = t.pressure_constant;
    o->rest_density = t.rest_density;
    o->damping_factor = t.damping_factor;
    o->viscosity_coefficient = t.viscosity_coefficient;
    o->surface_tension_treshold = t.surface_tension_treshold;
    o->surface_tension_coefficient = t.surface_tension_coefficient;
    o->poly6_kernel_volume = 4.0f / (PI_RS * powi_f32(h, 8));
    o->poly6_kernel_derivative = 24.0f / (PI_RS * powi_f32(h, 8));
    o->poly6_kernel_laplacian = 8.0f / (PI_RS * powi_f32(h, 8));
    o->spiky_kernel_derivative = 12.0f / (powi_f32(h, 4) * PI_RS);
    o->viscosity_kernel = 15.0f / (2.0f * PI_RS * powi_f32(h, 3));
    o->mouse_state = t.mouse_state;
    o->mouse_force_radius = t.mouse_force_radius;
    o->mouse_force_power = t.mouse_force_power;
    o->grid_w = grid_w;
    o->grid_h = grid_h;
    o->texture_size[0] = (float)settings.texture_size[0];
    o->texture_size[1] = (float)settings.texture_size[1];
}

// tiles from both ends of the key order while the neighbour passes are a few dozen waves of CTAs (kernels_sph.cuh
// tile_from_both_ends); measured: 1M dam break -13 % early, -21 % after 320 ticks; no gain on the 64M uniform box
constexpr uint32_t TILES_BOTH_ENDS_MAX = 16384;  // 4M particles: ~28 waves of 592 resident CTAs (balanced scenes pay ~2 % for the order: 1M-U, 8M-particle bands)
// ... and only while some cell is crowded: the scan publishes the largest cell population into pinned host memory every
// counted tick (read here without synchronising: it lags by the ticks in flight, which only delays the switch). A balanced
// scene (1M-U: largest cell ~8 of a mean of 4) keeps the index order, which is 2 % faster there.
constexpr uint32_t TILES_BOTH_ENDS_MIN_CELL = 10;  // (1M-D crosses it ~35 ticks in, 1M-U never: its largest cell stays at 8)
bool Sim::tiles_both_ends() const {
    if (tile_order >= 0) return tile_order != 0;
    return capacity / SPH_THREADS <= TILES_BOTH_ENDS_MAX && cell_max_host && cell_max_host[0] >= TILES_BOTH_ENDS_MIN_CELL;
}
DevParams Sim::make_params(const FluidSimUniform& u) const {
    DevParams p;
    memset(&p, 0, sizeof p);
    const float PI_WGSL = 3.14159265359f;  // funcs.wgsl:54
    p.n = n;
    p.grid_w = u.grid_w; p.grid_h = u.grid_h; p.num_cells = u.grid_w * u.grid_h;
    p.dt = u.delta;
    p.bounds_x = u.bounds[0]; p.bounds_y = u.bounds[1];
    p.half_x = u.bounds[0] * 0.5f; p.half_y = u.bounds[1] * 0.5f;
    p.h = u.smoothing_radius;
    p.h2 = u.smoothing_radius * u.smoothing_radius;
    p.mass = u.particle_mass;
    p.poly6_norm = 4.0f / (PI_WGSL * powf(u.smoothing_radius, 8.0f));  // funcs.wgsl:76
    p.pressure_constant = u.pressure_constant; p.rest_density = u.rest_density;
    p.spiky = u.spiky_kernel_derivative;
    p.visc_k = u.viscosity_kernel;
    p.visc_coef = u.viscosity_coefficient;
    p.two_h3 = 2.0f * u.smoothing_radius * u.smoothing_radius * u.smoothing_radius;
    p.gravity_dt_x = u.gravity[0] * u.delta; p.gravity_dt_y = u.gravity[1] * u.delta;
    p.damping = u.damping_factor;
    p.one_minus_damping = 1.0f - u.damping_factor;
    p.neg_damping = -1.0f * u.damping_factor;
    p.mouse_state = u.mouse_state;
    p.mouse_x = u.mouse_pos[0]; p.mouse_y = u.mouse_pos[1];
    p.mouse_radius = u.mouse_force_radius; p.mouse_power = u.mouse_force_power;
    p.mouse_state_f = (float)u.mouse_state;
    p.tex_w = u.texture_size[0]; p.tex_h = u.texture_size[1];
    p.tex_wu = settings.texture_size[0]; p.tex_hu = settings.texture_size[1];
    p.tex_len = p.tex_wu * p.tex_hu;
    p.p2w_x = (u.bounds[0] * 2.0f) / u.texture_size[0];
    p.p2w_y = (u.bounds[1] * 2.0f) / u.texture_size[1];
    p.inv_two_h3 = 1.0f / p.two_h3;
    p.inv_h2 = 1.0f / p.h2;
    p.exact_math = exact_math ? 1 : 0;
    {
        const float lo_h = 9.5367431640625e-07f /* 2^-20 */, hi_h = 1048576.0f /* 2^20 */, lo_b = 8.6736174e-19f /* 2^-60 */, hi_b = 6.338253e29f /* 2^99 */;
        p.key_fast_div = (p.h >= lo_h && p.h <= hi_h && p.half_x >= lo_b && p.half_x <= hi_b && p.half_y >= lo_b && p.half_y <= hi_b) ? 1 : 0;
    }
    return p;
}

void Sim::mark(const char* name) {
    if (!timing) return;
    if (n_marks >= MAX_MARKS) return;
    cudaEvent_t& ev = mark_events[timed_ticks % TIMING_RING][n_marks];
    if (!ev) cudaEventCreate(&ev);
    cudaEventRecord(ev, stream);
    mark_names[n_marks] = name;
    n_marks++;
}

// Launch with programmatic stream serialization: the kernel's CTAs may become resident while the previous kernel of the
// stream drains; every kernel launched this way starts with pdl_enter() (common.cuh), which waits for the previous grid
// to complete before touching memory. `pdl == false` is a plain launch.
template <typename... KArgs, typename... Args>
static void launch_pdl(bool pdl, void (*kernel)(KArgs...), dim3 grid, dim3 block, cudaStream_t stream, Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = 0;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// staging plans of the two neighbour passes (kernels_sph.cuh TilePlan), once per tick after the cell table is complete
void launch_tile_plans(Sim* s, const DevParams& p, const uint32_t* keys, bool pdl = false) {
    const uint32_t ntiles = (p.n + SPH_THREADS - 1) / SPH_THREADS;
    if (!ntiles) return;
    launch_pdl(pdl, k_tile_plans, dim3((ntiles + 127) / 128), dim3(128), s->stream, p, keys, s->cell_lb, ntiles,
               reinterpret_cast<TilePlan*>(s->tile_plans));
    s->launches++;
}

void launch_force(Sim* s, const DevParams& p, uint32_t blocks, float2* pos_out, float2* vel_out, bool pdl = false,
                  uint32_t* next_keys = nullptr, uint32_t* next_count = nullptr, const uint32_t* band_range = nullptr,
                  const uint32_t* goff_dev = nullptr, const BandSpec* band_spec = nullptr) {
    float2* da = s->debug_force ? s->dbg_accel : nullptr;
    float2* db = s->debug_force ? s->dbg_abs : nullptr;
    const uint32_t* keys = s->keys_sorted;
#define FLUID_LAUNCH_FORCE(DBG, EX, BE)                                                                                 \
    launch_pdl(pdl, k_force_integrate<DBG, EX, false, BE>, dim3(blocks), dim3(SPH_THREADS), s->stream, p, s->pos_s, s->pred_s, s->vel_s, \
               s->rho, keys, s->lb_view, s->quirk, s->force_field, s->d_frame_time, goff_dev, s->quirk + 2,             \
               reinterpret_cast<const TilePlan*>(s->tile_plans), pos_out, vel_out, da, db, next_keys, next_count, band_range, band_spec,      \
               (const uint4*)s->nb_head, (const uint2*)s->nb_mask)
#define FLUID_LAUNCH_FORCE_BAND(EX, BE)                                                                                 \
    launch_pdl(pdl, k_force_integrate<false, EX, true, BE>, dim3(blocks), dim3(SPH_THREADS), s->stream, p, s->pos_s, s->pred_s, s->vel_s, \
               s->rho, keys, s->lb_view, s->quirk, s->force_field, s->d_frame_time, goff_dev, s->quirk + 2,             \
               reinterpret_cast<const TilePlan*>(s->tile_plans), pos_out, vel_out, da, db, next_keys, next_count, band_range, band_spec,      \
               (const uint4*)s->nb_head, (const uint2*)s->nb_mask)
    const bool ex = s->exact_math, be = s->tiles_both_ends();
    if (band_range) {
        if (ex) { if (be) FLUID_LAUNCH_FORCE_BAND(true, true); else FLUID_LAUNCH_FORCE_BAND(true, false); }
        else { if (be) FLUID_LAUNCH_FORCE_BAND(false, true); else FLUID_LAUNCH_FORCE_BAND(false, false); }
    } else if (s->debug_force) {  // (diagnostic path: index order only)
        if (ex) FLUID_LAUNCH_FORCE(true, true, false); else FLUID_LAUNCH_FORCE(true, false, false);
    } else {
        if (ex) { if (be) FLUID_LAUNCH_FORCE(false, true, true); else FLUID_LAUNCH_FORCE(false, true, false); }
        else { if (be) FLUID_LAUNCH_FORCE(false, false, true); else FLUID_LAUNCH_FORCE(false, false, false); }
    }
#undef FLUID_LAUNCH_FORCE
#undef FLUID_LAUNCH_FORCE_BAND
    s->launches++;
}

template <int PASSES>
static void launch_hash(Sim* s, const DevParams& p, uint32_t* hist, uint32_t bump) {
    const uint32_t pairs = (p.n + 1) / 2;
    uint32_t blocks = (pairs + HASH_THREADS - 1) / HASH_THREADS;
    const uint32_t max_blocks = (uint32_t)s->num_sms * 8u;
    if (blocks > max_blocks) blocks = max_blocks;
    k_predict_key_hist<PASSES><<<blocks, HASH_THREADS, 0, s->stream>>>(p, s->pos_a, s->vel_a, s->keys_u, hist, s->d_frame_time, bump);
}

int Sim::enqueue_tick(const FluidTickSettings& t, bool hash_sort_only) {
    tick += 1;  // src/simulation.rs:460
    make_uniform(t, tick, &last_uniform);
    const DevParams p = make_params(last_uniform);
    n_marks = 0;
    mark("begin");

    // ---- sort path of this tick (kernels_count.cuh): counting sort over the cells unless a cell is too crowded ----
    const uint32_t gblocks = (n + GATHER_THREADS - 1) / GATHER_THREADS;
    const bool counted = sort_mode != FLUID_SORT_RADIX;
    bool counting = false;
    uint32_t* cursor = nullptr;
    uint32_t* count_next = nullptr;
    uint32_t bump_in_scan = 0u;
    const bool had_spec = spec_valid;
    spec_valid = false;  // consumed or dropped by this tick; set again below if this tick's force pass speculates
    if (counted) {
        // the two count arrays alternate: this tick's was cleared by the previous counted tick's scan (or at create)
        cell_count = cell_count_buf[cell_count_cur];
        count_next = cell_count_buf[cell_count_cur ^ 1];
        cursor = cell_count;  // the two-kernel scan turns the counts into the cursors in place
        cell_count_cur ^= 1;
        // did the previous tick's force pass already emit this tick's keys and counts (same delta, state untouched)?
        const bool hit = had_spec && memcmp(&spec_delta, &t.delta, sizeof(float)) == 0;
        if (hit) {
            std::swap(keys_u, keys_c);  // the speculative keys become this tick's unsorted keys
            bump_in_scan = 1u;
        } else {
            if (count_dirty) FLUID_CUDA(cudaMemsetAsync(cell_count, 0, (size_t)cell_tiles * CELL_SCAN_TILE * 4, stream));  // unused speculation
            const uint32_t pairs = (n + 1) / 2;
            uint32_t blocks = (pairs + HASH_THREADS - 1) / HASH_THREADS;
            const uint32_t max_blocks = (uint32_t)num_sms * 8u;
            if (blocks > max_blocks) blocks = max_blocks;
            launch_pdl(pdl, k_predict_key_count, dim3(blocks), dim3(HASH_THREADS), stream, p, pos_a, vel_a, keys_u, cell_count, d_frame_time);
            launches++;
        }
        count_dirty = false;
        mark("predict+key+hist");
        if (cell_tiles <= CELL_SCAN_SMALL_MAX_TILES) {
            launch_pdl(pdl, k_cell_scan_small, dim3(cell_tiles), dim3(CELL_SCAN_SMALL_THREADS), stream, cell_count, cell_cursor,
                       cell_lb, cell_max_dev, count_next, d_frame_time, bump_in_scan);
            launches += 1;
            cursor = cell_cursor;
        } else {
            launch_pdl(pdl, k_cell_reduce, dim3(cell_tiles), dim3(CELL_SCAN_THREADS), stream, cell_count, cell_tile_sum, cell_tile_max, cell_tile_sq,
                       d_frame_time, bump_in_scan);
            launch_pdl(pdl, k_cell_downsweep, dim3(cell_tiles), dim3(CELL_SCAN_THREADS), stream, cell_count, cell_lb, cell_tile_sum,
                       cell_tile_max, cell_tile_sq, cell_tiles, cell_max_dev, count_next);
            launches += 2;
        }
        if (sort_mode == FLUID_SORT_COUNTING) {
            counting = true;
        } else {
            // the published maximum lags the stream by the ticks in flight; after an upload it is unknown: wait once
            if (!cell_max_known) {
                FLUID_CUDA(cudaStreamSynchronize(stream));
                cell_max_known = true;
            }
            const uint32_t max_ok = std::min(std::max(n / 1024u, COUNT_SORT_MAX_CELL_LO), COUNT_SORT_MAX_CELL_HI);
            const uint32_t sq_bits = cell_max_host[1];
            float sum_sq;
            memcpy(&sum_sq, &sq_bits, 4);
            counting = cell_max_host[0] <= max_ok && sum_sq <= COUNT_SORT_MAX_MEAN_SQ * (float)n;
        }
    }
    if (counting) {
        launch_pdl(pdl, k_cell_slots, dim3((n + GATHER_THREADS * SLOT_ITEMS - 1) / (GATHER_THREADS * SLOT_ITEMS)), dim3(GATHER_THREADS),
                   stream, n, 0u, p.num_cells - 1u, keys_u, cursor, keys_b, idx_a, (const uint32_t*)nullptr);
        launches++;
        mark("sort");
        launch_pdl(pdl, k_rank_gather_cells, dim3((n + GATHER_THREADS * RANK_ITEMS - 1) / (GATHER_THREADS * RANK_ITEMS)),
                   dim3(GATHER_THREADS), stream, p, keys_b, idx_a, pos_a, vel_a, pos_s, pred_s, vel_s, start_indices, cell_lb, quirk,
                   d_frame_time, reinterpret_cast<TilePlan*>(tile_plans));
        launches++;
        keys_sorted = keys_b;
        idx_sorted = idx_b;  // filled on request: fluid_sim_download_permutation
        last_sort_path = FLUID_SORT_COUNTING;
    } else {
        const uint32_t passes = sort_passes;
        const uint32_t tiles = (n + SORT_TILE - 1) / SORT_TILE;
        uint32_t* hist = sort_scratch;
        uint32_t* counters = sort_scratch + MAX_PASSES * RADIX;
        uint32_t* status = counters + RADIX;
        // one memset clears histograms, tile tickets and the look-back status words of the passes in use
        FLUID_CUDA(cudaMemsetAsync(sort_scratch, 0, ((size_t)MAX_PASSES * RADIX + RADIX + (size_t)passes * tiles * RADIX) * 4, stream));

        const uint32_t bump = counted ? 0u : 1u;  // k_predict_key_count has already advanced frame_time
        switch (passes) {
            case 1: launch_hash<1>(this, p, hist, bump); break;
            case 2: launch_hash<2>(this, p, hist, bump); break;
            case 3: launch_hash<3>(this, p, hist, bump); break;
            default: launch_hash<4>(this, p, hist, bump); break;
        }
        launches++;
        if (!counted) mark("predict+key+hist");
        k_scan_histograms<<<passes, RADIX, 0, stream>>>(hist);
        launches++;

        const size_t sort_smem = (size_t)SORT_TILE * 2 * sizeof(uint32_t);
        const uint32_t* kin = keys_u;
        const uint32_t* vin = nullptr;
        uint32_t* kout = keys_b;
        uint32_t* vout = idx_a;
        for (uint32_t q = 0; q < passes; q++) {
            if (q == 0)
                k_onesweep_pass<SORT_ITEMS, true><<<tiles, SORT_THREADS, sort_smem, stream>>>(
                    kin, vin, kout, vout, hist + q * RADIX, counters + q, status + (size_t)q * tiles * RADIX, n, (int)(q * RADIX_BITS));
            else
                k_onesweep_pass<SORT_ITEMS, false><<<tiles, SORT_THREADS, sort_smem, stream>>>(
                    kin, vin, kout, vout, hist + q * RADIX, counters + q, status + (size_t)q * tiles * RADIX, n, (int)(q * RADIX_BITS));
            launches++;
            kin = kout; vin = vout;
            kout = (kout == keys_b) ? keys_c : keys_b;
            vout = (vout == idx_a) ? idx_b : idx_a;
        }
        keys_sorted = const_cast<uint32_t*>(kin);
        idx_sorted = const_cast<uint32_t*>(vin);
        mark("sort");

        k_gather_cells<<<gblocks, GATHER_THREADS, 0, stream>>>(p, keys_sorted, idx_sorted, pos_a, vel_a, pos_s, pred_s, vel_s,
                                                             start_indices, cell_lb, quirk, d_frame_time);
        launches++;
        last_sort_path = FLUID_SORT_RADIX;
    }
    mark("gather+cells");
    if (hash_sort_only) {
        // state stays consistent for download: positions/velocities unchanged, in sorted order
        FLUID_CUDA(cudaMemcpyAsync(pos_a, pos_s, (size_t)n * 8, cudaMemcpyDeviceToDevice, stream));
        FLUID_CUDA(cudaMemcpyAsync(vel_a, vel_s, (size_t)n * 8, cudaMemcpyDeviceToDevice, stream));
        mark("copy");
        if (timing) { marks_per_tick = n_marks; timed_ticks++; }
        FLUID_CUDA(cudaGetLastError());
        return FLUID_OK;
    }
    const uint32_t sblocks = (n + SPH_THREADS - 1) / SPH_THREADS;
    if (!counting) launch_tile_plans(this, p, keys_sorted, pdl);  // counting path: made by k_rank_gather_cells
    if (tiles_both_ends())
        launch_pdl(pdl, k_density<false, true>, dim3(sblocks), dim3(SPH_THREADS), stream, p, pred_s, keys_sorted, cell_lb, quirk, rho, quirk + 2,
                   d_frame_time, reinterpret_cast<const TilePlan*>(tile_plans), (const uint32_t*)nullptr, nb_head, nb_mask);
    else
        launch_pdl(pdl, k_density<false, false>, dim3(sblocks), dim3(SPH_THREADS), stream, p, pred_s, keys_sorted, cell_lb, quirk, rho, quirk + 2,
                   d_frame_time, reinterpret_cast<const TilePlan*>(tile_plans), (const uint32_t*)nullptr, nb_head, nb_mask);
    launches++;
    mark("density");
    // the force pass emits the next tick's keys (into keys_c: free on the counting path, and the radix passes of THIS tick are
    // done with it by now) and counts (into the count array the next counted tick will use, cleared by this tick's scan)
    const bool spec = speculate && counted && count_next != nullptr && keys_sorted != keys_c;  // (radix passes of an even count end in keys_c)
    launch_force(this, p, sblocks, pos_a, vel_a, pdl, spec ? keys_c : nullptr, spec ? count_next : nullptr);
    if (spec) { spec_valid = true; count_dirty = true; spec_delta = t.delta; }
    mark("force+integrate");
    if (timing) { marks_per_tick = n_marks; timed_ticks++; }
    FLUID_CUDA(cudaGetLastError());
    return FLUID_OK;
}

}  // namespace fluid

using namespace fluid;

extern "C" {

const char* fluid_last_error(void) { return g_last_error.c_str(); }
int fluid_abi_version(void) { return FLUID_B200_ABI_VERSION; }
int fluid_device_count(void) {
    int c = 0;
    cudaError_t e = cudaGetDeviceCount(&c);
    if (e != cudaSuccess) { cuda_fail(e, "cudaGetDeviceCount", __FILE__, __LINE__); return FLUID_ERR_NO_DEVICE; }
    return c;
}

int fluid_sim_upload_particles(FluidSim* h, const FluidParticle* src);
void fluid_sim_destroy(FluidSim* h);

}  // extern "C"

namespace fluid {
// the handle behind fluid_sim_create; a band of the multi-GPU step (multi.inl) is the same object with room for
// `capacity` particles and no initial lattice
int sim_create_internal(const FluidSimSettings* s, int device_ordinal, uint32_t capacity, bool lattice, Sim** out) {
    *out = nullptr;
    if (!(s->smoothing_radius > 0.0f) || !(s->size[0] > 0.0f) || !(s->size[1] > 0.0f))
        return invalid("fluid_sim_create: smoothing_radius and size must be positive");
    if (s->texture_size[0] == 0 || s->texture_size[1] == 0) return invalid("fluid_sim_create: texture_size must be non-zero");
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) {
        cudaGetLastError();
        set_error("fluid_sim_create: no CUDA device available (this library has no CPU fallback)");
        return FLUID_ERR_NO_DEVICE;
    }
    if (device_ordinal < 0 || device_ordinal >= count) { set_error("fluid_sim_create: device ordinal out of range"); return FLUID_ERR_NO_DEVICE; }
    FLUID_CUDA(cudaSetDevice(device_ordinal));
    cudaDeviceProp prop;
    FLUID_CUDA(cudaGetDeviceProperties(&prop, device_ordinal));
    if (prop.major < 10) { set_error("fluid_sim_create: device is not sm_100+ (library is built for sm_100a only)"); return FLUID_ERR_NO_DEVICE; }

    Sim* sim = new Sim();
    sim->device = device_ordinal;
    sim->num_sms = prop.multiProcessorCount;
    sim->settings = *s;
    sim->n = lattice ? s->particle_count : 0u;
    // src/simulation.rs:140-141
    sim->grid_w = (uint32_t)(size_t)ceilf(s->size[0] / s->smoothing_radius) + 2u;
    sim->grid_h = (uint32_t)(size_t)ceilf(s->size[1] / s->smoothing_radius) + 2u;
    const uint64_t cells64 = (uint64_t)sim->grid_w * sim->grid_h;
    if (cells64 >= (1ull << 31)) { delete sim; return invalid("fluid_sim_create: grid has too many cells for u32 keys"); }
    sim->num_cells = (uint32_t)cells64;
    sim->sort_passes = (bits_for(sim->num_cells - 1) + RADIX_BITS - 1) / RADIX_BITS;
    int rc = FLUID_OK;
    auto fail = [&](int code) { fluid_sim_destroy(reinterpret_cast<FluidSim*>(sim)); return code; };
    if (cudaStreamCreateWithFlags(&sim->stream, cudaStreamNonBlocking) != cudaSuccess) return fail(cuda_fail(cudaGetLastError(), "cudaStreamCreate", __FILE__, __LINE__));
    if ((rc = sim->alloc(capacity)) != FLUID_OK) return fail(rc);
    const size_t tex_len = (size_t)s->texture_size[0] * s->texture_size[1];
    // cell tables padded to whole scan tiles (kernels_count.cuh); the pad of cell_lb includes entry [num_cells]
    sim->cell_tiles = (uint32_t)(((size_t)sim->num_cells + 1 + CELL_SCAN_TILE - 1) / CELL_SCAN_TILE);
    const size_t cells_padded = (size_t)sim->cell_tiles * CELL_SCAN_TILE;
    if (cudaMalloc(&sim->start_indices, (size_t)sim->num_cells * 4) != cudaSuccess ||
        cudaMalloc(&sim->cell_lb, cells_padded * 4) != cudaSuccess ||
        cudaMalloc(&sim->cell_count_buf[0], cells_padded * 4) != cudaSuccess ||
        cudaMalloc(&sim->cell_count_buf[1], cells_padded * 4) != cudaSuccess ||
        (sim->cell_tiles <= CELL_SCAN_SMALL_MAX_TILES && cudaMalloc(&sim->cell_cursor, cells_padded * 4) != cudaSuccess) ||
        cudaMalloc(&sim->cell_tile_sum, (size_t)sim->cell_tiles * 4) != cudaSuccess ||
        cudaMalloc(&sim->cell_tile_max, (size_t)sim->cell_tiles * 4) != cudaSuccess ||
        cudaMalloc(&sim->cell_tile_sq, (size_t)sim->cell_tiles * 4) != cudaSuccess ||
        cudaHostAlloc((void**)&sim->cell_max_host, 8, cudaHostAllocMapped) != cudaSuccess ||
        cudaHostGetDevicePointer((void**)&sim->cell_max_dev, (void*)sim->cell_max_host, 0) != cudaSuccess ||
        cudaMalloc(&sim->quirk, 16) != cudaSuccess || cudaMalloc(&sim->d_frame_time, 4) != cudaSuccess ||
        cudaMalloc(&sim->force_field, tex_len * 8) != cudaSuccess)
        return fail(cuda_fail(cudaGetLastError(), "cudaMalloc(tables)", __FILE__, __LINE__));
    // zero-initialised like wgpu's un-mapped buffers (src/simulation.rs:204-218)
    cudaMemsetAsync(sim->start_indices, 0, (size_t)sim->num_cells * 4, sim->stream);
    cudaMemsetAsync(sim->cell_lb, 0, cells_padded * 4, sim->stream);
    cudaMemsetAsync(sim->cell_count_buf[0], 0, cells_padded * 4, sim->stream);
    cudaMemsetAsync(sim->cell_count_buf[1], 0, cells_padded * 4, sim->stream);
    sim->cell_max_host[0] = 0;
    sim->cell_max_host[1] = 0;
    sim->lb_view = sim->cell_lb;
    {
        const char* no_pdl = getenv("FLUID_NO_PDL");
        sim->pdl = !(no_pdl && no_pdl[0] == '1');
        const char* no_spec = getenv("FLUID_NO_SPECULATION");
        sim->speculate = !(no_spec && no_spec[0] == '1');
        const char* order = getenv("FLUID_TILE_ORDER");
        if (order && (order[0] == '0' || order[0] == '1')) sim->tile_order = order[0] - '0';
    }
    cudaMemsetAsync(sim->quirk, 0, 16, sim->stream);
    cudaMemsetAsync(sim->d_frame_time, 0, 4, sim->stream);
    cudaMemsetAsync(sim->force_field, 0, tex_len * 8, sim->stream);
    cudaFuncSetAttribute(k_onesweep_pass<SORT_ITEMS, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, SORT_TILE * 8);
    cudaFuncSetAttribute(k_onesweep_pass<SORT_ITEMS, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, SORT_TILE * 8);

    // init lattice, src/simulation.rs:147-163 (host arithmetic, uploaded once like create_buffer_init :168-172)
    if (lattice) {
        const uint32_t n = sim->n;
        std::vector<FluidParticle> host(n);
        const float ppr = sqrtf((float)n);
        const float ppc = ((float)n - 1.0f) / ppr + 1.0f;
        const size_t ppr_u = (size_t)ppr;
        for (uint32_t i = 0; i < n; i++) {
            const size_t xi = (size_t)i % ppr_u;
            const float x = ((float)xi - ppr * 0.5f + 0.5f) * s->particle_spacing;
            const float y = (floorf((float)i / ppr) - ppc * 0.5f + 0.5f) * s->particle_spacing;
            FluidParticle q;
            memset(&q, 0, sizeof q);
            q.position[0] = x; q.position[1] = y;
            q.predicted_position[0] = x; q.predicted_position[1] = y;
            host[i] = q;
        }
        if ((rc = fluid_sim_upload_particles(reinterpret_cast<FluidSim*>(sim), host.data())) != FLUID_OK) return fail(rc);
    }
    FluidTickSettings zero_tick;
    memset(&zero_tick, 0, sizeof zero_tick);
    sim->make_uniform(zero_tick, 0, &sim->last_uniform);
    if (cudaStreamSynchronize(sim->stream) != cudaSuccess) return fail(cuda_fail(cudaGetLastError(), "sync after create", __FILE__, __LINE__));
    *out = sim;
    return FLUID_OK;
}
}  // namespace fluid

extern "C" {

int fluid_sim_create(const FluidSimSettings* s, int device_ordinal, FluidSim** out) {
    if (!s || !out) return invalid("fluid_sim_create: null argument");
    *out = nullptr;
    if (s->particle_count < 2) return invalid("fluid_sim_create: particle_count < 2 (the reference panics: (0).ilog2(), src/simulation.rs:323-324)");
    Sim* sim = nullptr;
    int rc = sim_create_internal(s, device_ordinal, s->particle_count, true, &sim);
    if (rc != FLUID_OK) return rc;
    *out = reinterpret_cast<FluidSim*>(sim);
    return FLUID_OK;
}


void fluid_sim_destroy(FluidSim* h) {
    if (!h) return;
    Sim* s = reinterpret_cast<Sim*>(h);
    cudaSetDevice(s->device);
    if (s->stream) cudaStreamSynchronize(s->stream);
    void* bufs[] = {s->pos_a, s->vel_a, s->pos_s, s->vel_s, s->pred_s, s->rho, s->keys_u, s->keys_b, s->keys_c, s->idx_a,
                    s->idx_b, s->sort_scratch, s->tile_plans, s->nb_head, s->nb_mask, s->aos_stage, s->dbg_accel, s->dbg_abs, s->start_indices, s->cell_lb, s->quirk,
                    s->d_frame_time, s->force_field, s->jfa_a, s->jfa_b,
                    s->mask_dev, s->mask_flag, s->objects_dev, s->cell_count_buf[0], s->cell_count_buf[1], s->cell_tile_sum, s->cell_tile_max, s->cell_tile_sq, s->cell_cursor};
    for (void* b : bufs) if (b) cudaFree(b);
    for (void* b : {s->frame_f32, s->frame_u8, s->diag_out}) if (b) cudaFree(b);
    for (cudaEvent_t e : {s->ev_frame0, s->ev_frame1}) if (e) cudaEventDestroy(e);
    if (s->cell_max_host) cudaFreeHost((void*)s->cell_max_host);
    if (s->h2d_stream) { cudaStreamSynchronize(s->h2d_stream); cudaStreamDestroy(s->h2d_stream); }
    if (s->d2h_stream) { cudaStreamSynchronize(s->d2h_stream); cudaStreamDestroy(s->d2h_stream); }
    if (s->aos_in) cudaFree(s->aos_in);
    if (s->aos_out) cudaFree(s->aos_out);
    for (cudaEvent_t e : {s->ev_h2d_done, s->ev_in_free, s->ev_out_ready, s->ev_d2h_done}) if (e) cudaEventDestroy(e);
    for (auto& row : s->mark_events) for (auto& e : row) if (e) cudaEventDestroy(e);
    if (s->stream) cudaStreamDestroy(s->stream);
    cudaGetLastError();
    delete s;
}

#define SIM_OR_FAIL(h)                                                   \
    if (!(h)) return invalid("null FluidSim handle");                    \
    Sim* s = reinterpret_cast<Sim*>(h);                                  \
    FLUID_CUDA(cudaSetDevice(s->device));

#define NOT_BAND(s) if ((s)->band) return invalid("handle is a band of a FluidMulti: drive it through fluid_multi_*")

int fluid_sim_tick(FluidSim* h, const FluidTickSettings* t) {
    SIM_OR_FAIL(h);
    NOT_BAND(s);
    if (!t) return invalid("fluid_sim_tick: null settings");
    return s->enqueue_tick(*t, false);
}
int fluid_sim_tick_n(FluidSim* h, const FluidTickSettings* t, uint32_t count) {
    SIM_OR_FAIL(h);
    if (!t) return invalid("fluid_sim_tick_n: null settings");
    NOT_BAND(s);
    for (uint32_t k = 0; k < count; k++) {
        int rc = s->enqueue_tick(*t, false);
        if (rc != FLUID_OK) return rc;
    }
    return FLUID_OK;
}
int fluid_sim_tick_hash_sort_only(FluidSim* h, const FluidTickSettings* t) {
    SIM_OR_FAIL(h);
    if (!t) return invalid("fluid_sim_tick_hash_sort_only: null settings");
    NOT_BAND(s);
    return s->enqueue_tick(*t, true);
}
int fluid_sim_sync(FluidSim* h) {
    SIM_OR_FAIL(h);
    FLUID_CUDA(cudaStreamSynchronize(s->stream));
    return FLUID_OK;
}
uint32_t fluid_sim_tick_count(const FluidSim* h) { return h ? reinterpret_cast<const Sim*>(h)->tick : 0; }
int fluid_sim_set_tick_count(FluidSim* h, uint32_t tick) {
    SIM_OR_FAIL(h);
    s->tick = tick;
    FLUID_CUDA(cudaMemcpyAsync(s->d_frame_time, &s->tick, 4, cudaMemcpyHostToDevice, s->stream));
    FLUID_CUDA(cudaStreamSynchronize(s->stream));
    return FLUID_OK;
}

uint32_t fluid_sim_particle_count(const FluidSim* h) { return h ? reinterpret_cast<const Sim*>(h)->n : 0; }
uint32_t fluid_sim_grid_w(const FluidSim* h) { return h ? reinterpret_cast<const Sim*>(h)->grid_w : 0; }
uint32_t fluid_sim_grid_h(const FluidSim* h) { return h ? reinterpret_cast<const Sim*>(h)->grid_h : 0; }

static int ensure_stage(Sim* s) {
    if (!s->aos_stage) FLUID_CUDA(cudaMalloc(&s->aos_stage, ((size_t)s->capacity + 64) * sizeof(FluidParticle)));
    return FLUID_OK;
}

int fluid_sim_download_particles(FluidSim* h, FluidParticle* dst) {
    SIM_OR_FAIL(h);
    if (!dst) return invalid("fluid_sim_download_particles: null destination");
    int rc = ensure_stage(s);
    if (rc != FLUID_OK) return rc;
    if (s->n == 0) return FLUID_OK;
    k_soa_to_aos<<<(s->n + 255) / 256, 256, 0, s->stream>>>(s->aos_stage, s->n, s->pos_a + s->state_off, s->pred_s + s->sorted_off,
                                                          s->vel_a + s->state_off, s->rho + s->sorted_off, s->keys_view());
    s->launches++;
    FLUID_CUDA(cudaMemcpyAsync(dst, s->aos_stage, (size_t)s->n * sizeof(FluidParticle), cudaMemcpyDeviceToHost, s->stream));
    FLUID_CUDA(cudaStreamSynchronize(s->stream));
    return FLUID_OK;
}
int fluid_sim_upload_particles(FluidSim* h, const FluidParticle* src) {
    SIM_OR_FAIL(h);
    if (!src) return invalid("fluid_sim_upload_particles: null source");
    int rc = ensure_stage(s);
    if (rc != FLUID_OK) return rc;
    FLUID_CUDA(cudaMemcpyAsync(s->aos_stage, src, (size_t)s->n * sizeof(FluidParticle), cudaMemcpyHostToDevice, s->stream));
    s->cell_max_known = false;
    s->spec_valid = false;
    if (s->n == 0) return FLUID_OK;
    k_aos_to_soa<<<(s->n + 255) / 256, 256, 0, s->stream>>>(s->aos_stage, s->n, s->pos_a + s->state_off, s->pred_s + s->sorted_off,
                                                          s->vel_a + s->state_off, s->rho + s->sorted_off, s->keys_view());
    s->launches++;
    FLUID_CUDA(cudaGetLastError());
    // the source buffer may be pageable: the copy above is then already complete on return; for pinned
    // memory the caller must not modify `src` before the next fluid_sim_sync / download.
    return FLUID_OK;
}
static int ensure_pipeline(Sim* s) {
    if (!s->h2d_stream) {
        FLUID_CUDA(cudaStreamCreateWithFlags(&s->h2d_stream, cudaStreamNonBlocking));
        FLUID_CUDA(cudaStreamCreateWithFlags(&s->d2h_stream, cudaStreamNonBlocking));
        cudaEvent_t* evs[] = {&s->ev_h2d_done, &s->ev_in_free, &s->ev_out_ready, &s->ev_d2h_done};
        for (cudaEvent_t* e : evs) FLUID_CUDA(cudaEventCreateWithFlags(e, cudaEventDisableTiming));
    }
    if (!s->aos_in) {  // first use, or dropped by fluid_sim_reserve
        FLUID_CUDA(cudaMalloc(&s->aos_in, ((size_t)s->capacity + 64) * sizeof(FluidParticle)));
        FLUID_CUDA(cudaMalloc(&s->aos_out, ((size_t)s->capacity + 64) * sizeof(FluidParticle)));
        s->in_used = s->out_used = false;
    }
    return FLUID_OK;
}
int fluid_sim_upload_particles_async(FluidSim* h, const FluidParticle* src) {
    SIM_OR_FAIL(h);
    NOT_BAND(s);
    if (!src) return invalid("fluid_sim_upload_particles_async: null source");
    int rc = ensure_pipeline(s);
    if (rc != FLUID_OK) return rc;
    // the copy stream may overwrite the staging buffer once the previous transposition out of it has run
    if (s->in_used) FLUID_CUDA(cudaStreamWaitEvent(s->h2d_stream, s->ev_in_free, 0));
    FLUID_CUDA(cudaMemcpyAsync(s->aos_in, src, (size_t)s->n * sizeof(FluidParticle), cudaMemcpyHostToDevice, s->h2d_stream));
    FLUID_CUDA(cudaEventRecord(s->ev_h2d_done, s->h2d_stream));
    FLUID_CUDA(cudaStreamWaitEvent(s->stream, s->ev_h2d_done, 0));
    k_aos_to_soa<<<(s->n + 255) / 256, 256, 0, s->stream>>>(s->aos_in, s->n, s->pos_a, s->pred_s, s->vel_a, s->rho, s->keys_view());
    s->launches++;
    FLUID_CUDA(cudaEventRecord(s->ev_in_free, s->stream));
    s->in_used = true;
    s->cell_max_known = false;
    s->spec_valid = false;
    FLUID_CUDA(cudaGetLastError());
    return FLUID_OK;
}
int fluid_sim_download_particles_async(FluidSim* h, FluidParticle* dst) {
    SIM_OR_FAIL(h);
    NOT_BAND(s);
    if (!dst) return invalid("fluid_sim_download_particles_async: null destination");
    int rc = ensure_pipeline(s);
    if (rc != FLUID_OK) return rc;
    // the transposition may overwrite the staging buffer once the previous device-to-host copy out of it is done
    if (s->out_used) FLUID_CUDA(cudaStreamWaitEvent(s->stream, s->ev_d2h_done, 0));
    k_soa_to_aos<<<(s->n + 255) / 256, 256, 0, s->stream>>>(s->aos_out, s->n, s->pos_a, s->pred_s, s->vel_a, s->rho, s->keys_view());
    s->launches++;
    FLUID_CUDA(cudaEventRecord(s->ev_out_ready, s->stream));
    FLUID_CUDA(cudaStreamWaitEvent(s->d2h_stream, s->ev_out_ready, 0));
    FLUID_CUDA(cudaMemcpyAsync(dst, s->aos_out, (size_t)s->n * sizeof(FluidParticle), cudaMemcpyDeviceToHost, s->d2h_stream));
    FLUID_CUDA(cudaEventRecord(s->ev_d2h_done, s->d2h_stream));
    s->out_used = true;
    FLUID_CUDA(cudaGetLastError());
    return FLUID_OK;
}
int fluid_sim_sync_transfers(FluidSim* h) {
    SIM_OR_FAIL(h);
    if (s->h2d_stream) FLUID_CUDA(cudaStreamSynchronize(s->h2d_stream));
    FLUID_CUDA(cudaStreamSynchronize(s->stream));
    if (s->d2h_stream) FLUID_CUDA(cudaStreamSynchronize(s->d2h_stream));
    return FLUID_OK;
}
int fluid_sim_download_start_indices(FluidSim* h, uint32_t* dst) {
    SIM_OR_FAIL(h);
    if (!dst) return invalid("null destination");
    FLUID_CUDA(cudaMemcpyAsync(dst, s->start_indices, (size_t)s->num_cells * 4, cudaMemcpyDeviceToHost, s->stream));
    FLUID_CUDA(cudaStreamSynchronize(s->stream));
    return FLUID_OK;
}
int fluid_sim_upload_start_indices(FluidSim* h, const uint32_t* src) {
    SIM_OR_FAIL(h);
    if (!src) return invalid("null source");
    FLUID_CUDA(cudaMemcpyAsync(s->start_indices, src, (size_t)s->num_cells * 4, cudaMemcpyHostToDevice, s->stream));
    FLUID_CUDA(cudaStreamSynchronize(s->stream));
    return FLUID_OK;
}
int fluid_sim_upload_force_field(FluidSim* h, const float* xy) {
    SIM_OR_FAIL(h);
    if (!xy) return invalid("null source");
    const size_t len = (size_t)s->settings.texture_size[0] * s->settings.texture_size[1];
    FLUID_CUDA(cudaMemcpyAsync(s->force_field, xy, len * 8, cudaMemcpyHostToDevice, s->stream));
    FLUID_CUDA(cudaStreamSynchronize(s->stream));
    return FLUID_OK;
}
int fluid_sim_download_force_field(FluidSim* h, float* xy) {
    SIM_OR_FAIL(h);
    if (!xy) return invalid("null destination");
    const size_t len = (size_t)s->settings.texture_size[0] * s->settings.texture_size[1];
    FLUID_CUDA(cudaMemcpyAsync(xy, s->force_field, len * 8, cudaMemcpyDeviceToHost, s->stream));
    FLUID_CUDA(cudaStreamSynchronize(s->stream));
    return FLUID_OK;
}
int fluid_sim_get_uniform(const FluidSim* h, FluidSimUniform* out) {
    if (!h || !out) return invalid("null argument");
    *out = reinterpret_cast<const Sim*>(h)->last_uniform;
    return FLUID_OK;
}
int fluid_sim_download_unsorted_keys(FluidSim* h, uint32_t* dst) {
    SIM_OR_FAIL(h);
    if (!dst) return invalid("null destination");
    FLUID_CUDA(cudaMemcpyAsync(dst, s->keys_u, (size_t)s->n * 4, cudaMemcpyDeviceToHost, s->stream));
    FLUID_CUDA(cudaStreamSynchronize(s->stream));
    return FLUID_OK;
}
int fluid_sim_download_permutation(FluidSim* h, uint32_t* dst) {
    SIM_OR_FAIL(h);
    if (!dst) return invalid("null destination");
    if (!s->idx_sorted) return invalid("fluid_sim_download_permutation: no tick has run yet");
    if (!s->band && s->last_sort_path == FLUID_SORT_COUNTING) {
        k_rank_permutation<<<(s->n + GATHER_THREADS - 1) / GATHER_THREADS, GATHER_THREADS, 0, s->stream>>>(
            s->n, s->num_cells, s->keys_b, s->idx_a, s->cell_lb, s->idx_b);
        s->launches++;
    }
    FLUID_CUDA(cudaMemcpyAsync(dst, s->idx_sorted, (size_t)s->n * 4, cudaMemcpyDeviceToHost, s->stream));
    FLUID_CUDA(cudaStreamSynchronize(s->stream));
    return FLUID_OK;
}
int fluid_sim_enable_debug_force(FluidSim* h, int enable) {
    SIM_OR_FAIL(h);
    if (enable && !s->dbg_accel) {
        FLUID_CUDA(cudaMalloc(&s->dbg_accel, ((size_t)s->capacity + 64) * 8));
        FLUID_CUDA(cudaMalloc(&s->dbg_abs, ((size_t)s->capacity + 64) * 8));
    }
    s->debug_force = enable != 0;
    return FLUID_OK;
}
int fluid_sim_download_debug_force(FluidSim* h, float* accel_xy, float* abs_xy) {
    SIM_OR_FAIL(h);
    if (!s->dbg_accel) return invalid("debug force was never enabled");
    if (accel_xy) FLUID_CUDA(cudaMemcpyAsync(accel_xy, s->dbg_accel, (size_t)s->n * 8, cudaMemcpyDeviceToHost, s->stream));
    if (abs_xy) FLUID_CUDA(cudaMemcpyAsync(abs_xy, s->dbg_abs, (size_t)s->n * 8, cudaMemcpyDeviceToHost, s->stream));
    FLUID_CUDA(cudaStreamSynchronize(s->stream));
    return FLUID_OK;
}
int fluid_sim_device_view(FluidSim* h, FluidSimDeviceView* out) {
    SIM_OR_FAIL(h);
    if (!out) return invalid("null argument");
    out->particle_count = s->n; out->capacity = s->capacity;
    out->position = s->pos_a + s->state_off; out->predicted_position = s->pred_s + s->sorted_off;
    out->velocity = s->vel_a + s->state_off; out->density = s->rho + s->sorted_off; out->grid = s->keys_view();
    out->start_indices = s->start_indices; out->cell_lower_bound = s->cell_lb; out->force_field = s->force_field;
    return FLUID_OK;
}
int fluid_sim_state_changed(FluidSim* h) {
    SIM_OR_FAIL(h);
    s->spec_valid = false;
    s->cell_max_known = false;
    return FLUID_OK;
}
void* fluid_sim_stream(FluidSim* h) { return h ? (void*)reinterpret_cast<Sim*>(h)->stream : nullptr; }

int fluid_sim_enable_timing(FluidSim* h, int enable) {
    SIM_OR_FAIL(h);
    s->timing = enable != 0;
    s->n_marks = 0;
    s->marks_per_tick = 0;
    s->timed_ticks = 0;
    return FLUID_OK;
}
int fluid_sim_last_timing(FluidSim* h, const char** names, float* ms, int cap) {
    SIM_OR_FAIL(h);
    FLUID_CUDA(cudaStreamSynchronize(s->stream));
    if (s->timed_ticks == 0) return 0;
    const uint64_t ticks = s->timed_ticks < (uint64_t)TIMING_RING ? s->timed_ticks : (uint64_t)TIMING_RING;
    int out = 0;
    for (int k = 1; k < s->marks_per_tick && out < cap; k++, out++) {
        double sum = 0.0;
        for (uint64_t t = 0; t < ticks; t++) {
            const uint64_t slot = (s->timed_ticks - 1 - t) % TIMING_RING;
            float dt = 0.f;
            FLUID_CUDA(cudaEventElapsedTime(&dt, s->mark_events[slot][k - 1], s->mark_events[slot][k]));
            sum += dt;
        }
        if (names) names[out] = s->mark_names[k];
        if (ms) ms[out] = (float)(sum / (double)ticks);
    }
    // device time between the last mark of a tick and the first mark of the next one: zero while the host keeps the
    // stream fed, the host's lag otherwise
    if (ticks >= 2 && out < cap && s->marks_per_tick >= 2) {
        double sum = 0.0;
        for (uint64_t t = 0; t + 1 < ticks; t++) {
            const uint64_t later = (s->timed_ticks - 1 - t) % TIMING_RING, earlier = (s->timed_ticks - 2 - t) % TIMING_RING;
            float dt = 0.f;
            FLUID_CUDA(cudaEventElapsedTime(&dt, s->mark_events[earlier][s->marks_per_tick - 1], s->mark_events[later][0]));
            sum += dt;
        }
        if (names) names[out] = "between ticks";
        if (ms) ms[out] = (float)(sum / (double)(ticks - 1));
        out++;
    }
    return out;
}
uint64_t fluid_sim_launch_count(const FluidSim* h) { return h ? reinterpret_cast<const Sim*>(h)->launches : 0; }

int fluid_sim_set_math_mode(FluidSim* h, int mode) {
    SIM_OR_FAIL(h);
    if (mode != FLUID_MATH_EXACT && mode != FLUID_MATH_FAST) return invalid("fluid_sim_set_math_mode: unknown mode");
    s->exact_math = (mode == FLUID_MATH_EXACT);
    return FLUID_OK;
}
int fluid_sim_get_math_mode(const FluidSim* h) {
    return (h && !reinterpret_cast<const Sim*>(h)->exact_math) ? FLUID_MATH_FAST : FLUID_MATH_EXACT;
}

int fluid_sim_set_sort_mode(FluidSim* h, int mode) {
    SIM_OR_FAIL(h);
    if (mode != FLUID_SORT_AUTO && mode != FLUID_SORT_RADIX && mode != FLUID_SORT_COUNTING)
        return invalid("fluid_sim_set_sort_mode: unknown mode");
    s->sort_mode = mode;
    s->cell_max_known = false;
    s->spec_valid = false;
    s->count_dirty = false;
    // both count arrays clean again, whatever ran before (ticks in FLUID_SORT_RADIX mode do not touch them)
    const size_t bytes = (size_t)s->cell_tiles * CELL_SCAN_TILE * 4;
    FLUID_CUDA(cudaMemsetAsync(s->cell_count_buf[0], 0, bytes, s->stream));
    FLUID_CUDA(cudaMemsetAsync(s->cell_count_buf[1], 0, bytes, s->stream));
    return FLUID_OK;
}
int fluid_sim_get_sort_mode(const FluidSim* h) { return h ? reinterpret_cast<const Sim*>(h)->sort_mode : FLUID_SORT_AUTO; }
int fluid_sim_last_sort_path(const FluidSim* h) { return h ? reinterpret_cast<const Sim*>(h)->last_sort_path : FLUID_SORT_RADIX; }

int fluid_sim_reserve(FluidSim* h, uint32_t capacity) {
    SIM_OR_FAIL(h);
    NOT_BAND(s);
    if (capacity <= s->capacity) return FLUID_OK;
    FLUID_CUDA(cudaStreamSynchronize(s->stream));
    if (s->h2d_stream) { FLUID_CUDA(cudaStreamSynchronize(s->h2d_stream)); FLUID_CUDA(cudaStreamSynchronize(s->d2h_stream)); }
    // The new arrays are allocated FIRST and the old ones freed only after everything succeeded: a failed cudaMalloc
    // (out of memory at a large capacity) leaves the handle exactly as it was, live state included.
    void** members[] = {(void**)&s->pos_a, (void**)&s->vel_a, (void**)&s->pos_s, (void**)&s->vel_s, (void**)&s->pred_s, (void**)&s->rho,
                        (void**)&s->keys_u, (void**)&s->keys_b, (void**)&s->keys_c, (void**)&s->idx_a, (void**)&s->idx_b,
                        (void**)&s->sort_scratch, (void**)&s->aos_stage, (void**)&s->dbg_accel, (void**)&s->dbg_abs, (void**)&s->tile_plans,
                        (void**)&s->nb_head, (void**)&s->nb_mask, (void**)&s->aos_in, (void**)&s->aos_out};
    constexpr int NM = sizeof(members) / sizeof(members[0]);
    void* old[NM];
    for (int k = 0; k < NM; k++) { old[k] = *members[k]; *members[k] = nullptr; }
    const uint32_t old_capacity = s->capacity, old_max_tiles = s->max_tiles, n = s->n;
    const size_t old_scratch_words = s->sort_scratch_words;
    uint32_t* const old_keys_sorted = s->keys_sorted;
    uint32_t* const old_idx_sorted = s->idx_sorted;
    const bool dbg = s->debug_force;
    int rc = s->alloc(capacity);
    if (rc == FLUID_OK && dbg) {
        s->debug_force = false;
        rc = fluid_sim_enable_debug_force(h, 1);
    }
    if (rc == FLUID_OK && old_keys_sorted) {
        cudaError_t e = cudaSuccess;
        const float2 *o_pos = (const float2*)old[0], *o_vel = (const float2*)old[1], *o_pred = (const float2*)old[4];
        if (e == cudaSuccess) e = cudaMemcpyAsync(s->pos_a, o_pos, (size_t)n * 8, cudaMemcpyDeviceToDevice, s->stream);
        if (e == cudaSuccess) e = cudaMemcpyAsync(s->vel_a, o_vel, (size_t)n * 8, cudaMemcpyDeviceToDevice, s->stream);
        if (e == cudaSuccess) e = cudaMemcpyAsync(s->pred_s, o_pred, (size_t)n * 8, cudaMemcpyDeviceToDevice, s->stream);
        if (e == cudaSuccess) e = cudaMemcpyAsync(s->rho, old[5], (size_t)n * 4, cudaMemcpyDeviceToDevice, s->stream);
        if (e == cudaSuccess) e = cudaMemcpyAsync(s->keys_sorted, old_keys_sorted, (size_t)n * 4, cudaMemcpyDeviceToDevice, s->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(s->stream);
        if (e != cudaSuccess) rc = cuda_fail(e, "fluid_sim_reserve: copy into the new arrays", __FILE__, __LINE__);
    }
    if (rc != FLUID_OK) {
        // roll back: drop whatever the new allocation produced, put the old arrays back
        for (int k = 0; k < NM; k++) { if (*members[k]) cudaFree(*members[k]); *members[k] = old[k]; }
        s->capacity = old_capacity; s->max_tiles = old_max_tiles; s->sort_scratch_words = old_scratch_words;
        s->keys_sorted = old_keys_sorted; s->idx_sorted = old_idx_sorted; s->debug_force = dbg;
        cudaGetLastError();
        return rc;
    }
    for (int k = 0; k < NM; k++) if (old[k]) cudaFree(old[k]);
    s->in_used = s->out_used = false;  // (the pipelined-transfer staging buffers are re-created on next use)
    s->idx_sorted = nullptr;
    s->cell_max_known = false;
    s->spec_valid = false;
    return FLUID_OK;
}
int fluid_sim_set_particle_count(FluidSim* h, uint32_t n) {
    SIM_OR_FAIL(h);
    NOT_BAND(s);
    if (n < 2 || n > s->capacity) return invalid("fluid_sim_set_particle_count: need 2 <= n <= capacity");
    s->n = n;
    s->cell_max_known = false;
    s->spec_valid = false;
    return FLUID_OK;
}

}  // extern "C"

#include "multi.inl"
