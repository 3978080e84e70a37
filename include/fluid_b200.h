/*
 * fluid_b200.h — C ABI of libfluid_b200.so, the B200 (sm_100a) backend for the
 * per-frame fluid step of rookieCookies/gpu-fluid-simulation.
 *
 * Every entry point replaces one piece of the reference's Rust `FluidSimulation`
 * API (reference `src/simulation.rs`) or one host-side step the reference performs
 * around it; the reference interface each one stands in for is cited beside it.
 * Plain pointers and sizes only; no CUDA, torch or C++ types cross this boundary.
 *
 * Conventions
 *   - every function returns FLUID_OK (0) or a negative FluidStatus;
 *     fluid_last_error() gives the thread-local message of the last failure.
 *   - a FluidSim handle owns all of its device memory and one CUDA stream; it is
 *     NOT thread-safe (one caller thread at a time, like the reference, which only
 *     touches FluidSimulation from the winit thread).
 *   - the caller owns every host pointer it passes; none is retained past the call.
 *   - there is no CPU fallback: without a usable CUDA device create() fails.
 */
#ifndef FLUID_B200_H
#define FLUID_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FLUID_B200_ABI_VERSION 1

typedef enum FluidStatus {
    FLUID_OK = 0,
    FLUID_ERR_INVALID_ARGUMENT = -1, /* null pointer, particle_count < 2, bad sizes */
    FLUID_ERR_CUDA = -2,             /* a CUDA runtime call failed; see fluid_last_error */
    FLUID_ERR_NO_DEVICE = -3,        /* no CUDA device / ordinal out of range */
    FLUID_ERR_UNSUPPORTED = -4       /* e.g. image larger than the JFA seed encoding allows */
} FluidStatus;

/* reference: `ParticleInstance`, src/simulation.rs:126-135 (WGSL twin funcs.wgsl:1-8).
 * 32 bytes, align 4. The particle buffer is physically in key order after a tick. */
typedef struct FluidParticle {
    float position[2];
    float predicted_position[2];
    float velocity[2];
    float density;
    uint32_t grid;
} FluidParticle;

/* reference: `SimulationUniform`, src/simulation.rs:53-90 (WGSL twin funcs.wgsl:17-51). 120 bytes. */
typedef struct FluidSimUniform {
    float delta;
    uint32_t particle_count;
    float sqr_radius;
    uint32_t frame_time;
    float gravity[2];
    float bounds[2];
    float mouse_pos[2];
    float smoothing_radius;
    float particle_mass;
    float pressure_constant;
    float rest_density;
    float damping_factor;
    float viscosity_coefficient;
    float surface_tension_treshold;
    float surface_tension_coefficient;
    float poly6_kernel_volume;
    float poly6_kernel_derivative;
    float poly6_kernel_laplacian;
    float spiky_kernel_derivative;
    float viscosity_kernel;
    int32_t mouse_state;
    float mouse_force_radius;
    float mouse_force_power;
    uint32_t grid_w;
    uint32_t grid_h;
    float texture_size[2];
} FluidSimUniform;

/* reference: `SimulationSettings`, src/simulation.rs:95-104 (same field order). */
typedef struct FluidSimSettings {
    uint32_t particle_count;
    float particle_spacing;
    float smoothing_radius;
    float size[2];
    uint32_t texture_size[2];
} FluidSimSettings;

/* reference: `TickSettings`, src/simulation.rs:107-122 (same field order). */
typedef struct FluidTickSettings {
    float delta;
    float gravity[2];
    float mass;
    float pressure_constant;
    float rest_density;
    float damping_factor;
    float viscosity_coefficient;
    float surface_tension_treshold;
    float surface_tension_coefficient;
    float mouse_force_radius;
    float mouse_force_power;
    float mouse_pos[2];
    int32_t mouse_state;
} FluidTickSettings;

typedef struct FluidSim FluidSim; /* opaque */

const char* fluid_last_error(void);
int fluid_abi_version(void);
/* number of CUDA devices visible; negative FluidStatus on failure. */
int fluid_device_count(void);

/* ---- construction / destruction ------------------------------------------------ */

/* reference: `FluidSimulation::new(&Device, SimulationSettings)`, src/simulation.rs:139.
 * Builds the init lattice of src/simulation.rs:147-163 bit for bit, zero-initialises
 * start_indices (src/simulation.rs:204-209) and the force field (:213-218).
 * particle_count < 2 is rejected (the reference panics at src/simulation.rs:323-324). */
int fluid_sim_create(const FluidSimSettings* settings, int device_ordinal, FluidSim** out);
/* reference: Drop of FluidSimulation (rebuilt on restart, src/renderer.rs:873-875). */
void fluid_sim_destroy(FluidSim* sim);

/* ---- stepping -------------------------------------------------------------------- */

/* reference: `FluidSimulation::tick`, src/simulation.rs:459-539. Increments the tick
 * counter first (:460), derives the uniform (:470-497) and enqueues the whole step
 * (predict, key, sort, cell-start, density, force+integrate) on the handle's stream.
 * Asynchronous, mirroring the reference's record-then-submit (src/main.rs:226). */
int fluid_sim_tick(FluidSim* sim, const FluidTickSettings* tick);
/* `n` consecutive ticks with the same settings (reference: the Running/Render loops,
 * src/main.rs:137-147,199-201), enqueued back to back from one call (eight kernels per tick, chained by programmatic
 * dependent launch; no host synchronisation in between except the first tick after an upload). */
int fluid_sim_tick_n(FluidSim* sim, const FluidTickSettings* tick, uint32_t n);
/* reference: `device.poll(Wait)`, src/main.rs:79. */
int fluid_sim_sync(FluidSim* sim);
/* reference: `pub tick: u32`, src/simulation.rs:12. */
uint32_t fluid_sim_tick_count(const FluidSim* sim);
int fluid_sim_set_tick_count(FluidSim* sim, uint32_t tick);

/* Arithmetic of the force pass (the reference's WGSL leaves division / sqrt accuracy to the driver: 2.5 ulp).
 * FLUID_MATH_EXACT (default): IEEE-rounded operations in the shader's order, bit-identical to the CPU restatement.
 * FLUID_MATH_FAST: hardware reciprocal / rsqrt and FMA contraction in the pressure+viscosity pair terms; keys,
 * sorted order, cell tables, predicted positions and densities are unchanged, velocity/position agree with
 * EXACT within 1e-5 (relative to sum|terms|). */
#define FLUID_MATH_EXACT 0
#define FLUID_MATH_FAST 1
int fluid_sim_set_math_mode(FluidSim* sim, int mode);
int fluid_sim_get_math_mode(const FluidSim* sim);

/* How the tick produces the key order (replaces sort.wgsl:27-51 + the schedule of src/simulation.rs:323-347 either way;
 * both give the identical, stable order, so every downstream value is the same bit for bit).
 * FLUID_SORT_COUNTING: counting sort over the cells (a key is a cell id, funcs.wgsl:216-218): per-cell counts, an
 *   exclusive scan that IS the cell table, slot assignment, then each cell's few entries ranked by source index.
 *   Cost grows with the square of a cell's population.
 * FLUID_SORT_RADIX: onesweep LSD radix passes over (key, index); cost independent of the distribution.
 * FLUID_SORT_AUTO (default): counting while the most crowded cell holds <= clamp(n / 1024, 64, 4096) particles and the
 *   squared cell populations sum to <= 128 n (both watched per tick without a sync; one stream sync on the first tick
 *   after an upload), radix otherwise. Slab mode always uses the radix passes.
 * fluid_sim_last_sort_path: which of the two the most recently enqueued tick used. */
#define FLUID_SORT_AUTO 0
#define FLUID_SORT_RADIX 1
#define FLUID_SORT_COUNTING 2
int fluid_sim_set_sort_mode(FluidSim* sim, int mode);
int fluid_sim_get_sort_mode(const FluidSim* sim);
int fluid_sim_last_sort_path(const FluidSim* sim);

/* BASELINE.json config 2: spatial hash + key sort + cell-start table only
 * (reference passes K1..K4: compute.wgsl:8-56, sort.wgsl:27-51). Increments tick. */
int fluid_sim_tick_hash_sort_only(FluidSim* sim, const FluidTickSettings* tick);

/* ---- state exposure (headless stand-ins for the wgpu bind-group getters,
 *      src/simulation.rs:542-564) ------------------------------------------------- */

uint32_t fluid_sim_particle_count(const FluidSim* sim);
uint32_t fluid_sim_grid_w(const FluidSim* sim);
uint32_t fluid_sim_grid_h(const FluidSim* sim);

/* binding 0 of simulation_bg: N x 32 B AoS, in the current (sorted) physical order. */
int fluid_sim_download_particles(FluidSim* sim, FluidParticle* dst);
int fluid_sim_upload_particles(FluidSim* sim, const FluidParticle* src);
/* Pipelined variants for callers that stream whole states through the device (the reference never does: its state
 * lives in wgpu buffers, src/simulation.rs:168-172; this is the boundary's host-buffer path made full duplex).
 * Copies run on the handle's own copy streams through separate staging buffers, ordered against the tick stream by
 * events: the upload of the next state overlaps the download of the previous result. `src` / `dst` should be pinned
 * (pageable memory degrades to synchronous copies) and must stay untouched until fluid_sim_sync_transfers returns or,
 * for `src`, until the next _upload_particles_async on the same handle has been issued AND a later sync returned.
 * Not available in slab mode. */
int fluid_sim_upload_particles_async(FluidSim* sim, const FluidParticle* src);
int fluid_sim_download_particles_async(FluidSim* sim, FluidParticle* dst);
int fluid_sim_sync_transfers(FluidSim* sim);
/* binding 1: grid_w*grid_h u32, reference semantics (never cleared, compute.wgsl:45-56). */
int fluid_sim_download_start_indices(FluidSim* sim, uint32_t* dst);
int fluid_sim_upload_start_indices(FluidSim* sim, const uint32_t* src);
/* binding 2: reference `queue.write_buffer(force_field_texture(), ..)`, src/renderer.rs:498-502.
 * tex_w*tex_h (x,y) float pairs. */
int fluid_sim_upload_force_field(FluidSim* sim, const float* xy);
int fluid_sim_download_force_field(FluidSim* sim, float* xy);
/* the uniform the last tick used (reference: simulation_settings_bg). */
int fluid_sim_get_uniform(const FluidSim* sim, FluidSimUniform* out);

/* debug / parity taps (not in the reference API; transients of its shaders). */
/* keys in particle order as written by create_spatial_lookup, before the sort. */
int fluid_sim_download_unsorted_keys(FluidSim* sim, uint32_t* dst);
/* pressure+viscosity acceleration per particle (compute.wgsl:90-93) and the sum of
 * |terms| that the cancellation-aware tolerance uses. Enable before the tick. */
int fluid_sim_enable_debug_force(FluidSim* sim, int enable);
int fluid_sim_download_debug_force(FluidSim* sim, float* accel_xy, float* abs_terms_xy);
/* sorted-order permutation of the last tick: sorted index -> index before the sort. */
int fluid_sim_download_permutation(FluidSim* sim, uint32_t* dst);

/* zero-copy accessors for device-side consumers (renderer kernels, NCCL halo exchange).
 * Pointers stay valid until destroy; contents are ordered on fluid_sim_stream(). */
typedef struct FluidSimDeviceView {
    uint32_t particle_count;
    uint32_t capacity;
    void* position;           /* float2[N]  current positions, sorted order */
    void* predicted_position; /* float2[N] */
    void* velocity;           /* float2[N] */
    void* density;            /* float[N] */
    void* grid;               /* uint32[N] sorted keys */
    void* start_indices;      /* uint32[grid_w*grid_h] reference-compatible table */
    void* cell_lower_bound;   /* uint32[grid_w*grid_h+1] dense clean table */
    void* force_field;        /* float2[tex_w*tex_h] */
} FluidSimDeviceView;
int fluid_sim_device_view(FluidSim* sim, FluidSimDeviceView* out);
/* A consumer that WRITES position or velocity through the pointers above (instead of fluid_sim_upload_particles) must
 * say so before the next tick: the library carries work from one tick into the next (the force pass emits the next
 * tick's keys and cell counts, the sort path is chosen from the previous ticks' cell statistics). */
int fluid_sim_state_changed(FluidSim* sim);
void* fluid_sim_stream(FluidSim* sim); /* cudaStream_t */

/* per-phase device timing (CUDA events on the handle's stream), averaged over the ticks
 * enqueued since timing was enabled (the most recent 256 at most). names/ms arrays hold up to
 * `cap` entries; returns the number of phases written. Synchronises the stream. */
int fluid_sim_enable_timing(FluidSim* sim, int enable);
int fluid_sim_last_timing(FluidSim* sim, const char** names, float* ms, int cap);
/* number of kernels this library has launched on the handle since creation. */
uint64_t fluid_sim_launch_count(const FluidSim* sim);

/* ---- capacity / particle count ----------------------------------------------------- */
int fluid_sim_reserve(FluidSim* sim, uint32_t capacity);
int fluid_sim_set_particle_count(FluidSim* sim, uint32_t n);

/* ---- slab (multi-GPU) mode: SURVEY.md §8(e) -------------------------------------------
 * One handle per rank owns a band of grid rows [row_bounds[rank], row_bounds[rank+1]). Keys stay global. The host
 * drives the phases below and performs the three exchanges between them (migration, halo pred+vel, halo rho) with
 * NCCL on views of the handle's device arrays (FluidSlabView); see gpu_fluid_simulation_b200/slab.py. The reference
 * has no multi-GPU code; the contract is that the gathered result equals a single-GPU tick of the same state. */
typedef struct FluidSlabView {
    uint32_t n;            /* particles currently owned */
    uint32_t pad;          /* state arrays: natives start at index pad; key-ordered arrays: owned start at pad */
    uint32_t capacity;
    uint32_t emig_cap;
    void* pos_state;       /* float2[pad + capacity + pad] */
    void* vel_state;
    void* pos_sorted;      /* float2[...] key order, [halo below | owned | halo above] around index pad */
    void* pred_sorted;
    void* vel_sorted;
    void* rho;             /* float[...] */
    void* keys_ext;        /* uint32[...] */
    void* emig_list;       /* uint64[emig_cap]: (destination rank << 32) | index into the state arrays; 0xff.. = unused */
    void* emig_count;      /* uint32; emig_count[8 + r] = emigrants bound for rank r */
    void* meta;            /* uint32[5]: n owned, first-row count, last-row count, key of first owned, its start_indices entry */
    void* start_indices;   /* uint32[grid_w*grid_h], global sorted indices for the cells this rank owns */
    void* dyn;             /* uint32[8]: hb, ha, ext_begin, ext_end, global_offset, quirk c0, quirk stale, error bits */
} FluidSlabView;
int fluid_slab_configure(FluidSim* sim, int rank, int world, const uint32_t* row_bounds /* world+1 */, uint32_t capacity,
                         uint32_t pad, uint32_t emig_cap);
int fluid_slab_view(FluidSim* sim, FluidSlabView* out);
/* phase 1: tick += 1; keys of the natives; emigrants listed and given the sentinel key */
int fluid_slab_begin_tick(FluidSim* sim, const FluidTickSettings* tick);
/* phase 2a: sort the emigrant list (destination rank, then source order) and gather (pos, vel) of the n_emigrants
 * into payload_dev (DEVICE pointer, float[n_emigrants*4]); 2b: place the received immigrants (DEVICE pointer,
 * ordered by source rank: n_from_below then n_from_above records) around the natives */
int fluid_slab_pack(FluidSim* sim, uint32_t n_emigrants, float* payload_dev);
int fluid_slab_unpack(FluidSim* sim, const float* recv_dev, uint32_t n_from_below, uint32_t n_from_above);
/* phase 3+4: keys of the immigrants (already copied in around the natives), sort, gather, band metadata */
int fluid_slab_sort(FluidSim* sim, uint32_t n_from_below, uint32_t n_from_above, uint32_t n_emigrants);
/* phase 6+7a: halo keys, cell tables, density. quirk_*: lowest occupied cell of the whole domain and its stale
 * start_indices entry translated to this rank's key-ordered index space (0xffffffff, 0 when not visible) */
int fluid_slab_cells_density(FluidSim* sim, uint32_t halo_below, uint32_t halo_above, uint32_t global_offset,
                             uint32_t quirk_c0, uint32_t quirk_stale_ext);
/* phases 6+7a without a host read: allmeta_dev = DEVICE pointer to world x 8 uint32 (every rank's `meta` words, as
 * gathered by the exchange layer). Halo rows must have been received as pad-sized windows (below: right-aligned at
 * the owned range; above: starting right after it); halo counts, global offset and the quirk are resolved on the
 * device. FluidSlabView.dyn[7] latches error bits (1: a halo row exceeds the pad). */
int fluid_slab_cells_density_dev(FluidSim* sim, const uint32_t* allmeta_dev);
/* phase 7b: force + integrate of the owned particles */
int fluid_slab_force(FluidSim* sim);

/* ---- jump-flood distance field (reference: shaders/jump_flood.wgsl:17-65; the host
 *      protocol is pinned in SURVEY.md §8 row a7) ---------------------------------- */

#define FLUID_JFA_INVALID 32767u

/* Full JFA: seeds_in/out are rg32uint images (w*h pairs of u32), host pointers.
 * jumps 2^(ceil(log2 max(w,h))-1) .. 1, strict ping-pong. */
int fluid_jfa_run(const uint32_t* seeds_in, uint32_t w, uint32_t h, uint32_t* seeds_out, int device_ordinal);
/* one pass with an explicit jump (reference entry point `floodfill`). */
int fluid_jfa_pass(const uint32_t* seeds_in, uint32_t w, uint32_t h, uint32_t jump, uint32_t* seeds_out, int device_ordinal);
/* seed image from a u8 mask with the seed rule of src/main.rs:413-438 (> 128, border fallback). */
int fluid_jfa_seed_from_mask(const uint8_t* mask, uint32_t w, uint32_t h, uint32_t* seeds_out, int device_ordinal);
/* seed image from predicted positions with the texel mapping of compute.wgsl:127-128. */
int fluid_jfa_seed_from_particles(const FluidParticle* particles, uint32_t n, const float bounds[2],
                                  uint32_t w, uint32_t h, uint32_t* seeds_out, int device_ordinal);
/* field = -(pixel - seed) as float pairs (sign / zero rule of src/main.rs:493-511). */
int fluid_jfa_to_force_field(const uint32_t* seeds, uint32_t w, uint32_t h, float* field_xy, int device_ordinal);

/* Device-resident pipeline: particles of `sim` -> seeds -> JFA -> field, written into
 * caller-provided host buffers if non-null; when into_sim != 0 and (w,h) equals the
 * sim's texture size the field is also installed as the sim's force field.
 * ms_out (optional) receives the device time of the JFA passes only. */
int fluid_sim_jfa_from_particles(FluidSim* sim, uint32_t w, uint32_t h, uint32_t* seeds_out,
                                 float* field_xy_out, int into_sim, float* ms_out);

/* ---- obstacle objects -> mask -> seeds -> JFA -> force field ----------------
 * FluidObject = the reference's storage-buffer record (src/renderer.rs:82-90, image_shader.wgsl:6-11):
 * kind 0 = circle, radius = bitcast<f32>(pad); kind 1 = rectangle, rotation = bitcast<f32>(pad),
 * extents = bitcast<vec2<f32>>(pad2.xy). inv_proj = the inverse projection of ObjectUniform
 * (src/renderer.rs:72-79, :449-453), 16 floats, column-major (glam Mat4). Mask: u8, 0 inside an object,
 * 255 elsewhere (image_shader.wgsl:41-67), row 0 = top of the framebuffer. */
typedef struct FluidObject {
    float position[2];
    uint32_t kind;
    uint32_t pad;
    uint32_t pad2[4];
} FluidObject;
/* mask only, HOST buffers: mirrors the reference's object render pass + texture readback. */
int fluid_raster_objects(const FluidObject* objects, uint32_t n, const float inv_proj[16], uint32_t w, uint32_t h,
                         uint8_t* mask_out, int device_ordinal);
/* Device-resident pipeline at the sim's texture size: rasterise -> seed rule of src/main.rs:413-438 -> JFA ->
 * field installed as the sim's force field (what src/renderer.rs:497-547 does through a readback, a CPU thread
 * running generate_smooth_gradient_field and an 8 MiB upload, one or more frames late). Optional host outputs
 * (any may be NULL): mask (w*h u8), seeds (w*h rg32uint), field (w*h float pairs), device ms of the chain.
 * With every output NULL the call only enqueues work on the sim's stream. */
int fluid_sim_field_from_objects(FluidSim* sim, const FluidObject* objects, uint32_t n, const float inv_proj[16],
                                 uint8_t* mask_out, uint32_t* seeds_out, float* field_xy_out, float* ms_out);
/* the same from a caller-supplied u8 mask at texture size (the reference's `Image`, src/main.rs:379-399). */
int fluid_sim_field_from_mask(FluidSim* sim, const uint8_t* mask, uint32_t* seeds_out, float* field_xy_out, float* ms_out);

/* ---- headless fluid render --------------------------------------------------
 * fluid_shader.wgsl fs_main (the live root copy, src/renderer.rs:162) evaluated per pixel of a w x h target
 * (reference: RENDER_DIMS 960 x 540, src/renderer.rs:15) over the sim's sorted particles and start_indices
 * table; inv_proj as for the objects. HOST outputs, either may be NULL: rgba_f32_out = the fragment's
 * vec4<f32> (w*h*4 floats); rgba8_out = that fragment ALPHA_BLENDed over the cleared opaque-black target
 * (src/renderer.rs:178, :437) as UNORM8 (w*h*4 bytes). Row 0 = top of the framebuffer. */
int fluid_sim_render(FluidSim* sim, const float inv_proj[16], uint32_t w, uint32_t h, float* rgba_f32_out,
                     uint8_t* rgba8_out, float* ms_out);

/* ---- dormant physics --------------------------------------------------------
 * calculate_surface_tension (compute.wgsl:300-498; poly6_kernel_gradient / _laplacian funcs.wgsl:81-98) evaluated
 * for every particle of the sim's CURRENT sorted state (predicted positions, densities, keys and start_indices of
 * the last tick, treshold / coefficient / mass / frame_time of that tick's uniform). The reference's call is
 * commented out (compute.wgsl:92), so the result is NOT applied to the particles: force_xy_out (HOST, 2 floats
 * per particle in sorted order) is a diagnostic, bit-identical to the shader arithmetic's CPU restatement. */
int fluid_sim_surface_tension(FluidSim* sim, float* force_xy_out);
/* CursorMoved mapping of src/main.rs:287-303: window pixel -> world position for FluidTickSettings.mouse_pos
 * (ndc = cursor/window*2-1, clip = (ndc.x, -ndc.y, 0, 1), world = (inv_proj*clip).xy / w). Host arithmetic only. */
int fluid_mouse_world_pos(const float cursor_xy[2], const float window_size[2], const float inv_proj[16], float world_xy_out[2]);

/* ---- device self-tests ------------------------------------------------------------ */
/* The exact-math force pass computes the 9 divisions of a neighbour pair (compute.wgsl:206-223, :283-288,
 * funcs.wgsl:112-123) from shared, refined reciprocals. This sweeps every significand of the divisor at 13
 * exponents with per_b numerators each and counts quotients that differ from the correctly rounded one
 * (must be 0). inexact_reciprocals (optional) reports how many refined reciprocals differ from RN(1/b)
 * (informational); first_bad (optional, 4 floats) = a, b, wanted, got of one mismatch. */
int fluid_selftest_division(int device_ordinal, uint32_t seed, int per_b, uint64_t* tested, uint64_t* mismatches,
                            uint64_t* inexact_reciprocals, float* first_bad);

#ifdef __cplusplus
}
#endif
#endif /* FLUID_B200_H */
