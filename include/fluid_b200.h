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

#define FLUID_B200_ABI_VERSION 2

typedef enum FluidStatus {
    FLUID_OK = 0,
    FLUID_ERR_INVALID_ARGUMENT = -1, /* null pointer, particle_count < 2, bad sizes */
    FLUID_ERR_CUDA = -2,             /* a CUDA runtime call failed; see fluid_last_error */
    FLUID_ERR_NO_DEVICE = -3,        /* no CUDA device / ordinal out of range */
    FLUID_ERR_UNSUPPORTED = -4,      /* e.g. image larger than the JFA seed encoding allows; NCCL not loadable */
    FLUID_ERR_OVERFLOW = -5          /* multi-GPU step: an exchange window or a band's capacity overflowed (particles lost) */
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
 *   after an upload), radix otherwise. The bands of the multi-GPU step always count (ranked by origin, see fluid_multi_*).
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
 * for `src`, until the next _upload_particles_async on the same handle has been issued AND a later sync returned. */
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

/* ---- the multi-GPU step: SURVEY.md §8(e) ------------------------------------------------------------------------------
 * The domain is cut into bands of grid rows, one band per GPU ("rank"); keys stay global. The reference has no multi-GPU
 * code; its caller makes one call per frame, `FluidSimulation::tick` (src/simulation.rs:459, from src/renderer.rs:415-418),
 * and so does a caller of this block: fluid_multi_tick enqueues the whole step of every band this process drives — kernels
 * and the tick's single exchange (the boundary rows' particles as ghosts + the particles that changed bands, in one
 * fixed-size window per pair of ranks) — and returns without reading anything back: counts, offsets and halo sizes stay on
 * the device. The contract: the bands, concatenated in rank order, are bit for bit the state a single FluidSim holds after
 * the same ticks (same keys, same order inside a cell, same floats, same start_indices).
 *   fluid_multi_create       one process drives all ranks (one host thread; what a Rust `FluidSimulation::new_multi` binds)
 *   fluid_multi_create_rank  one rank per process (torchrun / mpirun style); the ranks share fluid_multi_unique_id's bytes
 * Errors that only the device can see (an exchange window or a band's capacity overflowed) are latched there and reported
 * as FLUID_ERR_OVERFLOW by the next call that synchronises (fluid_multi_sync, downloads). */
#define FLUID_TRANSPORT_AUTO 0   /* one process: NCCL when every rank has its own GPU, LOCAL when ranks share one; one rank per process: PEER if CUDA IPC works between all ranks, else NCCL */
#define FLUID_TRANSPORT_NCCL 1   /* one ncclGroup of send/recv windows per tick (NCCL over NVLink / NVSwitch) */
#define FLUID_TRANSPORT_LOCAL 2  /* ranks of ONE process: pull kernels over peer-mapped (or the same GPU's) memory, ordered by events */
#define FLUID_TRANSPORT_PEER 3   /* one rank per process (fluid_multi_create_rank): the other ranks' outboxes mapped through CUDA IPC; each rank
                                    pulls the records that are there over NVLink once the sender's header carries the tick's exchange number
                                    (NCCL is still used to set up and for the size all-gather of an upload) */
#define FLUID_MULTI_MAX_RANKS 8
#define FLUID_MULTI_ID_BYTES 128 /* sizeof(ncclUniqueId) */
typedef struct FluidMultiOptions {
    uint32_t row_bounds[FLUID_MULTI_MAX_RANKS + 1]; /* rank r owns grid rows [row_bounds[r], row_bounds[r+1]); all zero: equal bands */
    uint32_t capacity;     /* particles a rank can own; 0: particle_count / ranks * 1.25 + 4096 */
    uint32_t window_near;  /* records per exchange window between adjacent ranks (2 ghost rows + migrants each way); 0: max(8 x average row, 8192) */
    uint32_t window_far;   /* records per window between non-adjacent ranks (long jumps: the force-field push-out, compute.wgsl:131-136); 0: 1024 */
    int32_t transport;     /* FLUID_TRANSPORT_* */
    uint32_t reserved[4];
} FluidMultiOptions;
typedef struct FluidMulti FluidMulti; /* opaque */

/* reference: `FluidSimulation::new` (src/simulation.rs:139) on n_ranks GPUs: devices[r] = CUDA ordinal of rank r (the same
 * ordinal may repeat with FLUID_TRANSPORT_LOCAL: several bands on one GPU). Builds the init lattice of :147-163. opt may be NULL. */
int fluid_multi_create(const FluidSimSettings* settings, int n_ranks, const int* devices, const FluidMultiOptions* opt, FluidMulti** out);
/* one rank per process: rank 0 makes the id, the launcher hands the same bytes to every rank. The bands start empty. */
int fluid_multi_unique_id(uint8_t out[FLUID_MULTI_ID_BYTES]);
int fluid_multi_create_rank(const FluidSimSettings* settings, int rank, int world, const uint8_t id[FLUID_MULTI_ID_BYTES], int device,
                            const FluidMultiOptions* opt, FluidMulti** out);
void fluid_multi_destroy(FluidMulti* m); /* FLUID_TRANSPORT_PEER: collective over the ranks, like creation (the ranks read each other's memory) */
/* reference: `FluidSimulation::tick` (src/simulation.rs:459-539); asynchronous on every band's stream */
int fluid_multi_tick(FluidMulti* m, const FluidTickSettings* tick);
int fluid_multi_tick_n(FluidMulti* m, const FluidTickSettings* tick, uint32_t n);
/* reference: `device.poll(Wait)` (src/main.rs:79); returns FLUID_ERR_OVERFLOW if a rank latched an error */
int fluid_multi_sync(FluidMulti* m);
uint32_t fluid_multi_tick_count(const FluidMulti* m);
int fluid_multi_set_tick_count(FluidMulti* m, uint32_t tick);
int fluid_multi_set_math_mode(FluidMulti* m, int mode);
int fluid_multi_world(const FluidMulti* m);
int fluid_multi_local_bands(const FluidMulti* m); /* bands this process drives: world, or 1 */
int fluid_multi_transport(const FluidMulti* m);
int fluid_multi_row_bounds(const FluidMulti* m, uint32_t* out /* world + 1 */);
/* binding 0 of simulation_bg for the whole domain: N x 32 B AoS in the global array order (upload; it fixes the order of
 * particles that share a cell, as on one GPU) / in global key order (download). Handles that drive all ranks. */
int fluid_multi_upload_particles(FluidMulti* m, const FluidParticle* all, uint32_t n);
int fluid_multi_download_particles(FluidMulti* m, FluidParticle* dst);
/* one rank per process. upload: COLLECTIVE; each rank passes any subset of the scene (the first tick moves every particle to
 * the rank that owns its row); origins = the particles' indices in the global array, NULL = the ranks' subsets
 * concatenated in rank order. download: the band in key order, its size and its offset in the global key order. */
int fluid_multi_upload_band(FluidMulti* m, const FluidParticle* band, uint32_t n, const uint32_t* origins);
int fluid_multi_download_band(FluidMulti* m, int local_band, FluidParticle* dst, uint32_t cap, uint32_t* n_out, uint32_t* global_offset_out);
/* Pipelined variants for one-rank-per-process handles (as fluid_sim_upload / _download_particles_async: own copy streams and
 * staging buffers, the upload of the next state overlaps the download of the previous result). Nothing is read back by the
 * host: the upload's band sizes are all-gathered on the device (origins = rank-order concatenation). download copies the
 * band's first `cap` particles (key order) and {size, global offset} into info[2]; src, dst and info should be pinned; valid
 * after fluid_multi_sync_transfers (which also reports latched overflow errors). */
int fluid_multi_upload_band_async(FluidMulti* m, const FluidParticle* src, uint32_t n);
int fluid_multi_download_band_async(FluidMulti* m, FluidParticle* dst, uint32_t cap, uint32_t* info);
int fluid_multi_sync_transfers(FluidMulti* m);
/* synchronises; counts[r] = particles rank r owns (every rank tracks all of them on the device) */
int fluid_multi_particle_counts(FluidMulti* m, uint32_t* counts /* world */);
/* bindings 1 and 2 for the whole domain (start_indices holds GLOBAL sorted indices; one-rank-per-process handles return
 * their own table: valid for the rows they own) */
int fluid_multi_download_start_indices(FluidMulti* m, uint32_t* dst);
int fluid_multi_upload_start_indices(FluidMulti* m, const uint32_t* src);
int fluid_multi_upload_force_field(FluidMulti* m, const float* xy);
/* per-phase device timing of one local band (as fluid_sim_last_timing); kernels launched by all local bands */
int fluid_multi_enable_timing(FluidMulti* m, int enable);
int fluid_multi_last_timing(FluidMulti* m, int local_band, const char** names, float* ms, int cap);
uint64_t fluid_multi_launch_count(const FluidMulti* m);
void* fluid_multi_stream(FluidMulti* m, int local_band); /* cudaStream_t */
/* the FluidSim behind a local band, for read-only queries (device view, uniform); it cannot be ticked on its own */
FluidSim* fluid_multi_band_sim(FluidMulti* m, int local_band);
/* row bounds that give every rank (nearly) the same number of particles, from the positions (scenes whose rows are
 * unevenly filled: the dam break); out: world + 1 entries for FluidMultiOptions.row_bounds. Host arithmetic only. */
int fluid_multi_balanced_bounds(const FluidSimSettings* settings, const FluidParticle* particles, uint32_t n, int world, uint32_t* out);

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

/* fluid_jfa_run for callers that may not allocate (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc). NO COUNTERPART IN THE REFERENCE.
 * Same result as fluid_jfa_run, but the two ping-pong images and the staging image live in one file-scope
 * `static __device__` array of FLUID_JFA_STATIC_MAX_PIXELS * 16 bytes (16 MiB) that every GPU which loads the library
 * holds until exit; the call makes no cudaMalloc / cudaFree / stream or event creation. Images with
 * w * h > FLUID_JFA_STATIC_MAX_PIXELS are rejected with FLUID_ERR_UNSUPPORTED (use fluid_jfa_run).
 *   seeds_in / seeds_out  rg32uint images; device, managed or mapped pinned memory is used where it lies, plain host
 *                         memory is staged through the static array (the call then returns after the copy back);
 *   stream                cudaStream_t the work is enqueued on (NULL = the default stream);
 *   wait                  nonzero: cudaStreamSynchronize(stream) before returning; zero with device pointers: the call
 *                         only enqueues (nothing is read back by the host in between, so it can be captured).
 * All calls on one GPU SHARE the array: two calls whose kernels overlap on the device (different streams without an
 * ordering between them, or several threads) race and produce garbage; serialise them. The array keeps the previous
 * call's contents; nothing in it needs to start at zero (every byte read by a call was written by that call). */
#define FLUID_JFA_STATIC_MAX_PIXELS (1u << 20)
int fluid_jfa_run_static(const uint32_t* seeds_in, uint32_t w, uint32_t h, uint32_t* seeds_out, int device_ordinal,
                         void* stream /* cudaStream_t */, int wait);

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
/* The per-frame form: only the blended UNORM8 image, into rgba8_out (pinned memory for an asynchronous copy).
 * srgb != 0 encodes the colour channels with the sRGB transfer function, as the *Srgb surface format the reference
 * prefers does on store (src/renderer.rs:137-140: `find(|f| f.is_srgb())`) — those are the bytes the reference's window
 * shows; srgb == 0 (and fluid_sim_render) store the linear values. wait == 0 only enqueues the frame and its copy on
 * the sim's stream (complete after fluid_sim_sync). The device images are kept between frames. */
int fluid_sim_render_frame(FluidSim* sim, const float inv_proj[16], uint32_t w, uint32_t h, uint8_t* rgba8_out, int srgb, int wait);

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
