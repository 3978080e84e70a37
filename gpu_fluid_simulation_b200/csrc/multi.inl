// multi.inl — the multi-GPU fluid step behind the C ABI (fluid_multi_*, include/fluid_b200.h): SURVEY.md §8(e).
//
// The reference's caller makes ONE call per frame, FluidSimulation::tick (src/simulation.rs:459, from
// src/renderer.rs:415-418); so does a caller of this file: fluid_multi_tick enqueues, for every band the process
// drives, the kernels of kernels_band.cuh and the one exchange of the tick, and returns. The host reads nothing back:
// particle counts, offsets and halo sizes live on the device (BandDyn), grids are sized by capacity.
//
// Two ways to run it:
//   * fluid_multi_create        one process drives all ranks (the shape a Rust host would use: one thread, no spawn);
//   * fluid_multi_create_rank   one rank per process (torchrun-style launchers); the ranks share an NCCL unique id.
// Transports of the exchange:
//   * FLUID_TRANSPORT_NCCL      one ncclGroup of fixed-size send/recv windows per tick (NCCL over NVLink / NVSwitch);
//   * FLUID_TRANSPORT_LOCAL     ranks of one process: a pull kernel per window over peer-mapped memory (or plain device
//                               memory when the bands share a GPU: that is how the 1-GPU tests run 2..8 ranks), ordered
//                               by CUDA events.
// NCCL is resolved with dlopen at first use: a process that already loaded a libnccl.so.2 (PyTorch bundles one) keeps
// using that copy, a plain C / Rust host gets the system's.
// (textually included at the end of fluid_sim.cu: one translation unit, so the kernels are defined once)
#include <dlfcn.h>
#include <nccl.h>

#include "kernels_band.cuh"

namespace fluid {

struct NcclApi {
    void* handle = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    ncclResult_t (*GetVersion)(int*) = nullptr;
};
static NcclApi g_nccl;
static int nccl_load() {
    if (g_nccl.handle) return FLUID_OK;
    void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) { set_error(std::string("NCCL not found (dlopen libnccl.so.2): ") + dlerror()); return FLUID_ERR_UNSUPPORTED; }
#define FLUID_NCCL_SYM(field, name)                                                                    \
    *(void**)(&g_nccl.field) = dlsym(h, name);                                                         \
    if (!g_nccl.field) { set_error(std::string("NCCL symbol missing: ") + name); return FLUID_ERR_UNSUPPORTED; }
    FLUID_NCCL_SYM(GetUniqueId, "ncclGetUniqueId")
    FLUID_NCCL_SYM(CommInitRank, "ncclCommInitRank")
    FLUID_NCCL_SYM(CommInitAll, "ncclCommInitAll")
    FLUID_NCCL_SYM(CommDestroy, "ncclCommDestroy")
    FLUID_NCCL_SYM(GroupStart, "ncclGroupStart")
    FLUID_NCCL_SYM(GroupEnd, "ncclGroupEnd")
    FLUID_NCCL_SYM(Send, "ncclSend")
    FLUID_NCCL_SYM(Recv, "ncclRecv")
    FLUID_NCCL_SYM(AllGather, "ncclAllGather")
    FLUID_NCCL_SYM(GetErrorString, "ncclGetErrorString")
    FLUID_NCCL_SYM(GetVersion, "ncclGetVersion")
#undef FLUID_NCCL_SYM
    g_nccl.handle = h;
    return FLUID_OK;
}
static int nccl_fail(ncclResult_t r, const char* what) {
    set_error(std::string("NCCL error in ") + what + ": " + (g_nccl.GetErrorString ? g_nccl.GetErrorString(r) : "?"));
    return FLUID_ERR_CUDA;
}
#define FLUID_NCCL(call)                                                    \
    do {                                                                    \
        ncclResult_t r__ = (call);                                          \
        if (r__ != ncclSuccess) return ::fluid::nccl_fail(r__, #call);      \
    } while (0)

struct Band {
    Sim* sim = nullptr;
    BandRows rows{};
    uint32_t capacity = 0;     // owned particles the band can hold
    uint32_t state_cap = 0;    // natives + received records (length of every particle array)
    uint32_t key_lo = 0, last_bin = 0, bin_tiles = 0;
    uint32_t* origin = nullptr;
    BandDyn* dyn = nullptr;
    uint8_t* box_mem = nullptr;
    BandBoxes out{}, in{};
    BandBoxes out_alt{};               // PEER transport: the outboxes of odd exchanges (double buffering); else = out
    BandSpec* spec_dev_alt = nullptr;  // ... and the device-resident spec that points at them (else = spec_dev)
    void* peer_base[BAND_MAX_RANKS] = {};  // PEER: the other ranks' box memory, opened through CUDA IPC
    PeerBoxes peer_out[2] = {};            // PEER: rank s's outbox for this rank, by exchange parity
    uint32_t recv_slots = 0;
    bool explicit_origins = false;  // the next tick is the first after an upload: natives carry explicit origins
    cudaEvent_t ev_key = nullptr, ev_copied = nullptr;
    ncclComm_t comm = nullptr;
    uint32_t* scratch_word = nullptr;  // device word for the upload-time all-gather of the band sizes
    BandSpec* spec_dev = nullptr;      // rows, outboxes, counters: what a key pass needs, resident on the device
    // speculation (as on one GPU): the force pass of tick t runs tick t+1's key pass in its epilogue — keys, histogram,
    // copies for the other ranks, message headers — valid if t+1 uses the same delta and nothing replaced the state
    bool spec_valid = false, count_dirty = false;
    float spec_delta = 0.0f;
};

struct Multi {
    FluidSimSettings settings{};
    int world = 0;
    int transport = FLUID_TRANSPORT_NCCL;
    bool single_process = true;
    std::vector<Band*> bands;  // the bands this process drives (all of them, or one)
    uint32_t bounds[BAND_MAX_RANKS + 1] = {};
    uint32_t grid_w = 0, grid_h = 0;
    uint32_t window_near = 0, window_far = 0;
    uint32_t tick = 0;
    bool copies_pending = false;  // LOCAL transport: the previous tick's pull kernels have been enqueued
    uint32_t n_total = 0;
    uint32_t xseq = 0;  // exchanges so far (= ticks enqueued on this handle); never rewound, the same on every rank
};

int sim_create_internal(const FluidSimSettings* s, int device_ordinal, uint32_t capacity, bool lattice, Sim** out);

static void band_free(Band* b) {
    if (!b) return;
    if (b->sim) {
        cudaSetDevice(b->sim->device);
        if (b->sim->stream) cudaStreamSynchronize(b->sim->stream);
        if (b->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(b->comm);
        if (b->origin) cudaFree(b->origin);
        if (b->dyn) cudaFree(b->dyn);
        if (b->box_mem) cudaFree(b->box_mem);
        if (b->scratch_word) cudaFree(b->scratch_word);
        for (int r = 0; r < BAND_MAX_RANKS; r++) if (b->peer_base[r]) cudaIpcCloseMemHandle(b->peer_base[r]);
        if (b->spec_dev_alt && b->spec_dev_alt != b->spec_dev) cudaFree(b->spec_dev_alt);
        if (b->spec_dev) cudaFree(b->spec_dev);
        if (b->ev_key) cudaEventDestroy(b->ev_key);
        if (b->ev_copied) cudaEventDestroy(b->ev_copied);
        fluid_sim_destroy(reinterpret_cast<FluidSim*>(b->sim));
    }
    delete b;
}

static uint32_t round_up4(uint32_t v) { return (v + 3u) & ~3u; }

// Exchange windows of a rank: per other rank r, in rank order, [outbox | inbox | second outbox] of the same size. The layout is a
// function of (world, rank, window sizes) only, so that a rank can address another rank's boxes inside its IPC-mapped block.
static uint32_t band_window_cap(const Multi* m, int rank, int r) {
    return round_up4((r == rank - 1 || r == rank + 1) ? m->window_near : m->window_far);
}
static size_t band_box_slot_bytes(const Multi* m, int rank, int r) { return (band_box_bytes(band_window_cap(m, rank, r)) + 255) & ~(size_t)255; }
static size_t band_box_offset(const Multi* m, int rank, int r, int which /* 0 out, 1 in, 2 second out */) {
    size_t at = 0;
    for (int q = 0; q < r; q++) if (q != rank) at += 3 * band_box_slot_bytes(m, rank, q);
    return at + (size_t)which * band_box_slot_bytes(m, rank, r);
}
static size_t band_box_mem_bytes(const Multi* m, int rank) {
    size_t at = 0;
    for (int q = 0; q < m->world; q++) if (q != rank) at += 3 * band_box_slot_bytes(m, rank, q);
    return at;
}

// the band of `rank`: arrays, cell table rows, windows
static int band_create(Multi* m, int rank, int device, uint32_t capacity, Band** out) {
    Band* b = new Band();
    *out = b;
    b->rows.world = m->world;
    b->rows.rank = rank;
    for (int r = 0; r <= m->world; r++) b->rows.bounds[r] = m->bounds[r];
    const uint32_t lo = m->bounds[rank], hi = m->bounds[rank + 1], gh = m->grid_h, gw = m->grid_w;
    const uint32_t ext_lo = lo >= (uint32_t)BAND_GHOST_ROWS ? lo - BAND_GHOST_ROWS : 0u;
    const uint32_t ext_hi = std::min(hi + BAND_GHOST_ROWS, gh);
    const uint32_t table_lo = ext_lo >= 1u ? ext_lo - 1u : 0u;  // one always-empty row per side: the neighbour walks of the outermost ghost rows stay inside the table
    const uint32_t table_hi = std::min(ext_hi + 1u, gh);
    b->rows.table_lo = table_lo;
    b->key_lo = table_lo * gw;
    b->last_bin = (table_hi - table_lo) * gw;  // the bin of the sentinel key; its lower bound closes the table
    b->bin_tiles = (uint32_t)(((size_t)b->last_bin + 1 + CELL_SCAN_TILE - 1) / CELL_SCAN_TILE);  // entries 0 .. last_bin
    uint32_t recv = 0;
    for (int r = 0; r < m->world; r++) {
        if (r == rank) continue;
        const uint32_t cap = band_window_cap(m, rank, r);
        b->out.cap[r] = b->in.cap[r] = cap;
        recv += cap;
    }
    b->recv_slots = recv;
    b->capacity = capacity;
    b->state_cap = capacity + recv;
    int rc = sim_create_internal(&m->settings, device, b->state_cap, false, &b->sim);
    if (rc != FLUID_OK) return rc;
    Sim* s = b->sim;
    s->band = true;
    s->n = 0;
    s->keys_sorted = s->keys_b;
    s->lb_view = s->cell_lb - b->key_lo;  // indexed by global cell id; only the table's rows are ever read
    if (b->bin_tiles > s->cell_tiles) { set_error("band_create: table larger than the grid's"); return FLUID_ERR_INVALID_ARGUMENT; }
    FLUID_CUDA(cudaMalloc(&b->origin, ((size_t)b->state_cap + 64) * 4));
    FLUID_CUDA(cudaMalloc(&b->dyn, sizeof(BandDyn)));
    FLUID_CUDA(cudaMemsetAsync(b->dyn, 0, sizeof(BandDyn), s->stream));
    FLUID_CUDA(cudaMalloc(&b->scratch_word, 128));  // [0] this band's size, [4..12) the gathered sizes, [12..14) download info
    const size_t bytes = band_box_mem_bytes(m, rank);
    b->out_alt = b->out;  // (caps; the boxes follow)
    if (bytes) {
        FLUID_CUDA(cudaMalloc(&b->box_mem, bytes));
        FLUID_CUDA(cudaMemsetAsync(b->box_mem, 0, bytes, s->stream));
        for (int r = 0; r < m->world; r++) {
            if (r == rank) continue;
            b->out.box[r] = reinterpret_cast<uint32_t*>(b->box_mem + band_box_offset(m, rank, r, 0));
            b->in.box[r] = reinterpret_cast<uint32_t*>(b->box_mem + band_box_offset(m, rank, r, 1));
            b->out_alt.box[r] = reinterpret_cast<uint32_t*>(b->box_mem + band_box_offset(m, rank, r, 2));
        }
    }
    {
        BandSpec spec;
        memset(&spec, 0, sizeof spec);
        spec.rows = b->rows; spec.out = b->out; spec.dyn = b->dyn; spec.key_lo = b->key_lo; spec.last_bin = b->last_bin;
        FLUID_CUDA(cudaMalloc(&b->spec_dev, sizeof spec));
        FLUID_CUDA(cudaMemcpyAsync(b->spec_dev, &spec, sizeof spec, cudaMemcpyHostToDevice, s->stream));
        spec.out = b->out_alt;
        FLUID_CUDA(cudaMalloc(&b->spec_dev_alt, sizeof spec));
        FLUID_CUDA(cudaMemcpyAsync(b->spec_dev_alt, &spec, sizeof spec, cudaMemcpyHostToDevice, s->stream));
        FLUID_CUDA(cudaStreamSynchronize(s->stream));
    }
    FLUID_CUDA(cudaEventCreateWithFlags(&b->ev_key, cudaEventDisableTiming));
    FLUID_CUDA(cudaEventCreateWithFlags(&b->ev_copied, cudaEventDisableTiming));
    FLUID_CUDA(cudaStreamSynchronize(s->stream));
    return FLUID_OK;
}

// particles of one band in global array order -> the band's state arrays; origins explicit for the first tick
static int band_upload(Band* b, const FluidParticle* part, uint32_t n, const uint32_t* origins, uint32_t origin0) {
    Sim* s = b->sim;
    FLUID_CUDA(cudaSetDevice(s->device));
    if (n > b->capacity) { set_error("fluid_multi upload: a band receives more particles than its capacity"); return FLUID_ERR_INVALID_ARGUMENT; }
    s->n = n;
    s->sorted_off = 0;
    if (n) {
        int rc = fluid_sim_upload_particles(reinterpret_cast<FluidSim*>(s), part);
        if (rc != FLUID_OK) return rc;
        // the uploaded keys landed in keys_view() = keys_b; nothing reads them before the first tick rewrites them
        std::vector<uint32_t> org;
        if (!origins) {
            org.resize(n);
            for (uint32_t i = 0; i < n; i++) org[i] = origin0 + i;
            origins = org.data();
        }
        FLUID_CUDA(cudaMemcpyAsync(b->origin, origins, (size_t)n * 4, cudaMemcpyHostToDevice, s->stream));
        FLUID_CUDA(cudaStreamSynchronize(s->stream));
    }
    b->explicit_origins = true;
    if (b->spec_valid) b->count_dirty = true;  // the speculative histogram of the replaced state is dropped by the next tick
    b->spec_valid = false;
    return FLUID_OK;
}
// after every band's upload: the device-side counters (own range, global offset, every rank's count)
static int band_set_counts(Band* b, const uint32_t* n_all) {
    Sim* s = b->sim;
    FLUID_CUDA(cudaSetDevice(s->device));
    BandDyn d;
    memset(&d, 0, sizeof d);
    uint32_t goff = 0;
    for (int r = 0; r < b->rows.world; r++) { d.n_all[r] = n_all[r]; if (r < b->rows.rank) goff += n_all[r]; }
    d.own_begin = 0; d.own_end = n_all[b->rows.rank];
    d.dens_begin = 0; d.dens_end = d.own_end;
    d.goff = goff; d.goff_prev = goff;
    FLUID_CUDA(cudaMemcpyAsync(b->dyn, &d, sizeof d, cudaMemcpyHostToDevice, s->stream));
    FLUID_CUDA(cudaStreamSynchronize(s->stream));
    return FLUID_OK;
}

// ---- the tick, in two halves around the exchange ---------------------------------------------------------
// the spec (hence the outboxes) the key pass of exchange `seq` fills: PEER alternates between two sets
static BandSpec* band_spec_of(const Multi* m, const Band* b, uint32_t seq) {
    return (m->transport == FLUID_TRANSPORT_PEER && (seq & 1u)) ? b->spec_dev_alt : b->spec_dev;
}
static int band_enqueue_key(Multi* m, Band* b, const FluidTickSettings& t, DevParams* p_out) {
    Sim* s = b->sim;
    FLUID_CUDA(cudaSetDevice(s->device));
    s->tick = m->tick;
    s->make_uniform(t, s->tick, &s->last_uniform);
    s->last_uniform.particle_count = m->n_total;
    DevParams p = s->make_params(s->last_uniform);
    p.n = 0; p.own_begin = 0;
    *p_out = p;
    s->n_marks = 0;
    s->mark("begin");
    s->cell_count = s->cell_count_buf[s->cell_count_cur];
    // did the previous tick's force pass already run this tick's key pass (same delta, state untouched)?
    const bool hit = b->spec_valid && !b->explicit_origins && memcmp(&b->spec_delta, &t.delta, sizeof(float)) == 0;
    b->spec_valid = false;
    if (!hit) {
        if (b->count_dirty) {  // unused speculation: drop its histogram and its pushes into the outboxes
            FLUID_CUDA(cudaMemsetAsync(s->cell_count, 0, (size_t)b->bin_tiles * CELL_SCAN_TILE * 4, s->stream));
#ifndef FLUID_AB_NO_FORGET  // (A/B builds only: shows that tests/test_gpu_multi.py::test_changing_delta_* needs this launch)
            k_band_forget_pushes<<<1, 32, 0, s->stream>>>(b->dyn);
            s->launches++;
#endif
        }
        uint32_t blocks = (b->capacity + BAND_KEY_THREADS - 1) / BAND_KEY_THREADS;
        const uint32_t max_blocks = (uint32_t)s->num_sms * 8u;
        if (blocks > max_blocks) blocks = max_blocks;
        if (blocks == 0) blocks = 1;
        launch_pdl(s->pdl, k_band_key, dim3(blocks), dim3(BAND_KEY_THREADS), s->stream, p, (const BandSpec*)band_spec_of(m, b, m->xseq), (const float2*)s->pos_a,
                   (const float2*)s->vel_a, s->keys_u, s->cell_count, (const uint32_t*)(b->explicit_origins ? b->origin : nullptr));
        s->launches++;
    }
    b->count_dirty = false;
    launch_pdl(s->pdl, k_band_headers, dim3(1), dim3(32), s->stream, (const BandSpec*)band_spec_of(m, b, m->xseq), m->xseq);
    s->launches++;
    s->mark("predict+key+hist");
    FLUID_CUDA(cudaGetLastError());
    return FLUID_OK;
}

static int band_enqueue_rest(Multi* m, Band* b, const DevParams& p) {
    Sim* s = b->sim;
    FLUID_CUDA(cudaSetDevice(s->device));
    s->mark("exchange");
    uint32_t* count = s->cell_count;
    uint32_t* count_next = s->cell_count_buf[s->cell_count_cur ^ 1];
    s->cell_count_cur ^= 1;
    const uint32_t rblocks = std::max(1u, (b->recv_slots + BAND_KEY_THREADS - 1) / BAND_KEY_THREADS);
    const bool pdl = s->pdl;
    launch_pdl(pdl, k_band_recv, dim3(rblocks), dim3(BAND_KEY_THREADS), s->stream, p, b->rows, b->in, s->pos_a, s->vel_a, s->keys_u, b->origin, count,
               b->key_lo, b->dyn, b->state_cap, s->d_frame_time, s->quirk);
    launch_pdl(pdl, k_cell_reduce, dim3(b->bin_tiles), dim3(CELL_SCAN_THREADS), s->stream, (const uint32_t*)count, s->cell_tile_sum, s->cell_tile_max,
               s->cell_tile_sq, s->d_frame_time, 0u);
    launch_pdl(pdl, k_cell_downsweep, dim3(b->bin_tiles), dim3(CELL_SCAN_THREADS), s->stream, count, s->cell_lb, (const uint32_t*)s->cell_tile_sum,
               (const uint32_t*)s->cell_tile_max, (const float*)s->cell_tile_sq, b->bin_tiles, s->cell_max_dev, count_next);
    launch_pdl(pdl, k_cell_slots, dim3((b->state_cap + GATHER_THREADS * SLOT_ITEMS - 1) / (GATHER_THREADS * SLOT_ITEMS)), dim3(GATHER_THREADS), s->stream,
               b->state_cap, b->key_lo, b->last_bin, (const uint32_t*)s->keys_u, count, s->keys_b, s->idx_a, (const uint32_t*)&b->dyn->total);
    s->mark("sort");
    const uint32_t tiles = (b->state_cap + SPH_THREADS - 1) / SPH_THREADS;
    launch_pdl(pdl, k_band_rank_gather, dim3((b->state_cap + GATHER_THREADS * BAND_RANK_ITEMS - 1) / (GATHER_THREADS * BAND_RANK_ITEMS)), dim3(GATHER_THREADS),
               s->stream, p, b->rows, b->key_lo, b->last_bin, (const uint32_t*)s->keys_b, (const uint32_t*)s->idx_a, (const uint32_t*)s->cell_lb,
               (const float2*)s->pos_a, (const float2*)s->vel_a, (const uint32_t*)b->origin, b->explicit_origins ? 1 : 0, b->dyn, s->pos_s, s->pred_s,
               s->vel_s, s->start_indices, s->quirk, (const uint32_t*)s->d_frame_time, reinterpret_cast<TilePlan*>(s->tile_plans));
    s->mark("gather+cells");
    if (s->tiles_both_ends())
        launch_pdl(pdl, k_density<true, true>, dim3(tiles), dim3(SPH_THREADS), s->stream, p, (const float2*)s->pred_s, (const uint32_t*)s->keys_b, (const uint32_t*)s->lb_view,
               (const uint32_t*)s->quirk, s->rho, s->quirk + 2, (const uint32_t*)s->d_frame_time, reinterpret_cast<const TilePlan*>(s->tile_plans),
               (const uint32_t*)&b->dyn->dens_begin, s->nb_head, s->nb_mask);
    else
        launch_pdl(pdl, k_density<true, false>, dim3(tiles), dim3(SPH_THREADS), s->stream, p, (const float2*)s->pred_s, (const uint32_t*)s->keys_b, (const uint32_t*)s->lb_view,
               (const uint32_t*)s->quirk, s->rho, s->quirk + 2, (const uint32_t*)s->d_frame_time, reinterpret_cast<const TilePlan*>(s->tile_plans),
               (const uint32_t*)&b->dyn->dens_begin, s->nb_head, s->nb_mask);
    s->launches += 6;
    s->mark("density");
    s->keys_sorted = s->keys_b;
    // the force pass also runs the NEXT tick's key pass (keys into keys_u: the slot pass is done with it; histogram into the
    // count array the next tick will use, cleared by this tick's scan; copies into the outboxes: this tick's exchange has read them)
    const bool spec = s->speculate;
    if (spec && m->transport == FLUID_TRANSPORT_LOCAL && m->world > 1) {
        // the epilogue writes this band's outboxes: every destination must have pulled this tick's records out of them
        for (Band* d : m->bands) if (d != b) FLUID_CUDA(cudaStreamWaitEvent(s->stream, d->ev_copied, 0));
    }
    launch_force(s, p, tiles, s->pos_a, s->vel_a, pdl, spec ? s->keys_u : nullptr, spec ? count_next : nullptr, &b->dyn->own_begin, &b->dyn->goff,
                 spec ? band_spec_of(m, b, m->xseq + 1u) : nullptr);
    if (spec) { b->spec_valid = true; b->count_dirty = true; b->spec_delta = s->last_uniform.delta; }
    s->mark("force+integrate");
    if (s->timing) { s->marks_per_tick = s->n_marks; s->timed_ticks++; }
    b->explicit_origins = false;
    FLUID_CUDA(cudaGetLastError());
    return FLUID_OK;
}

constexpr uint64_t PEER_WAIT_LIMIT_NS = 30ull * 1000 * 1000 * 1000;  // a rank that is 30 s late is reported as lost
static int multi_exchange(Multi* m) {
    if (m->world == 1) return FLUID_OK;
    if (m->transport == FLUID_TRANSPORT_LOCAL) {
        // dest pulls from every source's outbox once that source's key pass is done
        for (Band* d : m->bands) {
            FLUID_CUDA(cudaSetDevice(d->sim->device));
            for (Band* s : m->bands) {
                if (s == d) continue;
                FLUID_CUDA(cudaStreamWaitEvent(d->sim->stream, s->ev_key, 0));
                const uint32_t cap = d->in.cap[s->rows.rank];
                k_band_copy_box<<<(std::max(cap, 16u) + 255) / 256, 256, 0, d->sim->stream>>>(
                    reinterpret_cast<const uint4*>(s->out.box[d->rows.rank]), reinterpret_cast<uint4*>(d->in.box[s->rows.rank]), cap);
                d->sim->launches++;
            }
            FLUID_CUDA(cudaEventRecord(d->ev_copied, d->sim->stream));
        }
        m->copies_pending = true;
        FLUID_CUDA(cudaGetLastError());
        return FLUID_OK;
    }
    if (m->transport == FLUID_TRANSPORT_PEER) {
        Band* b = m->bands[0];
        Sim* s = b->sim;
        const int me = b->rows.rank;
        uint32_t cap_max = 16;
        for (int r = 0; r < m->world; r++) if (r != me) cap_max = std::max(cap_max, b->in.cap[r]);
        const PeerBoxes& src = b->peer_out[m->xseq & 1u];
        k_band_wait_peers<<<1, 32, 0, s->stream>>>(src, m->world, me, m->xseq, b->dyn, PEER_WAIT_LIMIT_NS);
        k_band_pull<<<dim3((cap_max + 255) / 256, m->world), 256, 0, s->stream>>>(src, b->in, me);
        s->launches += 2;
        FLUID_CUDA(cudaGetLastError());
        return FLUID_OK;
    }
    FLUID_NCCL(g_nccl.GroupStart());
    for (Band* b : m->bands) {
        const int me = b->rows.rank;
        for (int r = 0; r < m->world; r++) {
            if (r == me) continue;
            const size_t bytes = band_box_bytes(b->out.cap[r]);
            FLUID_NCCL(g_nccl.Send(b->out.box[r], bytes, ncclChar, r, b->comm, b->sim->stream));
            FLUID_NCCL(g_nccl.Recv(b->in.box[r], bytes, ncclChar, r, b->comm, b->sim->stream));
        }
    }
    FLUID_NCCL(g_nccl.GroupEnd());
    return FLUID_OK;
}

static int multi_enqueue_tick(Multi* m, const FluidTickSettings& t) {
    m->tick += 1;  // src/simulation.rs:460
    m->xseq += 1;
    DevParams params[BAND_MAX_RANKS];
    if (m->transport == FLUID_TRANSPORT_LOCAL && m->copies_pending) {
        // a source may overwrite its outboxes once every destination has pulled the previous tick's records
        for (Band* s : m->bands) {
            FLUID_CUDA(cudaSetDevice(s->sim->device));
            for (Band* d : m->bands) if (d != s) FLUID_CUDA(cudaStreamWaitEvent(s->sim->stream, d->ev_copied, 0));
        }
    }
    size_t k = 0;
    for (Band* b : m->bands) {
        int rc = band_enqueue_key(m, b, t, &params[k++]);
        if (rc != FLUID_OK) return rc;
        if (m->transport == FLUID_TRANSPORT_LOCAL) FLUID_CUDA(cudaEventRecord(b->ev_key, b->sim->stream));
    }
    int rc = multi_exchange(m);
    if (rc != FLUID_OK) return rc;
    k = 0;
    for (Band* b : m->bands) {
        rc = band_enqueue_rest(m, b, params[k++]);
        if (rc != FLUID_OK) return rc;
    }
    return FLUID_OK;
}

static void default_bounds(uint32_t grid_h, int world, uint32_t* b) {
    // rows 1 .. grid_h-2 can hold particles; the empty border rows go to the edge ranks
    const uint32_t inner = grid_h > 2 ? grid_h - 2 : 1;
    const uint32_t per = (inner + world - 1) / world;
    b[0] = 0;
    for (int r = 1; r < world; r++) b[r] = std::min(1u + (uint32_t)r * per, grid_h);
    b[world] = grid_h;
}

static int multi_init_common(Multi* m, const FluidSimSettings* s, int world, const FluidMultiOptions* opt) {
    if (!s) { set_error("fluid_multi_create: null settings"); return FLUID_ERR_INVALID_ARGUMENT; }
    if (world < 1 || world > BAND_MAX_RANKS) { set_error("fluid_multi_create: 1 <= ranks <= 8"); return FLUID_ERR_INVALID_ARGUMENT; }
    if (s->particle_count < 2) { set_error("fluid_multi_create: particle_count < 2"); return FLUID_ERR_INVALID_ARGUMENT; }
    if (!(s->smoothing_radius > 0.0f) || !(s->size[0] > 0.0f) || !(s->size[1] > 0.0f)) { set_error("fluid_multi_create: smoothing_radius and size must be positive"); return FLUID_ERR_INVALID_ARGUMENT; }
    m->settings = *s;
    m->world = world;
    m->n_total = s->particle_count;
    m->grid_w = (uint32_t)(size_t)ceilf(s->size[0] / s->smoothing_radius) + 2u;  // src/simulation.rs:140-141
    m->grid_h = (uint32_t)(size_t)ceilf(s->size[1] / s->smoothing_radius) + 2u;
    bool have_bounds = false;
    if (opt) for (int r = 0; r <= world; r++) if (opt->row_bounds[r] != 0) have_bounds = true;
    if (have_bounds) {
        for (int r = 0; r <= world; r++) m->bounds[r] = opt->row_bounds[r];
        if (m->bounds[0] != 0 || m->bounds[world] != m->grid_h) { set_error("fluid_multi_create: row_bounds must start at 0 and end at grid_h"); return FLUID_ERR_INVALID_ARGUMENT; }
        for (int r = 0; r < world; r++) if (m->bounds[r] > m->bounds[r + 1]) { set_error("fluid_multi_create: row_bounds must be non-decreasing"); return FLUID_ERR_INVALID_ARGUMENT; }
    } else {
        default_bounds(m->grid_h, world, m->bounds);
    }
    const uint32_t rows_inner = std::max(1u, m->grid_h - 2u);
    const uint32_t per_row = (uint32_t)std::min<uint64_t>(((uint64_t)s->particle_count + rows_inner - 1) / rows_inner, 1u << 26);
    m->window_near = (opt && opt->window_near) ? opt->window_near : std::max(8u * per_row, 8192u);
    m->window_far = (opt && opt->window_far) ? opt->window_far : 1024u;
    m->transport = opt ? opt->transport : FLUID_TRANSPORT_AUTO;
    return FLUID_OK;
}
static uint32_t default_capacity(const Multi* m, const FluidMultiOptions* opt) {
    if (opt && opt->capacity) return opt->capacity;
    const uint64_t fair = ((uint64_t)m->n_total + m->world - 1) / m->world;
    return (uint32_t)std::min<uint64_t>(fair + fair / 4 + 4096, 0xfffffff0ull);
}

// grid row of a particle's position (host arithmetic of funcs.wgsl:212-214): only used to pick the band a particle is
// FIRST uploaded to; ownership is re-derived from the keys on every tick
static uint32_t host_row_of(const Multi* m, float y) {
    float half = m->settings.size[1] * 0.5f;
    if (fabsf(y) > half) y = half * (y > 0.f ? 1.f : -1.f);
    float f = floorf((y + half) / m->settings.smoothing_radius);
    if (!(f >= 0.f)) f = 0.f;
    uint32_t r = (uint32_t)std::min<double>(f, 4.0e9) + 1u;
    return std::min(r, m->grid_h - 1u);
}
static int host_owner(const Multi* m, uint32_t row) {
    int o = 0;
    for (int r = 1; r < m->world; r++) if (row >= m->bounds[r]) o = r;
    return o;
}

static int multi_upload_all(Multi* m, const FluidParticle* all, uint32_t n) {
    if (!m->single_process) { set_error("fluid_multi_upload_particles: one-rank-per-process handles upload their band (fluid_multi_upload_band)"); return FLUID_ERR_INVALID_ARGUMENT; }
    if (n != m->n_total) { set_error("fluid_multi_upload_particles: n differs from settings.particle_count"); return FLUID_ERR_INVALID_ARGUMENT; }
    std::vector<std::vector<FluidParticle>> part(m->world);
    std::vector<std::vector<uint32_t>> org(m->world);
    for (uint32_t i = 0; i < n; i++) {
        const int o = host_owner(m, host_row_of(m, all[i].position[1]));
        part[o].push_back(all[i]);
        org[o].push_back(i);
    }
    uint32_t n_all[BAND_MAX_RANKS] = {};
    for (int r = 0; r < m->world; r++) {
        n_all[r] = (uint32_t)part[r].size();
        int rc = band_upload(m->bands[r], part[r].data(), n_all[r], org[r].data(), 0);
        if (rc != FLUID_OK) return rc;
    }
    for (Band* b : m->bands) { int rc = band_set_counts(b, n_all); if (rc != FLUID_OK) return rc; }
    m->copies_pending = false;
    return FLUID_OK;
}

// PEER transport set-up (collective over the ranks of fluid_multi_create_rank): export this rank's box memory through CUDA
// IPC, all-gather the handles over NCCL, open the others', all-gather who succeeded; PEER only if everybody did.
static int band_open_peers(Multi* m, Band* b) {
    Sim* s = b->sim;
    const int me = b->rows.rank, world = m->world;
    FLUID_CUDA(cudaSetDevice(s->device));
    FLUID_CUDA(cudaStreamSynchronize(s->stream));  // the boxes are zeroed before anybody can look at them
    constexpr size_t HB = sizeof(cudaIpcMemHandle_t);
    static_assert(HB == 64, "cudaIpcMemHandle_t");
    uint8_t* dev = nullptr;
    FLUID_CUDA(cudaMalloc(&dev, HB * (BAND_MAX_RANKS + 1) + 64 * (BAND_MAX_RANKS + 1)));
    uint8_t* dev_ok = dev + HB * (BAND_MAX_RANKS + 1);
    cudaIpcMemHandle_t mine;
    memset(&mine, 0, sizeof mine);
    uint32_t ok = (b->box_mem && cudaIpcGetMemHandle(&mine, b->box_mem) == cudaSuccess) ? 1u : 0u;
    cudaGetLastError();
    std::vector<cudaIpcMemHandle_t> all(world);
    auto cleanup = [&](int code) { cudaFree(dev); return code; };
    if (cudaMemcpyAsync(dev, &mine, HB, cudaMemcpyHostToDevice, s->stream) != cudaSuccess) return cleanup(FLUID_ERR_CUDA);
    { ncclResult_t r = g_nccl.AllGather(dev, dev + HB, HB, ncclChar, b->comm, s->stream); if (r != ncclSuccess) return cleanup(nccl_fail(r, "ncclAllGather (IPC handles)")); }
    if (cudaMemcpyAsync(all.data(), dev + HB, HB * world, cudaMemcpyDeviceToHost, s->stream) != cudaSuccess || cudaStreamSynchronize(s->stream) != cudaSuccess) return cleanup(FLUID_ERR_CUDA);
    if (ok) {
        for (int r = 0; r < world && ok; r++) {
            if (r == me) continue;
            if (cudaIpcOpenMemHandle(&b->peer_base[r], all[r], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); b->peer_base[r] = nullptr; ok = 0; }
        }
    }
    uint32_t oks[BAND_MAX_RANKS] = {};
    if (cudaMemcpyAsync(dev_ok, &ok, 4, cudaMemcpyHostToDevice, s->stream) != cudaSuccess) return cleanup(FLUID_ERR_CUDA);
    { ncclResult_t r = g_nccl.AllGather(dev_ok, dev_ok + 64, 1, ncclUint32, b->comm, s->stream); if (r != ncclSuccess) return cleanup(nccl_fail(r, "ncclAllGather (IPC status)")); }
    if (cudaMemcpyAsync(oks, dev_ok + 64, 4 * world, cudaMemcpyDeviceToHost, s->stream) != cudaSuccess || cudaStreamSynchronize(s->stream) != cudaSuccess) return cleanup(FLUID_ERR_CUDA);
    bool everybody = true;
    for (int r = 0; r < world; r++) everybody = everybody && oks[r] == 1u;
    if (!everybody) {
        for (int r = 0; r < world; r++) if (b->peer_base[r]) { cudaIpcCloseMemHandle(b->peer_base[r]); b->peer_base[r] = nullptr; }
        return cleanup(FLUID_OK);  // stays on NCCL
    }
    for (int r = 0; r < world; r++) {
        if (r == me) continue;
        const uint8_t* base = static_cast<const uint8_t*>(b->peer_base[r]);
        b->peer_out[0].box[r] = reinterpret_cast<const uint32_t*>(base + band_box_offset(m, r, me, 0));
        b->peer_out[1].box[r] = reinterpret_cast<const uint32_t*>(base + band_box_offset(m, r, me, 2));
    }
    m->transport = FLUID_TRANSPORT_PEER;
    return cleanup(FLUID_OK);
}

static int multi_check_errors(Multi* m) {
    for (Band* b : m->bands) {
        FLUID_CUDA(cudaSetDevice(b->sim->device));
        uint32_t err = 0;
        FLUID_CUDA(cudaMemcpy(&err, &b->dyn->error, 4, cudaMemcpyDeviceToHost));
        if (err) {
            char buf[256];
            snprintf(buf, sizeof buf, "fluid_multi: rank %d latched error bits 0x%x (1: an exchange window overflowed, 2: band capacity exceeded, "
                     "4: particle count mismatch - particles were lost; recreate with larger windows / capacity; 8: a peer's records did not arrive "
                     "within 30 s)", b->rows.rank, err);
            set_error(buf);
            return FLUID_ERR_OVERFLOW;
        }
    }
    return FLUID_OK;
}

}  // namespace fluid

extern "C" {

#define MULTI_OR_FAIL(h)                                                                  \
    if (!(h)) { set_error("null FluidMulti handle"); return FLUID_ERR_INVALID_ARGUMENT; } \
    Multi* m = reinterpret_cast<Multi*>(h);

void fluid_multi_destroy(FluidMulti* h) {
    if (!h) return;
    Multi* m = reinterpret_cast<Multi*>(h);
    for (Band* b : m->bands) if (b && b->sim && b->sim->stream) { cudaSetDevice(b->sim->device); cudaStreamSynchronize(b->sim->stream); }
    if (m->transport == FLUID_TRANSPORT_PEER && !m->bands.empty() && m->bands[0] && m->bands[0]->comm && m->bands[0]->scratch_word) {
        // PEER: the other ranks read this rank's boxes through their IPC mappings; memory exported through CUDA IPC may only be
        // freed once the importers are done with it. Destruction is therefore collective, like creation: a barrier (an
        // all-gather of one word) after every rank has drained its own stream — its last pull included.
        Band* b = m->bands[0];
        uint32_t* w = reinterpret_cast<uint32_t*>(b->scratch_word);
        if (g_nccl.AllGather(w, w + 4, 1, ncclUint32, b->comm, b->sim->stream) == ncclSuccess) cudaStreamSynchronize(b->sim->stream);
    }
    for (Band* b : m->bands) band_free(b);
    delete m;
}

int fluid_multi_create(const FluidSimSettings* s, int n_ranks, const int* devices, const FluidMultiOptions* opt, FluidMulti** out) {
    if (!out || !devices) { set_error("fluid_multi_create: null argument"); return FLUID_ERR_INVALID_ARGUMENT; }
    *out = nullptr;
    Multi* m = new Multi();
    auto fail = [&](int code) { fluid_multi_destroy(reinterpret_cast<FluidMulti*>(m)); return code; };
    int rc = multi_init_common(m, s, n_ranks, opt);
    if (rc != FLUID_OK) return fail(rc);
    m->single_process = true;
    bool distinct = true;
    for (int a = 0; a < n_ranks; a++) for (int b2 = a + 1; b2 < n_ranks; b2++) if (devices[a] == devices[b2]) distinct = false;
    if (m->transport == FLUID_TRANSPORT_AUTO) m->transport = (distinct && n_ranks > 1) ? FLUID_TRANSPORT_NCCL : FLUID_TRANSPORT_LOCAL;
    if (m->transport == FLUID_TRANSPORT_NCCL && !distinct) { set_error("fluid_multi_create: NCCL needs one GPU per rank (use FLUID_TRANSPORT_LOCAL to share a device)"); return fail(FLUID_ERR_INVALID_ARGUMENT); }
    // the reference's init lattice (src/simulation.rs:147-163); it is split over the bands below
    const uint32_t n = s->particle_count;
    std::vector<FluidParticle> host(n);
    {
        const float ppr = sqrtf((float)n);
        const float ppc = ((float)n - 1.0f) / ppr + 1.0f;
        const size_t ppr_u = (size_t)ppr;
        for (uint32_t i = 0; i < n; i++) {
            const size_t xi = (size_t)i % ppr_u;
            FluidParticle q;
            memset(&q, 0, sizeof q);
            q.position[0] = ((float)xi - ppr * 0.5f + 0.5f) * s->particle_spacing;
            q.position[1] = (floorf((float)i / ppr) - ppc * 0.5f + 0.5f) * s->particle_spacing;
            q.predicted_position[0] = q.position[0]; q.predicted_position[1] = q.position[1];
            host[i] = q;
        }
    }
    uint32_t cap = default_capacity(m, opt);
    if (!(opt && opt->capacity)) {
        // the lattice need not fill the box (the default scene occupies the middle 60 %): size the bands for it
        std::vector<uint32_t> cnt(n_ranks, 0);
        for (uint32_t i = 0; i < n; i++) cnt[host_owner(m, host_row_of(m, host[i].position[1]))]++;
        const uint64_t mx = *std::max_element(cnt.begin(), cnt.end());
        cap = (uint32_t)std::min<uint64_t>(std::max<uint64_t>(cap, mx + mx / 2 + 4096), std::max<uint64_t>(n, 2));
    }
    for (int r = 0; r < n_ranks; r++) {
        Band* b = nullptr;
        rc = band_create(m, r, devices[r], cap, &b);
        m->bands.push_back(b);
        if (rc != FLUID_OK) return fail(rc);
    }
    if (m->transport == FLUID_TRANSPORT_LOCAL) {
        for (int a = 0; a < n_ranks; a++)
            for (int b2 = 0; b2 < n_ranks; b2++) {
                if (devices[a] == devices[b2]) continue;
                int can = 0;
                cudaDeviceCanAccessPeer(&can, devices[a], devices[b2]);
                if (!can) { set_error("fluid_multi_create: FLUID_TRANSPORT_LOCAL needs peer access between the devices"); return fail(FLUID_ERR_UNSUPPORTED); }
                cudaSetDevice(devices[a]);
                cudaError_t e = cudaDeviceEnablePeerAccess(devices[b2], 0);
                if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) return fail(cuda_fail(e, "cudaDeviceEnablePeerAccess", __FILE__, __LINE__));
                cudaGetLastError();
            }
    } else if (n_ranks > 1) {
        if ((rc = nccl_load()) != FLUID_OK) return fail(rc);
        ncclComm_t comms[BAND_MAX_RANKS];
        ncclResult_t r = g_nccl.CommInitAll(comms, n_ranks, devices);
        if (r != ncclSuccess) return fail(nccl_fail(r, "ncclCommInitAll"));
        for (int k = 0; k < n_ranks; k++) m->bands[k]->comm = comms[k];
    }
    if ((rc = multi_upload_all(m, host.data(), n)) != FLUID_OK) return fail(rc);
    *out = reinterpret_cast<FluidMulti*>(m);
    return FLUID_OK;
}

int fluid_multi_unique_id(uint8_t out[FLUID_MULTI_ID_BYTES]) {
    if (!out) { set_error("fluid_multi_unique_id: null argument"); return FLUID_ERR_INVALID_ARGUMENT; }
    int rc = nccl_load();
    if (rc != FLUID_OK) return rc;
    static_assert(sizeof(ncclUniqueId) == FLUID_MULTI_ID_BYTES, "ncclUniqueId size");
    ncclUniqueId id;
    FLUID_NCCL(g_nccl.GetUniqueId(&id));
    memcpy(out, &id, sizeof id);
    return FLUID_OK;
}

int fluid_multi_create_rank(const FluidSimSettings* s, int rank, int world, const uint8_t id_bytes[FLUID_MULTI_ID_BYTES], int device,
                            const FluidMultiOptions* opt, FluidMulti** out) {
    if (!out || (!id_bytes && world > 1)) { set_error("fluid_multi_create_rank: null argument"); return FLUID_ERR_INVALID_ARGUMENT; }
    *out = nullptr;
    if (rank < 0 || rank >= world) { set_error("fluid_multi_create_rank: bad rank"); return FLUID_ERR_INVALID_ARGUMENT; }
    Multi* m = new Multi();
    auto fail = [&](int code) { fluid_multi_destroy(reinterpret_cast<FluidMulti*>(m)); return code; };
    int rc = multi_init_common(m, s, world, opt);
    if (rc != FLUID_OK) return fail(rc);
    m->single_process = false;
    if (m->transport == FLUID_TRANSPORT_LOCAL && world > 1) { set_error("fluid_multi_create_rank: FLUID_TRANSPORT_LOCAL needs all ranks in one process"); return fail(FLUID_ERR_INVALID_ARGUMENT); }
    const char* no_peer = getenv("FLUID_MULTI_NO_PEER");
    const bool want_peer = world > 1 && (m->transport == FLUID_TRANSPORT_PEER || (m->transport == FLUID_TRANSPORT_AUTO && !(no_peer && no_peer[0] == '1')));
    const bool need_peer = m->transport == FLUID_TRANSPORT_PEER;
    m->transport = FLUID_TRANSPORT_NCCL;
    Band* b = nullptr;
    rc = band_create(m, rank, device, default_capacity(m, opt), &b);
    m->bands.push_back(b);
    if (rc != FLUID_OK) return fail(rc);
    if (world > 1) {
        if ((rc = nccl_load()) != FLUID_OK) return fail(rc);
        ncclUniqueId id;
        memcpy(&id, id_bytes, sizeof id);
        cudaSetDevice(device);
        ncclResult_t r = g_nccl.CommInitRank(&b->comm, world, id, rank);
        if (r != ncclSuccess) return fail(nccl_fail(r, "ncclCommInitRank"));
        if (want_peer) {
            rc = band_open_peers(m, b);  // collective: every rank takes the same decision
            if (rc != FLUID_OK) return fail(rc);
            if (need_peer && m->transport != FLUID_TRANSPORT_PEER) { set_error("fluid_multi_create_rank: FLUID_TRANSPORT_PEER asked for, but CUDA IPC between the ranks is not available"); return fail(FLUID_ERR_UNSUPPORTED); }
        }
    }
    // every band starts empty: the caller uploads its band (fluid_multi_upload_band)
    uint32_t zeros[BAND_MAX_RANKS] = {};
    if ((rc = band_set_counts(b, zeros)) != FLUID_OK) return fail(rc);
    *out = reinterpret_cast<FluidMulti*>(m);
    return FLUID_OK;
}

int fluid_multi_upload_particles(FluidMulti* h, const FluidParticle* all, uint32_t n) {
    MULTI_OR_FAIL(h);
    if (!all) { set_error("fluid_multi_upload_particles: null source"); return FLUID_ERR_INVALID_ARGUMENT; }
    return multi_upload_all(m, all, n);
}

// collective over the ranks: every rank passes its own particles (any subset of the scene; the first tick moves each
// particle to the rank that owns its row). origins: the particles' indices in the global array (they fix the order of
// particles that share a cell, as the array order does on one GPU); null = rank-order concatenation.
int fluid_multi_upload_band(FluidMulti* h, const FluidParticle* band, uint32_t n, const uint32_t* origins) {
    MULTI_OR_FAIL(h);
    if (m->single_process) { set_error("fluid_multi_upload_band: this handle drives all ranks (use fluid_multi_upload_particles)"); return FLUID_ERR_INVALID_ARGUMENT; }
    if (!band && n) { set_error("fluid_multi_upload_band: null source"); return FLUID_ERR_INVALID_ARGUMENT; }
    Band* b = m->bands[0];
    Sim* s = b->sim;
    FLUID_CUDA(cudaSetDevice(s->device));
    uint32_t n_all[BAND_MAX_RANKS] = {};
    n_all[b->rows.rank] = n;
    if (m->world > 1) {
        uint32_t* gathered = reinterpret_cast<uint32_t*>(b->scratch_word) + 4;
        FLUID_CUDA(cudaMemcpyAsync(b->scratch_word, &n, 4, cudaMemcpyHostToDevice, s->stream));
        FLUID_NCCL(g_nccl.AllGather(b->scratch_word, gathered, 1, ncclUint32, b->comm, s->stream));
        FLUID_CUDA(cudaMemcpyAsync(n_all, gathered, (size_t)m->world * 4, cudaMemcpyDeviceToHost, s->stream));
        FLUID_CUDA(cudaStreamSynchronize(s->stream));
    }
    uint64_t total = 0;
    uint32_t goff = 0;
    for (int r = 0; r < m->world; r++) { total += n_all[r]; if (r < b->rows.rank) goff += n_all[r]; }
    if (total != m->n_total) { set_error("fluid_multi_upload_band: the bands do not add up to settings.particle_count"); return FLUID_ERR_INVALID_ARGUMENT; }
    int rc = band_upload(b, band, n, origins, goff);
    if (rc != FLUID_OK) return rc;
    return band_set_counts(b, n_all);
}

int fluid_multi_tick(FluidMulti* h, const FluidTickSettings* t) {
    MULTI_OR_FAIL(h);
    if (!t) { set_error("fluid_multi_tick: null settings"); return FLUID_ERR_INVALID_ARGUMENT; }
    return multi_enqueue_tick(m, *t);
}
int fluid_multi_tick_n(FluidMulti* h, const FluidTickSettings* t, uint32_t count) {
    MULTI_OR_FAIL(h);
    if (!t) { set_error("fluid_multi_tick_n: null settings"); return FLUID_ERR_INVALID_ARGUMENT; }
    for (uint32_t k = 0; k < count; k++) {
        int rc = multi_enqueue_tick(m, *t);
        if (rc != FLUID_OK) return rc;
    }
    return FLUID_OK;
}
int fluid_multi_sync(FluidMulti* h) {
    MULTI_OR_FAIL(h);
    for (Band* b : m->bands) {
        FLUID_CUDA(cudaSetDevice(b->sim->device));
        FLUID_CUDA(cudaStreamSynchronize(b->sim->stream));
    }
    return multi_check_errors(m);
}
uint32_t fluid_multi_tick_count(const FluidMulti* h) { return h ? reinterpret_cast<const Multi*>(h)->tick : 0; }
int fluid_multi_set_tick_count(FluidMulti* h, uint32_t tick) {
    MULTI_OR_FAIL(h);
    m->tick = tick;
    for (Band* b : m->bands) { int rc = fluid_sim_set_tick_count(reinterpret_cast<FluidSim*>(b->sim), tick); if (rc != FLUID_OK) return rc; }
    return FLUID_OK;
}
int fluid_multi_world(const FluidMulti* h) { return h ? reinterpret_cast<const Multi*>(h)->world : 0; }
int fluid_multi_local_bands(const FluidMulti* h) { return h ? (int)reinterpret_cast<const Multi*>(h)->bands.size() : 0; }
int fluid_multi_transport(const FluidMulti* h) { return h ? reinterpret_cast<const Multi*>(h)->transport : 0; }
int fluid_multi_row_bounds(const FluidMulti* h, uint32_t* out /* world + 1 */) {
    if (!h || !out) { set_error("null argument"); return FLUID_ERR_INVALID_ARGUMENT; }
    const Multi* m = reinterpret_cast<const Multi*>(h);
    for (int r = 0; r <= m->world; r++) out[r] = m->bounds[r];
    return FLUID_OK;
}
FluidSim* fluid_multi_band_sim(FluidMulti* h, int local_band) {
    if (!h) return nullptr;
    Multi* m = reinterpret_cast<Multi*>(h);
    if (local_band < 0 || local_band >= (int)m->bands.size()) return nullptr;
    return reinterpret_cast<FluidSim*>(m->bands[local_band]->sim);
}

int fluid_multi_set_math_mode(FluidMulti* h, int mode) {
    MULTI_OR_FAIL(h);
    for (Band* b : m->bands) { int rc = fluid_sim_set_math_mode(reinterpret_cast<FluidSim*>(b->sim), mode); if (rc != FLUID_OK) return rc; }
    return FLUID_OK;
}

// synchronises; counts[r] = particles rank r owns (all ranks in either mode: every rank tracks the whole vector)
int fluid_multi_particle_counts(FluidMulti* h, uint32_t* counts /* world */) {
    MULTI_OR_FAIL(h);
    if (!counts) { set_error("null argument"); return FLUID_ERR_INVALID_ARGUMENT; }
    int rc = fluid_multi_sync(h);
    if (rc != FLUID_OK) return rc;
    Band* b = m->bands[0];
    FLUID_CUDA(cudaSetDevice(b->sim->device));
    BandDyn d;
    FLUID_CUDA(cudaMemcpy(&d, b->dyn, sizeof d, cudaMemcpyDeviceToHost));
    for (int r = 0; r < m->world; r++) counts[r] = d.n_all[r];
    return FLUID_OK;
}

// one band, in key order, into dst (capacity cap records); *n_out, *goff_out: its size and global sorted offset
static int band_download(Band* b, FluidParticle* dst, uint32_t cap, uint32_t* n_out, uint32_t* goff_out) {
    Sim* s = b->sim;
    FLUID_CUDA(cudaSetDevice(s->device));
    FLUID_CUDA(cudaStreamSynchronize(s->stream));
    BandDyn d;
    FLUID_CUDA(cudaMemcpy(&d, b->dyn, sizeof d, cudaMemcpyDeviceToHost));
    const uint32_t n = d.own_end - d.own_begin;
    if (n_out) *n_out = n;
    if (goff_out) *goff_out = d.goff;
    if (n > cap) { set_error("fluid_multi download: destination too small for the band"); return FLUID_ERR_INVALID_ARGUMENT; }
    s->n = n;
    s->sorted_off = d.own_begin;
    s->keys_sorted = s->keys_b;
    if (n == 0) return FLUID_OK;
    return fluid_sim_download_particles(reinterpret_cast<FluidSim*>(s), dst);
}

int fluid_multi_download_particles(FluidMulti* h, FluidParticle* dst) {
    MULTI_OR_FAIL(h);
    if (!dst) { set_error("fluid_multi_download_particles: null destination"); return FLUID_ERR_INVALID_ARGUMENT; }
    if (!m->single_process) { set_error("fluid_multi_download_particles: one-rank-per-process handles download their band (fluid_multi_download_band)"); return FLUID_ERR_INVALID_ARGUMENT; }
    int rc = fluid_multi_sync(h);
    if (rc != FLUID_OK) return rc;
    uint64_t at = 0;
    for (Band* b : m->bands) {
        uint32_t n = 0, goff = 0;
        // sizes first: the bands are concatenated in rank order = global key order
        FLUID_CUDA(cudaSetDevice(b->sim->device));
        BandDyn d;
        FLUID_CUDA(cudaMemcpy(&d, b->dyn, sizeof d, cudaMemcpyDeviceToHost));
        if (at + (d.own_end - d.own_begin) > m->n_total) { set_error("fluid_multi_download_particles: bands hold more particles than the scene"); return FLUID_ERR_OVERFLOW; }
        rc = band_download(b, dst + at, m->n_total - (uint32_t)at, &n, &goff);
        if (rc != FLUID_OK) return rc;
        at += n;
    }
    if (at != m->n_total) { set_error("fluid_multi_download_particles: particles were lost or duplicated"); return FLUID_ERR_OVERFLOW; }
    return FLUID_OK;
}
int fluid_multi_download_band(FluidMulti* h, int local_band, FluidParticle* dst, uint32_t cap, uint32_t* n_out, uint32_t* global_offset_out) {
    MULTI_OR_FAIL(h);
    if (local_band < 0 || local_band >= (int)m->bands.size()) { set_error("fluid_multi_download_band: bad band index"); return FLUID_ERR_INVALID_ARGUMENT; }
    if (!dst && cap) { set_error("fluid_multi_download_band: null destination"); return FLUID_ERR_INVALID_ARGUMENT; }
    int rc = multi_check_errors(m);
    if (rc != FLUID_OK) return rc;
    return band_download(m->bands[local_band], dst, cap, n_out, global_offset_out);
}

// start_indices (reference semantics, GLOBAL sorted indices): a rank maintains the rows it owns and its ghost rows
int fluid_multi_download_start_indices(FluidMulti* h, uint32_t* dst) {
    MULTI_OR_FAIL(h);
    if (!dst) { set_error("null destination"); return FLUID_ERR_INVALID_ARGUMENT; }
    int rc = fluid_multi_sync(h);
    if (rc != FLUID_OK) return rc;
    const size_t cells = (size_t)m->grid_w * m->grid_h;
    if (!m->single_process) return fluid_sim_download_start_indices(reinterpret_cast<FluidSim*>(m->bands[0]->sim), dst);
    std::vector<uint32_t> tmp(cells);
    for (Band* b : m->bands) {
        rc = fluid_sim_download_start_indices(reinterpret_cast<FluidSim*>(b->sim), tmp.data());
        if (rc != FLUID_OK) return rc;
        const size_t lo = (size_t)m->bounds[b->rows.rank] * m->grid_w, hi = (size_t)m->bounds[b->rows.rank + 1] * m->grid_w;
        memcpy(dst + lo, tmp.data() + lo, (hi - lo) * 4);
    }
    return FLUID_OK;
}
int fluid_multi_upload_start_indices(FluidMulti* h, const uint32_t* src) {
    MULTI_OR_FAIL(h);
    if (!src) { set_error("null source"); return FLUID_ERR_INVALID_ARGUMENT; }
    for (Band* b : m->bands) { int rc = fluid_sim_upload_start_indices(reinterpret_cast<FluidSim*>(b->sim), src); if (rc != FLUID_OK) return rc; }
    return FLUID_OK;
}
int fluid_multi_upload_force_field(FluidMulti* h, const float* xy) {
    MULTI_OR_FAIL(h);
    if (!xy) { set_error("null source"); return FLUID_ERR_INVALID_ARGUMENT; }
    for (Band* b : m->bands) { int rc = fluid_sim_upload_force_field(reinterpret_cast<FluidSim*>(b->sim), xy); if (rc != FLUID_OK) return rc; }
    return FLUID_OK;
}

int fluid_multi_enable_timing(FluidMulti* h, int enable) {
    MULTI_OR_FAIL(h);
    for (Band* b : m->bands) { int rc = fluid_sim_enable_timing(reinterpret_cast<FluidSim*>(b->sim), enable); if (rc != FLUID_OK) return rc; }
    return FLUID_OK;
}
int fluid_multi_last_timing(FluidMulti* h, int local_band, const char** names, float* ms, int cap) {
    MULTI_OR_FAIL(h);
    if (local_band < 0 || local_band >= (int)m->bands.size()) { set_error("bad band index"); return FLUID_ERR_INVALID_ARGUMENT; }
    return fluid_sim_last_timing(reinterpret_cast<FluidSim*>(m->bands[local_band]->sim), names, ms, cap);
}
uint64_t fluid_multi_launch_count(const FluidMulti* h) {
    if (!h) return 0;
    uint64_t c = 0;
    for (Band* b : reinterpret_cast<const Multi*>(h)->bands) c += b->sim->launches;
    return c;
}
void* fluid_multi_stream(FluidMulti* h, int local_band) {
    if (!h) return nullptr;
    Multi* m = reinterpret_cast<Multi*>(h);
    if (local_band < 0 || local_band >= (int)m->bands.size()) return nullptr;
    return (void*)m->bands[local_band]->sim->stream;
}

// ---- pipelined host transfers of a band: SURVEY.md §8(b) state exposure, full duplex (as fluid_sim_*_particles_async) ----
// One-rank-per-process handles (or world == 1). Nothing is read back by the host: the band sizes are all-gathered on
// the device. upload: COLLECTIVE; the ranks' subsets concatenated in rank order are the global array (origins implicit).
int fluid_multi_upload_band_async(FluidMulti* h, const FluidParticle* src, uint32_t n) {
    MULTI_OR_FAIL(h);
    if (m->single_process && m->world > 1) { set_error("fluid_multi_upload_band_async: one-rank-per-process handles only"); return FLUID_ERR_INVALID_ARGUMENT; }
    if (!src && n) { set_error("fluid_multi_upload_band_async: null source"); return FLUID_ERR_INVALID_ARGUMENT; }
    Band* b = m->bands[0];
    Sim* s = b->sim;
    FLUID_CUDA(cudaSetDevice(s->device));
    if (n > b->capacity) { set_error("fluid_multi_upload_band_async: more particles than the band's capacity"); return FLUID_ERR_INVALID_ARGUMENT; }
    int rc = ensure_pipeline(s);
    if (rc != FLUID_OK) return rc;
    if (s->in_used) FLUID_CUDA(cudaStreamWaitEvent(s->h2d_stream, s->ev_in_free, 0));
    if (n) FLUID_CUDA(cudaMemcpyAsync(s->aos_in, src, (size_t)n * sizeof(FluidParticle), cudaMemcpyHostToDevice, s->h2d_stream));
    FLUID_CUDA(cudaEventRecord(s->ev_h2d_done, s->h2d_stream));
    FLUID_CUDA(cudaStreamWaitEvent(s->stream, s->ev_h2d_done, 0));
    s->n = n;
    s->sorted_off = 0;
    s->keys_sorted = s->keys_b;
    if (n) {
        k_aos_to_soa<<<(n + 255) / 256, 256, 0, s->stream>>>(s->aos_in, n, s->pos_a, s->pred_s, s->vel_a, s->rho, s->keys_b);
        s->launches++;
    }
    FLUID_CUDA(cudaEventRecord(s->ev_in_free, s->stream));
    s->in_used = true;
    uint32_t* gathered = b->scratch_word + 4;
    k_band_set_word<<<1, 1, 0, s->stream>>>(m->world > 1 ? b->scratch_word : gathered + b->rows.rank, n);
    if (m->world > 1) FLUID_NCCL(g_nccl.AllGather(b->scratch_word, gathered, 1, ncclUint32, b->comm, s->stream));
    k_band_counts_from_gathered<<<1, 32, 0, s->stream>>>(b->dyn, gathered, b->rows.rank, m->world, m->n_total);
    s->launches += 2;
    b->explicit_origins = false;
    if (b->spec_valid) b->count_dirty = true;
    b->spec_valid = false;
    FLUID_CUDA(cudaGetLastError());
    return FLUID_OK;
}
// download: the band's first `cap` particles in key order into dst, {size, global offset} into info[2]; dst and info
// should be pinned; both are valid after fluid_multi_sync_transfers (size > cap: the tail was not copied).
int fluid_multi_download_band_async(FluidMulti* h, FluidParticle* dst, uint32_t cap, uint32_t* info) {
    MULTI_OR_FAIL(h);
    if (m->single_process && m->world > 1) { set_error("fluid_multi_download_band_async: one-rank-per-process handles only"); return FLUID_ERR_INVALID_ARGUMENT; }
    if (!dst || !info) { set_error("fluid_multi_download_band_async: null destination"); return FLUID_ERR_INVALID_ARGUMENT; }
    Band* b = m->bands[0];
    Sim* s = b->sim;
    FLUID_CUDA(cudaSetDevice(s->device));
    cap = std::min(cap, b->capacity);
    int rc = ensure_pipeline(s);
    if (rc != FLUID_OK) return rc;
    if (s->out_used) FLUID_CUDA(cudaStreamWaitEvent(s->stream, s->ev_d2h_done, 0));
    uint32_t* info_dev = b->scratch_word + 12;
    k_band_soa_to_aos<<<(std::max(cap, 1u) + 255) / 256, 256, 0, s->stream>>>(s->aos_out, cap, b->dyn, s->pos_a, s->pred_s, s->vel_a, s->rho, s->keys_b, info_dev);
    s->launches++;
    FLUID_CUDA(cudaEventRecord(s->ev_out_ready, s->stream));
    FLUID_CUDA(cudaStreamWaitEvent(s->d2h_stream, s->ev_out_ready, 0));
    if (cap) FLUID_CUDA(cudaMemcpyAsync(dst, s->aos_out, (size_t)cap * sizeof(FluidParticle), cudaMemcpyDeviceToHost, s->d2h_stream));
    FLUID_CUDA(cudaMemcpyAsync(info, info_dev, 8, cudaMemcpyDeviceToHost, s->d2h_stream));
    FLUID_CUDA(cudaEventRecord(s->ev_d2h_done, s->d2h_stream));
    s->out_used = true;
    FLUID_CUDA(cudaGetLastError());
    return FLUID_OK;
}
int fluid_multi_sync_transfers(FluidMulti* h) {
    MULTI_OR_FAIL(h);
    for (Band* b : m->bands) {
        int rc = fluid_sim_sync_transfers(reinterpret_cast<FluidSim*>(b->sim));
        if (rc != FLUID_OK) return rc;
    }
    return multi_check_errors(m);
}

// row bounds with (nearly) equal particle counts per band, from the particles' positions (host arithmetic): for scenes
// whose rows are unevenly filled (dam break). out: world + 1 entries for FluidMultiOptions.row_bounds.
int fluid_multi_balanced_bounds(const FluidSimSettings* s, const FluidParticle* particles, uint32_t n, int world, uint32_t* out) {
    if (!s || (!particles && n) || !out || world < 1 || world > BAND_MAX_RANKS) { set_error("fluid_multi_balanced_bounds: bad argument"); return FLUID_ERR_INVALID_ARGUMENT; }
    Multi tmp;
    tmp.settings = *s;
    tmp.grid_h = (uint32_t)(size_t)ceilf(s->size[1] / s->smoothing_radius) + 2u;
    std::vector<uint64_t> per_row(tmp.grid_h, 0);
    for (uint32_t i = 0; i < n; i++) per_row[host_row_of(&tmp, particles[i].position[1])]++;
    out[0] = 0;
    uint64_t acc = 0;
    int r = 1;
    for (uint32_t row = 0; row < tmp.grid_h && r < world; row++) {
        acc += per_row[row];
        while (r < world && acc * world >= (uint64_t)n * r) { out[r] = std::max(row + 1u, out[r - 1]); r++; }
    }
    for (; r < world; r++) out[r] = tmp.grid_h;
    out[world] = tmp.grid_h;
    for (int k = 1; k < world; k++) out[k] = std::min(out[k], tmp.grid_h);
    return FLUID_OK;
}

}  // extern "C"
