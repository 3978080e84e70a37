// slab.inl — C ABI of slab (multi-GPU) mode: one FluidSim handle per rank owns a band of grid rows.
// The exchanges themselves (NCCL over NVLink) are driven from the host language through
// torch.distributed on views of this handle's device arrays (gpu_fluid_simulation_b200/slab.py);
// this file holds the device passes in between. See kernels_slab.cuh for the per-tick sequence.
// (textually included at the end of fluid_sim.cu: one translation unit, so the kernels are defined once)
#include "kernels_slab.cuh"

#define SLAB_OR_FAIL(h)                                                                   \
    if (!(h)) { set_error("null FluidSim handle"); return FLUID_ERR_INVALID_ARGUMENT; }   \
    Sim* s = reinterpret_cast<Sim*>(h);                                                   \
    FLUID_CUDA(cudaSetDevice(s->device));                                                 \
    if (!s->slab) { set_error("handle is not in slab mode (call fluid_slab_configure)"); return FLUID_ERR_INVALID_ARGUMENT; }

template <int PASSES>
static void launch_slab_hash(Sim* s, const DevParams& p, uint32_t begin, uint32_t count, int detect, uint32_t* frame_time) {
    if (count == 0 && !frame_time) return;
    uint32_t blocks = (count + HASH_THREADS - 1) / HASH_THREADS;
    if (blocks == 0) blocks = 1;
    const uint32_t max_blocks = (uint32_t)s->num_sms * 8u;
    if (blocks > max_blocks) blocks = max_blocks;
    k_slab_key_hist<PASSES><<<blocks, HASH_THREADS, 0, s->stream>>>(
        p, s->pos_a + begin, s->vel_a + begin, count, s->keys_u + begin, s->sort_scratch, s->slab_rows, detect,
        begin, s->emig_list, s->emig_count, s->emig_count + 8, s->emig_cap, s->num_cells, frame_time);
    s->launches++;
}
static void slab_hash(Sim* s, const DevParams& p, uint32_t begin, uint32_t count, int detect, uint32_t* frame_time) {
    switch (s->sort_passes) {
        case 1: launch_slab_hash<1>(s, p, begin, count, detect, frame_time); break;
        case 2: launch_slab_hash<2>(s, p, begin, count, detect, frame_time); break;
        case 3: launch_slab_hash<3>(s, p, begin, count, detect, frame_time); break;
        default: launch_slab_hash<4>(s, p, begin, count, detect, frame_time); break;
    }
}

extern "C" {

int fluid_slab_configure(FluidSim* h, int rank, int world, const uint32_t* row_bounds, uint32_t capacity, uint32_t pad,
                         uint32_t emig_cap) {
    if (!h || !row_bounds) { set_error("fluid_slab_configure: null argument"); return FLUID_ERR_INVALID_ARGUMENT; }
    Sim* s = reinterpret_cast<Sim*>(h);
    FLUID_CUDA(cudaSetDevice(s->device));
    if (world < 1 || world > SLAB_MAX_RANKS || rank < 0 || rank >= world) { set_error("fluid_slab_configure: bad rank/world"); return FLUID_ERR_INVALID_ARGUMENT; }
    for (int r = 0; r < world; r++)
        if (row_bounds[r] > row_bounds[r + 1]) { set_error("fluid_slab_configure: row bounds must be non-decreasing"); return FLUID_ERR_INVALID_ARGUMENT; }
    pad = (pad + 63u) & ~63u;
    if (capacity < 2 || emig_cap == 0) { set_error("fluid_slab_configure: bad capacity"); return FLUID_ERR_INVALID_ARGUMENT; }
    FLUID_CUDA(cudaStreamSynchronize(s->stream));
    const bool dbg = s->debug_force;
    int rc = s->alloc(capacity + 2u * pad);  // contents dropped: the caller uploads its band afterwards
    if (rc != FLUID_OK) return rc;
    if (s->keys_ext) cudaFree(s->keys_ext);
    if (s->emig_list) cudaFree(s->emig_list);
    if (s->slab_meta) cudaFree(s->slab_meta);
    FLUID_CUDA(cudaMalloc(&s->keys_ext, ((size_t)capacity + 2u * pad + 64) * 4));
    FLUID_CUDA(cudaMemsetAsync(s->keys_ext, 0, ((size_t)capacity + 2u * pad + 64) * 4, s->stream));
    FLUID_CUDA(cudaMalloc(&s->emig_list, (size_t)emig_cap * 8));
    FLUID_CUDA(cudaMalloc(&s->slab_meta, 256));
    FLUID_CUDA(cudaMemsetAsync(s->slab_meta, 0, 256, s->stream));
    {
        // counting sort of the band (kernels_count.cuh): bins = the band's cells + one for the sentinel key of emigrants
        const uint32_t band_rows = row_bounds[rank + 1] - row_bounds[rank];
        s->slab_bins = band_rows * s->grid_w + 1u;
        s->slab_bin_tiles = (uint32_t)(((size_t)s->slab_bins + CELL_SCAN_TILE - 1) / CELL_SCAN_TILE);  // <= cell_tiles: the arrays of the whole grid hold it
        if (s->slab_bin_lb) cudaFree(s->slab_bin_lb);
        FLUID_CUDA(cudaMalloc(&s->slab_bin_lb, (size_t)s->slab_bin_tiles * CELL_SCAN_TILE * 4));
        const char* radix = getenv("FLUID_SLAB_RADIX");
        s->slab_counting = !(radix && radix[0] == '1');
        s->cell_max_known = false;
    }
    if (!s->slab_dyn) FLUID_CUDA(cudaMalloc(&s->slab_dyn, 64));
    FLUID_CUDA(cudaMemsetAsync(s->slab_dyn, 0, 64, s->stream));
    s->emig_count = s->slab_meta + 8;  // meta words 0..4, emigrant counter at word 8, per-destination counters at words 16..31
    s->emig_cap = emig_cap;
    s->slab = true;
    s->slab_pad = pad;
    s->slab_capacity = capacity;
    memset(&s->slab_rows, 0, sizeof s->slab_rows);
    for (int r = 0; r <= world; r++) s->slab_rows.bounds[r] = row_bounds[r];
    s->slab_rows.world = world;
    s->slab_rows.rank = rank;
    s->state_off = pad;
    s->sorted_off = pad;
    s->n = 0;
    // the sentinel key num_cells must be sortable: one more key value than a single GPU needs
    uint32_t bits = 0;
    while (bits < 32 && (s->num_cells >> bits) != 0) bits++;
    s->sort_passes = (bits + RADIX_BITS - 1) / RADIX_BITS;
    if (s->sort_passes < 1) s->sort_passes = 1;
    if (dbg) {
        FLUID_CUDA(cudaMalloc(&s->dbg_accel, ((size_t)capacity + 64) * 8));
        FLUID_CUDA(cudaMalloc(&s->dbg_abs, ((size_t)capacity + 64) * 8));
        s->debug_force = true;
    }
    FLUID_CUDA(cudaStreamSynchronize(s->stream));
    return FLUID_OK;
}

int fluid_slab_view(FluidSim* h, FluidSlabView* out) {
    SLAB_OR_FAIL(h);
    if (!out) { set_error("null argument"); return FLUID_ERR_INVALID_ARGUMENT; }
    out->n = s->n; out->pad = s->slab_pad; out->capacity = s->slab_capacity; out->emig_cap = s->emig_cap;
    out->pos_state = s->pos_a; out->vel_state = s->vel_a;
    out->pos_sorted = s->pos_s; out->pred_sorted = s->pred_s; out->vel_sorted = s->vel_s; out->rho = s->rho;
    out->keys_ext = s->keys_ext; out->emig_list = s->emig_list; out->emig_count = s->emig_count; out->meta = s->slab_meta;
    out->start_indices = s->start_indices;
    out->dyn = s->slab_dyn;
    return FLUID_OK;
}

// phase 1: tick += 1, uniform, clear sort scratch + emigrant list, keys of the natives with emigrant detection
int fluid_slab_begin_tick(FluidSim* h, const FluidTickSettings* t) {
    SLAB_OR_FAIL(h);
    if (!t) { set_error("null settings"); return FLUID_ERR_INVALID_ARGUMENT; }
    s->tick += 1;
    s->make_uniform(*t, s->tick, &s->last_uniform);
    s->slab_params = s->make_params(s->last_uniform);
    s->n_marks = 0;
    s->mark("begin");
    const uint32_t max_tiles = (s->slab_capacity + 2u * s->slab_pad + SORT_TILE - 1) / SORT_TILE;
    FLUID_CUDA(cudaMemsetAsync(s->sort_scratch, 0, ((size_t)MAX_PASSES * RADIX + RADIX + (size_t)s->sort_passes * max_tiles * RADIX) * 4, s->stream));
    FLUID_CUDA(cudaMemsetAsync(s->emig_count, 0, (8 + SLAB_MAX_RANKS) * 4, s->stream));
    slab_hash(s, s->slab_params, s->state_off, s->n, 1, s->d_frame_time);
    s->mark("predict+key+hist");
    FLUID_CUDA(cudaGetLastError());
    return FLUID_OK;
}

// phase 2a: sort the emigrant list (destination, then source order) and gather the payload
int fluid_slab_pack(FluidSim* h, uint32_t n_emigrants, float* payload_dev) {
    SLAB_OR_FAIL(h);
    if (n_emigrants > s->emig_cap) { set_error("fluid_slab_pack: more emigrants than emig_cap"); return FLUID_ERR_INVALID_ARGUMENT; }
    if (n_emigrants == 0) return FLUID_OK;
    if (!payload_dev) { set_error("fluid_slab_pack: null payload"); return FLUID_ERR_INVALID_ARGUMENT; }
    if (n_emigrants <= (uint32_t)EMIG_SORT_MAX) {
        k_emig_sort_small<<<1, EMIG_SORT_THREADS, 0, s->stream>>>(s->emig_list, n_emigrants);
        s->launches++;
    } else {
        uint32_t cap = 2;
        while (cap < n_emigrants) cap <<= 1;
        if (cap > s->emig_cap) { set_error("fluid_slab_pack: emig_cap must cover the next power of two of the emigrant count"); return FLUID_ERR_INVALID_ARGUMENT; }
        FLUID_CUDA(cudaMemsetAsync(s->emig_list + n_emigrants, 0xff, (size_t)(cap - n_emigrants) * 8, s->stream));
        for (uint32_t k = 2; k <= cap; k <<= 1)
            for (uint32_t j = k >> 1; j > 0; j >>= 1) {
                k_emig_sort_step<<<(cap + 255) / 256, 256, 0, s->stream>>>(s->emig_list, cap, k, j);
                s->launches++;
            }
    }
    k_emig_pack<<<(n_emigrants + 255) / 256, 256, 0, s->stream>>>(s->emig_list, n_emigrants, s->pos_a, s->vel_a,
                                                                   reinterpret_cast<float4*>(payload_dev));
    s->launches++;
    FLUID_CUDA(cudaGetLastError());
    return FLUID_OK;
}

// phase 2b: immigrants (ordered by source rank) into the slots around the natives
int fluid_slab_unpack(FluidSim* h, const float* recv_dev, uint32_t n_from_below, uint32_t n_from_above) {
    SLAB_OR_FAIL(h);
    if (n_from_below > s->slab_pad || n_from_above > s->slab_pad) { set_error("fluid_slab_unpack: more immigrants than the pad holds"); return FLUID_ERR_INVALID_ARGUMENT; }
    const uint32_t total = n_from_below + n_from_above;
    if (total == 0) return FLUID_OK;
    if (!recv_dev) { set_error("fluid_slab_unpack: null buffer"); return FLUID_ERR_INVALID_ARGUMENT; }
    k_immig_unpack<<<(total + 255) / 256, 256, 0, s->stream>>>(reinterpret_cast<const float4*>(recv_dev), n_from_below, n_from_above,
                                                               s->state_off - n_from_below, s->state_off + s->n, s->pos_a, s->vel_a);
    s->launches++;
    FLUID_CUDA(cudaGetLastError());
    return FLUID_OK;
}

// phases 3+4: keys of the immigrants, sort of [from below | natives | from above], gather into key order,
// band metadata. Afterwards the handle owns n_total - n_emigrants particles.
int fluid_slab_sort(FluidSim* h, uint32_t n_from_below, uint32_t n_from_above, uint32_t n_emigrants) {
    SLAB_OR_FAIL(h);
    if (n_from_below > s->slab_pad || n_from_above > s->slab_pad) { set_error("fluid_slab_sort: more immigrants than the pad holds"); return FLUID_ERR_INVALID_ARGUMENT; }
    const uint32_t begin = s->state_off - n_from_below;
    const uint32_t total = n_from_below + s->n + n_from_above;
    if (n_emigrants > s->n || total - n_emigrants > s->slab_capacity) { set_error("fluid_slab_sort: band overflows its capacity"); return FLUID_ERR_INVALID_ARGUMENT; }
    DevParams& p = s->slab_params;
    slab_hash(s, p, begin, n_from_below, 0, nullptr);
    slab_hash(s, p, s->state_off + s->n, n_from_above, 0, nullptr);
    // key order: counting sort over the band's cells (same stable order as the radix passes, kernels_count.cuh) unless
    // the cells are crowded. The statistics come from the previous tick's scan (the protocol's host read of the
    // migration counts has synchronised the stream since); the first tick after configure / a resize takes the radix
    // passes and only gathers the statistics.
    const uint32_t key_lo = s->slab_rows.bounds[s->slab_rows.rank] * s->grid_w;
    const uint32_t last_bin = s->slab_bins - 1u;
    bool counting = false;
    if (s->slab_counting && total > 0) {
        uint32_t* count = s->cell_count_buf[0];
        const size_t bin_bytes = (size_t)s->slab_bin_tiles * CELL_SCAN_TILE * 4;
        FLUID_CUDA(cudaMemsetAsync(count, 0, bin_bytes, s->stream));
        k_count_keys<<<(total + GATHER_THREADS - 1) / GATHER_THREADS, GATHER_THREADS, 0, s->stream>>>(s->keys_u + begin, total, key_lo, last_bin, count);
        if (s->cell_max_known) {
            const uint32_t max_ok = std::min(std::max(total / 1024u, COUNT_SORT_MAX_CELL_LO), COUNT_SORT_MAX_CELL_HI);
            const uint32_t sq_bits = s->cell_max_host[1];
            float sum_sq;
            memcpy(&sum_sq, &sq_bits, 4);
            counting = s->cell_max_host[0] <= max_ok && sum_sq <= COUNT_SORT_MAX_MEAN_SQ * (float)total;
        }
        k_cell_reduce<<<s->slab_bin_tiles, CELL_SCAN_THREADS, 0, s->stream>>>(count, s->cell_tile_sum, s->cell_tile_max, s->cell_tile_sq, s->d_frame_time, 0u);
        k_cell_downsweep<<<s->slab_bin_tiles, CELL_SCAN_THREADS, 0, s->stream>>>(count, s->slab_bin_lb, s->cell_tile_sum, s->cell_tile_max,
                                                                               s->cell_tile_sq, s->slab_bin_tiles, s->cell_max_dev,
                                                                               s->cell_count_buf[1]);
        s->launches += 3;
        s->cell_max_known = true;
    }
    s->n = total - n_emigrants;  // sentinel keys sort to the tail and are dropped here
    p.n = s->n;
    p.own_begin = s->sorted_off;
    if (counting) {
        k_cell_slots<<<(total + GATHER_THREADS * SLOT_ITEMS - 1) / (GATHER_THREADS * SLOT_ITEMS), GATHER_THREADS, 0, s->stream>>>(
            total, key_lo, last_bin, s->keys_u + begin, s->cell_count_buf[0], s->keys_b, s->idx_a);
        s->launches++;
        s->keys_sorted = s->keys_b;
        s->idx_sorted = nullptr;
        s->mark("sort");
        k_slab_rank_gather<<<(total + GATHER_THREADS - 1) / GATHER_THREADS, GATHER_THREADS, 0, s->stream>>>(
            p, total, key_lo, last_bin, s->keys_b, s->idx_a, s->slab_bin_lb, s->pos_a + begin, s->vel_a + begin, s->pos_s, s->pred_s,
            s->vel_s, s->keys_ext, s->quirk + 2, s->d_frame_time);
        s->launches++;
        s->last_sort_path = FLUID_SORT_COUNTING;
    } else {
        const uint32_t passes = s->sort_passes;
        uint32_t* hist = s->sort_scratch;
        uint32_t* counters = s->sort_scratch + MAX_PASSES * RADIX;
        uint32_t* status = counters + RADIX;
        k_scan_histograms<<<passes, RADIX, 0, s->stream>>>(hist);
        s->launches++;
        const uint32_t tiles = (total + SORT_TILE - 1) / SORT_TILE;
        const size_t sort_smem = (size_t)SORT_TILE * 2 * sizeof(uint32_t);
        const uint32_t* kin = s->keys_u + begin;
        const uint32_t* vin = nullptr;
        uint32_t* kout = s->keys_b;
        uint32_t* vout = s->idx_a;
        for (uint32_t q = 0; q < passes && tiles > 0; q++) {
            if (q == 0)
                k_onesweep_pass<SORT_ITEMS, true><<<tiles, SORT_THREADS, sort_smem, s->stream>>>(
                    kin, vin, kout, vout, hist + q * RADIX, counters + q, status + (size_t)q * tiles * RADIX, total, (int)(q * RADIX_BITS));
            else
                k_onesweep_pass<SORT_ITEMS, false><<<tiles, SORT_THREADS, sort_smem, s->stream>>>(
                    kin, vin, kout, vout, hist + q * RADIX, counters + q, status + (size_t)q * tiles * RADIX, total, (int)(q * RADIX_BITS));
            s->launches++;
            kin = kout; vin = vout;
            kout = (kout == s->keys_b) ? s->keys_c : s->keys_b;
            vout = (vout == s->idx_a) ? s->idx_b : s->idx_a;
        }
        s->keys_sorted = const_cast<uint32_t*>(kin);
        s->idx_sorted = const_cast<uint32_t*>(vin);
        s->mark("sort");
        if (s->n) {
            k_slab_gather<<<(s->n + GATHER_THREADS - 1) / GATHER_THREADS, GATHER_THREADS, 0, s->stream>>>(
                p, s->keys_sorted, s->idx_sorted, s->pos_a + begin, s->vel_a + begin, s->pos_s, s->pred_s, s->vel_s, s->keys_ext,
                s->quirk + 2, s->d_frame_time);
            s->launches++;
        }
        s->last_sort_path = FLUID_SORT_RADIX;
    }
    const uint32_t row_first = s->slab_rows.bounds[s->slab_rows.rank], row_end = s->slab_rows.bounds[s->slab_rows.rank + 1];
    k_slab_meta<<<1, 32, 0, s->stream>>>(s->keys_ext + s->sorted_off, s->n, s->grid_w, row_first, row_end > row_first ? row_end - 1 : row_first,
                                         s->start_indices, s->num_cells, s->slab_meta);
    s->launches++;
    s->mark("gather");
    FLUID_CUDA(cudaGetLastError());
    return FLUID_OK;
}

// phases 6+7a: halo keys, cell tables over [halo | owned | halo], density of the owned particles
int fluid_slab_cells_density(FluidSim* h, uint32_t halo_below, uint32_t halo_above, uint32_t global_offset, uint32_t quirk_c0,
                             uint32_t quirk_stale_ext) {
    SLAB_OR_FAIL(h);
    if (halo_below > s->slab_pad || halo_above > s->slab_pad) { set_error("fluid_slab_cells_density: halo larger than the pad"); return FLUID_ERR_INVALID_ARGUMENT; }
    DevParams& p = s->slab_params;
    p.ext_begin = s->sorted_off - halo_below;
    p.ext_end = s->sorted_off + s->n + halo_above;
    p.global_offset = global_offset;
    s->slab_goff_dev = nullptr;
    const uint32_t halos = halo_below + halo_above;
    if (halos) {
        k_slab_halo_keys<<<(halos + 255) / 256, 256, 0, s->stream>>>(p, s->pred_s, s->keys_ext, p.ext_begin, halo_below,
                                                                     s->sorted_off + s->n, halo_above, s->vel_s, s->quirk + 2, s->d_frame_time);
        s->launches++;
    }
    const uint32_t ext = p.ext_end - p.ext_begin;
    // table entries the neighbour passes of this band can read: rows row_first-1 .. row_end (+ the entry that ends row_end)
    const uint32_t brow0 = s->slab_rows.bounds[s->slab_rows.rank], brow1 = s->slab_rows.bounds[s->slab_rows.rank + 1];
    const uint32_t fill_lo = (brow0 > 0 ? brow0 - 1u : 0u) * s->grid_w;
    const uint32_t fill_hi = (uint32_t)std::min<uint64_t>((uint64_t)(brow1 + 1u) * s->grid_w + 1u, (uint64_t)s->num_cells + 1u);
    if (ext) {
        k_slab_build_cells<<<(ext + GATHER_THREADS - 1) / GATHER_THREADS, GATHER_THREADS, 0, s->stream>>>(
            p, s->keys_ext, s->start_indices, s->cell_lb, s->quirk, quirk_c0, quirk_stale_ext, fill_lo, fill_hi, nullptr, 0u);
    } else if (fill_hi > fill_lo) {
        k_slab_fill_table<<<(fill_hi - fill_lo + 255) / 256, 256, 0, s->stream>>>(s->cell_lb, fill_lo, fill_hi - fill_lo, p.ext_begin);
    }
    s->launches++;
    s->mark("cells");
    if (s->n) {
        launch_tile_plans(s, p, s->keys_ext);
        k_density<<<(s->n + SPH_THREADS - 1) / SPH_THREADS, SPH_THREADS, 0, s->stream>>>(p, s->pred_s, s->keys_ext, s->cell_lb, s->quirk, s->rho, s->quirk + 2, s->d_frame_time,
                                                                                         reinterpret_cast<const TilePlan*>(s->tile_plans));
        s->launches++;
    }
    s->mark("density");
    FLUID_CUDA(cudaGetLastError());
    return FLUID_OK;
}

// phases 6+7a with the band metadata of every rank still on the device (the all_gather output): no host read between
// the sort and the force pass. allmeta_dev: world x 8 words (n, first-row count, last-row count, first key, its table
// entry, 3 unused). Halo rows were received as fixed pad-sized windows: below right-aligned at the owned range,
// above starting right after it. Errors (a halo row larger than the pad) are latched in the handle and returned by
// fluid_slab_error_flags on the next host read.
int fluid_slab_cells_density_dev(FluidSim* h, const uint32_t* allmeta_dev) {
    SLAB_OR_FAIL(h);
    if (!allmeta_dev) { set_error("fluid_slab_cells_density_dev: null metadata"); return FLUID_ERR_INVALID_ARGUMENT; }
    DevParams& p = s->slab_params;
    p.ext_begin = 0; p.ext_end = 0; p.global_offset = 0;  // resolved on the device
    SlabDyn* dyn = reinterpret_cast<SlabDyn*>(s->slab_dyn);
    s->slab_goff_dev = &dyn->global_offset;
    const uint32_t pad = s->slab_pad;
    k_slab_resolve<<<1, 32, 0, s->stream>>>(allmeta_dev, s->slab_rows.world, s->slab_rows.rank, s->sorted_off, s->n, pad, dyn, s->quirk);
    k_slab_halo_keys_dev<<<(2u * pad + 255) / 256, 256, 0, s->stream>>>(p, s->pred_s, s->keys_ext, dyn, s->sorted_off, s->n, pad, s->vel_s, s->quirk + 2,
                                                                  s->d_frame_time);
    const uint32_t brow0 = s->slab_rows.bounds[s->slab_rows.rank], brow1 = s->slab_rows.bounds[s->slab_rows.rank + 1];
    const uint32_t fill_lo = (brow0 > 0 ? brow0 - 1u : 0u) * s->grid_w;
    const uint32_t fill_hi = (uint32_t)std::min<uint64_t>((uint64_t)(brow1 + 1u) * s->grid_w + 1u, (uint64_t)s->num_cells + 1u);
    if (fill_hi > fill_lo)
        k_slab_fill_if_empty<<<(fill_hi - fill_lo + 255) / 256, 256, 0, s->stream>>>(s->cell_lb, fill_lo, fill_hi - fill_lo, dyn);
    const uint32_t span = s->n + 2u * pad;
    k_slab_build_cells<<<(span + GATHER_THREADS - 1) / GATHER_THREADS, GATHER_THREADS, 0, s->stream>>>(
        p, s->keys_ext, s->start_indices, s->cell_lb, s->quirk, 0u, 0u, fill_lo, fill_hi, dyn, s->sorted_off - pad);
    s->launches += 4;
    s->mark("cells");
    if (s->n) {
        launch_tile_plans(s, p, s->keys_ext);
        k_density<<<(s->n + SPH_THREADS - 1) / SPH_THREADS, SPH_THREADS, 0, s->stream>>>(p, s->pred_s, s->keys_ext, s->cell_lb, s->quirk, s->rho, s->quirk + 2, s->d_frame_time,
                                                                                         reinterpret_cast<const TilePlan*>(s->tile_plans));
        s->launches++;
    }
    s->mark("density");
    FLUID_CUDA(cudaGetLastError());
    return FLUID_OK;
}

// phase 7b: force + integrate over the owned particles; the new state lands at state_off in key order
int fluid_slab_force(FluidSim* h) {
    SLAB_OR_FAIL(h);
    if (s->n) {
        // densities of the halo rows arrived after the density pass: the last input of the tick's wild stamp
        const DevParams& p = s->slab_params;
        const bool dev = s->slab_goff_dev != nullptr;
        const uint32_t hb = dev ? s->slab_pad : s->sorted_off - p.ext_begin, ha = dev ? s->slab_pad : p.ext_end - (s->sorted_off + s->n);
        if (hb + ha) {
            k_slab_check_rho_halo<<<(hb + ha + 255) / 256, 256, 0, s->stream>>>(
                s->rho, p.ext_begin, hb, s->sorted_off + s->n, ha, dev ? reinterpret_cast<const uint32_t*>(s->slab_dyn) : nullptr,
                s->sorted_off, s->quirk + 2, s->d_frame_time);
            s->launches++;
        }
    }
    if (s->n) launch_force(s, s->slab_params, (s->n + SPH_THREADS - 1) / SPH_THREADS, s->pos_a + s->state_off, s->vel_a + s->state_off);
    s->mark("force+integrate");
    if (s->timing) { s->marks_per_tick = s->n_marks; s->timed_ticks++; }
    FLUID_CUDA(cudaGetLastError());
    return FLUID_OK;
}

}  // extern "C"
