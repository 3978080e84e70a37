// kernels_sph.cuh — the two neighbour passes: density (K5) and force + integrate (K6).
//
// Reference: calculate_density (compute.wgsl:59-74 -> funcs.wgsl:157-203) and move_particle
// (compute.wgsl:79-157 -> calculate_pressure_force :160-235, calculate_viscosity_force :238-299).
//
// B200 shape. Keys are row-major (funcs.wgsl:216-218) and the particle arrays are in key order, so
// the 3 cells (x-1, x, x+1) of one grid row are ONE contiguous span of the sorted arrays, and a
// CTA that owns 256 consecutive sorted particles needs, for each of the three grid rows around it,
// one contiguous span too. The CTA stages those spans (predicted position; plus velocity and density
// for the force pass) in shared memory with 1-D TMA bulk copies (cp.async.bulk -> mbarrier), then
// each thread walks its own three sub-spans out of shared memory in ascending index order — which is
// exactly the reference's visiting order (rows y-1, y, y+1; cells x-1, x, x+1; ascending index in a
// cell), so every fp32 sum is accumulated in the reference's order.
//   * density walks 3x3 cells, not the reference's 7x7: the outer rings only ever add +0.0
//     (cell size == h); tests compare against the 7x7 CPU restatement.
//   * pressure and viscosity are accumulated in ONE walk with separate accumulators: same order, same
//     bits as the reference's two walks, half the traffic.
//   * move_particle reads neighbour velocity from the sorted snapshot and writes the new state to a
//     different buffer: Jacobi semantics by construction (the reference has a data race here,
//     SURVEY.md §3.3).
// A CTA whose particles straddle grid rows and whose neighbourhood does not fit the staging buffer
// (a row boundary inside a long row) falls back to walking the same spans straight from global/L2.
#pragma once
#include "common.cuh"
#include "band_common.cuh"

namespace fluid {

#ifndef FLUID_SPH_THREADS
#define FLUID_SPH_THREADS 256
#endif
#if FLUID_SPH_THREADS == 256
#define FLUID_LIST_STRIDE 512
#define FLUID_LIST_STRIDE_STR "512"
#elif FLUID_SPH_THREADS == 128
#define FLUID_LIST_STRIDE 256
#define FLUID_LIST_STRIDE_STR "256"
#else
#error "FLUID_SPH_THREADS must be 128 or 256"
#endif
constexpr int SPH_THREADS = FLUID_SPH_THREADS;  // sorted particles per CTA of the neighbour passes (= tile of the staging plans)
#ifndef FLUID_SPH_CAP
#define FLUID_SPH_CAP 1920  // the most that fits 48 KB of static shared memory in the force pass (20 B per candidate + the 10 KB list); 4 CTAs per SM either way
#endif
constexpr int SPH_CAP = FLUID_SPH_CAP;  // staged particles per CTA (3 row spans of ~256+2 cells, rounded)

constexpr int SPH_MAX_ROWS = 2;  // CTA row groups that get their own three staged spans

struct SpanPlan {
    // base[g*3 + r]: for a thread whose cell row is row_first + g, and neighbour row r (0..2 = y-1, y, y+1):
    // global index of the first staged particle of that span minus its offset in the staging buffer,
    // so staging index = global index - base. In the single-interval case all entries are equal.
    int base[SPH_MAX_ROWS * 3];
    uint32_t row_first;
    int staged;  // 1: spans are in shared memory; 0: walk global memory
};

// Per-tile staging plan, computed once per tick by k_tile_plans (one thread per tile of SPH_THREADS sorted particles)
// and consumed by thread 0 of the density and force CTAs: which global index intervals the tile needs and whether
// they fit. Keeping the key / table look-ups out of the CTAs shortens their start-up chain from three dependent
// global-memory latencies (keys -> table -> TMA) to two (plan -> TMA).
//   * tile within <= SPH_MAX_ROWS grid rows: per row group g, the cells [x_lo-1, x_hi+1] of rows y-1, y, y+1,
//     where [x_lo, x_hi] bounds the tile's particles of that row (3 spans per group);
//   * more rows (short rows): one span from (first row - 1, x_first - 1) to (last row + 1, x_last + 1);
//   * if the spans exceed SPH_CAP particles the CTA walks global memory instead.
struct __align__(16) TilePlan {
    uint32_t lo[SPH_MAX_ROWS * 3];   // first particle of each span (multiple of 4: 16-byte granularity for every stream)
    uint32_t len[SPH_MAX_ROWS * 3];  // particles per span (multiple of 4)
    uint32_t row_first;
    uint32_t flags;                  // bit 0: staged, bit 1: grouped
    uint32_t pad[2];
};
static_assert(sizeof(TilePlan) == 64, "TilePlan is read as four 16-byte words");

// [i0, i1): the tile's sorted particles (i1 > i0); `plans + tile` receives the plan
__device__ __forceinline__ void make_tile_plan_range(const DevParams& p, const uint32_t* __restrict__ keys, const uint32_t* __restrict__ lb,
                                                     uint32_t tile, uint32_t i0, uint32_t i1, TilePlan* __restrict__ plans) {
    constexpr int NI = SPH_MAX_ROWS * 3;
    const uint32_t kf = min(__ldg(keys + i0), p.num_cells - 1u), kl = min(__ldg(keys + i1 - 1u), p.num_cells - 1u);
    const uint32_t yf = kf / p.grid_w, xa = kf - yf * p.grid_w;
    const uint32_t yl = kl / p.grid_w, xb = kl - yl * p.grid_w;
    const uint32_t nrows = yl - yf + 1u;
    uint32_t lo[NI], hi[NI];
#pragma unroll
    for (int t = 0; t < NI; t++) { lo[t] = 0; hi[t] = 0; }
    const bool grouped = nrows <= (uint32_t)SPH_MAX_ROWS;
    if (grouped) {
#pragma unroll
        for (int g = 0; g < SPH_MAX_ROWS; g++) {
            if ((uint32_t)g < nrows) {
                const uint32_t y = yf + (uint32_t)g;
                const uint32_t x_lo = (g == 0) ? xa : 0u;
                const uint32_t x_hi = ((uint32_t)g == nrows - 1u) ? xb : p.grid_w - 1u;
#pragma unroll
                for (int r = 0; r < 3; r++) {
                    const uint32_t row = y + (uint32_t)r - 1u;  // wraps for y == 0 -> fails the range test
                    if (row < p.grid_h) {
                        lo[g * 3 + r] = __ldg(lb + row * p.grid_w + (x_lo > 0u ? x_lo - 1u : 0u));
                        hi[g * 3 + r] = __ldg(lb + row * p.grid_w + min(x_hi + 2u, p.grid_w));
                    }
                }
            }
        }
    } else {
        const uint32_t ca = yf > 0u ? (yf - 1u) * p.grid_w + (xa > 0u ? xa - 1u : 0u) : 0u;
        const uint32_t cb = (yl + 1u < p.grid_h) ? (yl + 1u) * p.grid_w + min(xb + 2u, p.grid_w) : p.num_cells;
        lo[0] = __ldg(lb + ca);
        hi[0] = __ldg(lb + cb);
    }
    TilePlan out;
    uint32_t total = 0;
#pragma unroll
    for (int t = 0; t < NI; t++) {
        lo[t] &= ~3u;
        hi[t] = (hi[t] + 3u) & ~3u;
        if (hi[t] < lo[t]) hi[t] = lo[t];
        out.lo[t] = lo[t];
        out.len[t] = hi[t] - lo[t];
        total += hi[t] - lo[t];
    }
    out.row_first = yf;
    out.flags = (total <= (uint32_t)SPH_CAP ? 1u : 0u) | (grouped ? 2u : 0u);
    out.pad[0] = 0; out.pad[1] = 0;
    const uint4* src = reinterpret_cast<const uint4*>(&out);
    uint4* dst = reinterpret_cast<uint4*>(plans + tile);
#pragma unroll
    for (int q = 0; q < 4; q++) dst[q] = src[q];
}
__device__ __forceinline__ void make_tile_plan(const DevParams& p, const uint32_t* __restrict__ keys, const uint32_t* __restrict__ lb,
                                               uint32_t tile, TilePlan* __restrict__ plans) {
    const uint32_t i0 = p.own_begin + tile * SPH_THREADS;
    make_tile_plan_range(p, keys, lb, tile, i0, min(i0 + SPH_THREADS, p.own_begin + p.n), plans);
}

__global__ void __launch_bounds__(128)
k_tile_plans(const DevParams p, const uint32_t* __restrict__ keys, const uint32_t* __restrict__ lb, uint32_t ntiles,
             TilePlan* __restrict__ plans) {
    pdl_enter();
    const uint32_t tile = blockIdx.x * blockDim.x + threadIdx.x;
    if (tile < ntiles) make_tile_plan(p, keys, lb, tile, plans);
}

// Which tile does CTA b of a neighbour pass compute? With gravity along y the fluid piles up in the first or the last grid
// rows, i.e. at one END of the row-major key order, and a crowded tile costs up to 40x an ordinary one (candidates per
// particle = the population of 3x3 cells): in index order the heaviest CTAs would start last and the pass would end on a
// few SMs (1M dam break after 320 ticks: a schedule model gives 1.6x the balanced makespan, a run with the order reversed
// confirmed it). So on grids of few waves the CTAs take tiles alternately from both ends towards the middle: whichever end
// is crowded starts at once, each front still walks neighbouring tiles (L2 reuse of the rows between them). On large grids
// (hundreds of waves) the tail does not matter and the single front keeps the most reuse: the host leaves the order alone
// there (BOTH_ENDS is a template parameter of the two kernels, chosen by Sim::tiles_both_ends(): the index-order kernels are compiled as before). A list of heavy tiles built by the plan makers and taken first by the CTAs (two tiles
// per CTA) was built and measured: the second pass per CTA cost 9 % on the uniform box and won less than this on the dam
// break; dropped.
__device__ __forceinline__ uint32_t tile_from_both_ends(uint32_t b, uint32_t first, uint32_t count) {
    return first + ((b & 1u) ? count - 1u - (b >> 1) : (b >> 1));
}

// thread 0 of a CTA: read the tile's plan, publish the staging bases, issue the copies.
// arrays[k] / smem[k] / elem_bytes[k]: the SoA streams to stage (1 for density, 3 for force).
template <int NARR>
__device__ __forceinline__ void stage_from_plan(const TilePlan* __restrict__ tp, const void* const (&arrays)[NARR],
                                                void* const (&smem)[NARR], const uint32_t (&elem_bytes)[NARR], uint64_t* bar,
                                                SpanPlan* plan) {
    constexpr int NI = SPH_MAX_ROWS * 3;
    const uint4* src = reinterpret_cast<const uint4*>(tp);
    const uint4 w0 = __ldg(src), w1 = __ldg(src + 1), w2 = __ldg(src + 2), w3 = __ldg(src + 3);
    const uint32_t lo[NI] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y};
    const uint32_t len[NI] = {w1.z, w1.w, w2.x, w2.y, w2.z, w2.w};
    static_assert(NI == 6, "TilePlan word layout");
    plan->row_first = w3.x;
    const uint32_t flags = w3.y;
    if (!(flags & 1u)) {
        plan->staged = 0;
        return;
    }
    plan->staged = 1;
    uint32_t total = 0, off[NI];
#pragma unroll
    for (int t = 0; t < NI; t++) { off[t] = total; total += len[t]; }
#pragma unroll
    for (int t = 0; t < NI; t++) plan->base[t] = (flags & 2u) ? (int)lo[t] - (int)off[t] : (int)lo[0];
    uint32_t bytes = 0;
#pragma unroll
    for (int k = 0; k < NARR; k++) bytes += total * elem_bytes[k];
    mbar_arrive_expect_tx(bar, bytes);
#pragma unroll
    for (int t = 0; t < NI; t++) {
        if (len[t] == 0) continue;
#pragma unroll
        for (int k = 0; k < NARR; k++) {
            bulk_g2s(reinterpret_cast<char*>(smem[k]) + (size_t)off[t] * elem_bytes[k],
                     reinterpret_cast<const char*>(arrays[k]) + (size_t)lo[t] * elem_bytes[k], len[t] * elem_bytes[k], bar);
        }
    }
}

// staging-buffer bases for a thread in cell row `cy`
__device__ __forceinline__ void thread_bases(const SpanPlan& plan, uint32_t cy, int (&base)[3]) {
    uint32_t g = cy - plan.row_first;
    if (g >= (uint32_t)SPH_MAX_ROWS) g = 0;  // single-interval plans: all entries equal
#pragma unroll
    for (int r = 0; r < 3; r++) base[r] = plan.base[g * 3 + r];
}

// this thread's three spans [a[r], b[r]) in global sorted-index space
__device__ __forceinline__ uint32_t thread_spans(const DevParams& p, uint32_t key, const uint32_t* __restrict__ lb,
                                                 uint32_t c0, uint32_t s_stale, uint32_t (&a)[3], uint32_t (&b)[3]) {
    const uint32_t k = min(key, p.num_cells - 1u);
    const uint32_t cy = k / p.grid_w, cx = k - cy * p.grid_w;
#pragma unroll
    for (int r = 0; r < 3; r++) {
        const uint32_t row = cy + (uint32_t)r - 1u;
        a[r] = 0; b[r] = 0;
        if (row < p.grid_h) {  // [survey-pinned] rows outside the grid are empty
            const uint32_t ca = row * p.grid_w + (cx > 0u ? cx - 1u : 0u);
            const uint32_t cb = row * p.grid_w + min(cx + 2u, p.grid_w);  // column grid_w is the next row's empty column 0
            uint32_t lo = __ldg(lb + ca), hi = __ldg(lb + cb);
            // lowest occupied cell: the reference walks it from its stale start_indices entry
            // (only that cell: if the stale entry is past its end the cell reads as empty, the cells after it
            // in the span are unaffected)
            if (ca <= c0 && c0 < cb) lo = min(max(lo, s_stale), __ldg(lb + c0 + 1u));
            a[r] = lo; b[r] = hi;
        }
    }
    return cy;
}

// ------------------------------------------------------------------------------------------------
// K5 density
// ------------------------------------------------------------------------------------------------
// funcs.wgsl:186-192: one candidate's contribution, added in visiting order
__device__ __forceinline__ void density_candidate(const DevParams& p, float2 q, float px, float py, float& density) {
    const float dx = fsub(q.x, px), dy = fsub(q.y, py);
    const float r2 = fadd(fmul(dx, dx), fmul(dy, dy));
    if (!(r2 > p.h2)) {  // funcs.wgsl:74: r2 > h2 -> 0 (+0 leaves the sum unchanged); NaN falls through, as there
        const float diff = fsub(p.h2, r2);
        const float k = fmul(fmul(fmul(p.poly6_norm, diff), diff), diff);
        density = fadd(density, fmul(p.mass, k));
    }
}

// fallback: straight from global memory
__device__ __forceinline__ float density_walk_global(const DevParams& p, const float2* __restrict__ pred,
                                                     const uint32_t (&a)[3], const uint32_t (&b)[3], float px, float py) {
    float density = 0.0f;
#pragma unroll
    for (int r = 0; r < 3; r++)
        for (uint32_t j = a[r]; j < b[r]; j++) density_candidate(p, __ldg(pred + j), px, py, density);
    return density;
}

// Inner loop of the staged walk, hand-scheduled, TWO candidates per iteration on the packed f32x2 instructions of
// sm_100a (FADD2 / FMUL2: two IEEE-rounded fp32 results per issue slot — the walk is issue-bound, not FP-pipe-bound):
// 2 LDS.64, 2 FADD2 (dx, dy of A and of B), 2 FMUL2, 2 FADD (r2), then packed across the two candidates: FADD2
// (h2 - r2), 4 FMUL2 (norm * d * d * d, mass *), 2 FSETP, the two predicated FADDs into the sum in visiting order,
// address bump, 2 ISETP, BRA = 21 instructions per two candidates (the scalar loop took 15 per candidate). An odd
// span ends with B predicated off (its load stays inside the padded staging buffer). The compiler's own version
// unrolled each span x4 / x2 / x1 with separate remainder loops, so lanes of a warp sat in three different loops per
// span (ncu: 27.7 warp instructions per particle for ~36 candidates).
// Tried and dropped (round 1): re-assigning particles to threads by neighbour count to even out the warps of the
// force pass — with the scan it made the candidate walk diverge (5.7 -> 5.9 ms), with a scanner/evaluator split it
// needed CTA-wide barriers per round (7.2 ms).
// The loop also records which candidates it accepted — the same test on the same operands the force pass's candidate scan
// would repeat (compute.wgsl:202 == funcs.wgsl:74 on identical r2) — as a bit string, first candidate in the HIGHEST
// position reached (shift left by two per iteration, +2 for A, +1 for B: three integer instructions per two candidates);
// density_walk_staged turns it into a mask with candidate c at bit c.
__device__ __forceinline__ void density_span(uint32_t addr, uint32_t end, float px, float py, float h2, float norm, float mass,
                                             float& density, uint32_t& hits) {
    asm volatile(
        "{\n"
        ".reg .pred ia, ib, vb, go;\n"
        ".reg .f32 sa, sb, ka, kb;\n"
        ".reg .b64 me, pa, pb, s2, d2, k2, hh, nn, mm;\n"
        ".reg .u32 nx;\n"
        "setp.lt.u32 go, %1, %3;\n"
        "@!go bra DENS_DONE;\n"
        "mov.b64 me, {%4, %5};\n"
        "mov.b64 hh, {%6, %6};\n"
        "mov.b64 nn, {%7, %7};\n"
        "mov.b64 mm, {%8, %8};\n"
        "DENS_LOOP:\n"
        ".pragma \"nounroll\";\n"       // ptxas would re-create the x4 / x2 / x1 remainder structure
        "ld.shared.b64 pa, [%1];\n"
        "ld.shared.b64 pb, [%1+8];\n"
        "add.u32 nx, %1, 8;\n"
        "sub.rn.f32x2 pa, pa, me;\n"
        "sub.rn.f32x2 pb, pb, me;\n"
        "mul.rn.f32x2 pa, pa, pa;\n"
        "mul.rn.f32x2 pb, pb, pb;\n"
        "mov.b64 {sa, ka}, pa;\n"
        "mov.b64 {sb, kb}, pb;\n"
        "add.rn.f32 sa, sa, ka;\n"
        "add.rn.f32 sb, sb, kb;\n"
        "setp.lt.u32 vb, nx, %3;\n"
        "mov.b64 s2, {sa, sb};\n"
        "sub.rn.f32x2 d2, hh, s2;\n"
        "mul.rn.f32x2 k2, nn, d2;\n"    // funcs.wgsl:76: norm * d * d * d, left to right
        "mul.rn.f32x2 k2, k2, d2;\n"
        "mul.rn.f32x2 k2, k2, d2;\n"
        "mul.rn.f32x2 k2, mm, k2;\n"    // funcs.wgsl:189-192: mass * kernel
        "setp.leu.f32 ia, sa, %6;\n"    // funcs.wgsl:74: r2 > h2 -> 0; NaN falls through, as there
        "setp.leu.and.f32 ib, sb, %6, vb;\n"
        "mov.b64 {ka, kb}, k2;\n"
        "shl.b32 %2, %2, 2;\n"
        "@ia add.rn.f32 %0, %0, ka;\n"
        "@ia add.u32 %2, %2, 2;\n"
        "@ib add.rn.f32 %0, %0, kb;\n"
        "@ib add.u32 %2, %2, 1;\n"
        "add.u32 %1, %1, 16;\n"
        "setp.lt.u32 go, %1, %3;\n"
        "@go bra DENS_LOOP;\n"
        "DENS_DONE:\n"
        "}\n"
        : "+f"(density), "+r"(addr), "+r"(hits)
        : "r"(end), "f"(px), "f"(py), "f"(h2), "f"(norm), "f"(mass));
}

// staged: the three row spans in visiting order, explicit 32-bit shared addresses
__device__ __forceinline__ float density_walk_staged(const DevParams& p, uint32_t sp_addr, const int (&base)[3],
                                                     const uint32_t (&a)[3], const uint32_t (&b)[3], float px, float py,
                                                     uint32_t (&mask)[3], bool& masks_ok) {
    float density = 0.0f;
    masks_ok = true;
#pragma unroll
    for (int r = 0; r < 3; r++) {
        uint32_t hits = 0;
        density_span(sp_addr + (a[r] - (uint32_t)base[r]) * 8u, sp_addr + (b[r] - (uint32_t)base[r]) * 8u, px, py, p.h2,
                     p.poly6_norm, p.mass, density, hits);
        // the loop took ceil(len / 2) iterations: candidate c sits at bit 2 * iterations - 1 - c
        const uint32_t len = b[r] - a[r];
        const uint32_t even = (len + 1u) & ~1u;
        masks_ok = masks_ok && len <= 32u;
        mask[r] = len ? (__brev(hits) >> (32u - min(even, 32u))) : 0u;
    }
    return density;
}

// What the density pass hands to the force pass per particle (24 B): the starts of its three row spans and, per span, the
// candidates with !(r2 > h2) as a bit mask — so that the force pass neither walks the key -> cell-table chain again nor
// repeats the distance test on ~36 candidates to find its ~12 neighbours (compute.wgsl:160-299 walks the 9 cells twice
// more; here they are walked once, by the density pass). head.x bit 31: masks not usable (a span longer than 32
// candidates, or a tile that was not staged): the force pass scans as before; bit 30: the tile-plan row group.
constexpr uint32_t NB_FALLBACK = 0x80000000u, NB_GROUP = 0x40000000u, NB_INDEX = 0x3fffffffu;

template <bool BAND, bool BOTH_ENDS>
__global__ void __launch_bounds__(SPH_THREADS)
k_density(const DevParams p, const float2* __restrict__ pred_s, const uint32_t* __restrict__ keys,
          const uint32_t* __restrict__ lb, const uint32_t* __restrict__ quirk, float* __restrict__ rho,
          uint32_t* __restrict__ wild_stamp, const uint32_t* __restrict__ frame_time,
          const TilePlan* __restrict__ plans,
          const uint32_t* __restrict__ band_range /* multi-GPU band: {first, end} of the particles to compute, on the device; tiles count from 0. Else null */,
          uint4* __restrict__ nb_head, uint2* __restrict__ nb_mask /* per particle: span starts + acceptance masks for the force pass */) {
    __shared__ __align__(128) float2 s_pred[SPH_CAP + 2];  // + pad: the pair loop may load one record past a span
    __shared__ uint64_t s_bar;
    __shared__ SpanPlan s_plan;
    pdl_enter();
    // band mode (kernels_band.cuh): the key-ordered arrays hold [ghost rows | owned | ghost rows]; density is computed for
    // the owned particles and one ghost row on each side, a range only the device knows
    // (BAND is a template parameter so that the single-GPU kernel is compiled exactly as it was without band support)
    const uint32_t r_begin = BAND ? __ldg(band_range) : p.own_begin;
    const uint32_t r_end = BAND ? __ldg(band_range + 1) : p.own_begin + p.n;
    uint32_t tile = blockIdx.x;
    if (BOTH_ENDS) {
        // band: the tiles that hold [r_begin, r_end) (tiles count from sorted index 0), the CTAs beyond them have nothing to do
        const uint32_t t_first = BAND ? r_begin / SPH_THREADS : 0u;
        const uint32_t t_count = BAND ? (r_end > r_begin ? (r_end + SPH_THREADS - 1u) / SPH_THREADS - t_first : 0u) : gridDim.x;
        if (BAND && blockIdx.x >= t_count) return;  // (whole CTA)
        tile = tile_from_both_ends(blockIdx.x, t_first, t_count);
    }
    const uint32_t i0 = (BAND ? 0u : p.own_begin) + tile * SPH_THREADS;
    if (BAND && !BOTH_ENDS && (i0 + SPH_THREADS <= r_begin || i0 >= r_end)) return;  // (whole CTA)
    if (threadIdx.x == 0) {
        mbar_init(&s_bar, 1);
        mbar_fence_init();
        const void* const arrays[1] = {pred_s};
        void* const smem[1] = {s_pred};
        const uint32_t eb[1] = {8u};
        stage_from_plan<1>(plans + tile, arrays, smem, eb, &s_bar, &s_plan);
    }
    const uint32_t i = i0 + threadIdx.x;
    const bool valid = BAND ? (i >= r_begin && i < r_end) : (i < r_end);
    uint32_t a[3] = {0, 0, 0}, b[3] = {0, 0, 0};
    float2 me = make_float2(0.f, 0.f);
    uint32_t cy = 0;
    if (valid) {
        cy = thread_spans(p, __ldg(keys + i), lb, __ldg(quirk), __ldg(quirk + 1), a, b);
        me = __ldg(pred_s + i);
    }
    __syncthreads();
    const int staged = s_plan.staged;
    float d;
    uint32_t mask[3] = {0, 0, 0};
    bool masks_ok = false;
    if (staged) {
        mbar_wait(&s_bar, 0);
        int base[3];
        thread_bases(s_plan, valid ? cy : s_plan.row_first, base);
        d = density_walk_staged(p, smem_u32_pinned(s_pred), base, a, b, me.x, me.y, mask, masks_ok);
    } else {
        d = density_walk_global(p, pred_s, a, b, me.x, me.y);
    }
    // funcs.wgsl:202 max(density, EPSILON); compute.wgsl:70 max(.., 0.1)
    if (valid) {
        uint32_t g = cy - s_plan.row_first;
        if (g >= (uint32_t)SPH_MAX_ROWS) g = 0;
        nb_head[i] = make_uint4((a[0] & NB_INDEX) | (masks_ok ? 0u : NB_FALLBACK) | (g ? NB_GROUP : 0u), a[1], a[2], mask[0]);
        nb_mask[i] = make_uint2(mask[1], mask[2]);
        const float dens = fmaxf(fmaxf(d, 1.19209290e-07f), 0.1f);
        rho[i] = dens;
        if (wild_stamp && !(dens <= DIV_DEN_HI)) *wild_stamp = __ldg(frame_time);  // common.cuh tame_component
    }
}

// ------------------------------------------------------------------------------------------------
// K6 force + integrate
// ------------------------------------------------------------------------------------------------
// funcs.wgsl:129-142
__device__ __forceinline__ float rand_f32(uint32_t& s) {
    uint32_t x = s;
    x ^= x << 13; x ^= x >> 17; x ^= x << 5;
    s = x;
    return fmul(__uint2float_rn(x), 2.3283064365386962890625e-10f);  // exact: / 4294967296.0
}
// funcs.wgsl:112-123
__device__ __forceinline__ float viscosity_kernel(const DevParams& p, float r) {
    if (r <= p.h) {
        if (r == 0.0f) return p.visc_k;
        const float A = fdiv(-fmul(fmul(r, r), r), p.two_h3);
        const float B = fdiv(fmul(r, r), p.h2);
        const float C = fdiv(p.h, fmul(2.0f, r));
        return fmul(p.visc_k, fsub(fadd(fadd(A, B), C), 1.0f));
    }
    return 0.0f;
}

struct ForceAcc {
    float px, py;  // pressure sum
    float vx, vy;  // viscosity sum (before * viscosity_coefficient)
    float ax, ay;  // sum of |terms| (debug only)
};

// one accepted neighbour: pressure term (compute.wgsl:206-223) and viscosity term (:283-288), each added
// to its own accumulator in visiting order.
//   EXACT: every operation IEEE-rounded in the reference's order (bit-identical to the CPU restatement).
//   !EXACT ("fast"): one MUFU.RSQ gives 1/r (so r = sq/r', dir = d/r', h/(2r)) and one MUFU.RCP gives 1/rho_j;
//   FMA contraction allowed; constants' reciprocals hoisted. Agrees with EXACT to a few ulp per term
//   (tests bound it by 1e-5 of sum|terms|); branch decisions (r2 cut-off, r == 0, r <= h) are unchanged.
// per-thread loop invariants of the exact pair body: refined reciprocals of the two constant denominators
struct PairConsts {
    float y_two_h3, y_h2;
    bool fast_div;    // h2, two_h3 in the shared-reciprocal range (else every pair takes the __fdiv_rn body)
    bool tn_bounded;  // h c_s |k_p| (2^20 + |rho_0|) <= 2^98: no pressure numerator of a tame tick can exceed 2^100
};
__device__ __forceinline__ PairConsts make_pair_consts(const DevParams& p) {
    PairConsts c;
    c.y_two_h3 = rcp_refined(p.two_h3);
    c.y_h2 = rcp_refined(p.h2);
    c.fast_div = p.two_h3 >= DIV_DEN_LO && p.two_h3 <= DIV_DEN_HI && p.h2 >= DIV_DEN_LO && p.h2 <= DIV_DEN_HI &&
                 p.h >= DIV_DEN_LO && p.h <= DIV_DEN_HI;
    c.tn_bounded = fabsf(p.h * p.spiky) * fabsf(p.pressure_constant) * (DIV_DEN_HI + fabsf(p.rest_density)) <= 3.1691265e29f;
    return c;
}

// the pair body with IEEE division throughout (== the CPU restatement operation for operation): taken by pairs whose
// operands fall outside the shared-reciprocal range, including the coincident-particle RNG branch
struct SlowConsts { float h, h2, two_h3, visc_k; };  // by value: a DevParams reference would force a local copy
struct SlowTerms { float tx, ty, ux, uy; uint32_t seed; };  // by value: references would go through the stack
__device__ __noinline__ SlowTerms force_pair_slow_call(const SlowConsts p, float dx, float dy, float dst, float k, float shared_p,
                                                       float dvx, float dvy, float nd, uint32_t rand_seed) {
    float tx, ty, ux, uy;
    float sdirx, sdiry;
    if (dst == 0.0f) {  // compute.wgsl:211-212
        const float rx = rand_f32(rand_seed);
        const float ry = rand_f32(rand_seed);
        const float len = fsqrt(fadd(fmul(rx, rx), fmul(ry, ry)));
        sdirx = fdiv(rx, len); sdiry = fdiv(ry, len);
    } else {
        sdirx = fdiv(dx, dst); sdiry = fdiv(dy, dst);
    }
    tx = fdiv(fmul(fmul(sdirx, k), shared_p), nd);
    ty = fdiv(fmul(fmul(sdiry, k), shared_p), nd);
    float skv = 0.0f;  // funcs.wgsl:112-123
    if (dst <= p.h) {
        if (dst == 0.0f) skv = p.visc_k;
        else {
            const float A = fdiv(-fmul(fmul(dst, dst), dst), p.two_h3);
            const float B = fdiv(fmul(dst, dst), p.h2);
            const float C = fdiv(p.h, fmul(2.0f, dst));
            skv = fmul(p.visc_k, fsub(fadd(fadd(A, B), C), 1.0f));
        }
    }
    ux = fmul(fdiv(dvx, nd), skv);
    uy = fmul(fdiv(dvy, nd), skv);
    return SlowTerms{tx, ty, ux, uy, rand_seed};
}
__device__ __forceinline__ void force_pair_slow(const SlowConsts p, float dx, float dy, float dst, float k, float shared_p,
                                                float dvx, float dvy, float nd, uint32_t& rand_seed, float& tx, float& ty,
                                                float& ux, float& uy) {
    const SlowTerms t = force_pair_slow_call(p, dx, dy, dst, k, shared_p, dvx, dvy, nd, rand_seed);
    tx = t.tx; ty = t.ty; ux = t.ux; uy = t.uy; rand_seed = t.seed;
}

template <bool DEBUG, bool EXACT>
__device__ __forceinline__ void force_pair(const DevParams& p, const PairConsts& c, float dx, float dy, float sq, float2 nv,
                                           float nd, float vx, float vy, float pressure, uint32_t& rand_seed, ForceAcc& acc) {
    float tx, ty, ux, uy;
    if (EXACT) {
        // Every quotient below is the correctly rounded one (== the CPU restatement's `/`). The common case
        // computes the 9 divisions of a pair from 2 refined reciprocals (1/r, 1/rho_j; common.cuh div_by) and
        // the two hoisted constant ones; operands outside the range where that is provably exact (r < 2^-20
        // including the coincident-particle RNG branch, tiny non-zero or huge numerators) take __fdiv_rn.
        const float dst = fsqrt(sq);
        const float np = fmul(p.pressure_constant, fsub(nd, p.rest_density));
        const float shared_p = fmul(fadd(pressure, np), 0.5f);
        const float dvx = fsub(nv.x, vx), dvy = fsub(nv.y, vy);
        const bool in_h = dst <= p.h;
        const float k = in_h ? fmul(-fsub(p.h, dst), p.spiky) : 0.0f;  // funcs.wgsl:101-109
        const float yr = rcp_refined(dst), yd = rcp_refined(nd);
        const float dirx = div_by(dx, dst, yr), diry = div_by(dy, dst, yr);
        const float tnx = fmul(fmul(dirx, k), shared_p), tny = fmul(fmul(diry, k), shared_p);
        tx = div_by(tnx, nd, yd);
        ty = div_by(tny, nd, yd);
        const float r2 = fmul(dst, dst);
        const float A = div_by(-fmul(r2, dst), p.two_h3, c.y_two_h3);
        const float B = div_by(r2, p.h2, c.y_h2);
        const float C = fmul(0.5f, div_by(p.h, dst, yr));  // h / (2 r): exact power-of-two scaling of h / r
        const float kv = in_h ? fmul(p.visc_k, fsub(fadd(fadd(A, B), C), 1.0f)) : 0.0f;  // funcs.wgsl:112-123
        ux = fmul(div_by(dvx, nd, yd), kv);  // compute.wgsl:288
        uy = fmul(div_by(dvy, nd, yd), kv);
        const int safe = (int)c.fast_div & (int)(sq >= DIV_DEN_LO * DIV_DEN_LO) & (int)(nd >= DIV_DEN_LO) &
                         (int)(nd <= DIV_DEN_HI) & div_num_lo_ok(dx) & div_num_lo_ok(dy) & div_num_ok(dvx) &
                         div_num_ok(dvy) & div_num_ok(tnx) & div_num_ok(tny);
        if (!safe) force_pair_slow(SlowConsts{p.h, p.h2, p.two_h3, p.visc_k}, dx, dy, dst, k, shared_p, dvx, dvy, nd, rand_seed, tx, ty, ux, uy);
        acc.px = fadd(acc.px, tx);
        acc.py = fadd(acc.py, ty);
        acc.vx = fadd(acc.vx, ux);
        acc.vy = fadd(acc.vy, uy);
    } else {
        float dst, dirx, diry, half_h_over_r;
        if (sq == 0.0f) {
            const float rx = rand_f32(rand_seed);
            const float ry = rand_f32(rand_seed);
            const float il = rsqrt_approx(rx * rx + ry * ry);
            dirx = rx * il; diry = ry * il;  // (0,0) draws give NaN here as in the reference; the guard zeroes it
            dst = 0.0f; half_h_over_r = 0.0f;
        } else {
            const float ir = rsqrt_approx(sq);
            dst = sq * ir;
            dirx = dx * ir; diry = dy * ir;
            half_h_over_r = 0.5f * p.h * ir;
        }
        const float ind = rcp_approx(nd);
        const float np = p.pressure_constant * (nd - p.rest_density);
        const float k = (dst <= p.h) ? -(p.h - dst) * p.spiky : 0.0f;
        const float w = k * ((pressure + np) * 0.5f) * ind;
        tx = dirx * w; ty = diry * w;
        acc.px += tx; acc.py += ty;
        float kv = 0.0f;
        if (dst <= p.h) {
            const float r2 = dst * dst;
            kv = (dst == 0.0f) ? p.visc_k : p.visc_k * (((-(r2 * dst) * p.inv_two_h3) + r2 * p.inv_h2 + half_h_over_r) - 1.0f);
        }
        const float wv = ind * kv;
        ux = (nv.x - vx) * wv; uy = (nv.y - vy) * wv;
        acc.vx += ux; acc.vy += uy;
    }
    if (DEBUG) {
        acc.ax += fabsf(tx) + fabsf(ux) * p.visc_coef;
        acc.ay += fabsf(ty) + fabsf(uy) * p.visc_coef;
    }
}

// Two accepted neighbours (A, then B in visiting order) per call, every operation of the exact body packed across
// the two (common.cuh f32x2): the same IEEE-rounded operations in the same order per neighbour, so the same bits, at
// about half the issue slots. hasB = false evaluates A twice and drops the copy. A neighbour outside the
// shared-reciprocal range takes force_pair_slow; the terms are added A first, then B (compute.wgsl:206-223, :283-288).
template <bool DEBUG>
__device__ __forceinline__ void force_pair2(const DevParams& p, const PairConsts& c, float dxA, float dyA, float dxB, float dyB,
                                            float2 nvA, float2 nvB, float ndA, float ndB, bool hasB, bool tame, float vx, float vy,
                                            float pressure, uint32_t& rand_seed, ForceAcc& acc) {
    const f32x2 dx = pk2(dxA, dxB), dy = pk2(dyA, dyB), nd = pk2(ndA, ndB);
    const float dvxA = fsub(nvA.x, vx), dvyA = fsub(nvA.y, vy), dvxB = fsub(nvB.x, vx), dvyB = fsub(nvB.y, vy);
    const f32x2 dvx = pk2(dvxA, dvxB), dvy = pk2(dvyA, dvyB);
    // ptxas (12.9) contracts mul.rn.f32x2 + add.rn.f32x2 into FFMA2 even though both carry .rn (and under -fmad=false);
    // a sum whose addend is an inexact packed product is therefore taken with scalar FADDs (add2_of_products)
    const f32x2 sq = add2_of_products(mul2(dx, dx), mul2(dy, dy));
    const f32x2 dst = sqrt2_inrange(sq);
    const f32x2 np = mul2(bc2(p.pressure_constant), sub2(nd, bc2(p.rest_density)));
    const f32x2 shared_p = mul2(add2_of_products(bc2(pressure), np), bc2(0.5f));
    const bool inA = lo2(dst) <= p.h, inB = hi2(dst) <= p.h;
    const f32x2 kraw = mul2(sub2(bc2(p.h), dst), bc2(-p.spiky));  // -(h - r) * c_s, funcs.wgsl:101-109
    const float kA = inA ? lo2(kraw) : 0.0f, kB = inB ? hi2(kraw) : 0.0f;
    const f32x2 k = pk2(kA, kB);
    const f32x2 yr = rcp_refined2(dst), yd = rcp_refined2(nd);
    const f32x2 dirx = div_by2(dx, dst, yr), diry = div_by2(dy, dst, yr);
    const f32x2 tnx = mul2(mul2(dirx, k), shared_p), tny = mul2(mul2(diry, k), shared_p);
    const f32x2 tx2 = div_by2(tnx, nd, yd), ty2 = div_by2(tny, nd, yd);
    const f32x2 r2 = mul2(dst, dst);
    const f32x2 r3 = mul2(r2, dst);
    const f32x2 A = div_by2(neg2(r3), bc2(p.two_h3), bc2(c.y_two_h3));
    const f32x2 B = div_by2(r2, bc2(p.h2), bc2(c.y_h2));
    // (A + B) + h / (2 r): the halving is exact (h / r >= 1 here), so fusing it into the addition rounds the same sum
    const f32x2 ABC = fma2(div_by2(bc2(p.h), dst, yr), bc2(0.5f), add2(A, B));
    const f32x2 kvraw = mul2(bc2(p.visc_k), sub2(ABC, bc2(1.0f)));  // funcs.wgsl:112-123
    const float kvA = inA ? lo2(kvraw) : 0.0f, kvB = inB ? hi2(kvraw) : 0.0f;
    const f32x2 kv = pk2(kvA, kvB);
    const f32x2 ux2 = mul2(div_by2(dvx, nd, yd), kv), uy2 = mul2(div_by2(dvy, nd, yd), kv);  // compute.wgsl:288
    float txA = lo2(tx2), tyA = lo2(ty2), uxA = lo2(ux2), uyA = lo2(uy2);
    float txB = hi2(tx2), tyB = hi2(ty2), uxB = hi2(ux2), uyB = hi2(uy2);
    // range tests of the shared-reciprocal quotients (common.cuh div_by). With tame inputs (the tick's stamp) only
    // r >= 2^-20 and the lower bound of the pressure numerators are left to test per pair; their upper bound is
    // c.tn_bounded: |dir| <= 1, |k| <= h c_s, |shared pressure| <= |k_p| (2^20 + |rho_0|).
    int safeA = (int)c.fast_div & (int)(lo2(sq) >= DIV_DEN_LO * DIV_DEN_LO) & div_num_lo_ok(lo2(tnx)) & div_num_lo_ok(lo2(tny));
    int safeB = (int)c.fast_div & (int)(hi2(sq) >= DIV_DEN_LO * DIV_DEN_LO) & div_num_lo_ok(hi2(tnx)) & div_num_lo_ok(hi2(tny));
    if (!(tame && c.tn_bounded)) {
        safeA &= (int)(ndA >= DIV_DEN_LO) & (int)(ndA <= DIV_DEN_HI) & div_num_lo_ok(dxA) & div_num_lo_ok(dyA) & div_num_ok(dvxA) &
                 div_num_ok(dvyA) & div_num_ok(lo2(tnx)) & div_num_ok(lo2(tny));
        safeB &= (int)(ndB >= DIV_DEN_LO) & (int)(ndB <= DIV_DEN_HI) & div_num_lo_ok(dxB) & div_num_lo_ok(dyB) & div_num_ok(dvxB) &
                 div_num_ok(dvyB) & div_num_ok(hi2(tnx)) & div_num_ok(hi2(tny));
    }
    if (!(safeA & safeB)) {
        if (!safeA) {
            const float sqA = fadd(fmul(dxA, dxA), fmul(dyA, dyA));
            const float dstA = fsqrt(sqA);
            const float kk = (dstA <= p.h) ? fmul(-fsub(p.h, dstA), p.spiky) : 0.0f;
            const float sp = fmul(fadd(pressure, fmul(p.pressure_constant, fsub(ndA, p.rest_density))), 0.5f);
            force_pair_slow(SlowConsts{p.h, p.h2, p.two_h3, p.visc_k}, dxA, dyA, dstA, kk, sp, dvxA, dvyA, ndA, rand_seed, txA, tyA, uxA, uyA);
        }
        if (hasB && !safeB) {
            const float sqB = fadd(fmul(dxB, dxB), fmul(dyB, dyB));
            const float dstB = fsqrt(sqB);
            const float kk = (dstB <= p.h) ? fmul(-fsub(p.h, dstB), p.spiky) : 0.0f;
            const float sp = fmul(fadd(pressure, fmul(p.pressure_constant, fsub(ndB, p.rest_density))), 0.5f);
            force_pair_slow(SlowConsts{p.h, p.h2, p.two_h3, p.visc_k}, dxB, dyB, dstB, kk, sp, dvxB, dvyB, ndB, rand_seed, txB, tyB, uxB, uyB);
        }
    }
    acc.px = fadd(acc.px, txA);
    acc.py = fadd(acc.py, tyA);
    acc.vx = fadd(acc.vx, uxA);
    acc.vy = fadd(acc.vy, uyA);
    if (DEBUG) {
        acc.ax += fabsf(txA) + fabsf(uxA) * p.visc_coef;
        acc.ay += fabsf(tyA) + fabsf(uyA) * p.visc_coef;
    }
    if (hasB) {
        acc.px = fadd(acc.px, txB);
        acc.py = fadd(acc.py, tyB);
        acc.vx = fadd(acc.vx, uxB);
        acc.vy = fadd(acc.vy, uyB);
        if (DEBUG) {
            acc.ax += fabsf(txB) + fabsf(uxB) * p.visc_coef;
            acc.ay += fabsf(tyB) + fabsf(uyB) * p.visc_coef;
        }
    }
}

// the "fast" body for two neighbours at once (FLUID_MATH_FAST: one MUFU.RSQ and one MUFU.RCP per neighbour, contraction
// allowed, a few ulp per term; branch decisions as in the exact body). A coincident neighbour (r == 0: RNG direction)
// goes through the scalar body.
template <bool DEBUG>
__device__ __forceinline__ void force_pair2_fast(const DevParams& p, const PairConsts& c, float dxA, float dyA, float dxB, float dyB,
                                                 float2 nvA, float2 nvB, float ndA, float ndB, bool hasB, float vx, float vy,
                                                 float pressure, uint32_t& rand_seed, ForceAcc& acc) {
    const f32x2 dx = pk2(dxA, dxB), dy = pk2(dyA, dyB), nd = pk2(ndA, ndB);
    const f32x2 dvx = sub2(pk2(nvA.x, nvB.x), bc2(vx)), dvy = sub2(pk2(nvA.y, nvB.y), bc2(vy));
    const f32x2 sq = fma2(dx, dx, mul2(dy, dy));
    const float sqA = lo2(sq), sqB = hi2(sq);
    if (sqA == 0.0f || (hasB && sqB == 0.0f)) {
        force_pair<DEBUG, false>(p, c, dxA, dyA, sqA, nvA, ndA, vx, vy, pressure, rand_seed, acc);
        if (hasB) force_pair<DEBUG, false>(p, c, dxB, dyB, sqB, nvB, ndB, vx, vy, pressure, rand_seed, acc);
        return;
    }
    const f32x2 ir = pk2(rsqrt_approx(sqA), rsqrt_approx(sqB));
    const f32x2 ind = pk2(rcp_approx(ndA), rcp_approx(ndB));
    const f32x2 dst = mul2(sq, ir);
    const f32x2 dirx = mul2(dx, ir), diry = mul2(dy, ir);
    const bool inA = lo2(dst) <= p.h, inB = hi2(dst) <= p.h;
    const f32x2 np = mul2(bc2(p.pressure_constant), sub2(nd, bc2(p.rest_density)));
    const f32x2 kraw = mul2(sub2(bc2(p.h), dst), bc2(-p.spiky));
    const f32x2 k = pk2(inA ? lo2(kraw) : 0.0f, inB ? hi2(kraw) : 0.0f);
    const f32x2 w = mul2(mul2(k, mul2(add2(bc2(pressure), np), bc2(0.5f))), ind);
    const f32x2 tx2 = mul2(dirx, w), ty2 = mul2(diry, w);
    const f32x2 r2 = mul2(dst, dst);
    const f32x2 poly = sub2(fma2(bc2(0.5f * p.h), ir, fma2(r2, bc2(p.inv_h2), mul2(mul2(r2, dst), bc2(-p.inv_two_h3)))), bc2(1.0f));
    const f32x2 kvraw = mul2(bc2(p.visc_k), poly);
    const f32x2 kv = pk2(inA ? lo2(kvraw) : 0.0f, inB ? hi2(kvraw) : 0.0f);
    const f32x2 wv = mul2(ind, kv);
    const f32x2 ux2 = mul2(dvx, wv), uy2 = mul2(dvy, wv);
    acc.px += lo2(tx2); acc.py += lo2(ty2); acc.vx += lo2(ux2); acc.vy += lo2(uy2);
    if (DEBUG) {
        acc.ax += fabsf(lo2(tx2)) + fabsf(lo2(ux2)) * p.visc_coef;
        acc.ay += fabsf(lo2(ty2)) + fabsf(lo2(uy2)) * p.visc_coef;
    }
    if (hasB) {
        acc.px += hi2(tx2); acc.py += hi2(ty2); acc.vx += hi2(ux2); acc.vy += hi2(uy2);
        if (DEBUG) {
            acc.ax += fabsf(hi2(tx2)) + fabsf(hi2(ux2)) * p.visc_coef;
            acc.ay += fabsf(hi2(ty2)) + fabsf(hi2(uy2)) * p.visc_coef;
        }
    }
}

// fallback walk straight from global memory (CTAs whose neighbourhood does not fit the staging buffer)
template <bool DEBUG, bool EXACT>
__device__ __forceinline__ void force_walk_global(const DevParams& p, const float2* __restrict__ pred,
                                                  const float2* __restrict__ vel, const float* __restrict__ rho,
                                                  const uint32_t (&a)[3], const uint32_t (&b)[3], uint32_t self, float px,
                                                  float py, float vx, float vy, float pressure, uint32_t& rand_seed,
                                                  ForceAcc& acc) {
    const PairConsts pc = make_pair_consts(p);
#pragma unroll
    for (int r = 0; r < 3; r++) {
        for (uint32_t j = a[r]; j < b[r]; j++) {
            if (j == self) continue;  // compute.wgsl:195
            const float2 q = pred[j];
            const float dx = fsub(q.x, px), dy = fsub(q.y, py);
            const float sq = fadd(fmul(dx, dx), fmul(dy, dy));
            if (sq > p.h2) continue;  // compute.wgsl:202 (sqr_radius == h*h)
            force_pair<DEBUG, EXACT>(p, pc, dx, dy, sq, vel[j], rho[j], vx, vy, pressure, rand_seed, acc);
        }
    }
}

// Staged walk, in two alternating warp-synchronous phases so that the expensive pair evaluation runs with
// (nearly) all lanes active: only ~1/3 of the candidates in the 3x3 cells are within h, and evaluating
// inside the candidate loop would execute the whole pair body for a warp whenever ANY lane accepts.
// Phase A scans candidates (cheap distance test) and appends accepted staging-buffer indices to a
// per-thread list of up to FORCE_LIST entries, in visiting order; phase B evaluates the list in that
// order. Visiting order, hence every fp32 sum, is unchanged. All shared accesses use explicit 32-bit
// shared-window addresses so the loops carry no address arithmetic beyond one add.
#ifndef FLUID_FORCE_LIST
#define FLUID_FORCE_LIST 20
#endif
#ifndef FLUID_FORCE_MIN_CTAS
#define FLUID_FORCE_MIN_CTAS 4
#endif
constexpr int FORCE_LIST = FLUID_FORCE_LIST;

// Phase-A inner loop, hand-scheduled: scans candidates [addr, end) of the staged predicted positions and appends
// the shared-window address (its low 16 bits: static shared memory is < 64 KB) of every candidate with
// !(r2 > h2) (compute.wgsl:202) to this thread's list column, until the span or the list is exhausted.
// 13 instructions per candidate (was 19 through the compiler): LDS.64, 2 FADD, 2 FMUL, FADD, FSETP, predicated
// STS + list bump, address bump, 2 ISETP, BRA.
__device__ __forceinline__ void scan_span(uint32_t& addr, uint32_t end, uint32_t& wr, uint32_t wr_end, float px, float py,
                                          float h2) {
    asm volatile(
        "{\n"
        ".reg .pred hit, go;\n"
        ".reg .f32 x, y, s;\n"
        ".reg .b16 lo;\n"
        "setp.lt.u32 go, %0, %2;\n"
        "setp.lt.and.u32 go, %1, %3, go;\n"
        "@!go bra SCAN_DONE;\n"
        "SCAN_LOOP:\n"
        "ld.shared.v2.f32 {x, y}, [%0];\n"
        "sub.rn.f32 x, x, %4;\n"
        "sub.rn.f32 y, y, %5;\n"
        "mul.rn.f32 x, x, x;\n"
        "mul.rn.f32 y, y, y;\n"
        "add.rn.f32 s, x, y;\n"
        "setp.leu.f32 hit, s, %6;\n"
        "cvt.u16.u32 lo, %0;\n"
        "@hit st.shared.u16 [%1], lo;\n"
        "@hit add.u32 %1, %1, " FLUID_LIST_STRIDE_STR ";\n"
        "add.u32 %0, %0, 8;\n"
        "setp.lt.u32 go, %0, %2;\n"
        "setp.lt.and.u32 go, %1, %3, go;\n"
        "@go bra SCAN_LOOP;\n"
        "SCAN_DONE:\n"
        "}\n"
        : "+r"(addr), "+r"(wr)
        : "r"(end), "r"(wr_end), "f"(px), "f"(py), "f"(h2)
        : "memory");
}
// the same scan, two candidates per iteration on packed subtract / multiply (see density_span), while at least two
// candidates and two list slots remain; scan_span finishes the span. 21 instructions per two candidates.
__device__ __forceinline__ void scan_span2(uint32_t& addr, uint32_t end, uint32_t& wr, uint32_t wr_end, float px, float py,
                                           float h2) {
    asm volatile(
        "{\n"
        ".reg .pred ha, hb, go;\n"
        ".reg .f32 sa, sb, ta, tb;\n"
        ".reg .b64 me, pa, pb;\n"
        ".reg .b16 la, lb;\n"
        ".reg .u32 nx, e2, w2;\n"
        "add.u32 e2, %2, -8;\n"
        "add.u32 w2, %3, -" FLUID_LIST_STRIDE_STR ";\n"
        "setp.lt.s32 go, %0, e2;\n"          // shared-window addresses are < 2^31: signed compare survives end - 8 < 0
        "setp.lt.and.s32 go, %1, w2, go;\n"
        "@!go bra SCAN2_DONE;\n"
        "mov.b64 me, {%4, %5};\n"
        "SCAN2_LOOP:\n"
        ".pragma \"nounroll\";\n"
        "ld.shared.b64 pa, [%0];\n"
        "ld.shared.b64 pb, [%0+8];\n"
        "add.u32 nx, %0, 8;\n"
        "sub.rn.f32x2 pa, pa, me;\n"
        "sub.rn.f32x2 pb, pb, me;\n"
        "mul.rn.f32x2 pa, pa, pa;\n"
        "mul.rn.f32x2 pb, pb, pb;\n"
        "mov.b64 {sa, ta}, pa;\n"
        "mov.b64 {sb, tb}, pb;\n"
        "add.rn.f32 sa, sa, ta;\n"
        "add.rn.f32 sb, sb, tb;\n"
        "setp.leu.f32 ha, sa, %6;\n"
        "setp.leu.f32 hb, sb, %6;\n"
        "cvt.u16.u32 la, %0;\n"
        "cvt.u16.u32 lb, nx;\n"
        "@ha st.shared.u16 [%1], la;\n"
        "@ha add.u32 %1, %1, " FLUID_LIST_STRIDE_STR ";\n"
        "@hb st.shared.u16 [%1], lb;\n"
        "@hb add.u32 %1, %1, " FLUID_LIST_STRIDE_STR ";\n"
        "add.u32 %0, %0, 16;\n"
        "setp.lt.s32 go, %0, e2;\n"
        "setp.lt.and.s32 go, %1, w2, go;\n"
        "@go bra SCAN2_LOOP;\n"
        "SCAN2_DONE:\n"
        "}\n"
        : "+r"(addr), "+r"(wr)
        : "r"(end), "r"(wr_end), "f"(px), "f"(py), "f"(h2)
        : "memory");
}
static_assert(SPH_THREADS * 2 == FLUID_LIST_STRIDE, "scan_span takes the list column stride as a literal");

template <bool DEBUG, bool EXACT>
__device__ __forceinline__ void force_walk_staged(const DevParams& p, uint32_t sp_addr, uint32_t sv_addr, uint32_t sr_addr,
                                                  uint32_t list_addr /* this thread's column of [FORCE_LIST][SPH_THREADS] u16 */,
                                                  const int (&base)[3], const uint32_t (&a)[3], const uint32_t (&b)[3],
                                                  uint32_t self, float px, float py, float vx, float vy, float pressure,
                                                  bool tame, uint32_t& rand_seed, ForceAcc& acc) {
    const PairConsts pc = make_pair_consts(p);
    // four spans in visiting order: row y-1; row y before self; row y after self (compute.wgsl:195 skips self by
    // index; a particle lies inside its own row's span unless the stale-entry quirk moved the span start past it); row y+1
    const bool self_in = self >= a[1] && self < b[1];
    const uint32_t A1 = sp_addr + (a[1] - (uint32_t)base[1]) * 8u, E1 = sp_addr + (b[1] - (uint32_t)base[1]) * 8u;
    const uint32_t S1 = self_in ? sp_addr + (self - (uint32_t)base[1]) * 8u : A1;  // end of the part before self
    const uint32_t R1 = self_in ? S1 + 8u : A1;                                    // start of the part after self
    const uint32_t A2 = sp_addr + (a[2] - (uint32_t)base[2]) * 8u, E2 = sp_addr + (b[2] - (uint32_t)base[2]) * 8u;
    uint32_t addr = sp_addr + (a[0] - (uint32_t)base[0]) * 8u, end = sp_addr + (b[0] - (uint32_t)base[0]) * 8u;
    const uint32_t vel_delta = sv_addr - sp_addr, rho_base = sr_addr - (sp_addr >> 1);
    const uint32_t wr_end = list_addr + FORCE_LIST * SPH_THREADS * 2u;
    int r = 0;
    bool more = true;
    while (__any_sync(0xffffffffu, more)) {
        uint32_t wr = list_addr;
        // ---- phase A: scan ----
        while (more) {
            scan_span2(addr, end, wr, wr_end, px, py, p.h2);
            scan_span(addr, end, wr, wr_end, px, py, p.h2);
            if (addr < end) break;  // list full: evaluate it, then resume this span
            if (r == 0) { addr = A1; end = S1; r = 1; }
            else if (r == 1) { addr = R1; end = E1; r = 2; }
            else if (r == 2) { addr = A2; end = E2; r = 3; }
            else more = false;
        }
        __syncwarp();
        // ---- phase B: evaluate ----
        const uint32_t cnt_bytes = wr - list_addr;
        const uint32_t max_bytes = __reduce_max_sync(0xffffffffu, cnt_bytes);
        // two list entries per iteration (force_pair2 / force_pair2_fast); an odd list ends with B = A, dropped
        for (uint32_t k = 0; k < max_bytes; k += SPH_THREADS * 4u) {
            if (k < cnt_bytes) {
                const bool hasB = k + SPH_THREADS * 2u < cnt_bytes;
                const uint32_t eA = lds_u16(list_addr + k);
                const uint32_t eB = hasB ? lds_u16(list_addr + k + SPH_THREADS * 2u) : eA;
                const float2 qA = lds_f2(eA), qB = lds_f2(eB);
                const float2 nvA = lds_f2(eA + vel_delta), nvB = lds_f2(eB + vel_delta);
                const float ndA = lds_f((eA >> 1) + rho_base), ndB = lds_f((eB >> 1) + rho_base);
                if (EXACT)
                    force_pair2<DEBUG>(p, pc, fsub(qA.x, px), fsub(qA.y, py), fsub(qB.x, px), fsub(qB.y, py), nvA, nvB, ndA, ndB,
                                       hasB, tame, vx, vy, pressure, rand_seed, acc);
                else
                    force_pair2_fast<DEBUG>(p, pc, fsub(qA.x, px), fsub(qA.y, py), fsub(qB.x, px), fsub(qB.y, py), nvA, nvB, ndA, ndB,
                                            hasB, vx, vy, pressure, rand_seed, acc);
            }
        }
        __syncwarp();
    }
}

// Staged walk over the neighbours the density pass has already found (nb_head / nb_mask, see k_density): the set bits of the
// three row masks, in ascending order = the reference's visiting order, two per iteration through force_pair2. No candidate
// scan, no accepted list, no key -> cell-table look-ups: ~8 integer instructions per neighbour instead of ~31 per neighbour
// for scanning three candidates to find it. The particle itself (accepted by the density pass, skipped here by index,
// compute.wgsl:195) lies in its own row's span.
template <bool DEBUG, bool EXACT>
__device__ __forceinline__ void force_walk_masked(const DevParams& p, uint32_t sp_addr, uint32_t sv_addr, uint32_t sr_addr,
                                                  const int (&base)[3], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t m0,
                                                  uint32_t m1, uint32_t m2, uint32_t self, float px, float py, float vx, float vy,
                                                  float pressure, bool tame, uint32_t& rand_seed, ForceAcc& acc) {
    const PairConsts pc = make_pair_consts(p);
    if (self - a1 < 32u) m1 &= ~(1u << (self - a1));  // (self < a1 wraps: the stale-entry quirk moved the span start past it)
    const uint32_t addr1 = sp_addr + (a1 - (uint32_t)base[1]) * 8u, addr2 = sp_addr + (a2 - (uint32_t)base[2]) * 8u;
    const uint32_t vel_delta = sv_addr - sp_addr, rho_base = sr_addr - (sp_addr >> 1);
    const uint32_t iters = ((uint32_t)(__popc(m0) + __popc(m1) + __popc(m2)) + 1u) >> 1;
    const uint32_t max_iters = __reduce_max_sync(0xffffffffu, iters);
    const uint32_t addr0 = sp_addr + (a0 - (uint32_t)base[0]) * 8u;
    // next set bit of (m0, m1, m2) in that order -> its staging address. Branch-free on purpose: with a loop over the rows
    // the lanes of a warp switch rows in different iterations and the warp pays for every lane's switch (ncu, first
    // version: 62 warp instructions per iteration at 12 active threads for two pops, against 130 for the pair body).
    auto pop = [&](uint32_t& e) -> bool {
        const bool p0 = m0 != 0u, p1 = m1 != 0u;
        const uint32_t sel = p0 ? m0 : (p1 ? m1 : m2);
        const uint32_t cur = p0 ? addr0 : (p1 ? addr1 : addr2);
        const uint32_t bit = (uint32_t)__ffs((int)sel) - 1u;  // (sel == 0: bit = 0xffffffff, e unused)
        const uint32_t rest = sel & (sel - 1u);
        e = cur + bit * 8u;
        m0 = p0 ? rest : m0;
        m1 = (!p0 && p1) ? rest : m1;
        m2 = (!p0 && !p1) ? rest : m2;
        return sel != 0u;
    };
    for (uint32_t it = 0; it < max_iters; it++) {
        if (it < iters) {
            uint32_t eA = 0, eB = 0;
            pop(eA);
            const bool hasB = pop(eB);
            if (!hasB) eB = eA;
            const float2 qA = lds_f2(eA), qB = lds_f2(eB);
            const float2 nvA = lds_f2(eA + vel_delta), nvB = lds_f2(eB + vel_delta);
            const float ndA = lds_f((eA >> 1) + rho_base), ndB = lds_f((eB >> 1) + rho_base);
            if (EXACT)
                force_pair2<DEBUG>(p, pc, fsub(qA.x, px), fsub(qA.y, py), fsub(qB.x, px), fsub(qB.y, py), nvA, nvB, ndA, ndB,
                                   hasB, tame, vx, vy, pressure, rand_seed, acc);
            else
                force_pair2_fast<DEBUG>(p, pc, fsub(qA.x, px), fsub(qA.y, py), fsub(qB.x, px), fsub(qB.y, py), nvA, nvB, ndA, ndB,
                                        hasB, vx, vy, pressure, rand_seed, acc);
        }
    }
}

template <bool DEBUG, bool EXACT, bool BAND, bool BOTH_ENDS>
__global__ void __launch_bounds__(SPH_THREADS, FLUID_FORCE_MIN_CTAS)
k_force_integrate(const DevParams p, const float2* __restrict__ pos_s, const float2* __restrict__ pred_s,
                  const float2* __restrict__ vel_s, const float* __restrict__ rho, const uint32_t* __restrict__ keys,
                  const uint32_t* __restrict__ lb, const uint32_t* __restrict__ quirk,
                  const float2* __restrict__ force_field, const uint32_t* __restrict__ frame_time,
                  const uint32_t* __restrict__ global_offset_dev /* multi-GPU band: the global sorted index of its first owned particle, on the device; else null */,
                  const uint32_t* __restrict__ wild_stamp /* null: always take the per-pair range tests */,
                  const TilePlan* __restrict__ plans,
                  float2* __restrict__ pos_out, float2* __restrict__ vel_out, float2* __restrict__ dbg_accel,
                  float2* __restrict__ dbg_abs,
                  uint32_t* __restrict__ next_keys /* null: do not speculate */, uint32_t* __restrict__ next_count,
                  const uint32_t* __restrict__ band_range /* multi-GPU band: {first, end} of the owned particles, on the device; else null */,
                  const BandSpec* __restrict__ band_spec /* multi-GPU band: run the next tick's key pass (band_common.cuh) into next_keys / next_count; else null */,
                  const uint4* __restrict__ nb_head, const uint2* __restrict__ nb_mask /* from k_density: span starts + acceptance masks */) {
    __shared__ __align__(128) float2 s_pred[SPH_CAP];
    __shared__ __align__(128) float2 s_vel[SPH_CAP];
    __shared__ __align__(128) float s_rho[SPH_CAP];
    __shared__ uint16_t s_list[FORCE_LIST * SPH_THREADS];
    __shared__ uint64_t s_bar;
    __shared__ SpanPlan s_plan;
    pdl_enter();
    // band mode: owned particles are [first, end) of arrays that also hold ghost rows on both sides (see k_density)
    // (BAND is a template parameter so that the single-GPU kernel is compiled exactly as it was without band support)
    const uint32_t r_begin = BAND ? __ldg(band_range) : p.own_begin;
    const uint32_t r_end = BAND ? __ldg(band_range + 1) : p.own_begin + p.n;
    uint32_t tile = blockIdx.x;
    const uint32_t t_first = (BOTH_ENDS && BAND) ? r_begin / SPH_THREADS : 0u;
    const uint32_t t_count = !BOTH_ENDS ? 0u : (BAND ? (r_end > r_begin ? (r_end + SPH_THREADS - 1u) / SPH_THREADS - t_first : 0u) : gridDim.x);
    if (BOTH_ENDS) {
        if (BAND && blockIdx.x >= t_count) return;  // (whole CTA)
        tile = tile_from_both_ends(blockIdx.x, t_first, t_count);
    }
    const uint32_t i0 = (BAND ? 0u : p.own_begin) + tile * SPH_THREADS;
    if (BAND && !BOTH_ENDS && (i0 + SPH_THREADS <= r_begin || i0 >= r_end)) return;  // (whole CTA)
    if (threadIdx.x == 0) {
        mbar_init(&s_bar, 1);
        mbar_fence_init();
        const void* const arrays[3] = {pred_s, vel_s, rho};
        void* const smem[3] = {s_pred, s_vel, s_rho};
        const uint32_t eb[3] = {8u, 8u, 4u};
        stage_from_plan<3>(plans + tile, arrays, smem, eb, &s_bar, &s_plan);
    }
    const uint32_t i = i0 + threadIdx.x;
    const bool valid = BAND ? (i >= r_begin && i < r_end) : (i < r_end);
    float2 pos = make_float2(0.f, 0.f), pred = pos, vel = pos;
    float dens = 1.0f;
    uint4 nh = make_uint4(0u, 0u, 0u, 0u);
    uint2 nm = make_uint2(0u, 0u);
    if (valid) {
        nh = __ldg(nb_head + i);
        nm = __ldg(nb_mask + i);
        pos = __ldg(pos_s + i);
        pred = __ldg(pred_s + i);
        vel = __ldg(vel_s + i);
        dens = __ldg(rho + i);
    }
    const uint32_t ft = __ldg(frame_time);
    const bool tame = wild_stamp != nullptr && __ldg(wild_stamp) != ft;  // common.cuh tame_component
    const uint32_t goff = p.global_offset + (global_offset_dev ? __ldg(global_offset_dev) : 0u);
    uint32_t rand_seed = (i - r_begin + goff) * 12u + ft * 69u;  // compute.wgsl:161 (global sorted index; i - r_begin = index in the ghost-free state arrays)
    const float pressure = fmul(p.pressure_constant, fsub(dens, p.rest_density));
    ForceAcc acc = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    __syncthreads();
    const bool staged = s_plan.staged != 0;
    // the density pass's masks serve a warp unless one of its particles has a span longer than 32 candidates
    if (staged && !__any_sync(0xffffffffu, valid && (nh.x & NB_FALLBACK))) {
        mbar_wait(&s_bar, 0);
        int base[3];
        const uint32_t g = (nh.x & NB_GROUP) ? 1u : 0u;
#pragma unroll
        for (int r = 0; r < 3; r++) base[r] = s_plan.base[g * 3 + r];
        force_walk_masked<DEBUG, EXACT>(p, smem_u32_pinned(s_pred), smem_u32_pinned(s_vel), smem_u32_pinned(s_rho), base, nh.x & NB_INDEX, nh.y,
                                        nh.z, nh.w, nm.x, nm.y, i, pred.x, pred.y, vel.x, vel.y, pressure, tame, rand_seed, acc);
    } else {
        uint32_t a[3] = {0, 0, 0}, b[3] = {0, 0, 0};
        uint32_t cy = 0;
        if (valid) cy = thread_spans(p, __ldg(keys + i), lb, __ldg(quirk), __ldg(quirk + 1), a, b);
        if (staged) {
            mbar_wait(&s_bar, 0);
            int base[3];
            thread_bases(s_plan, valid ? cy : s_plan.row_first, base);
            force_walk_staged<DEBUG, EXACT>(p, smem_u32_pinned(s_pred), smem_u32_pinned(s_vel), smem_u32_pinned(s_rho), smem_u32_pinned(s_list + threadIdx.x),
                                            base, a, b, i, pred.x, pred.y, vel.x, vel.y, pressure, tame, rand_seed, acc);
        } else {
            force_walk_global<DEBUG, EXACT>(p, pred_s, vel_s, rho, a, b, i, pred.x, pred.y, vel.x, vel.y, pressure, rand_seed, acc);
        }
    }
    float nx = 0.0f, ny = 0.0f, vx = 0.0f, vy = 0.0f;
    // BOTH_ENDS: the particle's index is derived again from the block index — held across the walk it would cost a register
    // of the 64 the kernel is capped at (index order needs none: blockIdx.x is free to re-read; the opaque copy keeps the
    // compiler from merging this with the computation at the top)
    uint32_t out_i = i - r_begin;
    if (BOTH_ENDS) {
        uint32_t b_again = blockIdx.x;
        asm volatile("" : "+r"(b_again));
        out_i = (BAND ? 0u : p.own_begin) + tile_from_both_ends(b_again, t_first, t_count) * SPH_THREADS + threadIdx.x - r_begin;
    }
    if (valid) {
        // compute.wgsl:90-96
        const float fvx = fmul(acc.vx, p.visc_coef), fvy = fmul(acc.vy, p.visc_coef);
        const float ax = fadd(acc.px, fvx), ay = fadd(acc.py, fvy);
        if (DEBUG) {
            dbg_accel[out_i] = make_float2(ax, ay);
            dbg_abs[out_i] = make_float2(acc.ax, acc.ay);
        }
        vx = fadd(vel.x, fmul(fdiv(ax, dens), p.dt));
        vy = fadd(vel.y, fmul(fdiv(ay, dens), p.dt));
        vx = fadd(vx, p.gravity_dt_x);
        vy = fadd(vy, p.gravity_dt_y);
        // compute.wgsl:99-108 mouse force
        if (p.mouse_state != 0) {
            const float dx = fsub(p.mouse_x, pred.x), dy = fsub(p.mouse_y, pred.y);
            const float dist = fsqrt(fadd(fmul(dx, dx), fmul(dy, dy)));
            if (dist <= p.mouse_radius) {
                const float dirx = fdiv(fdiv(dx, dist), dist), diry = fdiv(fdiv(dy, dist), dist);
                const float ratio = fdiv(dist, p.mouse_radius);
                vx = fadd(vx, fmul(fmul(fmul(dirx, p.mouse_power), p.mouse_state_f), ratio));
                vy = fadd(vy, fmul(fmul(fmul(diry, p.mouse_power), p.mouse_state_f), ratio));
            }
        }
        // compute.wgsl:113-116 NaN guard
        if (!(vx == vx && vy == vy)) { vx = 0.0f; vy = 0.0f; }
        // compute.wgsl:118-122 speed clamp
        const float speed = fsqrt(fadd(fmul(vx, vx), fmul(vy, vy)));
        if (speed > 500.0f) {
            vx = fmul(fdiv(vx, speed), 500.0f);
            vy = fmul(fdiv(vy, speed), 500.0f);
        }
        nx = fadd(pos.x, fmul(vx, p.dt));
        ny = fadd(pos.y, fmul(vy, p.dt));
        // compute.wgsl:127-140 force-field push-out (texel looked up at the PREDICTED position)
        {
            const float uvx = fadd(fmul(fdiv(pred.x, p.bounds_x), 1.0f), 0.5f);
            const float uvy = fadd(fmul(fdiv(pred.y, p.bounds_y), 1.0f), 0.5f);
            const uint32_t tx = f2u(fmul(uvx, p.tex_w)), ty = f2u(fmul(uvy, p.tex_h));
            uint32_t tidx = ty * p.tex_wu + tx;
            if (tidx >= p.tex_len) tidx = p.tex_len - 1u;  // [survey-pinned] out-of-range texel -> last texel
            const float2 f = __ldg(force_field + tidx);
            if (f.x != 0.0f || f.y != 0.0f) {
                const float fwx = fmul(f.x, p.p2w_x), fwy = fmul(f.y, p.p2w_y);
                const float len = fsqrt(fadd(fmul(f.x, f.x), fmul(f.y, f.y)));
                const float nnx = fdiv(f.x, len), nny = fdiv(f.y, len);
                nx = fadd(nx, fwx);
                ny = fadd(ny, fwy);
                const float vn = fadd(fmul(vx, nnx), fmul(vy, nny));
                vx = fsub(vx, fmul(fmul(p.one_minus_damping, vn), nnx));
                vy = fsub(vy, fmul(fmul(p.one_minus_damping, vn), nny));
            }
        }
        // compute.wgsl:143-153 walls
        if (fabsf(nx) > p.half_x) { nx = fmul(p.half_x, fsign(nx)); vx = fmul(vx, p.neg_damping); }
        if (fabsf(ny) > p.half_y) { ny = fmul(p.half_y, fsign(ny)); vy = fmul(vy, p.neg_damping); }
        pos_out[out_i] = make_float2(nx, ny);
        vel_out[out_i] = make_float2(vx, vy);
    }
    // Speculation for the next tick (single GPU, counting path): with the same delta its K1+K2 (compute.wgsl:8-42) see
    // exactly this position and velocity, so the key and the per-cell count are emitted here — same functions, same
    // bits as k_predict_key_count — and the next tick starts at the cell scan. The host drops the speculation when the
    // next tick's delta differs or the state was replaced (fluid_sim.cu enqueue_tick).
    if (BAND && band_spec) {
        // a band of the multi-GPU step: the same, plus the copies for the other ranks (k_band_key)
        const float2 q = predict_position(p, make_float2(nx, ny), make_float2(vx, vy));
        uint32_t row;
        const uint32_t k = band_key_and_count(p, band_spec->rows, q, rcp_refined(p.h), valid, band_spec->key_lo, next_count, row);
        if (valid) next_keys[out_i] = k;
        band_push_copies(band_spec->rows, band_spec->out, band_spec->dyn, valid, row, make_float2(nx, ny), make_float2(vx, vy), goff + out_i);
    } else if (next_keys) {
        uint32_t k = 0;
        if (valid) {
            const float2 q = predict_position(p, make_float2(nx, ny), make_float2(vx, vy));
            k = p.key_fast_div ? cell_key_shared(p, q, rcp_refined(p.h)) : cell_key(p, q);
            next_keys[out_i] = k;
        }
        atomic_count_run(next_count, min(k, p.num_cells - 1u), valid);
    }
}

}  // namespace fluid
