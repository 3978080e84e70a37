// wgsl_ref_driver.cpp — dispatch loops around the reference's shaders as transpiled by oracle/transpile_wgsl.py
// (TEST INFRASTRUCTURE; built into oracle/_ref/libwgsl_ref.so, which is git-ignored).
//
// The shader bodies come from oracle/_ref/wgsl_*.inc, GENERATED from /root/reference/*.wgsl at build time and never
// committed. This file only supplies what the reference's HOST does around them, restated from the Rust:
//   * one invocation per global id over ceil(n / workgroup_size) workgroups (src/simulation.rs:463-464, :512-537);
//   * the bitonic schedule of sort.wgsl's Settings (src/simulation.rs:323-347) and its dispatch size (:518-524);
//   * the pass order K1 K2 K3.. K4 K5 K6 (:512-537).
// Invocations of one dispatch run in parallel (OpenMP) — every pass except move_particle only writes its own
// element / disjoint pairs; move_particle's stores go to `out_particles` (the Jacobi pin, see transpile_wgsl.py).
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "wgsl_prelude.h"

#include "wgsl_compute.inc"
#include "wgsl_jfa.inc"
#include "wgsl_sort.inc"

static_assert(sizeof(wgsl_compute::ParticleInstance) == 32, "ParticleInstance must be the reference's 32-byte record");
static_assert(sizeof(wgsl_compute::Uniforms) == 120, "Uniforms must be the reference's 120-byte SimulationUniform");
static_assert(sizeof(wgsl_sort::ParticleInstance) == 32, "sort.wgsl ParticleInstance");
static_assert(sizeof(wgsl_sort::Settings) == 16 && sizeof(wgsl_jfa::Settings) == 12, "Settings layouts");

namespace {

template <class F> void dispatch(uint32_t invocations, F&& f) {
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < (int64_t)invocations; i++) f((uint32_t)i);
}
uint32_t groups(uint32_t n, uint32_t wg) { return (n + wg - 1) / wg; }

void bind(const void* uniform, void* particles, uint32_t n, uint32_t* start_indices, uint32_t cells, float* texture, uint32_t texels) {
    using namespace wgsl_compute;
    memcpy(&u, uniform, sizeof u);
    in_particles.data = reinterpret_cast<ParticleInstance*>(particles);
    in_particles.len = n;
    out_particles = in_particles;
    wgsl_compute::start_indices.data = start_indices;
    wgsl_compute::start_indices.len = cells;
    wgsl_compute::texture.data = reinterpret_cast<wgsl::vec2<float>*>(texture);
    wgsl_compute::texture.len = texels;
}

}  // namespace

extern "C" {

int wgslref_abi_version(void) { return 1; }
void wgslref_set_num_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}
int wgslref_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

// the bitonic network of src/simulation.rs:323-347 over sort.wgsl's `sort` entry point (@workgroup_size(128),
// ceil(next_pow2(n) / 2 / 128) workgroups per step, src/simulation.rs:518-524)
void wgslref_sort(void* particles, uint32_t n) {
    using namespace wgsl_sort;
    values.data = reinterpret_cast<ParticleInstance*>(particles);
    values.len = n;
    uint32_t pow2 = 1;
    while (pow2 < n) pow2 <<= 1;
    const uint32_t num_pairs = pow2 / 2;
    uint32_t num_stages = 0;
    while ((1u << num_stages) < num_pairs * 2) num_stages++;  // (num_pairs * 2).ilog2()
    for (uint32_t stage = 0; stage < num_stages; stage++)
        for (uint32_t step = 0; step <= stage; step++) {
            const uint32_t gw = 1u << (stage - step);
            settings.group_width = gw;
            settings.group_height = 2 * gw - 1;
            settings.step_index = step;
            settings.num_values = n;
            dispatch(groups(num_pairs, 128) * 128, [&](uint32_t i) { sort(wgsl::vec3<uint32_t>{i, 0u, 0u}); });
        }
}

// one pass each (names = the WGSL entry points bound at src/simulation.rs:315-319)
#define WGSLREF_PASS(NAME)                                                                                                     \
    void wgslref_##NAME(const void* uniform, void* particles, uint32_t n, uint32_t* table, uint32_t cells, float* tex, uint32_t texels) { \
        bind(uniform, particles, n, table, cells, tex, texels);                                                                \
        dispatch(groups(n, 256) * 256, [&](uint32_t i) { wgsl_compute::NAME(wgsl::vec3<uint32_t>{i, 0u, 0u}); });           \
    }
WGSLREF_PASS(predict_next_position)
WGSLREF_PASS(create_spatial_lookup)
WGSLREF_PASS(compute_start_indices)
WGSLREF_PASS(calculate_density)
#undef WGSLREF_PASS
// move_particle with the Jacobi pin: reads see `particles`, stores land in `out` (may not alias)
void wgslref_move_particle(const void* uniform, void* particles, void* out, uint32_t n, uint32_t* table, uint32_t cells, float* tex,
                           uint32_t texels) {
    bind(uniform, particles, n, table, cells, tex, texels);
    wgsl_compute::out_particles.data = reinterpret_cast<wgsl_compute::ParticleInstance*>(out);
    dispatch(groups(n, 256) * 256, [&](uint32_t i) { wgsl_compute::move_particle(wgsl::vec3<uint32_t>{i, 0u, 0u}); });
}

// FluidSimulation::tick (src/simulation.rs:459-539) for an already derived uniform: K1, K2, K3 x S(S+1)/2, K4, K5, K6.
// stop_after: 0 = whole tick, 4 = after compute_start_indices (BASELINE config 2). ms (optional): 6 phase times.
void wgslref_tick(const void* uniform, void* particles, uint32_t n, uint32_t* table, uint32_t cells, float* tex, uint32_t texels,
                  int stop_after, float* ms) {
    auto now = [] {
#ifdef _OPENMP
        return omp_get_wtime();
#else
        return 0.0;
#endif
    };
    double t[7];
    t[0] = now();
    wgslref_predict_next_position(uniform, particles, n, table, cells, tex, texels); t[1] = now();
    wgslref_create_spatial_lookup(uniform, particles, n, table, cells, tex, texels); t[2] = now();
    wgslref_sort(particles, n); t[3] = now();
    wgslref_compute_start_indices(uniform, particles, n, table, cells, tex, texels); t[4] = now();
    t[5] = t[6] = t[4];
    if (stop_after != 4) {
        wgslref_calculate_density(uniform, particles, n, table, cells, tex, texels); t[5] = now();
        std::vector<wgsl_compute::ParticleInstance> out(n);
        wgslref_move_particle(uniform, particles, out.data(), n, table, cells, tex, texels);
        memcpy(particles, out.data(), (size_t)n * 32);
        t[6] = now();
    }
    if (ms) for (int k = 0; k < 6; k++) ms[k] = (float)((t[k + 1] - t[k]) * 1e3);
}

// one jump-flood pass (shaders/jump_flood.wgsl `floodfill`): seeds_in / seeds_out are rg32uint images
void wgslref_jfa_pass(uint32_t* seeds_in, uint32_t w, uint32_t h, uint32_t jump, uint32_t* seeds_out) {
    using namespace wgsl_jfa;
    settings.width = w; settings.height = h; settings.jump = jump;
    in_values.data = seeds_in; in_values.w = w; in_values.h = h;
    out_values.data = seeds_out; out_values.w = w; out_values.h = h;
#pragma omp parallel for schedule(static)
    for (int64_t y = 0; y < (int64_t)h; y++)
        for (uint32_t x = 0; x < w; x++) floodfill(wgsl::vec3<uint32_t>{x, (uint32_t)y, 0u});
}

}  // extern "C"
