// fluid_oracle.cpp — CPU ORACLE (TEST INFRASTRUCTURE, NOT PRODUCT CODE).
//
// A plain C++ restatement of the per-frame fluid step of
// rookieCookies/gpu-fluid-simulation, function for function, each citing the
// reference file:line it follows. Only tests/, __graft_entry__.smoke() and
// bench.py's cpu_baseline / --impl reference legs may load this library, and only
// as the checker or the timed CPU baseline. The product (libfluid_b200.so) never
// links, loads or calls it.
//
// PARITY: the reference has no tests, golden vectors or fixtures and cannot be built or run in this
// environment (no rustc/cargo, no graphics API for the WGSL), so no reference-held vector pins this file. It is
// pinned by (a) a second, MECHANICAL derivation: oracle/transpile_wgsl.py rewrites the reference's WGSL text into C++
// (oracle/_ref/libwgsl_ref.so), and tests/test_wgsl_ref.py compares every pass, whole ticks, the bitonic network
// (incl. its tie order) and the jump-flood pass bit for bit with this restatement; (b) the restatement rules of
// IEEE-754 binary32, round-to-nearest-even, no FMA contraction (build with -ffp-contract=off),
// correctly rounded / and sqrt, pow(h,8) evaluated with powf, out-of-range cells empty, out-of-range texel index
// clamped to the last texel, Jacobi reads in move_particle; and (c) the self-checks in tests/test_oracle.py (bitonic
// keys == sorted keys, 7x7 vs 3x3 density walk, fp64 brute force, default-scene invariants, JFA vs brute force).
// What stays restated by hand only: the Rust host parts (uniform derivation, init lattice, bitonic schedule,
// generate_smooth_gradient_field) and the two render shaders (render_oracle.cpp).
//
// Build: see oracle/Makefile (g++ -O2 -ffp-contract=off -fopenmp -shared -fPIC).

#include <cmath>
#include <cstdint>
#include <cstring>
#include <cstdlib>
#include <vector>
#include <algorithm>
#include <chrono>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "../include/fluid_b200.h"

typedef FluidParticle P;
typedef FluidSimUniform U;

namespace {

// funcs.wgsl:54-55
const float PI_F = 3.14159265359f;
const float EPSILON_F = 1.19209290e-07f;

struct V2 { float x, y; };
inline V2 v2(float x, float y) { V2 r = {x, y}; return r; }

// Rust f32::powi -> compiler-rt __powisf2 (square-and-multiply), src/simulation.rs:486-490
inline float powi_f32(float a, int b) {
    float r = 1.0f;
    for (;;) {
        if (b & 1) r *= a;
        b /= 2;
        if (b == 0) break;
        a *= a;
    }
    return r;
}

inline float sign_f(float x) { return x > 0.0f ? 1.0f : (x < 0.0f ? -1.0f : 0.0f); }  // WGSL sign()

// WGSL f32 -> u32 conversion (value conversion: truncate toward zero, clamp to range, NaN -> 0)
inline uint32_t f2u(float f) {
    if (!(f > 0.0f)) return 0u;
    if (f >= 4294967296.0f) return 4294967295u;
    return (uint32_t)f;
}

// funcs.wgsl:212-214  xy_of_point
inline void xy_of_point(const U& u, float px, float py, uint32_t& cx, uint32_t& cy) {
    float hx = u.bounds[0] * 0.5f, hy = u.bounds[1] * 0.5f;
    float fx = floorf((px + hx) / u.smoothing_radius);
    float fy = floorf((py + hy) / u.smoothing_radius);
    cx = f2u(fx) + 1u;
    cy = f2u(fy) + 1u;
}
// funcs.wgsl:216-218  grid_pos_to_id  (u32 wrapping arithmetic)
inline uint32_t grid_pos_to_id(const U& u, uint32_t x, uint32_t y) { return y * u.grid_w + x; }

// [survey-pinned] a cell with x outside [0,grid_w) or y outside [0,grid_h) is empty
inline bool cell_in_range(const U& u, int64_t x, int64_t y) {
    return x >= 0 && y >= 0 && x < (int64_t)u.grid_w && y < (int64_t)u.grid_h;
}

// funcs.wgsl:72-78  poly6_kernel; `norm` = 4/(PI*pow(h,8)) hoisted (h is uniform)
inline float poly6_kernel(float h, float r2, float norm) {
    float h2 = h * h;
    if (r2 > h2) return 0.0f;
    float diff = h2 - r2;
    return norm * diff * diff * diff;
}
// funcs.wgsl:101-109
inline float spiky_kernel_derivative(const U& u, float h, float r) {
    if (r <= h) {
        float v = h - r;
        return -v * u.spiky_kernel_derivative;
    }
    return 0.0f;
}
// funcs.wgsl:112-123
inline float viscosity_kernel(const U& u, float h, float r) {
    if (r <= h) {
        float c = u.viscosity_kernel;
        if (r == 0.0f) return c;
        return c * ((-(r * r * r) / (2.0f * h * h * h)) + ((r * r) / (h * h)) + (h / (2.0f * r)) - 1.0f);
    }
    return 0.0f;
}
// funcs.wgsl:129-149
inline uint32_t xorshift32(uint32_t& s) {
    uint32_t x = s;
    x ^= x << 13; x ^= x >> 17; x ^= x << 5;
    s = x;
    return x;
}
inline float rand_f32(uint32_t& s) { return (float)xorshift32(s) / 4294967296.0f; }
// funcs.wgsl:152-154
inline float calculate_pressure(const U& u, float d) { return u.pressure_constant * (d - u.rest_density); }

}  // namespace

extern "C" {

int oracle_abi_version(void) { return 1; }
int oracle_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
void oracle_set_num_threads(int n) {
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

// src/simulation.rs:140-141, 466-467
void oracle_grid_dims(const FluidSimSettings* s, uint32_t* gw, uint32_t* gh) {
    *gw = (uint32_t)(size_t)ceilf(s->size[0] / s->smoothing_radius) + 2u;
    *gh = (uint32_t)(size_t)ceilf(s->size[1] / s->smoothing_radius) + 2u;
}

// src/simulation.rs:147-163  init lattice
void oracle_init_lattice(const FluidSimSettings* s, P* out) {
    uint32_t n = s->particle_count;
    float ppr = sqrtf((float)n);
    float ppc = ((float)n - 1.0f) / ppr + 1.0f;
    size_t ppr_u = (size_t)ppr;
    for (uint32_t i = 0; i < n; i++) {
        size_t xi = (size_t)i % ppr_u;
        float x = ((float)xi - ppr * 0.5f + 0.5f) * s->particle_spacing;
        float y = (floorf((float)i / ppr) - ppc * 0.5f + 0.5f) * s->particle_spacing;
        P p; memset(&p, 0, sizeof p);
        p.position[0] = x; p.position[1] = y;
        p.predicted_position[0] = x; p.predicted_position[1] = y;
        out[i] = p;
    }
}

// src/simulation.rs:470-497  uniform derivation; `tick` is the already-incremented counter (:460,:474)
void oracle_make_uniform(const FluidSimSettings* s, const FluidTickSettings* t, uint32_t tick, U* o) {
    const float PI_RS = 3.14159274101257324219f;  // std::f32::consts::PI
    float h = s->smoothing_radius;
    memset(o, 0, sizeof *o);
    o->delta = t->delta;
    o->particle_count = s->particle_count;
    o->sqr_radius = h * h;
    o->frame_time = tick;
    o->gravity[0] = t->gravity[0]; o->gravity[1] = t->gravity[1];
    o->bounds[0] = s->size[0]; o->bounds[1] = s->size[1];
    o->mouse_pos[0] = t->mouse_pos[0]; o->mouse_pos[1] = t->mouse_pos[1];
    o->smoothing_radius = h;
    o->particle_mass = t->mass;
    o->pressure_constant = t->pressure_constant;
    o->rest_density = t->rest_density;
    o->damping_factor = t->damping_factor;
    o->viscosity_coefficient = t->viscosity_coefficient;
    o->surface_tension_treshold = t->surface_tension_treshold;
    o->surface_tension_coefficient = t->surface_tension_coefficient;
    o->poly6_kernel_volume = 4.0f / (PI_RS * powi_f32(h, 8));
    o->poly6_kernel_derivative = 24.0f / (PI_RS * powi_f32(h, 8));
    o->poly6_kernel_laplacian = 8.0f / (PI_RS * powi_f32(h, 8));
    o->spiky_kernel_derivative = 12.0f / (powi_f32(h, 4) * PI_RS);
    o->viscosity_kernel = 15.0f / (2.0f * PI_RS * powi_f32(h, 3));
    o->mouse_state = t->mouse_state;
    o->mouse_force_radius = t->mouse_force_radius;
    o->mouse_force_power = t->mouse_force_power;
    oracle_grid_dims(s, &o->grid_w, &o->grid_h);
    o->texture_size[0] = (float)s->texture_size[0];
    o->texture_size[1] = (float)s->texture_size[1];
}

// compute.wgsl:8-30  predict_next_position
void oracle_predict_next_position(const U* u, P* p) {
    const uint32_t n = u->particle_count;
    const float hx = u->bounds[0] * 0.5f, hy = u->bounds[1] * 0.5f;
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < (int64_t)n; i++) {
        float qx = p[i].position[0] + p[i].velocity[0] * u->delta;
        float qy = p[i].position[1] + p[i].velocity[1] * u->delta;
        if (fabsf(qx) > hx) qx = hx * sign_f(qx);
        if (fabsf(qy) > hy) qy = hy * sign_f(qy);
        p[i].predicted_position[0] = qx;
        p[i].predicted_position[1] = qy;
    }
}

// compute.wgsl:33-42  create_spatial_lookup  (cell_of_point funcs.wgsl:206-209)
void oracle_create_spatial_lookup(const U* u, P* p) {
    const uint32_t n = u->particle_count;
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < (int64_t)n; i++) {
        uint32_t cx, cy;
        xy_of_point(*u, p[i].predicted_position[0], p[i].predicted_position[1], cx, cy);
        p[i].grid = grid_pos_to_id(*u, cx, cy);
    }
}

// sort.wgsl:27-51 with the schedule of src/simulation.rs:323-347: the faithful bitonic network
void oracle_sort_bitonic(P* p, uint32_t n) {
    if (n < 2) return;
    uint32_t np2 = 1; while (np2 < n) np2 <<= 1;
    uint32_t num_pairs = np2 / 2;
    uint32_t num_stages = 0; while ((1u << num_stages) < num_pairs * 2) num_stages++;
    uint32_t groups = (num_pairs + 127) / 128;
    int64_t invocations = (int64_t)groups * 128;
    for (uint32_t stage = 0; stage < num_stages; stage++) {
        for (uint32_t step = 0; step <= stage; step++) {
            uint32_t gw = 1u << (stage - step);
            uint32_t gh = 2 * gw - 1;
#pragma omp parallel for schedule(static)
            for (int64_t i = 0; i < invocations; i++) {
                uint32_t ii = (uint32_t)i;
                uint32_t hh = ii & (gw - 1u);
                uint32_t lo = hh + (gh + 1u) * (ii / gw);
                uint32_t hi = lo + (step == 0 ? gh - 2u * hh : (gh + 1u) / 2u);
                if (hi >= n) continue;
                if (p[lo].grid > p[hi].grid) { P t = p[lo]; p[lo] = p[hi]; p[hi] = t; }
            }
        }
    }
}

// Stable ascending sort by `grid` (the tie order an LSD radix sort produces). perm_out (optional)
// receives sorted index -> index before the sort.
void oracle_sort_stable(P* p, uint32_t n, uint32_t* perm_out) {
    std::vector<uint32_t> idx(n);
    for (uint32_t i = 0; i < n; i++) idx[i] = i;
    std::stable_sort(idx.begin(), idx.end(), [&](uint32_t a, uint32_t b) { return p[a].grid < p[b].grid; });
    std::vector<P> tmp(p, p + n);
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < (int64_t)n; i++) p[i] = tmp[idx[i]];
    if (perm_out) memcpy(perm_out, idx.data(), (size_t)n * 4);
}

// Fast stable counting sort by `grid` for the timed CPU baseline ("fair CPU variant").
// Same output as oracle_sort_stable.
void oracle_sort_counting(P* p, uint32_t n, uint32_t num_cells) {
    std::vector<uint32_t> cnt((size_t)num_cells + 1, 0);
    for (uint32_t i = 0; i < n; i++) { uint32_t k = p[i].grid; if (k >= num_cells) k = num_cells - 1; cnt[k + 1]++; }
    for (uint32_t c = 0; c < num_cells; c++) cnt[c + 1] += cnt[c];
    std::vector<P> tmp(p, p + n);
    for (uint32_t i = 0; i < n; i++) { uint32_t k = tmp[i].grid; if (k >= num_cells) k = num_cells - 1; p[cnt[k]++] = tmp[i]; }
}

// compute.wgsl:45-56  compute_start_indices (i == 0 returns early; table never cleared)
void oracle_compute_start_indices(const U* u, const P* p, uint32_t* start_indices) {
    const uint32_t n = u->particle_count;
    const uint64_t cells = (uint64_t)u->grid_w * u->grid_h;
#pragma omp parallel for schedule(static)
    for (int64_t i = 1; i < (int64_t)n; i++) {
        if (p[i].grid != p[i - 1].grid && p[i].grid < cells) start_indices[p[i].grid] = (uint32_t)i;
    }
}

// funcs.wgsl:157-203  calculate_density_at_point; walk_radius = 3 is the reference's 7x7 walk
static float density_at_point(const U& u, const P* p, const uint32_t* start_indices, float px, float py,
                              int walk_radius, float poly6_norm) {
    float density = 0.0f;
    uint32_t ucx, ucy;
    xy_of_point(u, px, py, ucx, ucy);
    int32_t cx = (int32_t)ucx, cy = (int32_t)ucy;
    for (int oy = -walk_radius; oy <= walk_radius; oy++) {
        for (int ox = -walk_radius; ox <= walk_radius; ox++) {
            int64_t x = (int64_t)cx + ox, y = (int64_t)cy + oy;
            if (!cell_in_range(u, x, y)) continue;
            uint32_t id = grid_pos_to_id(u, (uint32_t)x, (uint32_t)y);
            uint32_t s = start_indices[id];
            while (true) {
                if (s >= u.particle_count) break;
                const P& nb = p[s];
                if (nb.grid != id) break;
                s += 1;
                float dx = nb.predicted_position[0] - px, dy = nb.predicted_position[1] - py;
                float r2 = dx * dx + dy * dy;
                float k = poly6_kernel(u.smoothing_radius, r2, poly6_norm);
                float mul = 1.0f;
                density += u.particle_mass * k * mul;
            }
        }
    }
    return fmaxf(density, EPSILON_F);
}

float oracle_poly6_norm(float h) { return 4.0f / (PI_F * powf(h, 8.0f)); }

// compute.wgsl:59-74  calculate_density
void oracle_calculate_density(const U* u, P* p, const uint32_t* start_indices, int walk_radius) {
    const uint32_t n = u->particle_count;
    const float norm = oracle_poly6_norm(u->smoothing_radius);
#pragma omp parallel for schedule(dynamic, 1024)
    for (int64_t i = 0; i < (int64_t)n; i++) {
        float d = density_at_point(*u, p, start_indices, p[i].predicted_position[0], p[i].predicted_position[1],
                                   walk_radius, norm);
        p[i].density = fmaxf(d, 0.1f);
    }
}

// compute.wgsl:160-235  calculate_pressure_force. abs_terms (optional) accumulates sum |term|.
// index_offset: added to the particle's array index to form the id the shader seeds its RNG with. 0 for the
// reference's single array; in the slab tests the array is [halo | band | halo] of a larger sorted array.
static V2 pressure_force(const U& u, const P* p, const uint32_t* start_indices, uint32_t pid, V2* abs_terms,
                         uint32_t index_offset = 0) {
    uint32_t rand_seed = (pid + index_offset) * 12u + u.frame_time * 69u;
    const P& me = p[pid];
    float pressure = calculate_pressure(u, me.density);
    float px = me.predicted_position[0], py = me.predicted_position[1];
    V2 f = v2(0.0f, 0.0f);
    uint32_t ucx, ucy;
    xy_of_point(u, px, py, ucx, ucy);
    int32_t cx = (int32_t)ucx, cy = (int32_t)ucy;
    for (int oy = -1; oy <= 1; oy++) {
        for (int ox = -1; ox <= 1; ox++) {
            int64_t x = (int64_t)cx + ox, y = (int64_t)cy + oy;
            if (!cell_in_range(u, x, y)) continue;
            uint32_t id = grid_pos_to_id(u, (uint32_t)x, (uint32_t)y);
            uint32_t s = start_indices[id];
            while (true) {
                if (s >= u.particle_count) break;
                const P& nb = p[s];
                if (nb.grid != id) break;
                uint32_t i = s;
                s += 1;
                if (i == pid) continue;
                float dx = nb.predicted_position[0] - px, dy = nb.predicted_position[1] - py;
                float sq = dx * dx + dy * dy;
                if (sq > u.sqr_radius) continue;
                float dst = sqrtf(sq);
                float dirx, diry;
                if (dst == 0.0f) {
                    float rx = rand_f32(rand_seed);
                    float ry = rand_f32(rand_seed);
                    float len = sqrtf(rx * rx + ry * ry);
                    dirx = rx / len; diry = ry / len;
                } else {
                    dirx = dx / dst; diry = dy / dst;
                }
                float nd = nb.density;
                float np = calculate_pressure(u, nb.density);
                float k = spiky_kernel_derivative(u, u.smoothing_radius, dst);
                float shared = (pressure + np) * 0.5f;
                float tx = dirx * k * shared / nd;
                float ty = diry * k * shared / nd;
                f.x += tx; f.y += ty;
                if (abs_terms) { abs_terms->x += fabsf(tx); abs_terms->y += fabsf(ty); }
            }
        }
    }
    return f;
}

// compute.wgsl:238-299  calculate_viscosity_force
static V2 viscosity_force(const U& u, const P* p, const uint32_t* start_indices, uint32_t pid, V2* abs_terms) {
    const P& me = p[pid];
    float px = me.predicted_position[0], py = me.predicted_position[1];
    V2 f = v2(0.0f, 0.0f);
    uint32_t ucx, ucy;
    xy_of_point(u, px, py, ucx, ucy);
    int32_t cx = (int32_t)ucx, cy = (int32_t)ucy;
    for (int oy = -1; oy <= 1; oy++) {
        for (int ox = -1; ox <= 1; ox++) {
            int64_t x = (int64_t)cx + ox, y = (int64_t)cy + oy;
            if (!cell_in_range(u, x, y)) continue;
            uint32_t id = grid_pos_to_id(u, (uint32_t)x, (uint32_t)y);
            uint32_t s = start_indices[id];
            while (true) {
                if (s >= u.particle_count) break;
                const P& nb = p[s];
                if (nb.grid != id) break;
                uint32_t i = s;
                s += 1;
                if (i == pid) continue;
                float dx = nb.predicted_position[0] - px, dy = nb.predicted_position[1] - py;
                float sq = dx * dx + dy * dy;
                if (sq > u.sqr_radius) continue;
                float dst = sqrtf(sq);
                float nd = nb.density;
                float k = viscosity_kernel(u, u.smoothing_radius, dst);
                float tx = (nb.velocity[0] - me.velocity[0]) / nd * k;
                float ty = (nb.velocity[1] - me.velocity[1]) / nd * k;
                f.x += tx; f.y += ty;
                if (abs_terms) {
                    abs_terms->x += fabsf(tx) * u.viscosity_coefficient;
                    abs_terms->y += fabsf(ty) * u.viscosity_coefficient;
                }
            }
        }
    }
    f.x = f.x * u.viscosity_coefficient;
    f.y = f.y * u.viscosity_coefficient;
    return f;
}

// compute.wgsl:79-157  move_particle with JACOBI semantics [survey-pinned, SURVEY.md §3.3]:
// every neighbour read sees the state as it was when the pass started (`in`), results go to `out`.
// accel_out / abs_out (optional, 2 floats per particle): pressure+viscosity sum and sum of |terms|.
static void move_particle_range(const U* up, const P* in, P* out, const uint32_t* start_indices, const float* texture,
                                float* accel_out, float* abs_out, uint32_t begin, uint32_t end, uint32_t index_offset);

void oracle_move_particle(const U* up, const P* in, P* out, const uint32_t* start_indices, const float* texture,
                          float* accel_out, float* abs_out) {
    move_particle_range(up, in, out, start_indices, texture, accel_out, abs_out, 0, up->particle_count, 0);
}
// slab tests: move only particles [begin, end) of the array (the rest are halo copies owned by other ranks);
// RNG ids are array index + index_offset (wrapping u32), i.e. the particle's index in the global sorted array.
void oracle_move_particle_range(const U* up, const P* in, P* out, const uint32_t* start_indices, const float* texture,
                                uint32_t begin, uint32_t end, uint32_t index_offset) {
    move_particle_range(up, in, out, start_indices, texture, nullptr, nullptr, begin, end, index_offset);
}

static void move_particle_range(const U* up, const P* in, P* out, const uint32_t* start_indices, const float* texture,
                                float* accel_out, float* abs_out, uint32_t begin, uint32_t end, uint32_t index_offset) {
    const U& u = *up;
    const uint64_t tex_len = (uint64_t)f2u(u.texture_size[0]) * f2u(u.texture_size[1]);
#pragma omp parallel for schedule(dynamic, 1024)
    for (int64_t ii = (int64_t)begin; ii < (int64_t)end; ii++) {
        uint32_t id = (uint32_t)ii;
        P particle = in[id];
        V2 abs_terms = v2(0.0f, 0.0f);
        V2 pr = pressure_force(u, in, start_indices, id, abs_out ? &abs_terms : nullptr, index_offset);
        V2 vi = viscosity_force(u, in, start_indices, id, abs_out ? &abs_terms : nullptr);
        float ax = pr.x + vi.x, ay = pr.y + vi.y;
        if (accel_out) { accel_out[2 * ii] = ax; accel_out[2 * ii + 1] = ay; }
        if (abs_out) { abs_out[2 * ii] = abs_terms.x; abs_out[2 * ii + 1] = abs_terms.y; }

        float vx = particle.velocity[0], vy = particle.velocity[1];
        vx += (ax / particle.density) * u.delta;
        vy += (ay / particle.density) * u.delta;
        vx += u.gravity[0] * u.delta;
        vy += u.gravity[1] * u.delta;

        // compute.wgsl:99-108 mouse force
        if (u.mouse_state != 0) {
            float dx = u.mouse_pos[0] - particle.predicted_position[0];
            float dy = u.mouse_pos[1] - particle.predicted_position[1];
            float dist = sqrtf(dx * dx + dy * dy);
            if (dist <= u.mouse_force_radius) {
                float dirx = dx / dist / dist, diry = dy / dist / dist;
                float ratio = dist / u.mouse_force_radius;
                vx += dirx * u.mouse_force_power * (float)u.mouse_state * ratio;
                vy += diry * u.mouse_force_power * (float)u.mouse_state * ratio;
            }
        }
        // compute.wgsl:113-116 NaN guard
        if (!(vx == vx && vy == vy)) { vx = 0.0f; vy = 0.0f; }
        // compute.wgsl:118-122 speed clamp
        const float max_speed = 500.0f;
        float speed = sqrtf(vx * vx + vy * vy);
        if (speed > max_speed) { vx = (vx / speed) * max_speed; vy = (vy / speed) * max_speed; }

        float posx = particle.position[0] + vx * u.delta;
        float posy = particle.position[1] + vy * u.delta;

        // compute.wgsl:127-140 force-field push-out
        float uvx = (particle.predicted_position[0] / u.bounds[0] * 1.0f) + 0.5f;
        float uvy = (particle.predicted_position[1] / u.bounds[1] * 1.0f) + 0.5f;
        uint32_t tx = f2u(uvx * u.texture_size[0]);
        uint32_t ty = f2u(uvy * u.texture_size[1]);
        uint32_t tidx = ty * f2u(u.texture_size[0]) + tx;
        if ((uint64_t)tidx >= tex_len) tidx = (uint32_t)(tex_len - 1);  // [survey-pinned] OOB -> last texel
        float fx = texture[2 * (size_t)tidx], fy = texture[2 * (size_t)tidx + 1];
        float p2wx = (u.bounds[0] * 2.0f) / u.texture_size[0];
        float p2wy = (u.bounds[1] * 2.0f) / u.texture_size[1];
        float fwx = fx * p2wx, fwy = fy * p2wy;
        if (fx != 0.0f || fy != 0.0f) {
            float len = sqrtf(fx * fx + fy * fy);
            float nx = fx / len, ny = fy / len;
            posx += fwx; posy += fwy;
            float vn = vx * nx + vy * ny;
            vx -= (1.0f - u.damping_factor) * vn * nx;
            vy -= (1.0f - u.damping_factor) * vn * ny;
        }

        // compute.wgsl:143-153 walls
        float hx = u.bounds[0] * 0.5f, hy = u.bounds[1] * 0.5f;
        if (fabsf(posx) > hx) { posx = hx * sign_f(posx); vx *= -1.0f * u.damping_factor; }
        if (fabsf(posy) > hy) { posy = hy * sign_f(posy); vy *= -1.0f * u.damping_factor; }

        particle.position[0] = posx; particle.position[1] = posy;
        particle.velocity[0] = vx; particle.velocity[1] = vy;
        out[id] = particle;
    }
}

// src/simulation.rs:459-539  one tick: K1 -> K2 -> sort -> K4 -> K5 -> K6.
// sort_mode: 0 = stable (LSD-radix tie order), 1 = faithful bitonic network, 2 = counting (fast, == stable).
// stop_after: 0 = full tick, 1 = stop after compute_start_indices (BASELINE config 2).
// phase_ms (optional, 6 floats): predict, key, sort, cell-start, density, move.
void oracle_tick(const U* u, P* p, uint32_t* start_indices, const float* texture, int sort_mode, int walk_radius,
                 int stop_after, float* phase_ms) {
    using clk = std::chrono::steady_clock;
    auto t0 = clk::now();
    auto lap = [&](int k) {
        auto t1 = clk::now();
        if (phase_ms) phase_ms[k] = std::chrono::duration<float, std::milli>(t1 - t0).count();
        t0 = t1;
    };
    const uint32_t n = u->particle_count;
    oracle_predict_next_position(u, p); lap(0);
    oracle_create_spatial_lookup(u, p); lap(1);
    if (sort_mode == 1) oracle_sort_bitonic(p, n);
    else if (sort_mode == 2) oracle_sort_counting(p, n, u->grid_w * u->grid_h);
    else oracle_sort_stable(p, n, nullptr);
    lap(2);
    oracle_compute_start_indices(u, p, start_indices); lap(3);
    if (stop_after == 1) { if (phase_ms) { phase_ms[4] = 0; phase_ms[5] = 0; } return; }
    oracle_calculate_density(u, p, start_indices, walk_radius); lap(4);
    std::vector<P> snapshot(p, p + n);
    oracle_move_particle(u, snapshot.data(), p, start_indices, texture, nullptr, nullptr); lap(5);
}

// ---------------------------------------------------------------------------------------------
// Jump flood (shaders/jump_flood.wgsl:17-65). Images are rg32uint: 2 u32 per pixel.
// ---------------------------------------------------------------------------------------------

// shaders/jump_flood.wgsl:61-65
static inline int32_t sq_dist(int32_t ax, int32_t ay, int32_t bx, int32_t by) {
    int32_t dx = ax - bx, dy = ay - by;
    return dx * dx + dy * dy;
}

// shaders/jump_flood.wgsl:17-58  one `floodfill` pass
void oracle_jfa_pass(const uint32_t* in, uint32_t w, uint32_t h, uint32_t jump, uint32_t* out) {
    const int32_t j = (int32_t)jump;
    const int32_t offs[8][2] = {{-j, -j}, {0, -j}, {j, -j}, {-j, 0}, {j, 0}, {-j, j}, {0, j}, {j, j}};
#pragma omp parallel for schedule(static)
    for (int64_t yy = 0; yy < (int64_t)h; yy++) {
        for (uint32_t xx = 0; xx < w; xx++) {
            int32_t x = (int32_t)xx, y = (int32_t)yy;
            size_t pi = ((size_t)yy * w + xx) * 2;
            uint32_t bx = in[pi], by = in[pi + 1];
            int32_t best_d = sq_dist(x, y, (int32_t)bx, (int32_t)by);
            for (int k = 0; k < 8; k++) {
                int32_t nx = x + offs[k][0], ny = y + offs[k][1];
                if (nx < 0 || ny < 0 || nx >= (int32_t)w || ny >= (int32_t)h) continue;
                size_t ni = ((size_t)ny * w + (size_t)nx) * 2;
                uint32_t sx = in[ni], sy = in[ni + 1];
                if (sx > w || sy > w) continue;  // `width` used for both, `>` not `>=` (jump_flood.wgsl:48)
                int32_t d = sq_dist(x, y, (int32_t)sx, (int32_t)sy);
                if (d < best_d) { bx = sx; by = sy; best_d = d; }
            }
            out[pi] = bx; out[pi + 1] = by;
        }
    }
}

// [survey-pinned] host protocol, SURVEY.md §8 row a7: jumps 2^(ceil(log2 max(w,h))-1) .. 1, strict ping-pong.
uint32_t oracle_jfa_first_jump(uint32_t w, uint32_t h) {
    uint32_t m = w > h ? w : h;
    uint32_t l = 0; while ((1u << l) < m) l++;
    return l == 0 ? 0u : (1u << (l - 1));
}
void oracle_jfa_run(const uint32_t* seeds_in, uint32_t w, uint32_t h, uint32_t* seeds_out) {
    size_t len = (size_t)w * h * 2;
    std::vector<uint32_t> a(seeds_in, seeds_in + len), b(len);
    uint32_t* src = a.data(); uint32_t* dst = b.data();
    for (uint32_t jump = oracle_jfa_first_jump(w, h); jump >= 1; jump >>= 1) {
        oracle_jfa_pass(src, w, h, jump, dst);
        std::swap(src, dst);
    }
    memcpy(seeds_out, src, len * 4);
}

// seed rule of src/main.rs:413-438 expressed as a JFA seed image: seed pixels hold their own
// coordinate, all others FLUID_JFA_INVALID.
void oracle_jfa_seed_from_mask(const uint8_t* mask, uint32_t w, uint32_t h, uint32_t* seeds) {
    bool has_white = false;
    size_t npx = (size_t)w * h;
    for (size_t i = 0; i < npx; i++) if (mask[i] > 128) { has_white = true; break; }
    for (uint32_t y = 0; y < h; y++) for (uint32_t x = 0; x < w; x++) {
        size_t i = (size_t)y * w + x;
        bool seed = has_white ? (mask[i] > 128) : (y == h - 1 || y == 0 || x == w - 1 || x == 0);
        seeds[2 * i] = seed ? x : FLUID_JFA_INVALID;
        seeds[2 * i + 1] = seed ? y : FLUID_JFA_INVALID;
    }
}

// texel mapping of compute.wgsl:127-128 applied to predicted positions; clamp to tex-1 [survey-pinned §8d]
void oracle_jfa_seed_from_particles(const P* p, uint32_t n, const float* bounds, uint32_t w, uint32_t h,
                                    uint32_t* seeds) {
    size_t npx = (size_t)w * h;
    for (size_t i = 0; i < npx; i++) { seeds[2 * i] = FLUID_JFA_INVALID; seeds[2 * i + 1] = FLUID_JFA_INVALID; }
    for (uint32_t i = 0; i < n; i++) {
        float uvx = (p[i].predicted_position[0] / bounds[0] * 1.0f) + 0.5f;
        float uvy = (p[i].predicted_position[1] / bounds[1] * 1.0f) + 0.5f;
        uint32_t x = f2u(uvx * (float)w), y = f2u(uvy * (float)h);
        if (x > w - 1) x = w - 1;
        if (y > h - 1) y = h - 1;
        size_t k = (size_t)y * w + x;
        seeds[2 * k] = x; seeds[2 * k + 1] = y;
    }
}

// field = -(pixel - seed) with the zero rule of src/main.rs:493-511; invalid pixels -> 0 [survey-pinned]
void oracle_jfa_to_force_field(const uint32_t* seeds, uint32_t w, uint32_t h, float* field) {
#pragma omp parallel for schedule(static)
    for (int64_t y = 0; y < (int64_t)h; y++) for (uint32_t x = 0; x < w; x++) {
        size_t i = (size_t)y * w + x;
        uint32_t sx = seeds[2 * i], sy = seeds[2 * i + 1];
        float gx = 0.0f, gy = 0.0f;
        if (!(sx >= w || sy >= h)) {
            float dx = (float)x - (float)sx, dy = (float)y - (float)sy;
            float len = sqrtf(dx * dx + dy * dy);
            gx = len > 1e-6f ? dx : 0.0f;
            gy = len > 1e-6f ? dy : 0.0f;
        }
        field[2 * i] = -gx; field[2 * i + 1] = -gy;
    }
}

// src/main.rs:403-515  generate_smooth_gradient_field (the reference's own CPU routine), line for line.
// nearest_out (optional): w*h (x,y) u32 pairs of the nearest source found.
void oracle_generate_smooth_gradient_field(const uint8_t* img, uint32_t width_u, uint32_t height_u, float* field,
                                           uint32_t* nearest_out) {
    const size_t width = width_u, height = height_u;
    std::vector<float> dist(width * height, 3.40282347e+38f);
    std::vector<size_t> nsx(width * height, 0), nsy(width * height, 0);
    bool has_white = false;
    for (size_t y = 0; y < height; y++) for (size_t x = 0; x < width; x++) {
        if (img[y * width + x] > 128) { dist[y * width + x] = 0.0f; nsx[y * width + x] = x; nsy[y * width + x] = y; has_white = true; }
    }
    if (!has_white) {
        for (size_t y = 0; y < height; y++) for (size_t x = 0; x < width; x++) {
            if (y == height - 1 || y == 0 || x == width - 1 || x == 0) {
                dist[y * width + x] = 0.0f; nsx[y * width + x] = x; nsy[y * width + x] = y;
            }
        }
    }
    auto squared_dist = [](size_t x1, size_t y1, size_t x2, size_t y2) -> float {
        float dx = (float)x1 - (float)x2, dy = (float)y1 - (float)y2;
        return dx * dx + dy * dy;
    };
    for (size_t y = 0; y < height; y++) for (size_t x = 0; x < width; x++) {
        const size_t nb[4][2] = {{x - 1, y}, {x - 1, y - 1}, {x, y - 1}, {x + 1, y - 1}};  // usize wrapping_sub
        for (int k = 0; k < 4; k++) {
            size_t nx = nb[k][0], ny = nb[k][1];
            if (nx < width && ny < height) {
                size_t cxs = nsx[ny * width + nx], cys = nsy[ny * width + nx];
                float cd = squared_dist(x, y, cxs, cys);
                if (cd < dist[y * width + x]) { dist[y * width + x] = cd; nsx[y * width + x] = cxs; nsy[y * width + x] = cys; }
            }
        }
    }
    for (size_t y = height; y-- > 0;) for (size_t x = width; x-- > 0;) {
        const size_t nb[4][2] = {{x + 1, y}, {x + 1, y + 1}, {x, y + 1}, {x - 1, y + 1}};
        for (int k = 0; k < 4; k++) {
            size_t nx = nb[k][0], ny = nb[k][1];
            if (nx < width && ny < height) {
                size_t cxs = nsx[ny * width + nx], cys = nsy[ny * width + nx];
                float cd = squared_dist(x, y, cxs, cys);
                if (cd < dist[y * width + x]) { dist[y * width + x] = cd; nsx[y * width + x] = cxs; nsy[y * width + x] = cys; }
            }
        }
    }
    for (size_t y = 0; y < height; y++) for (size_t x = 0; x < width; x++) {
        size_t sx = nsx[y * width + x], sy = nsy[y * width + x];
        float dx = (float)x - (float)sx, dy = (float)y - (float)sy;
        float len = sqrtf(dx * dx + dy * dy);
        float gx = len > 1e-6f ? dx : 0.0f, gy = len > 1e-6f ? dy : 0.0f;
        field[2 * (y * width + x)] = -gx; field[2 * (y * width + x) + 1] = -gy;
        if (nearest_out) { nearest_out[2 * (y * width + x)] = (uint32_t)sx; nearest_out[2 * (y * width + x) + 1] = (uint32_t)sy; }
    }
}

// Brute-force checks used by tests (fp64, O(N^2)): density with poly6 and the pressure+viscosity sum.
void oracle_bruteforce_density_f64(const U* u, const P* p, double* out) {
    const uint32_t n = u->particle_count;
    const double h = u->smoothing_radius, h2 = h * h;
    const double norm = 4.0 / (3.14159265358979323846 * pow(h, 8.0));
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < (int64_t)n; i++) {
        double d = 0.0;
        for (uint32_t j = 0; j < n; j++) {
            double dx = (double)p[j].predicted_position[0] - (double)p[i].predicted_position[0];
            double dy = (double)p[j].predicted_position[1] - (double)p[i].predicted_position[1];
            double r2 = dx * dx + dy * dy;
            if (r2 <= h2) { double df = h2 - r2; d += (double)u->particle_mass * norm * df * df * df; }
        }
        out[i] = d;
    }
}
void oracle_bruteforce_accel_f64(const U* u, const P* p, double* out_xy) {
    const uint32_t n = u->particle_count;
    const double h = u->smoothing_radius, h2 = (double)u->sqr_radius;
    const double cs = u->spiky_kernel_derivative, cv = u->viscosity_kernel;
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < (int64_t)n; i++) {
        double fx = 0, fy = 0;
        double Pi = (double)u->pressure_constant * ((double)p[i].density - (double)u->rest_density);
        for (uint32_t j = 0; j < n; j++) {
            if ((int64_t)j == i) continue;
            double dx = (double)p[j].predicted_position[0] - (double)p[i].predicted_position[0];
            double dy = (double)p[j].predicted_position[1] - (double)p[i].predicted_position[1];
            double r2 = dx * dx + dy * dy;
            if (r2 > h2 || r2 == 0.0) continue;
            double r = sqrt(r2);
            double Pj = (double)u->pressure_constant * ((double)p[j].density - (double)u->rest_density);
            double k = -(h - r) * cs;
            double sh = (Pi + Pj) * 0.5;
            fx += dx / r * k * sh / p[j].density;
            fy += dy / r * k * sh / p[j].density;
            double kv = cv * ((-(r * r * r) / (2.0 * h * h * h)) + ((r * r) / (h * h)) + (h / (2.0 * r)) - 1.0);
            fx += (double)u->viscosity_coefficient * ((double)p[j].velocity[0] - (double)p[i].velocity[0]) / p[j].density * kv;
            fy += (double)u->viscosity_coefficient * ((double)p[j].velocity[1] - (double)p[i].velocity[1]) / p[j].density * kv;
        }
        out_xy[2 * i] = fx; out_xy[2 * i + 1] = fy;
    }
}

}  // extern "C"
