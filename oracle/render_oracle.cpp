// render_oracle.cpp — CPU ORACLE (TEST INFRASTRUCTURE, NOT PRODUCT CODE), part 2: the reference's per-pixel
// shaders restated as plain C++ loops. Same rules as fluid_oracle.cpp (IEEE binary32, no contraction, correctly
// rounded / and sqrt). PARITY UNPINNED for the same reason: the WGSL cannot run here.
//
//   oracle_rasterize_objects  image_shader.wgsl:32-85 (vs_main, fs_main, point_in_rect) -> R8Uint mask
//   oracle_render_fluid       fluid_shader.wgsl:18-103 (the live root copy: velocity scale 0.01, no density > 50 branch)
//                             -> the fragment's vec4<f32> per pixel, before blending
//
// Pinned choices where WGSL leaves latitude: the fragment's interpolated clip position is evaluated at the pixel
// centre, cx = ((px+0.5)/w - 0.5)*2, cy = -(((py+0.5)/h - 0.5)*2) (framebuffer row 0 = clip y +1);
// inv_proj * vec4(cx, cy, 0, 1) is evaluated per row as (m0r*cx + m1r*cy) + m3r; cos(-rot) and sin(-rot) come
// from libm's cosf/sinf, once per object.
#include <cmath>
#include <cstdint>
#include <cstring>

#include "../include/fluid_b200.h"

extern "C" {

void oracle_rasterize_objects(const FluidObject* objs, uint32_t n, const float* m /* inv_proj, 16 floats column-major */,
                              uint32_t w, uint32_t h, uint8_t* mask) {
#pragma omp parallel for schedule(static)
    for (int64_t py = 0; py < (int64_t)h; py++) {
        for (uint32_t px = 0; px < w; px++) {
            // vs_main :35 (position - 0.5) * 2 at the pixel centre; fs_main :43-45
            const float cx = (((float)px + 0.5f) / (float)w - 0.5f) * 2.0f;
            const float cy = -((((float)py + 0.5f) / (float)h - 0.5f) * 2.0f);
            const float ix = (m[0] * cx + m[4] * cy) + m[12];
            const float iy = (m[1] * cx + m[5] * cy) + m[13];
            const float iw = (m[3] * cx + m[7] * cy) + m[15];
            const float wx = ix / iw, wy = iy / iw;
            uint8_t out = 255;  // :66
            for (uint32_t i = 0; i < n && out; i++) {  // :47
                const FluidObject& o = objs[i];
                const float lx = wx - o.position[0], ly = wy - o.position[1];
                float f;
                memcpy(&f, &o.pad, 4);
                if (o.kind == 0u) {  // :50-55
                    const float dst = sqrtf(lx * lx + ly * ly);
                    if (dst < f) out = 0;
                } else if (o.kind == 1u) {  // :56-61 and point_in_rect :70-85
                    float ex, ey;
                    memcpy(&ex, &o.pad2[0], 4);
                    memcpy(&ey, &o.pad2[1], 4);
                    const float c = cosf(-f), s = sinf(-f);
                    const float rx = lx * c - ly * s;
                    const float ry = lx * s + ly * c;
                    const float hx = ex * 0.5f, hy = ey * 0.5f;
                    if (rx >= -hx && ry >= -hy && rx <= hx && ry <= hy) out = 0;
                }
            }
            mask[(size_t)py * w + px] = out;
        }
    }
}


// ---- fluid_shader.wgsl ------------------------------------------------------------------------------------------
namespace {
inline float clampf(float x, float lo, float hi) { return fminf(fmaxf(x, lo), hi); }
// WGSL smoothstep(e0, e1, x): t = clamp((x - e0) / (e1 - e0), 0, 1); t * t * (3 - 2 t)
inline float smoothstepf(float e0, float e1, float x) {
    const float t = clampf((x - e0) / (e1 - e0), 0.0f, 1.0f);
    return t * t * (3.0f - 2.0f * t);
}
// WGSL mix(a, b, t) = a * (1 - t) + b * t
inline float mixf(float a, float b, float t) { return a * (1.0f - t) + b * t; }
// WGSL f32 -> u32: truncate, clamp, NaN -> 0
inline uint32_t f2u_sat(float f) {
    if (!(f > 0.0f)) return 0u;
    if (f >= 4294967296.0f) return 0xffffffffu;
    return (uint32_t)f;
}
}  // namespace

// particles: the simulation's particle buffer after a tick (sorted AoS), start_indices: its table (never cleared,
// src/simulation.rs:204-209); u: the tick's uniform. Pinned: pixel-centre clip coordinates and the matrix rows as in
// oracle_rasterize_objects; cells whose id is outside the table are empty; expf / logf / sqrtf from libm.
// out: w*h RGBA float, the value fs_main returns (ALPHA_BLENDING over the cleared black target then gives
// rgb * a, 1: src/renderer.rs:178, :437).
void oracle_render_fluid(const FluidSimUniform* u, const FluidParticle* p, const uint32_t* start_indices, const float* m,
                         uint32_t w, uint32_t h, float* out) {
    const uint32_t num_cells = u->grid_w * u->grid_h;
    const float half_r2 = u->sqr_radius / 2.0f;  // :66
    const float log_norm = logf(1.0f + 5.0f);    // :82
#pragma omp parallel for schedule(dynamic, 4)
    for (int64_t py = 0; py < (int64_t)h; py++) {
        for (uint32_t px = 0; px < w; px++) {
            const float cx = (((float)px + 0.5f) / (float)w - 0.5f) * 2.0f;
            const float cy = -((((float)py + 0.5f) / (float)h - 0.5f) * 2.0f);
            const float ix = (m[0] * cx + m[4] * cy) + m[12];
            const float iy = (m[1] * cx + m[5] * cy) + m[13];
            const float iw = (m[3] * cx + m[7] * cy) + m[15];
            const float wx = ix / iw, wy = iy / iw;  // :29-32
            float density = 0.0f, velocity_factor = 0.0f;
            // :38 xy_of_point (funcs.wgsl:212-214), then vec2<i32>()
            const int32_t cellx = (int32_t)(f2u_sat(floorf((wx + u->bounds[0] * 0.5f) / u->smoothing_radius)) + 1u);
            const int32_t celly = (int32_t)(f2u_sat(floorf((wy + u->bounds[1] * 0.5f) / u->smoothing_radius)) + 1u);
            for (int oy = -2; oy < 3; oy++) {
                for (int ox = -2; ox < 3; ox++) {
                    const uint32_t x = (uint32_t)(cellx + ox), y = (uint32_t)(celly + oy);  // :41-42 (wraps)
                    const uint32_t id = y * u->grid_w + x;                                    // :44
                    if (id >= num_cells) continue;                                            // [pinned] outside the table: empty
                    uint32_t start_index = start_indices[id];
                    while (true) {
                        if (start_index >= u->particle_count) break;   // :49
                        if (p[start_index].grid != id) break;          // :53
                        const uint32_t i = start_index;
                        start_index += 1;
                        const float offx = p[i].predicted_position[0] - wx, offy = p[i].predicted_position[1] - wy;
                        const float r2 = offx * offx + offy * offy;    // :64 dot
                        const float contrib = expf(-r2 / half_r2);     // :66
                        density += contrib;
                        const float vx = p[i].velocity[0], vy = p[i].velocity[1];
                        velocity_factor += contrib * sqrtf(vx * vx + vy * vy);  // :68
                    }
                }
            }
            velocity_factor = velocity_factor * 0.01f;                                    // :79
            velocity_factor = logf(1.0f + 5.0f * velocity_factor) / log_norm;             // :82
            velocity_factor = clampf(velocity_factor, 0.0f, 1.0f);                        // :83
            const float interior = smoothstepf(0.5f, 1.5f, density);                      // :86
            float edge = smoothstepf(0.7f, 1.0f, density) - smoothstepf(1.0f, 1.5f, density);  // :89
            edge = edge * (1.0f + velocity_factor * 2.0f);                                // :90
            const float br = mixf(0.0f, 1.0f, velocity_factor) * interior;                // :93
            const float bg = mixf(0.5f, 0.0f, velocity_factor) * interior;
            const float bb = mixf(1.0f, 0.0f, velocity_factor) * interior;
            float* o = out + ((size_t)py * w + px) * 4;
            o[0] = br + edge; o[1] = bg + edge; o[2] = bb + edge;                         // :94-96
            o[3] = clampf(interior, 0.0f, 1.0f);                                          // :99
        }
    }
}


// ---- dormant physics: surface tension (compute.wgsl:300-498; the call at :92 is commented out in the reference) ----
// per particle: f = calculate_surface_tension(particle.predicted_position), NOT applied to the state (as in the
// reference). Neighbour walks as in the shader (start_indices entry, stop at the first foreign key; self is not
// skipped); cells outside the table are empty [pinned]; pow(h, 8) with powf; PI = 3.14159265359 (funcs.wgsl:54).
// Note: poly6_kernel_gradient is handed the UNIT direction (compute.wgsl:440), so for h <= 1 it returns 0 for every
// neighbour (r_len = 1 >= h) and the whole term vanishes at the reference's h = 0.2.
}  // extern "C"
namespace {
const float ST_PI = 3.14159265359f;
inline uint32_t st_xorshift32(uint32_t& s) { uint32_t x = s; x ^= x << 13; x ^= x >> 17; x ^= x << 5; s = x; return x; }
inline float st_rand_f32(uint32_t& s) { return (float)st_xorshift32(s) / 4294967296.0f; }
template <class F>
inline void st_walk(const FluidSimUniform* u, const FluidParticle* p, const uint32_t* start_indices, float px, float py, F&& body) {
    const uint32_t num_cells = u->grid_w * u->grid_h;
    const int32_t cellx = (int32_t)(f2u_sat(floorf((px + u->bounds[0] * 0.5f) / u->smoothing_radius)) + 1u);
    const int32_t celly = (int32_t)(f2u_sat(floorf((py + u->bounds[1] * 0.5f) / u->smoothing_radius)) + 1u);
    for (int oy = -1; oy <= 1; oy++)
        for (int ox = -1; ox <= 1; ox++) {
            const uint32_t x = (uint32_t)(cellx + ox), y = (uint32_t)(celly + oy);
            const uint32_t id = y * u->grid_w + x;
            if (id >= num_cells) continue;
            uint32_t si = start_indices[id];
            while (si < u->particle_count && p[si].grid == id) { body(p[si]); si++; }
        }
}
}  // namespace

extern "C" {
void oracle_surface_tension(const FluidSimUniform* u, const FluidParticle* p, const uint32_t* start_indices, float* out_xy) {
    const float h = u->smoothing_radius;
    const float h8 = powf(h, 8.0f);
    const float c_grad = -24.0f / (ST_PI * h8), c_lap = 8.0f / (ST_PI * h8);
#pragma omp parallel for schedule(dynamic, 256)
    for (int64_t id = 0; id < (int64_t)u->particle_count; id++) {
        const float px = p[id].predicted_position[0], py = p[id].predicted_position[1];
        // calculate_colour_field_gradient :410-498
        uint32_t seed = f2u_sat(px) * 324u + u->frame_time * 5632u;
        float nx = 0.0f, ny = 0.0f;
        st_walk(u, p, start_indices, px, py, [&](const FluidParticle& q) {
            const float ox = q.predicted_position[0] - px, oy = q.predicted_position[1] - py;
            const float sq = ox * ox + oy * oy;
            if (sq > u->sqr_radius) return;
            const float dst = sqrtf(sq);
            float dx, dy;
            if (dst == 0.0f) {
                const float rx = st_rand_f32(seed), ry = st_rand_f32(seed);
                const float len = sqrtf(rx * rx + ry * ry);
                dx = rx / len; dy = ry / len;
            } else { dx = ox / dst; dy = oy / dst; }
            // poly6_kernel_gradient(h, dir) funcs.wgsl:81-88
            const float r_len = sqrtf(dx * dx + dy * dy);
            float kx = 0.0f, ky = 0.0f;
            if (!(r_len >= h || r_len == 0.0f)) {
                const float diff2 = h * h - r_len * r_len;
                kx = c_grad * diff2 * diff2 * dx; ky = c_grad * diff2 * diff2 * dy;
            }
            const float w = u->particle_mass / q.density;
            nx += w * kx; ny += w * ky;
        });
        const float n_len = sqrtf(nx * nx + ny * ny);
        float fx = 0.0f, fy = 0.0f;
        if (n_len > u->surface_tension_treshold) {  // compute.wgsl:304
            float lap = 0.0f;                        // calculate_colour_field_laplacian :316-405
            st_walk(u, p, start_indices, px, py, [&](const FluidParticle& q) {
                const float ox = q.predicted_position[0] - px, oy = q.predicted_position[1] - py;
                const float sq = ox * ox + oy * oy;
                if (sq > u->sqr_radius) return;
                const float dst = sqrtf(sq);
                float k = 0.0f;
                if (!(dst > h)) { const float h2 = h * h, r2 = dst * dst; k = c_lap * (h2 - r2) * (3.0f * h2 - 4.0f * r2); }  // funcs.wgsl:91-98
                lap += (u->particle_mass / q.density) * k;
            });
            const float k = (-lap) / (n_len + 1e-6f);
            const float s = -u->surface_tension_coefficient * k;
            fx = s * (nx / n_len); fy = s * (ny / n_len);
        }
        out_xy[2 * id] = fx; out_xy[2 * id + 1] = fy;
    }
}

}  // extern "C"
