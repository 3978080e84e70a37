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

}  // extern "C"
