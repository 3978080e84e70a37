// dormant.cu — SURVEY.md §8 row f4: the reference's dormant surface-tension term.
//
// Reference: calculate_surface_tension, calculate_colour_field_gradient, calculate_colour_field_laplacian
// (compute.wgsl:300-498), poly6_kernel_gradient / poly6_kernel_laplacian (funcs.wgsl:81-98). The call in
// move_particle is commented out (compute.wgsl:92), so the reference never applies this force; this file evaluates
// it per particle from the state of the last tick into a separate buffer — a diagnostic with exactly the shader's
// arithmetic (IEEE-rounded, reference order, bit-identical to the CPU restatement of the shader) that can be
// added to the acceleration once the reference enables the line. Not on the per-tick path, so it is written for
// clarity: one thread per particle walking the reference's start_indices table, like the shader.
// (poly6_kernel_gradient receives the UNIT direction, compute.wgsl:440, so with h <= 1 the term is identically 0.)
#include <cmath>

#include "common.cuh"
#include "sim_internal.h"

namespace fluid {

struct StParams {
    float half_bx, half_by, h, h2, sqr_radius, mass, threshold, coefficient, c_grad, c_lap;
    uint32_t grid_w, num_cells, n, frame_time;
};

__device__ __forceinline__ float st_rand(uint32_t& s) {  // funcs.wgsl:126-149
    uint32_t x = s;
    x ^= x << 13; x ^= x >> 17; x ^= x << 5;
    s = x;
    return fmul(__uint2float_rn(x), 2.3283064365386962890625e-10f);  // exact: / 4294967296.0
}

__global__ void __launch_bounds__(128)
k_surface_tension(const StParams q, const float2* __restrict__ pred, const float* __restrict__ rho,
                  const uint32_t* __restrict__ keys, const uint32_t* __restrict__ start_indices, float2* __restrict__ out) {
    const uint32_t id = blockIdx.x * blockDim.x + threadIdx.x;
    if (id >= q.n) return;
    const float2 pt = __ldg(pred + id);
    const int cellx = (int)(f2u(floorf(fdiv(fadd(pt.x, q.half_bx), q.h))) + 1u);  // xy_of_point, funcs.wgsl:212-214
    const int celly = (int)(f2u(floorf(fdiv(fadd(pt.y, q.half_by), q.h))) + 1u);
    uint32_t seed = f2u(pt.x) * 324u + q.frame_time * 5632u;  // compute.wgsl:411
    float nx = 0.0f, ny = 0.0f, lap = 0.0f;
    // both sums walk the same neighbours in the same order; they are independent, so one walk computes both
    // (the Laplacian is only USED when |n| exceeds the threshold, as in the shader)
    for (int oy = -1; oy <= 1; oy++) {
        for (int ox = -1; ox <= 1; ox++) {
            const uint32_t x = (uint32_t)(cellx + ox), y = (uint32_t)(celly + oy);
            const uint32_t cid = y * q.grid_w + x;
            if (cid >= q.num_cells) continue;  // [pinned] outside the table: empty
            uint32_t i = __ldg(start_indices + cid);
            while (i < q.n && __ldg(keys + i) == cid) {
                const float2 np = __ldg(pred + i);
                const float nd = __ldg(rho + i);
                i++;
                const float offx = fsub(np.x, pt.x), offy = fsub(np.y, pt.y);
                const float sq = fadd(fmul(offx, offx), fmul(offy, offy));
                if (sq > q.sqr_radius) continue;
                const float dst = fsqrt(sq);
                float dx, dy;
                if (dst == 0.0f) {
                    const float rx = st_rand(seed), ry = st_rand(seed);
                    const float len = fsqrt(fadd(fmul(rx, rx), fmul(ry, ry)));
                    dx = fdiv(rx, len); dy = fdiv(ry, len);
                } else {
                    dx = fdiv(offx, dst); dy = fdiv(offy, dst);
                }
                const float w = fdiv(q.mass, nd);
                const float r_len = fsqrt(fadd(fmul(dx, dx), fmul(dy, dy)));  // poly6_kernel_gradient, funcs.wgsl:81-88
                float kx = 0.0f, ky = 0.0f;
                if (!(r_len >= q.h || r_len == 0.0f)) {
                    const float diff2 = fsub(fmul(q.h, q.h), fmul(r_len, r_len));
                    const float c = fmul(fmul(q.c_grad, diff2), diff2);
                    kx = fmul(c, dx); ky = fmul(c, dy);
                }
                nx = fadd(nx, fmul(w, kx));
                ny = fadd(ny, fmul(w, ky));
                float kl = 0.0f;  // poly6_kernel_laplacian, funcs.wgsl:91-98
                if (!(dst > q.h)) {
                    const float r2 = fmul(dst, dst);
                    kl = fmul(fmul(q.c_lap, fsub(q.h2, r2)), fsub(fmul(3.0f, q.h2), fmul(4.0f, r2)));
                }
                lap = fadd(lap, fmul(w, kl));
            }
        }
    }
    const float n_len = fsqrt(fadd(fmul(nx, nx), fmul(ny, ny)));
    float fx = 0.0f, fy = 0.0f;
    if (n_len > q.threshold) {  // compute.wgsl:304-308
        const float k = fdiv(-lap, fadd(n_len, 1e-6f));
        const float s = fmul(-q.coefficient, k);
        fx = fmul(s, fdiv(nx, n_len));
        fy = fmul(s, fdiv(ny, n_len));
    }
    out[id] = make_float2(fx, fy);
}

}  // namespace fluid

using namespace fluid;

extern "C" int fluid_sim_surface_tension(FluidSim* h_, float* force_xy_out) {
    if (!h_ || !force_xy_out) { set_error("fluid_sim_surface_tension: null argument"); return FLUID_ERR_INVALID_ARGUMENT; }
    Sim* s = reinterpret_cast<Sim*>(h_);
    if (s->band) { set_error("fluid_sim_surface_tension: not available on a band of a FluidMulti"); return FLUID_ERR_UNSUPPORTED; }
    FLUID_CUDA(cudaSetDevice(s->device));
    if (s->n == 0) return FLUID_OK;
    const FluidSimUniform& u = s->last_uniform;
    if (u.particle_count != s->n || (u.frame_time == 0 && u.delta == 0.0f)) {  // the create-time uniform
        set_error("fluid_sim_surface_tension: no tick has run on this handle yet"); return FLUID_ERR_INVALID_ARGUMENT; }
    StParams q;
    q.h = u.smoothing_radius;
    q.h2 = q.h * q.h;
    q.sqr_radius = u.sqr_radius;
    q.half_bx = u.bounds[0] * 0.5f; q.half_by = u.bounds[1] * 0.5f;
    q.mass = u.particle_mass; q.threshold = u.surface_tension_treshold; q.coefficient = u.surface_tension_coefficient;
    const float h8 = powf(q.h, 8.0f);  // pow(h, 8.0) of the shader, evaluated once on the host [pinned]
    const float PI_WGSL = 3.14159265359f;  // funcs.wgsl:54
    q.c_grad = -24.0f / (PI_WGSL * h8);
    q.c_lap = 8.0f / (PI_WGSL * h8);
    q.grid_w = s->grid_w; q.num_cells = s->num_cells; q.n = s->n; q.frame_time = u.frame_time;
    const size_t need = (size_t)s->n * sizeof(float2);
    if (need > s->diag_bytes) {  // grown on demand, kept
        FLUID_CUDA(cudaStreamSynchronize(s->stream));
        if (s->diag_out) cudaFree(s->diag_out);
        s->diag_out = nullptr; s->diag_bytes = 0;
        FLUID_CUDA(cudaMalloc(&s->diag_out, need));
        s->diag_bytes = need;
    }
    float2* d_out = reinterpret_cast<float2*>(s->diag_out);
    k_surface_tension<<<(s->n + 127) / 128, 128, 0, s->stream>>>(q, s->pred_s, s->rho, s->keys_view(), s->start_indices, d_out);
    s->launches++;
    cudaMemcpyAsync(force_xy_out, d_out, (size_t)s->n * sizeof(float2), cudaMemcpyDeviceToHost, s->stream);
    FLUID_CUDA(cudaStreamSynchronize(s->stream));
    FLUID_CUDA(cudaGetLastError());
    return FLUID_OK;
}

// src/main.rs:287-303 (glam: Mat4 * Vec4 = sum of columns scaled by the components, in x,y,z,w order)
extern "C" int fluid_mouse_world_pos(const float cursor_xy[2], const float window_size[2], const float inv_proj[16], float world_xy_out[2]) {
    if (!cursor_xy || !window_size || !inv_proj || !world_xy_out) { set_error("fluid_mouse_world_pos: null argument"); return FLUID_ERR_INVALID_ARGUMENT; }
    const volatile float nx = (cursor_xy[0] / window_size[0]) * 2.0f - 1.0f;
    const volatile float ny = (cursor_xy[1] / window_size[1]) * 2.0f - 1.0f;
    const float c[4] = {nx, -ny, 0.0f, 1.0f};
    float r[4];
    for (int k = 0; k < 4; k++) {
        volatile float a = inv_proj[0 + k] * c[0];
        volatile float b = inv_proj[4 + k] * c[1];
        volatile float d = inv_proj[8 + k] * c[2];
        volatile float e = inv_proj[12 + k] * c[3];
        volatile float ab = a + b;
        volatile float abd = ab + d;
        r[k] = abd + e;
    }
    world_xy_out[0] = r[0] / r[3];
    world_xy_out[1] = r[1] / r[3];
    return FLUID_OK;
}
