// render.cu — SURVEY.md §8 row f2: headless fluid render.
//
// Reference: fluid_shader.wgsl:18-103 (the live root copy included by src/renderer.rs:162), drawn as one quad over
// the RENDER_DIMS = 960x540 target (src/renderer.rs:15) with ALPHA_BLENDING over a target cleared to opaque black
// (:178, :437). Per pixel: world position from inv_proj, a 5x5-cell walk over the SORTED particle buffer through
// the reference's start_indices table (exactly as the shader does: start at start_indices[id], stop at the first
// particle whose key differs — so the never-cleared table's stale entries behave as in the reference), a Gaussian
// splat of density and speed, and a colour ramp. It is the only consumer of the boundary's "sorted AoS + start
// table" contract, so it exercises the drop-in. No graphics API: the image is a device buffer (RGBA32F fragment
// output, optionally the blended RGBA8 image) copied to the host on request.
//
// Arithmetic: the splat sum runs in the shader's order with IEEE-rounded + * / sqrt; exp and log are CUDA's expf /
// logf (<= 2 ulp), so parity with the CPU restatement is to a stated tolerance (tests: 2e-5 absolute on colour),
// not bit-exact. Pixel-centre clip coordinates and matrix rows as pinned in field.inl.
#include "common.cuh"
#include "sim_internal.h"

namespace fluid {

__device__ __forceinline__ float clampf(float x, float lo, float hi) { return fminf(fmaxf(x, lo), hi); }
__device__ __forceinline__ float smoothstepf(float e0, float e1, float x) {
    const float t = clampf(fdiv(fsub(x, e0), fsub(e1, e0)), 0.0f, 1.0f);
    return fmul(fmul(t, t), fsub(3.0f, fmul(2.0f, t)));
}
__device__ __forceinline__ float mixf(float a, float b, float t) { return fadd(fmul(a, fsub(1.0f, t)), fmul(b, t)); }

struct RenderParams {
    float m00, m01, m03, m10, m11, m13, m30, m31, m33;  // m<c><r> of the column-major inv_proj
    float half_bx, half_by, h, half_r2, log_norm;
    uint32_t grid_w, num_cells, n, w, h_px;
    int srgb;  // encode the UNORM8 image as an *Srgb surface would on store
};
// linear -> sRGB transfer function (IEC 61966-2-1), what a Bgra8UnormSrgb / Rgba8UnormSrgb render target applies
__device__ __forceinline__ float linear_to_srgb(float c) {
    return c <= 0.0031308f ? 12.92f * c : 1.055f * powf(c, 1.0f / 2.4f) - 0.055f;
}

__global__ void __launch_bounds__(256)
k_render_fluid(const RenderParams q, const float2* __restrict__ pred, const float2* __restrict__ vel,
               const uint32_t* __restrict__ keys, const uint32_t* __restrict__ start_indices, float4* __restrict__ out,
               uchar4* __restrict__ out_rgba8) {
    // 16x16 pixel tiles: neighbouring threads walk the same cells (L1-resident)
    const uint32_t px = blockIdx.x * 16u + (threadIdx.x & 15u);
    const uint32_t py = blockIdx.y * 16u + (threadIdx.x >> 4);
    if (px >= q.w || py >= q.h_px) return;
    const float cx = fmul(fsub(fdiv(fadd((float)px, 0.5f), (float)q.w), 0.5f), 2.0f);
    const float cy = -fmul(fsub(fdiv(fadd((float)py, 0.5f), (float)q.h_px), 0.5f), 2.0f);
    const float ix = fadd(fadd(fmul(q.m00, cx), fmul(q.m10, cy)), q.m30);
    const float iy = fadd(fadd(fmul(q.m01, cx), fmul(q.m11, cy)), q.m31);
    const float iw = fadd(fadd(fmul(q.m03, cx), fmul(q.m13, cy)), q.m33);
    const float wx = fdiv(ix, iw), wy = fdiv(iy, iw);  // fluid_shader.wgsl:29-32
    float density = 0.0f, velocity_factor = 0.0f;
    const int cellx = (int)(f2u(floorf(fdiv(fadd(wx, q.half_bx), q.h))) + 1u);  // :38, funcs.wgsl:212-214
    const int celly = (int)(f2u(floorf(fdiv(fadd(wy, q.half_by), q.h))) + 1u);
    for (int oy = -2; oy < 3; oy++) {
        for (int ox = -2; ox < 3; ox++) {
            const uint32_t x = (uint32_t)(cellx + ox), y = (uint32_t)(celly + oy);  // :41-42
            const uint32_t id = y * q.grid_w + x;                                    // :44
            if (id >= q.num_cells) continue;                                         // [pinned] outside the table: empty
            uint32_t i = __ldg(start_indices + id);
            while (i < q.n && __ldg(keys + i) == id) {                               // :49, :53
                const float2 pp = __ldg(pred + i), vv = __ldg(vel + i);
                const float offx = fsub(pp.x, wx), offy = fsub(pp.y, wy);
                const float r2 = fadd(fmul(offx, offx), fmul(offy, offy));           // :64
                const float contrib = expf(fdiv(-r2, q.half_r2));                    // :66
                density = fadd(density, contrib);
                velocity_factor = fadd(velocity_factor, fmul(contrib, fsqrt(fadd(fmul(vv.x, vv.x), fmul(vv.y, vv.y)))));  // :68
                i++;
            }
        }
    }
    velocity_factor = fmul(velocity_factor, 0.01f);                                           // :79
    velocity_factor = fdiv(logf(fadd(1.0f, fmul(5.0f, velocity_factor))), q.log_norm);        // :82
    velocity_factor = clampf(velocity_factor, 0.0f, 1.0f);                                    // :83
    const float interior = smoothstepf(0.5f, 1.5f, density);                                  // :86
    float edge = fsub(smoothstepf(0.7f, 1.0f, density), smoothstepf(1.0f, 1.5f, density));    // :89
    edge = fmul(edge, fadd(1.0f, fmul(velocity_factor, 2.0f)));                               // :90
    const float r = fadd(fmul(mixf(0.0f, 1.0f, velocity_factor), interior), edge);            // :93-96
    const float g = fadd(fmul(mixf(0.5f, 0.0f, velocity_factor), interior), edge);
    const float b = fadd(fmul(mixf(1.0f, 0.0f, velocity_factor), interior), edge);
    const float a = clampf(interior, 0.0f, 1.0f);                                             // :99
    const size_t o = (size_t)py * q.w + px;
    if (out) out[o] = make_float4(r, g, b, a);
    if (out_rgba8) {
        // ALPHA_BLENDING over (0,0,0,1): rgb = src.rgb * src.a, a = 1; UNORM8 store (clamp, round to nearest)
        float br = clampf(r * a, 0.0f, 1.0f), bg = clampf(g * a, 0.0f, 1.0f), bb = clampf(b * a, 0.0f, 1.0f);
        if (q.srgb) { br = linear_to_srgb(br); bg = linear_to_srgb(bg); bb = linear_to_srgb(bb); }
        out_rgba8[o] = make_uchar4((unsigned char)__float2uint_rn(br * 255.0f), (unsigned char)__float2uint_rn(bg * 255.0f),
                                   (unsigned char)__float2uint_rn(bb * 255.0f), 255);
    }
}

}  // namespace fluid

using namespace fluid;

static int render_common(FluidSim* h_, const float inv_proj[16], uint32_t w, uint32_t h, float* rgba_f32_out, uint8_t* rgba8_out,
                         float* ms_out, int srgb, bool sync) {
    if (!h_ || !inv_proj || (!rgba_f32_out && !rgba8_out)) { set_error("fluid_sim_render: null argument"); return FLUID_ERR_INVALID_ARGUMENT; }
    if (w == 0 || h == 0 || w > 16384 || h > 16384) { set_error("fluid_sim_render: bad image size"); return FLUID_ERR_INVALID_ARGUMENT; }
    Sim* s = reinterpret_cast<Sim*>(h_);
    if (s->band) { set_error("fluid_sim_render: not available on a band of a FluidMulti (gather the bands first)"); return FLUID_ERR_UNSUPPORTED; }
    FLUID_CUDA(cudaSetDevice(s->device));
    const size_t npx = (size_t)w * h;
    if (npx > s->frame_px) {  // grown on demand, kept for the following frames
        FLUID_CUDA(cudaStreamSynchronize(s->stream));
        if (s->frame_f32) cudaFree(s->frame_f32);
        if (s->frame_u8) cudaFree(s->frame_u8);
        s->frame_f32 = s->frame_u8 = nullptr; s->frame_px = 0;
        FLUID_CUDA(cudaMalloc(&s->frame_f32, npx * sizeof(float4)));
        FLUID_CUDA(cudaMalloc(&s->frame_u8, npx * sizeof(uchar4)));
        s->frame_px = npx;
    }
    float4* d_f = rgba_f32_out ? reinterpret_cast<float4*>(s->frame_f32) : nullptr;
    uchar4* d_8 = rgba8_out ? reinterpret_cast<uchar4*>(s->frame_u8) : nullptr;
    RenderParams q;
    const float* m = inv_proj;
    q.m00 = m[0]; q.m01 = m[1]; q.m03 = m[3]; q.m10 = m[4]; q.m11 = m[5]; q.m13 = m[7]; q.m30 = m[12]; q.m31 = m[13]; q.m33 = m[15];
    q.half_bx = s->settings.size[0] * 0.5f; q.half_by = s->settings.size[1] * 0.5f;
    q.h = s->settings.smoothing_radius;
    q.half_r2 = (s->settings.smoothing_radius * s->settings.smoothing_radius) / 2.0f;
    q.log_norm = logf(1.0f + 5.0f);
    q.grid_w = s->grid_w; q.num_cells = s->num_cells; q.n = s->n; q.w = w; q.h_px = h;
    q.srgb = srgb ? 1 : 0;
    if (ms_out) {
        if (!s->ev_frame0) { FLUID_CUDA(cudaEventCreate(&s->ev_frame0)); FLUID_CUDA(cudaEventCreate(&s->ev_frame1)); }
        cudaEventRecord(s->ev_frame0, s->stream);
    }
    dim3 grid((w + 15) / 16, (h + 15) / 16);
    k_render_fluid<<<grid, 256, 0, s->stream>>>(q, s->pred_s, s->vel_a, s->keys_view(), s->start_indices, d_f, d_8);
    s->launches++;
    if (ms_out) cudaEventRecord(s->ev_frame1, s->stream);
    if (rgba_f32_out) cudaMemcpyAsync(rgba_f32_out, d_f, npx * sizeof(float4), cudaMemcpyDeviceToHost, s->stream);
    if (rgba8_out) cudaMemcpyAsync(rgba8_out, d_8, npx * sizeof(uchar4), cudaMemcpyDeviceToHost, s->stream);
    if (sync || ms_out) {
        FLUID_CUDA(cudaStreamSynchronize(s->stream));
        if (ms_out) { float ms = 0.f; cudaEventElapsedTime(&ms, s->ev_frame0, s->ev_frame1); *ms_out = ms; }
    }
    FLUID_CUDA(cudaGetLastError());
    return FLUID_OK;
}

extern "C" int fluid_sim_render(FluidSim* h_, const float inv_proj[16], uint32_t w, uint32_t h, float* rgba_f32_out,
                                uint8_t* rgba8_out, float* ms_out) {
    return render_common(h_, inv_proj, w, h, rgba_f32_out, rgba8_out, ms_out, 0, true);
}
extern "C" int fluid_sim_render_frame(FluidSim* h_, const float inv_proj[16], uint32_t w, uint32_t h, uint8_t* rgba8_out, int srgb, int wait) {
    return render_common(h_, inv_proj, w, h, nullptr, rgba8_out, nullptr, srgb, wait != 0);
}
