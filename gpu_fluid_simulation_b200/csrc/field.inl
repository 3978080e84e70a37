// field.cu — SURVEY.md §8 row f1: object mask -> seeds -> jump flood -> force field, end to end on the device.
//
// Reference path being replaced (src/renderer.rs:464-547): render the obstacle mask with image_shader.wgsl
// (fs_main :41-67, point_in_rect :70-85) into an R8Uint 1024^2 texture, copy it to a staging buffer, map it,
// hand the bytes to a spawned CPU thread running generate_smooth_gradient_field (src/main.rs:403-515), receive the
// 8 MiB field over a channel one or more frames later and write_buffer it into the simulation's `texture` binding,
// which compute.wgsl:127-140 reads. Here the mask is rasterised by a CUDA kernel into a headless u8 device buffer
// (no graphics API), seeded with the rule of src/main.rs:413-438, flooded with the JFA kernels of jfa.cu and
// converted in place into the sim's force-field buffer: same frame, no host round trip.
//
// Pinned arithmetic of the mask (WGSL leaves interpolation, sin/cos and distance accuracy to the implementation;
// the CPU restatement used by the tests pins the same choices): pixel-centre clip coordinates
// cx = ((px+0.5)/w - 0.5)*2, cy = -(((py+0.5)/h - 0.5)*2) (framebuffer row 0 is clip y = +1); inv_pos rows
// evaluated as (m0r*cx + m1r*cy) + m3r (z = 0 contributes exactly nothing, w = 1); IEEE binary32 without
// contraction, correctly rounded / and sqrt; cos(-rot), sin(-rot) evaluated once per object on the host (libm).
// (textually included by jfa.cu: one translation unit, so the JFA kernels are defined once)
#include <cmath>

namespace fluid {

// an object with its per-object constants hoisted (image_shader.wgsl:50-62, :75-76, :83)
struct PreparedObject {
    float px, py;
    uint32_t kind;  // 0 circle, 1 rotated rectangle, anything else: ignored (as in the shader)
    float a, b;     // circle: radius, -; rect: cos(-rot), sin(-rot)
    float hx, hy;   // rect: half extents
    uint32_t pad;
};
constexpr int RASTER_MAX_SMEM_OBJECTS = 128;

__global__ void __launch_bounds__(256)
k_raster_objects(const PreparedObject* __restrict__ objs, uint32_t n, float m00, float m01, float m03, float m10, float m11,
                 float m13, float m30, float m31, float m33, uint32_t w, uint32_t h, uint8_t* __restrict__ mask) {
    // m<c><r>: column c, row r of inv_proj (column-major mat4x4, image_shader.wgsl:15,44)
    __shared__ PreparedObject s_obj[RASTER_MAX_SMEM_OBJECTS];
    const uint32_t px = blockIdx.x * 64u + (threadIdx.x & 63u);
    const uint32_t py = blockIdx.y * 4u + (threadIdx.x >> 6);
    const bool in = px < w && py < h;
    // vs_main :35 at the pixel centre, then inv_proj * clip (:44), then the perspective divide (:45)
    const float cx = fmul(fsub(fdiv(fadd((float)px, 0.5f), (float)w), 0.5f), 2.0f);
    const float cy = -fmul(fsub(fdiv(fadd((float)py, 0.5f), (float)h), 0.5f), 2.0f);
    const float ix = fadd(fadd(fmul(m00, cx), fmul(m10, cy)), m30);
    const float iy = fadd(fadd(fmul(m01, cx), fmul(m11, cy)), m31);
    const float iw = fadd(fadd(fmul(m03, cx), fmul(m13, cy)), m33);
    const float wx = fdiv(ix, iw), wy = fdiv(iy, iw);
    uint32_t out = 255u;  // :66
    for (uint32_t base = 0; base < n; base += RASTER_MAX_SMEM_OBJECTS) {
        const uint32_t cnt = min((uint32_t)RASTER_MAX_SMEM_OBJECTS, n - base);
        __syncthreads();
        for (uint32_t i = threadIdx.x; i < cnt; i += blockDim.x) s_obj[i] = objs[base + i];
        __syncthreads();
        for (uint32_t i = 0; i < cnt && out; i++) {
            const PreparedObject o = s_obj[i];
            const float lx = fsub(wx, o.px), ly = fsub(wy, o.py);
            if (o.kind == 0u) {  // :50-55 distance(world_pos, obj.pos) < radius
                const float dst = fsqrt(fadd(fmul(lx, lx), fmul(ly, ly)));
                if (dst < o.a) out = 0u;
            } else if (o.kind == 1u) {  // :56-61, point_in_rect :70-85
                const float rx = fsub(fmul(lx, o.a), fmul(ly, o.b));
                const float ry = fadd(fmul(lx, o.b), fmul(ly, o.a));
                if (rx >= -o.hx && ry >= -o.hy && rx <= o.hx && ry <= o.hy) out = 0u;
            }
        }
    }
    if (in) mask[(size_t)py * w + px] = (uint8_t)out;
}

static std::vector<PreparedObject> prepare_objects(const FluidObject* objects, uint32_t n) {
    std::vector<PreparedObject> v(n);
    for (uint32_t i = 0; i < n; i++) {
        const FluidObject& o = objects[i];
        PreparedObject& q = v[i];
        memset(&q, 0, sizeof q);
        q.px = o.position[0]; q.py = o.position[1]; q.kind = o.kind;
        float f;
        memcpy(&f, &o.pad, 4);  // bitcast<f32>(obj.pad): radius (:52) or rotation (:57)
        if (o.kind == 0u) {
            q.a = f;
        } else if (o.kind == 1u) {
            float ex, ey;
            memcpy(&ex, &o.pad2[0], 4); memcpy(&ey, &o.pad2[1], 4);  // bitcast<vec2<f32>>(obj.pad2.xy) :58
            q.a = cosf(-f); q.b = sinf(-f);                            // :75-76
            q.hx = ex * 0.5f; q.hy = ey * 0.5f;                        // :83
        }
    }
    return v;
}

static void launch_raster(const PreparedObject* d_objs, uint32_t n, const float* m, uint32_t w, uint32_t h, uint8_t* d_mask,
                          cudaStream_t st) {
    dim3 grid((w + 63) / 64, (h + 3) / 4);
    k_raster_objects<<<grid, 256, 0, st>>>(d_objs, n, m[0], m[1], m[3], m[4], m[5], m[7], m[12], m[13], m[15], w, h, d_mask);
}

}  // namespace fluid

using namespace fluid;

extern "C" {

// mask only, host buffers (parity harness; the reference's object render pass, src/renderer.rs:464-491)
int fluid_raster_objects(const FluidObject* objects, uint32_t n, const float inv_proj[16], uint32_t w, uint32_t h,
                         uint8_t* mask_out, int device_ordinal) {
    if ((!objects && n) || !inv_proj || !mask_out) { set_error("fluid_raster_objects: null argument"); return FLUID_ERR_INVALID_ARGUMENT; }
    int rc = check_image(w, h); if (rc) return rc;
    rc = select_device(device_ordinal); if (rc) return rc;
    const size_t npx = (size_t)w * h;
    std::vector<PreparedObject> prep = prepare_objects(objects, n);
    PreparedObject* d_objs = nullptr;
    uint8_t* d_mask = nullptr;
    FLUID_CUDA(cudaMalloc(&d_objs, sizeof(PreparedObject) * (size_t)(n ? n : 1)));
    FLUID_CUDA(cudaMalloc(&d_mask, npx));
    if (n) cudaMemcpy(d_objs, prep.data(), sizeof(PreparedObject) * n, cudaMemcpyHostToDevice);
    launch_raster(d_objs, n, inv_proj, w, h, d_mask, 0);
    cudaError_t e = cudaMemcpy(mask_out, d_mask, npx, cudaMemcpyDeviceToHost);
    cudaFree(d_objs); cudaFree(d_mask);
    FLUID_CUDA(e);
    FLUID_CUDA(cudaGetLastError());
    return FLUID_OK;
}

// shared tail of the two device-resident pipelines: mask (device, texture size) -> seeds -> JFA -> sim force field
static int field_from_device_mask(Sim* s, uint8_t* mask_out, uint32_t* seeds_out, float* field_xy_out, float* ms_out) {
    const uint32_t w = s->settings.texture_size[0], h = s->settings.texture_size[1];
    const size_t npx = (size_t)w * h;
    if (s->jfa_pixels != npx) {
        if (s->jfa_a) cudaFree(s->jfa_a);
        if (s->jfa_b) cudaFree(s->jfa_b);
        s->jfa_a = s->jfa_b = nullptr;
        FLUID_CUDA(cudaMalloc(&s->jfa_a, npx * 4));
        FLUID_CUDA(cudaMalloc(&s->jfa_b, npx * 4));
        s->jfa_pixels = npx;
    }
    const unsigned blocks = (unsigned)((npx + 255) / 256);
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    if (ms_out) { FLUID_CUDA(cudaEventCreate(&e0)); FLUID_CUDA(cudaEventCreate(&e1)); cudaEventRecord(e0, s->stream); }
    FLUID_CUDA(cudaMemsetAsync(s->mask_flag, 0, 4, s->stream));
    k_mask_has_white<<<blocks, 256, 0, s->stream>>>(s->mask_dev, npx, s->mask_flag);
    k_seed_from_mask<<<blocks, 256, 0, s->stream>>>(s->mask_dev, w, h, s->mask_flag, s->jfa_a);
    s->launches += 2;
    uint32_t* res = jfa_run_device(s->jfa_a, s->jfa_b, w, h, s->stream, &s->launches, false, h <= w + 1u);
    k_to_force_field<<<blocks, 256, 0, s->stream>>>(res, w, h, s->force_field);
    s->launches++;
    if (ms_out) cudaEventRecord(e1, s->stream);
    void* io = nullptr;
    if (seeds_out) {
        FLUID_CUDA(cudaMalloc(&io, npx * 8));
        k_unpack_rg32<<<blocks, 256, 0, s->stream>>>(res, reinterpret_cast<uint2*>(io), npx);
        s->launches++;
        cudaMemcpyAsync(seeds_out, io, npx * 8, cudaMemcpyDeviceToHost, s->stream);
    }
    if (mask_out) cudaMemcpyAsync(mask_out, s->mask_dev, npx, cudaMemcpyDeviceToHost, s->stream);
    if (field_xy_out) cudaMemcpyAsync(field_xy_out, s->force_field, npx * 8, cudaMemcpyDeviceToHost, s->stream);
    cudaError_t e = cudaSuccess;
    if (ms_out || io || mask_out || field_xy_out) e = cudaStreamSynchronize(s->stream);  // otherwise fully asynchronous
    if (io) cudaFree(io);
    if (ms_out) {
        float ms = 0.f;
        if (e == cudaSuccess) cudaEventElapsedTime(&ms, e0, e1);
        *ms_out = ms;
        cudaEventDestroy(e0); cudaEventDestroy(e1);
    }
    FLUID_CUDA(e);
    FLUID_CUDA(cudaGetLastError());
    return FLUID_OK;
}

static int ensure_mask(Sim* s) {
    const size_t npx = (size_t)s->settings.texture_size[0] * s->settings.texture_size[1];
    if (!s->mask_dev) FLUID_CUDA(cudaMalloc(&s->mask_dev, npx));
    if (!s->mask_flag) FLUID_CUDA(cudaMalloc(&s->mask_flag, 256));
    return FLUID_OK;
}

int fluid_sim_field_from_objects(FluidSim* h_, const FluidObject* objects, uint32_t n, const float inv_proj[16],
                                 uint8_t* mask_out, uint32_t* seeds_out, float* field_xy_out, float* ms_out) {
    if (!h_ || (!objects && n) || !inv_proj) { set_error("fluid_sim_field_from_objects: null argument"); return FLUID_ERR_INVALID_ARGUMENT; }
    Sim* s = reinterpret_cast<Sim*>(h_);
    int rc = check_image(s->settings.texture_size[0], s->settings.texture_size[1]); if (rc) return rc;
    FLUID_CUDA(cudaSetDevice(s->device));
    rc = ensure_mask(s); if (rc) return rc;
    std::vector<PreparedObject> prep = prepare_objects(objects, n);
    if (n > s->objects_cap) {
        if (s->objects_dev) cudaFree(s->objects_dev);
        s->objects_dev = nullptr;
        FLUID_CUDA(cudaMalloc(&s->objects_dev, sizeof(PreparedObject) * (size_t)n));
        s->objects_cap = n;
    }
    if (n) {
        // pageable source: the copy is staged before the call returns, so `prep` may die at scope exit
        FLUID_CUDA(cudaMemcpyAsync(s->objects_dev, prep.data(), sizeof(PreparedObject) * n, cudaMemcpyHostToDevice, s->stream));
    }
    launch_raster(reinterpret_cast<const PreparedObject*>(s->objects_dev), n, inv_proj, s->settings.texture_size[0],
                  s->settings.texture_size[1], s->mask_dev, s->stream);
    s->launches++;
    return field_from_device_mask(s, mask_out, seeds_out, field_xy_out, ms_out);
}

// the same with a caller-supplied u8 mask (the reference's `Image`, src/main.rs:379-399): e.g. a video frame
int fluid_sim_field_from_mask(FluidSim* h_, const uint8_t* mask, uint32_t* seeds_out, float* field_xy_out, float* ms_out) {
    if (!h_ || !mask) { set_error("fluid_sim_field_from_mask: null argument"); return FLUID_ERR_INVALID_ARGUMENT; }
    Sim* s = reinterpret_cast<Sim*>(h_);
    int rc = check_image(s->settings.texture_size[0], s->settings.texture_size[1]); if (rc) return rc;
    FLUID_CUDA(cudaSetDevice(s->device));
    rc = ensure_mask(s); if (rc) return rc;
    const size_t npx = (size_t)s->settings.texture_size[0] * s->settings.texture_size[1];
    FLUID_CUDA(cudaMemcpyAsync(s->mask_dev, mask, npx, cudaMemcpyHostToDevice, s->stream));
    return field_from_device_mask(s, nullptr, seeds_out, field_xy_out, ms_out);
}

}  // extern "C"
