// jfa.cu — jump-flood distance / surface field.
//
// Reference: one pass = `floodfill`, shaders/jump_flood.wgsl:17-58 (sq_dist :61-65). The shader is
// compiled but never dispatched by the reference (src/simulation.rs:423-426); the host protocol
// (seed encoding, INVALID = 32767, jump schedule, ping-pong, conversion to the vec2<f32> force
// field) is the one pinned in SURVEY.md §8 row a7. The live consumer of the field is
// compute.wgsl:127-140; the CPU routine this replaces is generate_smooth_gradient_field,
// src/main.rs:403-515.
//
// B200 shape: seeds are held on the device as packed ushort2 (x | y << 16, 4 B/pixel — half the
// reference's rg32uint traffic) and expanded to rg32uint only at the boundary. Large jumps run a
// row-coalesced gather kernel (4 pixels / thread, 16-byte loads); the final jumps (<= TAIL_JUMP) are
// fused in ONE kernel that stages a tile plus halo in shared memory and runs those passes there,
// each pass reading only the previous pass's values (strict ping-pong, as the shader's two bindings).
#include <cstring>
#include <mutex>
#include <vector>

#include "common.cuh"
#include "sim_internal.h"

namespace fluid {

constexpr uint32_t JFA_INVALID_PACKED = (FLUID_JFA_INVALID << 16) | FLUID_JFA_INVALID;

__device__ __forceinline__ int sq_dist(int ax, int ay, int bx, int by) {  // jump_flood.wgsl:61-65
    const int dx = ax - bx, dy = ay - by;
    return dx * dx + dy * dy;
}

// consider neighbour seed `n` for pixel (x, y): jump_flood.wgsl:47-54
__device__ __forceinline__ void consider(uint32_t n, int x, int y, uint32_t width, uint32_t& best, int& best_d) {
    const uint32_t sx = n & 0xffffu, sy = n >> 16;
    if (sx > width || sy > width) return;  // `width` for both, `>` not `>=` (jump_flood.wgsl:48)
    const int d = sq_dist(x, y, (int)sx, (int)sy);
    if (d < best_d) { best = n; best_d = d; }
}

// Canonical seed images: every value is either JFA_INVALID_PACKED or a seed the shader would never skip
// (sx <= width and sy <= width). Then the skip test of jump_flood.wgsl:48 can be dropped: an INVALID neighbour acts
// as a phantom seed at (32767, 32767), whose distance from any pixel (x, y <= 16383) exceeds the distance to every
// admissible seed (|x - sx| <= 16384 <= 32767 - x, never with equality on both axes) and ties with an INVALID
// incumbent (same phantom point, strict <) — so it never replaces anything, exactly like the skip. Everything the
// library generates itself (mask / particle seeds on images with h <= w + 1) is canonical; caller-supplied seed
// images are checked while they are packed and fall back to the general kernels otherwise.
__device__ __forceinline__ void consider_fast(uint32_t n, int x, int y, uint32_t& best, int& best_d) {
    const int d = sq_dist(x, y, (int)(n & 0xffffu), (int)(n >> 16));
    if (d < best_d) { best = n; best_d = d; }
}

// ---- generic pass: any jump, 1 pixel per thread ------------------------------------------------
__global__ void __launch_bounds__(256) k_jfa_pass_scalar(const uint32_t* __restrict__ in, uint32_t* __restrict__ out,
                                                         uint32_t w, uint32_t h, int j) {
    const uint32_t x = blockIdx.x * 64u + (threadIdx.x & 63u);
    const uint32_t y = blockIdx.y * 4u + (threadIdx.x >> 6);
    if (x >= w || y >= h) return;
    const int xi = (int)x, yi = (int)y;
    uint32_t best = __ldg(in + (size_t)y * w + x);
    int best_d = sq_dist(xi, yi, (int)(best & 0xffffu), (int)(best >> 16));
#pragma unroll
    for (int k = 0; k < 8; k++) {  // fixed order (-j,-j) (0,-j) (j,-j) (-j,0) (j,0) (-j,j) (0,j) (j,j)
        const int kk = k < 4 ? k : k + 1;
        const int nx = xi + (kk % 3 - 1) * j, ny = yi + (kk / 3 - 1) * j;
        if (nx < 0 || ny < 0 || nx >= (int)w || ny >= (int)h) continue;
        consider(__ldg(in + (size_t)ny * w + (size_t)nx), xi, yi, w, best, best_d);
    }
    out[(size_t)y * w + x] = best;
}

// ---- vector pass: w % 4 == 0 and j % 4 == 0; 4 pixels per thread, uint4 loads -------------------
template <bool CANON>
__global__ void __launch_bounds__(256) k_jfa_pass_vec4(const uint4* __restrict__ in, uint4* __restrict__ out,
                                                       uint32_t w, uint32_t h, int j) {
    const uint32_t w4 = w >> 2;
    const uint32_t x4 = blockIdx.x * 64u + (threadIdx.x & 63u);
    const uint32_t y = blockIdx.y * 4u + (threadIdx.x >> 6);
    if (x4 >= w4 || y >= h) return;
    const int yi = (int)y, x0 = (int)(x4 << 2), j4 = j >> 2;
    uint4 own = __ldg(in + (size_t)y * w4 + x4);
    uint32_t best[4] = {own.x, own.y, own.z, own.w};
    int best_d[4];
#pragma unroll
    for (int c = 0; c < 4; c++) best_d[c] = sq_dist(x0 + c, yi, (int)(best[c] & 0xffffu), (int)(best[c] >> 16));
#pragma unroll
    for (int k = 0; k < 8; k++) {
        const int kk = k < 4 ? k : k + 1;
        const int nx4 = (int)x4 + (kk % 3 - 1) * j4, ny = yi + (kk / 3 - 1) * j;
        if (nx4 < 0 || ny < 0 || nx4 >= (int)w4 || ny >= (int)h) continue;  // whole 4-group in or out (w, j multiples of 4)
        const uint4 nb = __ldg(in + (size_t)ny * w4 + (size_t)nx4);
        if (CANON) {
            consider_fast(nb.x, x0 + 0, yi, best[0], best_d[0]);
            consider_fast(nb.y, x0 + 1, yi, best[1], best_d[1]);
            consider_fast(nb.z, x0 + 2, yi, best[2], best_d[2]);
            consider_fast(nb.w, x0 + 3, yi, best[3], best_d[3]);
        } else {
            consider(nb.x, x0 + 0, yi, w, best[0], best_d[0]);
            consider(nb.y, x0 + 1, yi, w, best[1], best_d[1]);
            consider(nb.z, x0 + 2, yi, w, best[2], best_d[2]);
            consider(nb.w, x0 + 3, yi, w, best[3], best_d[3]);
        }
    }
    out[(size_t)y * w4 + x4] = make_uint4(best[0], best[1], best[2], best[3]);
}

// ---- window pass: jumps 1 and 2 (not multiples of 4) on canonical images with w % 4 == 0 -----------------------
// 4 pixels per thread like the vector pass; per neighbour row the three aligned uint4 around the thread's own group
// form a 12-pixel window and the neighbours at x -+ J are picked from it with compile-time indices. Groups outside
// the image are replaced by INVALID, which never wins on a canonical image (= the shader's out-of-image skip).
template <int J>
__global__ void __launch_bounds__(256) k_jfa_pass_win(const uint4* __restrict__ in, uint4* __restrict__ out,
                                                      uint32_t w, uint32_t h) {
    const uint32_t w4 = w >> 2;
    const uint32_t x4 = blockIdx.x * 64u + (threadIdx.x & 63u);
    const uint32_t y = blockIdx.y * 4u + (threadIdx.x >> 6);
    if (x4 >= w4 || y >= h) return;
    const int yi = (int)y, x0 = (int)(x4 << 2);
    const uint4 inv4 = make_uint4(JFA_INVALID_PACKED, JFA_INVALID_PACKED, JFA_INVALID_PACKED, JFA_INVALID_PACKED);
    const uint4 own = __ldg(in + (size_t)y * w4 + x4);
    uint32_t best[4] = {own.x, own.y, own.z, own.w};
    int best_d[4];
#pragma unroll
    for (int c = 0; c < 4; c++) best_d[c] = sq_dist(x0 + c, yi, (int)(best[c] & 0xffffu), (int)(best[c] >> 16));
#pragma unroll
    for (int r = -1; r <= 1; r++) {  // rows y-J, y, y+J; inside a row: x-J, x (not for the own row), x+J (jump_flood.wgsl:38-46)
        const int ny = yi + r * J;
        if (ny < 0 || ny >= (int)h) continue;
        const uint4* row = in + (size_t)ny * w4;
        const uint4 lq = x4 > 0u ? __ldg(row + x4 - 1u) : inv4;
        const uint4 cq = r == 0 ? own : __ldg(row + x4);
        const uint4 rq = x4 + 1u < w4 ? __ldg(row + x4 + 1u) : inv4;
        const uint32_t win[12] = {lq.x, lq.y, lq.z, lq.w, cq.x, cq.y, cq.z, cq.w, rq.x, rq.y, rq.z, rq.w};
#pragma unroll
        for (int c = 0; c < 4; c++) {
            consider_fast(win[4 + c - J], x0 + c, yi, best[c], best_d[c]);
            if (r != 0) consider_fast(win[4 + c], x0 + c, yi, best[c], best_d[c]);
            consider_fast(win[4 + c + J], x0 + c, yi, best[c], best_d[c]);
        }
    }
    out[(size_t)y * w4 + x4] = make_uint4(best[0], best[1], best[2], best[3]);
}

// ---- fused tail: jumps J0, J0/2, .., 1 inside shared memory -------------------------------------
// Tile TX x TY of output pixels plus a halo of H = 2*J0 - 1 (sum of the jumps) on every side.
// After each in-smem pass the region whose neighbours were all available shrinks by that jump;
// image borders do not shrink it (out-of-image neighbours are skipped by the shader too).
template <int J0, int TX, int TY>
__global__ void __launch_bounds__(256) k_jfa_tail(const uint32_t* __restrict__ in, uint32_t* __restrict__ out,
                                                  uint32_t w, uint32_t h) {
    constexpr int H = 2 * J0 - 1;
    constexpr int SX = TX + 2 * H, SY = TY + 2 * H;
    extern __shared__ uint32_t s_buf[];  // 2 * SX * SY
    uint32_t* cur = s_buf;
    uint32_t* nxt = s_buf + SX * SY;
    const int ox = (int)blockIdx.x * TX - H, oy = (int)blockIdx.y * TY - H;  // image coords of smem (0,0)
    for (int t = threadIdx.x; t < SX * SY; t += 256) {
        const int sx = t % SX, sy = t / SX;
        const int gx = ox + sx, gy = oy + sy;
        uint32_t v = JFA_INVALID_PACKED;
        if (gx >= 0 && gy >= 0 && gx < (int)w && gy < (int)h) v = __ldg(in + (size_t)gy * w + (size_t)gx);
        cur[t] = v;
    }
    __syncthreads();
    int margin = 0;  // pixels of the staged region (from each non-image edge) that are no longer exact
#pragma unroll
    for (int j = J0; j >= 1; j >>= 1) {
        margin += j;
        for (int t = threadIdx.x; t < SX * SY; t += 256) {
            const int sx = t % SX, sy = t / SX;
            const int gx = ox + sx, gy = oy + sy;
            uint32_t best = cur[t];
            const bool inside_img = gx >= 0 && gy >= 0 && gx < (int)w && gy < (int)h;
            // exact only if every in-image neighbour lies inside the staged region
            const bool lo_x_ok = (sx - j >= 0) || (gx - j < 0), hi_x_ok = (sx + j < SX) || (gx + j >= (int)w);
            const bool lo_y_ok = (sy - j >= 0) || (gy - j < 0), hi_y_ok = (sy + j < SY) || (gy + j >= (int)h);
            if (inside_img && lo_x_ok && hi_x_ok && lo_y_ok && hi_y_ok) {
                int best_d = sq_dist(gx, gy, (int)(best & 0xffffu), (int)(best >> 16));
#pragma unroll
                for (int k = 0; k < 8; k++) {
                    const int kk = k < 4 ? k : k + 1;
                    const int dx = (kk % 3 - 1) * j, dy = (kk / 3 - 1) * j;
                    const int nx = gx + dx, ny = gy + dy;
                    if (nx < 0 || ny < 0 || nx >= (int)w || ny >= (int)h) continue;
                    consider(cur[(sy + dy) * SX + (sx + dx)], gx, gy, w, best, best_d);
                }
            }
            nxt[t] = best;
        }
        __syncthreads();
        uint32_t* tmp = cur; cur = nxt; nxt = tmp;
    }
    (void)margin;
    for (int t = threadIdx.x; t < TX * TY; t += 256) {
        const int tx = t % TX, ty = t / TX;
        const int gx = (int)blockIdx.x * TX + tx, gy = (int)blockIdx.y * TY + ty;
        if (gx < (int)w && gy < (int)h) out[(size_t)gy * w + (size_t)gx] = cur[(ty + H) * SX + (tx + H)];
    }
}

// ---- boundary conversions -----------------------------------------------------------------------
// noncanon (optional): set to 1 if some value is neither INVALID nor a seed the shader never skips (see consider_fast)
__global__ void k_pack_rg32(const uint2* __restrict__ rg, uint32_t* __restrict__ packed, size_t npx, uint32_t width,
                            uint32_t* __restrict__ noncanon) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= npx) return;
    uint2 v = __ldg(rg + i);
    // coordinates above 65535 cannot be packed; they can only be "invalid" markers (> width), keep them invalid
    uint32_t x = v.x > 0xffffu ? 0xffffu : v.x, y = v.y > 0xffffu ? 0xffffu : v.y;
    const uint32_t pk = x | (y << 16);
    packed[i] = pk;
    if (noncanon && pk != JFA_INVALID_PACKED && (x > width || y > width)) *noncanon = 1u;
}
__global__ void k_unpack_rg32(const uint32_t* __restrict__ packed, uint2* __restrict__ rg, size_t npx) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= npx) return;
    uint32_t v = __ldg(packed + i);
    rg[i] = make_uint2(v & 0xffffu, v >> 16);
}
__global__ void k_fill_u32(uint32_t* __restrict__ p, uint32_t v, size_t n) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}
// any pixel > 128 ? (src/main.rs:413-423)
__global__ void k_mask_has_white(const uint8_t* __restrict__ mask, size_t npx, uint32_t* __restrict__ flag) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    bool white = i < npx && mask[i] > 128;
    if (__any_sync(0xffffffffu, white) && lane_id() == 0) atomicOr(flag, 1u);
}
// seed rule of src/main.rs:413-438
__global__ void k_seed_from_mask(const uint8_t* __restrict__ mask, uint32_t w, uint32_t h, const uint32_t* __restrict__ flag,
                                 uint32_t* __restrict__ packed) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (size_t)w * h) return;
    const uint32_t x = (uint32_t)(i % w), y = (uint32_t)(i / w);
    const bool has_white = __ldg(flag) != 0u;
    const bool seed = has_white ? (mask[i] > 128) : (y == h - 1 || y == 0 || x == w - 1 || x == 0);
    packed[i] = seed ? (x | (y << 16)) : JFA_INVALID_PACKED;
}
// texel mapping of compute.wgsl:127-128 on the predicted position, clamped to tex-1 [survey-pinned §8d]
__global__ void k_seed_from_pred(const float2* __restrict__ pred, uint32_t n, float bx, float by, uint32_t w, uint32_t h,
                                 uint32_t* __restrict__ packed) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float2 q = __ldg(pred + i);
    const float uvx = fadd(fmul(fdiv(q.x, bx), 1.0f), 0.5f), uvy = fadd(fmul(fdiv(q.y, by), 1.0f), 0.5f);
    uint32_t x = f2u(fmul(uvx, (float)w)), y = f2u(fmul(uvy, (float)h));
    x = min(x, w - 1u); y = min(y, h - 1u);
    packed[(size_t)y * w + x] = x | (y << 16);  // every writer of a pixel writes the same value
}
__global__ void k_seed_from_aos(const FluidParticle* __restrict__ p, uint32_t n, float bx, float by, uint32_t w, uint32_t h,
                                uint32_t* __restrict__ packed) {
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float qx = p[i].predicted_position[0], qy = p[i].predicted_position[1];
    const float uvx = fadd(fmul(fdiv(qx, bx), 1.0f), 0.5f), uvy = fadd(fmul(fdiv(qy, by), 1.0f), 0.5f);
    uint32_t x = f2u(fmul(uvx, (float)w)), y = f2u(fmul(uvy, (float)h));
    x = min(x, w - 1u); y = min(y, h - 1u);
    packed[(size_t)y * w + x] = x | (y << 16);
}
// field = -(pixel - seed), zero when |.| <= 1e-6 (src/main.rs:493-511); still-invalid pixels -> 0
__global__ void k_to_force_field(const uint32_t* __restrict__ packed, uint32_t w, uint32_t h, float2* __restrict__ field) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (size_t)w * h) return;
    const uint32_t x = (uint32_t)(i % w), y = (uint32_t)(i / w);
    const uint32_t v = __ldg(packed + i);
    const uint32_t sx = v & 0xffffu, sy = v >> 16;
    float gx = 0.0f, gy = 0.0f;
    if (!(sx >= w || sy >= h)) {
        const float dx = fsub((float)x, (float)sx), dy = fsub((float)y, (float)sy);
        const float len = fsqrt(fadd(fmul(dx, dx), fmul(dy, dy)));
        gx = len > 1e-6f ? dx : 0.0f;
        gy = len > 1e-6f ? dy : 0.0f;
    }
    field[i] = make_float2(-gx, -gy);
}

// ---- host-side schedule ---------------------------------------------------------------------------
constexpr int TAIL_JUMP = 4;  // jumps 4, 2, 1 fused in shared memory (halo 7)
constexpr int TAIL_TX = 64, TAIL_TY = 32;

static uint32_t first_jump(uint32_t w, uint32_t h) {  // 2^(ceil(log2 max(w,h)) - 1)
    uint32_t m = w > h ? w : h, l = 0;
    while ((1u << l) < m) l++;
    return l == 0 ? 0u : (1u << (l - 1));
}

static void launch_pass(const uint32_t* in, uint32_t* out, uint32_t w, uint32_t h, uint32_t j, cudaStream_t st, bool canonical) {
    const bool vec = (w & 3u) == 0;
    if (vec && (j & 3u) == 0) {
        dim3 grid(((w >> 2) + 63) / 64, (h + 3) / 4);
        if (canonical) k_jfa_pass_vec4<true><<<grid, 256, 0, st>>>(reinterpret_cast<const uint4*>(in), reinterpret_cast<uint4*>(out), w, h, (int)j);
        else k_jfa_pass_vec4<false><<<grid, 256, 0, st>>>(reinterpret_cast<const uint4*>(in), reinterpret_cast<uint4*>(out), w, h, (int)j);
    } else if (vec && canonical && (j == 1u || j == 2u)) {
        dim3 grid(((w >> 2) + 63) / 64, (h + 3) / 4);
        if (j == 1u) k_jfa_pass_win<1><<<grid, 256, 0, st>>>(reinterpret_cast<const uint4*>(in), reinterpret_cast<uint4*>(out), w, h);
        else k_jfa_pass_win<2><<<grid, 256, 0, st>>>(reinterpret_cast<const uint4*>(in), reinterpret_cast<uint4*>(out), w, h);
    } else {
        dim3 grid((w + 63) / 64, (h + 3) / 4);
        k_jfa_pass_scalar<<<grid, 256, 0, st>>>(in, out, w, h, (int)j);
    }
}

// runs the whole schedule on packed device buffers a (input) / b (scratch); returns the buffer holding the result.
// *launches is incremented per kernel. canonical: see consider_fast. fuse_tail: run the jumps <= TAIL_JUMP in the
// shared-memory tile kernel instead of one global pass each — measured SLOWER on B200 (ncu, 4096^2: k_jfa_tail 609 us
// for 3 jumps at 85 % issue-active against 47-79 us per global pass, whose reads are L2 hits anyway: the passes are
// instruction bound, ~125 thread instructions per pixel, so tiling buys nothing and its halo costs extra work), so
// every caller passes false; the kernel stays for images where that balance differs.
uint32_t* jfa_run_device(uint32_t* a, uint32_t* b, uint32_t w, uint32_t h, cudaStream_t st, uint64_t* launches, bool fuse_tail,
                         bool canonical) {
    uint32_t* src = a;
    uint32_t* dst = b;
    static bool attr_set = false;
    constexpr int H = 2 * TAIL_JUMP - 1;
    constexpr size_t tail_smem = (size_t)2 * (TAIL_TX + 2 * H) * (TAIL_TY + 2 * H) * 4;
    if (fuse_tail && !attr_set) {
        cudaFuncSetAttribute(k_jfa_tail<TAIL_JUMP, TAIL_TX, TAIL_TY>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)tail_smem);
        attr_set = true;
    }
    for (uint32_t j = first_jump(w, h); j >= 1; j >>= 1) {
        if (fuse_tail && j == (uint32_t)TAIL_JUMP) {
            dim3 grid((w + TAIL_TX - 1) / TAIL_TX, (h + TAIL_TY - 1) / TAIL_TY);
            k_jfa_tail<TAIL_JUMP, TAIL_TX, TAIL_TY><<<grid, 256, tail_smem, st>>>(src, dst, w, h);
            if (launches) (*launches)++;
            uint32_t* t = src; src = dst; dst = t;
            break;
        }
        launch_pass(src, dst, w, h, j, st, canonical);
        if (launches) (*launches)++;
        uint32_t* t = src; src = dst; dst = t;
    }
    return src;
}

struct JfaScratch {
    uint32_t *a = nullptr, *b = nullptr;
    void* io = nullptr;
    cudaStream_t st = nullptr;
    ~JfaScratch() {
        if (a) cudaFree(a);
        if (b) cudaFree(b);
        if (io) cudaFree(io);
        if (st) cudaStreamDestroy(st);
    }
};

static int check_image(uint32_t w, uint32_t h) {
    if (w == 0 || h == 0) { set_error("jfa: empty image"); return FLUID_ERR_INVALID_ARGUMENT; }
    if (w > 16384 || h > 16384) { set_error("jfa: image larger than 16384 (INVALID = 32767 must exceed width, SURVEY.md row a7)"); return FLUID_ERR_UNSUPPORTED; }
    return FLUID_OK;
}
static int select_device(int device_ordinal) {
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || count == 0) { cudaGetLastError(); set_error("jfa: no CUDA device (no CPU fallback)"); return FLUID_ERR_NO_DEVICE; }
    if (device_ordinal < 0 || device_ordinal >= count) { set_error("jfa: device ordinal out of range"); return FLUID_ERR_NO_DEVICE; }
    FLUID_CUDA(cudaSetDevice(device_ordinal));
    return FLUID_OK;
}

}  // namespace fluid

using namespace fluid;

extern "C" {

int fluid_jfa_run(const uint32_t* seeds_in, uint32_t w, uint32_t h, uint32_t* seeds_out, int device_ordinal) {
    if (!seeds_in || !seeds_out) { set_error("fluid_jfa_run: null argument"); return FLUID_ERR_INVALID_ARGUMENT; }
    int rc = check_image(w, h); if (rc) return rc;
    rc = select_device(device_ordinal); if (rc) return rc;
    const size_t npx = (size_t)w * h;
    JfaScratch s;
    FLUID_CUDA(cudaStreamCreateWithFlags(&s.st, cudaStreamNonBlocking));
    FLUID_CUDA(cudaMalloc(&s.a, npx * 4)); FLUID_CUDA(cudaMalloc(&s.b, npx * 4)); FLUID_CUDA(cudaMalloc(&s.io, npx * 8));
    FLUID_CUDA(cudaMemcpyAsync(s.io, seeds_in, npx * 8, cudaMemcpyHostToDevice, s.st));
    const unsigned blocks = (unsigned)((npx + 255) / 256);
    FLUID_CUDA(cudaMemsetAsync(s.b, 0, 4, s.st));  // s.b doubles as the non-canonical flag before the first pass overwrites it
    k_pack_rg32<<<blocks, 256, 0, s.st>>>(reinterpret_cast<const uint2*>(s.io), s.a, npx, w, s.b);
    uint32_t noncanon = 1;
    FLUID_CUDA(cudaMemcpyAsync(&noncanon, s.b, 4, cudaMemcpyDeviceToHost, s.st));
    FLUID_CUDA(cudaStreamSynchronize(s.st));
    uint32_t* res = jfa_run_device(s.a, s.b, w, h, s.st, nullptr, false, noncanon == 0);
    k_unpack_rg32<<<blocks, 256, 0, s.st>>>(res, reinterpret_cast<uint2*>(s.io), npx);
    FLUID_CUDA(cudaMemcpyAsync(seeds_out, s.io, npx * 8, cudaMemcpyDeviceToHost, s.st));
    FLUID_CUDA(cudaStreamSynchronize(s.st));
    FLUID_CUDA(cudaGetLastError());
    return FLUID_OK;
}

int fluid_jfa_pass(const uint32_t* seeds_in, uint32_t w, uint32_t h, uint32_t jump, uint32_t* seeds_out, int device_ordinal) {
    if (!seeds_in || !seeds_out || jump == 0) { set_error("fluid_jfa_pass: bad argument"); return FLUID_ERR_INVALID_ARGUMENT; }
    int rc = check_image(w, h); if (rc) return rc;
    rc = select_device(device_ordinal); if (rc) return rc;
    const size_t npx = (size_t)w * h;
    JfaScratch s;
    FLUID_CUDA(cudaStreamCreateWithFlags(&s.st, cudaStreamNonBlocking));
    FLUID_CUDA(cudaMalloc(&s.a, npx * 4)); FLUID_CUDA(cudaMalloc(&s.b, npx * 4)); FLUID_CUDA(cudaMalloc(&s.io, npx * 8));
    FLUID_CUDA(cudaMemcpyAsync(s.io, seeds_in, npx * 8, cudaMemcpyHostToDevice, s.st));
    const unsigned blocks = (unsigned)((npx + 255) / 256);
    FLUID_CUDA(cudaMemsetAsync(s.b, 0, 4, s.st));
    k_pack_rg32<<<blocks, 256, 0, s.st>>>(reinterpret_cast<const uint2*>(s.io), s.a, npx, w, s.b);
    uint32_t noncanon = 1;
    FLUID_CUDA(cudaMemcpyAsync(&noncanon, s.b, 4, cudaMemcpyDeviceToHost, s.st));
    FLUID_CUDA(cudaStreamSynchronize(s.st));
    launch_pass(s.a, s.b, w, h, jump, s.st, noncanon == 0);
    k_unpack_rg32<<<blocks, 256, 0, s.st>>>(s.b, reinterpret_cast<uint2*>(s.io), npx);
    FLUID_CUDA(cudaMemcpyAsync(seeds_out, s.io, npx * 8, cudaMemcpyDeviceToHost, s.st));
    FLUID_CUDA(cudaStreamSynchronize(s.st));
    FLUID_CUDA(cudaGetLastError());
    return FLUID_OK;
}

int fluid_jfa_seed_from_mask(const uint8_t* mask, uint32_t w, uint32_t h, uint32_t* seeds_out, int device_ordinal) {
    if (!mask || !seeds_out) { set_error("fluid_jfa_seed_from_mask: null argument"); return FLUID_ERR_INVALID_ARGUMENT; }
    int rc = check_image(w, h); if (rc) return rc;
    rc = select_device(device_ordinal); if (rc) return rc;
    const size_t npx = (size_t)w * h;
    JfaScratch s;
    uint8_t* dmask = nullptr;
    FLUID_CUDA(cudaStreamCreateWithFlags(&s.st, cudaStreamNonBlocking));
    FLUID_CUDA(cudaMalloc(&s.a, npx * 4)); FLUID_CUDA(cudaMalloc(&s.b, 256)); FLUID_CUDA(cudaMalloc(&s.io, npx * 8));
    FLUID_CUDA(cudaMalloc(&dmask, npx));
    cudaMemcpyAsync(dmask, mask, npx, cudaMemcpyHostToDevice, s.st);
    cudaMemsetAsync(s.b, 0, 4, s.st);
    const unsigned blocks = (unsigned)((npx + 255) / 256);
    k_mask_has_white<<<blocks, 256, 0, s.st>>>(dmask, npx, s.b);
    k_seed_from_mask<<<blocks, 256, 0, s.st>>>(dmask, w, h, s.b, s.a);
    k_unpack_rg32<<<blocks, 256, 0, s.st>>>(s.a, reinterpret_cast<uint2*>(s.io), npx);
    cudaMemcpyAsync(seeds_out, s.io, npx * 8, cudaMemcpyDeviceToHost, s.st);
    cudaError_t e = cudaStreamSynchronize(s.st);
    cudaFree(dmask);
    FLUID_CUDA(e);
    FLUID_CUDA(cudaGetLastError());
    return FLUID_OK;
}

int fluid_jfa_seed_from_particles(const FluidParticle* particles, uint32_t n, const float bounds[2], uint32_t w, uint32_t h,
                                  uint32_t* seeds_out, int device_ordinal) {
    if (!particles || !bounds || !seeds_out) { set_error("fluid_jfa_seed_from_particles: null argument"); return FLUID_ERR_INVALID_ARGUMENT; }
    int rc = check_image(w, h); if (rc) return rc;
    rc = select_device(device_ordinal); if (rc) return rc;
    const size_t npx = (size_t)w * h;
    JfaScratch s;
    FluidParticle* dp = nullptr;
    FLUID_CUDA(cudaStreamCreateWithFlags(&s.st, cudaStreamNonBlocking));
    FLUID_CUDA(cudaMalloc(&s.a, npx * 4)); FLUID_CUDA(cudaMalloc(&s.io, npx * 8));
    FLUID_CUDA(cudaMalloc(&dp, (size_t)(n ? n : 1) * sizeof(FluidParticle)));
    cudaMemcpyAsync(dp, particles, (size_t)n * sizeof(FluidParticle), cudaMemcpyHostToDevice, s.st);
    const unsigned blocks = (unsigned)((npx + 255) / 256);
    k_fill_u32<<<blocks, 256, 0, s.st>>>(s.a, JFA_INVALID_PACKED, npx);
    if (n) k_seed_from_aos<<<(n + 255) / 256, 256, 0, s.st>>>(dp, n, bounds[0], bounds[1], w, h, s.a);
    k_unpack_rg32<<<blocks, 256, 0, s.st>>>(s.a, reinterpret_cast<uint2*>(s.io), npx);
    cudaMemcpyAsync(seeds_out, s.io, npx * 8, cudaMemcpyDeviceToHost, s.st);
    cudaError_t e = cudaStreamSynchronize(s.st);
    cudaFree(dp);
    FLUID_CUDA(e);
    FLUID_CUDA(cudaGetLastError());
    return FLUID_OK;
}

int fluid_jfa_to_force_field(const uint32_t* seeds, uint32_t w, uint32_t h, float* field_xy, int device_ordinal) {
    if (!seeds || !field_xy) { set_error("fluid_jfa_to_force_field: null argument"); return FLUID_ERR_INVALID_ARGUMENT; }
    int rc = check_image(w, h); if (rc) return rc;
    rc = select_device(device_ordinal); if (rc) return rc;
    const size_t npx = (size_t)w * h;
    JfaScratch s;
    FLUID_CUDA(cudaStreamCreateWithFlags(&s.st, cudaStreamNonBlocking));
    FLUID_CUDA(cudaMalloc(&s.a, npx * 4)); FLUID_CUDA(cudaMalloc(&s.io, npx * 8));
    FLUID_CUDA(cudaMemcpyAsync(s.io, seeds, npx * 8, cudaMemcpyHostToDevice, s.st));
    const unsigned blocks = (unsigned)((npx + 255) / 256);
    k_pack_rg32<<<blocks, 256, 0, s.st>>>(reinterpret_cast<const uint2*>(s.io), s.a, npx, w, nullptr);
    k_to_force_field<<<blocks, 256, 0, s.st>>>(s.a, w, h, reinterpret_cast<float2*>(s.io));
    FLUID_CUDA(cudaMemcpyAsync(field_xy, s.io, npx * 8, cudaMemcpyDeviceToHost, s.st));
    FLUID_CUDA(cudaStreamSynchronize(s.st));
    FLUID_CUDA(cudaGetLastError());
    return FLUID_OK;
}

int fluid_sim_jfa_from_particles(FluidSim* h_, uint32_t w, uint32_t h, uint32_t* seeds_out, float* field_xy_out, int into_sim,
                                 float* ms_out) {
    if (!h_) { set_error("null FluidSim handle"); return FLUID_ERR_INVALID_ARGUMENT; }
    Sim* s = reinterpret_cast<Sim*>(h_);
    int rc = check_image(w, h); if (rc) return rc;
    FLUID_CUDA(cudaSetDevice(s->device));
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
    k_fill_u32<<<blocks, 256, 0, s->stream>>>(s->jfa_a, JFA_INVALID_PACKED, npx);
    k_seed_from_pred<<<(s->n + 255) / 256, 256, 0, s->stream>>>(s->pred_s, s->n, s->settings.size[0], s->settings.size[1], w, h, s->jfa_a);
    s->launches += 2;
    cudaEvent_t e0, e1;
    FLUID_CUDA(cudaEventCreate(&e0)); FLUID_CUDA(cudaEventCreate(&e1));
    cudaEventRecord(e0, s->stream);
    // seeds from particles are their own in-image coordinates: canonical unless the shader's `> width` test can reject
    // a row index (images taller than wide)
    uint32_t* res = jfa_run_device(s->jfa_a, s->jfa_b, w, h, s->stream, &s->launches, false, h <= w + 1u);
    cudaEventRecord(e1, s->stream);
    void* io = nullptr;
    const bool want_io = seeds_out || field_xy_out;
    if (want_io) FLUID_CUDA(cudaMalloc(&io, npx * 8));
    if (seeds_out) {
        k_unpack_rg32<<<blocks, 256, 0, s->stream>>>(res, reinterpret_cast<uint2*>(io), npx);
        cudaMemcpyAsync(seeds_out, io, npx * 8, cudaMemcpyDeviceToHost, s->stream);
        s->launches++;
    }
    const bool install = into_sim && w == s->settings.texture_size[0] && h == s->settings.texture_size[1];
    if (install) {
        k_to_force_field<<<blocks, 256, 0, s->stream>>>(res, w, h, s->force_field);
        s->launches++;
        if (field_xy_out) cudaMemcpyAsync(field_xy_out, s->force_field, npx * 8, cudaMemcpyDeviceToHost, s->stream);
    } else if (field_xy_out) {
        k_to_force_field<<<blocks, 256, 0, s->stream>>>(res, w, h, reinterpret_cast<float2*>(io));
        s->launches++;
        cudaMemcpyAsync(field_xy_out, io, npx * 8, cudaMemcpyDeviceToHost, s->stream);
    }
    cudaError_t e = cudaStreamSynchronize(s->stream);
    if (io) cudaFree(io);
    float ms = 0.f;
    if (e == cudaSuccess) cudaEventElapsedTime(&ms, e0, e1);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    if (ms_out) *ms_out = ms;
    FLUID_CUDA(e);
    FLUID_CUDA(cudaGetLastError());
    if (into_sim && !install) { set_error("fluid_sim_jfa_from_particles: into_sim requested but (w,h) != texture_size"); return FLUID_ERR_INVALID_ARGUMENT; }
    return FLUID_OK;
}

}  // extern "C"

// ---- fluid_jfa_run_static: the jump flood for callers that may not allocate ---------------------------------------
// No counterpart in the reference (see include/fluid_b200.h).
// The ping-pong images and the boundary staging image of fluid_jfa_run live in ONE file-scope device array instead of
// three cudaMalloc'ed buffers. Layout (bytes): [0, 4M) packed image a | [4M, 8M) packed image b | [8M, 16M) rg32uint
// staging image, M = FLUID_JFA_STATIC_MAX_PIXELS. Every GPU that loads the library holds the array until exit.
// In HOST code `g_jfa_static` is only a handle: its device address comes from cudaGetSymbolAddress, once per device,
// with that device current (jfa_static_base); kernels are given that address, never the symbol.
namespace fluid {

constexpr size_t JFA_STATIC_BYTES = (size_t)FLUID_JFA_STATIC_MAX_PIXELS * 16;
static __device__ __align__(256) unsigned char g_jfa_static[JFA_STATIC_BYTES];

constexpr int JFA_STATIC_MAX_DEVICES = 64;
static std::mutex g_jfa_static_mutex;
static unsigned char* g_jfa_static_base[JFA_STATIC_MAX_DEVICES] = {};

// device address of g_jfa_static on `device` (which must be the current device)
static int jfa_static_base(int device, unsigned char** out) {
    if (device < 0 || device >= JFA_STATIC_MAX_DEVICES) { set_error("fluid_jfa_run_static: device ordinal out of range"); return FLUID_ERR_NO_DEVICE; }
    std::lock_guard<std::mutex> lock(g_jfa_static_mutex);
    if (!g_jfa_static_base[device]) {
        void* p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_jfa_static);
        if (e != cudaSuccess || !p) {
            cudaGetLastError();
            set_error(std::string("fluid_jfa_run_static: cudaGetSymbolAddress(g_jfa_static) failed: ") + cudaGetErrorString(e));
            return FLUID_ERR_CUDA;
        }
        g_jfa_static_base[device] = static_cast<unsigned char*>(p);
    }
    *out = g_jfa_static_base[device];
    return FLUID_OK;
}

// can a kernel running on the current device dereference p? (device / managed memory, or mapped pinned host memory;
// *dev receives the address to hand to the kernel)
static bool jfa_device_reachable(const void* p, const void** dev) {
    cudaPointerAttributes a{};
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
    if (a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged) { *dev = p; return true; }
    if (a.type == cudaMemoryTypeHost && a.devicePointer) { *dev = a.devicePointer; return true; }
    return false;
}

}  // namespace fluid

extern "C" int fluid_jfa_run_static(const uint32_t* seeds_in, uint32_t w, uint32_t h, uint32_t* seeds_out, int device_ordinal,
                                    void* stream, int wait) {
    if (!seeds_in || !seeds_out) { set_error("fluid_jfa_run_static: null argument"); return FLUID_ERR_INVALID_ARGUMENT; }
    int rc = check_image(w, h); if (rc) return rc;
    const size_t npx = (size_t)w * h;
    if (npx > (size_t)FLUID_JFA_STATIC_MAX_PIXELS) {
        set_error("fluid_jfa_run_static: image of " + std::to_string(w) + " x " + std::to_string(h) + " = " + std::to_string(npx) +
                  " pixels exceeds FLUID_JFA_STATIC_MAX_PIXELS = " + std::to_string(FLUID_JFA_STATIC_MAX_PIXELS) +
                  " (the static scratch is fixed at build time; use fluid_jfa_run for larger images)");
        return FLUID_ERR_UNSUPPORTED;
    }
    rc = select_device(device_ordinal); if (rc) return rc;
    unsigned char* base = nullptr;
    rc = jfa_static_base(device_ordinal, &base); if (rc) return rc;
    uint32_t* a = reinterpret_cast<uint32_t*>(base);
    uint32_t* b = reinterpret_cast<uint32_t*>(base + (size_t)FLUID_JFA_STATIC_MAX_PIXELS * 4);
    uint2* io = reinterpret_cast<uint2*>(base + (size_t)FLUID_JFA_STATIC_MAX_PIXELS * 8);
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const unsigned blocks = (unsigned)((npx + 255) / 256);

    // inputs / outputs the kernels can reach are used where they lie; plain host memory goes through the staging image
    const void* in_dev = nullptr;
    const void* out_dev = nullptr;
    const bool in_direct = jfa_device_reachable(seeds_in, &in_dev);
    const bool out_direct = jfa_device_reachable(seeds_out, &out_dev);
    if (!in_direct) {
        FLUID_CUDA(cudaMemcpyAsync(io, seeds_in, npx * 8, cudaMemcpyHostToDevice, st));
        in_dev = io;
    }
    // Nothing in the scratch has to start at zero: a[0, npx) is written whole by the pack kernel, every pass writes every
    // pixel of its destination before the next pass reads it, and the staging image is written whole before it is read.
    // The canonical-image fast path of fluid_jfa_run needs a flag read back by the host between two launches; this entry
    // point never waits on the device before its last launch, so it always runs the general kernels (same results).
    k_pack_rg32<<<blocks, 256, 0, st>>>(static_cast<const uint2*>(in_dev), a, npx, w, nullptr);
    uint32_t* res = jfa_run_device(a, b, w, h, st, nullptr, false, false);
    k_unpack_rg32<<<blocks, 256, 0, st>>>(res, out_direct ? static_cast<uint2*>(const_cast<void*>(out_dev)) : io, npx);
    FLUID_CUDA(cudaGetLastError());
    if (!out_direct) FLUID_CUDA(cudaMemcpyAsync(seeds_out, io, npx * 8, cudaMemcpyDeviceToHost, st));
    if (wait || !out_direct || !in_direct) FLUID_CUDA(cudaStreamSynchronize(st));
    return FLUID_OK;
}

#include "field.inl"
