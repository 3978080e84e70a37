// render_oracle.cpp — CPU ORACLE (TEST INFRASTRUCTURE, NOT PRODUCT CODE), part 2: the reference's per-pixel
// shaders restated as plain C++ loops. Same rules as fluid_oracle.cpp (IEEE binary32, no contraction, correctly
// rounded / and sqrt). PARITY UNPINNED for the same reason: the WGSL cannot run here.
//
//   oracle_rasterize_objects  image_shader.wgsl:32-85 (vs_main, fs_main, point_in_rect) -> R8Uint mask
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

}  // extern "C"
