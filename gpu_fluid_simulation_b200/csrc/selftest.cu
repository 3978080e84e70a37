// selftest.cu — device self-tests of arithmetic identities the exact-math kernels rely on.
//
// fluid_selftest_division: the shared-reciprocal quotient of common.cuh (rcp_refined + div_by), which the
// force pass uses for the 9 divisions of compute.wgsl:206-223 / :283-288 / funcs.wgsl:112-123, must equal
// the correctly rounded quotient (__fdiv_rn, == the CPU restatement's `/`) for every operand pair in its
// stated range: b in [2^-20, 2^20], a zero or |a| in [2^-100, 2^100]. The sweep takes EVERY significand of
// b (2^23) at several exponents, and per b a set of numerators: random ones over the whole exponent range,
// plus the adversarial ones a = RN(b * (q + ulp/2)) and a = RN(b * q) whose quotients sit next to rounding
// boundaries.
#include "common.cuh"

namespace fluid {

__device__ __forceinline__ uint32_t mix32(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
    return x;
}

__global__ void k_selftest_division(uint32_t seed, int exp_b, int per_b, unsigned long long* counters, float* first_bad) {
    const uint32_t m = blockIdx.x * blockDim.x + threadIdx.x;  // significand of b
    if (m >= (1u << 23)) return;
    const float b = __uint_as_float(((uint32_t)(exp_b + 127) << 23) | m);
    const float y = rcp_refined(b);
    unsigned long long bad = 0, bad_rcp = (y != __frcp_rn(b));
    uint32_t s = mix32(m ^ seed);
    for (int k = 0; k < per_b; k++) {
        s = mix32(s + 0x9e3779b9u);
        float a;
        const int kind = k & 3;
        if (kind == 0) {  // random numerator over the whole admitted exponent range, random sign
            const int ea = (int)(mix32(s ^ 0x51u) % 201u) - 100;
            a = __uint_as_float((s & 0x807fffffu) | ((uint32_t)(ea + 127) << 23));
        } else {
            // quotient q with a random significand and small exponent, numerator placed next to a boundary
            const int eq = (int)(mix32(s ^ 0x77u) % 61u) - 30;
            const float q = __uint_as_float((s & 0x007fffffu) | ((uint32_t)(eq + 127) << 23));
            if (kind == 1) a = __fmul_rn(b, q);                                              // just off a representable q
            else if (kind == 2) a = __fmaf_rn(b, q, __fmul_rn(__fmul_rn(b, q), 5.9604645e-8f));  // ~ b*(q + ulp/2)
            else a = __fmaf_rn(b, q, -__fmul_rn(__fmul_rn(b, q), 2.9802322e-8f));             // ~ b*(q - ulp/4)
            if (s & 0x80000000u) a = -a;
        }
        const float want = __fdiv_rn(a, b);
        const float got = div_by(a, b, y);
        if (__float_as_uint(want) != __float_as_uint(got)) {
            if (!bad && atomicAdd(&counters[3], 1ull) == 0) { first_bad[0] = a; first_bad[1] = b; first_bad[2] = want; first_bad[3] = got; }
            bad++;
        }
    }
    // warp-aggregate the counters
    for (int o = 16; o > 0; o >>= 1) {
        bad += __shfl_xor_sync(0xffffffffu, bad, o);
        bad_rcp += __shfl_xor_sync(0xffffffffu, bad_rcp, o);
    }
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(&counters[0], (unsigned long long)per_b * 32ull);
        if (bad) atomicAdd(&counters[1], bad);
        if (bad_rcp) atomicAdd(&counters[2], bad_rcp);
    }
}

}  // namespace fluid

extern "C" int fluid_selftest_division(int device_ordinal, uint32_t seed, int per_b, uint64_t* tested, uint64_t* mismatches,
                                       uint64_t* inexact_reciprocals, float* first_bad /* a, b, want, got */) {
    using namespace fluid;
    if (!tested || !mismatches || per_b <= 0) { set_error("fluid_selftest_division: bad arguments"); return FLUID_ERR_INVALID_ARGUMENT; }
    FLUID_CUDA(cudaSetDevice(device_ordinal));
    unsigned long long* d_counters = nullptr;
    float* d_bad = nullptr;
    FLUID_CUDA(cudaMalloc(&d_counters, 4 * sizeof(unsigned long long)));
    FLUID_CUDA(cudaMalloc(&d_bad, 4 * sizeof(float)));
    FLUID_CUDA(cudaMemset(d_counters, 0, 4 * sizeof(unsigned long long)));
    FLUID_CUDA(cudaMemset(d_bad, 0, 4 * sizeof(float)));
    const int exps[] = {-20, -17, -11, -4, -3, -1, 0, 1, 2, 6, 7, 13, 19};  // b = 2^e * 1.m; 2^20 itself is the m = 0 end point
    for (int e : exps) k_selftest_division<<<(1u << 23) / 256, 256>>>(seed + (uint32_t)e * 977u, e, per_b, d_counters, d_bad);
    FLUID_CUDA(cudaGetLastError());
    FLUID_CUDA(cudaDeviceSynchronize());
    unsigned long long h[4];
    float hb[4];
    FLUID_CUDA(cudaMemcpy(h, d_counters, sizeof(h), cudaMemcpyDeviceToHost));
    FLUID_CUDA(cudaMemcpy(hb, d_bad, sizeof(hb), cudaMemcpyDeviceToHost));
    cudaFree(d_counters);
    cudaFree(d_bad);
    *tested = h[0];
    *mismatches = h[1];
    if (inexact_reciprocals) *inexact_reciprocals = h[2];
    if (first_bad) for (int i = 0; i < 4; i++) first_bad[i] = hb[i];
    return FLUID_OK;
}
