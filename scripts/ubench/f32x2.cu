// micro-benchmark: issue rate of scalar FADD/FMUL/FFMA against the packed FADD2/FMUL2/FFMA2 of sm_100a.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/ubench_f32x2 scripts/ubench/f32x2.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

#define ITERS 4096
template <int MODE, int CHAINS>
__global__ void __launch_bounds__(256) k(float* out, float seed) {
    float a[CHAINS * 2];
#pragma unroll
    for (int c = 0; c < CHAINS * 2; c++) a[c] = seed + threadIdx.x * 1e-3f + c;
    const float m = 1.0000001f, b = 1e-7f;
    unsigned long long mm, bb;
    asm("mov.b64 %0, {%1, %1};" : "=l"(mm) : "f"(m));
    asm("mov.b64 %0, {%1, %1};" : "=l"(bb) : "f"(b));
#pragma unroll 1
    for (int it = 0; it < ITERS; it++) {
#pragma unroll
        for (int c = 0; c < CHAINS; c++) {
            if (MODE == 0) {  // scalar fma: 2 instr for 2 values
                a[2 * c] = __fmaf_rn(a[2 * c], m, b);
                a[2 * c + 1] = __fmaf_rn(a[2 * c + 1], m, b);
            } else if (MODE == 1) {  // packed fma: 1 instr for 2 values
                unsigned long long v;
                asm("mov.b64 %0, {%1, %2};" : "=l"(v) : "f"(a[2 * c]), "f"(a[2 * c + 1]));
                asm("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(v) : "l"(mm), "l"(bb));
                asm("mov.b64 {%0, %1}, %2;" : "=f"(a[2 * c]), "=f"(a[2 * c + 1]) : "l"(v));
            } else if (MODE == 2) {  // scalar add
                a[2 * c] = __fadd_rn(a[2 * c], b);
                a[2 * c + 1] = __fadd_rn(a[2 * c + 1], b);
            } else if (MODE == 3) {  // packed add
                unsigned long long v;
                asm("mov.b64 %0, {%1, %2};" : "=l"(v) : "f"(a[2 * c]), "f"(a[2 * c + 1]));
                asm("add.rn.f32x2 %0, %0, %1;" : "+l"(v) : "l"(bb));
                asm("mov.b64 {%0, %1}, %2;" : "=f"(a[2 * c]), "=f"(a[2 * c + 1]) : "l"(v));
            } else if (MODE == 4) {  // scalar mul
                a[2 * c] = __fmul_rn(a[2 * c], m);
                a[2 * c + 1] = __fmul_rn(a[2 * c + 1], m);
            } else if (MODE == 5) {  // packed mul
                unsigned long long v;
                asm("mov.b64 %0, {%1, %2};" : "=l"(v) : "f"(a[2 * c]), "f"(a[2 * c + 1]));
                asm("mul.rn.f32x2 %0, %0, %1;" : "+l"(v) : "l"(mm));
                asm("mov.b64 {%0, %1}, %2;" : "=f"(a[2 * c]), "=f"(a[2 * c + 1]) : "l"(v));
            } else if (MODE == 6) {  // mixed: packed fma + as many integer ALU ops (issue-slot competition)
                unsigned long long v;
                asm("mov.b64 %0, {%1, %2};" : "=l"(v) : "f"(a[2 * c]), "f"(a[2 * c + 1]));
                asm("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(v) : "l"(mm), "l"(bb));
                asm("mov.b64 {%0, %1}, %2;" : "=f"(a[2 * c]), "=f"(a[2 * c + 1]) : "l"(v));
            }
        }
    }
    float s = 0;
#pragma unroll
    for (int c = 0; c < CHAINS * 2; c++) s += a[c];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int MODE, int CHAINS>
void run(const char* name, float* d) {
    int sms = 148, ctas = sms * 4;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<MODE, CHAINS><<<ctas, 256>>>(d, 1.0f);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    for (int r = 0; r < 5; r++) k<MODE, CHAINS><<<ctas, 256>>>(d, 1.0f);
    cudaEventRecord(e1);
    cudaDeviceSynchronize();
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    ms /= 5;
    const double lane_ops = (double)ctas * 256 * ITERS * CHAINS * 2;
    printf("%-14s chains=%d  %.3f ms  %.1f Glane-ops/s  (%.1f lane-ops/clk/SM at 1.9 GHz)\n", name, CHAINS, ms, lane_ops / ms * 1e-6,
           lane_ops / (ms * 1e-3) / 148 / 1.9e9);
}

int main() {
    float* d;
    cudaMalloc(&d, 148 * 4 * 256 * sizeof(float));
    run<0, 8>("ffma", d);  run<1, 8>("ffma2", d);
    run<2, 8>("fadd", d);  run<3, 8>("fadd2", d);
    run<4, 8>("fmul", d);  run<5, 8>("fmul2", d);
    run<0, 4>("ffma", d);  run<1, 4>("ffma2", d);
    cudaError_t e = cudaDeviceSynchronize();
    printf("status: %s\n", cudaGetErrorString(e));
    return 0;
}
