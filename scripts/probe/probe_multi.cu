// probe_multi.cu — what the multi-GPU transport of libfluid_b200 can rely on, measured on the box:
//   1. peer access in one process: pull kernel (P2P loads) bandwidth/latency, flag ping-pong latency between two GPUs
//   2. NCCL from a plain C++ program: grouped send/recv latency for window-sized messages
//   3. CUDA IPC between two processes: memory handle open + flag signalling
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o probe_multi probe_multi.cu -lnccl
#include <cuda_runtime.h>
#include <nccl.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <unistd.h>
#include <sys/wait.h>
#include <chrono>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA fail %s line %d: %s\n", #x, __LINE__, cudaGetErrorString(e)); exit(1);} } while (0)
#define NK(x) do { ncclResult_t e = (x); if (e != ncclSuccess) { printf("NCCL fail %s line %d: %s\n", #x, __LINE__, ncclGetErrorString(e)); exit(1);} } while (0)

__global__ void k_pull(const uint4* __restrict__ src, uint4* __restrict__ dst, size_t n16) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x) dst[i] = src[i];
}
__global__ void k_signal(volatile uint32_t* flag, uint32_t v) { __threadfence_system(); *flag = v; }
__global__ void k_wait(volatile uint32_t* flag, uint32_t v) {
    while (true) {
        uint32_t x;
        asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(x) : "l"(flag));
        if ((int32_t)(x - v) >= 0) break;
        __nanosleep(100);
    }
}
static double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

static void probe_peer() {
    int can01 = 0, can10 = 0;
    CK(cudaDeviceCanAccessPeer(&can01, 0, 1)); CK(cudaDeviceCanAccessPeer(&can10, 1, 0));
    printf("[peer] canAccess 0->1 %d, 1->0 %d\n", can01, can10);
    if (!can01 || !can10) return;
    CK(cudaSetDevice(0)); CK(cudaDeviceEnablePeerAccess(1, 0));
    CK(cudaSetDevice(1)); CK(cudaDeviceEnablePeerAccess(0, 0));
    const size_t bytes = 1300000 / 16 * 16;
    void *a0, *b1; uint32_t *f0, *f1;
    cudaStream_t s0, s1;
    CK(cudaSetDevice(0)); CK(cudaMalloc(&a0, 64 << 20)); CK(cudaMalloc(&f0, 256)); CK(cudaMemset(f0, 0, 256)); CK(cudaStreamCreateWithFlags(&s0, cudaStreamNonBlocking));
    CK(cudaSetDevice(1)); CK(cudaMalloc(&b1, 64 << 20)); CK(cudaMalloc(&f1, 256)); CK(cudaMemset(f1, 0, 256)); CK(cudaStreamCreateWithFlags(&s1, cudaStreamNonBlocking));
    CK(cudaDeviceSynchronize());
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    for (size_t sz : {bytes, (size_t)(64 << 20)}) {
        for (int blocks : {8, 32, 148}) {
            for (int w = 0; w < 3; w++) k_pull<<<blocks, 256, 0, s1>>>((const uint4*)a0, (uint4*)b1, sz / 16);
            CK(cudaEventRecord(e0, s1));
            const int it = 20;
            for (int i = 0; i < it; i++) k_pull<<<blocks, 256, 0, s1>>>((const uint4*)a0, (uint4*)b1, sz / 16);
            CK(cudaEventRecord(e1, s1)); CK(cudaStreamSynchronize(s1));
            float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
            printf("[peer] pull kernel GPU1<-GPU0 %zu B, %d blocks: %.2f us per call, %.1f GB/s\n", sz, blocks, 1e3 * ms / it, sz * it / (ms * 1e-3) / 1e9);
        }
        CK(cudaEventRecord(e0, s1));
        for (int i = 0; i < 20; i++) CK(cudaMemcpyPeerAsync(b1, 1, a0, 0, sz, s1));
        CK(cudaEventRecord(e1, s1)); CK(cudaStreamSynchronize(s1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        printf("[peer] cudaMemcpyPeerAsync %zu B: %.2f us per call, %.1f GB/s\n", sz, 1e3 * ms / 20, sz * 20 / (ms * 1e-3) / 1e9);
    }
    // flag ping-pong: GPU0 signals f1 (on GPU1) with value 2i+1, GPU1 waits, signals f0 with 2i+2, GPU0 waits
    const int rounds = 200;
    double t0 = now();
    for (int i = 0; i < rounds; i++) {
        CK(cudaSetDevice(0));
        k_signal<<<1, 1, 0, s0>>>(f1, 2 * i + 1);
        k_wait<<<1, 1, 0, s0>>>(f0, 2 * i + 2);
        CK(cudaSetDevice(1));
        k_wait<<<1, 1, 0, s1>>>(f1, 2 * i + 1);
        k_signal<<<1, 1, 0, s1>>>(f0, 2 * i + 2);
    }
    CK(cudaSetDevice(0)); CK(cudaStreamSynchronize(s0)); CK(cudaSetDevice(1)); CK(cudaStreamSynchronize(s1));
    printf("[peer] flag ping-pong (signal kernel + wait kernel each way): %.2f us per round trip (%d rounds, includes launch)\n", 1e6 * (now() - t0) / rounds, rounds);
    // same-device two-stream flag wait (single-GPU multi-band tests): does a spinning kernel let another stream run?
    CK(cudaSetDevice(0));
    cudaStream_t s0b; CK(cudaStreamCreateWithFlags(&s0b, cudaStreamNonBlocking));
    CK(cudaMemset(f0, 0, 4)); CK(cudaDeviceSynchronize());
    k_wait<<<1, 1, 0, s0>>>(f0, 7);
    usleep(2000);
    k_signal<<<1, 1, 0, s0b>>>(f0, 7);
    t0 = now();
    cudaError_t q = cudaErrorNotReady;
    while (now() - t0 < 2.0 && (q = cudaStreamQuery(s0)) == cudaErrorNotReady) usleep(100);
    printf("[peer] same-device cross-stream spin wait: %s\n", q == cudaSuccess ? "released (ok)" : "STUCK after 2 s");
    if (q != cudaSuccess) { uint32_t v = 7; cudaMemcpyAsync((void*)f0, &v, 4, cudaMemcpyHostToDevice, s0b); cudaDeviceSynchronize(); }
}

static void probe_nccl() {
    int devs[2] = {0, 1};
    ncclComm_t comms[2];
    double t0 = now();
    NK(ncclCommInitAll(comms, 2, devs));
    int ver; ncclGetVersion(&ver);
    printf("[nccl] version %d, ncclCommInitAll(2) %.2f s\n", ver, now() - t0);
    void* sb[2]; void* rb[2]; cudaStream_t st[2]; cudaEvent_t e0, e1;
    for (int r = 0; r < 2; r++) { CK(cudaSetDevice(r)); CK(cudaMalloc(&sb[r], 8 << 20)); CK(cudaMalloc(&rb[r], 8 << 20)); CK(cudaStreamCreateWithFlags(&st[r], cudaStreamNonBlocking)); }
    CK(cudaSetDevice(0)); CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    for (size_t bytes : {(size_t)64, (size_t)1300000, (size_t)(8 << 20)}) {
        const int it = 50;
        for (int phase = 0; phase < 2; phase++) {
            if (phase == 1) { CK(cudaSetDevice(0)); CK(cudaEventRecord(e0, st[0])); }
            double th = now();
            for (int i = 0; i < it; i++) {
                NK(ncclGroupStart());
                for (int r = 0; r < 2; r++) { NK(ncclSend(sb[r], bytes, ncclChar, 1 - r, comms[r], st[r])); NK(ncclRecv(rb[r], bytes, ncclChar, 1 - r, comms[r], st[r])); }
                NK(ncclGroupEnd());
            }
            double host = now() - th;
            if (phase == 1) {
                CK(cudaSetDevice(0)); CK(cudaEventRecord(e1, st[0]));
                for (int r = 0; r < 2; r++) { CK(cudaSetDevice(r)); CK(cudaStreamSynchronize(st[r])); }
                float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
                printf("[nccl] grouped send+recv %zu B each way: %.2f us per exchange on the device, %.2f us host enqueue\n", bytes, 1e3 * ms / it, 1e6 * host / it);
            } else for (int r = 0; r < 2; r++) { CK(cudaSetDevice(r)); CK(cudaStreamSynchronize(st[r])); }
        }
    }
    for (int r = 0; r < 2; r++) ncclCommDestroy(comms[r]);
}

static void probe_ipc() {
    int p2c[2], c2p[2];
    if (pipe(p2c) || pipe(c2p)) { printf("[ipc] pipe failed\n"); return; }
    pid_t pid = fork();
    if (pid == 0) {  // child: GPU 1
        CK(cudaSetDevice(1));
        cudaIpcMemHandle_t hd, hf;
        if (read(p2c[0], &hd, sizeof hd) != sizeof hd || read(p2c[0], &hf, sizeof hf) != sizeof hf) _exit(3);
        void* data = nullptr; uint32_t* flags = nullptr;
        cudaError_t e = cudaIpcOpenMemHandle(&data, hd, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) { printf("[ipc] child: cudaIpcOpenMemHandle failed: %s\n", cudaGetErrorString(e)); char c = 'F'; (void)!write(c2p[1], &c, 1); _exit(0); }
        CK(cudaIpcOpenMemHandle((void**)&flags, hf, cudaIpcMemLazyEnablePeerAccess));
        void* mine; CK(cudaMalloc(&mine, 4 << 20));
        cudaStream_t s; CK(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
        char c = 'R'; (void)!write(c2p[1], &c, 1);
        // wait for parent's flag[0] = 1 (written by a kernel on GPU0 into its own memory), pull 1 MB, check, signal flag[1] = 1 (P2P store)
        k_wait<<<1, 1, 0, s>>>(flags, 1);
        k_pull<<<32, 256, 0, s>>>((const uint4*)data, (uint4*)mine, (1 << 20) / 16);
        k_signal<<<1, 1, 0, s>>>(flags + 1, 1);
        CK(cudaStreamSynchronize(s));
        uint32_t first[4]; CK(cudaMemcpy(first, mine, 16, cudaMemcpyDeviceToHost));
        printf("[ipc] child pulled over IPC mapping: first words %08x %08x (want deadbeef)\n", first[0], first[1]);
        // ping-pong latency across processes: child waits 2i+2 on flags[0], signals 2i+3 on flags[1]
        for (int i = 0; i < 200; i++) { k_wait<<<1, 1, 0, s>>>(flags, 2 * i + 2); k_signal<<<1, 1, 0, s>>>(flags + 1, 2 * i + 3); }
        CK(cudaStreamSynchronize(s));
        _exit(0);
    }
    CK(cudaSetDevice(0));
    void* data; uint32_t* flags;
    CK(cudaMalloc(&data, 4 << 20)); CK(cudaMalloc(&flags, 256)); CK(cudaMemset(flags, 0, 256));
    CK(cudaMemset(data, 0, 4 << 20));
    cudaIpcMemHandle_t hd, hf;
    cudaError_t e = cudaIpcGetMemHandle(&hd, data);
    if (e != cudaSuccess) { printf("[ipc] cudaIpcGetMemHandle failed: %s\n", cudaGetErrorString(e)); kill(pid, 9); return; }
    CK(cudaIpcGetMemHandle(&hf, flags));
    (void)!write(p2c[1], &hd, sizeof hd); (void)!write(p2c[1], &hf, sizeof hf);
    char c = 0; (void)!read(c2p[0], &c, 1);
    if (c != 'R') { printf("[ipc] child could not open the handle\n"); waitpid(pid, nullptr, 0); return; }
    cudaStream_t s; CK(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    uint32_t pat[2] = {0xdeadbeefu, 0xdeadbeefu};
    CK(cudaMemcpyAsync(data, pat, 8, cudaMemcpyHostToDevice, s));
    k_signal<<<1, 1, 0, s>>>(flags, 1);
    k_wait<<<1, 1, 0, s>>>(flags + 1, 1);
    CK(cudaStreamSynchronize(s));
    double t0 = now();
    for (int i = 0; i < 200; i++) { k_signal<<<1, 1, 0, s>>>(flags, 2 * i + 2); k_wait<<<1, 1, 0, s>>>(flags + 1, 2 * i + 3); }
    CK(cudaStreamSynchronize(s));
    printf("[ipc] cross-process flag ping-pong: %.2f us per round trip\n", 1e6 * (now() - t0) / 200);
    int status = 0; waitpid(pid, &status, 0);
    printf("[ipc] child exit status %d\n", WEXITSTATUS(status));
}

int main(int argc, char** argv) {
    int n = 0; CK(cudaGetDeviceCount(&n));
    printf("devices: %d\n", n);
    if (n < 2) { printf("need 2 GPUs\n"); return 0; }
    const char* what = argc > 1 ? argv[1] : "all";
    if (!strcmp(what, "ipc") || !strcmp(what, "all")) { probe_ipc(); }   // fork BEFORE this process creates a context: done inside (cudaGetDeviceCount does not create one)
    if (!strcmp(what, "peer") || !strcmp(what, "all")) probe_peer();
    if (!strcmp(what, "nccl") || !strcmp(what, "all")) probe_nccl();
    return 0;
}
