// Throughput probe: FFMA vs FFMA2 (fma.rn.f32x2) and how many issue slots FFMA2 leaves free.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/ffma2_probe tools/ffma2_probe.cu
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float lo, float hi) { u64 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void upk(u64 v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) { u64 d; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ float fma1(float a, float b, float c) { float d; asm volatile("fma.rn.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c)); return d; }
__device__ __forceinline__ float fmin3v(float a, float b, float c) { float d; asm volatile("min.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c)); return d; }
__device__ __forceinline__ float fmin2v(float a, float b) { float d; asm volatile("min.f32 %0, %1, %2;" : "=f"(d) : "f"(a), "f"(b)); return d; }

// MODE 0: 16 FFMA chains   1: 8 FFMA2 chains (same flops)   2: 8 FFMA2 + 8 min   3: 16 FFMA + 8 min
// MODE 4: 8 FFMA2 + 4 min3  5: 8 FFMA2 + 8 FFMA
template <int MODE>
__global__ void __launch_bounds__(256) probe(float* out, int iters, float s) {
    float a[16];
    u64 p[8];
    float mn[8];
    for (int i = 0; i < 16; i++) a[i] = threadIdx.x * 0.001f + i;
    for (int i = 0; i < 8; i++) { p[i] = pk(a[2 * i], a[2 * i + 1]); mn[i] = 3e38f; }
    const u64 S = pk(s, s), T = pk(0.5f, 0.25f);
    for (int t = 0; t < iters; t++) {
        if (MODE == 0 || MODE == 3) {
#pragma unroll
            for (int i = 0; i < 16; i++) a[i] = fma1(a[i], s, 0.5f);
        }
        if (MODE == 1 || MODE == 2 || MODE == 4 || MODE == 5) {
#pragma unroll
            for (int i = 0; i < 8; i++) p[i] = fma2(p[i], S, T);
        }
        if (MODE == 5) {
#pragma unroll
            for (int i = 0; i < 8; i++) a[i] = fma1(a[i], s, 0.5f);
        }
        if (MODE == 2) {
#pragma unroll
            for (int i = 0; i < 8; i++) { float lo, hi; upk(p[i], lo, hi); mn[i] = fmin2v(mn[i], lo); }
        }
        if (MODE == 3) {
#pragma unroll
            for (int i = 0; i < 8; i++) mn[i] = fmin2v(mn[i], a[i]);
        }
        if (MODE == 4) {
#pragma unroll
            for (int i = 0; i < 4; i++) { float lo, hi; upk(p[i], lo, hi); mn[i] = fmin3v(mn[i], lo, hi); }
        }
    }
    float r = 0.f;
    for (int i = 0; i < 16; i++) r += a[i];
    for (int i = 0; i < 8; i++) { float lo, hi; upk(p[i], lo, hi); r += lo + hi + mn[i]; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = r;
}

template <int MODE>
void run(const char* name, double flop_per_iter, float* d) {
    const int iters = 20000, blocks = 148 * 8;
    probe<MODE><<<blocks, 256>>>(d, 100, 0.999f);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0);
    probe<MODE><<<blocks, 256>>>(d, iters, 0.999f);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms;
    cudaEventElapsedTime(&ms, e0, e1);
    double fl = flop_per_iter * iters * (double)blocks * 256;
    printf("%-28s %8.3f ms  %7.2f TFLOP/s\n", name, ms, fl / ms * 1e-9);
}

int main() {
    float* d;
    cudaMalloc(&d, 148 * 8 * 256 * 4);
    run<0>("16 FFMA", 32, d);
    run<1>("8 FFMA2", 32, d);
    run<2>("8 FFMA2 + 8 FMNMX", 32, d);
    run<3>("16 FFMA + 8 FMNMX", 32, d);
    run<4>("8 FFMA2 + 4 FMNMX3", 32, d);
    run<5>("8 FFMA2 + 8 FFMA", 48, d);
    cudaError_t e = cudaDeviceSynchronize();
    printf("%s\n", cudaGetErrorString(e));
    return 0;
}
