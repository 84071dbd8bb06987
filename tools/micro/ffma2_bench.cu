// Microbenchmark: FFMA vs FFMA2 (fma.rn.f32x2) issue / pipe throughput on sm_100a, alone and mixed with integer ALU work.
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o ffma2_bench ffma2_bench.cu
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float a, float b) { u64 r; asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ u64 ffma2(u64 a, u64 b, u64 c) { u64 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ float lo(u64 v) { float a, b; asm("mov.b64 {%0,%1}, %2;" : "=f"(a), "=f"(b) : "l"(v)); return a + b; }

template <int MODE, int ALU>
__global__ void __launch_bounds__(256) k(float* out, int iters, float s, int m)
{
    float x = threadIdx.x * 1e-3f;
    int q = threadIdx.x + m;
    int q0 = q, q1 = q ^ 5, q2 = q + 9, q3 = q * 3;
    if (MODE == 0) {
        float a[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) a[i] = x + i;
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int i = 0; i < 16; ++i) a[i] = fmaf(a[i], s, x);
            if (ALU) {
#pragma unroll
                for (int i = 0; i < ALU; ++i) { q0 = (q0 ^ m) + q1; q1 = (q1 ^ m) + q2; q2 = (q2 ^ m) + q3; q3 = (q3 ^ m) + q0; }
            }
        }
        float r = 0;
#pragma unroll
        for (int i = 0; i < 16; ++i) r += a[i];
        out[blockIdx.x * blockDim.x + threadIdx.x] = r + q0 + q1 + q2 + q3;
    } else {
        u64 a[8];
        const u64 s2 = pk(s, s), x2 = pk(x, x);
#pragma unroll
        for (int i = 0; i < 8; ++i) a[i] = pk(x + i, x - i);
        for (int it = 0; it < iters; ++it) {
#pragma unroll
            for (int i = 0; i < 8; ++i) a[i] = ffma2(a[i], s2, x2);
            if (ALU) {
#pragma unroll
                for (int i = 0; i < ALU; ++i) { q0 = (q0 ^ m) + q1; q1 = (q1 ^ m) + q2; q2 = (q2 ^ m) + q3; q3 = (q3 ^ m) + q0; }
            }
        }
        float r = 0;
#pragma unroll
        for (int i = 0; i < 8; ++i) r += lo(a[i]);
        out[blockIdx.x * blockDim.x + threadIdx.x] = r + q0 + q1 + q2 + q3;
    }
}

template <int MODE, int ALU>
void run(const char* name)
{
    float* out; cudaMalloc(&out, 148 * 8 * 256 * 4);
    const int iters = 20000;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<MODE, ALU><<<148 * 8, 256>>>(out, 100, 0.999f, 3);
    cudaEventRecord(e0);
    k<MODE, ALU><<<148 * 8, 256>>>(out, iters, 0.999f, 3);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double fma = 16.0 * iters * 148 * 8 * 256;
    printf("%-28s %8.3f ms  %7.2f TFLOP/s fp32  (%d int ops per 16 fma)\n", name, ms, 2 * fma / ms * 1e-9, ALU * 8);
    cudaFree(out);
}
int main()
{
    run<0, 0>("FFMA");  run<1, 0>("FFMA2");
    run<0, 1>("FFMA + 8 ALU");  run<1, 1>("FFMA2 + 8 ALU");
    run<0, 2>("FFMA + 16 ALU"); run<1, 2>("FFMA2 + 16 ALU");
    run<0, 4>("FFMA + 32 ALU"); run<1, 4>("FFMA2 + 32 ALU");
    cudaError_t e = cudaDeviceSynchronize();
    printf("status: %s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
