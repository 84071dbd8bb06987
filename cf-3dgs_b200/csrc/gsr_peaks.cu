// Measured ceilings for the roofline of the blend kernels ("measure with an FMA microbenchmark on the box").
// The blend kernels are bound by instruction issue on the FP32 pipe, the special-function unit and shared memory, none
// of which MEASURED_PEAKS.json covers (it holds the HBM copy bandwidth and a bf16 GEMM).  gsr_measure_peaks() runs four
// short kernels on the caller's stream and times them with its own CUDA events:
//   0  FFMA   : 16 independent fma chains per thread          -> FP32 TFLOP/s (2 flop per lane-FMA) = the issue-slot
//               ceiling too: one warp instruction per scheduler per clock
//   1  FFMA2  : 8 independent fma.rn.f32x2 chains per thread   -> packed FP32 TFLOP/s (4 flop per lane-instruction)
//   2  MUFU   : 8 independent ex2.approx chains per thread     -> G ex2/s
//   3  LDS    : ld.shared.v4 over a conflict-free 16 KB window -> shared-memory GB/s
// All four fill every SM with 8 CTAs of 256 threads and run long enough (a few hundred microseconds) to sit at the
// clock the GPU sustains for FP32 work.  No reference counterpart: measurement infrastructure of bench.py.
#include <cstdint>

#include "../../include/gsr.h"
#include "gsr_f32x2.cuh"
#include "gsr_kernels.h"

namespace gsr {
void set_last_error(const char* msg);

namespace {

template <int KIND>
__global__ void __launch_bounds__(256) peak_kernel(float* __restrict__ out, int iters, float s)
{
    __shared__ float4 win[1024];   // 16 KB
    const float x = threadIdx.x * 1e-3f + 1.0f;
    float r = 0.f;
    if (KIND == 0) {
        float a[16];
#pragma unroll
        for (int i = 0; i < 16; i++) a[i] = x + i;
        for (int it = 0; it < iters; it++) {
#pragma unroll
            for (int i = 0; i < 16; i++) a[i] = __fmaf_rn(a[i], s, x);
        }
#pragma unroll
        for (int i = 0; i < 16; i++) r += a[i];
    } else if (KIND == 1) {
        f32x2 a[8];
        const f32x2 s2 = bc(s), x2 = bc(x);
#pragma unroll
        for (int i = 0; i < 8; i++) a[i] = pk(x + i, x - i);
        for (int it = 0; it < iters; it++) {
#pragma unroll
            for (int i = 0; i < 8; i++) a[i] = fma2(a[i], s2, x2);
        }
#pragma unroll
        for (int i = 0; i < 8; i++) r += lo(a[i]) + hi(a[i]);
    } else if (KIND == 2) {
        float a[8];
#pragma unroll
        for (int i = 0; i < 8; i++) a[i] = -x * (i + 1) * 0.01f;
        for (int it = 0; it < iters; it++) {
#pragma unroll
            for (int i = 0; i < 8; i++) asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(a[i]));
#pragma unroll
            for (int i = 0; i < 8; i++) a[i] = -a[i];     // keeps the argument in (-1, 0]: one FADD-class op per ex2
        }
#pragma unroll
        for (int i = 0; i < 8; i++) r += a[i];
    } else {
        for (int i = threadIdx.x; i < 1024; i += 256) win[i] = make_float4(x, s, x, s);
        __syncthreads();
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        int p = threadIdx.x;
        for (int it = 0; it < iters; it++) {
#pragma unroll
            for (int i = 0; i < 8; i++) {
                const float4 v = win[(p + i * 128) & 1023];     // consecutive lanes, consecutive float4: conflict free
                acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
            }
            p = (p + 7) & 1023;
        }
        r = acc.x + acc.y + acc.z + acc.w;
    }
    if (r == 12345.678f) out[blockIdx.x * blockDim.x + threadIdx.x] = r;    // never true: keeps the work alive
}

}  // namespace
}  // namespace gsr

using namespace gsr;

extern "C" int gsr_measure_peaks(double* out4, void* stream)
{
    cudaStream_t s = (cudaStream_t)stream;
    if (!out4) { set_last_error("gsr_measure_peaks: out4 is NULL"); return GSR_ERR_INVALID; }
    int dev = 0, n_sm = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev);
    float* sink = nullptr;
    const int grid = n_sm * 8, block = 256;
    if (cudaMalloc(&sink, sizeof(float) * (size_t)grid * block) != cudaSuccess) {
        set_last_error("gsr_measure_peaks: cudaMalloc failed");
        return GSR_ERR_CUDA;
    }
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int iters[4] = {4000, 4000, 1500, 3000};
    // work per thread and iteration: FFMA 16 x 2 flop, FFMA2 8 x 4 flop, MUFU 8 ex2, LDS 8 x 16 bytes
    const double per_thread_iter[4] = {32.0, 32.0, 8.0, 128.0};
    for (int k = 0; k < 4; k++) {
        double best = 0.0;
        for (int rep = 0; rep < 4; rep++) {     // rep 0 warms the clocks up
            cudaEventRecord(e0, s);
            if (k == 0) peak_kernel<0><<<grid, block, 0, s>>>(sink, iters[k], 0.999f);
            if (k == 1) peak_kernel<1><<<grid, block, 0, s>>>(sink, iters[k], 0.999f);
            if (k == 2) peak_kernel<2><<<grid, block, 0, s>>>(sink, iters[k], 0.999f);
            if (k == 3) peak_kernel<3><<<grid, block, 0, s>>>(sink, iters[k], 0.999f);
            cudaEventRecord(e1, s);
            cudaEventSynchronize(e1);
            float ms = 0.f;
            cudaEventElapsedTime(&ms, e0, e1);
            const double rate = per_thread_iter[k] * iters[k] * (double)grid * block / (ms * 1e-3);
            if (rep > 0 && rate > best) best = rate;
        }
        out4[k] = best;    // flop/s, flop/s, ex2/s, bytes/s
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(sink);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { set_last_error(cudaGetErrorString(e)); return GSR_ERR_CUDA; }
    return GSR_OK;
}
