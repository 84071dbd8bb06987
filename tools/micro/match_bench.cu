// Microbenchmark: throughput of MATCH.ANY as a function of the number of distinct values in the warp, against the
// 8-ballot formulation of the same mask.  Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o match_bench match_bench.cu
#include <cstdio>
#include <cuda_runtime.h>
template <int MODE>
__global__ void __launch_bounds__(256) k(unsigned* out, int iters, int distinct)
{
    const unsigned lane = threadIdx.x & 31;
    unsigned acc = 0;
    unsigned v = (lane * 2654435761u >> 7) % distinct;
    for (int it = 0; it < iters; ++it) {
        const unsigned d = (v + it) & 255u;     // same number of distinct values every iteration
        unsigned m;
        if (MODE == 0) {
            m = __match_any_sync(0xffffffffu, d);
        } else {
            m = 0xffffffffu;
#pragma unroll
            for (int b = 0; b < 8; ++b) {
                const unsigned bal = __ballot_sync(0xffffffffu, (d >> b) & 1u);
                m &= ((d >> b) & 1u) ? bal : ~bal;
            }
        }
        acc += __popc(m & ((1u << lane) - 1u)) + (m >> 31);
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}
template <int MODE>
void run(const char* name, int distinct)
{
    unsigned* out; cudaMalloc(&out, 148 * 8 * 256 * 4);
    const int iters = 4000;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<MODE><<<148 * 8, 256>>>(out, 10, distinct);
    cudaEventRecord(e0);
    k<MODE><<<148 * 8, 256>>>(out, iters, distinct);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double warp_ops = (double)iters * 148 * 8 * 8;
    printf("%-8s distinct %2d: %7.3f ms  %6.2f G warp-ops/s  (%.1f cycles per op per SM sub-partition at 1.9 GHz)\n", name, distinct, ms,
           warp_ops / ms * 1e-6, ms * 1e-3 * 1.9e9 / (warp_ops / (148 * 4)));
    cudaFree(out);
}
int main()
{
    for (int d : {1, 2, 4, 8, 16, 32}) run<0>("match", d);
    for (int d : {1, 32}) run<1>("ballot8", d);
    cudaError_t e = cudaDeviceSynchronize();
    printf("status: %s\n", cudaGetErrorString(e));
    return e != cudaSuccess;
}
