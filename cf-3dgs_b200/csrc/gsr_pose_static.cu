// gsr_pose_gradient_static(): the rigid-pose gradient reduction for callers that must not allocate.
//
// Who this is for.  gsr_backward() can already produce the pose gradient, but it wants a caller-provided
// scratch buffer, and some callers have no way to make one:
//   * a tracking / SLAM front-end that drives this library from C on a thread with a fixed time slice:
//     cudaMalloc and cudaFree take the driver's process-wide lock and synchronise the device, so one call in
//     the pose loop (300 iterations per frame pair in CF-3DGS) can stall every other stream for milliseconds;
//   * a process whose GPU memory was claimed up front -- a framework's caching allocator that has grown to
//     the card's limit, or a fixed MPS / MIG partition: there a cudaMalloc of even a few hundred KB can fail,
//     and the host language may have no handle on the framework's allocator to ask it instead.
// For them the scratch lives inside the library image, in the file-scope __device__ array below.  The entry
// point performs no cudaMalloc, cudaFree or any other allocating call; inputs and the 6-float output are
// still the caller's.
//
// What it costs.  Every GPU that loads libgsr_b200.so carries the array (384 KiB) until the process exits,
// whether or not this entry point is ever called.  The array bounds the problem size:
// P <= GSR_POSE_STATIC_MAX_POINTS (4,194,304 points); larger P is refused with GSR_ERR_INVALID.
//
// Sharing.  There is one array per device and every call on that device uses it.  It still holds whatever
// the previous call left there; that is harmless here because nothing is accumulated into it -- stage 1
// overwrites slots [0, ceil(P/256)) and stage 2 reads exactly those slots, so there is nothing that has to be
// zeroed first.  Two calls that overlap on the same GPU (different streams, or different host threads) race on
// the array and give wrong sums; serialise them on one stream or with events.
//
// First call.  The first call on a device resolves the array's device address with cudaGetSymbolAddress
// (in host code the name `g_pose_partials` is only a stand-in for the symbol, not an address a kernel could
// use) and caches it.  Make that first call during start-up, with the device you will use current, and before
// beginning any stream capture.  The address is cached until the process exits: a cudaDeviceReset after the first
// call leaves a stale address behind (nothing here can notice a reset), so do not reset a device this has run on.
//
// Inputs.  means3D are the points the rasterizer saw and dL_dmeans3D the gradient with respect to them.  After the
// rigid entry points (gsr_backward_rigid with [R | t]) those are p' = R p + t and g' = R g for the g that call
// returned, not the raw p and g.
#include <cstdio>
#include <mutex>

#include "../../include/gsr.h"
#include "gsr_kernels.h"

namespace gsr {
void set_last_error(const char* msg);   // gsr_api.cu
}

namespace {

constexpr int PS_THREADS = 256;
constexpr int PS_MAX_BLOCKS = GSR_POSE_STATIC_MAX_POINTS / PS_THREADS;   // 16384
constexpr int PS_MAX_DEVICES = 64;

// [block][6] partial sums: 16384 * 6 floats = 384 KiB
static __device__ __align__(256) float g_pose_partials[PS_MAX_BLOCKS * 6];

float* g_partials_addr[PS_MAX_DEVICES] = {nullptr};
std::mutex g_setup_mutex;

__device__ __forceinline__ float block_sum(float v, float* red)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < PS_THREADS / 32; w++) t += red[w];
    __syncthreads();
    return t;
}

// stage 1: per-block sums of g and p x g, written (not accumulated) to partials[block][0..5].
// `partials` is the address cudaGetSymbolAddress returned for g_pose_partials on this device.
__global__ void __launch_bounds__(PS_THREADS)
pose_partials_kernel(int P, const float* __restrict__ means3D, const float* __restrict__ dL_dmeans3D,
                     float* __restrict__ partials)
{
    __shared__ float red[PS_THREADS / 32];
    const int i = blockIdx.x * PS_THREADS + threadIdx.x;
    float px = 0.f, py = 0.f, pz = 0.f, gx = 0.f, gy = 0.f, gz = 0.f;
    if (i < P) {
        px = means3D[3 * i]; py = means3D[3 * i + 1]; pz = means3D[3 * i + 2];
        gx = dL_dmeans3D[3 * i]; gy = dL_dmeans3D[3 * i + 1]; gz = dL_dmeans3D[3 * i + 2];
    }
    const float v0 = block_sum(gx, red), v1 = block_sum(gy, red), v2 = block_sum(gz, red);
    const float w0 = block_sum(py * gz - pz * gy, red);
    const float w1 = block_sum(pz * gx - px * gz, red);
    const float w2 = block_sum(px * gy - py * gx, red);
    if (threadIdx.x == 0) {
        float* o = partials + (size_t)blockIdx.x * 6;
        o[0] = v0; o[1] = v1; o[2] = v2; o[3] = w0; o[4] = w1; o[5] = w2;
    }
}

// stage 2: one block adds the nblk partials in double.  Device code may name the array directly.
__global__ void __launch_bounds__(PS_THREADS)
pose_total_kernel(int nblk, float* __restrict__ out)
{
    __shared__ double red[PS_THREADS / 32];
    for (int c = 0; c < 6; c++) {
        double v = 0.0;
        for (int i = threadIdx.x; i < nblk; i += PS_THREADS) v += (double)g_pose_partials[(size_t)i * 6 + c];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
        __syncthreads();
        if (threadIdx.x == 0) {
            double t = 0.0;
            for (int w = 0; w < PS_THREADS / 32; w++) t += red[w];
            out[c] = (float)t;
        }
        __syncthreads();
    }
}

int fail(int code, const char* msg)
{
    gsr::set_last_error(msg);
    return code;
}

}  // namespace

extern "C" int gsr_pose_gradient_static(int P, const float* means3D, const float* dL_dmeans3D, float* dL_dpose,
                                        void* stream)
{
    cudaStream_t s = (cudaStream_t)stream;
    char msg[256];
    if (P < 0) return fail(GSR_ERR_INVALID, "gsr_pose_gradient_static: P is negative");
    if (P > GSR_POSE_STATIC_MAX_POINTS) {
        snprintf(msg, sizeof(msg),
                 "gsr_pose_gradient_static: P=%d exceeds the built-in scratch (GSR_POSE_STATIC_MAX_POINTS=%d); "
                 "use gsr_backward with a caller-provided pose_scratch instead", P, GSR_POSE_STATIC_MAX_POINTS);
        return fail(GSR_ERR_INVALID, msg);
    }
    if (!dL_dpose) return fail(GSR_ERR_INVALID, "gsr_pose_gradient_static: dL_dpose is NULL");
    if (P > 0 && (!means3D || !dL_dmeans3D))
        return fail(GSR_ERR_INVALID, "gsr_pose_gradient_static: means3D / dL_dmeans3D is NULL");

    // one-time set-up per device: turn the symbol into an address kernels can take
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= PS_MAX_DEVICES) {
        snprintf(msg, sizeof(msg), "gsr_pose_gradient_static: cudaGetDevice: %s (device %d)", cudaGetErrorString(e), dev);
        return fail(GSR_ERR_CUDA, msg);
    }
    float* partials = nullptr;
    {
        std::lock_guard<std::mutex> lock(g_setup_mutex);
        if (!g_partials_addr[dev]) {
            float* p = nullptr;
            e = cudaGetSymbolAddress((void**)&p, g_pose_partials);
            if (e != cudaSuccess || !p) {
                snprintf(msg, sizeof(msg), "gsr_pose_gradient_static: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
                return fail(GSR_ERR_CUDA, msg);
            }
            g_partials_addr[dev] = p;
        }
        partials = g_partials_addr[dev];
    }

    const int nblk = (P + PS_THREADS - 1) / PS_THREADS;
    if (nblk > 0) pose_partials_kernel<<<nblk, PS_THREADS, 0, s>>>(P, means3D, dL_dmeans3D, partials);
    pose_total_kernel<<<1, PS_THREADS, 0, s>>>(nblk, dL_dpose);   // nblk == 0 writes six zeros
    e = cudaGetLastError();
    if (e != cudaSuccess) {
        snprintf(msg, sizeof(msg), "gsr_pose_gradient_static: launch: %s", cudaGetErrorString(e));
        return fail(GSR_ERR_CUDA, msg);
    }
    return GSR_OK;
}
