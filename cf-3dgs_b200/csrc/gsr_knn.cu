// Mean squared distance to the three nearest neighbours of every point: the reference's simple-knn `distCUDA2`
// (submodules/simple-knn/spatial.cu:15-26 -> simple_knn.cu:185-221), used once per Gaussian-model initialisation to
// size the initial splats (scene/gaussian_model_cf.py:194, scene/gaussian_model.py:130).
//
// What the reference computes: for every point the three smallest values of
//   d2 = fma(dz, dz, fma(dx, dx, rn(dy*dy)))          (its sm_100a SASS; d = other - self, self excluded by index)
// over all other points, then rn(rn(rn(b0 + b1) + b2) / 3).  Its Morton boxes only prune the search, conservatively up
// to rounding, so the result is the exact 3-nearest-neighbour answer (tests/: brute force on the CPU).
//
// Reference: two blocking D2H copies for the bounding box, cudaMalloc + five thrust::device_vector per call, CUB sort,
// a 10-barrier tree reduction per box, and a search in which every thread tests every box and then walks each
// surviving box through `points[indices[i]]` -- one dependent, uncoalesced gather per candidate.
// Here: no allocation and no host synchronisation (caller-owned scratch, bounding box kept on the device), the
// library's own radix sort on the 30-bit Morton codes, points re-ordered ONCE into Morton order as float4 (xyz, index),
// and a block-cooperative search: the 256 threads of a block own 256 consecutive sorted points (spatially compact), the
// block first compacts the list of candidate boxes against its own bounding box grown by its largest rejection radius
// (boxes are tested 256 at a time, one per thread), then stages each candidate box in shared memory with coalesced
// 16-byte loads and lets every thread that still needs it scan it from there.
//
// Each thread still performs the reference's own sequence of decisions: same Morton codes (same bounding box rule,
// same arithmetic), same stable sort, same 1024-point boxes visited in ascending order, the same box-distance and
// `dist > reject || dist > best[2]` tests with the same roundings.  The block-level candidate list only removes boxes
// that every thread of the block would have skipped (its test is the thread test relaxed by 0.1 %).  So even in the
// razor-edge cases where rounding makes the reference's pruning drop a true neighbour, the result is the reference's.
#include <cfloat>

#include "gsr_kernels.h"

namespace gsr {

constexpr int KNN_BOX = 1024;      // points per box (the reference's BOX_SIZE, simple_knn.cu:12)
constexpr int KNN_THREADS = 256;   // a block searches for a quarter of a box
constexpr int KNN_MAX_CAND = 1024; // candidate boxes kept per pass of the compaction

struct KnnView {
    uint32_t* bbox;        // 8 words: min xyz, max xyz as order-preserving uint32
    uint32_t* codes;       // P Morton codes
    uint32_t* keys[2];     // sort ping-pong
    uint32_t* vals[2];
    uint32_t* hist;
    uint32_t* digit_total;
    float4* sorted;        // P points in Morton order: x, y, z, bits(original index)
    float4* boxes;         // 2 per box: (min xyz, -), (max xyz, -)
};

size_t knn_layout(char* base, int P, KnnView& v)
{
    char* p = base;
    const size_t n = (size_t)(P > 0 ? P : 1);
    const size_t nblk = (size_t)div_up((int64_t)n, SORT_CHUNK);
    const size_t nbox = (size_t)div_up((int64_t)n, KNN_BOX);
    carve(p, v.bbox, 8);
    carve(p, v.codes, n);
    carve(p, v.keys[0], n);
    carve(p, v.keys[1], n);
    carve(p, v.vals[0], n);
    carve(p, v.vals[1], n);
    carve(p, v.hist, (size_t)RADIX * nblk);
    carve(p, v.digit_total, SORT_AUX_WORDS);
    carve(p, v.sorted, n);
    carve(p, v.boxes, 2 * nbox);
    return (size_t)(p - base) + 256;
}

size_t knn_scratch_bytes(int P)
{
    KnnView v;
    return knn_layout(nullptr, P, v);
}

// float <-> uint32 that preserves order (so atomicMin/atomicMax work on floats of either sign)
__device__ __forceinline__ uint32_t f2ord(float f)
{
    const uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ord2f(uint32_t o)
{
    return __uint_as_float((o & 0x80000000u) ? (o & 0x7fffffffu) : ~o);
}

// The reference reduces with init = (0,0,0) (simple_knn.cu:192-200), so its box always contains the origin.  Only the
// Morton codes depend on that, not the result; mirrored anyway so both implementations partition space alike.
__global__ void knn_bbox_init_kernel(uint32_t* bbox)
{
    if (threadIdx.x < 8) bbox[threadIdx.x] = f2ord(0.f);
}

__global__ void __launch_bounds__(256)
knn_bbox_kernel(int P, const float* __restrict__ pts, uint32_t* __restrict__ bbox)
{
    float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < P; i += gridDim.x * blockDim.x) {
#pragma unroll
        for (int c = 0; c < 3; c++) {
            const float v = pts[3 * (size_t)i + c];
            mn[c] = fminf(mn[c], v);
            mx[c] = fmaxf(mx[c], v);
        }
    }
#pragma unroll
    for (int c = 0; c < 3; c++) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            mn[c] = fminf(mn[c], __shfl_xor_sync(0xffffffffu, mn[c], o));
            mx[c] = fmaxf(mx[c], __shfl_xor_sync(0xffffffffu, mx[c], o));
        }
    }
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int c = 0; c < 3; c++) {
            atomicMin(&bbox[c], f2ord(mn[c]));
            atomicMax(&bbox[4 + c], f2ord(mx[c]));
        }
    }
}

__device__ __forceinline__ uint32_t spread10(uint32_t x)   // simple_knn.cu:44-51
{
    x = (x | (x << 16)) & 0x030000FFu;
    x = (x | (x << 8)) & 0x0300F00Fu;
    x = (x | (x << 4)) & 0x030C30C3u;
    x = (x | (x << 2)) & 0x09249249u;
    return x;
}

__global__ void __launch_bounds__(256)
knn_morton_kernel(int P, const float* __restrict__ pts, const uint32_t* __restrict__ bbox, uint32_t* __restrict__ codes)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    uint32_t q[3];
#pragma unroll
    for (int c = 0; c < 3; c++) {
        const float lo = ord2f(bbox[c]), hi = ord2f(bbox[4 + c]);
        // simple_knn.cu:55-57: ((v - min) / (max - min)) * 1023, truncated; clamped here so that a degenerate axis
        // (max == min) or a NaN gives a valid cell instead of an undefined float->uint conversion
        const float t = ((pts[3 * (size_t)i + c] - lo) / (hi - lo)) * 1023.0f;
        q[c] = (uint32_t)fminf(fmaxf(t, 0.f), 1023.f);   // fmaxf(NaN, 0) = 0
    }
    codes[i] = spread10(q[0]) | (spread10(q[1]) << 1) | (spread10(q[2]) << 2);
}

// sorted[i] = (point[order[i]], order[i]); one block = one box of 1024 points: its bounding box by warp shuffles
__global__ void __launch_bounds__(KNN_THREADS)
knn_gather_kernel(int P, const float* __restrict__ pts, const uint32_t* __restrict__ order, float4* __restrict__ sorted,
                  float4* __restrict__ boxes)
{
    __shared__ float s_mn[KNN_THREADS / 32][3], s_mx[KNN_THREADS / 32][3];
    float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
#pragma unroll
    for (int k = 0; k < KNN_BOX / KNN_THREADS; k++) {
        const int i = blockIdx.x * KNN_BOX + k * KNN_THREADS + threadIdx.x;
        if (i < P) {
            const uint32_t src = order[i];
            const float x = pts[3 * (size_t)src], y = pts[3 * (size_t)src + 1], z = pts[3 * (size_t)src + 2];
            sorted[i] = make_float4(x, y, z, __uint_as_float(src));
            mn[0] = fminf(mn[0], x); mx[0] = fmaxf(mx[0], x);
            mn[1] = fminf(mn[1], y); mx[1] = fmaxf(mx[1], y);
            mn[2] = fminf(mn[2], z); mx[2] = fmaxf(mx[2], z);
        }
    }
#pragma unroll
    for (int c = 0; c < 3; c++) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            mn[c] = fminf(mn[c], __shfl_xor_sync(0xffffffffu, mn[c], o));
            mx[c] = fmaxf(mx[c], __shfl_xor_sync(0xffffffffu, mx[c], o));
        }
    }
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) {
#pragma unroll
        for (int c = 0; c < 3; c++) { s_mn[warp][c] = mn[c]; s_mx[warp][c] = mx[c]; }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
#pragma unroll
        for (int w = 1; w < KNN_THREADS / 32; w++) {
#pragma unroll
            for (int c = 0; c < 3; c++) { mn[c] = fminf(mn[c], s_mn[w][c]); mx[c] = fmaxf(mx[c], s_mx[w][c]); }
        }
        boxes[2 * blockIdx.x] = make_float4(mn[0], mn[1], mn[2], 0.f);
        boxes[2 * blockIdx.x + 1] = make_float4(mx[0], mx[1], mx[2], 0.f);
    }
}

// squared distance with the reference's rounding sequence (other - self; y squared and rounded, then x and z fused:
// nvcc rounds the right product of `dx*dx + dy*dy` and fuses the left, simple_knn.cu:131-134)
__device__ __forceinline__ float knn_d2(float ox, float oy, float oz, float sx, float sy, float sz)
{
    const float dx = __fadd_rn(ox, -sx), dy = __fadd_rn(oy, -sy), dz = __fadd_rn(oz, -sz);
    return __fmaf_rn(dz, dz, __fmaf_rn(dx, dx, __fmul_rn(dy, dy)));
}

// keep the three smallest of {b0 <= b1 <= b2, d}; a NaN distance is never inserted (simple_knn.cu:135-142 compares
// `knn[j] > dist`, false for NaN)
__device__ __forceinline__ void knn_insert(float d, float& b0, float& b1, float& b2)
{
    d = (d == d) ? d : FLT_MAX;
    const float t1 = fmaxf(b0, d);
    b0 = fminf(b0, d);
    const float t2 = fmaxf(b1, t1);
    b1 = fminf(b1, t1);
    b2 = fminf(b2, t2);
}

// squared distance from a point to a box (simple_knn.cu:119-129), 0 inside; same predicates and the same rounding
// sequence as the reference's build (ey squared and rounded, then ex and ez fused)
__device__ __forceinline__ float knn_box_d2(float4 bmin, float4 bmax, float x, float y, float z)
{
    float ex = 0.f, ey = 0.f, ez = 0.f;
    if (x < bmin.x || x > bmax.x) ex = fminf(fabsf(__fadd_rn(x, -bmin.x)), fabsf(__fadd_rn(x, -bmax.x)));
    if (y < bmin.y || y > bmax.y) ey = fminf(fabsf(__fadd_rn(y, -bmin.y)), fabsf(__fadd_rn(y, -bmax.y)));
    if (z < bmin.z || z > bmax.z) ez = fminf(fabsf(__fadd_rn(z, -bmin.z)), fabsf(__fadd_rn(z, -bmax.z)));
    return __fmaf_rn(ez, ez, __fmaf_rn(ex, ex, __fmul_rn(ey, ey)));
}

__global__ void __launch_bounds__(KNN_THREADS)
knn_search_kernel(int P, const float4* __restrict__ sorted, const float4* __restrict__ boxes, int nbox,
                  float* __restrict__ mean_d2)
{
    __shared__ float4 s_pts[KNN_BOX];
    __shared__ uint32_t s_cand[KNN_MAX_CAND];
    __shared__ int s_ncand;
    __shared__ int s_wc[KNN_THREADS / 32];
    __shared__ uint32_t s_reject_bits;
    __shared__ uint32_t s_q[6];     // the block's own bounding box (order-preserving encoding)
    const int i = blockIdx.x * KNN_THREADS + threadIdx.x;
    const bool active = i < P;
    const unsigned lane = threadIdx.x & 31;
    float4 me = make_float4(0.f, 0.f, 0.f, 0.f);
    if (active) me = sorted[i];

    // rejection radius: third-best among the +-3 neighbours along the Morton order (simple_knn.cu:153-161)
    float b0 = FLT_MAX, b1 = FLT_MAX, b2 = FLT_MAX;
    if (active) {
        for (int k = max(0, i - 3); k <= min(P - 1, i + 3); k++) {
            if (k == i) continue;
            const float4 o = sorted[k];
            knn_insert(knn_d2(o.x, o.y, o.z, me.x, me.y, me.z), b0, b1, b2);
        }
    }
    const float reject = b2;   // FLT_MAX when P < 4: nothing is pruned
    b0 = b1 = b2 = FLT_MAX;

    // block-wide: the largest rejection radius and the bounding box of the block's points
    if (threadIdx.x == 0) s_reject_bits = 0u;
    if (threadIdx.x < 3) { s_q[threadIdx.x] = 0xffffffffu; s_q[3 + threadIdx.x] = 0u; }
    __syncthreads();
    if (active) {
        atomicMax(&s_reject_bits, __float_as_uint(reject));   // reject >= 0: bit patterns order like the values
        atomicMin(&s_q[0], f2ord(me.x)); atomicMax(&s_q[3], f2ord(me.x));
        atomicMin(&s_q[1], f2ord(me.y)); atomicMax(&s_q[4], f2ord(me.y));
        atomicMin(&s_q[2], f2ord(me.z)); atomicMax(&s_q[5], f2ord(me.z));
    }
    __syncthreads();
    const float blk_reject = __uint_as_float(s_reject_bits);
    const float qx0 = ord2f(s_q[0]), qy0 = ord2f(s_q[1]), qz0 = ord2f(s_q[2]);
    const float qx1 = ord2f(s_q[3]), qy1 = ord2f(s_q[4]), qz1 = ord2f(s_q[5]);

    int next = 0;   // first box not yet considered
    while (next < nbox) {
        // candidate boxes, in ascending order: gap between the two boxes within the largest radius (shrunk by 0.1 %
        // so rounding can only add candidates).  A NaN anywhere keeps the box.
        if (threadIdx.x == 0) s_ncand = 0;
        __syncthreads();
        while (next < nbox && s_ncand <= KNN_MAX_CAND - KNN_THREADS) {
            const int mine = next + threadIdx.x;
            bool take = false;
            if (mine < nbox) {
                const float4 cmin = boxes[2 * mine], cmax = boxes[2 * mine + 1];
                const float gx = fmaxf(0.f, fmaxf(cmin.x - qx1, qx0 - cmax.x));
                const float gy = fmaxf(0.f, fmaxf(cmin.y - qy1, qy0 - cmax.y));
                const float gz = fmaxf(0.f, fmaxf(cmin.z - qz1, qz0 - cmax.z));
                take = !((gx * gx + gy * gy + gz * gz) * 0.999f > blk_reject);
            }
            // append in thread order so the list stays ascending
            const unsigned m = __ballot_sync(0xffffffffu, take);
            const int warp = (int)(threadIdx.x >> 5);
            if (lane == 0) s_wc[warp] = __popc(m);
            __syncthreads();
            int base = s_ncand, tot = 0;
#pragma unroll
            for (int w = 0; w < KNN_THREADS / 32; w++) {
                const int c = s_wc[w];
                if (w < warp) base += c;
                tot += c;
            }
            if (take) s_cand[base + __popc(m & ((1u << lane) - 1u))] = (uint32_t)mine;
            __syncthreads();            // everyone has read s_ncand and s_wc
            if (threadIdx.x == 0) s_ncand += tot;
            __syncthreads();
            next += KNN_THREADS;
        }
        const int ncand = s_ncand;

        for (int c = 0; c < ncand; c++) {
            const int cb = (int)s_cand[c];
            const float4 cmin = boxes[2 * cb], cmax = boxes[2 * cb + 1];
            bool need = false;
            if (active) {
                const float d = knn_box_d2(cmin, cmax, me.x, me.y, me.z);
                need = fminf(reject, b2) >= d;   // simple_knn.cu:171-172: skip if dist > reject || dist > best[2]
            }
            if (!__syncthreads_or(need)) continue;   // (also: the previous box is fully consumed)
            const int j0 = cb * KNN_BOX;
            const int cnt = min(KNN_BOX, P - j0);
            for (int k = threadIdx.x; k < cnt; k += KNN_THREADS) s_pts[k] = sorted[j0 + k];
            __syncthreads();
            if (need) {
                const int self = i - j0;   // position of this thread's own point in the box, if it is there
#pragma unroll 4
                for (int k = 0; k < cnt; k++) {
                    const float4 o = s_pts[k];
                    const float d = knn_d2(o.x, o.y, o.z, me.x, me.y, me.z);
                    knn_insert(k == self ? FLT_MAX : d, b0, b1, b2);
                }
            }
        }
        __syncthreads();
    }
    if (active) mean_d2[__float_as_uint(me.w)] = __fdiv_rn(__fadd_rn(__fadd_rn(b0, b1), b2), 3.0f);
}

int launch_knn_mean_dist2(int P, const float* points, float* mean_d2, void* scratch, cudaStream_t s)
{
    if (P <= 0) return 0;
    KnnView v;
    knn_layout((char*)scratch, P, v);
    const int nbox = div_up(P, KNN_BOX);
    knn_bbox_init_kernel<<<1, 32, 0, s>>>(v.bbox);
    knn_bbox_kernel<<<min(div_up(P, 256), 148 * 8), 256, 0, s>>>(P, points, v.bbox);
    knn_morton_kernel<<<div_up(P, 256), 256, 0, s>>>(P, points, v.bbox, v.codes);
    const int cur = radix_sort_pairs(v.keys, v.vals, P, 0, 30, v.hist, v.digit_total, s, nullptr, v.codes);
    knn_gather_kernel<<<nbox, KNN_THREADS, 0, s>>>(P, points, v.vals[cur], v.sorted, v.boxes);
    knn_search_kernel<<<div_up(P, KNN_THREADS), KNN_THREADS, 0, s>>>(P, v.sorted, v.boxes, nbox, mean_d2);
    return 0;
}

}  // namespace gsr
