// Shared device/host helpers for the B200 (sm_100a) Gaussian-splatting rasterizer.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>

namespace gsr {

constexpr int TILE = 16;            // reference BLOCK_X = BLOCK_Y (cuda_rasterizer/config.h:16-17)
constexpr int TILE_PIX = TILE * TILE;
constexpr uint32_t CULLED_KEY = 0xFFFFFFFFu;  // depth key of an invisible Gaussian (sorts last)

// ---- radix sort geometry -------------------------------------------------------------------
constexpr int SORT_THREADS = 256;
#ifndef GSR_SORT_ITEMS
#define GSR_SORT_ITEMS 8
#endif
constexpr int SORT_ITEMS = GSR_SORT_ITEMS;              // items per thread
constexpr int SORT_CHUNK = SORT_THREADS * SORT_ITEMS;  // items per block
constexpr int SCAN_ITEMS = 8;                           // depth ranks per thread of the instance-count sums kernel
constexpr int SCAN_CHUNK = SORT_THREADS * SCAN_ITEMS;  // = 8 emit blocks of 256 ranks
constexpr int RADIX_BITS = 8;
constexpr int RADIX = 1 << RADIX_BITS;
constexpr int SORT_MAX_PASSES = 4;
constexpr int SORT_AUX_WORDS = RADIX * SORT_MAX_PASSES + 64;  // per-pass global digit histograms + block tickets

inline int div_up(int64_t a, int64_t b) { return (int)((a + b - 1) / b); }
inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// ---- workspaces ----------------------------------------------------------------------------
// Geometry workspace: per-Gaussian SoA records (ours, not the reference's layout).
struct GeomView {
    uint32_t* depth_key;   // [P] float bits of view depth, CULLED_KEY if invisible
    uint32_t* tiles;       // [P] tiles touched
    float2* xy;            // [P] pixel centre
    float4* conic_op;      // [P] conic a,b,c + opacity
    float4* rgb_depth;     // [P] colour (SH-evaluated or copied from colors_precomp) + view depth
    uint2* rect;           // [P] x: min.x | max.x << 16, y: min.y | max.y << 16
    uint8_t* clamped;      // [P] bit ch set if SH colour channel ch was clamped at 0
    uint32_t* sort_keys[2];  // [P] ping-pong depth keys
    uint32_t* sort_vals[2];  // [P] ping-pong Gaussian ids; sort_vals[0] ends up holding the depth order
    uint32_t* offsets;     // [P, first 8*nblk(P) used] sums of tiles[] per 256 depth ranks (the emit kernel scans the rest itself)
    uint32_t* hist;        // [RADIX * nblk(P)] per-block digit counts
    uint32_t* digit_total; // [SORT_AUX_WORDS] per-pass global digit histograms, block tickets
    uint32_t* block_sums;  // [nblk(P)+1]
    unsigned long long* counters;  // [4] 0: total instances (bit 62: prefilter violation)
};

struct ImageView {
    uint2* ranges;        // [tiles]
    uint32_t* n_contrib;  // [W*H]
};

struct BinView {
    uint32_t* keys[2];  // [R] tile ids (ping-pong)
    uint32_t* vals[2];  // [R] Gaussian ids (ping-pong); vals[0] ends up as the sorted point list
    uint32_t* hist;     // [RADIX * nblk(R)]
    uint32_t* digit_total;  // [SORT_AUX_WORDS]
};

template <typename T>
inline void carve(char*& p, T*& out, size_t count)
{
    p = (char*)align_up((size_t)p, 256);
    out = (T*)p;
    p += count * sizeof(T);
}

inline size_t geom_layout(char* base, int P, GeomView& g)
{
    char* p = base;
    const size_t n = (size_t)P;
    const size_t nblk = (size_t)div_up(P, SORT_CHUNK);
    carve(p, g.depth_key, n);
    carve(p, g.tiles, n);
    carve(p, g.xy, n);
    carve(p, g.conic_op, n);
    carve(p, g.rgb_depth, n);
    carve(p, g.rect, n);
    carve(p, g.clamped, n);
    carve(p, g.sort_keys[0], n);
    carve(p, g.sort_keys[1], n);
    carve(p, g.sort_vals[0], n);
    carve(p, g.sort_vals[1], n);
    carve(p, g.offsets, n);
    carve(p, g.hist, (size_t)RADIX * nblk);
    carve(p, g.digit_total, SORT_AUX_WORDS);
    carve(p, g.block_sums, (size_t)div_up(P, SCAN_CHUNK) + 1);
    carve(p, g.counters, 4);
    return (size_t)(p - base) + 256;
}

inline size_t image_layout(char* base, int W, int H, ImageView& v)
{
    char* p = base;
    const size_t tiles = (size_t)div_up(W, TILE) * div_up(H, TILE);
    carve(p, v.ranges, tiles);
    carve(p, v.n_contrib, (size_t)W * H);
    return (size_t)(p - base) + 256;
}

inline size_t bin_layout(char* base, int64_t R, BinView& b)
{
    char* p = base;
    const size_t n = (size_t)(R > 0 ? R : 1);
    const size_t nblk = (size_t)div_up((int64_t)n, SORT_CHUNK);
    carve(p, b.keys[0], n);
    carve(p, b.keys[1], n);
    carve(p, b.vals[0], n);
    carve(p, b.vals[1], n);
    carve(p, b.hist, (size_t)RADIX * nblk);
    carve(p, b.digit_total, SORT_AUX_WORDS);
    return (size_t)(p - base) + 256;
}

// ---- small device helpers ------------------------------------------------------------------
#ifdef __CUDACC__
__device__ __forceinline__ unsigned lane_id() { return threadIdx.x & 31; }

// Tile ranges (reference: identifyTileRanges, rasterizer_impl.cu:116-138) are written by atomics inside the last
// pass of the tile sort: every block knows where its own run of a tile id lands, the smallest start and the largest
// end over the blocks are the tile's range.  So that zero-initialised memory works for both with atomicMax, .x holds
// the COMPLEMENT of the start; a tile nobody touched decodes to the reference's (0, 0).
__device__ __forceinline__ uint2 load_range(const uint2* __restrict__ ranges, int tile)
{
    uint2 r = ranges[tile];
    r.x = r.y ? ~r.x : 0u;
    return r;
}

// Block-wide exclusive scan of one value per thread (blockDim.x == 256). smem: 8 uint32.
__device__ __forceinline__ uint32_t block_exclusive_scan_256(uint32_t v, uint32_t* warp_sums, uint32_t* total)
{
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= (unsigned)o) inc += t;
    }
    if (lane == 31) warp_sums[warp] = inc;
    __syncthreads();
    uint32_t base = 0, tot = 0;
#pragma unroll
    for (int w = 0; w < 8; w++) {
        const uint32_t s = warp_sums[w];
        if ((unsigned)w < warp) base += s;
        tot += s;
    }
    if (total) *total = tot;
    __syncthreads();
    return base + inc - v;
}
#endif

}  // namespace gsr
