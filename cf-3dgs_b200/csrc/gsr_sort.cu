// Binning: the (tile|depth) key sort of the reference (rasterizer_impl.cu:70-138,278-319) executed as a
// split-key LSD radix sort.
//
// The reference emits one 64-bit key (tile << 32 | depth bits) per tile instance and runs a stable LSD
// radix sort over all R instances on 32+bit key bits (CUB, ~6 passes of 24 B/instance each).  Every
// instance of a Gaussian carries the same low 32 bits, and a stable sort on those bits commutes with
// the duplication: so the low-32-bit digit passes run here on the P Gaussians BEFORE they are duplicated,
// instances are then emitted in depth order, and only the tile digits (10-13 bits) are sorted over the R
// instances.  The 64-bit keys are never materialised in the hot path; the result (order within each tile:
// depth bits ascending, ties by Gaussian index = emission order) is identical to the reference's stable
// sort, and gsr_debug_dump() rebuilds the sorted key array for bit-exact comparison.
//
// Each radix pass is histogram -> row scan -> stable scatter (three kernels, no inter-block waiting).  Ranks
// inside a block come from warp-level multi-split: __match_any_sync groups the lanes holding the same digit, the
// group's lowest lane bumps a warp-private counter, and popc(mask & lanemask_lt) orders the lanes; warp counters
// are then scanned across the 8 warps.  Items leave through shared memory in block-sorted order so a warp's
// stores form a few contiguous runs.
// (A single-kernel "onesweep" pass with decoupled look-back measured 5-15 % slower on B200 at these sizes -- 1M keys /
// 6M instances, look-back latency dominates 2048-item chunks -- and was removed.)
#include "gsr_kernels.h"
#include "gsr_launch.cuh"

namespace gsr {

// In the sync-free forward the instance count lives only on the device: kernels launched for `cap` items read the
// real count (bit 62 of the counter is the prefilter flag) and clamp it to the capacity of the workspace.
__device__ __forceinline__ int64_t device_count(const unsigned long long* n_dev, int64_t cap)
{
    if (!n_dev) return cap;
    const unsigned long long v = *n_dev & ((1ull << 62) - 1ull);
    return v < (unsigned long long)cap ? (int64_t)v : cap;
}

// ---------------------------------------------------------------------------------------------
// pass kernel 1: per-block digit histogram, written digit-major: hist[d * nblk + block]
__global__ void __launch_bounds__(SORT_THREADS)
radix_hist_kernel(const uint32_t* __restrict__ keys, int64_t n_cap, const unsigned long long* __restrict__ n_dev,
                  int shift, uint32_t mask, uint32_t* __restrict__ hist, int nblk)
{
    __shared__ uint32_t sh[RADIX];
    sh[threadIdx.x] = 0;
    pdl_wait();
    pdl_trigger();
    const int64_t n = device_count(n_dev, n_cap);
    __syncthreads();
    const int64_t base = (int64_t)blockIdx.x * SORT_CHUNK;
    const unsigned lane = threadIdx.x & 31;
    // all loads first: written inline with the counting, each load was followed by its own use (eight round trips to
    // memory per thread, 72 % of this kernel's stall samples)
    uint32_t k[SORT_ITEMS];
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; i++) {
        const int64_t idx = base + (int64_t)i * SORT_THREADS + threadIdx.x;
        k[i] = idx < n ? __ldg(keys + idx) : 0u;
    }
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; i++) {
        const int64_t idx = base + (int64_t)i * SORT_THREADS + threadIdx.x;
        const bool act = idx < n;
        const uint32_t d = act ? ((k[i] >> shift) & mask) : (uint32_t)RADIX;
        // high digits (tile rows, float exponents) are often identical across a warp: count those with one atomic,
        // everything else with plain shared atomics (few same-address collisions for mixed digits)
        const uint32_t d0 = __shfl_sync(0xffffffffu, d, 0);
        if (__all_sync(0xffffffffu, d == d0)) {
            if (lane == 0 && d0 < (uint32_t)RADIX) atomicAdd(&sh[d0], 32u);
        } else if (act) {
            atomicAdd(&sh[d], 1u);
        }
    }
    __syncthreads();
    hist[(size_t)threadIdx.x * nblk + blockIdx.x] = sh[threadIdx.x];
}

// pass kernel 2: block d turns row d of hist into its exclusive prefix and records the row total.  The row is walked
// in tiles of 1024 entries: one coalesced 16-byte load per thread, a block scan, a running carry.
__global__ void __launch_bounds__(SORT_THREADS)
radix_rowscan_kernel(uint32_t* __restrict__ hist, int nblk, uint32_t* __restrict__ digit_total)
{
    __shared__ uint32_t warp_sums[8];
    pdl_wait();
    pdl_trigger();
    uint32_t* row = hist + (size_t)blockIdx.x * nblk;
    // rows start at multiples of nblk words: 16-byte accesses only when that keeps them aligned
    const bool vec = (nblk & 3) == 0;
    uint32_t carry = 0;
    for (int t0 = 0; t0 < nblk; t0 += SORT_THREADS * 4) {
        const int i0 = t0 + threadIdx.x * 4;
        uint32_t v[4] = {0u, 0u, 0u, 0u};
        if (vec && i0 + 3 < nblk) {
            const uint4 q = *reinterpret_cast<const uint4*>(row + i0);
            v[0] = q.x; v[1] = q.y; v[2] = q.z; v[3] = q.w;
        } else {
#pragma unroll
            for (int c = 0; c < 4; c++)
                if (i0 + c < nblk) v[c] = row[i0 + c];
        }
        uint32_t total;
        uint32_t run = carry + block_exclusive_scan_256(v[0] + v[1] + v[2] + v[3], warp_sums, &total);
        uint32_t o[4];
#pragma unroll
        for (int c = 0; c < 4; c++) { o[c] = run; run += v[c]; }
        if (vec && i0 + 3 < nblk) {
            *reinterpret_cast<uint4*>(row + i0) = make_uint4(o[0], o[1], o[2], o[3]);
        } else {
#pragma unroll
            for (int c = 0; c < 4; c++)
                if (i0 + c < nblk) row[i0 + c] = o[c];
        }
        carry += total;   // (the scan ends with a barrier, so warp_sums can be reused by the next tile)
    }
    if (threadIdx.x == 0) digit_total[blockIdx.x] = carry;
}

// pass kernel 3: stable scatter.  Items are first placed at their block-local sorted position in shared memory, then
// written out in that order, so a warp's stores fall into a few contiguous runs (one per digit) instead of 32 sectors.
#ifndef GSR_SCATTER_MIN_BLOCKS
#define GSR_SCATTER_MIN_BLOCKS 4
#endif
// BALLOT_BITS > 0: the lanes holding the same digit are found with one ballot per digit bit instead of MATCH.ANY.
// Measured on B200 (tools/micro/match_bench.cu): MATCH.ANY costs ~7 cycles per DISTINCT value in the warp (226 cycles
// for 32 distinct values), eight ballots a flat ~109 -- and the digits a warp sees here are mostly distinct (low tile
// bits of consecutive instances, low mantissa bits of depths).
template <int BALLOT_BITS, bool FULL>
__device__ __forceinline__ void scatter_block(const uint32_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                                              uint32_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out, int count,
                                              int shift, uint32_t mask, uint32_t my_gbase, uint2* __restrict__ ranges,
                                              uint32_t (*wcount)[RADIX + 1], uint32_t* goff, uint32_t* warp_sums,
                                              uint32_t* skey, uint32_t* sval)
{
    // FULL: all SORT_CHUNK items of the block exist (every block but the last): no bounds logic anywhere below
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t block_base = blockIdx.x * (uint32_t)SORT_CHUNK;       // item counts are below 2^31 (GSR_ERR_OVERFLOW)
    const uint32_t wfirst = warp * (SORT_ITEMS * 32) + lane;              // block-local index of this lane's first item
    uint32_t key[SORT_ITEMS], rank[SORT_ITEMS];
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; i++) {
        const uint32_t li = wfirst + i * 32;
        key[i] = (FULL || (int)li < count) ? keys_in[block_base + li] : 0u;
    }
    // Ranking, written so that the eight items' latencies overlap instead of adding up: all group masks first, then one
    // shared-memory atomic per digit group (the group leader fetches-and-adds the warp-private counter; a warp's
    // shared-memory instructions execute in order, so item i's groups are counted before item i+1's), then the
    // broadcasts.  Nothing in one item's chain waits for another item's result.
    unsigned grp[SORT_ITEMS];
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; i++) {
        const bool act = FULL || (int)(wfirst + i * 32) < count;
        const uint32_t d = act ? ((key[i] >> shift) & mask) : (uint32_t)RADIX;
        if (BALLOT_BITS > 0) {
            unsigned same = 0xffffffffu;
            if (!FULL) {
                same = __ballot_sync(0xffffffffu, act);
                if (!act) same = ~same;
            }
#pragma unroll
            for (int b = 0; b < BALLOT_BITS; b++) {
                const bool bit = (d >> b) & 1u;
                const unsigned bal = __ballot_sync(0xffffffffu, bit);
                if (bit) same &= bal; else same &= ~bal;
            }
            grp[i] = same;
        } else {
            grp[i] = __match_any_sync(0xffffffffu, d);
        }
    }
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; i++) {
        const bool act = FULL || (int)(wfirst + i * 32) < count;
        const uint32_t d = act ? ((key[i] >> shift) & mask) : (uint32_t)RADIX;
        rank[i] = 0;
        if (lane == (unsigned)(__ffs(grp[i]) - 1)) rank[i] = atomicAdd(&wcount[warp][d], (uint32_t)__popc(grp[i]));
        __syncwarp();
    }
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; i++)
        rank[i] = __shfl_sync(0xffffffffu, rank[i], __ffs(grp[i]) - 1) + (uint32_t)__popc(grp[i] & ((1u << lane) - 1u));
    __syncthreads();
    // per digit: total over the block, its block-local start (scan over digits), then warp bases
    {
        uint32_t total = 0;
#pragma unroll
        for (int w = 0; w < SORT_THREADS / 32; w++) total += wcount[w][threadIdx.x];
        uint32_t run = block_exclusive_scan_256(total, warp_sums, nullptr);
        goff[threadIdx.x] = my_gbase - run;
#pragma unroll
        for (int w = 0; w < SORT_THREADS / 32; w++) {
            const uint32_t c = wcount[w][threadIdx.x];
            wcount[w][threadIdx.x] = run;
            run += c;
        }
    }
    __syncthreads();
    // the values are only needed now: loading them here instead of next to the keys keeps 8 registers free during the
    // ranking loop (occupancy is what hides its shared-memory latency)
    uint32_t val[SORT_ITEMS];
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; i++) {
        const uint32_t li = wfirst + i * 32;
        val[i] = vals_in ? ((FULL || (int)li < count) ? vals_in[block_base + li] : 0u) : block_base + li;   // no value array: the item index
    }
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; i++) {
        if (FULL || (int)(wfirst + i * 32) < count) {
            const uint32_t d = (key[i] >> shift) & mask;
            const uint32_t lp = wcount[warp][d] + rank[i];
            skey[lp] = key[i];
            sval[lp] = val[i];
        }
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < SORT_ITEMS; i++) {
        const int t = i * SORT_THREADS + threadIdx.x;
        if (FULL || t < count) {
            const uint32_t k = skey[t];
            const uint32_t pos = goff[(k >> shift) & mask] + (uint32_t)t;
            keys_out[pos] = k;
            vals_out[pos] = sval[t];
            // last pass of the tile sort: the block's items with one tile id sit next to each other (equal last digit, and
            // inside it the order the earlier passes left: ascending lower bits), so the two ends of each such run give the
            // tile's range -- see load_range() in gsr_common.cuh
            if (ranges) {
                if (t == 0 || skey[t - 1] != k) atomicMax(&ranges[k].x, ~pos);
                if (t == count - 1 || skey[t + 1] != k) atomicMax(&ranges[k].y, pos + 1u);
            }
        }
    }
}

template <int BALLOT_BITS>
__global__ void __launch_bounds__(SORT_THREADS, GSR_SCATTER_MIN_BLOCKS)
radix_scatter_kernel(const uint32_t* __restrict__ keys_in, const uint32_t* __restrict__ vals_in,
                     uint32_t* __restrict__ keys_out, uint32_t* __restrict__ vals_out, int64_t n_cap,
                     const unsigned long long* __restrict__ n_dev, int shift,
                     uint32_t mask, const uint32_t* __restrict__ hist, int nblk,
                     const uint32_t* __restrict__ digit_total, uint2* __restrict__ ranges)
{
    __shared__ uint32_t wcount[SORT_THREADS / 32][RADIX + 1];
    __shared__ uint32_t goff[RADIX];   // global base of digit d for this block minus its block-local start
    __shared__ uint32_t warp_sums[8];
    __shared__ uint32_t skey[SORT_CHUNK + 1], sval[SORT_CHUNK];
    for (int i = threadIdx.x; i < (SORT_THREADS / 32) * (RADIX + 1); i += SORT_THREADS) (&wcount[0][0])[i] = 0;
    pdl_wait();
    pdl_trigger();
    const int64_t n = device_count(n_dev, n_cap);
    const int count = (int)max((int64_t)0, min((int64_t)SORT_CHUNK, n - (int64_t)blockIdx.x * SORT_CHUNK));
    // global base of each digit for this block = exclusive scan of digit totals + row prefix
    uint32_t my_gbase;
    {
        const uint32_t tot = digit_total[threadIdx.x];
        const uint32_t ex = block_exclusive_scan_256(tot, warp_sums, nullptr);  // contains __syncthreads
        my_gbase = ex + hist[(size_t)threadIdx.x * nblk + blockIdx.x];
    }
    if (count == SORT_CHUNK)
        scatter_block<BALLOT_BITS, true>(keys_in, vals_in, keys_out, vals_out, count, shift, mask, my_gbase, ranges, wcount,
                                         goff, warp_sums, skey, sval);
    else if (count > 0)
        scatter_block<BALLOT_BITS, false>(keys_in, vals_in, keys_out, vals_out, count, shift, mask, my_gbase, ranges, wcount,
                                          goff, warp_sums, skey, sval);
}

int radix_sort_pairs(uint32_t* keys[2], uint32_t* vals[2], int64_t n, int begin_bit, int end_bit,
                     uint32_t* hist, uint32_t* digit_total, cudaStream_t s, const unsigned long long* n_dev,
                     const uint32_t* first_keys, uint2* ranges)
{
    int cur = 0;
    if (n <= 0) return cur;
    const int nblk = div_up(n, SORT_CHUNK);
    for (int shift = begin_bit; shift < end_bit; shift += RADIX_BITS) {
        const int bits = min(RADIX_BITS, end_bit - shift);
        const uint32_t mask = (1u << bits) - 1u;
        // first_keys: the first pass reads its keys there and takes the item index as the value, so the caller
        // does not have to materialise (key, iota) pairs first
        const bool first = first_keys && shift == begin_bit;
        const uint32_t* kin = first ? first_keys : keys[cur];
        const uint32_t* vin = first ? nullptr : vals[cur];
        launch_pdl(radix_hist_kernel, nblk, SORT_THREADS, 0, s, kin, n, n_dev, shift, mask, hist, nblk);
        launch_pdl(radix_rowscan_kernel, RADIX, SORT_THREADS, 0, s, hist, nblk, digit_total);
        // the sign/exponent byte of a depth takes few distinct values (MATCH is cheap there); every other digit is
        // close to all-distinct within a warp
        const bool ballots = !(shift == 24 && end_bit == 32);
#define GSR_SCATTER(B)                                                                                                \
    launch_pdl(radix_scatter_kernel<B>, nblk, SORT_THREADS, 0, s, kin, vin, keys[cur ^ 1], vals[cur ^ 1], n, n_dev, shift, \
               mask, (const uint32_t*)hist, nblk, (const uint32_t*)digit_total,                                       \
               (shift + RADIX_BITS >= end_bit) ? ranges : (uint2*)nullptr)
        if (!ballots) GSR_SCATTER(0);
        else if (bits <= 5) GSR_SCATTER(5);
        else GSR_SCATTER(8);
#undef GSR_SCATTER
        cur ^= 1;
    }
    return cur;
}

// ---------------------------------------------------------------------------------------------
__global__ void init_sort_input_kernel(const uint32_t* __restrict__ depth_key, uint32_t* __restrict__ keys,
                                       uint32_t* __restrict__ vals, int P)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < P) {
        keys[i] = depth_key[i];
        vals[i] = (uint32_t)i;
    }
}

void launch_init_sort_input(const GeomView& g, int P, cudaStream_t s)
{
    init_sort_input_kernel<<<div_up(P, 256), 256, 0, s>>>(g.depth_key, g.sort_keys[0], g.sort_vals[0], P);
}

// ---------------------------------------------------------------------------------------------
// Instance offsets = exclusive prefix sum over depth order of tiles[order[i]]  (reference: cub InclusiveSum over
// Gaussian order, rasterizer_impl.cu:278; here over depth order because instances are emitted in depth order).
// Two levels of partial sums are all that is stored: per 2048 ranks (block_sums) and per 256 ranks (group_sums, one
// per warp of this kernel = one per block of the emit kernel).  The emit kernel adds up the sums in front of its
// 256 ranks itself and scans its own ranks in shared memory, so no per-rank offset array is written or read and the
// second scan kernel is gone.
static_assert(SCAN_ITEMS * 32 == 256, "one warp of the sums kernel covers the 256 ranks of one emit block");

__global__ void __launch_bounds__(SORT_THREADS)
scan_block_sums_kernel(const uint32_t* __restrict__ tiles, const uint32_t* __restrict__ order, int P,
                       uint32_t* __restrict__ block_sums, uint32_t* __restrict__ group_sums)
{
    __shared__ uint32_t warp_sums[8];
    const int base = blockIdx.x * SCAN_CHUNK + threadIdx.x * SCAN_ITEMS;
    uint32_t sum = 0;
    pdl_wait();
    pdl_trigger();
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; i++)
        if (base + i < P) sum += tiles[order[base + i]];
    uint32_t wsum = sum;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) wsum += __shfl_xor_sync(0xffffffffu, wsum, o);
    if ((threadIdx.x & 31) == 0) {
        warp_sums[threadIdx.x >> 5] = wsum;
        group_sums[blockIdx.x * 8 + (threadIdx.x >> 5)] = wsum;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t total = 0;
#pragma unroll
        for (int w = 0; w < 8; w++) total += warp_sums[w];
        block_sums[blockIdx.x] = total;
    }
}

void launch_offsets_scan(const uint32_t* tiles, const uint32_t* order, uint32_t* group_sums, uint32_t* block_sums,
                         int P, cudaStream_t s)
{
    const int nblk = div_up(P, SCAN_CHUNK);
    launch_pdl(scan_block_sums_kernel, nblk, SORT_THREADS, 0, s, tiles, order, P, block_sums, group_sums);
}

// ---------------------------------------------------------------------------------------------
// Instance emit in depth order (reference duplicateWithKeys, rasterizer_impl.cu:70-111, which emits in Gaussian
// order with one thread looping over each Gaussian's tiles).  key = tile id only; the depth half of the reference key
// is implied by the order.  Load-balanced: a block owns 256 consecutive depth ranks, whose instances form ONE
// contiguous output range; its threads walk that range with stride 256 and find the owning Gaussian of each slot by
// binary search over the block's 257 offsets in shared memory, so every store is coalesced and a splat covering
// hundreds of tiles costs no more per instance than one covering four.
constexpr int EMIT_THREADS = 256;

__global__ void __launch_bounds__(EMIT_THREADS)
emit_kernel(const uint32_t* __restrict__ order, const uint32_t* __restrict__ tiles,
            const uint32_t* __restrict__ group_sums, const uint32_t* __restrict__ block_sums,
            const uint2* __restrict__ rect, int P, int grid_x,
            uint32_t* __restrict__ tile_keys, uint32_t* __restrict__ tile_vals, uint32_t cap)
{
    __shared__ uint32_t warp_sums[8];
    __shared__ uint32_t s_off[EMIT_THREADS + 1];
    __shared__ uint32_t s_gid[EMIT_THREADS];
    __shared__ uint2 s_rect[EMIT_THREADS];
    const int r = blockIdx.x * EMIT_THREADS + threadIdx.x;
    uint32_t off = 0xffffffffu, gid = 0, n = 0;
    uint2 rc = make_uint2(0u, 0u);
    pdl_wait();
    pdl_trigger();
    if (r < P) {
        gid = order[r];
        n = tiles[gid];
        rc = rect[gid];
    }
    {
        // instances in front of this block: whole 2048-rank chunks, then the 256-rank groups of its own chunk
        const int chunk = blockIdx.x >> 3;
        uint32_t before = 0;
        for (int i = threadIdx.x; i < chunk; i += EMIT_THREADS) before += block_sums[i];
        if ((int)threadIdx.x < (int)(blockIdx.x & 7)) before += group_sums[chunk * 8 + threadIdx.x];
        uint32_t block_base;
        block_exclusive_scan_256(before, warp_sums, &block_base);
        const uint32_t mine = block_exclusive_scan_256(n, warp_sums, nullptr) + block_base;
        if (r < P) off = mine;
    }
    s_off[threadIdx.x] = off;
    s_gid[threadIdx.x] = gid;
    s_rect[threadIdx.x] = rc;
    // end of the block's range = offset + count of its last valid rank
    const int last = min(EMIT_THREADS, P - blockIdx.x * EMIT_THREADS) - 1;
    if ((int)threadIdx.x == last) s_off[EMIT_THREADS] = off + n;
    __syncthreads();
    const uint32_t begin = s_off[0];
    const uint32_t end = min(s_off[EMIT_THREADS], cap);   // cap: only the sync-free path can run out of room
    // four consecutive instances per thread: one binary search, then a walk; 16-byte stores when the quad is whole
    for (uint32_t p0 = (begin & ~3u) + 4u * threadIdx.x; p0 < end; p0 += 4u * EMIT_THREADS) {
        const uint32_t first = max(p0, begin);
        // last k in [0, last] with s_off[k] <= first
        int lo = 0, hi = last;
        while (lo < hi) {
            const int mid = (lo + hi + 1) >> 1;
            if (s_off[mid] <= first) lo = mid; else hi = mid - 1;
        }
        uint32_t key[4], val[4];
        uint32_t next_off = s_off[lo + 1];     // s_off[last + 1] = end of the block's range
        uint2 q = s_rect[lo];
        uint32_t gid = s_gid[lo];
        uint32_t x0 = q.x & 0xffffu, w = (q.x >> 16) - x0, y0 = q.y & 0xffffu;
        uint32_t local = first - s_off[lo];
        uint32_t row = local / w, col = local - row * w;
#pragma unroll
        for (int c = 0; c < 4; c++) {
            const uint32_t p = p0 + c;
            key[c] = 0u; val[c] = 0u;
            if (p < first || p >= end) continue;
            while (p >= next_off) {            // next Gaussian with at least one tile (zero-tile ranks repeat the offset)
                lo++;
                next_off = s_off[lo + 1];
                q = s_rect[lo]; gid = s_gid[lo];
                x0 = q.x & 0xffffu; w = (q.x >> 16) - x0; y0 = q.y & 0xffffu;
                row = 0; col = 0;
            }
            key[c] = (y0 + row) * (uint32_t)grid_x + x0 + col;
            val[c] = gid;
            if (++col == w) { col = 0; row++; }
        }
        if (p0 >= begin && p0 + 3 < end) {
            *reinterpret_cast<uint4*>(tile_keys + p0) = make_uint4(key[0], key[1], key[2], key[3]);
            *reinterpret_cast<uint4*>(tile_vals + p0) = make_uint4(val[0], val[1], val[2], val[3]);
        } else {
#pragma unroll
            for (int c = 0; c < 4; c++)
                if (p0 + c >= begin && p0 + c < end) { tile_keys[p0 + c] = key[c]; tile_vals[p0 + c] = val[c]; }
        }
    }
}

void launch_emit(const GeomView& g, const uint32_t* order, int P, int grid_x, uint32_t* tile_keys,
                 uint32_t* tile_vals, int64_t cap, cudaStream_t s)
{
    launch_pdl(emit_kernel, div_up(P, EMIT_THREADS), EMIT_THREADS, 0, s, order, (const uint32_t*)g.tiles,
               (const uint32_t*)g.offsets, (const uint32_t*)g.block_sums, (const uint2*)g.rect, P, grid_x, tile_keys,
               tile_vals, (uint32_t)cap);
}

// sync-free forward: publish (count, overflow) for the host to look at whenever it likes.  The status words only ever
// grow: the count is max-ed and the flags are OR-ed in, so a status buffer shared by many calls (several views, graph
// replays) still shows an overflow of any one of them until the caller zeroes it.  Calls are stream-ordered, so plain
// read-modify-write is enough for calls on one stream; atomics make concurrent streams safe as well.
__global__ void publish_status_kernel(const unsigned long long* __restrict__ counters, long long cap, long long* __restrict__ status)
{
    pdl_wait();
    pdl_trigger();
    const unsigned long long raw = counters[0];
    const unsigned long long n = raw & ((1ull << 62) - 1ull);
    atomicMax(reinterpret_cast<unsigned long long*>(status), n);
    const unsigned long long flags = (n > (unsigned long long)cap ? 1ull : 0ull) | ((raw >> 62) ? 2ull : 0ull);
    if (flags) atomicOr(reinterpret_cast<unsigned long long*>(status + 1), flags);
}

void launch_publish_status(const unsigned long long* counters, int64_t cap, long long* status, cudaStream_t s)
{
    launch_pdl(publish_status_kernel, 1, 1, 0, s, counters, (long long)cap, status);
}

__global__ void build_keys_kernel(const uint32_t* __restrict__ sorted_tiles, const uint32_t* __restrict__ point_list,
                                  const uint32_t* __restrict__ depth_key, int64_t R_cap,
                                  const unsigned long long* __restrict__ n_dev, uint64_t* __restrict__ keys_out)
{
    // after a sync-free forward only the first (device-side) count entries of the lists are defined
    const int64_t R = device_count(n_dev, R_cap);
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < R) keys_out[i] = ((uint64_t)sorted_tiles[i] << 32) | depth_key[point_list[i]];
    else if (i < R_cap) keys_out[i] = 0;
}

void launch_build_keys(const uint32_t* sorted_tiles, const uint32_t* point_list, const uint32_t* depth_key,
                       int64_t R, const unsigned long long* n_dev, uint64_t* keys_out, cudaStream_t s)
{
    if (R > 0) build_keys_kernel<<<div_up(R, 256), 256, 0, s>>>(sorted_tiles, point_list, depth_key, R, n_dev, keys_out);
}

}  // namespace gsr
