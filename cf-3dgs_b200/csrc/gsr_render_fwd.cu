// Forward tile blend.  Semantics: reference forward.cu:261-381 (renderCUDA): per pixel, front to back,
//   power > 0 -> skip;  alpha = min(0.99, o*exp(power)) < 1/255 -> skip;  T*(1-alpha) < 1e-4 -> pixel done;
//   C += c*alpha*T, weight += alpha*T, D += depth*alpha*T;  out = C + T*bg, alpha = weight, depth = D,
//   n_contrib = 1-based list position of the last blended entry.
// All per-pair arithmetic uses the reference's rounding sequence (gsr_math.cuh: blend_power; expf is the
// full-accuracy CUDA expf the reference calls), so the images are bit-identical given identical inputs.
//
// What differs is how much work is skipped: entries are staged 256 at a time as packed records (one 16-B
// gather each for geometry, conic and colour+depth), the staging thread rejects entries that cannot reach
// alpha >= 1/255 anywhere in the tile, and each warp (8x8 pixels, two per lane) rejects entries for its own
// footprint with one lane per entry + a ballot, then walks only the surviving bits.  Rejected pairs are exactly
// pairs the reference would have skipped at its alpha test.
#include "gsr_blend.cuh"
#include "gsr_f32x2.cuh"
#include "gsr_kernels.h"
#include "gsr_launch.cuh"

namespace gsr {

// Two pixels per lane, FP32 pairs (gsr_f32x2.cuh).  128 threads per tile; each warp owns an 8x8
// footprint and lane (lx, ly) blends pixels (lx, ly) and (lx, ly + 4) of it.  The entry walk, the record loads, dx and
// the scalar factors dx*A, dx*B are shared by the two pixels; dy, the exponent, alpha, the transmittance and the five
// accumulators are FFMA2 / FMUL2 / FADD2 pairs.  Per-pixel control flow becomes selects: a pixel that skips an entry
// (power > 0, alpha < 1/255, or already finished) blends it with alpha = 0, which leaves T and every accumulator
// bit-unchanged (T*1, C + 0), and the entry that would drop T below 1e-4 is blended with alpha = 0 too and latches the
// pixel's `done` predicate -- so each pixel sees exactly the reference's sequence of rounded operations.
#ifndef GSR_FWDP_MIN_BLOCKS
#define GSR_FWDP_MIN_BLOCKS 8
#endif
constexpr int PK_WARPS = 4;
constexpr int PK_THREADS = 32 * PK_WARPS;
constexpr int PK_FOOT = 8;

__global__ void __launch_bounds__(PK_THREADS, GSR_FWDP_MIN_BLOCKS)
render_forward_packed_kernel(int W, int H, int grid_x, const uint2* __restrict__ ranges,
                             const uint32_t* __restrict__ point_list, GeomView g, const float* __restrict__ bg,
                             float* __restrict__ out_color, float* __restrict__ out_depth,
                             float* __restrict__ out_alpha, uint32_t* __restrict__ n_contrib)
{
    __shared__ Staged sm;
    const int tile = blockIdx.x;
    const int tile_x = tile % grid_x, tile_y = tile / grid_x;
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;   // footprint index 0..3
    const int fx0 = tile_x * TILE + (warp & 1) * PK_FOOT, fy0 = tile_y * TILE + (warp >> 1) * PK_FOOT;
    const int pix_x = fx0 + (lane & 7), pix_ya = fy0 + (lane >> 3), pix_yb = pix_ya + 4;
    const float pxf = (float)pix_x;
    const f32x2 py2 = pk((float)pix_ya, (float)pix_yb);
    const float foot_x0 = (float)fx0, foot_y0 = (float)fy0;
    const float tile_x0 = (float)(tile_x * TILE), tile_y0 = (float)(tile_y * TILE);

    pdl_wait();
    pdl_trigger();
    const uint2 range = load_range(ranges, tile);
    const int total = (int)(range.y - range.x);
    const bool inside_a = pix_x < W && pix_ya < H, inside_b = pix_x < W && pix_yb < H;
    // A finished pixel (outside the image, or T*(1-alpha) < 1e-4 reached) carries its transmittance with the sign
    // flipped: T < 0 <=> done, |T| is the value the background term needs.  No separate flag to keep in a register.
    // (kept as two floats and packed at its uses: ptxas then updates the pair in place)
    float2 T = {inside_a ? 1.0f : -1.0f, inside_b ? 1.0f : -1.0f};
    f32x2 C0 = bc(0.f), C1 = bc(0.f), C2 = bc(0.f), Wt = bc(0.f), Dacc = bc(0.f);
    uint32_t last_a = 0, last_b = 0;
    const auto all_done = [&]() { return (T.x < 0.f) && (T.y < 0.f); };

    for (int base = 0; base < total; base += BATCH) {
        if (__syncthreads_count(all_done()) == PK_THREADS) break;
        const int n = min(BATCH, total - base);
        for (int k = threadIdx.x; k < n; k += PK_THREADS)
            stage_entry(sm, k, point_list[range.x + base + k], g, tile_x0, tile_y0);
        __syncthreads();

        for (int c0 = 0; c0 < n; c0 += 32) {
            if (__all_sync(0xffffffffu, all_done())) break;
            const int e = c0 + lane;
            bool keep = false;
            if (e < n) {
                const float thr = sm.thr[e];
                if (!(thr < 0.f)) {
                    const float4 r0 = sm.rec0[e];
                    const float4 r1 = sm.rec1[e];
                    keep = !reject_rect(r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, thr, foot_x0,
                                        foot_x0 + (PK_FOOT - 1), foot_y0, foot_y0 + (PK_FOOT - 1));
                }
            }
            unsigned live = __ballot_sync(0xffffffffu, keep);
            const uint32_t pos0 = (uint32_t)(base + c0 + 1);
            while (live) {
                const int b = __ffs(live) - 1;
                live &= live - 1;
                const int j = c0 + b;
                const float4 r0 = sm.rec0[j];
                const float4 r1 = sm.rec1[j];
                // blend_power() for both pixels: t = fma(dx, dx*A, dy*(dy*C)); power = fma(t, -0.5, -(dy*(dx*B)))
                const float dx = __fadd_rn(r0.x, -pxf);
                const f32x2 dy = sub2(bc(r0.y), py2);
                const float dxA = __fmul_rn(dx, r1.x), ndxB = __fmul_rn(dx, -r1.y);
                const f32x2 t = fma2(bc(dx), bc(dxA), mul2(dy, mul2(dy, bc(r1.z))));
                const f32x2 power = fma2(t, bc(-0.5f), mul2(dy, bc(ndxB)));
                float pa, pb, ala, alb, tta, ttb;
                const float Ta = T.x, Tb = T.y;
                const f32x2 Tp = pk(Ta, Tb);
                unpk(power, pa, pb);
                unpk(mul2(bc(r1.w), exp2x(pa, pb)), ala, alb);
                ala = fminf(0.99f, ala);
                alb = fminf(0.99f, alb);
                // (a finished pixel, T < 0, needs no test of its own: T*(1-alpha) < 0 sends it down the `fin` path below,
                //  which zeroes its alpha and keeps -|T|)
                const bool skip_a = (pa > 0.0f) || (ala < 1.0f / 255.0f);
                const bool skip_b = (pb > 0.0f) || (alb < 1.0f / 255.0f);
                ala = skip_a ? 0.f : ala;
                alb = skip_b ? 0.f : alb;
                unpk(mul2(Tp, sub2(bc(1.f), pk(ala, alb))), tta, ttb);
                // forward.cu:352-356: the entry that would take T below 1e-4 ends the pixel and is not blended.
                // (a finished pixel has T < 0 and alpha = 0 here, so it lands in the same branch and keeps -|T|)
                const bool fin_a = tta < 0.0001f, fin_b = ttb < 0.0001f;
                ala = fin_a ? 0.f : ala;
                alb = fin_b ? 0.f : alb;
                const f32x2 al = pk(ala, alb);
                const float4 r2 = sm.rec2[j];
                C0 = fma2(Tp, mul2(al, bc(r2.x)), C0);
                C1 = fma2(Tp, mul2(al, bc(r2.y)), C1);
                C2 = fma2(Tp, mul2(al, bc(r2.z)), C2);
                Wt = fma2(Tp, al, Wt);
                Dacc = fma2(Tp, mul2(al, bc(r2.w)), Dacc);
                T.x = fin_a ? -fabsf(Ta) : tta;
                T.y = fin_b ? -fabsf(Tb) : ttb;
                const uint32_t pos = pos0 + (uint32_t)b;
                last_a = (skip_a || fin_a) ? last_a : pos;
                last_b = (skip_b || fin_b) ? last_b : pos;
            }
        }
    }

    const size_t plane = (size_t)W * H;
    float Ta, Tb, c0a, c0b, c1a, c1b, c2a, c2b, wa, wb, da, db;
    Ta = fabsf(T.x); Tb = fabsf(T.y); unpk(C0, c0a, c0b); unpk(C1, c1a, c1b); unpk(C2, c2a, c2b); unpk(Wt, wa, wb); unpk(Dacc, da, db);
    if (inside_a) {
        const size_t pid = (size_t)W * pix_ya + pix_x;
        n_contrib[pid] = last_a;
        out_color[pid] = __fmaf_rn(bg[0], Ta, c0a);
        out_color[plane + pid] = __fmaf_rn(bg[1], Ta, c1a);
        out_color[2 * plane + pid] = __fmaf_rn(bg[2], Ta, c2a);
        out_alpha[pid] = wa;
        out_depth[pid] = da;
    }
    if (inside_b) {
        const size_t pid = (size_t)W * pix_yb + pix_x;
        n_contrib[pid] = last_b;
        out_color[pid] = __fmaf_rn(bg[0], Tb, c0b);
        out_color[plane + pid] = __fmaf_rn(bg[1], Tb, c1b);
        out_color[2 * plane + pid] = __fmaf_rn(bg[2], Tb, c2b);
        out_alpha[pid] = wb;
        out_depth[pid] = db;
    }
}

// Exhaustive check of exp2x() against expf() over a range of float bit patterns (test hook).
__global__ void expf_check_kernel(uint32_t first_bits, uint32_t count, unsigned long long* mismatches)
{
    unsigned long long bad = 0;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (uint64_t)gridDim.x * blockDim.x) {
        const float x = __uint_as_float(first_bits + (uint32_t)i);
        float a, b;
        unpk(exp2x(x, x * 0.5f), a, b);
        const float ra = expf(x), rb = expf(x * 0.5f);
        bad += (__float_as_uint(a) != __float_as_uint(ra)) + (__float_as_uint(b) != __float_as_uint(rb));
    }
    if (bad) atomicAdd(mismatches, bad);
}

void launch_expf_check(uint32_t first_bits, uint32_t count, unsigned long long* mismatches, cudaStream_t s)
{
    cudaMemsetAsync(mismatches, 0, sizeof(unsigned long long), s);
    expf_check_kernel<<<148 * 8, 256, 0, s>>>(first_bits, count, mismatches);
}

// Work counters for roofline accounting (see gsr_debug_pair_counts in include/gsr.h).
__global__ void __launch_bounds__(BLEND_THREADS)
pair_counts_kernel(int W, int H, int grid_x, const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list,
                   GeomView g, const uint32_t* __restrict__ n_contrib, unsigned long long* __restrict__ counts)
{
    __shared__ Staged sm;
    const int tile = blockIdx.x;
    const int tile_x = tile % grid_x, tile_y = tile / grid_x;
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int pix_x = tile_x * TILE + (warp & 1) * FOOT_W + (lane & 7), pix_y = tile_y * TILE + (warp >> 1) * FOOT_H + (lane >> 3);
    const bool inside = pix_x < W && pix_y < H;
    const float pxf = (float)pix_x, pyf = (float)pix_y;
    const uint2 range = load_range(ranges, tile);
    const int total = (int)(range.y - range.x);
    const int last = inside ? (int)n_contrib[(size_t)W * pix_y + pix_x] : 0;
    unsigned long long in_range = 0, blended = 0, tile_rej = 0;
    for (int base = 0; base < total; base += BATCH) {
        __syncthreads();
        const int n = min(BATCH, total - base);
        if ((int)threadIdx.x < n) {
            stage_entry(sm, threadIdx.x, point_list[range.x + base + threadIdx.x], g, (float)(tile_x * TILE), (float)(tile_y * TILE));
            if (sm.thr[threadIdx.x] < 0.f) tile_rej++;
        }
        __syncthreads();
        for (int j = 0; j < n; j++) {
            if (base + j + 1 > last) break;
            in_range++;
            const float4 r0 = sm.rec0[j];
            const float4 r1 = sm.rec1[j];
            const float power = blend_power(__fadd_rn(r0.x, -pxf), __fadd_rn(r0.y, -pyf), r1.x, r1.y, r1.z);
            if (power > 0.0f) continue;
            const float alpha = fminf(0.99f, __fmul_rn(r1.w, expf(power)));
            if (alpha < 1.0f / 255.0f) continue;
            blended++;
        }
    }
    atomicAdd(&counts[0], in_range);
    atomicAdd(&counts[1], blended);
    if (threadIdx.x == 0) atomicAdd(&counts[2], (unsigned long long)total * TILE_PIX);
    atomicAdd(&counts[3], tile_rej);
}

void launch_pair_counts(int W, int H, const uint2* ranges, const uint32_t* point_list, const GeomView& g,
                        const uint32_t* n_contrib, unsigned long long* counts, cudaStream_t s)
{
    const int gx = div_up(W, TILE), gy = div_up(H, TILE);
    cudaMemsetAsync(counts, 0, 4 * sizeof(unsigned long long), s);
    pair_counts_kernel<<<gx * gy, BLEND_THREADS, 0, s>>>(W, H, gx, ranges, point_list, g, n_contrib, counts);
}

void launch_render_forward(int W, int H, const uint2* ranges, const uint32_t* point_list, const GeomView& g,
                           const float* bg, float* out_color, float* out_depth, float* out_alpha,
                           uint32_t* n_contrib, cudaStream_t s)
{
    const int gx = div_up(W, TILE), gy = div_up(H, TILE);
    launch_pdl(render_forward_packed_kernel, gx * gy, PK_THREADS, 0, s, W, H, gx, ranges, point_list, g, bg, out_color,
               out_depth, out_alpha, n_contrib);
}

}  // namespace gsr
