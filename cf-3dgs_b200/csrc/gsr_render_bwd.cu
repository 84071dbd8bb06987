// Backward tile blend.  Semantics: reference backward.cu:415-601 (renderCUDA): per pixel, back to front from
// n_contrib, T restarts at T_final = 1 - alpha_out and is unwound with T = T / (1 - alpha) -- the same
// recurrence, not an algebraic equivalent, so nearly opaque pixels behave like the reference.
//
// The reference issues 10 global float atomics per blended (pixel, Gaussian) pair.  Here the 32 lanes of a warp always
// look at the same Gaussian, so the partials are summed across the warp first: every lane adds its two pixels and
// drops its 9 (10 with a depth cotangent) partials into a warp-private shared-memory slab ([value][lane], rows padded
// to 36 floats so both the scalar stores and the 128-bit loads are bank-conflict free); once three entries are parked,
// lane L sums slab row L (eight LDS.128) and adds it into a table shared by the CTA's 4 warps -- ~30 issue slots per
// entry instead of the ~100 a 5-step shuffle butterfly over 10 values costs (ncu: the butterfly was 70 % of the first
// kernel's instructions; same-address shared atomics from the contributing lanes measured slower still).  The CTA
// flushes one 48-byte record per touched Gaussian with three 16-byte vector reductions (red.global.add.v4.f32; SASS
// REDG.E.ADD.F32x4.FTZ: the vector form flushes denormal operands and results to zero -- partial sums below 1.2e-38 are
// dropped, far under the 1e-4 relative bar; the reference's scalar atomicAdd, REDG.E.ADD.F32.FTZ.RN, flushes as well).
// Entries that cannot reach alpha >= 1/255 in the tile / in the warp's 8x8 footprint are rejected exactly as in the
// forward.
//
// Gradient record per Gaussian (grad_scratch, 12 floats):
//   0 mean2D.x  1 mean2D.y  2 conic.x  3 conic.y | 4 conic.w  5 opacity  6 colour.r  7 colour.g |
//   8 colour.b  9 depth  10,11 unused
#include "gsr_blend.cuh"
#include "gsr_f32x2.cuh"
#include "gsr_kernels.h"
#include "gsr_launch.cuh"

namespace gsr {

#ifndef GSR_BWD_BATCH
#define GSR_BWD_BATCH 128      // tile-list entries staged per round
#endif
constexpr int NGRAD = 12;
constexpr int BWD_BATCH = GSR_BWD_BATCH;
constexpr int PKB_THREADS = 128;   // packed kernel: 4 warps, two pixels per lane
constexpr int RED_STRIDE = 36;  // floats per value row of the warp slab (32 lanes + pad; 144 B keeps LDS.128 aligned)

// ---------------------------------------------------------------------------------------------------------------------
// Packed variant: two pixels per lane, FP32 pairs (gsr_f32x2.cuh), 128 threads per tile, one 8x8 footprint per warp
// (lane (lx, ly) owns rows ly and ly + 4 of it) -- the mirror image of render_forward_packed_kernel.
//  * power, exp and alpha use the forward's exact operation sequence, so the "did this pair blend" decision is the
//    forward's, bit for bit.
//  * Per-pixel control flow is a select on alpha: a pixel that does not blend the entry runs it with alpha = 0, which
//    is an exact no-op for the recurrences (T/(1-0) = T, accum + 0*(c - accum) = accum) and zeroes every partial.
//  * The reference's lazy `accum_rec = last_alpha*last_color + (1-last_alpha)*accum_rec` becomes the equivalent eager
//    update accum += alpha*(c - accum) at the end of the entry: no last_alpha / last_color state, and it reuses the
//    difference the opacity gradient needs anyway.
//  * Factors that depend only on the entry (opacity, -1/2, W/2, H/2) are applied once per (tile, Gaussian) in the
//    batch flush instead of once per pixel pair.
//  * What is reduced per entry are the moments of m0 = G * dL/dalpha over the pixel offsets (sum m0, m0 dx, m0 dy, m0 dx^2,
//    m0 dx dy, m0 dy^2); the mean and conic gradients are linear in them with per-entry coefficients.
//  * The two pixels of a lane are added before the slab reduction, so that reduction runs once per 64 pixels.
//  * PARAMS = false (pose-only fits): no opacity / colour partials, five (six with depth) values per entry instead of nine.
#ifndef GSR_BWDP_MIN_BLOCKS
#define GSR_BWDP_MIN_BLOCKS 6
#endif

template <bool HAS_DEPTH, bool HAS_ALPHA, bool PARAMS>
__global__ void __launch_bounds__(PKB_THREADS, GSR_BWDP_MIN_BLOCKS)
render_backward_packed_kernel(int W, int H, int grid_x, const uint2* __restrict__ ranges,
                              const uint32_t* __restrict__ point_list, GeomView g, const float* __restrict__ bg,
                              const float* __restrict__ out_alpha, const uint32_t* __restrict__ n_contrib,
                              const float* __restrict__ dL_dcolor, const float* __restrict__ dL_ddepth,
                              const float* __restrict__ dL_dalpha, float* __restrict__ grad)
{
    // PARAMS = false: the caller wants no opacity / colour / SH gradient and the colour does not feed the mean gradient
    // (SH degree 0 or precomputed colours) -- a pose-only fit.  Only the five geometric partials (+ depth) are reduced.
    constexpr int NV = 5 + (PARAMS ? 4 : 0) + (HAS_DEPTH ? 1 : 0);   // partials per entry that are reduced
    constexpr int GRP = (32 / NV) < 4 ? (32 / NV) : 4;               // pending entries per slab reduction (8-bit slots in pj)
    constexpr int ROWS = GRP * NV;
    static_assert(ROWS <= 32, "one lane sums one slab row");
    static_assert(BWD_BATCH <= PKB_THREADS, "one staging / flush thread per entry");
    __shared__ StagedT<BWD_BATCH> sm;
    __shared__ uint32_t sm_gid[BWD_BATCH];
    __shared__ __align__(16) float acc[BWD_BATCH][NGRAD];
    __shared__ __align__(16) float red[PKB_THREADS / 32][ROWS * RED_STRIDE];
    __shared__ int s_start;

    const int tile = blockIdx.x;
    const int tile_x = tile % grid_x, tile_y = tile / grid_x;
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int fx0 = tile_x * TILE + (warp & 1) * 8, fy0 = tile_y * TILE + (warp >> 1) * 8;
    const int pix_x = fx0 + (lane & 7), pix_ya = fy0 + (lane >> 3), pix_yb = pix_ya + 4;
    const bool in_a = pix_x < W && pix_ya < H, in_b = pix_x < W && pix_yb < H;
    const float pxf = (float)pix_x;
    const f32x2 py2 = pk((float)pix_ya, (float)pix_yb);
    const float foot_x0 = (float)fx0, foot_y0 = (float)fy0;
    const float tile_x0 = (float)(tile_x * TILE), tile_y0 = (float)(tile_y * TILE);

    pdl_wait();
    pdl_trigger();
    const uint2 range = load_range(ranges, tile);
    const int total = (int)(range.y - range.x);
    if (total == 0) return;

    const size_t plane = (size_t)W * H;
    const size_t pid_a = (size_t)W * pix_ya + pix_x, pid_b = (size_t)W * pix_yb + pix_x;
    const float Tfa = in_a ? (1.f - out_alpha[pid_a]) : 0.f, Tfb = in_b ? (1.f - out_alpha[pid_b]) : 0.f;
    // loop-carried pairs are kept as two floats and packed at their uses: ptxas then updates them in place (as 64-bit
    // values it computes into a temporary pair and copies back)
    float2 T = {Tfa, Tfb};
    const int last_a = in_a ? (int)n_contrib[pid_a] : 0, last_b = in_b ? (int)n_contrib[pid_b] : 0;
    const auto load2 = [&](const float* p) { return pk(in_a ? p[pid_a] : 0.f, in_b ? p[pid_b] : 0.f); };
    const f32x2 dLp0 = load2(dL_dcolor), dLp1 = load2(dL_dcolor + plane), dLp2 = load2(dL_dcolor + 2 * plane);
    const f32x2 dLd = HAS_DEPTH ? load2(dL_ddepth) : bc(0.f);
    const f32x2 dLa = HAS_ALPHA ? load2(dL_dalpha) : bc(0.f);
    // -T_final * (bg . dL_dpixel): the background term of dL/dalpha is this times 1/(1-alpha)
    const f32x2 nTf_bg = mul2(pk(-Tfa, -Tfb), fma2(bc(bg[2]), dLp2, fma2(bc(bg[1]), dLp1, mul2(bc(bg[0]), dLp0))));
    float2 ar0 = {0.f, 0.f}, ar1 = {0.f, 0.f}, ar2 = {0.f, 0.f}, adr = {0.f, 0.f}, aar = {0.f, 0.f};   // colour / depth / alpha behind

    if (threadIdx.x == 0) s_start = 0;
    __syncthreads();
    const int warp_last = __reduce_max_sync(0xffffffffu, max(last_a, last_b));
    if (lane == 0 && warp_last > 0) atomicMax(&s_start, warp_last);
    __syncthreads();
    const int start = s_start;              // process positions start-1 ... 0

    float* const slab = red[warp];
    const int my_grp = (int)lane / NV, my_comp = (int)lane - my_grp * NV;
    const float4* const my_row = reinterpret_cast<const float4*>(slab + lane * RED_STRIDE);
    int pending = 0;
    uint32_t pj = 0;    // slots of the pending entries, 8 bits each, newest in the low byte
    float* park = slab + lane;   // where this lane parks the partials of the next entry

    auto flush_slab = [&]() {
        __syncwarp();
        if ((int)lane < pending * NV) {
            // 32 floats -> 1: packed adds on the register pairs the 16-byte loads deliver
            f32x2 s2 = bc(0.f);
#pragma unroll
            for (int q = 0; q < 8; q += 2) {
                const float4 a = my_row[q], b = my_row[q + 1];
                s2 = add2(s2, add2(add2(pk(a.x, a.y), pk(a.z, a.w)), add2(pk(b.x, b.y), pk(b.z, b.w))));
            }
            const float sacc = lo(s2) + hi(s2);
            const int jj = (int)((pj >> (8 * (pending - 1 - my_grp))) & 0xffu);
            // column of the gradient record this partial belongs to: 0-4 geometry, 5 opacity, 6-8 colour, 9 depth
            const int col = (my_comp < 5 || PARAMS) ? my_comp : 9;
            atomicAdd(&acc[jj][col], sacc);
        }
        __syncwarp();
        pending = 0;
        park = slab + lane;
    };

    for (int base = 0; base < start; base += BWD_BATCH) {
        __syncthreads();   // previous flush finished
        const int n = min(BWD_BATCH, start - base);
        for (int k = threadIdx.x; k < BWD_BATCH * NGRAD / 4; k += PKB_THREADS)
            reinterpret_cast<float4*>(&acc[0][0])[k] = make_float4(0.f, 0.f, 0.f, 0.f);
        if ((int)threadIdx.x < n) {
            const uint32_t gid = point_list[range.x + (start - 1 - base - threadIdx.x)];
            sm_gid[threadIdx.x] = gid;
            stage_entry(sm, threadIdx.x, gid, g, tile_x0, tile_y0);
        }
        __syncthreads();

        for (int c0 = 0; c0 < n; c0 += 32) {
            const int chunk_first_pos = start - 1 - base - c0;   // highest position in the chunk
            if (chunk_first_pos - 31 >= warp_last) continue;
            const int e = c0 + lane;
            bool keep = false;
            if (e < n && (start - 1 - base - e) < warp_last) {
                const float thr = sm.thr[e];
                if (!(thr < 0.f)) {
                    const float4 r0 = sm.rec0[e];
                    const float4 r1 = sm.rec1[e];
                    keep = !reject_rect(r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, r1.z, thr, foot_x0, foot_x0 + 7.f, foot_y0,
                                        foot_y0 + 7.f);
                }
            }
            unsigned live = __ballot_sync(0xffffffffu, keep);
            while (live) {
                const int j = c0 + __ffs(live) - 1;
                live &= live - 1;
                const int pos = start - 1 - base - j;
                const float4 r0 = sm.rec0[j];
                const float4 r1 = sm.rec1[j];
                // forward's sequence: t = fma(dx, dx*A, dy*(dy*C)); power = fma(t, -0.5, -(dy*(dx*B)))
                const float dx = __fadd_rn(r0.x, -pxf);
                const f32x2 dy = sub2(bc(r0.y), py2);
                const float dxA = __fmul_rn(dx, r1.x), ndxB = __fmul_rn(dx, -r1.y);
                const f32x2 t = fma2(bc(dx), bc(dxA), mul2(dy, mul2(dy, bc(r1.z))));
                const f32x2 power = fma2(t, bc(-0.5f), mul2(dy, bc(ndxB)));
                float pa, pb, ala, alb;
                unpk(power, pa, pb);
                // exp of min(power, 0): same value wherever the pair can blend, and finite everywhere else, so the zeroed
                // partials of a non-blending pixel stay zero (0 * inf would not)
                const f32x2 G = exp2x(fminf(pa, 0.f), fminf(pb, 0.f));
                unpk(mul2(bc(r1.w), G), ala, alb);
                ala = fminf(0.99f, ala);
                alb = fminf(0.99f, alb);
                const bool ca = (pos < last_a) && !(pa > 0.0f) && !(ala < 1.0f / 255.0f);
                const bool cb = (pos < last_b) && !(pb > 0.0f) && !(alb < 1.0f / 255.0f);
                if (!__any_sync(0xffffffffu, ca || cb)) continue;
                const f32x2 al = pk(ca ? ala : 0.f, cb ? alb : 0.f);

                // T <- T / (1 - alpha): reciprocal seed + one residual correction
                const f32x2 nm = sub2(al, bc(1.f));            // -(1 - alpha)
                float nma, nmb, rca, rcb;
                unpk(nm, nma, nmb);
                asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(rca) : "f"(-nma));
                asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(rcb) : "f"(-nmb));
                const f32x2 rc = pk(rca, rcb);
                const f32x2 Tin = pk(T.x, T.y);
                const f32x2 q = mul2(Tin, rc);
                const f32x2 Tn = fma2(fma2(nm, q, Tin), rc, q);
                unpk(Tn, T.x, T.y);
                const f32x2 aT = mul2(al, Tn);

                // dL/dalpha = T * sum_ch (c_ch - behind_ch) * dL/dpixel_ch  (+ background term); behind += alpha*(c - behind)
                const float4 r2 = sm.rec2[j];
                f32x2 diff = sub2(bc(r2.x), pk(ar0.x, ar0.y));
                f32x2 d = mul2(diff, dLp0);
                unpk(fma2(al, diff, pk(ar0.x, ar0.y)), ar0.x, ar0.y);
                diff = sub2(bc(r2.y), pk(ar1.x, ar1.y));
                d = fma2(diff, dLp1, d);
                unpk(fma2(al, diff, pk(ar1.x, ar1.y)), ar1.x, ar1.y);
                diff = sub2(bc(r2.z), pk(ar2.x, ar2.y));
                d = fma2(diff, dLp2, d);
                unpk(fma2(al, diff, pk(ar2.x, ar2.y)), ar2.x, ar2.y);
                if (HAS_DEPTH) {
                    diff = sub2(bc(r2.w), pk(adr.x, adr.y));
                    d = fma2(diff, dLd, d);
                    unpk(fma2(al, diff, pk(adr.x, adr.y)), adr.x, adr.y);
                }
                if (HAS_ALPHA) {
                    diff = sub2(bc(1.f), pk(aar.x, aar.y));
                    d = fma2(diff, dLa, d);
                    unpk(fma2(al, diff, pk(aar.x, aar.y)), aar.x, aar.y);
                }
                d = mul2(d, Tn);
                d = fma2(nTf_bg, rc, d);   // background term; nTf_bg is 0 for a black background
                {
                    float da, db;
                    unpk(d, da, db);
                    d = pk(ca ? da : 0.f, cb ? db : 0.f);
                }
                // The six geometric partials are the moments of m0 = G*d over the pixel offsets: the mean and conic
                // gradients are per-entry linear combinations of them (applied in the batch flush):
                //   0,1: sum m0*dx, m0*dy    2,3,4: sum m0*dx*dx, m0*dx*dy, m0*dy*dy    5: sum m0
                f32x2 v[10];
                v[5] = mul2(G, d);
                v[0] = mul2(v[5], bc(dx));
                v[1] = mul2(v[5], dy);
                v[2] = mul2(v[0], bc(dx));
                v[3] = mul2(v[0], dy);
                v[4] = mul2(v[1], dy);
                v[6] = mul2(aT, dLp0);
                v[7] = mul2(aT, dLp1);
                v[8] = mul2(aT, dLp2);
                v[9] = HAS_DEPTH ? mul2(aT, dLd) : bc(0.f);
                {
#pragma unroll
                    for (int k = 0; k < NV; k++) {
                        const f32x2 vk = v[(k < 5 || PARAMS) ? k : 9];
                        park[k * RED_STRIDE] = lo(vk) + hi(vk);
                    }
                    park += NV * RED_STRIDE;      // running pointer: next pending entry's rows
                    pj = (pj << 8) | (uint32_t)j;
                    pending++;
                    if (pending == GRP) flush_slab();
                }
            }
        }
        if (pending) flush_slab();
        __syncthreads();
        // flush: apply the per-entry factors, then one record per entry with three 16-byte vector reductions
        if ((int)threadIdx.x < n) {
            float4 a0 = reinterpret_cast<const float4*>(&acc[threadIdx.x][0])[0];
            float4 a1 = reinterpret_cast<const float4*>(&acc[threadIdx.x][0])[1];
            const float4 a2 = reinterpret_cast<const float4*>(&acc[threadIdx.x][0])[2];
            const bool nz = (a0.x != 0.f) | (a0.y != 0.f) | (a0.z != 0.f) | (a0.w != 0.f) | (a1.x != 0.f) |
                            (a1.y != 0.f) | (a1.z != 0.f) | (a1.w != 0.f) | (a2.x != 0.f) | (a2.y != 0.f);
            if (nz) {
                // dL_dG = o*d;  dG/ddelx = -G (A dx + B dy), dG/ddely = -G (C dy + B dx);  ddelx/dx = W/2, ddely/dy = H/2
                // dL/dconic = -1/2 o (sum m0 dx^2, sum m0 dx dy, sum m0 dy^2)
                const float4 co = sm.rec1[threadIdx.x];
                const float nh = -0.5f * co.w;
                const float s1x = a0.x, s1y = a0.y;
                a0.x = nh * W * (co.x * s1x + co.y * s1y);
                a0.y = nh * H * (co.z * s1y + co.y * s1x);
                a0.z *= nh;
                a0.w *= nh;
                a1.x *= nh;
                float4* dst = reinterpret_cast<float4*>(grad + (size_t)sm_gid[threadIdx.x] * NGRAD);
                atomicAdd(dst, a0);
                atomicAdd(dst + 1, a1);
                atomicAdd(dst + 2, a2);
            }
        }
    }
}

void launch_render_backward(int W, int H, const uint2* ranges, const uint32_t* point_list, const GeomView& g,
                            const float* bg, const float* out_alpha, const uint32_t* n_contrib,
                            const float* dL_dcolor, const float* dL_ddepth, const float* dL_dalpha,
                            float* grad_scratch, cudaStream_t s, bool want_params)
{
    const int gx = div_up(W, TILE), gy = div_up(H, TILE);
    const dim3 grid(gx * gy);
#define GSR_LAUNCH(D, A)                                                                                              \
    do {                                                                                                              \
        if (want_params)                                                                                              \
            launch_pdl(render_backward_packed_kernel<D, A, true>, grid, PKB_THREADS, 0, s,                            \
                W, H, gx, ranges, point_list, g, bg, out_alpha, n_contrib, dL_dcolor, dL_ddepth, dL_dalpha,           \
                grad_scratch);                                                                                        \
        else                                                                                                          \
            launch_pdl(render_backward_packed_kernel<D, A, false>, grid, PKB_THREADS, 0, s,                           \
                W, H, gx, ranges, point_list, g, bg, out_alpha, n_contrib, dL_dcolor, dL_ddepth, dL_dalpha,           \
                grad_scratch);                                                                                        \
    } while (0)
    if (dL_ddepth && dL_dalpha) GSR_LAUNCH(true, true);
    else if (dL_ddepth) GSR_LAUNCH(true, false);
    else if (dL_dalpha) GSR_LAUNCH(false, true);
    else GSR_LAUNCH(false, false);
#undef GSR_LAUNCH
}

}  // namespace gsr
