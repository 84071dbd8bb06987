// Forward per-Gaussian preprocess and the frustum test.
// Behaviour: reference forward.cu:155-256 (preprocessCUDA), auxiliary.h:139-164 (in_frustum),
// rasterizer_impl.cu:54-66 (checkFrustum).  Layout and fusion are ours: one pass produces packed
// 16-byte records for the blend kernels, the depth sort key, the tile rectangle, and the total
// instance count (block reduction + one 64-bit atomic per block) so no separate scan is needed to
// learn num_rendered.
#include "gsr_kernels.h"
#include "gsr_launch.cuh"
#include "gsr_math.cuh"

namespace gsr {

constexpr int PRE_THREADS = 128;
// SH rows are staged through shared memory when M == 16 and the active degree needs more than the DC term: the block's
// rows are one contiguous 24 KB span, moved with coalesced 128-bit loads; rows padded to 13 float4 keep the per-thread
// LDS.128 conflict free (same layout as the backward kernel).
constexpr int PRE_ROW4 = 13;

template <bool SLAB>
#ifndef GSR_PRE_MIN_BLOCKS
#define GSR_PRE_MIN_BLOCKS 8
#endif
__global__ void __launch_bounds__(PRE_THREADS, GSR_PRE_MIN_BLOCKS)
preprocess_kernel(PreprocessArgs a, GeomView g)
{
    __shared__ float sV[16], sP[16], sCam[3];
    __shared__ uint32_t warp_sums[PRE_THREADS / 32];
    __shared__ float4 slab[SLAB ? PRE_THREADS * PRE_ROW4 : 1];
    pdl_wait();
    pdl_trigger();
    if (SLAB) {
        const int first = blockIdx.x * PRE_THREADS;
        const int rows = min(PRE_THREADS, a.P - first);
        const float4* src = reinterpret_cast<const float4*>(a.shs + (size_t)first * 48);
        for (int q = threadIdx.x; q < rows * 12; q += PRE_THREADS) slab[(q / 12) * PRE_ROW4 + (q % 12)] = src[q];
    }
    if (threadIdx.x < 16) {
        sV[threadIdx.x] = a.viewmatrix[threadIdx.x];
        sP[threadIdx.x] = a.projmatrix[threadIdx.x];
    }
    if (threadIdx.x < 3) sCam[threadIdx.x] = a.campos[threadIdx.x];
    __syncthreads();

    const int idx = blockIdx.x * PRE_THREADS + threadIdx.x;
    uint32_t n_tiles = 0;
    bool violated = false;
    if (idx < a.P) {
        int radius = 0;
        uint32_t depth_key = CULLED_KEY;
        float px = a.means3D[3 * idx], py = a.means3D[3 * idx + 1], pz = a.means3D[3 * idx + 2];
        if (a.rigid) rigid_apply(a.rigid, px, py, pz);
        const float depth = xform_row(sV, 2, px, py, pz);
        if (!(depth <= 0.2f)) {  // auxiliary.h:154 (NaN depth is not culled there either)
            const float hw = xform_row(sP, 3, px, py, pz);
            const float hx = xform_row(sP, 0, px, py, pz);
            const float hy = xform_row(sP, 1, px, py, pz);
            const float pw = __frcp_rn(__fadd_rn(hw, 0.0000001f));
            const float projx = __fmul_rn(hx, pw), projy = __fmul_rn(hy, pw);
            float c3[6];
            if (a.cov3D_precomp) {
#pragma unroll
                for (int k = 0; k < 6; k++) c3[k] = a.cov3D_precomp[6 * (size_t)idx + k];
            } else {
                const float4 q = reinterpret_cast<const float4*>(a.rotations)[idx];
                cov3d_from_scale_rot(a.scales[3 * idx], a.scales[3 * idx + 1], a.scales[3 * idx + 2],
                                     a.scale_modifier, q, c3);
            }
            const float fy = __fdiv_rn((float)a.H, __fmul_rn(2.0f, a.tanfovy));
            const float fx = __fdiv_rn((float)a.W, __fmul_rn(2.0f, a.tanfovx));
            Cov2D cv;
            cov2d_chain(px, py, pz, fx, fy, a.tanfovx, a.tanfovy, c3, sV, cv);
            const float ca = __fadd_rn(cv.a, 0.3f), cb = cv.b, cc = __fadd_rn(cv.c, 0.3f);
            const float det = __fmaf_rn(ca, cc, -__fmul_rn(cb, cb));
            if (det != 0.0f) {
                const float det_inv = __frcp_rn(det);
                const float mid = __fmul_rn(__fadd_rn(ca, cc), 0.5f);
                const float sq = __fsqrt_rn(fmaxf(__fmaf_rn(mid, mid, -det), 0.1f));
                const float lam = fmaxf(__fadd_rn(mid, sq), __fadd_rn(mid, -sq));
                const int rad = __float2int_ru(__fmul_rn(3.f, __fsqrt_rn(lam)));
                const float pix_x = ndc2pix(projx, a.W), pix_y = ndc2pix(projy, a.H);
                const int gx = (a.W + TILE - 1) / TILE, gy = (a.H + TILE - 1) / TILE;
                const float rf = (float)rad;
                int x0 = __float2int_rz(__fmul_rn(__fadd_rn(pix_x, -rf), 0.0625f));
                int y0 = __float2int_rz(__fmul_rn(__fadd_rn(pix_y, -rf), 0.0625f));
                int x1 = __float2int_rz(__fmul_rn(__fadd_rn(__fadd_rn(__fadd_rn(pix_x, rf), 16.f), -1.f), 0.0625f));
                int y1 = __float2int_rz(__fmul_rn(__fadd_rn(__fadd_rn(__fadd_rn(pix_y, rf), 16.f), -1.f), 0.0625f));
                const uint32_t ux0 = min((uint32_t)gx, (uint32_t)max(0, x0));
                const uint32_t uy0 = min((uint32_t)gy, (uint32_t)max(0, y0));
                const uint32_t ux1 = min((uint32_t)gx, (uint32_t)max(0, x1));
                const uint32_t uy1 = min((uint32_t)gy, (uint32_t)max(0, y1));
                n_tiles = (ux1 - ux0) * (uy1 - uy0);
                if (n_tiles != 0) {
                    float cr, cg_, cb_;
                    uint8_t clamp_bits = 0;
                    if (a.colors_precomp) {
                        cr = a.colors_precomp[3 * idx];
                        cg_ = a.colors_precomp[3 * idx + 1];
                        cb_ = a.colors_precomp[3 * idx + 2];
                    } else {
                        float dx = px - sCam[0], dy = py - sCam[1], dz = pz - sCam[2];
                        const float len = sqrtf(dx * dx + dy * dy + dz * dz);
                        dx = dx / len; dy = dy / len; dz = dz / len;
                        float basis[16];
                        sh_basis(a.D, dx, dy, dz, basis);
                        const int ncoef = (a.D + 1) * (a.D + 1);
                        float r = 0.f, gg = 0.f, b = 0.f;
                        if (SLAB) {
#pragma unroll
                            for (int k = 0; k < 16; k++)
                                if (k >= ncoef) basis[k] = 0.f;
                            float acc3[3] = {0.f, 0.f, 0.f};
                            const float4* row = slab + threadIdx.x * PRE_ROW4;
#pragma unroll
                            for (int c = 0; c < 12; c++) {
                                const float4 v = row[c];
                                acc3[(4 * c) % 3] += basis[(4 * c) / 3] * v.x;
                                acc3[(4 * c + 1) % 3] += basis[(4 * c + 1) / 3] * v.y;
                                acc3[(4 * c + 2) % 3] += basis[(4 * c + 2) / 3] * v.z;
                                acc3[(4 * c + 3) % 3] += basis[(4 * c + 3) / 3] * v.w;
                            }
                            r = acc3[0]; gg = acc3[1]; b = acc3[2];
                        } else {
                            const float* sh = a.shs + (size_t)idx * a.M * 3;
#pragma unroll
                            for (int k = 0; k < 16; k++) {
                                if (k < ncoef) {
                                    r += basis[k] * sh[3 * k];
                                    gg += basis[k] * sh[3 * k + 1];
                                    b += basis[k] * sh[3 * k + 2];
                                }
                            }
                        }
                        r += 0.5f; gg += 0.5f; b += 0.5f;
                        clamp_bits = (r < 0.f ? 1 : 0) | (gg < 0.f ? 2 : 0) | (b < 0.f ? 4 : 0);
                        cr = fmaxf(r, 0.f); cg_ = fmaxf(gg, 0.f); cb_ = fmaxf(b, 0.f);
                    }
                    radius = rad;
                    depth_key = __float_as_uint(depth);
                    g.xy[idx] = make_float2(pix_x, pix_y);
                    g.conic_op[idx] = make_float4(__fmul_rn(cc, det_inv), -__fmul_rn(cb, det_inv),
                                                  __fmul_rn(ca, det_inv), a.opacities[idx]);
                    g.rgb_depth[idx] = make_float4(cr, cg_, cb_, depth);
                    g.rect[idx] = make_uint2(ux0 | (ux1 << 16), uy0 | (uy1 << 16));
                    g.clamped[idx] = clamp_bits;
                }
            }
        } else if (a.prefiltered) {
            violated = true;
        }
        a.radii[idx] = radius;
        g.depth_key[idx] = depth_key;
        g.tiles[idx] = n_tiles;
    }
    // total number of tile instances: warp shuffle -> block -> one atomic
    uint32_t v = n_tiles;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if ((threadIdx.x & 31) == 0) warp_sums[threadIdx.x >> 5] = v;
    const int any_violation = __syncthreads_or(violated);
    if (threadIdx.x == 0) {
        unsigned long long tot = 0;
#pragma unroll
        for (int w = 0; w < PRE_THREADS / 32; w++) tot += warp_sums[w];
        if (tot) atomicAdd(&g.counters[0], tot);
        if (any_violation) atomicOr(&g.counters[0], 1ull << 62);
    }
}

void launch_preprocess(const PreprocessArgs& a, const GeomView& g, cudaStream_t s)
{
    cudaMemsetAsync(g.counters, 0, 4 * sizeof(unsigned long long), s);
    chain_break();   // the kernel must not start before the counters are cleared
    const bool slab = a.shs && !a.colors_precomp && a.M == 16 && a.D >= 1 && (((uintptr_t)a.shs) & 15) == 0;
    if (slab) launch_pdl(preprocess_kernel<true>, div_up(a.P, PRE_THREADS), PRE_THREADS, 0, s, a, g);
    else launch_pdl(preprocess_kernel<false>, div_up(a.P, PRE_THREADS), PRE_THREADS, 0, s, a, g);
}

__global__ void mark_visible_kernel(int P, const float* __restrict__ means, const float* __restrict__ view,
                                    uint8_t* __restrict__ present)
{
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P) return;
    const float d = xform_row(view, 2, means[3 * idx], means[3 * idx + 1], means[3 * idx + 2]);
    present[idx] = (d <= 0.2f) ? 0 : 1;
}

void launch_mark_visible(int P, const float* means3D, const float* view, uint8_t* present, cudaStream_t s)
{
    mark_visible_kernel<<<div_up(P, 256), 256, 0, s>>>(P, means3D, view, present);
}

}  // namespace gsr
