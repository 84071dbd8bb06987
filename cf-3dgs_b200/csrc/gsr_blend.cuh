// Pieces shared by the forward and backward blend kernels.
#pragma once
#include "gsr_common.cuh"
#include "gsr_math.cuh"

namespace gsr {

constexpr int BLEND_THREADS = 256;  // one 16x16 tile per CTA, 8 warps
#ifndef GSR_FWD_BATCH
#define GSR_FWD_BATCH 256
#endif
constexpr int BATCH = GSR_FWD_BATCH;  // tile-list entries staged per round (forward)
// each warp owns an 8x4 pixel footprint of the tile (tighter than the reference's 16x2 strip, which makes
// the per-warp rejection test below bite more often)
constexpr int FOOT_W = 8, FOOT_H = 4;

// Per-entry staged record (shared memory).  rec0/rec1 are read by every pixel test, rec2 only when blending.
template <int N>
struct StagedT {
    float4 rec0[N];  // x, y, -B/C, -B/A
    float4 rec1[N];  // conic A, B, C, opacity
    float4 rec2[N];  // r, g, b, depth
    float thr[N];    // ln(255*opacity): alpha < 1/255 <=> q > thr ;  -1 = rejected for the whole tile
};
using Staged = StagedT<BATCH>;

// Conservative rejection: true only if alpha < 1/255 is guaranteed for EVERY pixel centre inside
// [x_lo,x_hi] x [y_lo,y_hi].  The reference skips such pairs one by one (forward.cu:347-349), so skipping
// them wholesale changes no output bit.  q(d) = 1/2 (A dx^2 + C dy^2) + B dx dy is minimised over the
// rectangle (convex: the minimum lies on the face(s) nearest to the centre) and compared with
// ln(255*opacity) plus a margin that bounds the rounding error of both evaluations.
__device__ __forceinline__ bool reject_rect(float X, float Y, float nBoC, float nBoA, float A, float B, float C,
                                            float thr, float x_lo, float x_hi, float y_lo, float y_hi)
{
    const float dxl = X - x_hi, dxh = X - x_lo;
    const float dyl = Y - y_hi, dyh = Y - y_lo;
    const float cx = fminf(fmaxf(0.f, dxl), dxh);
    const float cy = fminf(fmaxf(0.f, dyl), dyh);
    float qmin = 3.0e38f;
    if (cx != 0.f) {
        const float dy = fminf(fmaxf(nBoC * cx, dyl), dyh);
        qmin = 0.5f * (A * cx * cx + C * dy * dy) + B * cx * dy;
    }
    if (cy != 0.f) {
        const float dx = fminf(fmaxf(nBoA * cy, dxl), dxh);
        qmin = fminf(qmin, 0.5f * (A * dx * dx + C * cy * cy) + B * dx * cy);
    }
    if (cx == 0.f && cy == 0.f) qmin = 0.f;
    const float mx = fmaxf(fabsf(dxl), fabsf(dxh)), my = fmaxf(fabsf(dyl), fabsf(dyh));
    const float S = 0.5f * (A * mx * mx + C * my * my) + fabsf(B) * mx * my;
    return qmin > thr + (2e-6f * S + 1e-5f);
}

// Stage one tile-list entry.  Also applies the rejection test for the whole tile.
template <int N>
__device__ __forceinline__ void stage_entry(StagedT<N>& sm, int slot, uint32_t gid, const GeomView& g, float tile_x0,
                                            float tile_y0, float rect_h = (float)(TILE - 1))
{
    const float2 xy = g.xy[gid];
    const float4 co = g.conic_op[gid];
    const float4 cd = g.rgb_depth[gid];
    const bool pd = (co.x > 0.f) && (co.z > 0.f) && (co.x * co.z - co.y * co.y > 0.f);
    float thr = 3.0e38f;  // never reject unless the conic is positive definite
    float nBoC = 0.f, nBoA = 0.f;
    if (pd) {
        thr = logf(255.f * co.w);   // NaN/-inf for opacity <= 0: NaN compares false (kept), -inf rejects
        nBoC = -co.y / co.z;
        nBoA = -co.y / co.x;
        if (reject_rect(xy.x, xy.y, nBoC, nBoA, co.x, co.y, co.z, thr, tile_x0, tile_x0 + (TILE - 1), tile_y0,
                        tile_y0 + rect_h))
            thr = -1.f;
    }
    sm.rec0[slot] = make_float4(xy.x, xy.y, nBoC, nBoA);
    sm.rec1[slot] = co;
    sm.rec2[slot] = cd;
    sm.thr[slot] = thr;
}

}  // namespace gsr
