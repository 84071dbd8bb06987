// Fused Gaussian-model step: the per-iteration PyTorch work around the rasterizer in a
// CF-3DGS training step, as three kernels.
//   gaussian_activate : opacity = sigmoid(raw), scale = exp(raw), rotation = raw / max(|raw|, 1e-12)
//                       (scene/gaussian_model_cf.py:41-56,115-172: the activations behind get_opacity / get_scaling /
//                       get_rotation)
//   adam_features     : Adam on the (P, M, 3) SH tensor, one thread per float (perfectly coalesced); coefficient 0 uses
//                       the f_dc learning rate, the rest the f_rest rate (gaussian_model_cf.py:254-266)
//   adam_gaussians    : per Gaussian: chain rule through the three activations (the gradients arriving from the
//                       rasterizer are w.r.t. the ACTIVATED values), Adam on xyz / opacity / scaling / rotation, and
//                       the densification statistics of add_densification_stats + max_radii2D
//                       (gaussian_model_cf.py:716-719, trainer/cf3dgs_trainer.py:147-152)
// Adam follows torch.optim.Adam (no amsgrad, no weight decay): m = b1 m + (1-b1) g;  v = b2 v + (1-b2) g^2;
//   p -= (lr / (1 - b1^t)) * m / (sqrt(v) / sqrt(1 - b2^t) + eps).
#include <cmath>
#include <cstdint>

#include "../../include/gsr.h"
#include "gsr_kernels.h"
#include "gsr_launch.cuh"

namespace gsr {
void set_last_error(const char* msg);

namespace {

struct AdamCfg {
    float b1, b2, eps;
    float inv_bias1;       // 1 / (1 - b1^t)
    float inv_sqrt_bias2;  // 1 / sqrt(1 - b2^t)
};

// Every operation is written with its rounding (no contraction left to the compiler), so that the several kernels that
// inline this function -- one thread per float, per float4, per Gaussian, the shared multi-GPU step -- give the same bits
// for the same inputs.
__device__ __forceinline__ float adam_update(float p, float g, float& m, float& v, float lr, const AdamCfg& c)
{
    m = __fmaf_rn(c.b1, m, __fmul_rn(1.f - c.b1, g));
    v = __fmaf_rn(c.b2, v, __fmul_rn(__fmul_rn(1.f - c.b2, g), g));
    const float denom = __fmaf_rn(__fsqrt_rn(v), c.inv_sqrt_bias2, c.eps);
    return __fmaf_rn(-__fmul_rn(lr, c.inv_bias1), __fdiv_rn(m, denom), p);
}

__global__ void gaussian_activate_kernel(int P, const float* __restrict__ opacity_raw, const float* __restrict__ log_scale,
                                         const float* __restrict__ rot_raw, float* __restrict__ opacity,
                                         float* __restrict__ scale, float* __restrict__ rot)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    opacity[i] = 1.f / (1.f + expf(-opacity_raw[i]));
    scale[3 * i] = expf(log_scale[3 * i]);
    scale[3 * i + 1] = expf(log_scale[3 * i + 1]);
    scale[3 * i + 2] = expf(log_scale[3 * i + 2]);
    const float4 q = reinterpret_cast<const float4*>(rot_raw)[i];
    const float n = fmaxf(sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w), 1e-12f);
    reinterpret_cast<float4*>(rot)[i] = make_float4(q.x / n, q.y / n, q.z / n, q.w / n);
}

// One thread per SH float of the first upd_floats floats of the rows [row0, row0 + n / upd_floats).  The gradient may
// be packed to its first grad_floats floats per Gaussian (the active bands): floats between grad_floats and upd_floats
// have gradient zero, which still decays their moments and moves the parameter by whatever the moments hold, exactly
// like torch.optim.Adam on a zero grad.  Floats from upd_floats on are not touched at all: the caller guarantees that
// they never had a non-zero gradient, so their moments are zero and Adam's update of them is exactly zero.
// The moments of coefficient 0 (the three "dc" floats of a row) live in dense (P, 3) arrays of their own, m_dc / v_dc: at SH
// degree 0 -- every local model, the global one's first thousand iterations -- they are all the step touches, and as
// 12 bytes out of every 192-byte row of m / v they cost three partial sectors per Gaussian.  Slots 0..2 of a row of m / v
// are unused.
__global__ void adam_features_kernel(size_t row0, size_t n, int floats_per_gaussian, int grad_floats, int upd_floats,
                                     float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                                     float* __restrict__ v, float* __restrict__ m_dc, float* __restrict__ v_dc,
                                     float lr_dc, float lr_rest, AdamCfg c, const float* __restrict__ bias_dev)
{
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (bias_dev) { c.inv_bias1 = bias_dev[0]; c.inv_sqrt_bias2 = bias_dev[1]; }
    const size_t row = row0 + i / (size_t)upd_floats;
    const int k = (int)(i % (size_t)upd_floats);
    i = row * (size_t)floats_per_gaussian + k;
    const float gi = k < grad_floats ? g[row * (size_t)grad_floats + k] : 0.f;
    float* const mp = k < 3 ? m_dc + 3 * row + k : m + i;
    float* const vp = k < 3 ? v_dc + 3 * row + k : v + i;
    float mm = *mp, vv = *vp;
    p[i] = adam_update(p[i], gi, mm, vv, k < 3 ? lr_dc : lr_rest, c);
    *mp = mm; *vp = vv;
}

struct GaussianStepArgs {
    int P;
    float *xyz, *opacity_raw, *log_scale, *rot_raw;                 // parameters (updated in place)
    const float *d_xyz, *d_opacity, *d_scale, *d_rot;               // gradients w.r.t. xyz and the ACTIVATED values
    float *m_xyz, *v_xyz, *m_op, *v_op, *m_sc, *v_sc, *m_rot, *v_rot;
    float lr_xyz, lr_op, lr_sc, lr_rot;
    const float* d_means2D;   // optional (P,3)
    const int* radii;         // optional
    float *grad_accum, *denom, *max_radii2D;   // optional (P)
};

__global__ void adam_gaussians_kernel(GaussianStepArgs a, AdamCfg c, const float* __restrict__ bias_dev)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= a.P) return;
    if (bias_dev) { c.inv_bias1 = bias_dev[0]; c.inv_sqrt_bias2 = bias_dev[1]; }
#pragma unroll
    for (int k = 0; k < 3; k++) {
        const int j = 3 * i + k;
        float mm = a.m_xyz[j], vv = a.v_xyz[j];
        a.xyz[j] = adam_update(a.xyz[j], a.d_xyz[j], mm, vv, a.lr_xyz, c);
        a.m_xyz[j] = mm; a.v_xyz[j] = vv;
    }
    {   // opacity = sigmoid(raw): d raw = d o * o (1 - o)
        const float raw = a.opacity_raw[i];
        const float o = 1.f / (1.f + expf(-raw));
        const float g = a.d_opacity[i] * o * (1.f - o);
        float mm = a.m_op[i], vv = a.v_op[i];
        a.opacity_raw[i] = adam_update(raw, g, mm, vv, a.lr_op, c);
        a.m_op[i] = mm; a.v_op[i] = vv;
    }
#pragma unroll
    for (int k = 0; k < 3; k++) {   // scale = exp(raw): d raw = d s * s
        const int j = 3 * i + k;
        const float raw = a.log_scale[j];
        const float g = a.d_scale[j] * expf(raw);
        float mm = a.m_sc[j], vv = a.v_sc[j];
        a.log_scale[j] = adam_update(raw, g, mm, vv, a.lr_sc, c);
        a.m_sc[j] = mm; a.v_sc[j] = vv;
    }
    {   // rotation = q / max(|q|, eps): d q = (g - y (y . g)) / n
        const float4 q = reinterpret_cast<const float4*>(a.rot_raw)[i];
        const float4 gy = reinterpret_cast<const float4*>(a.d_rot)[i];
        const float n2 = q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w;
        const float n = sqrtf(n2);
        float4 g;
        if (n > 1e-12f) {
            const float inv = 1.f / n;
            const float4 y = make_float4(q.x * inv, q.y * inv, q.z * inv, q.w * inv);
            const float d = y.x * gy.x + y.y * gy.y + y.z * gy.z + y.w * gy.w;
            g = make_float4((gy.x - y.x * d) * inv, (gy.y - y.y * d) * inv, (gy.z - y.z * d) * inv, (gy.w - y.w * d) * inv);
        } else {
            g = make_float4(gy.x * 1e12f, gy.y * 1e12f, gy.z * 1e12f, gy.w * 1e12f);
        }
        float4 mm = reinterpret_cast<float4*>(a.m_rot)[i], vv = reinterpret_cast<float4*>(a.v_rot)[i];
        float4 o;
        o.x = adam_update(q.x, g.x, mm.x, vv.x, a.lr_rot, c);
        o.y = adam_update(q.y, g.y, mm.y, vv.y, a.lr_rot, c);
        o.z = adam_update(q.z, g.z, mm.z, vv.z, a.lr_rot, c);
        o.w = adam_update(q.w, g.w, mm.w, vv.w, a.lr_rot, c);
        reinterpret_cast<float4*>(a.rot_raw)[i] = o;
        reinterpret_cast<float4*>(a.m_rot)[i] = mm;
        reinterpret_cast<float4*>(a.v_rot)[i] = vv;
    }
    if (a.radii && a.grad_accum) {   // densification statistics for Gaussians visible in this view
        const int r = a.radii[i];
        if (r > 0) {
            const float gx = a.d_means2D[3 * i], gy2 = a.d_means2D[3 * i + 1];
            a.grad_accum[i] += sqrtf(gx * gx + gy2 * gy2);
            a.denom[i] += 1.f;
            if (a.max_radii2D) a.max_radii2D[i] = fmaxf(a.max_radii2D[i], (float)r);
        }
    }
}


// ---------------------------------------------------------------------------------------------------------------------
// The step shared between the replicas of a multi-GPU run (batched global views).  Every replica owns
// the rows [row0, row0 + rows) of the model.  For its rows it needs the SUM of all replicas' gradients and has to hand the
// updated parameters to ALL replicas.  With the gradient and parameter arrays of the replicas mapped behind one NVSwitch
// multicast address each, that is a load and a store: multimem.ld_reduce.add returns the sum over the replicas (added in
// the switch), multimem.st writes every replica's copy.  So the reduce-scatter, the optimiser step and the all-gather of
// the NCCL form are ONE pass over this replica's rows; gradients cross NVLink once (reduced), parameters once.
// MC = false: the same kernels on ordinary addresses (one process, or a test): load and store are plain.
template <bool MC>
__device__ __forceinline__ float4 ld_sum4(const float* p)
{
    float4 r;
    if (MC)
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                     : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p) : "memory");
    else
        r = *reinterpret_cast<const float4*>(p);
    return r;
}
template <bool MC>
__device__ __forceinline__ float ld_sum1(const float* p)
{
    float r;
    if (MC)
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.f32 %0, [%1];" : "=f"(r) : "l"(p) : "memory");
    else
        r = *p;
    return r;
}
template <bool MC>
__device__ __forceinline__ void st_all4(float* p, const float4 v)
{
    if (MC)
        asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};"
                     :: "l"(p), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
    else
        *reinterpret_cast<float4*>(p) = v;
}
template <bool MC>
__device__ __forceinline__ void st_all1(float* p, const float v)
{
    if (MC)
        asm volatile("multimem.st.relaxed.sys.global.f32 [%0], %1;" :: "l"(p), "f"(v) : "memory");
    else
        *p = v;
}

struct SharedStepArgs {
    int row0, rows;                                                   // this replica's rows (multiples of 4)
    const float *xyz, *opacity_raw, *log_scale, *rot_raw;             // this replica's copy (read)
    float *xyz_all, *opacity_raw_all, *log_scale_all, *rot_raw_all;   // every replica's copy (written)
    const float *d_xyz, *d_opacity, *d_scale, *d_rot;                 // sum over the replicas' gradients (read)
    float *m_xyz, *v_xyz, *m_op, *v_op, *m_sc, *v_sc, *m_rot, *v_rot; // this replica's moments
    float lr_xyz, lr_op, lr_sc, lr_rot;
};

__device__ __forceinline__ float4 adam_update4(const float4 p, const float4 g, float4& m, float4& v, float lr, const AdamCfg& c)
{
    float4 o;
    o.x = adam_update(p.x, g.x, m.x, v.x, lr, c);
    o.y = adam_update(p.y, g.y, m.y, v.y, lr, c);
    o.z = adam_update(p.z, g.z, m.z, v.z, lr, c);
    o.w = adam_update(p.w, g.w, m.w, v.w, lr, c);
    return o;
}

// The per-Gaussian block of the rows this replica owns, as four streams of float4: rotations (one quaternion each),
// then xyz, opacity and scale as flat element-wise ranges (their chain rules are per element).  One thread per float4.
template <bool MC>
__global__ void adam_shared_gaussians_kernel(SharedStepArgs a, AdamCfg c, const float* __restrict__ bias_dev)
{
    // The SH kernel behind this one touches other arrays: it is launched as a programmatic dependent and may start while
    // this grid is still running (it waits for this grid only at its very end, so that its completion implies ours).
    pdl_trigger();
    const int rows = a.rows;
    const int n_rot = rows, n_xyz = 3 * (rows / 4), n_op = rows / 4, n_sc = 3 * (rows / 4);
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_rot + n_xyz + n_op + n_sc) return;
    if (bias_dev) { c.inv_bias1 = bias_dev[0]; c.inv_sqrt_bias2 = bias_dev[1]; }
    if (t < n_rot) {   // rotation = q / max(|q|, eps): d q = (g - y (y . g)) / n
        const size_t f = (size_t)4 * (a.row0 + t);
        const float4 q = *reinterpret_cast<const float4*>(a.rot_raw + f);
        const float4 gy = ld_sum4<MC>(a.d_rot + f);
        const float n = sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w);
        float4 g;
        if (n > 1e-12f) {
            const float inv = 1.f / n;
            const float4 y = make_float4(q.x * inv, q.y * inv, q.z * inv, q.w * inv);
            const float d = y.x * gy.x + y.y * gy.y + y.z * gy.z + y.w * gy.w;
            g = make_float4((gy.x - y.x * d) * inv, (gy.y - y.y * d) * inv, (gy.z - y.z * d) * inv, (gy.w - y.w * d) * inv);
        } else {
            g = make_float4(gy.x * 1e12f, gy.y * 1e12f, gy.z * 1e12f, gy.w * 1e12f);
        }
        float4 mm = *reinterpret_cast<float4*>(a.m_rot + f), vv = *reinterpret_cast<float4*>(a.v_rot + f);
        st_all4<MC>(a.rot_raw_all + f, adam_update4(q, g, mm, vv, a.lr_rot, c));
        *reinterpret_cast<float4*>(a.m_rot + f) = mm;
        *reinterpret_cast<float4*>(a.v_rot + f) = vv;
        return;
    }
    t -= n_rot;
    if (t < n_xyz) {
        const size_t f = (size_t)3 * a.row0 + (size_t)4 * t;
        const float4 p = *reinterpret_cast<const float4*>(a.xyz + f);
        const float4 g = ld_sum4<MC>(a.d_xyz + f);
        float4 mm = *reinterpret_cast<float4*>(a.m_xyz + f), vv = *reinterpret_cast<float4*>(a.v_xyz + f);
        st_all4<MC>(a.xyz_all + f, adam_update4(p, g, mm, vv, a.lr_xyz, c));
        *reinterpret_cast<float4*>(a.m_xyz + f) = mm;
        *reinterpret_cast<float4*>(a.v_xyz + f) = vv;
        return;
    }
    t -= n_xyz;
    if (t < n_op) {    // opacity = sigmoid(raw): d raw = d o * o (1 - o)
        const size_t f = (size_t)a.row0 + (size_t)4 * t;
        const float4 p = *reinterpret_cast<const float4*>(a.opacity_raw + f);
        float4 g = ld_sum4<MC>(a.d_opacity + f);
        const float ox = 1.f / (1.f + expf(-p.x)), oy = 1.f / (1.f + expf(-p.y));
        const float oz = 1.f / (1.f + expf(-p.z)), ow = 1.f / (1.f + expf(-p.w));
        g = make_float4(g.x * ox * (1.f - ox), g.y * oy * (1.f - oy), g.z * oz * (1.f - oz), g.w * ow * (1.f - ow));
        float4 mm = *reinterpret_cast<float4*>(a.m_op + f), vv = *reinterpret_cast<float4*>(a.v_op + f);
        st_all4<MC>(a.opacity_raw_all + f, adam_update4(p, g, mm, vv, a.lr_op, c));
        *reinterpret_cast<float4*>(a.m_op + f) = mm;
        *reinterpret_cast<float4*>(a.v_op + f) = vv;
        return;
    }
    t -= n_op;
    {                  // scale = exp(raw): d raw = d s * s
        const size_t f = (size_t)3 * a.row0 + (size_t)4 * t;
        const float4 p = *reinterpret_cast<const float4*>(a.log_scale + f);
        float4 g = ld_sum4<MC>(a.d_scale + f);
        g = make_float4(g.x * expf(p.x), g.y * expf(p.y), g.z * expf(p.z), g.w * expf(p.w));
        float4 mm = *reinterpret_cast<float4*>(a.m_sc + f), vv = *reinterpret_cast<float4*>(a.v_sc + f);
        st_all4<MC>(a.log_scale_all + f, adam_update4(p, g, mm, vv, a.lr_sc, c));
        *reinterpret_cast<float4*>(a.m_sc + f) = mm;
        *reinterpret_cast<float4*>(a.v_sc + f) = vv;
    }
}

// SH rows [row0, row0 + rows); see adam_features_kernel for grad_floats / upd_floats.
//  VEC: one thread per float4 of a row's first ceil4(upd_floats) floats (floats_per_gaussian % 4 == 0, so parameters and
//       moments move as aligned float4).  Rounding upd_floats up touches floats that never had a gradient: their moments
//       are zero, Adam's update of them is exactly zero, they are rewritten with their own value.  The gradient comes as
//       one 16-byte sum when the packed row allows it (GVEC: grad_floats % 4 == 0), else float by float.
//  !VEC: one thread per float.
template <bool MC, bool VEC, bool GVEC>
__global__ void adam_shared_features_kernel(size_t row0, size_t n, int floats_per_gaussian, int grad_floats, int upd_floats,
                                            const float* __restrict__ p, float* p_all, const float* g,
                                            float* __restrict__ m, float* __restrict__ v, float* __restrict__ m_dc,
                                            float* __restrict__ v_dc, float lr_dc, float lr_rest,
                                            AdamCfg c, const float* __restrict__ bias_dev)
{
    constexpr int W = VEC ? 4 : 1;
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (bias_dev) { c.inv_bias1 = bias_dev[0]; c.inv_sqrt_bias2 = bias_dev[1]; }
    const size_t per_row = (size_t)((upd_floats + W - 1) / W);
    const size_t row = row0 + i / per_row;
    const int k = (int)(i % per_row) * W;
    i = row * (size_t)floats_per_gaussian + k;
    const float* grow = g + row * (size_t)grad_floats;
    if (VEC) {
        float4 gi = make_float4(0.f, 0.f, 0.f, 0.f);
        if (GVEC) {
            if (k < grad_floats) gi = ld_sum4<MC>(grow + k);
        } else {
            if (k < grad_floats) gi.x = ld_sum1<MC>(grow + k);
            if (k + 1 < grad_floats) gi.y = ld_sum1<MC>(grow + k + 1);
            if (k + 2 < grad_floats) gi.z = ld_sum1<MC>(grow + k + 2);
            if (k + 3 < grad_floats) gi.w = ld_sum1<MC>(grow + k + 3);
        }
        const float4 pp = *reinterpret_cast<const float4*>(p + i);
        float4 mm, vv, o;
        if (k == 0) {   // coefficient 0 (floats 0..2 of the row) keeps its moments in the dense arrays
            float* const md = m_dc + 3 * row;
            float* const vd = v_dc + 3 * row;
            mm = make_float4(md[0], md[1], md[2], m[i + 3]);
            vv = make_float4(vd[0], vd[1], vd[2], v[i + 3]);
            o.x = adam_update(pp.x, gi.x, mm.x, vv.x, lr_dc, c);
            o.y = adam_update(pp.y, gi.y, mm.y, vv.y, lr_dc, c);
            o.z = adam_update(pp.z, gi.z, mm.z, vv.z, lr_dc, c);
            o.w = adam_update(pp.w, gi.w, mm.w, vv.w, lr_rest, c);
            md[0] = mm.x; md[1] = mm.y; md[2] = mm.z; m[i + 3] = mm.w;
            vd[0] = vv.x; vd[1] = vv.y; vd[2] = vv.z; v[i + 3] = vv.w;
        } else {
            mm = *reinterpret_cast<float4*>(m + i);
            vv = *reinterpret_cast<float4*>(v + i);
            o = adam_update4(pp, gi, mm, vv, lr_rest, c);
            *reinterpret_cast<float4*>(m + i) = mm;
            *reinterpret_cast<float4*>(v + i) = vv;
        }
        st_all4<MC>(p_all + i, o);
    } else {
        const float gi = k < grad_floats ? ld_sum1<MC>(grow + k) : 0.f;
        float* const mp = k < 3 ? m_dc + 3 * row + k : m + i;
        float* const vp = k < 3 ? v_dc + 3 * row + k : v + i;
        float mm = *mp, vv = *vp;
        st_all1<MC>(p_all + i, adam_update(p[i], gi, mm, vv, k < 3 ? lr_dc : lr_rest, c));
        *mp = mm; *vp = vv;
    }
    pdl_wait();     // see adam_shared_gaussians_kernel
}

// SH degree 0 (grad_floats == upd_floats == 3: every local CF-3DGS model, and the global one's first thousand
// iterations): a block takes 256 rows, whose packed gradients are 192 consecutive float4 -- fetched as such (one
// 16-byte sum per thread instead of three 4-byte ones per row) and handed to the row's thread through shared memory.
template <bool MC>
__global__ void __launch_bounds__(256)
adam_shared_features_dc_kernel(int row0, int rows, int floats_per_gaussian, const float* __restrict__ p, float* p_all,
                               const float* g, float* __restrict__ m_dc, float* __restrict__ v_dc, float lr_dc,
                               AdamCfg c, const float* __restrict__ bias_dev)
{
    __shared__ __align__(16) float sg[768];
    const int first = blockIdx.x * 256;
    const int nr = min(256, rows - first);                  // a multiple of 4
    const float* gblock = g + (size_t)3 * (row0 + first);   // 16-byte aligned: row0 and first are multiples of 4
    if ((int)threadIdx.x < 3 * nr / 4)
        reinterpret_cast<float4*>(sg)[threadIdx.x] = ld_sum4<MC>(gblock + 4 * threadIdx.x);
    __syncthreads();
    if ((int)threadIdx.x >= nr) return;
    if (bias_dev) { c.inv_bias1 = bias_dev[0]; c.inv_sqrt_bias2 = bias_dev[1]; }
    const size_t row = (size_t)(row0 + first + threadIdx.x);
    const size_t i = row * (size_t)floats_per_gaussian;
    const float4 pp = *reinterpret_cast<const float4*>(p + i);
    float* const md = m_dc + 3 * row;      // dense: a warp's 32 rows are 384 consecutive bytes
    float* const vd = v_dc + 3 * row;
    float m0 = md[0], m1 = md[1], m2 = md[2], v0 = vd[0], v1 = vd[1], v2 = vd[2];
    float4 o;
    o.x = adam_update(pp.x, sg[3 * threadIdx.x], m0, v0, lr_dc, c);
    o.y = adam_update(pp.y, sg[3 * threadIdx.x + 1], m1, v1, lr_dc, c);
    o.z = adam_update(pp.z, sg[3 * threadIdx.x + 2], m2, v2, lr_dc, c);
    o.w = pp.w;                            // first float of coefficient 1: never had a gradient, rewritten as it is
    st_all4<MC>(p_all + i, o);
    md[0] = m0; md[1] = m1; md[2] = m2;
    vd[0] = v0; vd[1] = v1; vd[2] = v2;
    pdl_wait();     // see adam_shared_gaussians_kernel
}

// A one-coefficient SH tensor (M = 1: the dense coefficient-0 plane a model keeps while it is trained at SH degree 0,
// gsr_model.py shs()): parameters, gradients and the dense dc moments are all (P, 3), so the rows of a replica are one
// flat range of float4 -- the same stream as xyz.
template <bool MC>
__global__ void adam_shared_dc_plane_kernel(size_t f0, int n4, const float* __restrict__ p, float* p_all, const float* g,
                                            float* __restrict__ m_dc, float* __restrict__ v_dc, float lr_dc, AdamCfg c,
                                            const float* __restrict__ bias_dev)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t < n4) {
        if (bias_dev) { c.inv_bias1 = bias_dev[0]; c.inv_sqrt_bias2 = bias_dev[1]; }
        const size_t f = f0 + (size_t)4 * t;
        const float4 pp = *reinterpret_cast<const float4*>(p + f);
        const float4 gi = ld_sum4<MC>(g + f);
        float4 mm = *reinterpret_cast<float4*>(m_dc + f), vv = *reinterpret_cast<float4*>(v_dc + f);
        st_all4<MC>(p_all + f, adam_update4(pp, gi, mm, vv, lr_dc, c));
        *reinterpret_cast<float4*>(m_dc + f) = mm;
        *reinterpret_cast<float4*>(v_dc + f) = vv;
    }
    pdl_wait();     // see adam_shared_gaussians_kernel
}

// dst[i] = sum over the replicas of src[i], read through a multicast address (set-up check of the shared step)
__global__ void multicast_probe_kernel(const float* mc_src, float* __restrict__ dst, int n4)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n4) reinterpret_cast<float4*>(dst)[i] = ld_sum4<true>(mc_src + 4 * (size_t)i);
}

// add_densification_stats + max_radii2D alone (gaussian_model_cf.py:716-719): in the shared step they stay per replica
// (each replica saw its own view) and cover all P Gaussians.
__global__ void densify_stats_kernel(int P, const float* __restrict__ d_means2D, const int* __restrict__ radii,
                                     float* __restrict__ grad_accum, float* __restrict__ denom, float* __restrict__ max_radii2D)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const int r = radii[i];
    if (r > 0) {
        const float gx = d_means2D[3 * i], gy = d_means2D[3 * i + 1];
        grad_accum[i] += sqrtf(gx * gx + gy * gy);
        denom[i] += 1.f;
        if (max_radii2D) max_radii2D[i] = fmaxf(max_radii2D[i], (float)r);
    }
}

}  // namespace
}  // namespace gsr

using namespace gsr;

extern "C" int gsr_gaussian_activate(int P, const float* opacity_raw, const float* log_scale, const float* rot_raw,
                                     float* opacity, float* scale, float* rot, void* stream)
{
    if (P < 0 || (P > 0 && (!opacity_raw || !log_scale || !rot_raw || !opacity || !scale || !rot))) {
        set_last_error("gsr_gaussian_activate: bad argument");
        return GSR_ERR_INVALID;
    }
    if (P == 0) return GSR_OK;
    gaussian_activate_kernel<<<div_up(P, 256), 256, 0, (cudaStream_t)stream>>>(P, opacity_raw, log_scale, rot_raw, opacity,
                                                                             scale, rot);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { set_last_error(cudaGetErrorString(e)); return GSR_ERR_CUDA; }
    return GSR_OK;
}

// step counter kept on the device: [0] int64 step count, then two floats 1/(1-b1^t), 1/sqrt(1-b2^t)
__global__ void adam_tick_kernel(long long* step_state, float b1, float b2)
{
    const long long t = step_state[0] + 1;
    step_state[0] = t;
    float* f = reinterpret_cast<float*>(step_state + 1);
    f[0] = (float)(1.0 / (1.0 - pow((double)b1, (double)t)));
    f[1] = (float)(1.0 / sqrt(1.0 - pow((double)b2, (double)t)));
}

// what: bit 0 = the per-Gaussian block (xyz, opacity, scale, rotation, statistics), bit 1 = SH rows [row0, row0 + rows).
// tick: advance the device-side step counter first (once per optimiser step).
static int adam_step_impl(int P, int M, int M_grad, int M_upd, int what, int row0, int rows, int tick, int step, long long* step_state,
                          float beta1, float beta2, float eps,
                          const float* lrs_host, float* xyz, float* features, float* opacity_raw, float* log_scale,
                          float* rot_raw, const float* d_xyz, const float* d_features, const float* d_opacity,
                          const float* d_scale, const float* d_rot, float* state, const float* d_means2D,
                          const int* radii, float* grad_accum, float* denom, float* max_radii2D, void* stream)
{
    cudaStream_t s = (cudaStream_t)stream;
    if (P < 0 || M < 0 || (!step_state && step < 1) || !lrs_host) { set_last_error("gsr_gaussian_adam_step: bad argument"); return GSR_ERR_INVALID; }
    if (M_grad < 0 || M_grad > M || M_upd < 0 || M_upd > M || row0 < 0 || rows < 0 || (long long)row0 + rows > P) {
        set_last_error("gsr_gaussian_adam_step: M_grad <= M, M_upd <= M and the row range must lie inside P");
        return GSR_ERR_INVALID;
    }
    if (P == 0) return GSR_OK;
    if (!xyz || !opacity_raw || !log_scale || !rot_raw || !d_xyz || !d_opacity || !d_scale || !d_rot || !state ||
        (M > 0 && (!features || !d_features))) {
        set_last_error("gsr_gaussian_adam_step: required pointer is NULL");
        return GSR_ERR_INVALID;
    }
    if ((((uintptr_t)rot_raw) | ((uintptr_t)d_rot) | ((uintptr_t)state)) & 15) {
        set_last_error("gsr_gaussian_adam_step: rot_raw, d_rot and state must be 16-byte aligned");
        return GSR_ERR_INVALID;
    }
    AdamCfg c;
    c.b1 = beta1; c.b2 = beta2; c.eps = eps;
    c.inv_bias1 = c.inv_sqrt_bias2 = 1.f;
    const float* bias_dev = nullptr;
    if (step_state) {   // the step number lives on the device (the call can be captured in a CUDA graph and replayed)
        if (tick) adam_tick_kernel<<<1, 1, 0, s>>>(step_state, beta1, beta2);
        bias_dev = reinterpret_cast<const float*>(step_state + 1);
    } else {
        c.inv_bias1 = (float)(1.0 / (1.0 - std::pow((double)beta1, (double)step)));
        c.inv_sqrt_bias2 = (float)(1.0 / std::sqrt(1.0 - std::pow((double)beta2, (double)step)));
    }
    // state layout: [m | v] for rot(4P) xyz(3P) opacity(P) scale(3P) features-dc(3P) features(3MP), in that order.  The rotation moments
    // come first because the kernel moves them as float4: at 32*P bytes past a 16-byte aligned base everything behind
    // them may start at any 4-byte boundary (odd P), they may not.
    const size_t n = (size_t)P;
    float* st = state;
    GaussianStepArgs a;
    a.P = P;
    a.xyz = xyz; a.opacity_raw = opacity_raw; a.log_scale = log_scale; a.rot_raw = rot_raw;
    a.d_xyz = d_xyz; a.d_opacity = d_opacity; a.d_scale = d_scale; a.d_rot = d_rot;
    a.m_rot = st; st += 4 * n; a.v_rot = st; st += 4 * n;
    a.m_xyz = st; st += 3 * n; a.v_xyz = st; st += 3 * n;
    a.m_op = st; st += n; a.v_op = st; st += n;
    a.m_sc = st; st += 3 * n; a.v_sc = st; st += 3 * n;
    float* m_dc = st; st += 3 * n;
    float* v_dc = st; st += 3 * n;
    float* m_f = st; st += 3 * (size_t)M * n;
    float* v_f = st;
    a.lr_xyz = lrs_host[0]; a.lr_op = lrs_host[3]; a.lr_sc = lrs_host[4]; a.lr_rot = lrs_host[5];
    a.d_means2D = d_means2D; a.radii = radii; a.grad_accum = (d_means2D && radii && denom) ? grad_accum : nullptr;
    a.denom = denom; a.max_radii2D = max_radii2D;
    if (what & 1) adam_gaussians_kernel<<<div_up(P, 256), 256, 0, s>>>(a, c, bias_dev);
    if ((what & 2) && M_upd > 0 && rows > 0) {
        const int fpg = 3 * M, gf = 3 * M_grad, uf = 3 * M_upd;
        // rows of 16-byte multiples (M = 4, 16; gradient packed to 4 or 16 coefficients): one thread per float4 -- the
        // kernel of the shared step on plain addresses (0.23 ms against 0.30 ms for 1 M rows of 48 floats)
        if (!((fpg | gf) & 3) && !((((uintptr_t)features) | ((uintptr_t)d_features)) & 15)) {
            const size_t nf = (size_t)((uf + 3) / 4) * (size_t)rows;
            adam_shared_features_kernel<false, true, true><<<(unsigned)((nf + 255) / 256), 256, 0, s>>>(
                (size_t)row0, nf, fpg, gf, uf, features, features, d_features, m_f, v_f, m_dc, v_dc, lrs_host[1], lrs_host[2],
                c, bias_dev);
        } else {
            const size_t nf = (size_t)uf * (size_t)rows;
            adam_features_kernel<<<(unsigned)((nf + 255) / 256), 256, 0, s>>>((size_t)row0, nf, fpg, gf, uf, features,
                                                                          d_features, m_f, v_f, m_dc, v_dc, lrs_host[1],
                                                                          lrs_host[2], c, bias_dev);
        }
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { set_last_error(cudaGetErrorString(e)); return GSR_ERR_CUDA; }
    return GSR_OK;
}

extern "C" int gsr_gaussian_adam_step(int P, int M, int step, float beta1, float beta2, float eps, const float* lrs_host,
                                      float* xyz, float* features, float* opacity_raw, float* log_scale, float* rot_raw,
                                      const float* d_xyz, const float* d_features, const float* d_opacity,
                                      const float* d_scale, const float* d_rot, float* state, const float* d_means2D,
                                      const int* radii, float* grad_accum, float* denom, float* max_radii2D, void* stream)
{
    return adam_step_impl(P, M, M, M, 3, 0, P, 1, step, nullptr, beta1, beta2, eps, lrs_host, xyz, features, opacity_raw, log_scale, rot_raw,
                          d_xyz, d_features, d_opacity, d_scale, d_rot, state, d_means2D, radii, grad_accum, denom,
                          max_radii2D, stream);
}

extern "C" int gsr_gaussian_adam_step_dev(int P, int M, int64_t* step_state, float beta1, float beta2, float eps,
                                          const float* lrs_host, float* xyz, float* features, float* opacity_raw,
                                          float* log_scale, float* rot_raw, const float* d_xyz, const float* d_features,
                                          const float* d_opacity, const float* d_scale, const float* d_rot, float* state,
                                          const float* d_means2D, const int* radii, float* grad_accum, float* denom,
                                          float* max_radii2D, void* stream)
{
    if (!step_state) { set_last_error("gsr_gaussian_adam_step_dev: step_state is NULL"); return GSR_ERR_INVALID; }
    return adam_step_impl(P, M, M, M, 3, 0, P, 1, 0, reinterpret_cast<long long*>(step_state), beta1, beta2, eps, lrs_host, xyz, features,
                          opacity_raw, log_scale, rot_raw, d_xyz, d_features, d_opacity, d_scale, d_rot, state, d_means2D,
                          radii, grad_accum, denom, max_radii2D, stream);
}

extern "C" int gsr_gaussian_adam_step_part(int P, int M, int M_grad, int M_upd, int what, int row0, int rows, int tick,
                                           int64_t* step_state, float beta1, float beta2, float eps, const float* lrs_host,
                                           float* xyz, float* features, float* opacity_raw, float* log_scale, float* rot_raw,
                                           const float* d_xyz, const float* d_features, const float* d_opacity,
                                           const float* d_scale, const float* d_rot, float* state, const float* d_means2D,
                                           const int* radii, float* grad_accum, float* denom, float* max_radii2D,
                                           void* stream)
{
    if (!step_state) { set_last_error("gsr_gaussian_adam_step_part: step_state is NULL"); return GSR_ERR_INVALID; }
    return adam_step_impl(P, M, M_grad, M_upd, what, row0, rows, tick, 0, reinterpret_cast<long long*>(step_state), beta1, beta2, eps,
                          lrs_host, xyz, features, opacity_raw, log_scale, rot_raw, d_xyz, d_features, d_opacity, d_scale,
                          d_rot, state, d_means2D, radii, grad_accum, denom, max_radii2D, stream);
}

extern "C" size_t gsr_gaussian_adam_state_floats(int P, int M)
{
    return (size_t)2 * ((size_t)3 + 1 + 3 + 4 + 3 + (size_t)3 * M) * (size_t)(P > 0 ? P : 0);   // + 3: the dense dc moments
}

extern "C" int gsr_gaussian_adam_step_shared(int P_pad, int M, int M_grad, int M_upd, int row0, int rows, int multicast,
                                             int tick, int64_t* step_state, float beta1, float beta2, float eps,
                                             const float* lrs_host, const float* xyz, const float* features,
                                             const float* opacity_raw, const float* log_scale, const float* rot_raw,
                                             float* xyz_all, float* features_all, float* opacity_raw_all,
                                             float* log_scale_all, float* rot_raw_all, const float* d_xyz_sum,
                                             const float* d_features_sum, const float* d_opacity_sum,
                                             const float* d_scale_sum, const float* d_rot_sum, float* state, void* stream)
{
    cudaStream_t s = (cudaStream_t)stream;
    if (P_pad < 0 || M < 0 || !step_state || !lrs_host || M_grad < 0 || M_grad > M || M_upd < 0 || M_upd > M || row0 < 0 ||
        rows < 0 || (long long)row0 + rows > P_pad) {
        set_last_error("gsr_gaussian_adam_step_shared: bad size, or the row range lies outside P_pad");
        return GSR_ERR_INVALID;
    }
    if ((P_pad | row0 | rows) & 3) {
        set_last_error("gsr_gaussian_adam_step_shared: P_pad, row0 and rows must be multiples of 4 (the rows are moved as float4)");
        return GSR_ERR_INVALID;
    }
    if (P_pad == 0) return GSR_OK;
    if (!xyz || !opacity_raw || !log_scale || !rot_raw || !xyz_all || !opacity_raw_all || !log_scale_all || !rot_raw_all ||
        !d_xyz_sum || !d_opacity_sum || !d_scale_sum || !d_rot_sum || !state ||
        (M > 0 && (!features || !features_all || !d_features_sum))) {
        set_last_error("gsr_gaussian_adam_step_shared: required pointer is NULL");
        return GSR_ERR_INVALID;
    }
    const uintptr_t all_ptrs = (uintptr_t)xyz | (uintptr_t)features | (uintptr_t)opacity_raw | (uintptr_t)log_scale |
                               (uintptr_t)rot_raw | (uintptr_t)xyz_all | (uintptr_t)features_all | (uintptr_t)opacity_raw_all |
                               (uintptr_t)log_scale_all | (uintptr_t)rot_raw_all | (uintptr_t)d_xyz_sum |
                               (uintptr_t)d_features_sum | (uintptr_t)d_opacity_sum | (uintptr_t)d_scale_sum |
                               (uintptr_t)d_rot_sum | (uintptr_t)state;
    if (all_ptrs & 15) {
        set_last_error("gsr_gaussian_adam_step_shared: every array must be 16-byte aligned");
        return GSR_ERR_INVALID;
    }
    AdamCfg c;
    c.b1 = beta1; c.b2 = beta2; c.eps = eps;
    c.inv_bias1 = c.inv_sqrt_bias2 = 1.f;
    if (tick) adam_tick_kernel<<<1, 1, 0, s>>>(reinterpret_cast<long long*>(step_state), beta1, beta2);
    const float* bias_dev = reinterpret_cast<const float*>(reinterpret_cast<long long*>(step_state) + 1);
    // moments: the layout of gsr_gaussian_adam_step for P_pad Gaussians (every sub-array then starts 16-byte aligned)
    const size_t n = (size_t)P_pad;
    float* st = state;
    SharedStepArgs a;
    a.row0 = row0; a.rows = rows;
    a.xyz = xyz; a.opacity_raw = opacity_raw; a.log_scale = log_scale; a.rot_raw = rot_raw;
    a.xyz_all = xyz_all; a.opacity_raw_all = opacity_raw_all; a.log_scale_all = log_scale_all; a.rot_raw_all = rot_raw_all;
    a.d_xyz = d_xyz_sum; a.d_opacity = d_opacity_sum; a.d_scale = d_scale_sum; a.d_rot = d_rot_sum;
    a.m_rot = st; st += 4 * n; a.v_rot = st; st += 4 * n;
    a.m_xyz = st; st += 3 * n; a.v_xyz = st; st += 3 * n;
    a.m_op = st; st += n; a.v_op = st; st += n;
    a.m_sc = st; st += 3 * n; a.v_sc = st; st += 3 * n;
    float* m_dc = st; st += 3 * n;
    float* v_dc = st; st += 3 * n;
    float* m_f = st; st += 3 * (size_t)M * n;
    float* v_f = st;
    a.lr_xyz = lrs_host[0]; a.lr_op = lrs_host[3]; a.lr_sc = lrs_host[4]; a.lr_rot = lrs_host[5];
    if (rows > 0) {
        const int threads = rows + 3 * (rows / 4) + rows / 4 + 3 * (rows / 4);
        chain_break();      // ordinary launch: waits for everything before it in the stream
        if (multicast) launch_pdl(adam_shared_gaussians_kernel<true>, div_up(threads, 256), 256, 0, s, a, c, bias_dev);
        else launch_pdl(adam_shared_gaussians_kernel<false>, div_up(threads, 256), 256, 0, s, a, c, bias_dev);
    }
    if (M_upd > 0 && rows > 0) {
        const int fpg = 3 * M, gf = 3 * M_grad, uf = 3 * M_upd;
        const bool vec = !(fpg & 3), gvec = vec && !(gf & 3);
        const size_t nf = (size_t)(vec ? (uf + 3) / 4 : uf) * (size_t)rows;
        const unsigned blocks = (unsigned)((nf + 255) / 256);
#define GSR_FEAT(MC, VEC, GVEC)                                                                                     \
        launch_pdl(adam_shared_features_kernel<MC, VEC, GVEC>, blocks, 256, 0, s, (size_t)row0, nf, fpg, gf, uf,      \
                   features, features_all, d_features_sum, m_f, v_f, m_dc, v_dc, lrs_host[1], lrs_host[2], c, bias_dev)
        if (fpg == 3 && gf == 3 && uf == 3) {
            const int n4 = 3 * (rows / 4);
            if (multicast)
                launch_pdl(adam_shared_dc_plane_kernel<true>, div_up(n4, 256), 256, 0, s, (size_t)3 * row0, n4, features,
                           features_all, d_features_sum, m_dc, v_dc, lrs_host[1], c, bias_dev);
            else
                launch_pdl(adam_shared_dc_plane_kernel<false>, div_up(n4, 256), 256, 0, s, (size_t)3 * row0, n4, features,
                           features_all, d_features_sum, m_dc, v_dc, lrs_host[1], c, bias_dev);
        } else if (vec && gf == 3 && uf == 3 && fpg >= 4) {
            if (multicast)
                launch_pdl(adam_shared_features_dc_kernel<true>, div_up(rows, 256), 256, 0, s, row0, rows, fpg, features,
                           features_all, d_features_sum, m_dc, v_dc, lrs_host[1], c, bias_dev);
            else
                launch_pdl(adam_shared_features_dc_kernel<false>, div_up(rows, 256), 256, 0, s, row0, rows, fpg, features,
                           features_all, d_features_sum, m_dc, v_dc, lrs_host[1], c, bias_dev);
        } else
        if (multicast) { if (gvec) GSR_FEAT(true, true, true); else if (vec) GSR_FEAT(true, true, false); else GSR_FEAT(true, false, false); }
        else           { if (gvec) GSR_FEAT(false, true, true); else if (vec) GSR_FEAT(false, true, false); else GSR_FEAT(false, false, false); }
#undef GSR_FEAT
    }
    chain_break();
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { set_last_error(cudaGetErrorString(e)); return GSR_ERR_CUDA; }
    return GSR_OK;
}

extern "C" int gsr_densify_stats(int P, const float* d_means2D, const int* radii, float* grad_accum, float* denom,
                                 float* max_radii2D, void* stream)
{
    if (P < 0 || (P > 0 && (!d_means2D || !radii || !grad_accum || !denom))) {
        set_last_error("gsr_densify_stats: bad argument");
        return GSR_ERR_INVALID;
    }
    if (P == 0) return GSR_OK;
    densify_stats_kernel<<<div_up(P, 256), 256, 0, (cudaStream_t)stream>>>(P, d_means2D, radii, grad_accum, denom, max_radii2D);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { set_last_error(cudaGetErrorString(e)); return GSR_ERR_CUDA; }
    return GSR_OK;
}

extern "C" int gsr_multicast_probe(const float* mc_src, float* dst, int n, void* stream)
{
    if (!mc_src || !dst || n < 0 || (n & 3) || (((uintptr_t)mc_src | (uintptr_t)dst) & 15)) {
        set_last_error("gsr_multicast_probe: n must be a multiple of 4 and both pointers 16-byte aligned");
        return GSR_ERR_INVALID;
    }
    if (n == 0) return GSR_OK;
    multicast_probe_kernel<<<div_up(n / 4, 256), 256, 0, (cudaStream_t)stream>>>(mc_src, dst, n / 4);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { set_last_error(cudaGetErrorString(e)); return GSR_ERR_CUDA; }
    return GSR_OK;
}
