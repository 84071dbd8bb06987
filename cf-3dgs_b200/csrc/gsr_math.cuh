// Device math shared by the forward and backward per-Gaussian kernels.
//
// Rounding contract: every value that decides a tile key -- view depth, pixel centre,
// covariance, radius, tile rectangle -- is evaluated with explicitly rounded intrinsics
// (__fmul_rn / __fmaf_rn / __fadd_rn / __fdiv_rn / __fsqrt_rn, never re-contracted by nvcc) in the
// order the reference's sm_100a build evaluates them (read off its SASS; forward.cu:74-152,
// auxiliary.h:41-77).  Correctly rounded operations make the result a function of that order only,
// so keys, sorted order and ranges are bit-identical to the reference.  oracle/gsr_oracle.c states
// the same chains with fmaf().
#pragma once
#include "gsr_common.cuh"

namespace gsr {

// row r of (column-major 4x4 m) * (x, y, z, 1)
__device__ __forceinline__ float xform_row(const float* __restrict__ m, int r, float x, float y, float z)
{
    float t = __fmul_rn(y, m[4 + r]);
    t = __fmaf_rn(x, m[r], t);
    t = __fmaf_rn(z, m[8 + r], t);
    return __fadd_rn(t, m[12 + r]);
}

// p' = R p + t for a 3x4 row-major [R | t]; one fixed rounding sequence shared by forward and backward
__device__ __forceinline__ void rigid_apply(const float* __restrict__ T, float& x, float& y, float& z)
{
    const float nx = __fmaf_rn(T[2], z, __fmaf_rn(T[1], y, __fmaf_rn(T[0], x, T[3])));
    const float ny = __fmaf_rn(T[6], z, __fmaf_rn(T[5], y, __fmaf_rn(T[4], x, T[7])));
    const float nz = __fmaf_rn(T[10], z, __fmaf_rn(T[9], y, __fmaf_rn(T[8], x, T[11])));
    x = nx; y = ny; z = nz;
}

// Sigma = (S R)^T (S R) from scale * modifier and the un-normalised quaternion (w,x,y,z).
__device__ __forceinline__ void cov3d_from_scale_rot(float s0, float s1, float s2, float mod, float4 q, float* c3)
{
    const float sx = __fmul_rn(mod, s0), sy = __fmul_rn(mod, s1), sz = __fmul_rn(mod, s2);
    const float r = q.x, x = q.y, y = q.z, z = q.w;
    const float xz = __fmul_rn(x, z), rx = __fmul_rn(r, x), rz = __fmul_rn(r, z);
    const float yy = __fmul_rn(y, y), zz = __fmul_rn(z, z);
    const float A = __fmaf_rn(r, y, xz);
    const float B = __fmaf_rn(-r, y, xz);
    const float C = __fmaf_rn(y, z, -rx);
    const float Dd = __fmaf_rn(y, z, rx);
    const float E = __fmaf_rn(x, y, -rz);
    const float F = __fmaf_rn(x, y, rz);
    const float G = __fmaf_rn(x, x, yy);
    const float Hh = __fadd_rn(yy, zz);
    const float I = __fmaf_rn(x, x, zz);
    // M[c][k] = s_k * R[c][k] (glm column c)
    const float m00 = __fmul_rn(sx, __fadd_rn(1.f, -__fadd_rn(Hh, Hh)));
    const float m01 = __fmul_rn(sy, __fadd_rn(E, E));
    const float m02 = __fmul_rn(sz, __fadd_rn(A, A));
    const float m10 = __fmul_rn(sx, __fadd_rn(F, F));
    const float m11 = __fmul_rn(sy, __fadd_rn(1.f, -__fadd_rn(I, I)));
    const float m12 = __fmul_rn(sz, __fadd_rn(C, C));
    const float m20 = __fmul_rn(sx, __fadd_rn(B, B));
    const float m21 = __fmul_rn(sy, __fadd_rn(Dd, Dd));
    const float m22 = __fmul_rn(sz, __fadd_rn(1.f, -__fadd_rn(G, G)));
#define GSR_SIG(a0, a1, a2, b0, b1, b2) __fmaf_rn(a2, b2, __fmaf_rn(a0, b0, __fmul_rn(a1, b1)))
    c3[0] = GSR_SIG(m00, m01, m02, m00, m01, m02);
    c3[1] = GSR_SIG(m00, m01, m02, m10, m11, m12);
    c3[2] = GSR_SIG(m00, m01, m02, m20, m21, m22);
    c3[3] = GSR_SIG(m10, m11, m12, m10, m11, m12);
    c3[4] = GSR_SIG(m10, m11, m12, m20, m21, m22);
    c3[5] = GSR_SIG(m20, m21, m22, m20, m21, m22);
#undef GSR_SIG
}

struct Cov2D {
    float a, b, c;       // EWA covariance before the +0.3 low-pass
    float T0[3], T1[3];  // the two non-zero columns of glm T = W * J
    float tx, ty, tz;    // clamped view-space point
    float txtz, tytz;    // unclamped ratios (for the clamp masks in backward)
};

__device__ __forceinline__ void cov2d_chain(float px, float py, float pz, float fx, float fy, float tanx,
                                            float tany, const float* c3, const float* __restrict__ V, Cov2D& o)
{
    const float tx = xform_row(V, 0, px, py, pz);
    const float ty = xform_row(V, 1, px, py, pz);
    const float tz = xform_row(V, 2, px, py, pz);
    const float limx = __fmul_rn(1.3f, tanx), limy = __fmul_rn(1.3f, tany);
    o.txtz = __fdiv_rn(tx, tz);
    o.tytz = __fdiv_rn(ty, tz);
    o.tx = __fmul_rn(fminf(limx, fmaxf(-limx, o.txtz)), tz);
    o.ty = __fmul_rn(fminf(limy, fmaxf(-limy, o.tytz)), tz);
    o.tz = tz;
    const float tz2 = __fmul_rn(tz, tz);
    const float ja = __fdiv_rn(fx, tz), jb = __fdiv_rn(fy, tz);
    const float jc = -__fdiv_rn(__fmul_rn(o.tx, fx), tz2);
    const float jd = -__fdiv_rn(__fmul_rn(o.ty, fy), tz2);
#pragma unroll
    for (int r = 0; r < 3; r++) {
        o.T0[r] = __fmaf_rn(V[4 * r + 2], jc, __fmul_rn(V[4 * r], ja));
        o.T1[r] = __fmaf_rn(V[4 * r + 2], jd, __fmul_rn(V[4 * r + 1], jb));
    }
    // X[c][r] = sum_k T_r[k] * Vrk[k][c]
    const float v00 = c3[0], v01 = c3[1], v02 = c3[2], v11 = c3[3], v12 = c3[4], v22 = c3[5];
#define GSR_X(T, k0, k1, k2) __fmaf_rn(T[2], k2, __fmaf_rn(T[0], k0, __fmul_rn(T[1], k1)))
    const float x00 = GSR_X(o.T0, v00, v01, v02), x01 = GSR_X(o.T1, v00, v01, v02);
    const float x10 = GSR_X(o.T0, v01, v11, v12), x11 = GSR_X(o.T1, v01, v11, v12);
    const float x20 = GSR_X(o.T0, v02, v12, v22), x21 = GSR_X(o.T1, v02, v12, v22);
#undef GSR_X
    o.a = __fmaf_rn(x20, o.T0[2], __fmaf_rn(x00, o.T0[0], __fmul_rn(x10, o.T0[1])));
    o.b = __fmaf_rn(x21, o.T0[2], __fmaf_rn(x01, o.T0[0], __fmul_rn(x11, o.T0[1])));
    o.c = __fmaf_rn(x21, o.T1[2], __fmaf_rn(x01, o.T1[0], __fmul_rn(x11, o.T1[1])));
}

// auxiliary.h:41-44 is evaluated in double by the reference (1.0 / 0.5 literals).
__device__ __forceinline__ float ndc2pix(float v, int S)
{
    const double t = __fma_rn(__dadd_rn((double)v, 1.0), (double)S, -1.0);
    return __double2float_rn(__dmul_rn(t, 0.5));
}

// (tile|depth) falloff exponent with the reference's roundings (forward.cu:336-340).
__device__ __forceinline__ float blend_power(float dx, float dy, float A, float B, float C)
{
    const float t = __fmaf_rn(dx, __fmul_rn(dx, A), __fmul_rn(dy, __fmul_rn(dy, C)));
    return __fmaf_rn(t, -0.5f, -__fmul_rn(dy, __fmul_rn(dx, B)));
}

__device__ const float SH_C0 = 0.28209479177387814f;
__device__ const float SH_C1 = 0.4886025119029199f;
__device__ const float SH_C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                                   -1.0925484305920792f, 0.5462742152960396f};
__device__ const float SH_C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                                   0.3731763325901154f, -0.4570457994644658f, 1.445305721320277f,
                                   -0.5900435899266435f};

// Real SH basis up to degree 3 for unit direction (x,y,z): colour = sum_k b[k] * sh[k].
__device__ __forceinline__ void sh_basis(int deg, float x, float y, float z, float* b)
{
    b[0] = SH_C0;
    if (deg > 0) {
        b[1] = -SH_C1 * y; b[2] = SH_C1 * z; b[3] = -SH_C1 * x;
        if (deg > 1) {
            const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
            b[4] = SH_C2[0] * xy; b[5] = SH_C2[1] * yz; b[6] = SH_C2[2] * (2.0f * zz - xx - yy);
            b[7] = SH_C2[3] * xz; b[8] = SH_C2[4] * (xx - yy);
            if (deg > 2) {
                b[9] = SH_C3[0] * y * (3.0f * xx - yy);
                b[10] = SH_C3[1] * xy * z;
                b[11] = SH_C3[2] * y * (4.0f * zz - xx - yy);
                b[12] = SH_C3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy);
                b[13] = SH_C3[4] * x * (4.0f * zz - xx - yy);
                b[14] = SH_C3[5] * z * (xx - yy);
                b[15] = SH_C3[6] * x * (xx - 3.0f * yy);
            }
        }
    }
}

}  // namespace gsr
