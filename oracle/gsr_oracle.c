/*
 * gsr_oracle.c -- CPU restatement of the Gaussian-splatting rasterizer hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product path (cf-3dgs_b200/) may
 * import, link or call this file; it is the checker used by tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs.
 *
 * It restates, stage by stage, what the reference CUDA rasterizer computes
 * (all paths relative to /root/reference/submodules/diff-gaussian-rasterization):
 *   preprocess        cuda_rasterizer/forward.cu:155-256, auxiliary.h:41-56,139-164
 *   cov3D / cov2D     forward.cu:74-152
 *   SH -> RGB         forward.cu:20-71
 *   key emit          cuda_rasterizer/rasterizer_impl.cu:70-111
 *   sort              rasterizer_impl.cu:35-50,304-309 (stable, ascending, low 32+bit bits)
 *   tile ranges       rasterizer_impl.cu:116-138,311
 *   forward blend     forward.cu:261-381
 *   backward blend    backward.cu:415-601
 *   cov2D backward    backward.cu:144-274
 *   preprocess bwd    backward.cu:20-139,278-412
 *
 * Rounding: every value that decides a tile key (depth, pixel centre, radius,
 * tile rectangle) is evaluated with exactly the sequence of IEEE-754 roundings
 * the reference's sm_100a SASS performs (which products nvcc fused into FMAs was
 * read off `cuobjdump -sass`, see DESIGN.md "Rounding contract").  fmaf() is a
 * correctly rounded FMA, so those values are bit-exact on the CPU.  Compile with
 * -ffp-contract=off so gcc adds no fusions of its own.  expf() here is glibc's,
 * which differs from CUDA's expf in the last ulp: alpha-threshold decisions can
 * flip for a ~1e-6 fraction of (pixel, Gaussian) pairs; tests account for that.
 *
 * Parity pinning: the reference ships no golden vectors;
 * this oracle is pinned against the reference's own CUDA build (oracle/_ref)
 * through golden vectors that build produced on a B200 (tests/golden/*.npz, tests/test_golden.py); the CUDA path is
 * compared with the live reference build in tests/test_gpu_parity.py and tests/test_gpu_fuzz.py.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define TILE 16

static const float SH_C0 = 0.28209479177387814f;
static const float SH_C1 = 0.4886025119029199f;
static const float SH_C2[5] = {1.0925484305920792f, -1.0925484305920792f, 0.31539156525252005f,
                               -1.0925484305920792f, 0.5462742152960396f};
static const float SH_C3[7] = {-0.5900435899266435f, 2.890611442640554f, -0.4570457994644658f,
                               0.3731763325901154f, -0.4570457994644658f, 1.445305721320277f,
                               -0.5900435899266435f};

/* row r of (column-major 4x4) * (x,y,z,1): rn(y*m) -> fma(x,.) -> fma(z,.) -> + m[12+r]
 * (auxiliary.h:58-77 as compiled: SASS 0x0430-0x0470, 0x06b0-0x07a0) */
static inline float xform_row(const float* m, int r, float x, float y, float z)
{
    float t = y * m[4 + r];
    t = fmaf(x, m[r], t);
    t = fmaf(z, m[8 + r], t);
    return t + m[12 + r];
}

static inline float clampf(float v, float lo, float hi)
{
    /* FMNMX semantics: NaN loses */
    float t = fmaxf(lo, v);
    return fminf(hi, t);
}

/* forward.cu:118-152 with the roundings of SASS 0x0a50-0x0ef0 */
static void cov3d_from_scale_rot(const float* scale, float mod, const float* rot, float* out)
{
    const float sx = mod * scale[0], sy = mod * scale[1], sz = mod * scale[2];
    const float r = rot[0], x = rot[1], y = rot[2], z = rot[3];
    const float xz = x * z, rx = r * x, rz = r * z, yy = y * y, zz = z * z;
    const float A = fmaf(r, y, xz);    /* xz + ry */
    const float B = fmaf(-r, y, xz);   /* xz - ry */
    const float C = fmaf(y, z, -rx);   /* yz - rx */
    const float Dd = fmaf(y, z, rx);   /* yz + rx */
    const float E = fmaf(x, y, -rz);   /* xy - rz */
    const float F = fmaf(x, y, rz);    /* xy + rz */
    const float G = fmaf(x, x, yy);    /* xx + yy */
    const float Hh = yy + zz;          /* yy + zz : both products shared, plain add */
    const float I = fmaf(x, x, zz);    /* xx + zz */
    /* R[c][r], glm column-major */
    float R[3][3];
    R[0][0] = 1.f - (Hh + Hh); R[0][1] = E + E;           R[0][2] = A + A;
    R[1][0] = F + F;           R[1][1] = 1.f - (I + I);   R[1][2] = C + C;
    R[2][0] = B + B;           R[2][1] = Dd + Dd;         R[2][2] = 1.f - (G + G);
    float M[3][3];
    for (int c = 0; c < 3; c++) {
        M[c][0] = sx * R[c][0];
        M[c][1] = sy * R[c][1];
        M[c][2] = sz * R[c][2];
    }
#define SIG(i, j) fmaf(M[i][2], M[j][2], fmaf(M[i][0], M[j][0], M[i][1] * M[j][1]))
    out[0] = SIG(0, 0); out[1] = SIG(0, 1); out[2] = SIG(0, 2);
    out[3] = SIG(1, 1); out[4] = SIG(1, 2); out[5] = SIG(2, 2);
#undef SIG
}

/* forward.cu:74-113 with the roundings of SASS 0x0f60-0x1a90.
 * Also returns the 2x3 matrix T (T0 = column 0 of glm T, T1 = column 1). */
static void cov2d_chain(const float* p, float fx, float fy, float tanx, float tany,
                        const float* s /*cov3D[6]*/, const float* V,
                        float* cov /*a,b,c without the +0.3*/, float* T0, float* T1,
                        float* t_out /* clamped t */, float* txtz_out, float* tytz_out)
{
    const float tx = xform_row(V, 0, p[0], p[1], p[2]);
    const float ty = xform_row(V, 1, p[0], p[1], p[2]);
    const float tz = xform_row(V, 2, p[0], p[1], p[2]);
    const float limx = 1.3f * tanx, limy = 1.3f * tany;
    const float txtz = tx / tz, tytz = ty / tz;
    const float txc = clampf(txtz, -limx, limx) * tz;
    const float tyc = clampf(tytz, -limy, limy) * tz;
    const float tz2 = tz * tz;
    const float ja = fx / tz, jb = fy / tz;
    const float jc = -((txc * fx) / tz2);
    const float jd = -((tyc * fy) / tz2);
    for (int r = 0; r < 3; r++) {
        T0[r] = fmaf(V[4 * r + 2], jc, V[4 * r] * ja);
        T1[r] = fmaf(V[4 * r + 2], jd, V[4 * r + 1] * jb);
    }
    const float Vk[3][3] = {{s[0], s[1], s[2]}, {s[1], s[3], s[4]}, {s[2], s[4], s[5]}};
    float X[3][2]; /* X[c][r] */
    for (int c = 0; c < 3; c++) {
        X[c][0] = fmaf(T0[2], Vk[2][c], fmaf(T0[0], Vk[0][c], T0[1] * Vk[1][c]));
        X[c][1] = fmaf(T1[2], Vk[2][c], fmaf(T1[0], Vk[0][c], T1[1] * Vk[1][c]));
    }
    cov[0] = fmaf(X[2][0], T0[2], fmaf(X[0][0], T0[0], X[1][0] * T0[1]));
    cov[1] = fmaf(X[2][1], T0[2], fmaf(X[0][1], T0[0], X[1][1] * T0[1]));
    cov[2] = fmaf(X[2][1], T1[2], fmaf(X[0][1], T1[0], X[1][1] * T1[1]));
    if (t_out) { t_out[0] = txc; t_out[1] = tyc; t_out[2] = tz; }
    if (txtz_out) *txtz_out = txtz;
    if (tytz_out) *tytz_out = tytz;
}

/* auxiliary.h:41-44 -- evaluated in double, (v+1)*S-1 fused (SASS 0x1e40-0x1f30) */
static inline float ndc2pix(float v, int S)
{
    double t = fma((double)v + 1.0, (double)S, -1.0);
    return (float)(t * 0.5);
}

static inline int f2i_trunc(float f)
{
    if (f != f) return 0;
    if (f >= 2147483648.0f) return 2147483647;
    if (f <= -2147483648.0f) return (-2147483647 - 1);
    return (int)f;
}

/* auxiliary.h:46-56 */
static void tile_rect(float px, float py, int radius, int gx, int gy, uint32_t* rect)
{
    const float rf = (float)radius;
    int x0 = f2i_trunc((px - rf) * 0.0625f);
    int y0 = f2i_trunc((py - rf) * 0.0625f);
    int x1 = f2i_trunc((((px + rf) + 16.f) - 1.f) * 0.0625f);
    int y1 = f2i_trunc((((py + rf) + 16.f) - 1.f) * 0.0625f);
    x0 = x0 < 0 ? 0 : x0; y0 = y0 < 0 ? 0 : y0; x1 = x1 < 0 ? 0 : x1; y1 = y1 < 0 ? 0 : y1;
    rect[0] = (uint32_t)x0 < (uint32_t)gx ? (uint32_t)x0 : (uint32_t)gx;
    rect[1] = (uint32_t)y0 < (uint32_t)gy ? (uint32_t)y0 : (uint32_t)gy;
    rect[2] = (uint32_t)x1 < (uint32_t)gx ? (uint32_t)x1 : (uint32_t)gx;
    rect[3] = (uint32_t)y1 < (uint32_t)gy ? (uint32_t)y1 : (uint32_t)gy;
}

static void sh_to_rgb(int deg, const float* pos, const float* campos, const float* sh /*[M][3]*/,
                      float* rgb, uint8_t* clamped)
{
    float dx = pos[0] - campos[0], dy = pos[1] - campos[1], dz = pos[2] - campos[2];
    const float len = sqrtf(dx * dx + dy * dy + dz * dz);
    const float x = dx / len, y = dy / len, z = dz / len;
    for (int ch = 0; ch < 3; ch++) {
#define S(k) sh[(k) * 3 + ch]
        float res = SH_C0 * S(0);
        if (deg > 0) {
            res = res - SH_C1 * y * S(1) + SH_C1 * z * S(2) - SH_C1 * x * S(3);
            if (deg > 1) {
                const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
                res = res + SH_C2[0] * xy * S(4) + SH_C2[1] * yz * S(5) +
                      SH_C2[2] * (2.0f * zz - xx - yy) * S(6) + SH_C2[3] * xz * S(7) +
                      SH_C2[4] * (xx - yy) * S(8);
                if (deg > 2) {
                    res = res + SH_C3[0] * y * (3.0f * xx - yy) * S(9) + SH_C3[1] * xy * z * S(10) +
                          SH_C3[2] * y * (4.0f * zz - xx - yy) * S(11) +
                          SH_C3[3] * z * (2.0f * zz - 3.0f * xx - 3.0f * yy) * S(12) +
                          SH_C3[4] * x * (4.0f * zz - xx - yy) * S(13) +
                          SH_C3[5] * z * (xx - yy) * S(14) + SH_C3[6] * x * (xx - 3.0f * yy) * S(15);
                }
            }
        }
#undef S
        res += 0.5f;
        clamped[ch] = res < 0.f;
        rgb[ch] = res > 0.f ? res : 0.f;
    }
}

/* ------------------------------------------------------------------------- */
/* Stage 1: per-Gaussian preprocess.  Returns the number of tile instances.   */
/* Null conventions follow the reference: cov3D_pre / colors_pre / shs /      */
/* scales may be NULL.                                                        */
int64_t gsro_preprocess(int P, int D, int M, const float* means, const float* scales, float mod,
                        const float* rots, const float* opac, const float* shs,
                        const float* cov3D_pre, const float* colors_pre, const float* V,
                        const float* Pm, const float* campos, int W, int H, float tanx, float tany,
                        int* radii, float* xy, float* depths, float* cov3D, float* rgb,
                        float* conic_op, uint32_t* tiles, uint8_t* clamped, uint32_t* rects)
{
    const float fy = (float)H / (2.0f * tany);
    const float fx = (float)W / (2.0f * tanx);
    const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
    int64_t total = 0;
#pragma omp parallel for schedule(static) reduction(+ : total)
    for (int i = 0; i < P; i++) {
        radii[i] = 0;
        tiles[i] = 0;
        rects[4 * i] = rects[4 * i + 1] = rects[4 * i + 2] = rects[4 * i + 3] = 0;
        const float* p = means + 3 * i;
        const float depth = xform_row(V, 2, p[0], p[1], p[2]);
        if (depth <= 0.2f) continue; /* auxiliary.h:154 */
        const float hw = xform_row(Pm, 3, p[0], p[1], p[2]);
        const float hx = xform_row(Pm, 0, p[0], p[1], p[2]);
        const float hy = xform_row(Pm, 1, p[0], p[1], p[2]);
        const float pw = 1.0f / (hw + 0.0000001f);
        const float projx = hx * pw, projy = hy * pw;

        const float* c3;
        if (cov3D_pre) c3 = cov3D_pre + 6 * i;
        else {
            cov3d_from_scale_rot(scales + 3 * i, mod, rots + 4 * i, cov3D + 6 * i);
            c3 = cov3D + 6 * i;
        }
        float cov[3], T0[3], T1[3];
        cov2d_chain(p, fx, fy, tanx, tany, c3, V, cov, T0, T1, NULL, NULL, NULL);
        const float a = cov[0] + 0.3f, b = cov[1], c = cov[2] + 0.3f;
        const float det = fmaf(a, c, -(b * b));
        if (det == 0.0f) continue;
        const float det_inv = 1.f / det;
        const float con0 = c * det_inv, con1 = -(b * det_inv), con2 = a * det_inv;
        const float mid = (a + c) * 0.5f;
        const float q = fmaxf(fmaf(mid, mid, -det), 0.1f);
        const float sq = sqrtf(q);
        const float lam = fmaxf(mid + sq, mid - sq);
        const float rad_f = ceilf(3.f * sqrtf(lam));
        const int radius = f2i_trunc(rad_f);
        const float px = ndc2pix(projx, W), py = ndc2pix(projy, H);
        uint32_t rect[4];
        tile_rect(px, py, radius, gx, gy, rect);
        const uint32_t n = (rect[2] - rect[0]) * (rect[3] - rect[1]);
        if (n == 0) continue;
        if (!colors_pre) sh_to_rgb(D, p, campos, shs + (size_t)i * M * 3, rgb + 3 * i, clamped + 3 * i);
        depths[i] = depth;
        radii[i] = radius;
        xy[2 * i] = px; xy[2 * i + 1] = py;
        conic_op[4 * i] = con0; conic_op[4 * i + 1] = con1; conic_op[4 * i + 2] = con2;
        conic_op[4 * i + 3] = opac[i];
        tiles[i] = n;
        memcpy(rects + 4 * i, rect, sizeof(rect));
        total += n;
    }
    return total;
}

/* rasterizer_impl.cu:35-50 */
static uint32_t higher_msb(uint32_t n)
{
    uint32_t msb = sizeof(n) * 4, step = msb;
    while (step > 1) {
        step /= 2;
        if (n >> msb) msb += step; else msb -= step;
    }
    if (n >> msb) msb++;
    return msb;
}

/* Stage 2: emit (tile|depth) keys in Gaussian order, stable sort on the low
 * 32+bit bits, derive tile ranges.  rasterizer_impl.cu:70-138,290-319.
 * keys_unsorted/vals_unsorted may be NULL. */
int gsro_bin(int P, int W, int H, const int* radii, const uint32_t* rects, const float* depths,
             const uint32_t* tiles, int64_t R, uint64_t* keys_sorted, uint32_t* point_list,
             uint32_t* ranges /* 2 per tile */, uint64_t* keys_unsorted, uint32_t* vals_unsorted)
{
    const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
    uint64_t* k0 = (uint64_t*)malloc(sizeof(uint64_t) * (size_t)(R > 0 ? R : 1));
    uint64_t* k1 = (uint64_t*)malloc(sizeof(uint64_t) * (size_t)(R > 0 ? R : 1));
    uint32_t* v0 = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)(R > 0 ? R : 1));
    uint32_t* v1 = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)(R > 0 ? R : 1));
    if (!k0 || !k1 || !v0 || !v1) return -1;
    int64_t off = 0;
    for (int i = 0; i < P; i++) {
        if (radii[i] <= 0) continue;
        uint32_t dbits;
        memcpy(&dbits, depths + i, 4);
        for (uint32_t y = rects[4 * i + 1]; y < rects[4 * i + 3]; y++)
            for (uint32_t x = rects[4 * i]; x < rects[4 * i + 2]; x++) {
                k0[off] = ((uint64_t)(y * gx + x) << 32) | dbits;
                v0[off] = (uint32_t)i;
                off++;
            }
    }
    if (off != R) { free(k0); free(k1); free(v0); free(v1); return -2; }
    if (keys_unsorted) memcpy(keys_unsorted, k0, sizeof(uint64_t) * (size_t)R);
    if (vals_unsorted) memcpy(vals_unsorted, v0, sizeof(uint32_t) * (size_t)R);
    const int end_bit = 32 + (int)higher_msb((uint32_t)(gx * gy));
    /* LSD radix, 16-bit digits, stable */
    size_t* hist = (size_t*)malloc(sizeof(size_t) * 65537);
    for (int shift = 0; shift < end_bit; shift += 16) {
        int bits = end_bit - shift < 16 ? end_bit - shift : 16;
        uint64_t mask = ((uint64_t)1 << bits) - 1;
        memset(hist, 0, sizeof(size_t) * 65537);
        for (int64_t i = 0; i < R; i++) hist[((k0[i] >> shift) & mask) + 1]++;
        for (int d = 0; d < 65536; d++) hist[d + 1] += hist[d];
        for (int64_t i = 0; i < R; i++) {
            size_t pos = hist[(k0[i] >> shift) & mask]++;
            k1[pos] = k0[i];
            v1[pos] = v0[i];
        }
        uint64_t* tk = k0; k0 = k1; k1 = tk;
        uint32_t* tv = v0; v0 = v1; v1 = tv;
    }
    free(hist);
    memcpy(keys_sorted, k0, sizeof(uint64_t) * (size_t)R);
    memcpy(point_list, v0, sizeof(uint32_t) * (size_t)R);
    memset(ranges, 0, sizeof(uint32_t) * 2 * (size_t)gx * gy);
    for (int64_t i = 0; i < R; i++) {
        uint32_t cur = (uint32_t)(k0[i] >> 32);
        if (i == 0) ranges[2 * cur] = 0;
        else {
            uint32_t prev = (uint32_t)(k0[i - 1] >> 32);
            if (cur != prev) { ranges[2 * prev + 1] = (uint32_t)i; ranges[2 * cur] = (uint32_t)i; }
        }
        if (i == R - 1) ranges[2 * cur + 1] = (uint32_t)R;
    }
    free(k0); free(k1); free(v0); free(v1);
    return 0;
}

/* the (pixel, Gaussian) falloff with the reference's roundings, forward.cu:336-340,
 * SASS 0x0620-0x06a0 */
static inline float blend_power(float gx_, float gy_, float pxf, float pyf, float A, float B, float C,
                                float* dx_out, float* dy_out)
{
    const float dx = gx_ - pxf, dy = gy_ - pyf;
    const float t = fmaf(dx, dx * A, dy * (dy * C));
    *dx_out = dx; *dy_out = dy;
    return fmaf(t, -0.5f, -(dy * (dx * B)));
}

/* Stage 3: forward blend, forward.cu:261-381 (per pixel, front to back). */
void gsro_render(int W, int H, const uint32_t* ranges, const uint32_t* point_list, const float* xy,
                 const float* feat /*[P][3]*/, const float* depths, const float* conic_op,
                 const float* bg, float* out_color, float* out_depth, float* out_alpha,
                 uint32_t* n_contrib)
{
    const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
#pragma omp parallel for schedule(dynamic, 4)
    for (int tile = 0; tile < gx * gy; tile++) {
        const int tx = tile % gx, ty = tile / gx;
        const uint32_t r0 = ranges[2 * tile], r1 = ranges[2 * tile + 1];
        for (int ly = 0; ly < TILE; ly++)
            for (int lx = 0; lx < TILE; lx++) {
                const int px = tx * TILE + lx, py = ty * TILE + ly;
                if (px >= W || py >= H) continue;
                const float pxf = (float)px, pyf = (float)py;
                float T = 1.0f, C[3] = {0, 0, 0}, weight = 0, Dacc = 0;
                uint32_t contributor = 0, last = 0;
                for (uint32_t k = r0; k < r1; k++) {
                    contributor++;
                    const uint32_t g = point_list[k];
                    const float* co = conic_op + 4 * g;
                    float dx, dy;
                    const float power = blend_power(xy[2 * g], xy[2 * g + 1], pxf, pyf, co[0], co[1], co[2], &dx, &dy);
                    if (power > 0.0f) continue;
                    const float alpha = fminf(0.99f, co[3] * expf(power));
                    if (alpha < 1.0f / 255.0f) continue;
                    const float test_T = T * (1 - alpha);
                    if (test_T < 0.0001f) break;
                    for (int ch = 0; ch < 3; ch++) C[ch] = fmaf(T, alpha * feat[3 * g + ch], C[ch]);
                    weight = fmaf(T, alpha, weight);
                    Dacc = fmaf(T, alpha * depths[g], Dacc);
                    T = test_T;
                    last = contributor;
                }
                const size_t pid = (size_t)W * py + px;
                n_contrib[pid] = last;
                for (int ch = 0; ch < 3; ch++) out_color[(size_t)ch * H * W + pid] = fmaf(bg[ch], T, C[ch]);
                out_alpha[pid] = weight;
                out_depth[pid] = Dacc;
            }
    }
}

/* Stage 4: backward blend, backward.cu:415-601.  Per-Gaussian sums are kept in
 * double (the reference's float atomics have no defined order), written as float.
 * dL_dmean2D is [P][3] (z untouched), dL_dconic is [P][4] (x,y,w written). */
void gsro_render_backward(int P, int W, int H, const uint32_t* ranges, const uint32_t* point_list,
                          const float* xy, const float* feat, const float* depths,
                          const float* conic_op, const float* bg, const float* out_alpha,
                          const uint32_t* n_contrib, const float* dL_dpix, const float* dL_dpix_depth,
                          const float* dL_dpix_alpha, float* dL_dmean2D, float* dL_dconic,
                          float* dL_dopacity, float* dL_dcolors, float* dL_ddepths)
{
    const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
    double* acc = (double*)calloc((size_t)P * 10, sizeof(double));
    const float ddelx_dx = (float)(0.5 * W), ddely_dy = (float)(0.5 * H);
#pragma omp parallel for schedule(dynamic, 4)
    for (int tile = 0; tile < gx * gy; tile++) {
        const int tx = tile % gx, ty = tile / gx;
        const uint32_t r0 = ranges[2 * tile], r1 = ranges[2 * tile + 1];
        for (int ly = 0; ly < TILE; ly++)
            for (int lx = 0; lx < TILE; lx++) {
                const int px = tx * TILE + lx, py = ty * TILE + ly;
                if (px >= W || py >= H) continue;
                const size_t pid = (size_t)W * py + px;
                const float pxf = (float)px, pyf = (float)py;
                const float T_final = 1 - out_alpha[pid];
                float T = T_final;
                uint32_t contributor = r1 - r0;
                const uint32_t last_contributor = n_contrib[pid];
                float accum_rec[3] = {0, 0, 0}, dLp[3], last_color[3] = {0, 0, 0};
                for (int ch = 0; ch < 3; ch++) dLp[ch] = dL_dpix[(size_t)ch * H * W + pid];
                const float dLd = dL_dpix_depth[pid], dLa = dL_dpix_alpha[pid];
                float accum_depth_rec = 0, accum_alpha_rec = 0, last_alpha = 0, last_depth = 0;
                for (uint32_t k = r1; k-- > r0;) {
                    contributor--;
                    if (contributor >= last_contributor) continue;
                    const uint32_t g = point_list[k];
                    const float* co = conic_op + 4 * g;
                    float dx, dy;
                    const float power = blend_power(xy[2 * g], xy[2 * g + 1], pxf, pyf, co[0], co[1], co[2], &dx, &dy);
                    if (power > 0.0f) continue;
                    const float G = expf(power);
                    const float alpha = fminf(0.99f, co[3] * G);
                    if (alpha < 1.0f / 255.0f) continue;
                    T = T / (1.f - alpha);
                    const float aT = alpha * T;
                    float dL_dopa = 0.0f;
                    double* a = acc + (size_t)g * 10;
                    for (int ch = 0; ch < 3; ch++) {
                        const float c = feat[3 * g + ch];
                        accum_rec[ch] = last_alpha * last_color[ch] + (1.f - last_alpha) * accum_rec[ch];
                        last_color[ch] = c;
                        dL_dopa += (c - accum_rec[ch]) * dLp[ch];
                        const double v = (double)(aT * dLp[ch]);
#pragma omp atomic
                        a[6 + ch] += v;
                    }
                    const float c_d = depths[g];
                    accum_depth_rec = last_alpha * last_depth + (1.f - last_alpha) * accum_depth_rec;
                    last_depth = c_d;
                    dL_dopa += (c_d - accum_depth_rec) * dLd;
                    {
                        const double v = (double)(aT * dLd);
#pragma omp atomic
                        a[9] += v;
                    }
                    accum_alpha_rec = last_alpha + (1.f - last_alpha) * accum_alpha_rec;
                    dL_dopa += (1 - accum_alpha_rec) * dLa;
                    dL_dopa *= T;
                    last_alpha = alpha;
                    float bg_dot = 0;
                    for (int ch = 0; ch < 3; ch++) bg_dot += bg[ch] * dLp[ch];
                    dL_dopa += (-T_final / (1.f - alpha)) * bg_dot;

                    const float dL_dG = co[3] * dL_dopa;
                    const float gdx = G * dx, gdy = G * dy;
                    const float dG_ddelx = -gdx * co[0] - gdy * co[1];
                    const float dG_ddely = -gdy * co[2] - gdx * co[1];
                    const double v0 = (double)(dL_dG * dG_ddelx * ddelx_dx);
                    const double v1 = (double)(dL_dG * dG_ddely * ddely_dy);
                    const double v2 = (double)(-0.5f * gdx * dx * dL_dG);
                    const double v3 = (double)(-0.5f * gdx * dy * dL_dG);
                    const double v4 = (double)(-0.5f * gdy * dy * dL_dG);
                    const double v5 = (double)(G * dL_dopa);
#pragma omp atomic
                    a[0] += v0;
#pragma omp atomic
                    a[1] += v1;
#pragma omp atomic
                    a[2] += v2;
#pragma omp atomic
                    a[3] += v3;
#pragma omp atomic
                    a[4] += v4;
#pragma omp atomic
                    a[5] += v5;
                }
            }
    }
    for (int i = 0; i < P; i++) {
        const double* a = acc + (size_t)i * 10;
        dL_dmean2D[3 * i] = (float)a[0]; dL_dmean2D[3 * i + 1] = (float)a[1]; dL_dmean2D[3 * i + 2] = 0.f;
        dL_dconic[4 * i] = (float)a[2]; dL_dconic[4 * i + 1] = (float)a[3]; dL_dconic[4 * i + 2] = 0.f;
        dL_dconic[4 * i + 3] = (float)a[4];
        dL_dopacity[i] = (float)a[5];
        dL_dcolors[3 * i] = (float)a[6]; dL_dcolors[3 * i + 1] = (float)a[7]; dL_dcolors[3 * i + 2] = (float)a[8];
        dL_ddepths[i] = (float)a[9];
    }
    free(acc);
}

/* auxiliary.h:107-120 */
static void dnormvdv3(const float* v, const float* dv, float* out)
{
    const float sum2 = v[0] * v[0] + v[1] * v[1] + v[2] * v[2];
    const float invsum32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
    out[0] = ((+sum2 - v[0] * v[0]) * dv[0] - v[1] * v[0] * dv[1] - v[2] * v[0] * dv[2]) * invsum32;
    out[1] = (-v[0] * v[1] * dv[0] + (sum2 - v[1] * v[1]) * dv[1] - v[2] * v[1] * dv[2]) * invsum32;
    out[2] = (-v[0] * v[2] * dv[0] - v[1] * v[2] * dv[1] + (sum2 - v[2] * v[2]) * dv[2]) * invsum32;
}

/* Stage 5: backward of preprocess = computeCov2DCUDA (backward.cu:144-274) followed by
 * preprocessCUDA (backward.cu:346-412) incl. SH (20-139) and cov3D (278-341) backward.
 * All outputs are fully written (zeros for culled Gaussians).  Optional outputs may be NULL:
 * dL_dsh when shs==NULL, dL_dscale/dL_drot when scales==NULL.  dL_dcov3D is always written
 * (it is the returned gradient when cov3D was precomputed). */
void gsro_preprocess_backward(int P, int D, int M, const float* means, const int* radii,
                              const float* shs, const uint8_t* clamped, const float* scales,
                              const float* rots, float mod, const float* cov3Ds, const float* V,
                              const float* Pm, const float* campos, int W, int H, float tanx,
                              float tany, const float* dL_dmean2D /*[P][3]*/,
                              const float* dL_dconic /*[P][4]*/, float* dL_dcolor /*[P][3] in*/,
                              const float* dL_ddepth, float* dL_dmeans, float* dL_dcov3D,
                              float* dL_dsh, float* dL_dscale, float* dL_drot)
{
    const float fy = (float)H / (2.0f * tany);
    const float fx = (float)W / (2.0f * tanx);
#pragma omp parallel for schedule(static)
    for (int i = 0; i < P; i++) {
        float* dm = dL_dmeans + 3 * i;
        dm[0] = dm[1] = dm[2] = 0.f;
        for (int k = 0; k < 6; k++) dL_dcov3D[6 * i + k] = 0.f;
        if (dL_dsh) memset(dL_dsh + (size_t)i * M * 3, 0, sizeof(float) * (size_t)M * 3);
        if (dL_dscale) { dL_dscale[3 * i] = dL_dscale[3 * i + 1] = dL_dscale[3 * i + 2] = 0.f; }
        if (dL_drot) { for (int k = 0; k < 4; k++) dL_drot[4 * i + k] = 0.f; }
        if (!(radii[i] > 0)) continue;
        const float* p = means + 3 * i;
        const float* c3 = cov3Ds + 6 * i;
        /* ---- cov2D backward ---- */
        float cov[3], T0[3], T1[3], t[3], txtz, tytz;
        cov2d_chain(p, fx, fy, tanx, tany, c3, V, cov, T0, T1, t, &txtz, &tytz);
        const float limx = 1.3f * tanx, limy = 1.3f * tany;
        const float x_grad_mul = (txtz < -limx || txtz > limx) ? 0.f : 1.f;
        const float y_grad_mul = (tytz < -limy || tytz > limy) ? 0.f : 1.f;
        const float a = cov[0] + 0.3f, b = cov[1], c = cov[2] + 0.3f;
        const float dc0 = dL_dconic[4 * i], dc1 = dL_dconic[4 * i + 1], dc2 = dL_dconic[4 * i + 3];
        const float denom = a * c - b * b;
        float dL_da = 0, dL_db = 0, dL_dc = 0;
        const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
        float* dcv = dL_dcov3D + 6 * i;
        if (denom2inv != 0) {
            dL_da = denom2inv * (-c * c * dc0 + 2 * b * c * dc1 + (denom - a * c) * dc2);
            dL_dc = denom2inv * (-a * a * dc2 + 2 * a * b * dc1 + (denom - a * c) * dc0);
            dL_db = denom2inv * 2 * (b * c * dc0 - (denom + 2 * b * b) * dc1 + a * b * dc2);
            dcv[0] = (T0[0] * T0[0] * dL_da + T0[0] * T1[0] * dL_db + T1[0] * T1[0] * dL_dc);
            dcv[3] = (T0[1] * T0[1] * dL_da + T0[1] * T1[1] * dL_db + T1[1] * T1[1] * dL_dc);
            dcv[5] = (T0[2] * T0[2] * dL_da + T0[2] * T1[2] * dL_db + T1[2] * T1[2] * dL_dc);
            dcv[1] = 2 * T0[0] * T0[1] * dL_da + (T0[0] * T1[1] + T0[1] * T1[0]) * dL_db + 2 * T1[0] * T1[1] * dL_dc;
            dcv[2] = 2 * T0[0] * T0[2] * dL_da + (T0[0] * T1[2] + T0[2] * T1[0]) * dL_db + 2 * T1[0] * T1[2] * dL_dc;
            dcv[4] = 2 * T0[2] * T0[1] * dL_da + (T0[1] * T1[2] + T0[2] * T1[1]) * dL_db + 2 * T1[1] * T1[2] * dL_dc;
        }
        const float Vk[3][3] = {{c3[0], c3[1], c3[2]}, {c3[1], c3[3], c3[4]}, {c3[2], c3[4], c3[5]}};
        float dT0[3], dT1[3];
        for (int k = 0; k < 3; k++) {
            const float u0 = T0[0] * Vk[k][0] + T0[1] * Vk[k][1] + T0[2] * Vk[k][2];
            const float u1 = T1[0] * Vk[k][0] + T1[1] * Vk[k][1] + T1[2] * Vk[k][2];
            dT0[k] = 2 * u0 * dL_da + u1 * dL_db;
            dT1[k] = 2 * u1 * dL_dc + u0 * dL_db;
        }
        /* W[c][r] = V[4r+c] */
        const float dJ00 = V[0] * dT0[0] + V[4] * dT0[1] + V[8] * dT0[2];
        const float dJ02 = V[2] * dT0[0] + V[6] * dT0[1] + V[10] * dT0[2];
        const float dJ11 = V[1] * dT1[0] + V[5] * dT1[1] + V[9] * dT1[2];
        const float dJ12 = V[2] * dT1[0] + V[6] * dT1[1] + V[10] * dT1[2];
        const float tzi = 1.f / t[2], tz2 = tzi * tzi, tz3 = tz2 * tzi;
        const float dtx = x_grad_mul * -fx * tz2 * dJ02;
        const float dty = y_grad_mul * -fy * tz2 * dJ12;
        const float dtz = -fx * tz2 * dJ00 - fy * tz2 * dJ11 + (2 * fx * t[0]) * tz3 * dJ02 + (2 * fy * t[1]) * tz3 * dJ12;
        dm[0] = V[0] * dtx + V[1] * dty + V[2] * dtz;
        dm[1] = V[4] * dtx + V[5] * dty + V[6] * dtz;
        dm[2] = V[8] * dtx + V[9] * dty + V[10] * dtz;
        /* ---- projection backward ---- */
        const float hw = xform_row(Pm, 3, p[0], p[1], p[2]);
        const float mw = 1.0f / (hw + 0.0000001f);
        const float mul1 = (Pm[0] * p[0] + Pm[4] * p[1] + Pm[8] * p[2] + Pm[12]) * mw * mw;
        const float mul2 = (Pm[1] * p[0] + Pm[5] * p[1] + Pm[9] * p[2] + Pm[13]) * mw * mw;
        const float g2x = dL_dmean2D[3 * i], g2y = dL_dmean2D[3 * i + 1];
        dm[0] += (Pm[0] * mw - Pm[3] * mul1) * g2x + (Pm[1] * mw - Pm[3] * mul2) * g2y;
        dm[1] += (Pm[4] * mw - Pm[7] * mul1) * g2x + (Pm[5] * mw - Pm[7] * mul2) * g2y;
        dm[2] += (Pm[8] * mw - Pm[11] * mul1) * g2x + (Pm[9] * mw - Pm[11] * mul2) * g2y;
        /* ---- depth backward ---- */
        const float mul3 = V[2] * p[0] + V[6] * p[1] + V[10] * p[2] + V[14];
        dm[0] += (V[2] - V[3] * mul3) * dL_ddepth[i];
        dm[1] += (V[6] - V[7] * mul3) * dL_ddepth[i];
        dm[2] += (V[10] - V[11] * mul3) * dL_ddepth[i];
        /* ---- SH backward ---- */
        if (shs) {
            const float* sh = shs + (size_t)i * M * 3;
            float* dsh = dL_dsh + (size_t)i * M * 3;
            float dorig[3] = {p[0] - campos[0], p[1] - campos[1], p[2] - campos[2]};
            const float len = sqrtf(dorig[0] * dorig[0] + dorig[1] * dorig[1] + dorig[2] * dorig[2]);
            const float x = dorig[0] / len, y = dorig[1] / len, z = dorig[2] / len;
            float dRGB[3];
            for (int ch = 0; ch < 3; ch++) dRGB[ch] = dL_dcolor[3 * i + ch] * (clamped[3 * i + ch] ? 0.f : 1.f);
            float ddir[3] = {0, 0, 0};
            for (int ch = 0; ch < 3; ch++) {
#define S(k) sh[(k) * 3 + ch]
#define DS(k) dsh[(k) * 3 + ch]
                float dx_ = 0, dy_ = 0, dz_ = 0;
                DS(0) = SH_C0 * dRGB[ch];
                if (D > 0) {
                    DS(1) = -SH_C1 * y * dRGB[ch];
                    DS(2) = SH_C1 * z * dRGB[ch];
                    DS(3) = -SH_C1 * x * dRGB[ch];
                    dx_ = -SH_C1 * S(3); dy_ = -SH_C1 * S(1); dz_ = SH_C1 * S(2);
                    if (D > 1) {
                        const float xx = x * x, yy = y * y, zz = z * z, xy_ = x * y, yz = y * z, xz = x * z;
                        DS(4) = SH_C2[0] * xy_ * dRGB[ch];
                        DS(5) = SH_C2[1] * yz * dRGB[ch];
                        DS(6) = SH_C2[2] * (2.f * zz - xx - yy) * dRGB[ch];
                        DS(7) = SH_C2[3] * xz * dRGB[ch];
                        DS(8) = SH_C2[4] * (xx - yy) * dRGB[ch];
                        dx_ += SH_C2[0] * y * S(4) + SH_C2[2] * 2.f * -x * S(6) + SH_C2[3] * z * S(7) + SH_C2[4] * 2.f * x * S(8);
                        dy_ += SH_C2[0] * x * S(4) + SH_C2[1] * z * S(5) + SH_C2[2] * 2.f * -y * S(6) + SH_C2[4] * 2.f * -y * S(8);
                        dz_ += SH_C2[1] * y * S(5) + SH_C2[2] * 2.f * 2.f * z * S(6) + SH_C2[3] * x * S(7);
                        if (D > 2) {
                            DS(9) = SH_C3[0] * y * (3.f * xx - yy) * dRGB[ch];
                            DS(10) = SH_C3[1] * xy_ * z * dRGB[ch];
                            DS(11) = SH_C3[2] * y * (4.f * zz - xx - yy) * dRGB[ch];
                            DS(12) = SH_C3[3] * z * (2.f * zz - 3.f * xx - 3.f * yy) * dRGB[ch];
                            DS(13) = SH_C3[4] * x * (4.f * zz - xx - yy) * dRGB[ch];
                            DS(14) = SH_C3[5] * z * (xx - yy) * dRGB[ch];
                            DS(15) = SH_C3[6] * x * (xx - 3.f * yy) * dRGB[ch];
                            dx_ += (SH_C3[0] * S(9) * 3.f * 2.f * xy_ + SH_C3[1] * S(10) * yz + SH_C3[2] * S(11) * -2.f * xy_ +
                                    SH_C3[3] * S(12) * -3.f * 2.f * xz + SH_C3[4] * S(13) * (-3.f * xx + 4.f * zz - yy) +
                                    SH_C3[5] * S(14) * 2.f * xz + SH_C3[6] * S(15) * 3.f * (xx - yy));
                            dy_ += (SH_C3[0] * S(9) * 3.f * (xx - yy) + SH_C3[1] * S(10) * xz +
                                    SH_C3[2] * S(11) * (-3.f * yy + 4.f * zz - xx) + SH_C3[3] * S(12) * -3.f * 2.f * yz +
                                    SH_C3[4] * S(13) * -2.f * xy_ + SH_C3[5] * S(14) * -2.f * yz + SH_C3[6] * S(15) * -3.f * 2.f * xy_);
                            dz_ += (SH_C3[1] * S(10) * xy_ + SH_C3[2] * S(11) * 4.f * 2.f * yz +
                                    SH_C3[3] * S(12) * 3.f * (2.f * zz - xx - yy) + SH_C3[4] * S(13) * 4.f * 2.f * xz +
                                    SH_C3[5] * S(14) * (xx - yy));
                        }
                    }
                }
#undef S
#undef DS
                ddir[0] += dx_ * dRGB[ch]; ddir[1] += dy_ * dRGB[ch]; ddir[2] += dz_ * dRGB[ch];
            }
            float dmean_sh[3];
            dnormvdv3(dorig, ddir, dmean_sh);
            dm[0] += dmean_sh[0]; dm[1] += dmean_sh[1]; dm[2] += dmean_sh[2];
        }
        /* ---- cov3D backward ---- */
        if (scales) {
            const float* q = rots + 4 * i;
            const float r = q[0], x = q[1], y = q[2], z = q[3];
            /* R as a plain row-major rotation matrix Rm[row][col]; glm's R[c][r] above is its transpose */
            const float Rg[3][3] = {/* glm columns */
                {1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y)},
                {2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x)},
                {2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y)}};
            const float s[3] = {mod * scales[3 * i], mod * scales[3 * i + 1], mod * scales[3 * i + 2]};
            float Mg[3][3]; /* M = S*R, glm [c][r] */
            for (int cc = 0; cc < 3; cc++) for (int rr = 0; rr < 3; rr++) Mg[cc][rr] = s[rr] * Rg[cc][rr];
            const float dSg[3][3] = {{dcv[0], 0.5f * dcv[1], 0.5f * dcv[2]},
                                     {0.5f * dcv[1], dcv[3], 0.5f * dcv[4]},
                                     {0.5f * dcv[2], 0.5f * dcv[4], dcv[5]}};
            /* dL_dM = 2 * M * dL_dSigma  (glm: (A*B)[c][r] = sum_k A[k][r]*B[c][k]) */
            float dM[3][3];
            for (int cc = 0; cc < 3; cc++) for (int rr = 0; rr < 3; rr++)
                dM[cc][rr] = 2.0f * (Mg[0][rr] * dSg[cc][0] + Mg[1][rr] * dSg[cc][1] + Mg[2][rr] * dSg[cc][2]);
            /* Rt[c] = row c of R ; dMt[c] = row c of dL_dM:  Rt[c][k] = Rg[k][c], dMt[c][k] = dM[k][c] */
            float dMt[3][3];
            for (int cc = 0; cc < 3; cc++) for (int k = 0; k < 3; k++) dMt[cc][k] = dM[k][cc];
            for (int cc = 0; cc < 3; cc++)
                dL_dscale[3 * i + cc] = Rg[0][cc] * dMt[cc][0] + Rg[1][cc] * dMt[cc][1] + Rg[2][cc] * dMt[cc][2];
            for (int cc = 0; cc < 3; cc++) for (int k = 0; k < 3; k++) dMt[cc][k] *= s[cc];
            float* dq = dL_drot + 4 * i;
            dq[0] = 2 * z * (dMt[0][1] - dMt[1][0]) + 2 * y * (dMt[2][0] - dMt[0][2]) + 2 * x * (dMt[1][2] - dMt[2][1]);
            dq[1] = 2 * y * (dMt[1][0] + dMt[0][1]) + 2 * z * (dMt[2][0] + dMt[0][2]) + 2 * r * (dMt[1][2] - dMt[2][1]) - 4 * x * (dMt[2][2] + dMt[1][1]);
            dq[2] = 2 * x * (dMt[1][0] + dMt[0][1]) + 2 * r * (dMt[2][0] - dMt[0][2]) + 2 * z * (dMt[1][2] + dMt[2][1]) - 4 * y * (dMt[2][2] + dMt[0][0]);
            dq[3] = 2 * r * (dMt[0][1] - dMt[1][0]) + 2 * x * (dMt[2][0] + dMt[0][2]) + 2 * y * (dMt[1][2] + dMt[2][1]) - 4 * z * (dMt[1][1] + dMt[0][0]);
        }
    }
}

/* rasterizer_impl.cu:54-66 */
void gsro_mark_visible(int P, const float* means, const float* V, uint8_t* present)
{
    for (int i = 0; i < P; i++) {
        const float* p = means + 3 * i;
        present[i] = !(xform_row(V, 2, p[0], p[1], p[2]) <= 0.2f);
    }
}

int gsro_sort_bits(int W, int H)
{
    const int gx = (W + TILE - 1) / TILE, gy = (H + TILE - 1) / TILE;
    return 32 + (int)higher_msb((uint32_t)(gx * gy));
}

/* simple-knn distCUDA2 (submodules/simple-knn/simple_knn.cu:131-183, spatial.cu:15-26): mean of the three smallest
 * squared distances from every point to the others.  The reference's Morton boxes only prune its search, so the
 * restatement is the brute-force definition with the reference build's rounding sequence:
 *   d = other - self;  d2 = fma(dz, dz, fma(dx, dx, rn(dy*dy)));  mean = rn(rn(rn(b0 + b1) + b2) / 3)
 * (b0 <= b1 <= b2 start at FLT_MAX, simple_knn.cu:151; a point is excluded by index, so coincident points count). */
void gsro_knn_mean_dist2(int P, const float* pts, float* out)
{
#pragma omp parallel for schedule(static)
    for (int i = 0; i < P; i++) {
        const float sx = pts[3 * i], sy = pts[3 * i + 1], sz = pts[3 * i + 2];
        float best[3] = {3.402823466e+38f, 3.402823466e+38f, 3.402823466e+38f};
        for (int j = 0; j < P; j++) {
            if (j == i) continue;
            const float dx = pts[3 * j] - sx, dy = pts[3 * j + 1] - sy, dz = pts[3 * j + 2] - sz;
            float dist = fmaf(dz, dz, fmaf(dx, dx, dy * dy));
            for (int k = 0; k < 3; k++) {   /* updateKBest<3>, simple_knn.cu:131-145 */
                if (best[k] > dist) {
                    const float t = best[k];
                    best[k] = dist;
                    dist = t;
                }
            }
        }
        out[i] = ((best[0] + best[1]) + best[2]) / 3.0f;
    }
}
