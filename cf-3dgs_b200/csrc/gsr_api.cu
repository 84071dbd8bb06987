// extern "C" entry points declared in include/gsr.h: argument checks, workspace carving, stage ordering.
// Orchestration counterpart of the reference's CudaRasterizer::Rasterizer (rasterizer_impl.cu:197-447).
#include <cstdarg>
#include <cstdio>
#include <cstring>

#include "../../include/gsr.h"
#include "gsr_kernels.h"
#include "gsr_launch.cuh"

using namespace gsr;

namespace {

thread_local char g_err[512] = "";

int fail(int code, const char* fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

// After each stage: launch-error check always; full synchronise + check when debug (reference CHECK_CUDA).
#define GSR_STAGE(name)                                                                           \
    do {                                                                                          \
        cudaError_t e_ = cudaGetLastError();                                                      \
        if (e_ == cudaSuccess && debug) e_ = cudaStreamSynchronize(s);                            \
        if (e_ != cudaSuccess) return fail(GSR_ERR_CUDA, "%s: %s", name, cudaGetErrorString(e_)); \
    } while (0)

// ---- stage timing ------------------------------------------------------------------------------
enum Stage { ST_PREPROCESS, ST_DEPTH_SORT, ST_OFFSETS, ST_EMIT, ST_TILE_SORT, ST_RENDER_FWD,
             ST_GRAD_ZERO, ST_RENDER_BWD, ST_PREPROCESS_BWD, ST_COUNT };
const char* const kStageNames[ST_COUNT] = {"preprocess", "depth_sort", "offsets_scan", "emit", "tile_sort",
                                           "render_forward", "grad_zero", "render_backward",
                                           "preprocess_backward"};
bool g_prof_on = false;
cudaEvent_t g_ev[ST_COUNT][2];
bool g_ev_made = false;
bool g_ev_used[ST_COUNT];

struct StageTimer {
    int st; cudaStream_t s;
    // (an event record is a stream operation of its own: while profiling, stages do not overlap programmatically)
    StageTimer(int st_, cudaStream_t s_) : st(st_), s(s_) { if (g_prof_on) { cudaEventRecord(g_ev[st][0], s); chain_break(); } }
    ~StageTimer() { if (g_prof_on) { cudaEventRecord(g_ev[st][1], s); g_ev_used[st] = true; chain_break(); } }
};

// The event that marks "num_rendered has landed on the host": one per (host thread, device), created on first use and
// kept (two threads driving the same device must not record and wait on one shared event).
cudaError_t count_event(cudaEvent_t* out)
{
    constexpr int MAX_DEV = 64;
    thread_local cudaEvent_t ev[MAX_DEV] = {};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= MAX_DEV) return cudaErrorInvalidDevice;
    if (!ev[dev]) {
        e = cudaEventCreateWithFlags(&ev[dev], cudaEventDisableTiming);
        if (e != cudaSuccess) { ev[dev] = nullptr; return e; }
    }
    *out = ev[dev];
    return cudaSuccess;
}

int tile_bits(int tiles)
{
    int bits = 0;
    while ((1 << bits) < tiles) bits++;
    return bits < 1 ? 1 : bits;
}

}  // namespace

namespace gsr {
bool& pdl_chain_flag()
{
    thread_local bool chained = false;
    return chained;
}
// lets other translation units of the library report through gsr_last_error()
void set_last_error(const char* msg) { snprintf(g_err, sizeof(g_err), "%s", msg ? msg : ""); }
}  // namespace gsr

extern "C" {

int gsr_version(void) { return 1000; }

const char* gsr_last_error(void) { return g_err; }

size_t gsr_geometry_bytes(int P)
{
    GeomView g;
    return geom_layout(nullptr, P > 0 ? P : 1, g);
}

size_t gsr_image_bytes(int W, int H)
{
    ImageView v;
    return image_layout(nullptr, W, H, v);
}

size_t gsr_binning_bytes(int64_t num_rendered, int W, int H)
{
    (void)W; (void)H;
    BinView b;
    return bin_layout(nullptr, num_rendered, b);
}

size_t gsr_pose_scratch_bytes(int P) { return (size_t)pose_blocks(P > 0 ? P : 1) * 6 * sizeof(float); }

// ---- shared pieces of the two forward flavours --------------------------------------------------------
static int check_forward_args(int P, int sh_degree, int M, const float* means3D, const float* scales,
                              const float* rotations, const float* opacities, const float* shs,
                              const float* cov3D_precomp, const float* colors_precomp, const float* viewmatrix,
                              const float* projmatrix, const float* campos, int W, int H, const int* radii,
                              const void* geometry, size_t geometry_bytes)
{
    if (P < 0 || W <= 0 || H <= 0) return fail(GSR_ERR_INVALID, "bad sizes P=%d W=%d H=%d", P, W, H);
    if (P == 0) return GSR_OK;
    if (!means3D || !opacities || !viewmatrix || !projmatrix || !radii || !geometry)
        return fail(GSR_ERR_INVALID, "means3D, opacities, matrices, radii and geometry workspace are required");
    if ((shs == nullptr) == (colors_precomp == nullptr))
        return fail(GSR_ERR_INVALID, "provide exactly one of SHs or precomputed colors");
    if (((scales == nullptr) || (rotations == nullptr)) == (cov3D_precomp == nullptr))
        return fail(GSR_ERR_INVALID, "provide exactly one of scale/rotation pair or precomputed 3D covariance");
    if (shs && (sh_degree < 0 || sh_degree > 3 || (sh_degree + 1) * (sh_degree + 1) > M))
        return fail(GSR_ERR_INVALID, "sh_degree %d needs %d coefficients, M=%d", sh_degree,
                    (sh_degree + 1) * (sh_degree + 1), M);
    if (shs && !campos) return fail(GSR_ERR_INVALID, "campos is required with SHs");
    if (div_up(W, TILE) > 65535 || div_up(H, TILE) > 65535) return fail(GSR_ERR_INVALID, "image too large");
    if (gsr_geometry_bytes(P) > geometry_bytes)
        return fail(GSR_ERR_WORKSPACE, "geometry workspace too small: %zu < %zu", geometry_bytes, gsr_geometry_bytes(P));
    return GSR_OK;
}

// depth-digit passes of the key sort + instance offsets (stages 2-3)
static int depth_sort_and_offsets(int P, const GeomView& g, cudaStream_t s)
{
    {
        StageTimer t(ST_DEPTH_SORT, s);
        uint32_t* keys[2] = {g.sort_keys[0], g.sort_keys[1]};
        uint32_t* vals[2] = {g.sort_vals[0], g.sort_vals[1]};
        // first pass straight from the depth keys preprocess wrote, values = Gaussian index
        const int cur = radix_sort_pairs(keys, vals, P, 0, 32, g.hist, g.digit_total, s, nullptr, g.depth_key);
        if (cur != 0) {  // keep the invariant "sort_vals[0] holds the depth order"
            cudaMemcpyAsync(g.sort_vals[0], g.sort_vals[1], sizeof(uint32_t) * (size_t)P, cudaMemcpyDeviceToDevice, s);
            chain_break();
        }
    }
    { StageTimer t(ST_OFFSETS, s); launch_offsets_scan(g.tiles, g.sort_vals[0], g.offsets, g.block_sums, P, s); }
    return GSR_OK;
}

// stages 4-7.  n_layout sizes the binning workspace; with n_dev != nullptr the real count is read on the device.
static int bin_and_render(int P, int W, int H, int64_t n_layout, const unsigned long long* n_dev, const float* background,
                          void* geometry, void* binning, void* image, float* out_color, float* out_depth,
                          float* out_alpha, cudaStream_t s, int debug, bool ranges_zeroed)
{
    GeomView g;
    geom_layout((char*)geometry, P, g);
    ImageView iv;
    image_layout((char*)image, W, H, iv);
    BinView b;
    bin_layout((char*)binning, n_layout, b);
    const int gx = div_up(W, TILE), gy = div_up(H, TILE), tiles = gx * gy;
    // memsets are ordinary stream operations: one between two kernels would undo their programmatic overlap, so the
    // tile ranges are cleared ahead of the kernel chain
    if (!ranges_zeroed) cudaMemsetAsync(iv.ranges, 0, sizeof(uint2) * (size_t)tiles, s);
    if (n_layout > 0) {
        { StageTimer t(ST_EMIT, s); launch_emit(g, g.sort_vals[0], P, gx, b.keys[0], b.vals[0], n_layout, s); }
        GSR_STAGE("emit");
        {
            StageTimer t(ST_TILE_SORT, s);
            const int cur = radix_sort_pairs(b.keys, b.vals, n_layout, 0, tile_bits(tiles), b.hist, b.digit_total, s, n_dev,
                                             nullptr, iv.ranges);
            if (cur != 0) {  // invariant: keys[0]/vals[0] hold the sorted list
                cudaMemcpyAsync(b.keys[0], b.keys[1], sizeof(uint32_t) * (size_t)n_layout, cudaMemcpyDeviceToDevice, s);
                chain_break();
                cudaMemcpyAsync(b.vals[0], b.vals[1], sizeof(uint32_t) * (size_t)n_layout, cudaMemcpyDeviceToDevice, s);
                chain_break();
            }
        }
        GSR_STAGE("tile sort");
    }
    // (tile ranges: written by the last pass of the tile sort)
    {
        StageTimer t(ST_RENDER_FWD, s);
        launch_render_forward(W, H, iv.ranges, b.vals[0], g, background, out_color, out_depth, out_alpha, iv.n_contrib, s);
    }
    GSR_STAGE("render");
    return GSR_OK;
}

static void zero_images(int W, int H, float* out_color, float* out_depth, float* out_alpha, cudaStream_t s)
{
    cudaMemsetAsync(out_color, 0, sizeof(float) * 3 * (size_t)W * H, s);
    chain_break();
    cudaMemsetAsync(out_depth, 0, sizeof(float) * (size_t)W * H, s);
    chain_break();
    cudaMemsetAsync(out_alpha, 0, sizeof(float) * (size_t)W * H, s);
    chain_break();
}

int gsr_forward_prepare(int P, int sh_degree, int M, const float* means3D, const float* scales,
                        float scale_modifier, const float* rotations, const float* opacities,
                        const float* shs, const float* cov3D_precomp, const float* colors_precomp,
                        const float* viewmatrix, const float* projmatrix, const float* campos, int W, int H,
                        float tanfovx, float tanfovy, int prefiltered, int* radii, void* geometry,
                        size_t geometry_bytes, int64_t* num_rendered_host, void* stream, int debug)
{
    return gsr_forward_prepare_rigid(P, sh_degree, M, means3D, scales, scale_modifier, rotations, opacities, shs,
                                     cov3D_precomp, colors_precomp, viewmatrix, projmatrix, campos, W, H, tanfovx, tanfovy,
                                     prefiltered, radii, geometry, geometry_bytes, num_rendered_host, nullptr, stream, debug);
}

int gsr_forward_prepare_rigid(int P, int sh_degree, int M, const float* means3D, const float* scales,
                              float scale_modifier, const float* rotations, const float* opacities,
                              const float* shs, const float* cov3D_precomp, const float* colors_precomp,
                              const float* viewmatrix, const float* projmatrix, const float* campos, int W, int H,
                              float tanfovx, float tanfovy, int prefiltered, int* radii, void* geometry,
                              size_t geometry_bytes, int64_t* num_rendered_host, const float* rigid, void* stream,
                              int debug)
{
    chain_break();   // whatever ran before this call on the stream is not ours to overlap with
    cudaStream_t s = (cudaStream_t)stream;
    if (!num_rendered_host) return fail(GSR_ERR_INVALID, "num_rendered_host is NULL");
    *num_rendered_host = 0;
    const int rc = check_forward_args(P, sh_degree, M, means3D, scales, rotations, opacities, shs, cov3D_precomp,
                                      colors_precomp, viewmatrix, projmatrix, campos, W, H, radii, geometry, geometry_bytes);
    if (rc != GSR_OK) return rc;
    if (P == 0) return GSR_OK;  // rasterize_points.cu:83 -- nothing launched
    GeomView g;
    geom_layout((char*)geometry, P, g);

    PreprocessArgs a;
    a.P = P; a.D = sh_degree; a.M = M;
    a.means3D = means3D; a.scales = scales; a.scale_modifier = scale_modifier; a.rotations = rotations;
    a.opacities = opacities; a.shs = shs; a.cov3D_precomp = cov3D_precomp; a.colors_precomp = colors_precomp;
    a.viewmatrix = viewmatrix; a.projmatrix = projmatrix; a.campos = campos ? campos : viewmatrix;
    a.W = W; a.H = H; a.tanfovx = tanfovx; a.tanfovy = tanfovy; a.prefiltered = prefiltered; a.radii = radii;
    a.rigid = rigid;
    { StageTimer t(ST_PREPROCESS, s); launch_preprocess(a, g, s); }
    GSR_STAGE("preprocess");

    // "how many instances?" travels to the host (8 bytes, into the caller's -- ideally pinned -- word) while
    // the GPU keeps sorting; bit 62 of the counter flags a prefilter violation
    cudaEvent_t ev;
    cudaError_t e = count_event(&ev);
    if (e != cudaSuccess) return fail(GSR_ERR_CUDA, "event: %s", cudaGetErrorString(e));
    cudaMemcpyAsync(num_rendered_host, g.counters, sizeof(int64_t), cudaMemcpyDeviceToHost, s);
    chain_break();
    cudaEventRecord(ev, s);
    chain_break();

    depth_sort_and_offsets(P, g, s);   // low 32 key bits (depth) sorted over Gaussians, before duplication

    e = cudaEventSynchronize(ev);
    if (e != cudaSuccess) return fail(GSR_ERR_CUDA, "preprocess/count: %s", cudaGetErrorString(e));
    GSR_STAGE("depth sort + offsets");
    const unsigned long long raw = (unsigned long long)*num_rendered_host;
    *num_rendered_host = 0;
    if (raw >> 62) return fail(GSR_ERR_PREFILTER, "Point is filtered although prefiltered is set. This shouldn't happen!");
    if (raw > 0x7fffffffull) return fail(GSR_ERR_OVERFLOW, "%llu tile instances", raw);
    *num_rendered_host = (int64_t)raw;
    return GSR_OK;
}

int gsr_forward_render(int P, int W, int H, int64_t num_rendered, const float* background,
                       const float* colors_precomp, void* geometry, void* binning, size_t binning_bytes,
                       void* image, float* out_color, float* out_depth, float* out_alpha, void* stream, int debug)
{
    chain_break();   // whatever ran before this call on the stream is not ours to overlap with
    cudaStream_t s = (cudaStream_t)stream;
    (void)colors_precomp;
    if (W <= 0 || H <= 0) return fail(GSR_ERR_INVALID, "bad image size");
    if (!out_color || !out_depth || !out_alpha) return fail(GSR_ERR_INVALID, "outputs are required");
    if (P == 0) {  // reference returns zero images without launching (rasterize_points.cu:68-71,83)
        zero_images(W, H, out_color, out_depth, out_alpha, s);
        return GSR_OK;
    }
    if (!geometry || !binning || !image || !background)
        return fail(GSR_ERR_INVALID, "workspaces and background are required");
    if (gsr_binning_bytes(num_rendered, W, H) > binning_bytes)
        return fail(GSR_ERR_WORKSPACE, "binning workspace too small: %zu < %zu", binning_bytes,
                    gsr_binning_bytes(num_rendered, W, H));
    return bin_and_render(P, W, H, num_rendered, nullptr, background, geometry, binning, image, out_color, out_depth,
                          out_alpha, s, debug, false);
}

int gsr_forward_fused(int P, int sh_degree, int M, const float* means3D, const float* scales, float scale_modifier,
                      const float* rotations, const float* opacities, const float* shs, const float* cov3D_precomp,
                      const float* colors_precomp, const float* viewmatrix, const float* projmatrix,
                      const float* campos, int W, int H, float tanfovx, float tanfovy, const float* background,
                      int* radii, void* geometry, size_t geometry_bytes, void* binning, size_t binning_bytes,
                      int64_t capacity, void* image, float* out_color, float* out_depth, float* out_alpha,
                      int64_t* status, void* stream, int debug)
{
    return gsr_forward_fused_rigid(P, sh_degree, M, means3D, scales, scale_modifier, rotations, opacities, shs,
                                   cov3D_precomp, colors_precomp, viewmatrix, projmatrix, campos, W, H, tanfovx, tanfovy,
                                   background, radii, geometry, geometry_bytes, binning, binning_bytes, capacity, image,
                                   out_color, out_depth, out_alpha, status, nullptr, stream, debug);
}

int gsr_forward_fused_rigid(int P, int sh_degree, int M, const float* means3D, const float* scales, float scale_modifier,
                            const float* rotations, const float* opacities, const float* shs,
                            const float* cov3D_precomp, const float* colors_precomp, const float* viewmatrix,
                            const float* projmatrix, const float* campos, int W, int H, float tanfovx, float tanfovy,
                            const float* background, int* radii, void* geometry, size_t geometry_bytes, void* binning,
                            size_t binning_bytes, int64_t capacity, void* image, float* out_color, float* out_depth,
                            float* out_alpha, int64_t* status, const float* rigid, void* stream, int debug)
{
    chain_break();   // whatever ran before this call on the stream is not ours to overlap with
    cudaStream_t s = (cudaStream_t)stream;
    if (!out_color || !out_depth || !out_alpha || !status) return fail(GSR_ERR_INVALID, "outputs and status are required");
    if (capacity < 1 || capacity > 0x7fffffffll) return fail(GSR_ERR_INVALID, "capacity out of range");
    const int rc = check_forward_args(P, sh_degree, M, means3D, scales, rotations, opacities, shs, cov3D_precomp,
                                      colors_precomp, viewmatrix, projmatrix, campos, W, H, radii, geometry, geometry_bytes);
    if (rc != GSR_OK) return rc;
    if (P == 0) {
        zero_images(W, H, out_color, out_depth, out_alpha, s);   // status: nothing to add (count 0, no flag)
        return GSR_OK;
    }
    if (!binning || !image || !background) return fail(GSR_ERR_INVALID, "workspaces and background are required");
    if (gsr_binning_bytes(capacity, W, H) > binning_bytes)
        return fail(GSR_ERR_WORKSPACE, "binning workspace too small for capacity %lld: %zu < %zu", (long long)capacity,
                    binning_bytes, gsr_binning_bytes(capacity, W, H));
    GeomView g;
    geom_layout((char*)geometry, P, g);
    PreprocessArgs a;
    a.P = P; a.D = sh_degree; a.M = M;
    a.means3D = means3D; a.scales = scales; a.scale_modifier = scale_modifier; a.rotations = rotations;
    a.opacities = opacities; a.shs = shs; a.cov3D_precomp = cov3D_precomp; a.colors_precomp = colors_precomp;
    a.viewmatrix = viewmatrix; a.projmatrix = projmatrix; a.campos = campos ? campos : viewmatrix;
    a.W = W; a.H = H; a.tanfovx = tanfovx; a.tanfovy = tanfovy; a.prefiltered = 0; a.radii = radii;
    a.rigid = rigid;
    {
        ImageView iv;
        image_layout((char*)image, W, H, iv);
        cudaMemsetAsync(iv.ranges, 0, sizeof(uint2) * (size_t)div_up(W, TILE) * div_up(H, TILE), s);
        chain_break();
    }
    { StageTimer t(ST_PREPROCESS, s); launch_preprocess(a, g, s); }
    GSR_STAGE("preprocess");
    launch_publish_status(g.counters, capacity, (long long*)status, s);
    depth_sort_and_offsets(P, g, s);
    GSR_STAGE("depth sort + offsets");
    return bin_and_render(P, W, H, capacity, g.counters, background, geometry, binning, image, out_color, out_depth,
                          out_alpha, s, debug, true);
}

int gsr_backward(int P, int sh_degree, int M, int W, int H, int64_t num_rendered, const float* background,
                 const float* means3D, const int* radii, const float* colors_precomp, const float* scales,
                 const float* rotations, float scale_modifier, const float* cov3D_precomp,
                 const float* viewmatrix, const float* projmatrix, const float* campos, float tanfovx,
                 float tanfovy, const float* shs, const float* out_alpha, const float* dL_dout_color,
                 const float* dL_dout_depth, const float* dL_dout_alpha, const void* geometry,
                 const void* binning, const void* image, void* grad_scratch, float* dL_dmeans2D,
                 float* dL_dcolors, float* dL_dopacity, float* dL_dmeans3D, float* dL_dcov3D, float* dL_dsh,
                 float* dL_dscales, float* dL_drotations, float* dL_dpose, void* pose_scratch, void* stream,
                 int debug)
{
    return gsr_backward_rigid(P, sh_degree, M, W, H, num_rendered, background, means3D, radii, colors_precomp, scales,
                              rotations, scale_modifier, cov3D_precomp, viewmatrix, projmatrix, campos, tanfovx, tanfovy,
                              shs, out_alpha, dL_dout_color, dL_dout_depth, dL_dout_alpha, geometry, binning, image,
                              grad_scratch, dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, M,
                              dL_dscales, dL_drotations, dL_dpose, pose_scratch, nullptr, stream, debug);
}

int gsr_backward_rigid(int P, int sh_degree, int M, int W, int H, int64_t num_rendered, const float* background,
                       const float* means3D, const int* radii, const float* colors_precomp, const float* scales,
                       const float* rotations, float scale_modifier, const float* cov3D_precomp,
                       const float* viewmatrix, const float* projmatrix, const float* campos, float tanfovx,
                       float tanfovy, const float* shs, const float* out_alpha, const float* dL_dout_color,
                       const float* dL_dout_depth, const float* dL_dout_alpha, const void* geometry,
                       const void* binning, const void* image, void* grad_scratch, float* dL_dmeans2D,
                       float* dL_dcolors, float* dL_dopacity, float* dL_dmeans3D, float* dL_dcov3D, float* dL_dsh,
                       int dsh_M, float* dL_dscales, float* dL_drotations, float* dL_dpose, void* pose_scratch,
                       const float* rigid, void* stream, int debug)
{
    chain_break();   // whatever ran before this call on the stream is not ours to overlap with
    cudaStream_t s = (cudaStream_t)stream;
    if (P == 0) {
        if (dL_dpose) cudaMemsetAsync(dL_dpose, 0, 6 * sizeof(float), s);
        return GSR_OK;
    }
    if (!means3D || !radii || !viewmatrix || !projmatrix || !geometry || !binning || !image || !grad_scratch ||
        !out_alpha || !dL_dout_color || !background)
        return fail(GSR_ERR_INVALID, "required pointer is NULL");
    // every gradient output is optional: NULL = not wanted (a pose-only fit passes dL_dpose alone and saves the 236 bytes
    // per Gaussian the others cost); dL_dscales and dL_drotations come as a pair
    if (scales && !rotations) return fail(GSR_ERR_INVALID, "rotations are required with scales");
    if ((dL_dscales == nullptr) != (dL_drotations == nullptr))
        return fail(GSR_ERR_INVALID, "dL_dscales and dL_drotations are written together: pass both or neither");
    if (dL_dpose && !pose_scratch) return fail(GSR_ERR_INVALID, "pose_scratch is required with dL_dpose");
    if (shs && dL_dsh && (dsh_M < (sh_degree + 1) * (sh_degree + 1) || dsh_M > M))
        return fail(GSR_ERR_INVALID, "dsh_M=%d must hold the %d active coefficients and not exceed M=%d", dsh_M,
                    (sh_degree + 1) * (sh_degree + 1), M);
    GeomView g;
    geom_layout((char*)geometry, P, g);
    ImageView iv;
    image_layout((char*)image, W, H, iv);
    BinView b;
    bin_layout((char*)binning, num_rendered, b);

    { StageTimer t(ST_GRAD_ZERO, s); cudaMemsetAsync(grad_scratch, 0, (size_t)P * 12 * sizeof(float), s); }
    if (num_rendered > 0) {
        StageTimer t(ST_RENDER_BWD, s);
        // opacity / colour partials are needed when their gradients are wanted, or when the colour depends on the
        // position (SH above degree 0), because the mean gradient then has a term through dL/dRGB
        const bool want_params = dL_dopacity || dL_dcolors || dL_dsh || (shs && sh_degree > 0);
        launch_render_backward(W, H, iv.ranges, b.vals[0], g, background, out_alpha, iv.n_contrib, dL_dout_color,
                               dL_dout_depth, dL_dout_alpha, (float*)grad_scratch, s, want_params);
    }
    GSR_STAGE("render backward");
    PreprocessBwdArgs a;
    a.P = P; a.D = sh_degree; a.M = M;
    a.means3D = means3D; a.radii = radii; a.shs = shs; a.scales = scales; a.rotations = rotations;
    a.scale_modifier = scale_modifier; a.cov3D_precomp = cov3D_precomp; a.colors_precomp = colors_precomp;
    a.viewmatrix = viewmatrix; a.projmatrix = projmatrix; a.campos = campos ? campos : viewmatrix;
    a.W = W; a.H = H; a.tanfovx = tanfovx; a.tanfovy = tanfovy;
    a.grad_scratch = (const float*)grad_scratch;
    a.dL_dmeans2D = dL_dmeans2D; a.dL_dcolors = colors_precomp ? dL_dcolors : nullptr; a.dL_dopacity = dL_dopacity;
    a.dL_dmeans3D = dL_dmeans3D; a.dL_dcov3D = dL_dcov3D; a.dL_dsh = shs ? dL_dsh : nullptr; a.dsh_M = dsh_M;
    a.dL_dscales = scales ? dL_dscales : nullptr; a.dL_drotations = scales ? dL_drotations : nullptr;
    a.dL_dcov3D = cov3D_precomp ? dL_dcov3D : nullptr;
    a.dL_dpose = dL_dpose; a.pose_scratch = (float*)pose_scratch;
    a.rigid = rigid;
    { StageTimer t(ST_PREPROCESS_BWD, s); launch_preprocess_backward(a, g, s); }
    GSR_STAGE("preprocess backward");
    return GSR_OK;
}

int gsr_mark_visible(int P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                     uint8_t* present, void* stream)
{
    chain_break();   // whatever ran before this call on the stream is not ours to overlap with
    (void)projmatrix;
    cudaStream_t s = (cudaStream_t)stream;
    const int debug = 0;
    if (P == 0) return GSR_OK;
    if (!means3D || !viewmatrix || !present) return fail(GSR_ERR_INVALID, "required pointer is NULL");
    launch_mark_visible(P, means3D, viewmatrix, present, s);
    GSR_STAGE("mark_visible");
    return GSR_OK;
}

// ---- parity accessor ----------------------------------------------------------------------------
namespace {
__global__ void decode_ranges_kernel(const uint2* __restrict__ enc, int tiles, uint2* __restrict__ out)
{
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t < tiles) out[t] = load_range(enc, t);
}

__global__ void unpack_geom_kernel(int P, GeomView g, float* depths, float* means2D, float* conic_opacity, float* rgb,
                                   uint32_t* tiles_touched, uint8_t* clamped)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const bool vis = g.tiles[i] != 0;
    if (depths) depths[i] = vis ? __uint_as_float(g.depth_key[i]) : 0.f;
    if (means2D) {
        means2D[2 * i] = vis ? g.xy[i].x : 0.f;
        means2D[2 * i + 1] = vis ? g.xy[i].y : 0.f;
    }
    if (conic_opacity) {
        const float4 c = vis ? g.conic_op[i] : make_float4(0.f, 0.f, 0.f, 0.f);
        conic_opacity[4 * i] = c.x; conic_opacity[4 * i + 1] = c.y; conic_opacity[4 * i + 2] = c.z; conic_opacity[4 * i + 3] = c.w;
    }
    if (rgb) {
        const float4 c = vis ? g.rgb_depth[i] : make_float4(0.f, 0.f, 0.f, 0.f);
        rgb[3 * i] = c.x; rgb[3 * i + 1] = c.y; rgb[3 * i + 2] = c.z;
    }
    if (tiles_touched) tiles_touched[i] = g.tiles[i];
    if (clamped) {
        const uint8_t c = vis ? g.clamped[i] : 0;
        clamped[3 * i] = c & 1; clamped[3 * i + 1] = (c >> 1) & 1; clamped[3 * i + 2] = (c >> 2) & 1;
    }
}
}  // namespace

int gsr_debug_dump(int P, int W, int H, int64_t num_rendered, const void* geometry, const void* binning,
                   const void* image, uint64_t* keys_sorted, uint32_t* point_list, uint32_t* ranges,
                   uint32_t* n_contrib, float* depths, float* means2D, float* conic_opacity, float* rgb,
                   uint32_t* tiles_touched, uint8_t* clamped, void* stream)
{
    chain_break();   // whatever ran before this call on the stream is not ours to overlap with
    cudaStream_t s = (cudaStream_t)stream;
    const int debug = 1;
    if (P == 0) return GSR_OK;
    GeomView g;
    geom_layout((char*)geometry, P, g);
    ImageView iv;
    image_layout((char*)image, W, H, iv);
    BinView b;
    bin_layout((char*)binning, num_rendered, b);
    const int tiles = div_up(W, TILE) * div_up(H, TILE);
    if (keys_sorted && num_rendered > 0) launch_build_keys(b.keys[0], b.vals[0], g.depth_key, num_rendered, g.counters, keys_sorted, s);
    if (point_list && num_rendered > 0)
        cudaMemcpyAsync(point_list, b.vals[0], sizeof(uint32_t) * (size_t)num_rendered, cudaMemcpyDeviceToDevice, s);
        chain_break();
    if (ranges) decode_ranges_kernel<<<div_up(tiles, 256), 256, 0, s>>>(iv.ranges, tiles, reinterpret_cast<uint2*>(ranges));
    if (n_contrib) cudaMemcpyAsync(n_contrib, iv.n_contrib, sizeof(uint32_t) * (size_t)W * H, cudaMemcpyDeviceToDevice, s);
    unpack_geom_kernel<<<div_up(P, 256), 256, 0, s>>>(P, g, depths, means2D, conic_opacity, rgb, tiles_touched, clamped);
    GSR_STAGE("debug dump");
    return GSR_OK;
}

int gsr_profile_enable(int on)
{
    if (on && !g_ev_made) {
        for (int i = 0; i < ST_COUNT; i++)
            for (int k = 0; k < 2; k++) {
                cudaError_t e = cudaEventCreate(&g_ev[i][k]);
                if (e != cudaSuccess) return fail(GSR_ERR_CUDA, "event: %s", cudaGetErrorString(e));
            }
        g_ev_made = true;
    }
    for (int i = 0; i < ST_COUNT; i++) g_ev_used[i] = false;
    g_prof_on = on != 0;
    return GSR_OK;
}

int gsr_profile_count(void) { return ST_COUNT; }

const char* gsr_profile_name(int i) { return (i >= 0 && i < ST_COUNT) ? kStageNames[i] : ""; }

int gsr_profile_read(float* ms_out, int capacity)
{
    if (!ms_out) return fail(GSR_ERR_INVALID, "ms_out is NULL");
    for (int i = 0; i < ST_COUNT && i < capacity; i++) {
        ms_out[i] = -1.f;
        if (!g_ev_made || !g_ev_used[i]) continue;
        cudaError_t e = cudaEventSynchronize(g_ev[i][1]);
        if (e == cudaSuccess) e = cudaEventElapsedTime(&ms_out[i], g_ev[i][0], g_ev[i][1]);
        if (e != cudaSuccess) return fail(GSR_ERR_CUDA, "profile read: %s", cudaGetErrorString(e));
    }
    return GSR_OK;
}

int gsr_debug_pair_counts(int P, int W, int H, int64_t num_rendered, const void* geometry, const void* binning,
                          const void* image, uint64_t* counts, void* stream)
{
    chain_break();   // whatever ran before this call on the stream is not ours to overlap with
    cudaStream_t s = (cudaStream_t)stream;
    const int debug = 1;
    if (!counts) return fail(GSR_ERR_INVALID, "counts is NULL");
    if (P == 0) { cudaMemsetAsync(counts, 0, 4 * sizeof(uint64_t), s); return GSR_OK; }
    GeomView g;
    geom_layout((char*)geometry, P, g);
    ImageView iv;
    image_layout((char*)image, W, H, iv);
    BinView b;
    bin_layout((char*)binning, num_rendered, b);
    launch_pair_counts(W, H, iv.ranges, b.vals[0], g, iv.n_contrib, (unsigned long long*)counts, s);
    GSR_STAGE("pair counts");
    return GSR_OK;
}

size_t gsr_knn_scratch_bytes(int P) { return knn_scratch_bytes(P); }

int gsr_knn_mean_dist2(int P, const float* points, float* mean_dist2, void* scratch, size_t scratch_bytes, void* stream)
{
    chain_break();   // whatever ran before this call on the stream is not ours to overlap with
    cudaStream_t s = (cudaStream_t)stream;
    const int debug = 0;
    if (P < 0) return fail(GSR_ERR_INVALID, "P < 0");
    if (P == 0) return GSR_OK;
    if (!points || !mean_dist2 || !scratch) return fail(GSR_ERR_INVALID, "points, mean_dist2 and scratch are required");
    if (scratch_bytes < knn_scratch_bytes(P))
        return fail(GSR_ERR_WORKSPACE, "knn scratch too small: %zu < %zu", scratch_bytes, knn_scratch_bytes(P));
    launch_knn_mean_dist2(P, points, mean_dist2, scratch, s);
    GSR_STAGE("knn");
    return GSR_OK;
}

int gsr_debug_expf_check(uint32_t first_bits, uint32_t count, uint64_t* mismatches, void* stream)
{
    chain_break();   // whatever ran before this call on the stream is not ours to overlap with
    cudaStream_t s = (cudaStream_t)stream;
    const int debug = 1;
    if (!mismatches) return fail(GSR_ERR_INVALID, "mismatches is NULL");
    launch_expf_check(first_bits, count, (unsigned long long*)mismatches, s);
    GSR_STAGE("expf check");
    return GSR_OK;
}

}  // extern "C"
