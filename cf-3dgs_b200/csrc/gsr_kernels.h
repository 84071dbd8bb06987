// Internal launch interface between gsr_api.cu and the kernel translation units.
#pragma once
#include "gsr_common.cuh"

namespace gsr {

struct PreprocessArgs {
    const float* rigid = nullptr;   // optional 3x4 row-major [R | t]: means3D are moved by p' = R p + t first
    int P, D, M;
    const float* means3D;
    const float* scales;
    float scale_modifier;
    const float* rotations;
    const float* opacities;
    const float* shs;
    const float* cov3D_precomp;
    const float* colors_precomp;
    const float* viewmatrix;
    const float* projmatrix;
    const float* campos;
    int W, H;
    float tanfovx, tanfovy;
    int prefiltered;
    int* radii;
};

void launch_preprocess(const PreprocessArgs& a, const GeomView& g, cudaStream_t s);
void launch_mark_visible(int P, const float* means3D, const float* view, uint8_t* present, cudaStream_t s);

// Stable LSD radix passes over (u32 key, u32 value) pairs.  Sorts bits [begin_bit, end_bit) of the keys,
// RADIX_BITS per pass, ping-ponging between buffers 0 and 1.  Returns the index (0/1) of the buffer that
// holds the result.
int radix_sort_pairs(uint32_t* keys[2], uint32_t* vals[2], int64_t n, int begin_bit, int end_bit,
                     uint32_t* hist, uint32_t* digit_total, cudaStream_t s,
                     const unsigned long long* n_dev = nullptr, const uint32_t* first_keys = nullptr,
                     uint2* ranges = nullptr);   // ranges: the LAST pass also writes the tile ranges (load_range())

// partial sums of tiles[order[i]] per 2048 ranks (block_sums) and per 256 ranks (group_sums) for the emit kernel
void launch_offsets_scan(const uint32_t* tiles, const uint32_t* order, uint32_t* group_sums, uint32_t* block_sums,
                         int P, cudaStream_t s);
void launch_init_sort_input(const GeomView& g, int P, cudaStream_t s);
void launch_emit(const GeomView& g, const uint32_t* order, int P, int grid_x, uint32_t* tile_keys,
                 uint32_t* tile_vals, int64_t cap, cudaStream_t s);
void launch_publish_status(const unsigned long long* counters, int64_t cap, long long* status, cudaStream_t s);
void launch_build_keys(const uint32_t* sorted_tiles, const uint32_t* point_list, const uint32_t* depth_key,
                       int64_t R, const unsigned long long* n_dev, uint64_t* keys_out, cudaStream_t s);

void launch_render_forward(int W, int H, const uint2* ranges, const uint32_t* point_list, const GeomView& g,
                           const float* bg, float* out_color, float* out_depth, float* out_alpha,
                           uint32_t* n_contrib, cudaStream_t s);

size_t knn_scratch_bytes(int P);
int launch_knn_mean_dist2(int P, const float* points, float* mean_d2, void* scratch, cudaStream_t s);

void launch_expf_check(uint32_t first_bits, uint32_t count, unsigned long long* mismatches, cudaStream_t s);

void launch_pair_counts(int W, int H, const uint2* ranges, const uint32_t* point_list, const GeomView& g,
                        const uint32_t* n_contrib, unsigned long long* counts, cudaStream_t s);

void launch_render_backward(int W, int H, const uint2* ranges, const uint32_t* point_list, const GeomView& g,
                            const float* bg, const float* out_alpha, const uint32_t* n_contrib,
                            const float* dL_dcolor, const float* dL_ddepth, const float* dL_dalpha,
                            float* grad_scratch, cudaStream_t s, bool want_params = true);

struct PreprocessBwdArgs {
    const float* rigid = nullptr;   // same transform as the forward; dL_dmeans3D is then w.r.t. the raw points
    int P, D, M;
    const float* means3D;
    const int* radii;
    const float* shs;
    const float* scales;
    const float* rotations;
    float scale_modifier;
    const float* cov3D_precomp;
    const float* colors_precomp;
    const float* viewmatrix;
    const float* projmatrix;
    const float* campos;
    int W, H;
    float tanfovx, tanfovy;
    const float* grad_scratch;  // [P][12]
    float* dL_dmeans2D;
    float* dL_dcolors;
    float* dL_dopacity;
    float* dL_dmeans3D;
    float* dL_dcov3D;
    float* dL_dsh;
    int dsh_M;            // coefficients per Gaussian in dL_dsh (row stride / 3); >= (D+1)^2, == M unless the caller packs
    float* dL_dscales;
    float* dL_drotations;
    float* dL_dpose;      // optional [6]
    float* pose_scratch;  // [nblk][6] when dL_dpose != nullptr
};
void launch_preprocess_backward(const PreprocessBwdArgs& a, const GeomView& g, cudaStream_t s);
int pose_blocks(int P);

}  // namespace gsr
