// Shared declarations for the sm_100a rasterizer kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>

#define HTGS_TILE 16            // reference tile edge (cuda_rasterizer/config.h:16-17)
#define HTGS_TILE_PIX 256
#define HTGS_PRE_BLOCK 256      // Gaussians per preprocess / key-generation block

namespace htgs {

void set_error(const char* fmt, ...);

#define HTGS_CUDA_CHECK(expr)                                                              \
    do {                                                                                   \
        cudaError_t _e = (expr);                                                           \
        if (_e != cudaSuccess) {                                                           \
            ::htgs::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
            return 1;                                                                      \
        }                                                                                  \
    } while (0)

#define HTGS_LAUNCH_CHECK(name)                                                            \
    do {                                                                                   \
        cudaError_t _e = cudaGetLastError();                                               \
        if (_e != cudaSuccess) {                                                           \
            ::htgs::set_error("launch of %s failed: %s", name, cudaGetErrorString(_e));    \
            return 1;                                                                      \
        }                                                                                  \
    } while (0)

static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// ---- buffer layouts (opaque to callers; see htgs_*_layout) -----------------
struct GeomView {
    float2* means2D;
    float4* conic_opacity;
    float4* feat;
    uint8_t* clamped;
    uint32_t* tiles_touched;
    uint32_t* point_offsets;
    uint32_t* block_sums;
    uint32_t* num_rendered;
    size_t bytes;
};
struct BinView {
    uint64_t* keys_sorted;
    uint32_t* values_sorted;
    uint64_t* keys_alt;
    uint32_t* values_alt;
    uint8_t* reach;  // reach mask of every sorted instance, written by the forward blend for the backward (see blend_common.cuh)
    void* sort_scratch;
    size_t bytes;
};
struct ImgView {
    uint2* ranges;
    uint32_t* n_contrib;
    uint32_t* bin_capacity;  // [2] the instance count the binning buffer of the last forward was carved with, and a magic word
    size_t bytes;
};
GeomView geom_view(void* base, int P);
BinView bin_view(void* base, int64_t R, int W, int H);
ImgView img_view(void* base, int W, int H);

// ---- kernels' host launchers (one per .cu) ---------------------------------
int launch_preprocess(int P, int D, int M, const float* means3D, const float* scales, float scale_modifier,
                      const float* rotations, const float* opacities, const float* shs, const float* shs_rest,
                      const float* cov3D_precomp, const float* colors_precomp, const float* viewmatrix,
                      const float* projmatrix, const float* campos, int W, int H, float tanfovx, float tanfovy,
                      int prefiltered, const GeomView& g, int* radii, cudaStream_t st);
int launch_scan_block_sums(int P, const GeomView& g, cudaStream_t st);
int launch_mark_visible(int P, const float* means3D, const float* viewmatrix, uint8_t* present, cudaStream_t st);
int launch_make_keys(int P, int W, int H, const int* radii, const GeomView& g, uint64_t* keys, uint32_t* values,
                     int64_t capacity, cudaStream_t st);
int launch_tile_ranges(int64_t R, const uint32_t* R_dev, int W, int H, const uint64_t* keys_sorted, uint2* ranges,
                       cudaStream_t st);
int launch_blend_forward(int W, int H, const uint2* ranges, const uint32_t* point_list, const GeomView& g,
                         const float* bg, float* out_color, float* out_depth, float* out_alpha,
                         uint32_t* n_contrib, uint8_t* reach, uint32_t* bin_capacity, uint32_t capacity, cudaStream_t st);
int launch_blend_stats(int W, int H, const uint2* ranges, const uint32_t* point_list, const GeomView& g,
                       unsigned long long* stats, cudaStream_t st);
int launch_blend_backward(int P, int W, int H, const uint2* ranges, const uint32_t* point_list, const GeomView& g,
                          const float* bg, const float* alphas, const uint32_t* n_contrib, const float* dL_dpix,
                          const float* dL_ddepth, const float* dL_dalpha, float* gacc, const uint8_t* reach,
                          const uint32_t* bin_capacity, uint32_t capacity, cudaStream_t st);
int launch_geom_backward(int P, int D, int M, const float* means3D, const int* radii, const float* shs,
                         const float* shs_rest, const float* opacity_raw, float* dL_dsh_rest, const float* colors_precomp, const float* scales, float scale_modifier,
                         const float* rotations, const float* cov3D_precomp, const float* viewmatrix,
                         const float* projmatrix, const float* campos, int W, int H, float tanfovx, float tanfovy,
                         const GeomView& g, const float* gacc, float* dL_dmeans2D, float* dL_dopacity,
                         float* dL_dcolors, float* dL_dmeans3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscales,
                         float* dL_drotations, cudaStream_t st);

int launch_rgb_extract(int P, const float* gacc, const uint8_t* clamped, const float* campos, float* slot, bool multicast,
                       cudaStream_t st);
int join_rgb_extract(cudaStream_t st);

size_t sort_scratch_bytes(int64_t n);
int sort_pairs(int64_t n, int end_bit, uint64_t* k_in, uint32_t* v_in, uint64_t* k_out, uint32_t* v_out,
               void* scratch, cudaStream_t st, const uint32_t* n_dev = nullptr);
// number of passes sort_pairs will run for end_bit (decides ping-pong parity)
int sort_num_passes(int end_bit);

uint32_t higher_msb(uint32_t n);

// ---- optional per-stage timing (htgs_profile_*) ----
enum Stage { ST_PREPROCESS = 0, ST_SCAN, ST_MAKE_KEYS, ST_SORT, ST_TILE_RANGES, ST_BLEND_FWD, ST_ZERO_ACC, ST_BLEND_BWD,
             ST_GEOM_BWD, ST_COUNT };
struct StageTimer {  // records a CUDA event pair around a stage when profiling is on
    StageTimer(int stage, cudaStream_t st);
    ~StageTimer();
    int slot;
    cudaStream_t st;
};

}  // namespace htgs
