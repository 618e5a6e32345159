// extern "C" boundary of libhtgs.so (see include/htgs.h).  Host-side only:
// carves the caller-owned work buffers, orders the kernel launches on the
// caller's stream, reports errors through htgs_last_error().
#include "../../include/htgs.h"
#include "common.cuh"
#include <cstdarg>
#include <cstdio>
#include <cstring>

namespace htgs {

static thread_local char g_error[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_error, sizeof(g_error), fmt, ap);
    va_end(ap);
}

template <typename T>
static T* carve(char*& p, size_t count) {
    p = reinterpret_cast<char*>(align_up(reinterpret_cast<size_t>(p), 256));
    T* r = reinterpret_cast<T*>(p);
    p += count * sizeof(T);
    return r;
}

GeomView geom_view(void* base, int P) {
    GeomView g;
    char* p = reinterpret_cast<char*>(base);
    const size_t n = (size_t)P;
    const size_t nb = (n + HTGS_PRE_BLOCK - 1) / HTGS_PRE_BLOCK;
    g.means2D = carve<float2>(p, n);
    g.conic_opacity = carve<float4>(p, n);
    g.feat = carve<float4>(p, n);
    g.clamped = carve<uint8_t>(p, n);
    g.tiles_touched = carve<uint32_t>(p, n);
    g.point_offsets = carve<uint32_t>(p, n);
    g.block_sums = carve<uint32_t>(p, nb + 1);
    g.num_rendered = carve<uint32_t>(p, 4);
    g.bytes = (size_t)(p - reinterpret_cast<char*>(base)) + 256;
    return g;
}

BinView bin_view(void* base, int64_t R, int W, int H) {
    (void)W; (void)H;
    BinView b;
    char* p = reinterpret_cast<char*>(base);
    const size_t n = (size_t)(R > 0 ? R : 0);
    b.values_sorted = carve<uint32_t>(p, n);  // first: the backward only needs this one, at offset 0 for any R
    b.keys_sorted = carve<uint64_t>(p, n);
    b.keys_alt = carve<uint64_t>(p, n);
    b.values_alt = carve<uint32_t>(p, n);
    b.reach = carve<uint8_t>(p, n);
    b.sort_scratch = carve<char>(p, sort_scratch_bytes(R));
    b.bytes = (size_t)(p - reinterpret_cast<char*>(base)) + 256;
    return b;
}

ImgView img_view(void* base, int W, int H) {
    ImgView v;
    char* p = reinterpret_cast<char*>(base);
    const size_t tiles = (size_t)((W + HTGS_TILE - 1) / HTGS_TILE) * ((H + HTGS_TILE - 1) / HTGS_TILE);
    v.ranges = carve<uint2>(p, tiles);
    v.n_contrib = carve<uint32_t>(p, (size_t)W * H);
    v.bin_capacity = carve<uint32_t>(p, 2);
    v.bytes = (size_t)(p - reinterpret_cast<char*>(base)) + 256;
    return v;
}

// ---- per-stage timing: explicit, process-wide measurement aid (the one piece of global state;
// not thread-safe, off by default) ----
static bool g_prof_on = false;
static const int PROF_CAP = 8192;
static cudaEvent_t g_prof_ev[PROF_CAP][2];
static int g_prof_stage[PROF_CAP];
static int g_prof_n = 0, g_prof_created = 0;

StageTimer::StageTimer(int stage, cudaStream_t s) : slot(-1), st(s) {
    if (!g_prof_on || g_prof_n >= PROF_CAP) return;
    if (g_prof_n >= g_prof_created) {
        if (cudaEventCreate(&g_prof_ev[g_prof_n][0]) != cudaSuccess || cudaEventCreate(&g_prof_ev[g_prof_n][1]) != cudaSuccess) return;
        g_prof_created = g_prof_n + 1;
    }
    slot = g_prof_n++;
    g_prof_stage[slot] = stage;
    cudaEventRecord(g_prof_ev[slot][0], st);
}
StageTimer::~StageTimer() {
    if (slot >= 0) cudaEventRecord(g_prof_ev[slot][1], st);
}

static bool aligned256(const void* p) { return (reinterpret_cast<size_t>(p) & 255) == 0; }

}  // namespace htgs

using namespace htgs;

extern "C" {

int htgs_abi_version(void) { return HTGS_ABI_VERSION; }

void htgs_profile_enable(int on) {
    g_prof_on = on != 0;
    g_prof_n = 0;
}

int htgs_profile_read(double* ms_sum, int* counts, int cap) {
    for (int i = 0; i < cap; i++) { ms_sum[i] = 0.0; counts[i] = 0; }
    for (int i = 0; i < g_prof_n; i++) {
        if (cudaEventSynchronize(g_prof_ev[i][1]) != cudaSuccess) { set_error("htgs_profile_read: event sync failed"); return -1; }
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, g_prof_ev[i][0], g_prof_ev[i][1]) != cudaSuccess) { set_error("htgs_profile_read: elapsed failed"); return -1; }
        const int s = g_prof_stage[i];
        if (s < cap) { ms_sum[s] += ms; counts[s]++; }
    }
    g_prof_n = 0;
    return ST_COUNT;
}

int htgs_blend_stats(int P, int64_t R, int W, int H, const void* geom_buf, const void* binning_buf, const void* image_buf,
                     unsigned long long* stats3, void* stream) {
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    if (P <= 0 || !geom_buf || !image_buf || !stats3) { set_error("htgs_blend_stats: bad arguments"); return 1; }
    GeomView g = geom_view(const_cast<void*>(geom_buf), P);
    ImgView img = img_view(const_cast<void*>(image_buf), W, H);
    BinView b = bin_view(const_cast<void*>(binning_buf), R, W, H);
    HTGS_CUDA_CHECK(cudaMemsetAsync(stats3, 0, HTGS_BLEND_NSTATS * sizeof(unsigned long long), st));
    return launch_blend_stats(W, H, img.ranges, b.values_sorted, g, stats3, st);
}
const char* htgs_last_error(void) { return g_error; }

size_t htgs_geom_bytes(int P) { return geom_view(nullptr, P < 0 ? 0 : P).bytes; }
size_t htgs_image_bytes(int W, int H) { return img_view(nullptr, W, H).bytes; }
size_t htgs_binning_bytes(int64_t R, int W, int H) { return bin_view(nullptr, R, W, H).bytes; }
size_t htgs_backward_scratch_bytes(int P) { return align_up((size_t)(P < 0 ? 0 : P) * 12 * sizeof(float), 256) + 256; }
size_t htgs_sort_scratch_bytes(int64_t n) { return sort_scratch_bytes(n); }

int htgs_geom_layout(int P, size_t* offsets, int cap) {
    GeomView g = geom_view(nullptr, P);
    const size_t o[HTGS_GEOM_NFIELDS] = {(size_t)g.means2D, (size_t)g.conic_opacity, (size_t)g.feat, (size_t)g.clamped,
                                         (size_t)g.tiles_touched, (size_t)g.point_offsets, (size_t)g.block_sums,
                                         (size_t)g.num_rendered};
    int k = 0;
    for (; k < HTGS_GEOM_NFIELDS && k < cap; k++) offsets[k] = o[k];
    return k;
}
int htgs_binning_layout(int64_t R, int W, int H, size_t* offsets, int cap) {
    BinView b = bin_view(nullptr, R, W, H);
    const size_t o[HTGS_BIN_NFIELDS] = {(size_t)b.keys_sorted, (size_t)b.values_sorted, (size_t)b.keys_alt,
                                        (size_t)b.values_alt, (size_t)b.sort_scratch};
    int k = 0;
    for (; k < HTGS_BIN_NFIELDS && k < cap; k++) offsets[k] = o[k];
    return k;
}
int htgs_image_layout(int W, int H, size_t* offsets, int cap) {
    ImgView v = img_view(nullptr, W, H);
    const size_t o[HTGS_IMG_NFIELDS] = {(size_t)v.ranges, (size_t)v.n_contrib};
    int k = 0;
    for (; k < HTGS_IMG_NFIELDS && k < cap; k++) offsets[k] = o[k];
    return k;
}

static int forward_preprocess_impl(int P, int D, int M, const float* means3D, const float* scales, float scale_modifier,
                                   const float* rotations, const float* opacities, const float* shs, const float* shs_rest,
                                   const float* cov3D_precomp, const float* colors_precomp, const float* viewmatrix,
                                   const float* projmatrix, const float* campos, int W, int H, float tanfovx, float tanfovy,
                                   int prefiltered, void* geom_buf, int* radii, cudaStream_t st, const char* who) {
    if (P < 0 || W <= 0 || H <= 0) { set_error("%s: bad sizes P=%d W=%d H=%d", who, P, W, H); return 1; }
    if (P == 0) return 0;
    if (!means3D || !opacities || !viewmatrix || !projmatrix || !campos || !geom_buf || !radii) {
        set_error("%s: null required pointer", who);
        return 1;
    }
    if (!shs && !colors_precomp) {
        // the plugin layer raises before reaching here (diff_gaussian_rasterization/__init__.py:194-195)
        set_error("%s: need shs or colors_precomp", who);
        return 1;
    }
    if (!cov3D_precomp && (!scales || !rotations)) { set_error("%s: need scales+rotations or cov3D_precomp", who); return 1; }
    if (D < 0 || D > 3) { set_error("%s: sh degree %d outside 0..3", who, D); return 1; }
    if (shs && !colors_precomp && M < (D + 1) * (D + 1)) { set_error("%s: M=%d too small for degree %d", who, M, D); return 1; }
    if (!aligned256(geom_buf)) { set_error("%s: geom_buf must be 256-byte aligned", who); return 1; }
    if (rotations && (reinterpret_cast<size_t>(rotations) & 15)) { set_error("%s: rotations must be 16-byte aligned", who); return 1; }
    GeomView g = geom_view(geom_buf, P);
    { StageTimer t(ST_PREPROCESS, st);
    if (launch_preprocess(P, D, M, means3D, scales, scale_modifier, rotations, opacities, shs, shs_rest, cov3D_precomp, colors_precomp,
                          viewmatrix, projmatrix, campos, W, H, tanfovx, tanfovy, prefiltered, g, radii, st)) return 1; }
    { StageTimer t(ST_SCAN, st);
    if (launch_scan_block_sums(P, g, st)) return 1; }
    return 0;
}

int htgs_forward_preprocess(int P, int D, int M, const float* means3D, const float* scales, float scale_modifier,
                            const float* rotations, const float* opacities, const float* shs,
                            const float* cov3D_precomp, const float* colors_precomp, const float* viewmatrix,
                            const float* projmatrix, const float* campos, int W, int H, float tanfovx, float tanfovy,
                            int prefiltered, void* geom_buf, int* radii, int64_t* num_rendered_host, void* stream) {
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    if (!num_rendered_host) { set_error("htgs_forward_preprocess: null num_rendered_host"); return 1; }
    *num_rendered_host = 0;
    if (forward_preprocess_impl(P, D, M, means3D, scales, scale_modifier, rotations, opacities, shs, nullptr, cov3D_precomp,
                                colors_precomp, viewmatrix, projmatrix, campos, W, H, tanfovx, tanfovy, prefiltered, geom_buf,
                                radii, st, "htgs_forward_preprocess")) return 1;
    if (P == 0) return 0;
    GeomView g = geom_view(geom_buf, P);
    uint32_t r32 = 0;
    HTGS_CUDA_CHECK(cudaMemcpyAsync(&r32, g.num_rendered, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    HTGS_CUDA_CHECK(cudaStreamSynchronize(st));
    *num_rendered_host = (int64_t)r32;
    return 0;
}

int htgs_forward_preprocess_async(int P, int D, int M, const float* means3D, const float* scales, float scale_modifier,
                                  const float* rotations, const float* opacities, const float* shs,
                                  const float* cov3D_precomp, const float* colors_precomp, const float* viewmatrix,
                                  const float* projmatrix, const float* campos, int W, int H, float tanfovx, float tanfovy,
                                  int prefiltered, void* geom_buf, int* radii, uint32_t* num_rendered_pinned_host,
                                  void* stream) {
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    if (!num_rendered_pinned_host) { set_error("htgs_forward_preprocess_async: null num_rendered_pinned_host"); return 1; }
    if (forward_preprocess_impl(P, D, M, means3D, scales, scale_modifier, rotations, opacities, shs, nullptr, cov3D_precomp,
                                colors_precomp, viewmatrix, projmatrix, campos, W, H, tanfovx, tanfovy, prefiltered, geom_buf,
                                radii, st, "htgs_forward_preprocess_async")) return 1;
    if (P == 0) { *num_rendered_pinned_host = 0; return 0; }
    GeomView g = geom_view(geom_buf, P);
    HTGS_CUDA_CHECK(cudaMemcpyAsync(num_rendered_pinned_host, g.num_rendered, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    return 0;
}

int htgs_forward_preprocess_raw_async(int P, int D, const float* xyz, const float* scaling_raw, float scale_modifier,
                                      const float* rotation_raw, const float* opacity_raw, const float* features_dc,
                                      const float* features_rest, const float* viewmatrix, const float* projmatrix,
                                      const float* campos, int W, int H, float tanfovx, float tanfovy, int prefiltered,
                                      void* geom_buf, int* radii, uint32_t* num_rendered_pinned_host, void* stream) {
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    const char* who = "htgs_forward_preprocess_raw_async";
    if (!num_rendered_pinned_host) { set_error("%s: null num_rendered_pinned_host", who); return 1; }
    if (P > 0 && (!scaling_raw || !rotation_raw || !opacity_raw || !features_dc || !features_rest)) {
        set_error("%s: null parameter pointer", who);
        return 1;
    }
    if (reinterpret_cast<size_t>(features_rest) & 15) { set_error("%s: features_rest must be 16-byte aligned", who); return 1; }
    if (forward_preprocess_impl(P, D, 16, xyz, scaling_raw, scale_modifier, rotation_raw, opacity_raw, features_dc,
                                features_rest, nullptr, nullptr, viewmatrix, projmatrix, campos, W, H, tanfovx, tanfovy,
                                prefiltered, geom_buf, radii, st, who)) return 1;
    if (P == 0) { *num_rendered_pinned_host = 0; return 0; }
    GeomView g = geom_view(geom_buf, P);
    HTGS_CUDA_CHECK(cudaMemcpyAsync(num_rendered_pinned_host, g.num_rendered, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    return 0;
}

// R_or_capacity sizes the binning buffer; with count_on_device the kernels read the true count from
// the geometry buffer (written by the scan) and R_or_capacity is only an upper bound.
static int forward_render_impl(int P, int64_t R_or_capacity, bool count_on_device, int W, int H, const float* background,
                               const int* radii, void* geom_buf, void* binning_buf, void* image_buf, float* out_color,
                               float* out_depth, float* out_alpha, cudaStream_t st, const char* who) {
    const int64_t R = R_or_capacity;
    if (P <= 0) return 0;
    if (!background || !radii || !geom_buf || !image_buf || !out_color || !out_depth || !out_alpha || (R > 0 && !binning_buf)) {
        set_error("%s: null required pointer", who);
        return 1;
    }
    if (!aligned256(geom_buf) || !aligned256(image_buf) || (R > 0 && !aligned256(binning_buf))) {
        set_error("%s: work buffers must be 256-byte aligned", who);
        return 1;
    }
    if (R < 0 || R >= (int64_t(1) << 30)) { set_error("%s: R=%lld out of range", who, (long long)R); return 1; }
    GeomView g = geom_view(geom_buf, P);
    ImgView img = img_view(image_buf, W, H);
    BinView b = bin_view(binning_buf, R, W, H);
    const uint32_t* R_dev = count_on_device ? g.num_rendered : nullptr;
    const int gx = (W + HTGS_TILE - 1) / HTGS_TILE, gy = (H + HTGS_TILE - 1) / HTGS_TILE;
    if (R > 0) {
        const int end_bit = 32 + (int)higher_msb((uint32_t)(gx * gy));  // rasterizer_impl.cu:301-309
        const int passes = sort_num_passes(end_bit);
        // start in the buffer that makes the last pass land in keys_sorted/values_sorted
        uint64_t* k0 = (passes & 1) ? b.keys_alt : b.keys_sorted;
        uint32_t* v0 = (passes & 1) ? b.values_alt : b.values_sorted;
        uint64_t* k1 = (passes & 1) ? b.keys_sorted : b.keys_alt;
        uint32_t* v1 = (passes & 1) ? b.values_sorted : b.values_alt;
        { StageTimer t(ST_MAKE_KEYS, st);
        if (launch_make_keys(P, W, H, radii, g, k0, v0, R, st)) return 1; }
        { StageTimer t(ST_SORT, st);
        if (sort_pairs(R, end_bit, k0, v0, k1, v1, b.sort_scratch, st, R_dev)) return 1; }
    }
    { StageTimer t(ST_TILE_RANGES, st);
    if (launch_tile_ranges(R, R_dev, W, H, b.keys_sorted, img.ranges, st)) return 1; }
    { StageTimer t(ST_BLEND_FWD, st);
    if (launch_blend_forward(W, H, img.ranges, b.values_sorted, g, background, out_color, out_depth, out_alpha,
                             img.n_contrib, R > 0 ? b.reach : nullptr, img.bin_capacity, (uint32_t)R, st)) return 1; }
    return 0;
}

int htgs_forward_render(int P, int64_t R, int W, int H, const float* background, const int* radii, void* geom_buf,
                        void* binning_buf, void* image_buf, float* out_color, float* out_depth, float* out_alpha,
                        void* stream) {
    return forward_render_impl(P, R, false, W, H, background, radii, geom_buf, binning_buf, image_buf, out_color, out_depth,
                               out_alpha, reinterpret_cast<cudaStream_t>(stream), "htgs_forward_render");
}

int htgs_forward_render_capacity(int P, int64_t capacity, int W, int H, const float* background, const int* radii,
                                 void* geom_buf, void* binning_buf, void* image_buf, float* out_color, float* out_depth,
                                 float* out_alpha, void* stream) {
    if (capacity <= 0) { set_error("htgs_forward_render_capacity: capacity must be positive"); return 1; }
    return forward_render_impl(P, capacity, true, W, H, background, radii, geom_buf, binning_buf, image_buf, out_color,
                               out_depth, out_alpha, reinterpret_cast<cudaStream_t>(stream), "htgs_forward_render_capacity");
}

static int backward_impl(const char* who, const float* shs_rest, const float* opacity_raw, float* dL_dsh_rest,
                         float* rgb_slot, bool rgb_multicast,
                         int P, int D, int M, int64_t R, int W, int H, const float* background, const float* means3D,
                  const float* shs, const float* colors_precomp, const float* alphas, const float* scales,
                  float scale_modifier, const float* rotations, const float* cov3D_precomp, const float* viewmatrix,
                  const float* projmatrix, const float* campos, float tanfovx, float tanfovy, const int* radii,
                  const void* geom_buf, const void* binning_buf, const void* image_buf, const float* dL_dout_color,
                  const float* dL_dout_depth, const float* dL_dout_alpha, void* scratch, float* dL_dmeans2D,
                  float* dL_dopacity, float* dL_dcolors, float* dL_dmeans3D, float* dL_dcov3D, float* dL_dsh,
                  float* dL_dscales, float* dL_drotations, void* stream) {
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    if (P <= 0) return 0;
    if (!background || !means3D || !alphas || !viewmatrix || !projmatrix || !campos || !radii || !geom_buf || !image_buf ||
        !dL_dout_color || !dL_dout_depth || !dL_dout_alpha || !scratch || !dL_dmeans2D || !dL_dopacity || !dL_dmeans3D ||
        (R > 0 && !binning_buf)) {
        set_error("%s: null required pointer", who);
        return 1;
    }
    if (shs && !dL_dsh && !rgb_slot) { set_error("%s: dL_dsh required with shs", who); return 1; }
    if (rgb_slot && (!shs || (reinterpret_cast<size_t>(rgb_slot) & 15))) { set_error("%s: the colour-gradient slot needs the SH path and 16-byte alignment", who); return 1; }
    if (colors_precomp && !dL_dcolors) { set_error("%s: dL_dcolors required with colors_precomp", who); return 1; }
    if (scales && (!rotations || !dL_dscales || !dL_drotations)) { set_error("%s: scale/rotation outputs required", who); return 1; }
    if (!scales && !cov3D_precomp) { set_error("%s: need scales+rotations or cov3D_precomp", who); return 1; }
    if (cov3D_precomp && !dL_dcov3D) { set_error("%s: dL_dcov3D required with cov3D_precomp", who); return 1; }
    if (!aligned256(scratch) || !aligned256(geom_buf) || !aligned256(image_buf)) { set_error("%s: buffers must be 256-byte aligned", who); return 1; }
    if ((dL_drotations && (reinterpret_cast<size_t>(dL_drotations) & 15)) || (rotations && (reinterpret_cast<size_t>(rotations) & 15))) {
        set_error("%s: rotations / dL_drotations must be 16-byte aligned", who);
        return 1;
    }
    GeomView g = geom_view(const_cast<void*>(geom_buf), P);
    ImgView img = img_view(const_cast<void*>(image_buf), W, H);
    BinView b = bin_view(const_cast<void*>(binning_buf), R, W, H);
    float* gacc = reinterpret_cast<float*>(scratch);
    { StageTimer t(ST_ZERO_ACC, st);
    HTGS_CUDA_CHECK(cudaMemsetAsync(gacc, 0, (size_t)P * 12 * sizeof(float), st)); }
    if (R > 0) {
        StageTimer t(ST_BLEND_BWD, st);
        if (launch_blend_backward(P, W, H, img.ranges, b.values_sorted, g, background, alphas, img.n_contrib, dL_dout_color,
                                  dL_dout_depth, dL_dout_alpha, gacc, b.reach, img.bin_capacity, (uint32_t)R, st)) return 1;
    }
    // compact SH mode: the clamp-masked colour gradient leaves for its view slot as soon as the blend backward is done, on a
    // side stream beside the geometry backward (joined below)
    if (rgb_slot && launch_rgb_extract(P, gacc, g.clamped, campos, rgb_slot, rgb_multicast, st)) return 1;
    // the SH path feeds dL/dcolour into dL/dsh; only the precomputed-colour path returns it
    StageTimer t(ST_GEOM_BWD, st);
    if (launch_geom_backward(P, D, M, means3D, radii, shs, shs_rest, opacity_raw, dL_dsh_rest, colors_precomp, scales, scale_modifier, rotations, cov3D_precomp,
                             viewmatrix, projmatrix, campos, W, H, tanfovx, tanfovy, g, gacc, dL_dmeans2D, dL_dopacity,
                             dL_dcolors, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations, st)) return 1;
    if (rgb_slot && join_rgb_extract(st)) return 1;  // the stream the caller sees again orders everything, the push included
    return 0;
}

int htgs_backward(int P, int D, int M, int64_t R, int W, int H, const float* background, const float* means3D,
                  const float* shs, const float* colors_precomp, const float* alphas, const float* scales,
                  float scale_modifier, const float* rotations, const float* cov3D_precomp, const float* viewmatrix,
                  const float* projmatrix, const float* campos, float tanfovx, float tanfovy, const int* radii,
                  const void* geom_buf, const void* binning_buf, const void* image_buf, const float* dL_dout_color,
                  const float* dL_dout_depth, const float* dL_dout_alpha, void* scratch, float* dL_dmeans2D,
                  float* dL_dopacity, float* dL_dcolors, float* dL_dmeans3D, float* dL_dcov3D, float* dL_dsh,
                  float* dL_dscales, float* dL_drotations, void* stream) {
    return backward_impl("htgs_backward", nullptr, nullptr, nullptr, nullptr, false, P, D, M, R, W, H, background, means3D, shs, colors_precomp,
                         alphas, scales, scale_modifier, rotations, cov3D_precomp, viewmatrix, projmatrix, campos, tanfovx,
                         tanfovy, radii, geom_buf, binning_buf, image_buf, dL_dout_color, dL_dout_depth, dL_dout_alpha, scratch,
                         dL_dmeans2D, dL_dopacity, dL_dcolors, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations, stream);
}

int htgs_backward_compact(int P, int D, int M, int64_t R, int W, int H, const float* background, const float* means3D,
                          const float* shs, const float* alphas, const float* scales, float scale_modifier,
                          const float* rotations, const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix,
                          const float* campos, float tanfovx, float tanfovy, const int* radii, const void* geom_buf,
                          const void* binning_buf, const void* image_buf, const float* dL_dout_color, const float* dL_dout_depth,
                          const float* dL_dout_alpha, void* scratch, float* dL_dmeans2D, float* dL_dopacity, float* dL_dmeans3D,
                          float* dL_dcov3D, float* dL_dscales, float* dL_drotations, float* rgb_slot, int rgb_slot_is_multicast,
                          void* stream) {
    if (P > 0 && (!shs || !rgb_slot)) { set_error("htgs_backward_compact: shs and rgb_slot are required"); return 1; }
    return backward_impl("htgs_backward_compact", nullptr, nullptr, nullptr, rgb_slot, rgb_slot_is_multicast != 0, P, D, M, R, W, H,
                         background, means3D, shs, nullptr, alphas, scales, scale_modifier, rotations, cov3D_precomp, viewmatrix,
                         projmatrix, campos, tanfovx, tanfovy, radii, geom_buf, binning_buf, image_buf, dL_dout_color,
                         dL_dout_depth, dL_dout_alpha, scratch, dL_dmeans2D, dL_dopacity, nullptr, dL_dmeans3D, dL_dcov3D, nullptr,
                         dL_dscales, dL_drotations, stream);
}

int htgs_backward_raw(int P, int D, int64_t R, int W, int H, const float* background, const float* xyz,
                      const float* features_dc, const float* features_rest, const float* alphas, const float* scaling_raw,
                      float scale_modifier, const float* rotation_raw, const float* opacity_raw, const float* viewmatrix,
                      const float* projmatrix, const float* campos, float tanfovx, float tanfovy, const int* radii,
                      const void* geom_buf, const void* binning_buf, const void* image_buf, const float* dL_dout_color,
                      const float* dL_dout_depth, const float* dL_dout_alpha, void* scratch, float* dL_dmeans2D,
                      float* dL_dxyz, float* dL_dfeatures_dc, float* dL_dfeatures_rest, float* dL_dopacity_raw,
                      float* dL_dscaling_raw, float* dL_drotation_raw, void* stream) {
    const char* who = "htgs_backward_raw";
    if (P <= 0) return 0;
    if (!features_dc || !features_rest || !scaling_raw || !rotation_raw || !opacity_raw || !dL_dfeatures_dc ||
        !dL_dfeatures_rest || !dL_dopacity_raw || !dL_dscaling_raw || !dL_drotation_raw) {
        set_error("%s: null parameter / gradient pointer", who);
        return 1;
    }
    if ((reinterpret_cast<size_t>(features_rest) & 15) || (reinterpret_cast<size_t>(dL_dfeatures_rest) & 15)) {
        set_error("%s: features_rest and its gradient must be 16-byte aligned", who);
        return 1;
    }
    return backward_impl(who, features_rest, opacity_raw, dL_dfeatures_rest, nullptr, false, P, D, 16, R, W, H, background, xyz, features_dc,
                         nullptr, alphas, scaling_raw, scale_modifier, rotation_raw, nullptr, viewmatrix, projmatrix, campos,
                         tanfovx, tanfovy, radii, geom_buf, binning_buf, image_buf, dL_dout_color, dL_dout_depth, dL_dout_alpha,
                         scratch, dL_dmeans2D, dL_dopacity_raw, nullptr, dL_dxyz, nullptr, dL_dfeatures_dc, dL_dscaling_raw,
                         dL_drotation_raw, stream);
}

int htgs_mark_visible(int P, const float* means3D, const float* viewmatrix, const float* projmatrix, uint8_t* present,
                      void* stream) {
    (void)projmatrix;
    if (P <= 0) return 0;
    if (!means3D || !viewmatrix || !present) { set_error("htgs_mark_visible: null pointer"); return 1; }
    return launch_mark_visible(P, means3D, viewmatrix, present, reinterpret_cast<cudaStream_t>(stream));
}

int htgs_sort_pairs(int64_t n, int end_bit, uint64_t* keys_in, uint32_t* values_in, uint64_t* keys_out,
                    uint32_t* values_out, void* scratch, void* stream) {
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    if (n <= 0) return 0;
    if (!keys_in || !values_in || !keys_out || !values_out || !scratch) { set_error("htgs_sort_pairs: null pointer"); return 1; }
    if (sort_pairs(n, end_bit, keys_in, values_in, keys_out, values_out, scratch, st)) return 1;
    if ((sort_num_passes(end_bit) & 1) == 0) {  // even pass count: result sits in the input buffers
        HTGS_CUDA_CHECK(cudaMemcpyAsync(keys_out, keys_in, (size_t)n * sizeof(uint64_t), cudaMemcpyDeviceToDevice, st));
        HTGS_CUDA_CHECK(cudaMemcpyAsync(values_out, values_in, (size_t)n * sizeof(uint32_t), cudaMemcpyDeviceToDevice, st));
    }
    return 0;
}

}  // extern "C"
