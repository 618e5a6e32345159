// ref_shim_dgr.cu -- TEST INFRASTRUCTURE.  A thin extern "C" door onto the
// UNMODIFIED reference CUDA rasterizer (compiled by oracle/Makefile from
// /root/reference/submodules/diff-gaussian-rasterization/cuda_rasterizer/*.cu
// where those sources lie) so the GPU parity tests and the A/B timing harness
// can run it without its torch/pybind glue.  It plays the role of
// rasterize_points.cu: owns the three resizable byte buffers and forwards raw
// pointers to CudaRasterizer::Rasterizer (rasterizer.h:20-92).  Nothing here
// is part of the product.
#include <cuda_runtime.h>
#include <cstdint>
#include <cstring>
#include <functional>
#include "rasterizer.h"
#include "rasterizer_impl.h"

namespace {
struct Chunk {
    char* ptr = nullptr;
    size_t cap = 0;
    char* get(size_t n) {
        if (n > cap) {
            if (ptr) cudaFree(ptr);
            cap = n + n / 4 + 4096;
            if (cudaMalloc(&ptr, cap) != cudaSuccess) { ptr = nullptr; cap = 0; }
        }
        return ptr;
    }
};
Chunk g_geom, g_bin, g_img;
int g_P = 0, g_R = 0, g_W = 0, g_H = 0;
}  // namespace

extern "C" {

int ref_forward(int P, int D, int M, const float* bg, int W, int H, const float* means3D, const float* shs,
                const float* colors_precomp, const float* opacities, const float* scales, float scale_modifier,
                const float* rotations, const float* cov3D_precomp, const float* view, const float* proj,
                const float* campos, float tanfovx, float tanfovy, int prefiltered, float* out_color, float* out_depth,
                float* out_alpha, int* radii, int debug) {
    std::function<char*(size_t)> gf = [](size_t n) { return g_geom.get(n); };
    std::function<char*(size_t)> bf = [](size_t n) { return g_bin.get(n); };
    std::function<char*(size_t)> imf = [](size_t n) { return g_img.get(n); };
    int R = CudaRasterizer::Rasterizer::forward(gf, bf, imf, P, D, M, bg, W, H, means3D, shs, colors_precomp, opacities,
                                                scales, scale_modifier, rotations, cov3D_precomp, view, proj, campos,
                                                tanfovx, tanfovy, prefiltered != 0, out_color, out_depth, out_alpha, radii,
                                                debug != 0);
    g_P = P; g_R = R; g_W = W; g_H = H;
    return R;
}

void ref_backward(int P, int D, int M, int R, const float* bg, int W, int H, const float* means3D, const float* shs,
                  const float* colors_precomp, const float* alphas, const float* scales, float scale_modifier,
                  const float* rotations, const float* cov3D_precomp, const float* view, const float* proj,
                  const float* campos, float tanfovx, float tanfovy, const int* radii, const float* dL_dpix,
                  const float* dL_dpix_depth, const float* dL_dalphas, float* dL_dmean2D, float* dL_dconic,
                  float* dL_dopacity, float* dL_dcolor, float* dL_ddepth, float* dL_dmean3D, float* dL_dcov3D,
                  float* dL_dsh, float* dL_dscale, float* dL_drot, int debug) {
    CudaRasterizer::Rasterizer::backward(P, D, M, R, bg, W, H, means3D, shs, colors_precomp, alphas, scales,
                                         scale_modifier, rotations, cov3D_precomp, view, proj, campos, tanfovx, tanfovy,
                                         radii, g_geom.ptr, g_bin.ptr, g_img.ptr, dL_dpix, dL_dpix_depth, dL_dalphas,
                                         dL_dmean2D, dL_dconic, dL_dopacity, dL_dcolor, dL_ddepth, dL_dmean3D, dL_dcov3D,
                                         dL_dsh, dL_dscale, dL_drot, debug != 0);
}

void ref_mark_visible(int P, float* means3D, float* view, float* proj, bool* present) {
    CudaRasterizer::Rasterizer::markVisible(P, means3D, view, proj, present);
}

// Copies one array of the state of the last ref_forward into dst (device memory, `bytes` long).
// Returns the number of bytes the array holds, or -1 for an unknown name / short dst.
long long ref_get(const char* name, void* dst, size_t bytes) {
    if (!g_geom.ptr) return -1;
    char* gp = g_geom.ptr;
    CudaRasterizer::GeometryState geo = CudaRasterizer::GeometryState::fromChunk(gp, g_P);
    char* ip = g_img.ptr;
    CudaRasterizer::ImageState img = CudaRasterizer::ImageState::fromChunk(ip, (size_t)g_W * g_H);
    CudaRasterizer::BinningState bin;
    memset(&bin, 0, sizeof(bin));
    if (g_bin.ptr) { char* bp = g_bin.ptr; bin = CudaRasterizer::BinningState::fromChunk(bp, g_R); }
    const size_t P = g_P, R = g_R, tiles = (size_t)((g_W + 15) / 16) * ((g_H + 15) / 16);
    const void* src = nullptr;
    size_t n = 0;
    auto is = [&](const char* s) { return strcmp(name, s) == 0; };
    if (is("depths")) { src = geo.depths; n = P * 4; }
    else if (is("clamped")) { src = geo.clamped; n = P * 3; }
    else if (is("means2D")) { src = geo.means2D; n = P * 8; }
    else if (is("cov3D")) { src = geo.cov3D; n = P * 24; }
    else if (is("conic_opacity")) { src = geo.conic_opacity; n = P * 16; }
    else if (is("rgb")) { src = geo.rgb; n = P * 12; }
    else if (is("tiles_touched")) { src = geo.tiles_touched; n = P * 4; }
    else if (is("point_offsets")) { src = geo.point_offsets; n = P * 4; }
    else if (is("point_list")) { src = bin.point_list; n = R * 4; }
    else if (is("point_list_keys")) { src = bin.point_list_keys; n = R * 8; }
    else if (is("point_list_unsorted")) { src = bin.point_list_unsorted; n = R * 4; }
    else if (is("point_list_keys_unsorted")) { src = bin.point_list_keys_unsorted; n = R * 8; }
    else if (is("ranges")) { src = img.ranges; n = tiles * 8; }
    else if (is("n_contrib")) { src = img.n_contrib; n = (size_t)g_W * g_H * 4; }
    else return -1;
    if (n > bytes) return -1;
    if (n && cudaMemcpy(dst, src, n, cudaMemcpyDeviceToDevice) != cudaSuccess) return -1;
    return (long long)n;
}

void ref_release(void) {
    if (g_geom.ptr) cudaFree(g_geom.ptr);
    if (g_bin.ptr) cudaFree(g_bin.ptr);
    if (g_img.ptr) cudaFree(g_img.ptr);
    g_geom = Chunk(); g_bin = Chunk(); g_img = Chunk();
}
}
