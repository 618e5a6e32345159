// Merge-time kernels of the hierarchical trainer (second half; opt-in):
//   importance_accum_kernel   calc_importance (trainer/ht3dgs_trainer.py:1428-1462): per view,
//                             importance[i][k][c] += |dL/dsh[i][k][c]| for loss = image.sum().
//                             The reference obtains it from a full backward with abs() hooks on the feature tensors.
//                             dL/dsh is the outer product of the SH basis of the view direction with the clamp-masked
//                             colour gradient the blend backward leaves in its accumulator, so it is formed right
//                             there: no geometry backward, no gradient tensors, no hooks.
//   importance_keys_kernel +  the bottom-k choice of merge_two_3DGS (:232-238, :246-251): amax over the 48 entries,
//   mark_first_k_kernel       stable radix sort of (amax bits, index) with this library's own sort, mask of the first k.
//   transform_points_kernel   the homogeneous transform of the source model's positions (:254-257).
#include "../../include/htgs.h"
#include "sh_basis.cuh"

namespace htgs {

__global__ void __launch_bounds__(256)
importance_accum_kernel(int P, int D, int M, const float* __restrict__ means3D, const float* __restrict__ campos,
                        const int* __restrict__ radii, const uint8_t* __restrict__ clamped, const float* __restrict__ gacc,
                        float* __restrict__ importance) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P || !(radii[idx] > 0)) return;
    const float4 g = reinterpret_cast<const float4*>(gacc + (size_t)idx * 12)[2];  // dL/dcolour of this view
    const uint32_t cl = clamped[idx];
    const float r[3] = {(cl & 1) ? 0.f : fabsf(g.x), (cl & 2) ? 0.f : fabsf(g.y), (cl & 4) ? 0.f : fabsf(g.z)};
    if (r[0] == 0.f && r[1] == 0.f && r[2] == 0.f) return;
    const float ox = means3D[3 * (size_t)idx] - campos[0], oy = means3D[3 * (size_t)idx + 1] - campos[1],
                oz = means3D[3 * (size_t)idx + 2] - campos[2];
    const float inv = 1.0f / sqrtf(ox * ox + oy * oy + oz * oz);
    float w[16];
    const int ncoef = sh_basis(D, ox * inv, oy * inv, oz * inv, w);
    float* out = importance + (size_t)idx * M * 3;
#pragma unroll
    for (int k = 0; k < 16; k++) {
        if (k < ncoef && k < M) {
            const float a = fabsf(w[k]);
            out[3 * k] += a * r[0];
            out[3 * k + 1] += a * r[1];
            out[3 * k + 2] += a * r[2];
        }
    }
}

// key = bits of max_j importance[i][j] (non-negative floats order like their bit patterns), value = i
__global__ void __launch_bounds__(256)
importance_keys_kernel(int P, int n_per, const float* __restrict__ importance, uint64_t* __restrict__ keys,
                       uint32_t* __restrict__ values) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P) return;
    const float* row = importance + (size_t)idx * n_per;
    float m = 0.f;
    for (int j = 0; j < n_per; j++) m = fmaxf(m, fabsf(row[j]));
    keys[idx] = (uint64_t)__float_as_uint(m);
    values[idx] = (uint32_t)idx;
}

__global__ void __launch_bounds__(256)
mark_first_k_kernel(int64_t k, const uint32_t* __restrict__ order, uint8_t* __restrict__ mask) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < k) mask[order[i]] = 1;
}

// out = (T [x 1]^T)[:3] / (T [x 1]^T)[3],  T row-major 4x4 in device memory
__global__ void __launch_bounds__(256)
transform_points_kernel(int P, const float* __restrict__ xyz, const float* __restrict__ T, float* __restrict__ out) {
    __shared__ float s_T[16];
    if (threadIdx.x < 16) s_T[threadIdx.x] = T[threadIdx.x];
    __syncthreads();
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P) return;
    const float x = xyz[3 * (size_t)idx], y = xyz[3 * (size_t)idx + 1], z = xyz[3 * (size_t)idx + 2];
    float h[4];
#pragma unroll
    for (int r = 0; r < 4; r++) h[r] = fmaf(s_T[4 * r + 2], z, fmaf(s_T[4 * r + 1], y, fmaf(s_T[4 * r], x, s_T[4 * r + 3])));
    out[3 * (size_t)idx] = h[0] / h[3];
    out[3 * (size_t)idx + 1] = h[1] / h[3];
    out[3 * (size_t)idx + 2] = h[2] / h[3];
}

static inline size_t select_keys_bytes(int P) { return align_up((size_t)P * sizeof(uint64_t), 256); }
static inline size_t select_vals_bytes(int P) { return align_up((size_t)P * sizeof(uint32_t), 256); }

}  // namespace htgs

using namespace htgs;

extern "C" int htgs_importance_accumulate(int P, int D, int M, int64_t R, int W, int H, const float* background,
                                          const float* means3D, const float* campos, const float* alphas, const int* radii,
                                          const void* geom_buf, const void* binning_buf, const void* image_buf,
                                          const float* dL_dout_color, void* scratch, float* importance, void* stream) {
    const char* who = "htgs_importance_accumulate";
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    if (P <= 0) return 0;
    if (!background || !means3D || !campos || !alphas || !radii || !geom_buf || !image_buf || !dL_dout_color || !scratch || !importance ||
        (R > 0 && !binning_buf)) {
        set_error("%s: null required pointer", who);
        return 1;
    }
    if (D < 0 || D > 3 || M < (D + 1) * (D + 1)) { set_error("%s: bad SH degree %d / M %d", who, D, M); return 1; }
    if ((reinterpret_cast<size_t>(scratch) & 255) || (reinterpret_cast<size_t>(geom_buf) & 255) || (reinterpret_cast<size_t>(image_buf) & 255)) {
        set_error("%s: buffers must be 256-byte aligned", who);
        return 1;
    }
    GeomView g = geom_view(const_cast<void*>(geom_buf), P);
    ImgView img = img_view(const_cast<void*>(image_buf), W, H);
    BinView b = bin_view(const_cast<void*>(binning_buf), R, W, H);
    float* gacc = reinterpret_cast<float*>(scratch);
    HTGS_CUDA_CHECK(cudaMemsetAsync(gacc, 0, (size_t)P * 12 * sizeof(float), st));
    if (R <= 0) return 0;
    // loss = image.sum(): no upstream gradient for depth and alpha (NULL = zero)
    if (launch_blend_backward(P, W, H, img.ranges, b.values_sorted, g, background, alphas, img.n_contrib, dL_dout_color, nullptr,
                              nullptr, gacc, b.reach, img.bin_capacity, (uint32_t)R, st)) return 1;
    importance_accum_kernel<<<(P + 255) / 256, 256, 0, st>>>(P, D, M, means3D, campos, radii, g.clamped, gacc, importance);
    HTGS_LAUNCH_CHECK("importance_accum_kernel");
    return 0;
}

extern "C" size_t htgs_importance_select_workspace_bytes(int P) {
    if (P <= 0) return 256;
    return 2 * select_keys_bytes(P) + 2 * select_vals_bytes(P) + align_up(sort_scratch_bytes(P), 256) + 256;
}

extern "C" int htgs_importance_select(int P, int n_per, const float* importance, int64_t k, void* workspace, uint8_t* mask,
                                      void* stream) {
    const char* who = "htgs_importance_select";
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    if (P <= 0) return 0;
    if (!importance || !workspace || !mask || n_per <= 0) { set_error("%s: bad arguments", who); return 1; }
    if (k < 0 || k > P) { set_error("%s: k=%lld outside 0..P", who, (long long)k); return 1; }
    if (reinterpret_cast<size_t>(workspace) & 255) { set_error("%s: workspace must be 256-byte aligned", who); return 1; }
    char* p = reinterpret_cast<char*>(workspace);
    uint64_t* k0 = reinterpret_cast<uint64_t*>(p); p += select_keys_bytes(P);
    uint64_t* k1 = reinterpret_cast<uint64_t*>(p); p += select_keys_bytes(P);
    uint32_t* v0 = reinterpret_cast<uint32_t*>(p); p += select_vals_bytes(P);
    uint32_t* v1 = reinterpret_cast<uint32_t*>(p); p += select_vals_bytes(P);
    HTGS_CUDA_CHECK(cudaMemsetAsync(mask, 0, (size_t)P, st));
    if (k == 0) return 0;
    importance_keys_kernel<<<(P + 255) / 256, 256, 0, st>>>(P, n_per, importance, k0, v0);
    HTGS_LAUNCH_CHECK("importance_keys_kernel");
    if (sort_pairs(P, 32, k0, v0, k1, v1, p, st)) return 1;  // 4 passes (even): the sorted pairs end in (k0, v0)
    const uint32_t* order = (sort_num_passes(32) & 1) ? v1 : v0;
    mark_first_k_kernel<<<(unsigned)((k + 255) / 256), 256, 0, st>>>(k, order, mask);
    HTGS_LAUNCH_CHECK("mark_first_k_kernel");
    return 0;
}

extern "C" int htgs_transform_points(int P, const float* xyz, const float* transform, float* out, void* stream) {
    if (P <= 0) return 0;
    if (!xyz || !transform || !out) { set_error("htgs_transform_points: null pointer"); return 1; }
    transform_points_kernel<<<(P + 255) / 256, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(P, xyz, transform, out);
    HTGS_LAUNCH_CHECK("transform_points_kernel");
    return 0;
}
