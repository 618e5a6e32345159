// "Next" row 1 of SURVEY.md section 8(f): the per-step tail that follows the rasterizer backward in
// HTGaussianTrainer.train_step (trainer/ht3dgs_trainer.py:137-166).
//
//   densify_stats_kernel   max_radii2D[vis] = max(max_radii2D[vis], radii[vis]);
//                          xyz_gradient_accum[vis] += ||viewspace_grad[vis, :2]||; denom[vis] += 1
//                          (ht3dgs_trainer.py:143-148, scene/gaussian_model_ht.py:718-721).  The reference
//                          does this with three boolean-mask indexing ops, each of which launches a
//                          nonzero() and stalls the host; here it is one streaming kernel, no sync.
//   adam_multi_kernel      one launch for all parameter groups of torch.optim.Adam(l, lr=0.0, eps=1e-15)
//                          (scene/gaussian_model_ht.py:275-289): 4 reads + 3 writes of 4 B per element,
//                          HBM-bound.  Arithmetic follows torch's single-tensor Adam (lerp for the first
//                          moment, addcmul for the second, addcdiv for the update).
#include "../../include/htgs.h"
#include "common.cuh"

namespace htgs {

__global__ void densify_stats_kernel(int P, const int* __restrict__ radii, const float* __restrict__ grad_means2D,
                                     float* __restrict__ max_radii2D, float* __restrict__ grad_accum,
                                     float* __restrict__ denom) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const int r = radii[i];
    if (!(r > 0)) return;  // visibility_filter = radii > 0 (scene/gaussian_model_ht.py:892)
    max_radii2D[i] = fmaxf(max_radii2D[i], (float)r);
    const float gx = grad_means2D[3 * (size_t)i], gy = grad_means2D[3 * (size_t)i + 1];
    grad_accum[i] += sqrtf(gx * gx + gy * gy);
    denom[i] += 1.0f;
}

constexpr int ADAM_MAX_GROUPS = HTGS_ADAM_MAX_GROUPS;
constexpr int ADAM_THREADS = 256;
constexpr int ADAM_CHUNK = ADAM_THREADS * 16;  // elements per CTA

struct AdamLaunch {
    htgs_adam_group g[ADAM_MAX_GROUPS];
    float step_size[ADAM_MAX_GROUPS];  // lr / (1 - beta1^t)
    long long chunk_start[ADAM_MAX_GROUPS + 1];
    int n_groups;
};

// omb1 / omb2 are (1 - beta) formed in fp64 on the host, as torch forms its python-float scalars
// (1 - 0.999 in fp32 would be off by 1.3e-5 relative)
__device__ __forceinline__ void adam_one(float& p, float g, float& m, float& v, float omb1, float beta2, float omb2, float eps,
                                         float step_size, float bc2_sqrt) {
    m = m + omb1 * (g - m);                     // exp_avg.lerp_(grad, 1 - beta1)
    v = v * beta2 + omb2 * g * g;               // exp_avg_sq.mul_(beta2).addcmul_(grad, grad, value=1 - beta2)
    const float den = sqrtf(v) / bc2_sqrt + eps;  // (exp_avg_sq.sqrt() / bias_correction2_sqrt).add_(eps)
    p = p - step_size * (m / den);              // param.addcdiv_(exp_avg, denom, value=-step_size)
}

__global__ void __launch_bounds__(ADAM_THREADS)
adam_multi_kernel(const __grid_constant__ AdamLaunch a, float omb1, float beta2, float omb2, float eps, float bc2_sqrt) {
    int gi = 0;
#pragma unroll
    for (int k = 1; k < ADAM_MAX_GROUPS; k++)
        if (k < a.n_groups && (long long)blockIdx.x >= a.chunk_start[k]) gi = k;
    const htgs_adam_group G = a.g[gi];
    const float step_size = a.step_size[gi];
    const long long base = ((long long)blockIdx.x - a.chunk_start[gi]) * ADAM_CHUNK;
    const long long n = G.numel;
    const bool vec = ((reinterpret_cast<size_t>(G.param) | reinterpret_cast<size_t>(G.grad) | reinterpret_cast<size_t>(G.exp_avg) |
                       reinterpret_cast<size_t>(G.exp_avg_sq)) & 15) == 0;
    if (vec && base + ADAM_CHUNK <= n) {
        float4* p4 = reinterpret_cast<float4*>(G.param + base);
        const float4* g4 = reinterpret_cast<const float4*>(G.grad + base);
        float4* m4 = reinterpret_cast<float4*>(G.exp_avg + base);
        float4* v4 = reinterpret_cast<float4*>(G.exp_avg_sq + base);
#pragma unroll
        for (int r = 0; r < ADAM_CHUNK / (4 * ADAM_THREADS); r++) {
            const int i = r * ADAM_THREADS + threadIdx.x;
            float4 p = p4[i], g = g4[i], m = m4[i], v = v4[i];
            adam_one(p.x, g.x, m.x, v.x, omb1, beta2, omb2, eps, step_size, bc2_sqrt);
            adam_one(p.y, g.y, m.y, v.y, omb1, beta2, omb2, eps, step_size, bc2_sqrt);
            adam_one(p.z, g.z, m.z, v.z, omb1, beta2, omb2, eps, step_size, bc2_sqrt);
            adam_one(p.w, g.w, m.w, v.w, omb1, beta2, omb2, eps, step_size, bc2_sqrt);
            p4[i] = p; m4[i] = m; v4[i] = v;
        }
    } else {
        for (long long i = base + threadIdx.x; i < n && i < base + ADAM_CHUNK; i += ADAM_THREADS) {
            float p = G.param[i], m = G.exp_avg[i], v = G.exp_avg_sq[i];
            adam_one(p, G.grad[i], m, v, omb1, beta2, omb2, eps, step_size, bc2_sqrt);
            G.param[i] = p; G.exp_avg[i] = m; G.exp_avg_sq[i] = v;
        }
    }
}

}  // namespace htgs

using namespace htgs;

extern "C" int htgs_densification_stats(int P, const int* radii, const float* grad_means2D, float* max_radii2D,
                                        float* xyz_gradient_accum, float* denom, void* stream) {
    if (P <= 0) return 0;
    if (!radii || !grad_means2D || !max_radii2D || !xyz_gradient_accum || !denom) {
        set_error("htgs_densification_stats: null pointer");
        return 1;
    }
    densify_stats_kernel<<<(P + 255) / 256, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(P, radii, grad_means2D, max_radii2D,
                                                                                             xyz_gradient_accum, denom);
    HTGS_LAUNCH_CHECK("densify_stats_kernel");
    return 0;
}

extern "C" int htgs_adam_step(int n_groups, const htgs_adam_group* groups, float beta1, float beta2, float eps, long long step,
                              void* stream) {
    if (n_groups <= 0) return 0;
    if (n_groups > ADAM_MAX_GROUPS || !groups) { set_error("htgs_adam_step: need 1..%d groups", ADAM_MAX_GROUPS); return 1; }
    if (step < 1) { set_error("htgs_adam_step: step counts from 1"); return 1; }
    AdamLaunch a;
    a.n_groups = 0;
    long long chunks = 0;
    const double bc1 = 1.0 - pow((double)beta1, (double)step);
    const double bc2 = 1.0 - pow((double)beta2, (double)step);
    for (int k = 0; k < n_groups; k++) {
        const htgs_adam_group& g = groups[k];
        if (g.numel <= 0) continue;
        if (!g.param || !g.grad || !g.exp_avg || !g.exp_avg_sq) { set_error("htgs_adam_step: null pointer in group %d", k); return 1; }
        a.g[a.n_groups] = g;
        a.step_size[a.n_groups] = (float)((double)g.lr / bc1);
        a.chunk_start[a.n_groups] = chunks;
        chunks += (g.numel + ADAM_CHUNK - 1) / ADAM_CHUNK;
        a.n_groups++;
    }
    if (a.n_groups == 0) return 0;
    a.chunk_start[a.n_groups] = chunks;
    adam_multi_kernel<<<(unsigned)chunks, ADAM_THREADS, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
        a, (float)(1.0 - (double)beta1), beta2, (float)(1.0 - (double)beta2), eps, (float)sqrt(bc2));
    HTGS_LAUNCH_CHECK("adam_multi_kernel");
    return 0;
}
