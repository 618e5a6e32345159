// "Next" row 2 of SURVEY.md section 8(f): the image loss that produces dL/dcolour for the rasterizer
// backward,  loss = (1 - lambda) * mean|pred - gt| + lambda * (1 - SSIM(pred, gt))
// (trainer/losses.py:98-136; l1_loss :139-140; SSIM with an 11x11 sigma-1.5 Gaussian window, zero
// padding, per-channel, C1 = 0.01^2, C2 = 0.03^2, :147-200; lambda_dssim = 0.2 arguments/__init__.py:134).
//
// The reference evaluates it with five grouped conv2d calls plus a dozen elementwise kernels and lets
// autograd replay them.  Here: one forward kernel (separable 11-tap convolution of the five moments in
// shared memory, SSIM map, deterministic two-stage sums, three derivative maps) and one backward kernel
// (separable convolution of the derivative maps, combined with the L1 sign term).  HBM traffic: forward
// reads 2 images and writes 3 maps, backward reads 3 maps + 2 images and writes 1 image (8*C*H*W*4 B total).
#include "../../include/htgs.h"
#include "common.cuh"
#include <mutex>

namespace htgs {

constexpr int LT = 32;           // output tile edge (per CTA of 256 threads)
constexpr int LH = 5;            // window radius (11 taps)
constexpr int LR = LT + 2 * LH;  // 42: tile + halo
constexpr int PA = LR + 1;       // 43: odd row pitch of the input tiles  -> lane-per-row accesses are conflict-free
constexpr int PH = LT + 1;       // 33: odd row pitch of the row-filtered planes
constexpr int LTHREADS = 256;
constexpr int HS = 8;            // outputs per thread in the horizontal pass (18 inputs held in registers)
constexpr int VS = 4;            // outputs per thread in the vertical pass   (14 inputs held in registers)

// Both passes are register-blocked sliding windows: a thread loads HS+10 (VS+10) inputs once and forms HS (VS)
// outputs from registers, so shared-memory reads per multiply-add drop from 1 to 0.2-0.3 -- the first version of
// these kernels (one output per thread, one LDS per tap) was bound by shared-memory issue, not by HBM.
// Tap order per output is k = 0..10 ascending in both passes.

__constant__ float c_win[11];

static void win_host(float* w) {  // trainer/losses.py:147-151 (gaussian(11, 1.5), normalised)
    double g[11], s = 0.0;
    for (int x = 0; x < 11; x++) { g[x] = exp(-(double)((x - 5) * (x - 5)) / (2.0 * 1.5 * 1.5)); s += g[x]; }
    for (int x = 0; x < 11; x++) w[x] = (float)((float)g[x] / (float)s);
}

template <int N>
__device__ __forceinline__ void window_outputs(const float (&in)[N + 10], float (&out)[N]) {
#pragma unroll
    for (int j = 0; j < N; j++) {
        float acc = 0.f;
#pragma unroll
        for (int k = 0; k < 11; k++) acc = fmaf(c_win[k], in[j + k], acc);
        out[j] = acc;
    }
}

// forward: per (32x32 tile, channel) CTA of 256 threads
__global__ void __launch_bounds__(LTHREADS)
ssim_forward_kernel(int C, int H, int W, const float* __restrict__ pred, const float* __restrict__ gt,
                    float* __restrict__ dmaps /*[3][C][H][W] or null*/, double* __restrict__ partial /*[blocks][2]*/) {
    __shared__ float s_a[LR * PA], s_b[LR * PA];
    __shared__ float s_h[5][LR * PH];
    __shared__ double s_red[2][LTHREADS / 32];
    const int c = blockIdx.z, x0 = blockIdx.x * LT, y0 = blockIdx.y * LT;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const size_t plane = (size_t)H * W;
    const float* A = pred + (size_t)c * plane;
    const float* B = gt + (size_t)c * plane;
    {   // halo tile load, fully unrolled so that all global loads of a thread are in flight together
        constexpr int NIT = (LR * LR + LTHREADS - 1) / LTHREADS;
        float va[NIT], vb[NIT];
#pragma unroll
        for (int it = 0; it < NIT; it++) {
            const int i = tid + it * LTHREADS, r = i / LR, q = i - r * LR;
            const int y = y0 + r - LH, x = x0 + q - LH;
            const bool in = i < LR * LR && y >= 0 && y < H && x >= 0 && x < W;  // zero padding of F.conv2d(padding=5)
            va[it] = in ? __ldg(A + (size_t)y * W + x) : 0.f;
            vb[it] = in ? __ldg(B + (size_t)y * W + x) : 0.f;
        }
#pragma unroll
        for (int it = 0; it < NIT; it++) {
            const int i = tid + it * LTHREADS, r = i / LR, q = i - r * LR;
            if (i < LR * LR) { s_a[r * PA + q] = va[it]; s_b[r * PA + q] = vb[it]; }
        }
    }
    __syncthreads();
    {   // horizontal pass: 42 halo rows x four 8-column strips = 168 threads; consecutive threads take
        // consecutive rows, which the odd row pitch spreads over the banks
        const int strip = tid / LR, r = tid - strip * LR;
        if (tid < LR * (LT / HS)) {
            float a[HS + 10], b[HS + 10], t[HS + 10], o[HS];
#pragma unroll
            for (int k = 0; k < HS + 10; k++) { a[k] = s_a[r * PA + strip * HS + k]; b[k] = s_b[r * PA + strip * HS + k]; }
            float* dst = &s_h[0][r * PH + strip * HS];
            window_outputs<HS>(a, o);
#pragma unroll
            for (int j = 0; j < HS; j++) dst[j] = o[j];
            window_outputs<HS>(b, o);
#pragma unroll
            for (int j = 0; j < HS; j++) dst[LR * PH + j] = o[j];
#pragma unroll
            for (int k = 0; k < HS + 10; k++) t[k] = a[k] * a[k];
            window_outputs<HS>(t, o);
#pragma unroll
            for (int j = 0; j < HS; j++) dst[2 * LR * PH + j] = o[j];
#pragma unroll
            for (int k = 0; k < HS + 10; k++) t[k] = b[k] * b[k];
            window_outputs<HS>(t, o);
#pragma unroll
            for (int j = 0; j < HS; j++) dst[3 * LR * PH + j] = o[j];
#pragma unroll
            for (int k = 0; k < HS + 10; k++) t[k] = a[k] * b[k];
            window_outputs<HS>(t, o);
#pragma unroll
            for (int j = 0; j < HS; j++) dst[4 * LR * PH + j] = o[j];
        }
    }
    __syncthreads();
    // vertical pass: lane <-> column, warp <-> 4-row strip
    const int tx = lane, ty0 = warp * VS, x = x0 + tx;
    float v[5][VS];
#pragma unroll
    for (int p = 0; p < 5; p++) {
        float col[VS + 10];
#pragma unroll
        for (int k = 0; k < VS + 10; k++) col[k] = s_h[p][(ty0 + k) * PH + tx];
        window_outputs<VS>(col, v[p]);
    }
    double l1 = 0.0, ss = 0.0;
#pragma unroll
    for (int j = 0; j < VS; j++) {
        const int y = y0 + ty0 + j;
        if (x < W && y < H) {
            const float mu1 = v[0][j], mu2 = v[1][j], e11 = v[2][j], e22 = v[3][j], e12 = v[4][j];
            const float C1 = 0.01f * 0.01f, C2 = 0.03f * 0.03f;
            const float mu1s = mu1 * mu1, mu2s = mu2 * mu2, mu12 = mu1 * mu2;
            const float s1 = e11 - mu1s, s2 = e22 - mu2s, s12 = e12 - mu12;
            const float An = 2.f * mu12 + C1, Bn = 2.f * s12 + C2, Cd = mu1s + mu2s + C1, Dd = s1 + s2 + C2;
            const float m = (An * Bn) / (Cd * Dd);
            ss += (double)m;
            const float a = s_a[(ty0 + j + LH) * PA + tx + LH], b = s_b[(ty0 + j + LH) * PA + tx + LH];
            l1 += (double)fabsf(a - b);
            if (dmaps) {
                const float inv_cd = 1.f / (Cd * Dd);
                const float dm_dmu1 = 2.f * mu2 * Bn * inv_cd - 2.f * mu1 * m / Cd + 2.f * mu1 * m / Dd - 2.f * mu2 * An * inv_cd;
                const float dm_de11 = -m / Dd;
                const float dm_de12 = 2.f * An * inv_cd;
                const size_t o = (size_t)c * plane + (size_t)y * W + x;
                const size_t vol = (size_t)C * plane;
                dmaps[o] = dm_dmu1; dmaps[vol + o] = dm_de11; dmaps[2 * vol + o] = dm_de12;
            }
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { l1 += __shfl_xor_sync(0xffffffffu, l1, o); ss += __shfl_xor_sync(0xffffffffu, ss, o); }
    if (lane == 0) { s_red[0][warp] = l1; s_red[1][warp] = ss; }
    __syncthreads();
    if (tid == 0) {
        double a = 0.0, b = 0.0;
        for (int w = 0; w < LTHREADS / 32; w++) { a += s_red[0][w]; b += s_red[1][w]; }
        const size_t blk = ((size_t)blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
        partial[2 * blk] = a; partial[2 * blk + 1] = b;
    }
}

// second stage: fixed-order sum of the per-CTA partials (deterministic), writes loss, l1, dssim
__global__ void __launch_bounds__(1024) ssim_finish_kernel(int nblk, const double* __restrict__ partial, double n_elem,
                                                           float lambda, float* __restrict__ out3) {
    __shared__ double s[2][1024];
    double a = 0.0, b = 0.0;
    for (int i = threadIdx.x; i < nblk; i += 1024) { a += partial[2 * i]; b += partial[2 * i + 1]; }
    s[0][threadIdx.x] = a; s[1][threadIdx.x] = b;
    __syncthreads();
    for (int o = 512; o > 0; o >>= 1) {
        if ((int)threadIdx.x < o) { s[0][threadIdx.x] += s[0][threadIdx.x + o]; s[1][threadIdx.x] += s[1][threadIdx.x + o]; }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        const float l1 = (float)(s[0][0] / n_elem), dssim = (float)(1.0 - s[1][0] / n_elem);
        out3[0] = (1.f - lambda) * l1 + lambda * dssim;  // trainer/losses.py:107,123
        out3[1] = l1;
        out3[2] = dssim;
    }
}

__global__ void __launch_bounds__(LTHREADS)
ssim_backward_kernel(int C, int H, int W, const float* __restrict__ pred, const float* __restrict__ gt,
                     const float* __restrict__ dmaps, float lambda, float inv_n, const float* __restrict__ grad_loss,
                     float* __restrict__ dL_dpred) {
    __shared__ float s_m[3][LR * PA];
    __shared__ float s_h[3][LR * PH];
    const int c = blockIdx.z, x0 = blockIdx.x * LT, y0 = blockIdx.y * LT;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const size_t plane = (size_t)H * W, vol = (size_t)C * plane;
    {
        constexpr int NIT = (LR * LR + LTHREADS - 1) / LTHREADS;
        float vm[NIT][3];
#pragma unroll
        for (int it = 0; it < NIT; it++) {
            const int i = tid + it * LTHREADS, r = i / LR, q = i - r * LR;
            const int y = y0 + r - LH, x = x0 + q - LH;
            const bool in = i < LR * LR && y >= 0 && y < H && x >= 0 && x < W;
            const size_t o = (size_t)c * plane + (size_t)y * W + x;
#pragma unroll
            for (int k = 0; k < 3; k++) vm[it][k] = in ? __ldg(dmaps + k * vol + o) : 0.f;
        }
#pragma unroll
        for (int it = 0; it < NIT; it++) {
            const int i = tid + it * LTHREADS, r = i / LR, q = i - r * LR;
            if (i < LR * LR) {
#pragma unroll
                for (int k = 0; k < 3; k++) s_m[k][r * PA + q] = vm[it][k];
            }
        }
    }
    __syncthreads();
    {
        const int strip = tid / LR, r = tid - strip * LR;
        if (tid < LR * (LT / HS)) {
#pragma unroll
            for (int p = 0; p < 3; p++) {
                float a[HS + 10], o[HS];
#pragma unroll
                for (int k = 0; k < HS + 10; k++) a[k] = s_m[p][r * PA + strip * HS + k];
                window_outputs<HS>(a, o);
#pragma unroll
                for (int j = 0; j < HS; j++) s_h[p][r * PH + strip * HS + j] = o[j];
            }
        }
    }
    __syncthreads();
    const int tx = lane, ty0 = warp * VS, x = x0 + tx;
    float v[3][VS];
#pragma unroll
    for (int p = 0; p < 3; p++) {
        float col[VS + 10];
#pragma unroll
        for (int k = 0; k < VS + 10; k++) col[k] = s_h[p][(ty0 + k) * PH + tx];
        window_outputs<VS>(col, v[p]);
    }
    const float g = grad_loss ? grad_loss[0] : 1.f;
#pragma unroll
    for (int j = 0; j < VS; j++) {
        const int y = y0 + ty0 + j;
        if (x < W && y < H) {
            const size_t o = (size_t)c * plane + (size_t)y * W + x;
            const float a = pred[o], b = gt[o];
            const float dssim_map = v[0][j] + 2.f * a * v[1][j] + b * v[2][j];  // d(sum of ssim_map)/d pred
            const float sgn = (a > b) ? 1.f : ((a < b) ? -1.f : 0.f);           // d|a-b|/da, 0 at equality like torch.abs
            dL_dpred[o] = g * inv_n * ((1.f - lambda) * sgn - lambda * dssim_map);
        }
    }
}

// The 11 window taps live in __constant__ memory and are uploaded once per device, under a lock (host threads may
// race to the first call).  The upload is a blocking cudaMemcpyToSymbol: the first htgs_l1_ssim_* call on a device
// must therefore come outside any stream capture (documented in include/htgs.h).
static bool g_win_set[64] = {};
static std::mutex g_win_mutex;
static int ensure_window() {
    int dev = 0;
    HTGS_CUDA_CHECK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) { set_error("htgs_l1_ssim: device ordinal %d not supported (max 63)", dev); return 1; }
    std::lock_guard<std::mutex> lock(g_win_mutex);
    if (!g_win_set[dev]) {
        float w[11];
        win_host(w);
        HTGS_CUDA_CHECK(cudaMemcpyToSymbol(c_win, w, sizeof(w)));
        g_win_set[dev] = true;
    }
    return 0;
}

}  // namespace htgs

using namespace htgs;

extern "C" size_t htgs_l1_ssim_scratch_bytes(int C, int H, int W) {
    const size_t blocks = (size_t)((W + LT - 1) / LT) * ((H + LT - 1) / LT) * (size_t)(C > 0 ? C : 0);
    return align_up(blocks * 2 * sizeof(double), 256) + 256;
}

extern "C" int htgs_l1_ssim_forward(int C, int H, int W, const float* pred, const float* gt, float lambda_dssim,
                                    float* dmaps, void* scratch, float* out3, void* stream) {
    if (C <= 0 || H <= 0 || W <= 0) { set_error("htgs_l1_ssim_forward: bad shape"); return 1; }
    if (!pred || !gt || !scratch || !out3) { set_error("htgs_l1_ssim_forward: null pointer"); return 1; }
    if (ensure_window()) return 1;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    const dim3 grid((W + LT - 1) / LT, (H + LT - 1) / LT, C), block(LTHREADS);
    double* partial = reinterpret_cast<double*>(scratch);
    ssim_forward_kernel<<<grid, block, 0, st>>>(C, H, W, pred, gt, dmaps, partial);
    HTGS_LAUNCH_CHECK("ssim_forward_kernel");
    ssim_finish_kernel<<<1, 1024, 0, st>>>((int)(grid.x * grid.y * grid.z), partial, (double)C * H * W, lambda_dssim, out3);
    HTGS_LAUNCH_CHECK("ssim_finish_kernel");
    return 0;
}

extern "C" int htgs_l1_ssim_backward(int C, int H, int W, const float* pred, const float* gt, const float* dmaps,
                                     float lambda_dssim, const float* grad_loss, float* dL_dpred, void* stream) {
    if (C <= 0 || H <= 0 || W <= 0) { set_error("htgs_l1_ssim_backward: bad shape"); return 1; }
    if (!pred || !gt || !dmaps || !dL_dpred) { set_error("htgs_l1_ssim_backward: null pointer"); return 1; }
    if (ensure_window()) return 1;
    const dim3 grid((W + LT - 1) / LT, (H + LT - 1) / LT, C), block(LTHREADS);
    ssim_backward_kernel<<<grid, block, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
        C, H, W, pred, gt, dmaps, lambda_dssim, (float)(1.0 / ((double)C * H * W)), grad_loss, dL_dpred);
    HTGS_LAUNCH_CHECK("ssim_backward_kernel");
    return 0;
}
