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

namespace htgs {

constexpr int LT = 16;           // output tile edge
constexpr int LH = 5;            // window radius (11 taps)
constexpr int LR = LT + 2 * LH;  // 26: tile + halo

__constant__ float c_win[11];

static void win_host(float* w) {  // trainer/losses.py:147-151 (gaussian(11, 1.5), normalised)
    double g[11], s = 0.0;
    for (int x = 0; x < 11; x++) { g[x] = exp(-(double)((x - 5) * (x - 5)) / (2.0 * 1.5 * 1.5)); s += g[x]; }
    for (int x = 0; x < 11; x++) w[x] = (float)((float)g[x] / (float)s);
}

// forward: per (tile, channel) CTA of 256 threads
__global__ void __launch_bounds__(LT* LT)
ssim_forward_kernel(int C, int H, int W, const float* __restrict__ pred, const float* __restrict__ gt,
                    float* __restrict__ dmaps /*[3][C][H][W] or null*/, double* __restrict__ partial /*[blocks][2]*/) {
    __shared__ float s_a[LR][LR + 1], s_b[LR][LR + 1];
    __shared__ float s_h[5][LR][LT];
    __shared__ double s_red[2][LT * LT / 32];
    const int c = blockIdx.z, x0 = blockIdx.x * LT, y0 = blockIdx.y * LT;
    const int tid = threadIdx.y * LT + threadIdx.x;
    const size_t plane = (size_t)H * W;
    const float* A = pred + (size_t)c * plane;
    const float* B = gt + (size_t)c * plane;
    for (int i = tid; i < LR * LR; i += LT * LT) {
        const int r = i / LR, q = i - r * LR;
        const int y = y0 + r - LH, x = x0 + q - LH;
        const bool in = y >= 0 && y < H && x >= 0 && x < W;  // zero padding of F.conv2d(padding=5)
        s_a[r][q] = in ? A[(size_t)y * W + x] : 0.f;
        s_b[r][q] = in ? B[(size_t)y * W + x] : 0.f;
    }
    __syncthreads();
    for (int i = tid; i < LR * LT; i += LT * LT) {  // horizontal pass over all 26 rows
        const int r = i / LT, q = i - r * LT;
        float m1 = 0.f, m2 = 0.f, e11 = 0.f, e22 = 0.f, e12 = 0.f;
#pragma unroll
        for (int k = 0; k < 11; k++) {
            const float w = c_win[k], a = s_a[r][q + k], b = s_b[r][q + k];
            m1 = fmaf(w, a, m1); m2 = fmaf(w, b, m2);
            e11 = fmaf(w, a * a, e11); e22 = fmaf(w, b * b, e22); e12 = fmaf(w, a * b, e12);
        }
        s_h[0][r][q] = m1; s_h[1][r][q] = m2; s_h[2][r][q] = e11; s_h[3][r][q] = e22; s_h[4][r][q] = e12;
    }
    __syncthreads();
    const int tx = threadIdx.x, ty = threadIdx.y, x = x0 + tx, y = y0 + ty;
    double l1 = 0.0, ss = 0.0;
    if (x < W && y < H) {
        float mu1 = 0.f, mu2 = 0.f, e11 = 0.f, e22 = 0.f, e12 = 0.f;
#pragma unroll
        for (int k = 0; k < 11; k++) {
            const float w = c_win[k];
            mu1 = fmaf(w, s_h[0][ty + k][tx], mu1); mu2 = fmaf(w, s_h[1][ty + k][tx], mu2);
            e11 = fmaf(w, s_h[2][ty + k][tx], e11); e22 = fmaf(w, s_h[3][ty + k][tx], e22);
            e12 = fmaf(w, s_h[4][ty + k][tx], e12);
        }
        const float C1 = 0.01f * 0.01f, C2 = 0.03f * 0.03f;
        const float mu1s = mu1 * mu1, mu2s = mu2 * mu2, mu12 = mu1 * mu2;
        const float s1 = e11 - mu1s, s2 = e22 - mu2s, s12 = e12 - mu12;
        const float An = 2.f * mu12 + C1, Bn = 2.f * s12 + C2, Cd = mu1s + mu2s + C1, Dd = s1 + s2 + C2;
        const float m = (An * Bn) / (Cd * Dd);
        ss = (double)m;
        const float a = s_a[ty + LH][tx + LH], b = s_b[ty + LH][tx + LH];
        l1 = (double)fabsf(a - b);
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
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { l1 += __shfl_xor_sync(0xffffffffu, l1, o); ss += __shfl_xor_sync(0xffffffffu, ss, o); }
    if ((tid & 31) == 0) { s_red[0][tid >> 5] = l1; s_red[1][tid >> 5] = ss; }
    __syncthreads();
    if (tid == 0) {
        double a = 0.0, b = 0.0;
        for (int w = 0; w < LT * LT / 32; w++) { a += s_red[0][w]; b += s_red[1][w]; }
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

__global__ void __launch_bounds__(LT* LT)
ssim_backward_kernel(int C, int H, int W, const float* __restrict__ pred, const float* __restrict__ gt,
                     const float* __restrict__ dmaps, float lambda, float inv_n, const float* __restrict__ grad_loss,
                     float* __restrict__ dL_dpred) {
    __shared__ float s_m[3][LR][LR + 1];
    __shared__ float s_h[3][LR][LT];
    const int c = blockIdx.z, x0 = blockIdx.x * LT, y0 = blockIdx.y * LT;
    const int tid = threadIdx.y * LT + threadIdx.x;
    const size_t plane = (size_t)H * W, vol = (size_t)C * plane;
    for (int i = tid; i < LR * LR; i += LT * LT) {
        const int r = i / LR, q = i - r * LR;
        const int y = y0 + r - LH, x = x0 + q - LH;
        const bool in = y >= 0 && y < H && x >= 0 && x < W;
        const size_t o = (size_t)c * plane + (size_t)y * W + x;
#pragma unroll
        for (int k = 0; k < 3; k++) s_m[k][r][q] = in ? dmaps[k * vol + o] : 0.f;
    }
    __syncthreads();
    for (int i = tid; i < LR * LT; i += LT * LT) {
        const int r = i / LT, q = i - r * LT;
        float h0 = 0.f, h1 = 0.f, h2 = 0.f;
#pragma unroll
        for (int k = 0; k < 11; k++) {
            const float w = c_win[k];
            h0 = fmaf(w, s_m[0][r][q + k], h0); h1 = fmaf(w, s_m[1][r][q + k], h1); h2 = fmaf(w, s_m[2][r][q + k], h2);
        }
        s_h[0][r][q] = h0; s_h[1][r][q] = h1; s_h[2][r][q] = h2;
    }
    __syncthreads();
    const int tx = threadIdx.x, ty = threadIdx.y, x = x0 + tx, y = y0 + ty;
    if (x >= W || y >= H) return;
    float v0 = 0.f, v1 = 0.f, v2 = 0.f;
#pragma unroll
    for (int k = 0; k < 11; k++) {
        const float w = c_win[k];
        v0 = fmaf(w, s_h[0][ty + k][tx], v0); v1 = fmaf(w, s_h[1][ty + k][tx], v1); v2 = fmaf(w, s_h[2][ty + k][tx], v2);
    }
    const size_t o = (size_t)c * plane + (size_t)y * W + x;
    const float a = pred[o], b = gt[o];
    const float dssim_map = v0 + 2.f * a * v1 + b * v2;      // d(sum of ssim_map)/d pred
    const float sgn = (a > b) ? 1.f : ((a < b) ? -1.f : 0.f);  // d|a-b|/da, 0 at equality like torch.abs
    const float g = grad_loss ? grad_loss[0] : 1.f;
    dL_dpred[o] = g * inv_n * ((1.f - lambda) * sgn - lambda * dssim_map);
}

static bool g_win_set[64] = {};
static int ensure_window() {
    int dev = 0;
    HTGS_CUDA_CHECK(cudaGetDevice(&dev));
    if (!g_win_set[dev & 63]) {
        float w[11];
        win_host(w);
        HTGS_CUDA_CHECK(cudaMemcpyToSymbol(c_win, w, sizeof(w)));
        g_win_set[dev & 63] = true;
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
    const dim3 grid((W + LT - 1) / LT, (H + LT - 1) / LT, C), block(LT, LT);
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
    const dim3 grid((W + LT - 1) / LT, (H + LT - 1) / LT, C), block(LT, LT);
    ssim_backward_kernel<<<grid, block, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
        C, H, W, pred, gt, dmaps, lambda_dssim, (float)(1.0 / ((double)C * H * W)), grad_loss, dL_dpred);
    HTGS_LAUNCH_CHECK("ssim_backward_kernel");
    return 0;
}
