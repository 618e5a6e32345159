// K5: back-to-front gradient of the tile blend (colour + depth + alpha).
// Replaces BACKWARD::render / renderCUDA (cuda_rasterizer/backward.cu:415-601).
//
// The reference issues 10 global float atomics per blended (pixel, Gaussian)
// pair (backward.cu:555-598) and is bound by L2 atomic throughput.  Here the 10
// partial gradients of a Gaussian are
//   1. summed over the 32 pixels of a warp with a 12-shuffle transpose-reduce
//      (each halving step swaps half of the remaining values with the partner
//      lane, so lane pairs end up holding the warp total of one value each),
//   2. parked in a per-warp shared-memory slot (plain store, no atomics),
//   3. summed over the tile's 8 warps when the 128-Gaussian batch is done and
//      sent to HBM as three 128-bit vector reductions per (tile, Gaussian),
// i.e. 3 vector atomics per tile instance instead of 10 scalar ones per pixel
// pair.  Per-pixel arithmetic follows the reference (same thresholds, same
// T = T/(1-alpha) recurrence); only the summation order of the atomics differs.
// The 8x4 reach masks of blend_common.cuh skip Gaussians that cannot touch a
// warp's pixels, and whole batches behind the tile's last contributor.
//
// Accumulator layout gacc[P][12] (zeroed by the caller):
//   [0..3] dL/dmean2D.x, .y, dL/dopacity, dL/ddepth   [4..6] dL/dconic x,y,w
//   [8..10] dL/dcolour rgb                             [7],[11] unused
#include "blend_common.cuh"

namespace htgs {

constexpr int BWD_BATCH = 128;
constexpr int BWD_NV = 10;

// value order inside the kernel: 0 mean.x 1 mean.y 2 opacity 3 depth 4..6 conic x,y,w 7..9 rgb
struct Reduce10 {
    bool b4, b3, b2, b1, writer;
    int id;
    __device__ __forceinline__ void init(int lane) {
        b4 = lane & 16; b3 = lane & 8; b2 = lane & 4; b1 = lane & 2;
        id = (b4 ? 5 : 0) + (b3 ? 3 : 0) + (b2 ? 2 : 0) + (b1 ? 1 : 0);
        const bool valid = b3 ? !b2 : !(b2 && b1);
        writer = valid && !(lane & 1);
    }
    // returns, on writer lanes, the sum over the warp of v[id]
    __device__ __forceinline__ float run(const float (&v)[BWD_NV]) const {
        const unsigned full = 0xffffffffu;
        float a[5], c[3], d[2];
#pragma unroll
        for (int k = 0; k < 5; k++) {
            const float send = b4 ? v[k] : v[5 + k];
            const float keep = b4 ? v[5 + k] : v[k];
            a[k] = keep + __shfl_xor_sync(full, send, 16);
        }
#pragma unroll
        for (int k = 0; k < 3; k++) {
            const float hi = (3 + k < 5) ? a[(3 + k < 5) ? 3 + k : 0] : 0.f;
            const float send = b3 ? a[k] : hi;
            const float keep = b3 ? hi : a[k];
            c[k] = keep + __shfl_xor_sync(full, send, 8);
        }
#pragma unroll
        for (int k = 0; k < 2; k++) {
            const float hi = (k == 0) ? c[2] : 0.f;
            const float send = b2 ? c[k] : hi;
            const float keep = b2 ? hi : c[k];
            d[k] = keep + __shfl_xor_sync(full, send, 4);
        }
        const float send = b1 ? d[0] : d[1];
        const float keep = b1 ? d[1] : d[0];
        float e = keep + __shfl_xor_sync(full, send, 2);
        e += __shfl_xor_sync(full, e, 1);
        return e;
    }
};

__global__ void __launch_bounds__(HTGS_TILE_PIX)
blend_backward_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list, int W, int H,
                      const float* __restrict__ bg, const float2* __restrict__ means2D,
                      const float4* __restrict__ conic_opacity, const float4* __restrict__ feat,
                      const float* __restrict__ alphas, const uint32_t* __restrict__ n_contrib,
                      const float* __restrict__ dL_dpix, const float* __restrict__ dL_ddepth_pix,
                      const float* __restrict__ dL_dalpha_pix, float* __restrict__ gacc) {
    __shared__ float s_stage[BLEND_WARPS][BWD_BATCH][BWD_NV];  // 40 KB
    __shared__ float4 s_co[BWD_BATCH];
    __shared__ float4 s_feat[BWD_BATCH];
    __shared__ float2 s_xy[BWD_BATCH];
    __shared__ uint32_t s_id[BWD_BATCH];
    __shared__ uint8_t s_mask[BWD_BATCH];
    __shared__ uint32_t s_warp_last[BLEND_WARPS];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tile_x0 = blockIdx.x * HTGS_TILE, tile_y0 = blockIdx.y * HTGS_TILE;
    const int px = tile_x0 + warp_block_x(warp) + (lane & 7);
    const int py = tile_y0 + warp_block_y(warp) + (lane >> 3);
    const bool inside = px < W && py < H;
    const float pxf = (float)px, pyf = (float)py;
    const size_t pix = (size_t)py * W + px;
    const size_t plane = (size_t)H * W;

    const uint2 range = ranges[blockIdx.y * gridDim.x + blockIdx.x];
    const int n = (int)(range.y - range.x);

    // backward.cu:463-486
    const float T_final = inside ? 1.0f - alphas[pix] : 0.f;
    float T = T_final;
    const uint32_t last_contributor = inside ? n_contrib[pix] : 0u;
    float dLdp0 = 0.f, dLdp1 = 0.f, dLdp2 = 0.f, dLdd = 0.f, dLda = 0.f;
    if (inside) {
        dLdp0 = dL_dpix[pix];
        dLdp1 = dL_dpix[plane + pix];
        dLdp2 = dL_dpix[2 * plane + pix];
        dLdd = dL_ddepth_pix[pix];
        dLda = dL_dalpha_pix[pix];
    }
    const float bg_dot_dpixel = bg[0] * dLdp0 + bg[1] * dLdp1 + bg[2] * dLdp2;  // backward.cu:575-577
    float acc0 = 0.f, acc1 = 0.f, acc2 = 0.f, accd = 0.f, acca = 0.f;
    float lc0 = 0.f, lc1 = 0.f, lc2 = 0.f, ld = 0.f, last_alpha = 0.f;
    const float ddelx_dx = 0.5f * W, ddely_dy = 0.5f * H;  // backward.cu:490-491

    uint32_t wl = last_contributor;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) wl = max(wl, __shfl_xor_sync(0xffffffffu, wl, o));
    if (lane == 0) s_warp_last[warp] = wl;
    __syncthreads();
    const uint32_t warp_last = wl;
    uint32_t block_last = 0;
#pragma unroll
    for (int w = 0; w < BLEND_WARPS; w++) block_last = max(block_last, s_warp_last[w]);

    Reduce10 red;
    red.init(lane);

    for (int base = 0; base < n; base += BWD_BATCH) {
        const int cnt = min(BWD_BATCH, n - base);
        // list positions (0-based, front to back) covered by this batch: top .. top-cnt+1
        const int top = n - 1 - base;
        if (top - cnt + 1 >= (int)block_last) continue;  // nothing in this batch was blended by any pixel
        __syncthreads();
        if (tid < cnt) {
            const uint32_t g = point_list[range.y - 1 - (uint32_t)(base + tid)];  // back to front, backward.cu:502
            const float2 xy = means2D[g];
            const float4 co = conic_opacity[g];
            s_id[tid] = g;
            s_xy[tid] = xy;
            s_co[tid] = co;
            s_feat[tid] = feat[g];
            s_mask[tid] = (uint8_t)block_reach_mask(xy, co, tile_x0, tile_y0);
        }
        __syncthreads();
        for (int c0 = 0; c0 < cnt; c0 += 32) {
            const int j = c0 + lane;
            const bool mine = j < cnt && ((s_mask[j] >> warp) & 1) && (uint32_t)(top - j) < warp_last;
            uint32_t m = __ballot_sync(0xffffffffu, mine);
            while (m) {
                const int idx = c0 + __ffs(m) - 1;
                m &= m - 1;
                const uint32_t pos0 = (uint32_t)(top - idx);  // == `contributor` after the decrement (backward.cu:517)
                float v[BWD_NV];
#pragma unroll
                for (int k = 0; k < BWD_NV; k++) v[k] = 0.f;
                bool contributed = false;
                if (pos0 < last_contributor) {
                    const float2 xy = s_xy[idx];
                    const float4 co = s_co[idx];
                    const float dx = SB(xy.x, pxf), dy = SB(xy.y, pyf);
                    const float power = pair_power(dx, dy, co.x, co.y, co.z);
                    if (!(power > 0.0f)) {
                        const float G = expf(power);
                        const float alpha = fminf(0.99f, MU(co.w, G));
                        if (!(alpha < 1.0f / 255.0f)) {
                            contributed = true;
                            const float4 f = s_feat[idx];
                            T = __fdiv_rn(T, SB(1.f, alpha));
                            const float dchannel_dcolor = alpha * T;
                            const float one_m_la = 1.f - last_alpha;
                            float dL_dopa;
                            acc0 = last_alpha * lc0 + one_m_la * acc0; lc0 = f.x;
                            dL_dopa = (f.x - acc0) * dLdp0;
                            acc1 = last_alpha * lc1 + one_m_la * acc1; lc1 = f.y;
                            dL_dopa += (f.y - acc1) * dLdp1;
                            acc2 = last_alpha * lc2 + one_m_la * acc2; lc2 = f.z;
                            dL_dopa += (f.z - acc2) * dLdp2;
                            v[7] = dchannel_dcolor * dLdp0;
                            v[8] = dchannel_dcolor * dLdp1;
                            v[9] = dchannel_dcolor * dLdp2;
                            accd = last_alpha * ld + one_m_la * accd; ld = f.w;
                            dL_dopa += (f.w - accd) * dLdd;
                            v[3] = dchannel_dcolor * dLdd;
                            acca = last_alpha + one_m_la * acca;
                            dL_dopa += (1.f - acca) * dLda;
                            dL_dopa *= T;
                            last_alpha = alpha;
                            dL_dopa += (-T_final / (1.f - alpha)) * bg_dot_dpixel;
                            const float dL_dG = co.w * dL_dopa;
                            const float gdx = G * dx, gdy = G * dy;
                            const float dG_ddelx = -gdx * co.x - gdy * co.y;
                            const float dG_ddely = -gdy * co.z - gdx * co.y;
                            v[0] = dL_dG * dG_ddelx * ddelx_dx;
                            v[1] = dL_dG * dG_ddely * ddely_dy;
                            v[4] = -0.5f * gdx * dx * dL_dG;
                            v[5] = -0.5f * gdx * dy * dL_dG;
                            v[6] = -0.5f * gdy * dy * dL_dG;
                            v[2] = G * dL_dopa;
                        }
                    }
                }
                float total = 0.f;
                if (__any_sync(0xffffffffu, contributed)) total = red.run(v);
                if (red.writer) s_stage[warp][idx][red.id] = total;
            }
        }
        __syncthreads();
        if (tid < cnt) {
            float s[BWD_NV];
#pragma unroll
            for (int k = 0; k < BWD_NV; k++) s[k] = 0.f;
            const uint32_t mask = s_mask[tid];
            const uint32_t pos0 = (uint32_t)(top - tid);
#pragma unroll
            for (int w = 0; w < BLEND_WARPS; w++) {
                if (((mask >> w) & 1) && pos0 < s_warp_last[w]) {
#pragma unroll
                    for (int k = 0; k < BWD_NV; k++) s[k] += s_stage[w][tid][k];
                }
            }
            float4* dst = reinterpret_cast<float4*>(gacc + (size_t)s_id[tid] * 12);
            if (s[0] != 0.f || s[1] != 0.f || s[2] != 0.f || s[3] != 0.f) atomicAdd(dst, make_float4(s[0], s[1], s[2], s[3]));
            if (s[4] != 0.f || s[5] != 0.f || s[6] != 0.f) atomicAdd(dst + 1, make_float4(s[4], s[5], s[6], 0.f));
            if (s[7] != 0.f || s[8] != 0.f || s[9] != 0.f) atomicAdd(dst + 2, make_float4(s[7], s[8], s[9], 0.f));
        }
    }
}

int launch_blend_backward(int P, int W, int H, const uint2* ranges, const uint32_t* point_list, const GeomView& g,
                          const float* bg, const float* alphas, const uint32_t* n_contrib, const float* dL_dpix,
                          const float* dL_ddepth, const float* dL_dalpha, float* gacc, cudaStream_t st) {
    (void)P;
    const dim3 grid((W + HTGS_TILE - 1) / HTGS_TILE, (H + HTGS_TILE - 1) / HTGS_TILE);
    blend_backward_kernel<<<grid, HTGS_TILE_PIX, 0, st>>>(ranges, point_list, W, H, bg, g.means2D, g.conic_opacity, g.feat,
                                                          alphas, n_contrib, dL_dpix, dL_ddepth, dL_dalpha, gacc);
    HTGS_LAUNCH_CHECK("blend_backward_kernel");
    return 0;
}

}  // namespace htgs
