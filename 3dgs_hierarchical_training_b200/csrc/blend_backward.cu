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
// Instruction economy (the kernel is FP32-issue bound): a lane does not form the five geometric
// gradients itself.  With w = G * dL/dalpha it only accumulates the moments  w, w*dx, w*dy,
// w*dx*dx, w*dx*dy, w*dy*dy ; the conic/opacity factors, which are per-Gaussian constants, are
// applied once per (tile, Gaussian) when the batch is flushed.  Likewise the per-channel
// "accumulated colour behind me" recurrences of backward.cu:547-567 collapse into one scalar
// recurrence of their dot product with the pixel's upstream gradient (they only ever enter
// dL/dalpha through that dot product).  Same sums, re-associated; thresholds, alpha and the
// T = T/(1-alpha) chain are untouched.
//
// Accumulator layout gacc[P][12] (zeroed by the caller):
//   [0..3] dL/dmean2D.x, .y, dL/dopacity, dL/ddepth   [4..6] dL/dconic x,y,w
//   [8..10] dL/dcolour rgb                             [7],[11] unused
#include "blend_common.cuh"

namespace htgs {

constexpr int BWD_BATCH = 128;  // 96 / 112 / 64 per batch and 5-6 CTAs per SM (48 / 40 registers) all measured slower
constexpr int BWD_NV = 10;

// value order inside the kernel: 0 w 1 w*dx 2 w*dy 3 depth-grad 4..6 w*dx*dx, w*dx*dy, w*dy*dy 7..9 rgb-grad
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
        float ak[5], ar[5];
#pragma unroll
        for (int k = 0; k < 5; k++) {
            const float send = b4 ? v[k] : v[5 + k];
            ak[k] = b4 ? v[5 + k] : v[k];
            ar[k] = __shfl_xor_sync(full, send, 16);
        }
        up2(add2(pk2(ak[0], ak[1]), pk2(ar[0], ar[1])), a[0], a[1]);  // packed adds: one issue slot per pair
        up2(add2(pk2(ak[2], ak[3]), pk2(ar[2], ar[3])), a[2], a[3]);
        a[4] = ak[4] + ar[4];
        float ck[3], cr[3];
#pragma unroll
        for (int k = 0; k < 3; k++) {
            const float hi = (3 + k < 5) ? a[(3 + k < 5) ? 3 + k : 0] : 0.f;
            const float send = b3 ? a[k] : hi;
            ck[k] = b3 ? hi : a[k];
            cr[k] = __shfl_xor_sync(full, send, 8);
        }
        up2(add2(pk2(ck[0], ck[1]), pk2(cr[0], cr[1])), c[0], c[1]);
        c[2] = ck[2] + cr[2];
        float dk[2], dr[2];
#pragma unroll
        for (int k = 0; k < 2; k++) {
            const float hi = (k == 0) ? c[2] : 0.f;
            const float send = b2 ? c[k] : hi;
            dk[k] = b2 ? hi : c[k];
            dr[k] = __shfl_xor_sync(full, send, 4);
        }
        up2(add2(pk2(dk[0], dk[1]), pk2(dr[0], dr[1])), d[0], d[1]);
        const float send = b1 ? d[0] : d[1];
        const float keep = b1 ? d[1] : d[0];
        float e = keep + __shfl_xor_sync(full, send, 2);
        e += __shfl_xor_sync(full, e, 1);
        return e;
    }
};

// 4 CTAs per SM (64 registers): measured 1.12 ms against 1.27 ms at 3 CTAs (80 registers, 4 instructions fewer per visit)
constexpr int BWD_MIN_BLOCKS = 4;
__global__ void __launch_bounds__(HTGS_TILE_PIX, BWD_MIN_BLOCKS)
blend_backward_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list, int W, int H,
                      const float* __restrict__ bg, const float2* __restrict__ means2D,
                      const float4* __restrict__ conic_opacity, const float4* __restrict__ feat,
                      const float* __restrict__ alphas, const uint32_t* __restrict__ n_contrib,
                      const float* __restrict__ dL_dpix, const float* __restrict__ dL_ddepth_pix,
                      const float* __restrict__ dL_dalpha_pix, float* __restrict__ gacc) {
    __shared__ float s_stage[BLEND_WARPS][BWD_BATCH][BWD_NV];  // 40 KB
    __shared__ StagedGaussian s_g[BWD_BATCH];
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
    float d0 = 0.f, d1 = 0.f, d2 = 0.f, dd = 0.f, da = 0.f;
    if (inside) {
        d0 = dL_dpix[pix];
        d1 = dL_dpix[plane + pix];
        d2 = dL_dpix[2 * plane + pix];
        dd = dL_ddepth_pix[pix];
        da = dL_dalpha_pix[pix];
    }
    // -(T_final * bg . dL/dpixel): numerator of the background term, backward.cu:575-578
    const float bg_k = -T_final * (bg[0] * d0 + bg[1] * d1 + bg[2] * d2);
    const bool warp_has_bg = __any_sync(0xffffffffu, bg_k != 0.f);
    // A = sum_ch accum_rec[ch]*dL/dch (colour, depth, alpha channels); Cprev = the same dot product
    // for the previously visited Gaussian's own colour/depth/1
    float A = 0.f, Cprev = 0.f, last_alpha = 0.f, one_m_last = 1.f;  // one_m_last = 1 - last_alpha, carried instead of recomputed

    uint32_t wl = last_contributor;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) wl = max(wl, __shfl_xor_sync(0xffffffffu, wl, o));
    if (lane == 0) s_warp_last[warp] = wl;
    __syncthreads();
    const uint32_t warp_last = wl;
    uint32_t block_last = 0;
#pragma unroll
    for (int w = 0; w < BLEND_WARPS; w++) block_last = max(block_last, s_warp_last[w]);

    const uint64_t pxy = pk2(pxf, pyf), d0_d1 = pk2(d0, d1), d2_dd = pk2(d2, dd), da_0 = pk2(da, 0.f);
    Reduce10 red;
    red.init(lane);
    const uint32_t sg_base = smem_u32(s_g);
    const uint32_t stage_base = smem_u32(&s_stage[warp][0][0]) + (uint32_t)red.id * 4u;

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
            stage_gaussian(&s_g[tid], xy, co, feat[g]);
            s_mask[tid] = (uint8_t)block_reach_mask(xy, co, tile_x0, tile_y0);
        }
        __syncthreads();
        for (int c0 = 0; c0 < cnt; c0 += 32) {
            const int j = c0 + lane;
            const bool mine = j < cnt && ((s_mask[j] >> warp) & 1) && (uint32_t)(top - j) < warp_last;
            uint32_t m = __ballot_sync(0xffffffffu, mine);
            const uint32_t chunk_addr = sg_base + (uint32_t)c0 * STAGED_BYTES;
            const uint32_t chunk_stage = stage_base + (uint32_t)c0 * (BWD_NV * 4u);
            const uint32_t chunk_top = (uint32_t)(top - c0);
            while (m) {
                const int b = __ffs(m) - 1;
                m &= m - 1;
                const uint32_t a = chunk_addr + (uint32_t)b * STAGED_BYTES;
                const float4 q0 = lds128(a);       // mean.xy, conic.x, conic.z
                const float4 q1 = lds128(a + 16);  // conic.y, opacity, r, g
                const uint64_t dxy = sub2(pk2(q0.x, q0.y), pxy);   // (dx, dy)
                float dx, dy, tx, ty, my, mc;
                up2(dxy, dx, dy);
                up2(mul2(dxy, pk2(q0.z, q0.w)), tx, ty);                  // (dx*cx, dy*cz)
                up2(mul2(pk2(dy, dy), pk2(ty, MU(dx, q1.x))), my, mc);    // (dy*(dy*cz), dy*(dx*cy))
                const float power = FM(FM(dx, tx, my), -0.5f, -mc);       // pair_power()
                const float G = expf(power);
                const float alpha = fminf(0.99f, MU(q1.y, G));
                // `contributor` after its decrement is chunk_top - b (backward.cu:517-519)
                const bool live = (chunk_top - (uint32_t)b) < last_contributor && !(power > 0.0f) && !(alpha < 1.0f / 255.0f);
                // Everything a lane contributes is proportional to w = G*dL/dalpha or to aT = alpha*T: lanes that do
                // not blend this Gaussian carry w = aT = 0 and the ten products are formed once, outside the branch
                // (zeroing two scalars instead of ten values).
                float w = 0.f, aT = 0.f;
                if (live) {
                    const float2 q2 = lds64(a + 32);  // b, depth
                    // T = T / (1 - alpha), backward.cu:534.  One Newton step on the hardware reciprocal
                    // (<= 1 ulp from div.rn, 4 issue slots instead of 10); also feeds the background term.
                    const float one_m_a = SB(1.f, alpha);
                    float rinv;
                    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(rinv) : "f"(one_m_a));
                    rinv = fmaf(fmaf(-one_m_a, rinv, 1.f), rinv, rinv);
                    T = T * rinv;
                    aT = alpha * T;
                    // Cd = r*d0 + g*d1 + b*d2 + depth*dd + da as two packed chains
                    float c_lo, c_hi;
                    up2(fma2(pk2(q2.x, q2.y), d2_dd, fma2(pk2(q1.z, q1.w), d0_d1, da_0)), c_lo, c_hi);
                    const float Cd = c_lo + c_hi;
                    A = fmaf(last_alpha, Cprev, one_m_last * A);
                    Cprev = Cd;
                    last_alpha = alpha;
                    one_m_last = one_m_a;
                    float dL_dopa = (Cd - A) * T;
                    if (warp_has_bg) dL_dopa = fmaf(bg_k, rinv, dL_dopa);
                    w = G * dL_dopa;
                }
                float total = 0.f;
                if (__any_sync(0xffffffffu, live)) {
                    float v[BWD_NV];
                    const uint64_t wd = mul2(pk2(w, w), dxy);  // (w*dx, w*dy)
                    float wdx, wdy;
                    up2(wd, wdx, wdy);
                    v[0] = w;          // -> dL/dopacity, and scaled by opacity the common factor below
                    v[1] = wdx;        // first moments  -> dL/dmean2D
                    v[2] = wdy;
                    up2(mul2(wd, dxy), v[4], v[6]);  // second moments w*dx*dx, w*dy*dy -> dL/dconic
                    v[5] = wdx * dy;
                    const uint64_t aTaT = pk2(aT, aT);
                    up2(mul2(aTaT, d0_d1), v[7], v[8]);  // dL/dcolour r, g
                    up2(mul2(aTaT, d2_dd), v[9], v[3]);  // dL/dcolour b, dL/ddepth
                    total = red.run(v);
                }
                if (red.writer)
                    asm volatile("st.shared.f32 [%0], %1;" ::"r"(chunk_stage + (uint32_t)b * (BWD_NV * 4u)), "f"(total) : "memory");
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
            // moments -> gradients (backward.cu:582-598), per-Gaussian constants applied once here
            const float4 q0 = s_g[tid].q0;
            const float4 q1 = s_g[tid].q1;
            const float cx = q0.z, cz = q0.w, cy = q1.x, op = q1.y;
            const float ddelx_dx = 0.5f * W, ddely_dy = 0.5f * H;  // backward.cu:490-491
            const float gmx = -op * (cx * s[1] + cy * s[2]) * ddelx_dx;
            const float gmy = -op * (cz * s[2] + cy * s[1]) * ddely_dy;
            const float gcx = -0.5f * op * s[4], gcy = -0.5f * op * s[5], gcw = -0.5f * op * s[6];
            float4* dst = reinterpret_cast<float4*>(gacc + (size_t)s_id[tid] * 12);
            if (s[0] != 0.f || s[1] != 0.f || s[2] != 0.f || s[3] != 0.f) atomicAdd(dst, make_float4(gmx, gmy, s[0], s[3]));
            if (s[4] != 0.f || s[5] != 0.f || s[6] != 0.f) atomicAdd(dst + 1, make_float4(gcx, gcy, gcw, 0.f));
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
