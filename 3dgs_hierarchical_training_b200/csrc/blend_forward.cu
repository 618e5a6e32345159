// K4: front-to-back alpha blend of one 16x16 tile per CTA.
// Replaces FORWARD::render / renderCUDA (cuda_rasterizer/forward.cu:261-381).
//
// Bound: FP32 issue (ncu r01: issue slots 80% busy, DRAM 1.5%); the levers are testing
// fewer pairs and spending fewer instructions per test.  Gaussians are staged
// in shared memory in batches of 256 together with colour+depth (the reference
// re-fetches those from global memory per blended pair, forward.cu:359-361) and
// an 8-bit reach mask (blend_common.cuh); each warp walks only the Gaussians
// that can reach its 8x4 pixel block and leaves as soon as its 32 pixels are
// saturated.  Per-pixel arithmetic is the reference's, operation for operation,
// so colour/depth/alpha/n_contrib are bit-identical.
#include "blend_common.cuh"

namespace htgs {

constexpr int FWD_BATCH = 256;

// STATS=true is a measurement variant: same traversal, no image writes; it adds to stats[0] the
// pair tests the reference's one-thread-per-pixel loop would execute (list entries walked
// until the pixel saturates), to stats[1] the blended pairs, to stats[2] the pair tests this
// kernel executes after reach-mask culling.
// 6 CTAs per SM (40 registers, no spills): 0.514 ms against 0.538 ms at 4 CTAs and 0.525 ms at 8 (config 3)
#ifndef HTGS_FWD_MINB
#define HTGS_FWD_MINB 6
#endif
constexpr int FWD_MIN_BLOCKS = HTGS_FWD_MINB;
template <bool STATS>
__global__ void __launch_bounds__(HTGS_TILE_PIX, FWD_MIN_BLOCKS)
blend_forward_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list, int W, int H,
                     const float2* __restrict__ means2D, const float4* __restrict__ conic_opacity,
                     const float4* __restrict__ feat, const float* __restrict__ bg, float* __restrict__ out_color,
                     float* __restrict__ out_depth, float* __restrict__ out_alpha, uint32_t* __restrict__ n_contrib,
                     unsigned long long* __restrict__ stats, uint8_t* __restrict__ reach, uint32_t* __restrict__ bin_capacity,
                     uint32_t capacity) {
    __shared__ StagedGaussian s_g[FWD_BATCH];
    __shared__ uint8_t s_mask[FWD_BATCH];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (!STATS && bin_capacity != nullptr && tid == 0 && blockIdx.x == 0 && blockIdx.y == 0) {
        bin_capacity[0] = reach != nullptr ? capacity : 0xffffffffu;  // which carving the cached reach masks belong to
        bin_capacity[1] = REACH_CACHE_MAGIC;
    }
    const int tile_x0 = blockIdx.x * HTGS_TILE, tile_y0 = blockIdx.y * HTGS_TILE;
    const int px = tile_x0 + warp_block_x(warp) + (lane & 7);
    const int py = tile_y0 + warp_block_y(warp) + (lane >> 3);
    const bool inside = px < W && py < H;
    const float pxf = (float)px, pyf = (float)py;
    // A saturated pixel (forward.cu:351-355 `done`) is marked by a NEGATIVE transmittance: its test_T = T * (1 - alpha)
    // is then negative, hence below the 1e-4 threshold, and the pair is rejected by the comparison that is evaluated
    // anyway -- no separate flag to test and to set in every visit.  |T| is the value the reference keeps.
#define HTGS_DONE (T < 0.f)

    const uint2 range = ranges[blockIdx.y * gridDim.x + blockIdx.x];
    const int n = (int)(range.y - range.x);
    const uint32_t sg_base = smem_u32(s_g);

    float T = inside ? 1.0f : -1.0f, wsum = 0.f;
    // Scalar accumulators on purpose: ptxas if-converts the short blend block below, and a predicated packed FFMA2 then
    // lands in a temporary pair that two moves copy back; predicated scalar FFMAs accumulate in place.
    float C0 = 0.f, C1 = 0.f, C2 = 0.f, dsum = 0.f;
    const uint64_t pxy = pk2(pxf, pyf);
    uint32_t last_contributor = 0;
    int last_in_batch = -1;  // batch position of the last pair this pixel blended in the current batch (-1: none)
    uint32_t st_blended = 0, st_tested = 0, st_done_pos = 0;
    uint32_t st_hist[7] = {0, 0, 0, 0, 0, 0, 0};  // warp visits by number of blending lanes: 0, 1, 2-4, 5-8, 9-16, 17-24, 25-32

    for (int base = 0; base < n; base += FWD_BATCH) {
        // leave when every pixel of the tile is saturated (forward.cu:313-315)
        if (__syncthreads_count(HTGS_DONE) == HTGS_TILE_PIX) break;
        const int cnt = min(FWD_BATCH, n - base);
        if (tid < cnt) {
            const uint32_t g = point_list[range.x + base + tid];
            const float2 xy = means2D[g];
            const float4 co = conic_opacity[g];
            stage_gaussian(&s_g[tid], xy, co, feat[g]);
            const uint32_t rm = block_reach_mask(xy, co, tile_x0, tile_y0);
            s_mask[tid] = (uint8_t)rm;
            if (!STATS && reach != nullptr) reach[range.x + base + tid] = (uint8_t)rm;  // kept for the backward
            if (STATS && rm == 0) atomicAdd(&stats[10], 1ull);  // instances whose Gaussian cannot reach any pixel of the tile
        }
        __syncthreads();
        if (__all_sync(0xffffffffu, HTGS_DONE)) continue;
        for (int c0 = 0; c0 < cnt; c0 += 32) {
            // Lane L tests batch position c0 + 31 - L: the LEADING one of the ballot is then the next position to visit, and
            // the loop works on its bit index p directly (FLO, IMAD for the address) instead of BREV, FLO, ADD, IMAD.
            const int j = c0 + 31 - lane;
            const bool mine = j < cnt && ((s_mask[j] >> warp) & 1);
            uint32_t m = __ballot_sync(0xffffffffu, mine);
            const int c31 = c0 + 31;
            const uint32_t chunk_end = sg_base + (uint32_t)c31 * STAGED_BYTES;
            while (m) {
                int p;  // index of the leading one (bfind -> one FLO; written as 31 - __clz(m) nvcc rebuilds clz and undoes the saving)
                asm("bfind.u32 %0, %1;" : "=r"(p) : "r"(m));
                uint32_t below;  // the p bits under the leading one
                asm("bmsk.clamp.b32 %0, 0, %1;" : "=r"(below) : "r"(p));
                m &= below;
                const uint32_t a = chunk_end - (uint32_t)p * STAGED_BYTES;
                const float4 q0 = lds128(a);       // mean.xy, conic.y, conic.z
                const float4 q1 = lds128(a + 16);  // conic.x, opacity, r, g
                const uint64_t d2 = sub2(pk2(q0.x, q0.y), pxy);     // (dx, dy)
                float dx, dy, my, mc;
                up2(d2, dx, dy);
                // pair_power(): fma(fma(dx, dx*cx, dy*(dy*cz)), -0.5, -(dy*(dx*cy)))
                up2(mul2(pk2(dy, dy), mul2(d2, pk2(q0.z, q0.w))), mc, my);  // (dy*(dx*cy), dy*(dy*cz))
                const float power = FM(FM(dx, MU(dx, q1.x), my), -0.5f, -mc);
                const float alpha = fminf(0.99f, MU(q1.y, expf(power)));
                const float test_T = MU(T, SB(1.0f, alpha));
                // the reference's three rejections (forward.cu:340-355) as one predicate chain
                const bool live = !(power > 0.0f) && !(alpha < 1.0f / 255.0f);
                if (STATS) st_tested += HTGS_DONE ? 0u : 1u;
                const bool kill = live && (test_T < 0.0001f);  // also true again for a pixel that is already saturated
                if (STATS && kill && !HTGS_DONE) st_done_pos = (uint32_t)(base + c31 - p + 1);
                if (STATS) {
                    const int nl = __popc(__ballot_sync(0xffffffffu, live && !kill));
                    const int bucket = nl == 0 ? 0 : nl == 1 ? 1 : nl <= 4 ? 2 : nl <= 8 ? 3 : nl <= 16 ? 4 : nl <= 24 ? 5 : 6;
#pragma unroll
                    for (int k = 0; k < 7; k++) st_hist[k] += (k == bucket) ? 1u : 0u;
                }
                // skip = !live || test_T < 1e-4 serves both the saturation mark (live && skip) and the blend (!skip); written in
                // PTX because nvcc otherwise derives a second comparison (test_T >= 1e-4) for the mark
                int skip_i;
                asm("{\n\t.reg .pred l, s;\n\t.reg .f32 t;\n\tsetp.ne.b32 l, %3, 0;\n\tsetp.lt.or.f32 s, %2, 0f38D1B717, !l;\n\t"
                    "abs.f32 t, %0;\n\tneg.f32 t, t;\n\t@l selp.f32 %0, t, %0, s;\n\tselp.u32 %1, 1, 0, s;\n\t}"
                    : "+f"(T), "=r"(skip_i) : "f"(test_T), "r"((int)live));
                const bool skip = skip_i != 0;
                if (!skip) {
                    const float2 q2 = lds64(a + 32);  // b, depth
                    const uint64_t aa = pk2(alpha, alpha);
                    float ac0, ac1, ac2, ad;
                    up2(mul2(aa, pk2(q1.z, q1.w)), ac0, ac1);  // alpha * c, forward.cu:359-361
                    up2(mul2(aa, pk2(q2.x, q2.y)), ac2, ad);
                    C0 = FM(T, ac0, C0);                       // C += T * (alpha * c)
                    C1 = FM(T, ac1, C1);
                    C2 = FM(T, ac2, C2);
                    dsum = FM(T, ad, dsum);
                    wsum = FM(T, alpha, wsum);
                    T = test_T;
                    last_in_batch = c31 - p;
                    if (STATS) st_blended++;
                }
            }
            if (__all_sync(0xffffffffu, HTGS_DONE)) break;
        }
        if (last_in_batch >= 0) {  // forward.cu:365: position in the tile's list, counted from 1
            last_contributor = (uint32_t)(base + last_in_batch + 1);
            last_in_batch = -1;
        }
    }
    if (STATS) {
        if (inside) {
            atomicAdd(&stats[0], (unsigned long long)(st_done_pos ? st_done_pos : (uint32_t)n));
            atomicAdd(&stats[1], (unsigned long long)st_blended);
            atomicAdd(&stats[2], (unsigned long long)st_tested);
        }
        if (lane == 0) {
#pragma unroll
            for (int k = 0; k < 7; k++) atomicAdd(&stats[3 + k], (unsigned long long)st_hist[k]);
        }
        return;
    }
    T = fabsf(T);
#undef HTGS_DONE
    if (inside) {
        const size_t pix = (size_t)py * W + px;
        const size_t plane = (size_t)H * W;
        n_contrib[pix] = last_contributor;
        out_color[pix] = FM(T, bg[0], C0);
        out_color[plane + pix] = FM(T, bg[1], C1);
        out_color[2 * plane + pix] = FM(T, bg[2], C2);
        out_alpha[pix] = wsum;  // forward.cu:378: the accumulated weight, not 1-T
        out_depth[pix] = dsum;
    }
}

int launch_blend_forward(int W, int H, const uint2* ranges, const uint32_t* point_list, const GeomView& g,
                         const float* bg, float* out_color, float* out_depth, float* out_alpha,
                         uint32_t* n_contrib, uint8_t* reach, uint32_t* bin_capacity, uint32_t capacity, cudaStream_t st) {
    const dim3 grid((W + HTGS_TILE - 1) / HTGS_TILE, (H + HTGS_TILE - 1) / HTGS_TILE);
    blend_forward_kernel<false><<<grid, HTGS_TILE_PIX, 0, st>>>(ranges, point_list, W, H, g.means2D, g.conic_opacity,
                                                                g.feat, bg, out_color, out_depth, out_alpha, n_contrib,
                                                                nullptr, reach, bin_capacity, capacity);
    HTGS_LAUNCH_CHECK("blend_forward_kernel");
    return 0;
}

int launch_blend_stats(int W, int H, const uint2* ranges, const uint32_t* point_list, const GeomView& g,
                       unsigned long long* stats, cudaStream_t st) {
    const dim3 grid((W + HTGS_TILE - 1) / HTGS_TILE, (H + HTGS_TILE - 1) / HTGS_TILE);
    blend_forward_kernel<true><<<grid, HTGS_TILE_PIX, 0, st>>>(ranges, point_list, W, H, g.means2D, g.conic_opacity, g.feat,
                                                               nullptr, nullptr, nullptr, nullptr, nullptr, stats, nullptr, nullptr, 0u);
    HTGS_LAUNCH_CHECK("blend_forward_kernel<stats>");
    return 0;
}

}  // namespace htgs
