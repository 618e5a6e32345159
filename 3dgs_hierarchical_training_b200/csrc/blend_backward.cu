// K5: back-to-front gradient of the tile blend (colour + depth + alpha).
// Replaces BACKWARD::render / renderCUDA (cuda_rasterizer/backward.cu:415-601).
//
// The reference issues 10 global float atomics per blended (pixel, Gaussian) pair (backward.cu:555-598).  Here a
// (tile, Gaussian) instance ends in three 128-bit vector reductions, and the sum over the tile's pixels that precedes
// them is done in two steps that keep the per-pair instruction count low (the kernel is issue-bound):
//
//  1. Per pair a lane produces only TWO numbers: w = G * dL/dalpha and aT = alpha * T.  Everything the ten partial
//     gradients need beyond those is a per-PIXEL constant:
//        dL/dcolour_c, dL/ddepth   = sum_pix aT * (dL/dpixel_c | dL/ddepth_pix)
//        dL/dopacity, dL/dmean2D, dL/dconic  follow from the six moments  sum_pix w * {1, x, y, x^2, xy, y^2}
//     with (x, y) the pixel's position relative to the centre of the warp's 8x4 block; the moments about the Gaussian's
//     centre that the reference's formulas use (backward.cu:582-598) are recovered from them in the epilogue of every
//     tensor-core phase (dx = X - x with X the centre's offset from the block centre: plain polynomial expansion, done
//     while |X| is a few pixels -- one frame for the whole tile cost sub-pixel Gaussians four digits of dL/dconic).
//  2. So the sum over a warp's 32 pixels for 8 Gaussians at a time is a small matrix product
//        C[16 x 8] = A[16 x 64] * B[64 x 8],   A = per-pixel constants (6 monomial rows, 4 upstream-gradient rows),
//                                              B = the (w | aT) values of 32 pixels x 8 Gaussians,
//     which the warp runs on the tensor cores (mma.sync m16n8k8, TF32 inputs, FP32 accumulate) instead of a
//     shuffle tree: lanes park (w, aT) of each visited Gaussian in a per-warp shared-memory panel (two STS per visit),
//     and every 8 visits 20 HMMA instructions produce all ten sums of all eight Gaussians.  FP32 accuracy is kept by
//     splitting the FP32 operands into TF32 high and low parts (x = hi + lo, products hi*hi + hi*lo + lo*hi; the
//     monomials are small integers and exact): relative error ~2^-21 per term, far inside the 1e-4 gradient bar.
//     r01 used a 12-shuffle / 22-select transpose-reduce per visit (~55 of 118 instructions per visit; ncu showed the
//     ALU pipe, not FMA, as the busiest); this form needs ~16.
//  3. The 8 warps of the tile add their results (moments already about the Gaussian's centre) into one per-batch
//     shared-memory accumulator row per Gaussian (shared-memory atomics: 4 per lane per 8 visits), which the flush
//     scales by the per-Gaussian constants and sends to HBM as three red.global.add.v4.f32 per (tile, Gaussian).
//
// Per-pixel arithmetic follows the reference (same thresholds, same alpha, same T = T/(1-alpha) recurrence); only
// the summation order / association of the gradient sums differs.  The 8x4 reach masks of blend_common.cuh skip
// Gaussians that cannot touch a warp's pixels, and whole batches behind the tile's last contributor.
//
// Accumulator layout gacc[P][12] (zeroed by the caller):
//   [0..3] dL/dmean2D.x, .y, dL/dopacity, dL/ddepth   [4..6] dL/dconic x,y,w
//   [8..10] dL/dcolour rgb                             [7],[11] unused
#include "blend_common.cuh"

namespace htgs {

constexpr int BWD_BATCH = 128;   // staged Gaussians per batch (96 / 112 / 64 measured slower in r01)
constexpr int PH_N = 8;          // Gaussians per tensor-core phase (the N of m16n8k8)
constexpr int PH_STRIDE = 36;    // floats per panel row: 32 pixels + 4 pad => conflict-free STS.32 by lane and LDS.128 by fragment
constexpr int ACC_STRIDE = 12;   // per-Gaussian tile sums: [0..5] moments of w about the Gaussian's centre (1, -dx, -dy, dx dx, dx dy, dy dy)  [6..9] colour r,g,b + depth  [10..11] pad

// Per-warp shared memory of the reduction.
//   park[r][lane]   r = 0..7: w of visit r, r = 8..15: aT of visit r - 8; column = lane = pixel index inside the warp's
//                   8x4 block (y * 8 + x).  Rows are PH_STRIDE floats apart: STS.32 by lane and LDS.128 by fragment
//                   lane are both bank-conflict free.
//   afrag[ks][l]    the A operand (per-pixel constants), stored in FRAGMENT order: one float4 per (k-step, fragment lane
//                   l = g*4 + t) = (A[g][k0], A[g+8][k0], A[g][k1], A[g+8][k1]) for the two pixels k0 = t*8 + 2ks,
//                   k1 = k0 + 1 the lane covers in that k-step, so a single LDS.128 lands in the four consecutive
//                   registers HMMA wants.  Rows of A:
//                     0..5   monomials 1, x, y, xx, xy, yy of the pixel's position in the block (half-integers, exact in TF32)
//                     6, 7   zero
//                     8..11  TF32 high parts of dL/dpixel r, g, b and dL/ddepth
//                     12..15 the low parts (d - hi) of the same four
//                   With B = w the rows 0..5 give the six moments (rows 8..15 are ignored); with B = aT rows 8..11 +
//                   rows 12..15 give the colour / depth sums (rows 0..7 are ignored).
constexpr int PH_ROW_AT = 8, PH_PARK_ROWS = 16, PH_KSTEPS = 4;
struct __align__(16) WarpPhase {
    float park[PH_PARK_ROWS][PH_STRIDE];   // 2304 B
    float4 afrag[PH_KSTEPS][32];           // 2048 B
};
constexpr uint32_t PH_OFF_AFRAG = PH_PARK_ROWS * PH_STRIDE * 4;
constexpr uint32_t PH_OFF_CENTRE = 15 * PH_STRIDE * 4 + 32 * 4;  // park[15][32..33] (pad words): the block centre in pixel coordinates

__device__ __forceinline__ void sts32(uint32_t a, float v) { asm volatile("st.shared.f32 [%0], %1;" ::"r"(a), "f"(v) : "memory"); }
__device__ __forceinline__ void sts64(uint32_t a, float x, float y) { asm volatile("st.shared.v2.f32 [%0], {%1,%2};" ::"r"(a), "f"(x), "f"(y) : "memory"); }
__device__ __forceinline__ void sts32u(uint32_t a, uint32_t v) { asm volatile("st.shared.u32 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }

// D += A * B, m16n8k8, A row-major 16x8 (a0: row g, col t | a1: row g+8, col t | a2: row g, col t+4 | a3: row g+8, col t+4),
// B 8x8 (b0: row t, col g | b1: row t+4, col g), D 16x8 (d0,d1: row g, cols 2t,2t+1 | d2,d3: row g+8), g = lane>>2, t = lane&3.
__device__ __forceinline__ void mma_tf32(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
// first MMA of a chain: D = A * B (zero accumulator input, no register initialisation)
__device__ __forceinline__ void mma_tf32_first(float (&d)[4], const float4& a, uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%10,%10,%10,%10};"
                 : "=f"(d[0]), "=f"(d[1]), "=f"(d[2]), "=f"(d[3])
                 : "r"(__float_as_uint(a.x)), "r"(__float_as_uint(a.y)), "r"(__float_as_uint(a.z)), "r"(__float_as_uint(a.w)),
                   "r"(b0), "r"(b1), "f"(0.f));
}
__device__ __forceinline__ void mma_tf32(float (&d)[4], const float4& a, uint32_t b0, uint32_t b1) {
    mma_tf32(d, __float_as_uint(a.x), __float_as_uint(a.y), __float_as_uint(a.z), __float_as_uint(a.w), b0, b1);
}
// x = hi + lo exactly: hi = x rounded to nearest TF32 (13 low mantissa bits clear), so |lo| <= 2^-12 |x| and what the
// tensor core drops of lo is below 2^-23 |x|.  The rounding is the integer form (add half an ulp of the kept part, clear
// the rest): cvt.rna.tf32.f32 compiles to the same two instructions PLUS an infinity guard (FSETP + SEL) per value, 80
// instead of 48 instructions for the 16 values of a phase.  Without the guard a finite |x| >= 2^128 - 2^115 would round
// to infinity -- gradient terms are nowhere near that, and infinities / NaNs themselves pass through unchanged.
__device__ __forceinline__ void split_tf32(float x, uint32_t& hi, uint32_t& lo) {
    hi = (__float_as_uint(x) + 0x1000u) & 0xffffe000u;
    lo = __float_as_uint(x - __uint_as_float(hi));
}
// The B operand (parked w / aT, 16 values per lane and phase) is split by TRUNCATION instead: the tensor core ignores the
// 13 low mantissa bits of a TF32 input, so x itself serves as its own high part at no cost, and the low part is
// x - (x with those bits cleared) -- one LOP3 and half a packed FADD2 per value instead of three instructions.  |lo| is
// below 2^-10 |x| (2^-11 when rounding), so what the tensor core drops of lo is below 2^-20 |x| per element.
__device__ __forceinline__ void split_tf32_pair(float x0, float x1, uint32_t& lo0, uint32_t& lo1) {
    float l0, l1;
    up2(sub2(pk2(x0, x1), pk2(__uint_as_float(__float_as_uint(x0) & 0xffffe000u), __uint_as_float(__float_as_uint(x1) & 0xffffe000u))), l0, l1);
    lo0 = __float_as_uint(l0);
    lo1 = __float_as_uint(l1);
}

// One tensor-core phase of a warp: sums the parked (w, aT) of `cnt` <= 8 visits over the warp's 32 pixels against the
// per-pixel constants and adds the 10 x cnt results into the batch accumulator rows.  ph = shared address of the
// warp's WarpPhase, acc = shared address of s_acc.  Warp-collective (all 32 lanes, converged).
__device__ __forceinline__ void reduce_phase(uint32_t ph, uint32_t acc, uint32_t sg, int cnt, int lane, uint64_t slots) {
    __syncwarp();
    const int g = lane >> 2, t = lane & 3;
    const uint32_t aw = ph + (uint32_t)g * (PH_STRIDE * 4u) + (uint32_t)t * 32u;   // visit g, pixels t*8 .. t*8+7
    const uint32_t aa = aw + (uint32_t)PH_ROW_AT * (PH_STRIDE * 4u);
    const uint32_t af = ph + PH_OFF_AFRAG + (uint32_t)lane * 16u;
    // batch index of the Gaussian behind each of this lane's two visits: one byte per visit, the latest visit in the lowest
    // byte (one byte permute picks both: byte (cnt-1-2t) & 7 of the 64-bit queue into byte 0, its predecessor into byte 1)
    const bool v0 = 2 * t < cnt, v1 = 2 * t + 1 < cnt;
    const uint32_t i0 = (uint32_t)(cnt - 1 - 2 * t) & 7u, i1 = (i0 - 1u) & 7u;
    const uint32_t both = __byte_perm((uint32_t)slots, (uint32_t)(slots >> 32), i0 | (i1 << 4));
    const uint32_t s0 = both & 0xffu, s1 = (both >> 8) & 0xffu;
    // the two Gaussians' centres relative to the block centre -- loaded here so that the loads complete under the MMAs (the
    // phase is a chain of dependent latencies, not of issue slots)
    const float2 bc = lds64(ph + PH_OFF_CENTRE), m0 = lds64(sg + s0 * STAGED_BYTES), m1 = lds64(sg + s1 * STAGED_BYTES);
    const float X0 = m0.x - bc.x, Y0 = m0.y - bc.y, X1 = m1.x - bc.x, Y1 = m1.y - bc.y;
    float cw1[4], cw2[4], ca1[4], ca2[4];  // four independent chains: (w_hi, w_lo, aT_hi, aT_lo) against the same A
#pragma unroll
    for (int half = 0; half < 2; half++) {
        const float4 w4 = lds128(aw + half * 16), t4 = lds128(aa + half * 16);
        const float wv[4] = {w4.x, w4.y, w4.z, w4.w}, tv[4] = {t4.x, t4.y, t4.z, t4.w};
#pragma unroll
        for (int k = 0; k < 2; k++) {  // k-step 2*half + k: elements 2k (k-index t) and 2k+1 (k-index t+4)
            const float4 a = lds128(af + (uint32_t)(2 * half + k) * (32u * 16u));
            uint32_t wl0, wl1, tl0, tl1;
            const uint32_t wh0 = __float_as_uint(wv[2 * k]), wh1 = __float_as_uint(wv[2 * k + 1]);
            const uint32_t th0 = __float_as_uint(tv[2 * k]), th1 = __float_as_uint(tv[2 * k + 1]);
            split_tf32_pair(wv[2 * k], wv[2 * k + 1], wl0, wl1);
            split_tf32_pair(tv[2 * k], tv[2 * k + 1], tl0, tl1);
            if (half == 0 && k == 0) {
                mma_tf32_first(cw1, a, wh0, wh1);
                mma_tf32_first(cw2, a, wl0, wl1);
                mma_tf32_first(ca1, a, th0, th1);
                mma_tf32_first(ca2, a, tl0, tl1);
            } else {
                mma_tf32(cw1, a, wh0, wh1);
                mma_tf32(cw2, a, wl0, wl1);
                mma_tf32(ca1, a, th0, th1);
                mma_tf32(ca2, a, tl0, tl1);
            }
        }
    }
    // lane (g, t) holds rows g and g+8 of columns (visits) 2t and 2t+1
    float x0 = cw1[0] + cw2[0], x1 = cw1[1] + cw2[1];         // moment g about the warp's BLOCK centre (g < 6)
    float y0 = ca1[2] + ca2[2], y1 = ca1[3] + ca2[3];         // colour/depth row g+8: high-part rows for g < 4, low-part rows above
    y0 += __shfl_xor_sync(0xffffffffu, y0, 16);               // lane g+4 holds the low-part row of the same quantity
    y1 += __shfl_xor_sync(0xffffffffu, y1, 16);
    // Moments about the block centre -> moments about each Gaussian's own centre (d = mean - pixel, the quantities
    // backward.cu:582-598 sums), BEFORE they meet the other warps' sums: with X = mean - block centre
    //    sum w dx = X S0 - Sx,   sum w dx dx = X (X S0 - 2 Sx) + Sxx,   sum w dx dy = X (Y S0 - Sy) - Y Sx + Sxy,  ...
    // The cancellation in these differences grows with (X / sigma)^2; X is a few pixels here, where one frame for the
    // whole tile (round 2's first form: |X| up to 8 and more) cost sub-pixel Gaussians near a tile corner four digits of
    // their conic gradient (tools/grad_margin.py).
    float2 S0, Sx, Sy;  // rows 0..2 of this lane's two visits: held by lanes t, 4 + t, 8 + t
    S0.x = __shfl_sync(0xffffffffu, x0, t);
    S0.y = __shfl_sync(0xffffffffu, x1, t);
    Sx.x = __shfl_sync(0xffffffffu, x0, 4 + t);
    Sx.y = __shfl_sync(0xffffffffu, x1, 4 + t);
    Sy.x = __shfl_sync(0xffffffffu, x0, 8 + t);
    Sy.y = __shfl_sync(0xffffffffu, x1, 8 + t);
    // per-row form  r = own + A * (V * S0 - K * SK) - R * Y * Sx  (first moments are kept with the opposite sign,
    // sum w (pixel - mean); the flush flips it back):
    //   g:     0      1      2      3        4        5
    //   A      -      X      Y      X        X        Y
    //   V      0     -1     -1      X        Y        Y
    //   K,SK   0      0      0      2,Sx     1,Sy     2,Sy
    //   R      0      0      0      0        1        0
    const bool usesX = g == 1 || g == 3 || g == 4, first = g < 3, cross = g == 4, kx = g == 3;
    const float vc = g == 0 ? 0.f : -1.f, fK = g < 3 ? 0.f : (g == 4 ? 1.f : 2.f), fR = g == 4 ? 1.f : 0.f;
    {
        const float A = usesX ? X0 : Y0, V0 = cross ? Y0 : A, V = first ? vc : V0;
        x0 = fmaf(-fR, Y0 * Sx.x, fmaf(A, fmaf(V, S0.x, -fK * (kx ? Sx.x : Sy.x)), x0));
    }
    {
        const float A = usesX ? X1 : Y1, V0 = cross ? Y1 : A, V = first ? vc : V0;
        x1 = fmaf(-fR, Y1 * Sx.y, fmaf(A, fmaf(V, S0.y, -fK * (kx ? Sx.y : Sy.y)), x1));
    }
    float* accp = reinterpret_cast<float*>(__cvta_shared_to_generic((size_t)acc));
    if (g < 6) {
        if (v0 && x0 != 0.f) atomicAdd(accp + s0 * ACC_STRIDE + g, x0);
        if (v1 && x1 != 0.f) atomicAdd(accp + s1 * ACC_STRIDE + g, x1);
    }
    if (g < 4) {
        if (v0 && y0 != 0.f) atomicAdd(accp + s0 * ACC_STRIDE + 6 + g, y0);
        if (v1 && y1 != 0.f) atomicAdd(accp + s1 * ACC_STRIDE + 6 + g, y1);
    }
    __syncwarp();
}

// 4 CTAs per SM (64 registers; r01: 1.12 ms at 4 CTAs against 1.27 ms at 3); shared memory 46.9 KB per CTA
constexpr int BWD_MIN_BLOCKS = 4;
__global__ void __launch_bounds__(HTGS_TILE_PIX, BWD_MIN_BLOCKS)
blend_backward_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ point_list, int W, int H,
                      const float* __restrict__ bg, const float2* __restrict__ means2D,
                      const float4* __restrict__ conic_opacity, const float4* __restrict__ feat,
                      const float* __restrict__ alphas, const uint32_t* __restrict__ n_contrib,
                      const float* __restrict__ dL_dpix, const float* __restrict__ dL_ddepth_pix,
                      const float* __restrict__ dL_dalpha_pix, float* __restrict__ gacc, const uint8_t* __restrict__ reach,
                      const uint32_t* __restrict__ bin_capacity, uint32_t capacity) {
    __shared__ WarpPhase s_ph[BLEND_WARPS];                          // 35 072 B
    __shared__ __align__(16) float s_acc[BWD_BATCH][ACC_STRIDE];     //  6 144 B
    __shared__ StagedGaussian s_g[BWD_BATCH];                        //  6 144 B
    __shared__ uint32_t s_id[BWD_BATCH];
    __shared__ uint8_t s_mask[BWD_BATCH];
    __shared__ uint32_t s_warp_last[BLEND_WARPS];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tile_x0 = blockIdx.x * HTGS_TILE, tile_y0 = blockIdx.y * HTGS_TILE;
    const int lx = warp_block_x(warp) + (lane & 7), ly = warp_block_y(warp) + (lane >> 3);  // position inside the tile
    const int px = tile_x0 + lx, py = tile_y0 + ly;
    const bool inside = px < W && py < H;
    const float pxf = (float)px, pyf = (float)py;
    const size_t pix = (size_t)py * W + px;
    const size_t plane = (size_t)H * W;

    const uint2 range = ranges[blockIdx.y * gridDim.x + blockIdx.x];
    const int n = (int)(range.y - range.x);
    // reach masks cached by the forward blend are valid only for the carving they were written under (blend_common.cuh)
    const bool reach_cached = reach != nullptr && bin_capacity != nullptr && bin_capacity[0] == capacity &&
                              bin_capacity[1] == REACH_CACHE_MAGIC;

    // backward.cu:463-486
    const float T_final = inside ? 1.0f - alphas[pix] : 0.f;
    float T = T_final;
    const uint32_t last_contributor = inside ? n_contrib[pix] : 0u;
    float d0 = 0.f, d1 = 0.f, d2 = 0.f, dd = 0.f, da = 0.f;
    if (inside) {
        d0 = dL_dpix[pix];
        d1 = dL_dpix[plane + pix];
        d2 = dL_dpix[2 * plane + pix];
        if (dL_ddepth_pix) dd = dL_ddepth_pix[pix];   // NULL = no upstream gradient for that output (importance pass)
        if (dL_dalpha_pix) da = dL_dalpha_pix[pix];
    }
    // -(T_final * bg . dL/dpixel): numerator of the background term, backward.cu:575-578
    const float bg_k = -T_final * (bg[0] * d0 + bg[1] * d1 + bg[2] * d2);
    const bool warp_has_bg = __any_sync(0xffffffffu, bg_k != 0.f);
    // A = sum_ch accum_rec[ch]*dL/dch (colour, depth, alpha channels) as the NEXT blended Gaussian will see it: the
    // reference folds the previous Gaussian's colour in lazily (backward.cu:547-567, last_alpha / last_color carried to
    // the next iteration); folding it in right after use is the same fma on the same values and carries one float
    // instead of four
    // The alpha channel's "colour" is the constant 1, so its term da enters Cd - A only as a constant: it is folded into
    // the start value (A' = A - da obeys the same recurrence), which keeps da out of the per-visit sum.
    float A = -da;

    // This pixel's column of the A operand, scattered into fragment order: pixel p = lane sits in k-step (p & 7) >> 1,
    // fragment lanes g*4 + (p >> 3) for every row pair (g, g+8), element (p & 1) of the k-step's pixel pair.
    const uint32_t ph = smem_u32(&s_ph[warp]);
    {
        // pixel position relative to the centre of the warp's 8x4 BLOCK (half-integers in [-3.5, 3.5] x [-1.5, 1.5];
        // squares <= 12.25: exact in TF32).  The phase epilogue moves the moments from there to each Gaussian's centre.
        const float fx = (float)(lane & 7) - 3.5f, fy = (float)(lane >> 3) - 1.5f;
        if (lane == 0) sts64(ph + PH_OFF_CENTRE, pxf + 3.5f, pyf + 1.5f);  // lane 0 is the block's top-left pixel
        const float lo_row[8] = {1.f, fx, fy, fx * fx, fx * fy, fy * fy, 0.f, 0.f};  // A rows 0..7
        float hi_row[8];                                                               // A rows 8..15
        const float dv[4] = {d0, d1, d2, dd};
#pragma unroll
        for (int r = 0; r < 4; r++) {
            uint32_t hi, lo;
            split_tf32(dv[r], hi, lo);
            hi_row[r] = __uint_as_float(hi);
            hi_row[4 + r] = __uint_as_float(lo);
        }
        const uint32_t e = ph + PH_OFF_AFRAG + (uint32_t)((lane & 7) >> 1) * (32u * 16u) + (uint32_t)(lane >> 3) * 16u +
                           (uint32_t)(lane & 1) * 8u;
#pragma unroll
        for (int g = 0; g < 8; g++) {
            sts32(e + (uint32_t)g * 64u, lo_row[g]);       // .x / .z of fragment lane g*4 + t
            sts32(e + (uint32_t)g * 64u + 4u, hi_row[g]);  // .y / .w
        }
    }

    uint32_t wl = last_contributor;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) wl = max(wl, __shfl_xor_sync(0xffffffffu, wl, o));
    if (lane == 0) s_warp_last[warp] = wl;
    __syncthreads();
    const uint32_t warp_last = wl;
    uint32_t block_last = 0;
#pragma unroll
    for (int w = 0; w < BLEND_WARPS; w++) block_last = max(block_last, s_warp_last[w]);

    const uint64_t pxy = pk2(pxf, pyf), d0_d1 = pk2(d0, d1), d2_dd = pk2(d2, dd);
    // (the bases pass through an opaque asm so that ptxas keeps them in registers: it otherwise rebuilds the shared
    // window address -- S2UR CgaCtaId, UMOV, UIADD3, ULEA -- inside every visit)
    uint32_t sg_base, acc_base;
    asm volatile("mov.u32 %0, %1;" : "=r"(sg_base) : "r"(smem_u32(s_g)));
    asm volatile("mov.u32 %0, %1;" : "=r"(acc_base) : "r"(smem_u32(s_acc)));
    const uint32_t park = ph + (uint32_t)lane * 4u;                          // this lane's column of the (w | aT) panel

    for (int base = 0; base < n; base += BWD_BATCH) {
        const int cnt = min(BWD_BATCH, n - base);
        // list positions (0-based, front to back) covered by this batch: top .. top-cnt+1
        const int top = n - 1 - base;
        const int jmin = top - (int)last_contributor + 1;  // first batch position this pixel's blend reached (may be < 0)
        if (top - cnt + 1 >= (int)block_last) continue;  // nothing in this batch was blended by any pixel
        __syncthreads();
        if (tid < cnt) {
            const uint32_t at = range.y - 1 - (uint32_t)(base + tid);  // back to front, backward.cu:502
            const uint32_t g = point_list[at];
            const float2 xy = means2D[g];
            const float4 co = conic_opacity[g];
            s_id[tid] = g;
            stage_gaussian(&s_g[tid], xy, co, feat[g]);
            // Every list position below the tile's last contributor was staged by the forward, so its mask is in the cache;
            // positions above it are never live, whatever their (possibly unwritten) mask says.
            s_mask[tid] = reach_cached ? reach[at] : (uint8_t)block_reach_mask(xy, co, tile_x0, tile_y0);
            float4* a4 = reinterpret_cast<float4*>(&s_acc[tid][0]);
            a4[0] = a4[1] = a4[2] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        __syncthreads();
        int parked = 0;           // visits parked in the warp's panel since its last tensor-core phase (warp-uniform)
        uint64_t slots = 0;       // their batch indices, one byte each, newest in the lowest byte (warp-uniform)
        uint32_t park_at = park;  // where this lane parks the next visit
        for (int c0 = 0; c0 < cnt; c0 += 32) {
            // lane L tests batch position c0 + 31 - L: the LEADING one of the ballot is the next position to visit (see the
            // forward kernel)
            const int j = c0 + 31 - lane;
            const bool mine = j < cnt && ((s_mask[j] >> warp) & 1) && (uint32_t)(top - j) < warp_last;
            uint32_t m = __ballot_sync(0xffffffffu, mine);
            const int c31 = c0 + 31;
            const uint32_t chunk_end = sg_base + (uint32_t)c31 * STAGED_BYTES;
            const int pmax = c31 - jmin;  // this pixel blended the positions with bit index <= pmax
            while (m) {
                int p;
                uint32_t below;
                asm("bfind.u32 %0, %1;" : "=r"(p) : "r"(m));
                asm("bmsk.clamp.b32 %0, 0, %1;" : "=r"(below) : "r"(p));
                m &= below;
                const uint32_t a = chunk_end - (uint32_t)p * STAGED_BYTES;
                const float4 q0 = lds128(a);       // mean.xy, conic.y, conic.z
                const float4 q1 = lds128(a + 16);  // conic.x, opacity, r, g
                const uint64_t dxy = sub2(pk2(q0.x, q0.y), pxy);   // (dx, dy)
                float dx, dy, my, mc;
                up2(dxy, dx, dy);
                up2(mul2(pk2(dy, dy), mul2(dxy, pk2(q0.z, q0.w))), mc, my);  // (dy*(dx*cy), dy*(dy*cz))
                const float power = FM(FM(dx, MU(dx, q1.x), my), -0.5f, -mc);  // pair_power()
                const float G = expf(power);
                const float alpha = fminf(0.99f, MU(q1.y, G));
                // `contributor` after its decrement is top - (c0 + b) (backward.cu:517-519); it is below this pixel's last
                // contributor from batch position jmin on
                const bool live = p <= pmax && !(power > 0.0f) && !(alpha < 1.0f / 255.0f);
                // Everything a lane contributes is proportional to w = G*dL/dalpha or to aT = alpha*T; lanes that do
                // not blend this Gaussian park zeros.  Straight-line code with selects at the end: a warp visit without
                // any live lane is rare (3 % at config 3), and the branch around an 18-instruction block cost a
                // BSSY/BRA/BSYNC triple plus the zero defaults in every visit.
                const float2 q2 = lds64(a + 32);  // b, depth
                // T = T / (1 - alpha), backward.cu:534.  One Newton step on the hardware reciprocal (<= 1 ulp from div.rn, 4
                // issue slots instead of 10); also feeds the background term.  Without the Newton step (2 instructions
                // fewer) the error of T compounds along the blend chain and dL_dcov3D misses the element-wise bar at
                // config 5 (excess 1.04).
                const float one_m_a = SB(1.f, alpha);
                float rinv;
                asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(rinv) : "f"(one_m_a));
                rinv = fmaf(fmaf(-one_m_a, rinv, 1.f), rinv, rinv);
                const float Tn = T * rinv;
                // Cd = r*d0 + g*d1 + b*d2 + depth*dd as two packed chains (da: see A)
                float c_lo, c_hi;
                up2(fma2(pk2(q2.x, q2.y), d2_dd, mul2(pk2(q1.z, q1.w), d0_d1)), c_lo, c_hi);
                const float D = (c_lo + c_hi) - A;          // Cd - A
                float dL_dopa = D * Tn;
                const float An = fmaf(alpha, D, A);         // alpha Cd + (1 - alpha) A
                if (warp_has_bg) dL_dopa = fmaf(bg_k, rinv, dL_dopa);
                const float w = live ? G * dL_dopa : 0.f;
                const float aT = live ? alpha * Tn : 0.f;
                T = live ? Tn : T;
                A = live ? An : A;
                if (__any_sync(0xffffffffu, live)) {
                    sts32(park_at, w);
                    sts32(park_at + (uint32_t)(PH_ROW_AT * PH_STRIDE * 4), aT);
                    slots = (slots << 8) | (uint64_t)(uint32_t)(c31 - p);
                    park_at += (uint32_t)(PH_STRIDE * 4);
                    if (++parked == PH_N) {
                        reduce_phase(ph, acc_base, sg_base, PH_N, lane, slots);
                        parked = 0;
                        park_at = park;
                    }
                }
            }
        }
        if (parked) reduce_phase(ph, acc_base, sg_base, parked, lane, slots);
        __syncthreads();
        if (tid < cnt) {
            const float4* a4 = reinterpret_cast<const float4*>(&s_acc[tid][0]);
            const float4 m0 = a4[0], m1 = a4[1], m2 = a4[2];
            // moments about the Gaussian's own centre, d = mean - pixel (the phase epilogues delivered them in that frame,
            // the two first moments with the opposite sign)
            const float S0 = m0.x, s1 = -m0.y, s2 = -m0.z, s4 = m0.w, s5 = m1.x, s6 = m1.y;
            const float gr = m1.z, gg = m1.w, gb = m2.x, gdepth = m2.y;
            const float4 q0 = s_g[tid].q0;
            const float4 q1 = s_g[tid].q1;
            // moments -> gradients (backward.cu:582-598), per-Gaussian constants applied once here
            const float cy = q0.z, cz = q0.w, cx = q1.x, op = q1.y;
            const float ddelx_dx = 0.5f * W, ddely_dy = 0.5f * H;  // backward.cu:490-491
            const float gmx = -op * (cx * s1 + cy * s2) * ddelx_dx;
            const float gmy = -op * (cz * s2 + cy * s1) * ddely_dy;
            const float gcx = -0.5f * op * s4, gcy = -0.5f * op * s5, gcw = -0.5f * op * s6;
            float4* dst = reinterpret_cast<float4*>(gacc + (size_t)s_id[tid] * 12);
            if (S0 != 0.f || s1 != 0.f || s2 != 0.f || gdepth != 0.f) atomicAdd(dst, make_float4(gmx, gmy, S0, gdepth));
            if (s4 != 0.f || s5 != 0.f || s6 != 0.f) atomicAdd(dst + 1, make_float4(gcx, gcy, gcw, 0.f));
            if (gr != 0.f || gg != 0.f || gb != 0.f) atomicAdd(dst + 2, make_float4(gr, gg, gb, 0.f));
        }
    }
}

int launch_blend_backward(int P, int W, int H, const uint2* ranges, const uint32_t* point_list, const GeomView& g,
                          const float* bg, const float* alphas, const uint32_t* n_contrib, const float* dL_dpix,
                          const float* dL_ddepth, const float* dL_dalpha, float* gacc, const uint8_t* reach,
                          const uint32_t* bin_capacity, uint32_t capacity, cudaStream_t st) {
    (void)P;
    const dim3 grid((W + HTGS_TILE - 1) / HTGS_TILE, (H + HTGS_TILE - 1) / HTGS_TILE);
    blend_backward_kernel<<<grid, HTGS_TILE_PIX, 0, st>>>(ranges, point_list, W, H, bg, g.means2D, g.conic_opacity, g.feat,
                                                          alphas, n_contrib, dL_dpix, dL_ddepth, dL_dalpha, gacc, reach,
                                                          bin_capacity, capacity);
    HTGS_LAUNCH_CHECK("blend_backward_kernel");
    return 0;
}

}  // namespace htgs
