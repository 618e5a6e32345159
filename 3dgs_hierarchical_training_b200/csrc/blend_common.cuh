// Shared pieces of the forward and backward tile-blend kernels.
//
// Thread mapping: one CTA per 16x16 tile (as the reference, forward.cu:261),
// but each of the 8 warps owns an 8x4 pixel block instead of a 16x2 strip, and
// every staged Gaussian carries an 8-bit mask saying which of those blocks it
// can possibly reach.  A warp only visits the Gaussians whose bit is set.
// Skipping is exact, not approximate: a Gaussian is dropped for a block only
// when a conservative bound proves that EVERY pixel of the block would have
// hit the reference's `alpha < 1/255` rejection (forward.cu:348), so the set and
// order of blended contributions per pixel -- and hence every output bit -- is
// unchanged.
//
// Both kernels are FP32-issue bound (ncu: issue slots 73-80% busy), so the inner
// loops are written for instruction count: one 48-byte shared-memory record per
// staged Gaussian read with two (three when blending) 128-bit loads from an
// address formed by a single multiply-add, and one combined predicate instead of
// a branch per rejection test.
#pragma once
#include "gs_math.cuh"

namespace htgs {

constexpr int BLEND_WARPS = 8;
// The forward blend leaves the reach mask of every sorted (tile, Gaussian) instance it staged in the binning buffer, so
// that the backward need not evaluate it again.  Both sides locate that array by carving the binning buffer with an
// instance count; the forward records the count it carved with (plus this magic word) in the image buffer, and the
// backward uses the cached masks only when its own carving count matches -- otherwise (a caller that sized the buffer
// from a prediction and hands the backward the exact count, a buffer from an older build) it computes the masks itself.
constexpr uint32_t REACH_CACHE_MAGIC = 0x48544753u;

// pixel block of warp w inside the tile: 2 columns x 4 rows of 8x4 blocks
__device__ __forceinline__ int warp_block_x(int w) { return (w & 1) * 8; }
__device__ __forceinline__ int warp_block_y(int w) { return (w >> 1) * 4; }

// One staged Gaussian in shared memory (48 B, 16-byte aligned):
//   q0 = (mean.x, mean.y, conic.y, conic.z)   q1 = (conic.x, opacity, r, g)   q2 = (b, depth, -, -)
// conic.y and conic.z sit next to each other so that (dx*cy, dy*cz) is one packed multiply and dy * that pair the
// next one: the exponent costs FADD2, FMUL2, FMUL, FMUL2, FFMA, FFMA with no register move to form a pair.
struct __align__(16) StagedGaussian {
    float4 q0, q1, q2;
};
constexpr int STAGED_BYTES = 48;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ float4 lds128(uint32_t a) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a) : "memory");
    return v;
}
__device__ __forceinline__ float2 lds64(uint32_t a) {
    float2 v;
    asm volatile("ld.shared.v2.f32 {%0,%1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"(a) : "memory");
    return v;
}
// Packed FP32 pairs (sm_100 FADD2 / FMUL2 / FFMA2): one issue slot for two IEEE-rounded operations, bit-identical to
// the scalar ones.  The blend loops are issue-bound, so every pair formed is an instruction saved.  Only products
// that are NOT followed by a packed add are written as mul2: ptxas contracts mul.rn.f32x2 + add.rn.f32x2 into FFMA2
// even with explicit rounding modifiers, which would change results.
__device__ __forceinline__ uint64_t pk2(float a, float b) { uint64_t r; asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ void up2(uint64_t r, float& a, float& b) { asm("mov.b64 {%0,%1}, %2;" : "=f"(a), "=f"(b) : "l"(r)); }
__device__ __forceinline__ uint64_t fma2(uint64_t a, uint64_t b, uint64_t c) { uint64_t d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ uint64_t mul2(uint64_t a, uint64_t b) { uint64_t d; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ uint64_t sub2(uint64_t a, uint64_t b) { uint64_t d; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ uint64_t add2(uint64_t a, uint64_t b) { uint64_t d; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }

__device__ __forceinline__ void stage_gaussian(StagedGaussian* dst, float2 xy, float4 co, float4 f) {
    dst->q0 = make_float4(xy.x, xy.y, co.y, co.z);
    dst->q1 = make_float4(co.x, co.w, f.x, f.y);
    dst->q2 = make_float4(f.z, f.w, 0.f, 0.f);
}

// Minimum over the rectangle dx in [dxl,dxh], dy in [dyl,dyh] of the convex
// quadratic q = 0.5*(A dx^2 + C dy^2) + B dx dy  ( = -power ), whose unconstrained minimum 0 sits at the origin.
// Origin inside the rectangle: 0.  Otherwise the minimum lies on the boundary, and only on an edge that FACES the
// origin: from any point of another edge the straight segment towards the origin runs through the rectangle first (q
// decreases along it) and leaves through a facing edge.  So at most two edge minimisations are needed -- the nearest
// vertical edge when the origin is outside the x range, the nearest horizontal edge when it is outside the y range --
// instead of all four (the staging of a batch, where this runs for every Gaussian and warp block, was 12 % of the
// forward kernel's instructions with four).
__device__ __forceinline__ float quad_min_over_rect(float A, float B, float C, float nBiC, float nBiA, float dxl,
                                                    float dxh, float dyl, float dyh) {
    const bool x_out = dxl > 0.f || dxh < 0.f, y_out = dyl > 0.f || dyh < 0.f;
    if (!x_out && !y_out) return 0.f;
    float best = 3.0e38f;
    {
        const float ex = dxl > 0.f ? dxl : dxh;  // vertical edge nearest to the origin
        const float dy = fminf(dyh, fmaxf(dyl, nBiC * ex));
        const float v = 0.5f * (A * ex * ex + C * dy * dy) + B * ex * dy;
        if (x_out) best = v;
    }
    {
        const float ey = dyl > 0.f ? dyl : dyh;  // horizontal edge nearest to the origin
        const float dx = fminf(dxh, fmaxf(dxl, nBiA * ey));
        const float v = 0.5f * (A * dx * dx + C * ey * ey) + B * dx * ey;
        if (y_out) best = fminf(best, v);
    }
    return best;
}

// Bit w set  <=>  warp block w of tile (tile_x0, tile_y0) may receive a
// contribution from this Gaussian.  Conservative: margins cover the fp32
// rounding of both this bound and the per-pixel evaluation.
__device__ __forceinline__ uint32_t block_reach_mask(float2 xy, float4 co, int tile_x0, int tile_y0) {
    const float A = co.x, B = co.y, C = co.z, o = co.w;
    // alpha = min(0.99, o*G) with G <= 1 can never reach 1/255 when o < 1/255
    if (!(o >= 1.0f / 255.0f)) return (o != o) ? 0xffu : 0u;
    // not a finite, well-conditioned positive-definite conic -> do not cull
    if (!(A > 0.f) || !(C > 0.f) || !(A < 1e18f) || !(C < 1e18f) || !(fabsf(B) < 1e18f) ||
        !(fabsf(xy.x) < 1e9f) || !(fabsf(xy.y) < 1e9f) || !(A * C - B * B > 1e-12f * A * C))
        return 0xffu;
    // contributes only if -power <= ln(255*o)
    const float thr = __logf(255.0f * o) * 1.0001f + 1e-3f;
    const float nBiC = -B / C, nBiA = -B / A;
    uint32_t m = 0;
#pragma unroll
    for (int w = 0; w < BLEND_WARPS; w++) {
        const float x0 = (float)(tile_x0 + warp_block_x(w)), y0 = (float)(tile_y0 + warp_block_y(w));
        const float dxl = xy.x - (x0 + 7.f), dxh = xy.x - x0;
        const float dyl = xy.y - (y0 + 3.f), dyh = xy.y - y0;
        const float q = quad_min_over_rect(A, B, C, nBiC, nBiA, dxl, dxh, dyl, dyh);
        const float mx = fmaxf(fabsf(dxl), fabsf(dxh)), my = fmaxf(fabsf(dyl), fabsf(dyh));
        const float slack = 4e-6f * (A * mx * mx + C * my * my + fabsf(B) * mx * my);
        if (q <= thr + slack) m |= 1u << w;
    }
    return m;
}

}  // namespace htgs
