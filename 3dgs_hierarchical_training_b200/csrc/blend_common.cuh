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
#pragma once
#include "gs_math.cuh"

namespace htgs {

constexpr int BLEND_WARPS = 8;

// pixel block of warp w inside the tile: 2 columns x 4 rows of 8x4 blocks
__device__ __forceinline__ int warp_block_x(int w) { return (w & 1) * 8; }
__device__ __forceinline__ int warp_block_y(int w) { return (w >> 1) * 4; }

// Minimum over the rectangle dx in [dxl,dxh], dy in [dyl,dyh] of the convex
// quadratic q = 0.5*(A dx^2 + C dy^2) + B dx dy  ( = -power ).
__device__ __forceinline__ float quad_min_over_rect(float A, float B, float C, float dxl, float dxh, float dyl,
                                                    float dyh) {
    if (dxl <= 0.f && dxh >= 0.f && dyl <= 0.f && dyh >= 0.f) return 0.f;
    float best = 3.0e38f;
    const float invC = 1.f / C, invA = 1.f / A;
#pragma unroll
    for (int s = 0; s < 2; s++) {
        const float ex = s ? dxh : dxl;  // vertical edges dx = ex
        float dy = fminf(dyh, fmaxf(dyl, -B * ex * invC));
        best = fminf(best, 0.5f * (A * ex * ex + C * dy * dy) + B * ex * dy);
        const float ey = s ? dyh : dyl;  // horizontal edges dy = ey
        float dx = fminf(dxh, fmaxf(dxl, -B * ey * invA));
        best = fminf(best, 0.5f * (A * dx * dx + C * ey * ey) + B * dx * ey);
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
    uint32_t m = 0;
#pragma unroll
    for (int w = 0; w < BLEND_WARPS; w++) {
        const float x0 = (float)(tile_x0 + warp_block_x(w)), y0 = (float)(tile_y0 + warp_block_y(w));
        const float dxl = xy.x - (x0 + 7.f), dxh = xy.x - x0;
        const float dyl = xy.y - (y0 + 3.f), dyh = xy.y - y0;
        const float q = quad_min_over_rect(A, B, C, dxl, dxh, dyl, dyh);
        const float mx = fmaxf(fabsf(dxl), fabsf(dxh)), my = fmaxf(fabsf(dyl), fabsf(dyh));
        const float slack = 4e-6f * (A * mx * mx + C * my * my + fabsf(B) * mx * my);
        if (q <= thr + slack) m |= 1u << w;
    }
    return m;
}

}  // namespace htgs
