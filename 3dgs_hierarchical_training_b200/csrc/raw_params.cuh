// Caller-side parameter preparation folded into K1 / K7.
//
// The HT model keeps its parameters un-activated and split (scene/gaussian_model_ht.py:127-188):
//   scales    = exp(_scaling)                      rotations = F.normalize(_rotation)
//   opacities = sigmoid(_opacity)                  shs       = cat(_features_dc[P,1,3], _features_rest[P,15,3], dim=1)
// Every render call materialises those four tensors (the cat alone copies 192 B per Gaussian) and the
// autograd graph walks back through them (another 192 B per Gaussian for the split of dL/dshs).  The
// "raw" entry points take the un-activated, un-concatenated parameters, apply the activations in
// registers, and write the gradients with respect to the raw parameters directly.
//
// _features_rest rows are 180 bytes, so only every fourth row starts on a 16-byte boundary.  A row is
// therefore moved as ten aligned 128-bit accesses for the interior (words 4..43 of the aligned window
// that contains it -- always inside the row) plus up to eight predicated 32-bit accesses for the two
// edges, and rotated into place by a two-stage select on the row's phase (idx & 3).  No access ever
// touches a word outside the row, so neighbouring rows (other threads) are never read or written.
#pragma once
#include "gs_math.cuh"

namespace htgs {

constexpr int RAW_REST = 45;  // floats per _features_rest row (15 coefficients x RGB)

__device__ __forceinline__ float act_scale(float raw) { return expf(raw); }              // torch.exp
__device__ __forceinline__ float act_opacity(float raw) { return DV(1.0f, AD(1.0f, expf(-raw))); }  // torch.sigmoid

// F.normalize(q, dim=-1): q / max(||q||_2, 1e-12).  `inv_n` returns 1 / that denominator's value n itself.
__device__ __forceinline__ float4 act_rotation(float4 q, float& n) {
    n = __fsqrt_rn(AD(AD(AD(MU(q.x, q.x), MU(q.y, q.y)), MU(q.z, q.z)), MU(q.w, q.w)));
    n = fmaxf(n, 1e-12f);
    return make_float4(DV(q.x, n), DV(q.y, n), DV(q.z, n), DV(q.w, n));
}

// e[k] = rest[45*idx + k], k < nw (the remaining entries are 0).  `rest` must be 16-byte aligned.
__device__ __forceinline__ void load_rest_row(const float* __restrict__ rest, int idx, int nw, float (&e)[RAW_REST]) {
    const int o = idx & 3;  // (45 * idx) mod 4
    const float* win = rest + (size_t)RAW_REST * idx - o;  // aligned window, row = words [o, o + 45)
    float w[48];
#pragma unroll
    for (int j = 1; j <= 10; j++) {
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (4 * j < o + nw) v = __ldg(reinterpret_cast<const float4*>(win) + j);
        w[4 * j] = v.x; w[4 * j + 1] = v.y; w[4 * j + 2] = v.z; w[4 * j + 3] = v.w;
    }
#pragma unroll
    for (int j = 0; j < 4; j++) w[j] = (j >= o && j < o + nw) ? __ldg(win + j) : 0.f;
#pragma unroll
    for (int j = 44; j < 48; j++) w[j] = (j < o + nw) ? __ldg(win + j) : 0.f;
    float a[47];
#pragma unroll
    for (int k = 0; k < 47; k++) a[k] = (o & 1) ? w[k + 1] : w[k];
#pragma unroll
    for (int k = 0; k < RAW_REST; k++) e[k] = (o & 2) ? a[k + 2] : a[k];
}

// rest_out[45*idx + k] = e[k] for all 45 k.  `rest_out` must be 16-byte aligned.
__device__ __forceinline__ void store_rest_row(float* __restrict__ rest_out, int idx, const float (&e)[RAW_REST]) {
    const int o = idx & 3;
    float* win = rest_out + (size_t)RAW_REST * idx - o;
    float a[46];  // a[j] = e[j - (o & 1)]
#pragma unroll
    for (int j = 0; j < 46; j++) {
        const float lo = j >= 1 ? e[j - 1] : 0.f, hi = j < RAW_REST ? e[j] : 0.f;
        a[j] = (o & 1) ? lo : hi;
    }
    float w[48];  // w[j] = a[j - (o & 2)] = e[j - o]
#pragma unroll
    for (int j = 0; j < 48; j++) {
        const float lo = (j >= 2) ? a[j - 2] : 0.f, hi = j < 46 ? a[j] : 0.f;
        w[j] = (o & 2) ? lo : hi;
    }
#pragma unroll
    for (int j = 1; j <= 10; j++)
        reinterpret_cast<float4*>(win)[j] = make_float4(w[4 * j], w[4 * j + 1], w[4 * j + 2], w[4 * j + 3]);
#pragma unroll
    for (int j = 0; j < 4; j++)
        if (j >= o) win[j] = w[j];
#pragma unroll
    for (int j = 44; j < 48; j++)
        if (j < o + RAW_REST) win[j] = w[j];
}

}  // namespace htgs
