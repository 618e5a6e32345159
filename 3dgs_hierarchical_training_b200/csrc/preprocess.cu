// K1: per-Gaussian preprocess (cull, EWA covariance projection, conic, radius,
// tile rectangle, SH colour) fused with the first level of the tile-count scan.
// Replaces FORWARD::preprocess (cuda_rasterizer/forward.cu:155-256) and half of
// cub::DeviceScan::InclusiveSum (rasterizer_impl.cu:278).
//
// HBM-bound streaming kernel: one thread per Gaussian, 128-bit loads of the SH
// block (the dominant stream, 192 B/Gaussian) issued only for Gaussians that
// survive culling; cov3D is not stored (backward recomputes it bit-identically
// from scale/rotation), colour and depth are packed into one float4 so the
// blend kernels gather 40 B per instance.
#include "raw_params.cuh"
#include <cstdio>

namespace htgs {

// SH colour, forward.cu:20-71, in the operation order of the reference's SASS (sm_100a).
// NCOEF floats of `sh` are valid ((D+1)^2 * 3).
template <int D>
__device__ __forceinline__ void sh_to_rgb(const float* sh, float x, float y, float z, float* res) {
#pragma unroll
    for (int c = 0; c < 3; c++) res[c] = MU(sh[c], HTGS_SH_C0);
    if (D > 0) {
        const float k1 = MU(y, HTGS_SH_C1), k2 = MU(z, HTGS_SH_C1), k3 = MU(x, HTGS_SH_C1);
#pragma unroll
        for (int c = 0; c < 3; c++) {
            float t = FM(-k1, sh[3 + c], res[c]);
            t = FM(k2, sh[6 + c], t);
            res[c] = FM(-k3, sh[9 + c], t);
        }
    }
    if (D > 1) {
        const float xx = MU(x, x), yy = MU(y, y), zz = MU(z, z), xy = MU(x, y), yz = MU(y, z), xz = MU(x, z);
        const float zz2 = AD(zz, zz);
        const float k4 = MU(xy, HTGS_SH_C2_0);
        const float k5 = MU(yz, HTGS_SH_C2_1);
        const float k6 = MU(SB(SB(zz2, xx), yy), HTGS_SH_C2_2);
        const float k7 = MU(xz, HTGS_SH_C2_3);
        const float xx_yy = SB(xx, yy);
        const float k8 = MU(xx_yy, HTGS_SH_C2_4);
#pragma unroll
        for (int c = 0; c < 3; c++) {
            float t = res[c];
            t = FM(k4, sh[12 + c], t);
            t = FM(k5, sh[15 + c], t);
            t = FM(k6, sh[18 + c], t);
            t = FM(k7, sh[21 + c], t);
            t = FM(k8, sh[24 + c], t);
            res[c] = t;
        }
        if (D > 2) {
            const float k9 = MU(MU(y, HTGS_SH_C3_0), FM(xx, 3.0f, -yy));
            const float k10 = MU(z, MU(xy, HTGS_SH_C3_1));
            const float m4 = SB(FM(zz, 4.0f, -xx), yy);
            const float k11 = MU(MU(y, HTGS_SH_C3_2), m4);
            const float k12 = MU(MU(z, HTGS_SH_C3_3), FM(yy, -3.0f, FM(xx, -3.0f, zz2)));
            const float k13 = MU(MU(x, HTGS_SH_C3_4), m4);
            const float k14 = MU(MU(z, HTGS_SH_C3_5), xx_yy);
            const float k15 = MU(MU(x, HTGS_SH_C3_6), FM(yy, -3.0f, xx));
#pragma unroll
            for (int c = 0; c < 3; c++) {
                float t = res[c];
                t = FM(k9, sh[27 + c], t);
                t = FM(k10, sh[30 + c], t);
                t = FM(k11, sh[33 + c], t);
                t = FM(k12, sh[36 + c], t);
                t = FM(k13, sh[39 + c], t);
                t = FM(k14, sh[42 + c], t);
                t = FM(k15, sh[45 + c], t);
                res[c] = t;
            }
        }
    }
}

template <int D>
__device__ __forceinline__ void load_sh_and_eval(const float* __restrict__ shp, bool vec_ok, bool vec256_ok, float x, float y,
                                                 float z, float* res) {
    constexpr int NC = (D + 1) * (D + 1) * 3;
    constexpr int NV = (NC + 3) / 4;
    constexpr int NV8 = (NC + 7) / 8;
    float sh[NV8 * 8];
    if (vec256_ok) {  // rows of 48 floats starting 32-byte aligned: whole sectors per lane (reads stay inside the row)
#pragma unroll
        for (int i = 0; i < NV8; i++) ldg256_nc(shp + 8 * i, sh + 8 * i);
    } else if (vec_ok) {
        const float4* p4 = reinterpret_cast<const float4*>(shp);
#pragma unroll
        for (int i = 0; i < NV; i++) {
            float4 v = __ldg(p4 + i);
            sh[4 * i] = v.x; sh[4 * i + 1] = v.y; sh[4 * i + 2] = v.z; sh[4 * i + 3] = v.w;
        }
    } else {
#pragma unroll
        for (int i = 0; i < NC; i++) sh[i] = __ldg(shp + i);
    }
    sh_to_rgb<D>(sh, x, y, z, res);
}

// RAW = true: `scales`, `rotations`, `opacities` are the un-activated parameters, `shs` is _features_dc
// [P,1,3] and `shs_rest` is _features_rest [P,15,3] (raw_params.cuh); M is 16.
// 4 CTAs per SM (64 registers, no spills): 0.065 ms against 0.070 ms at the compiler's own 70 registers / 3 CTAs
constexpr int PRE_MB = 4;
template <bool RAW>
__global__ void __launch_bounds__(HTGS_PRE_BLOCK, PRE_MB)
preprocess_kernel(int P, int D, int M, const float* __restrict__ means3D, const float* __restrict__ scales,
                  float scale_modifier, const float* __restrict__ rotations, const float* __restrict__ opacities,
                  const float* __restrict__ shs, const float* __restrict__ shs_rest,
                  const float* __restrict__ cov3D_precomp,
                  const float* __restrict__ colors_precomp, const float* __restrict__ viewmatrix,
                  const float* __restrict__ projmatrix, const float* __restrict__ campos, int W, int H,
                  float tanfovx, float tanfovy, float focal_x, float focal_y, int gx, int gy, int prefiltered,
                  int* __restrict__ radii, float2* __restrict__ means2D, float4* __restrict__ conic_opacity,
                  float4* __restrict__ feat, uint8_t* __restrict__ clamped, uint32_t* __restrict__ tiles_touched,
                  uint32_t* __restrict__ block_sums) {
    __shared__ float s_cam[35];
    __shared__ uint32_t s_warp_sum[HTGS_PRE_BLOCK / 32];
    if (threadIdx.x < 16) s_cam[threadIdx.x] = viewmatrix[threadIdx.x];
    else if (threadIdx.x < 32) s_cam[threadIdx.x] = projmatrix[threadIdx.x - 16];
    else if (threadIdx.x < 35) s_cam[threadIdx.x] = campos[threadIdx.x - 32];
    __syncthreads();
    const float* vm = s_cam;
    const float* pm = s_cam + 16;

    const int idx = blockIdx.x * HTGS_PRE_BLOCK + threadIdx.x;
    uint32_t my_tiles = 0;
    int my_radius = 0;
    if (idx < P) {
        const float px = means3D[3 * idx], py = means3D[3 * idx + 1], pz = means3D[3 * idx + 2];
        // scale and rotation are requested together with the position, before the near-plane test that only 2 % of the
        // Gaussians fail: one memory round trip less on the dependent chain position -> covariance -> SH row
        float4 q = make_float4(1.f, 0.f, 0.f, 0.f);
        float s0 = 0.f, s1 = 0.f, s2 = 0.f;
        if (cov3D_precomp == nullptr) {
            q = reinterpret_cast<const float4*>(rotations)[idx];
            s0 = scales[3 * idx]; s1 = scales[3 * idx + 1]; s2 = scales[3 * idx + 2];
        }
        const float depth = xform_row(vm, 2, px, py, pz);
        do {
            if (depth <= 0.2f) {  // auxiliary.h:154
                if (prefiltered) {
                    printf("Point is filtered although prefiltered is set. This shouldn't happen!");
                    __trap();
                }
                break;
            }
            const float hx = xform_row(pm, 0, px, py, pz);
            const float hy = xform_row(pm, 1, px, py, pz);
            const float hw = xform_row(pm, 3, px, py, pz);
            const float p_w = __frcp_rn(AD(hw, 0.0000001f));
            const float projx = MU(hx, p_w), projy = MU(hy, p_w);

            float cov3[6];
            if (cov3D_precomp != nullptr) {
#pragma unroll
                for (int k = 0; k < 6; k++) cov3[k] = cov3D_precomp[6 * (size_t)idx + k];
            } else {
                if (RAW) {
                    float n;
                    q = act_rotation(q, n);
                    s0 = act_scale(s0); s1 = act_scale(s1); s2 = act_scale(s2);
                }
                cov3d_from_scale_rot(s0, s1, s2, scale_modifier, q, cov3);
            }
            const Ewa e = ewa_setup(px, py, pz, vm, focal_x, focal_y, tanfovx, tanfovy);
            float a, b, c;
            ewa_cov2d(e, cov3, a, b, c);
            const float det = FM(a, c, -MU(b, b));
            if (det == 0.0f) break;
            const float det_inv = __frcp_rn(det);
            const float conx = MU(c, det_inv), cony = MU(det_inv, -b), conz = MU(a, det_inv);
            const float mid = MU(AD(a, c), 0.5f);
            const float sq = __fsqrt_rn(fmaxf(FM(mid, mid, -det), 0.1f));
            const float lam = fmaxf(AD(mid, sq), SB(mid, sq));
            const int rad = (int)ceilf(MU(__fsqrt_rn(lam), 3.0f));
            const float pix_x = ndc2pix(projx, W), pix_y = ndc2pix(projy, H);
            uint2 rmin, rmax;
            get_rect(pix_x, pix_y, rad, gx, gy, rmin, rmax);
            const uint32_t nt = (rmax.x - rmin.x) * (rmax.y - rmin.y);
            if (nt == 0) break;

            float rgb[3];
            uint32_t cl = 0;
            if (colors_precomp == nullptr) {
                const float dx = SB(px, s_cam[32]), dy = SB(py, s_cam[33]), dz = SB(pz, s_cam[34]);
                const float len = __fsqrt_rn(FM(dz, dz, FM(dx, dx, MU(dy, dy))));
                const float x = DV(dx, len), y = DV(dy, len), z = DV(dz, len);
                float res[3];
                if (RAW) {
                    float sh[48];
                    sh[0] = __ldg(shs + 3 * (size_t)idx); sh[1] = __ldg(shs + 3 * (size_t)idx + 1); sh[2] = __ldg(shs + 3 * (size_t)idx + 2);
                    float e[RAW_REST];
#pragma unroll
                    for (int k = 0; k < RAW_REST; k++) e[k] = 0.f;
                    if (D > 0) load_rest_row(shs_rest, idx, 3 * (D + 1) * (D + 1) - 3, e);
#pragma unroll
                    for (int k = 0; k < RAW_REST; k++) sh[3 + k] = e[k];
                    switch (D) {
                        case 0: sh_to_rgb<0>(sh, x, y, z, res); break;
                        case 1: sh_to_rgb<1>(sh, x, y, z, res); break;
                        case 2: sh_to_rgb<2>(sh, x, y, z, res); break;
                        default: sh_to_rgb<3>(sh, x, y, z, res); break;
                    }
                } else {
                const float* shp = shs + (size_t)idx * M * 3;
                const bool vec_ok = ((M * 3) % 4 == 0) && ((reinterpret_cast<uintptr_t>(shs) & 15) == 0);
                const bool v256 = M == 16 && ((reinterpret_cast<uintptr_t>(shs) & 31) == 0);
                switch (D) {
                    case 0: load_sh_and_eval<0>(shp, vec_ok && M * 3 >= 4, v256, x, y, z, res); break;
                    case 1: load_sh_and_eval<1>(shp, vec_ok, v256, x, y, z, res); break;
                    case 2: load_sh_and_eval<2>(shp, vec_ok && M * 3 >= 28, v256, x, y, z, res); break;
                    default: load_sh_and_eval<3>(shp, vec_ok, v256, x, y, z, res); break;
                }
                }
#pragma unroll
                for (int ch = 0; ch < 3; ch++) {
                    const float v = AD(res[ch], 0.5f);
                    if (v < 0.0f) { cl |= 1u << ch; rgb[ch] = 0.0f; } else rgb[ch] = v;
                }
                clamped[idx] = (uint8_t)cl;
            } else {
                rgb[0] = colors_precomp[3 * (size_t)idx];
                rgb[1] = colors_precomp[3 * (size_t)idx + 1];
                rgb[2] = colors_precomp[3 * (size_t)idx + 2];
            }
            means2D[idx] = make_float2(pix_x, pix_y);
            conic_opacity[idx] = make_float4(conx, cony, conz, RAW ? act_opacity(opacities[idx]) : opacities[idx]);
            feat[idx] = make_float4(rgb[0], rgb[1], rgb[2], depth);
            my_tiles = nt;
            my_radius = rad;
        } while (false);
        radii[idx] = my_radius;
        tiles_touched[idx] = my_tiles;
    }
    // first level of the scan: this block's tile total
    uint32_t s = my_tiles;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0) s_warp_sum[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t t = 0;
#pragma unroll
        for (int w = 0; w < HTGS_PRE_BLOCK / 32; w++) t += s_warp_sum[w];
        block_sums[blockIdx.x] = t;
    }
}

// Second level: exclusive scan of the per-block totals, in place, by one block.
// block_sums[nb] receives the grand total R, also stored to num_rendered.
__global__ void __launch_bounds__(1024) scan_block_sums_kernel(int nb, uint32_t* __restrict__ block_sums,
                                                               uint32_t* __restrict__ num_rendered) {
    __shared__ uint32_t s_warp[32];
    __shared__ uint32_t s_carry;
    if (threadIdx.x == 0) s_carry = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int base = 0; base < nb; base += 1024) {
        const int i = base + threadIdx.x;
        const uint32_t v = i < nb ? block_sums[i] : 0;
        uint32_t inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t n = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += n;
        }
        if (lane == 31) s_warp[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            uint32_t w = s_warp[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t n = __shfl_up_sync(0xffffffffu, w, o);
                if (lane >= o) w += n;
            }
            s_warp[lane] = w;  // inclusive over warps
        }
        __syncthreads();
        const uint32_t warp_off = warp == 0 ? 0 : s_warp[warp - 1];
        const uint32_t carry = s_carry;
        if (i < nb) block_sums[i] = carry + warp_off + inc - v;
        __syncthreads();
        if (threadIdx.x == 1023) s_carry = carry + warp_off + inc;
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        block_sums[nb] = s_carry;
        *num_rendered = s_carry;
    }
}

// K8: checkFrustum (rasterizer_impl.cu:54-66)
__global__ void mark_visible_kernel(int P, const float* __restrict__ means3D, const float* __restrict__ viewmatrix,
                                    uint8_t* __restrict__ present) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P) return;
    const float z = xform_row(viewmatrix, 2, means3D[3 * idx], means3D[3 * idx + 1], means3D[3 * idx + 2]);
    present[idx] = z > 0.2f ? 1 : 0;
}

int launch_preprocess(int P, int D, int M, const float* means3D, const float* scales, float scale_modifier,
                      const float* rotations, const float* opacities, const float* shs, const float* shs_rest,
                      const float* cov3D_precomp,
                      const float* colors_precomp, const float* viewmatrix, const float* projmatrix,
                      const float* campos, int W, int H, float tanfovx, float tanfovy, int prefiltered,
                      const GeomView& g, int* radii, cudaStream_t st) {
    const float focal_y = H / (2.0f * tanfovy);  // rasterizer_impl.cu:223-224
    const float focal_x = W / (2.0f * tanfovx);
    const int gx = (W + HTGS_TILE - 1) / HTGS_TILE, gy = (H + HTGS_TILE - 1) / HTGS_TILE;
    const int nb = (P + HTGS_PRE_BLOCK - 1) / HTGS_PRE_BLOCK;
    if (shs_rest != nullptr)  // raw-parameter entry point
        preprocess_kernel<true><<<nb, HTGS_PRE_BLOCK, 0, st>>>(P, D, M, means3D, scales, scale_modifier, rotations, opacities,
                                                               shs, shs_rest, cov3D_precomp, colors_precomp, viewmatrix,
                                                               projmatrix, campos, W, H, tanfovx, tanfovy, focal_x, focal_y,
                                                               gx, gy, prefiltered, radii, g.means2D, g.conic_opacity, g.feat,
                                                               g.clamped, g.tiles_touched, g.block_sums);
    else
        preprocess_kernel<false><<<nb, HTGS_PRE_BLOCK, 0, st>>>(P, D, M, means3D, scales, scale_modifier, rotations, opacities,
                                                                shs, nullptr, cov3D_precomp, colors_precomp, viewmatrix,
                                                                projmatrix, campos, W, H, tanfovx, tanfovy, focal_x, focal_y,
                                                                gx, gy, prefiltered, radii, g.means2D, g.conic_opacity, g.feat,
                                                                g.clamped, g.tiles_touched, g.block_sums);
    HTGS_LAUNCH_CHECK("preprocess_kernel");
    return 0;
}

int launch_scan_block_sums(int P, const GeomView& g, cudaStream_t st) {
    const int nb = (P + HTGS_PRE_BLOCK - 1) / HTGS_PRE_BLOCK;
    scan_block_sums_kernel<<<1, 1024, 0, st>>>(nb, g.block_sums, g.num_rendered);
    HTGS_LAUNCH_CHECK("scan_block_sums_kernel");
    return 0;
}

int launch_mark_visible(int P, const float* means3D, const float* viewmatrix, uint8_t* present, cudaStream_t st) {
    mark_visible_kernel<<<(P + 255) / 256, 256, 0, st>>>(P, means3D, viewmatrix, present);
    HTGS_LAUNCH_CHECK("mark_visible_kernel");
    return 0;
}

}  // namespace htgs
