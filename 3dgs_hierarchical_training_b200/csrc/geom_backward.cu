// K6 + K7 fused: per-Gaussian backward of the projection.
// Replaces computeCov2DCUDA and BACKWARD::preprocessCUDA
// (cuda_rasterizer/backward.cu:144-274, :346-412, SH backward :20-139,
// scale/rotation backward :278-341) and the ten zero-filled gradient tensors of
// rasterize_points.cu:158-167.
//
// HBM-bound streaming kernel, one thread per Gaussian: reads the 48 B
// accumulator row written by the blend backward plus the Gaussian's parameters,
// recomputes cov3D / T instead of loading stored copies, and writes EVERY output
// element exactly once (zeros for culled Gaussians), so the caller allocates the
// outputs uninitialised and no memset pass over 300 B/Gaussian is needed.
#include "raw_params.cuh"

namespace htgs {

__device__ __forceinline__ void store3(float* p, float a, float b, float c) { p[0] = a; p[1] = b; p[2] = c; }

// Generic-layout fallback of the SH gradient store (any M): out[k*3 + c] = w[k] * g[c] for
// k < ncoef_active, 0 above.  All indices into w[] are compile-time so it stays in registers.
__device__ __forceinline__ void write_sh_grad(float* __restrict__ out, const float (&w)[16], int ncoef_active, int M,
                                              float g0, float g1, float g2) {
#pragma unroll
    for (int k = 0; k < 16; k++) {
        if (k < M) {
            const float wk = k < ncoef_active ? w[k] : 0.f;
            out[3 * k] = wk * g0;
            out[3 * k + 1] = wk * g1;
            out[3 * k + 2] = wk * g2;
        }
    }
    for (int e = 48; e < 3 * M; e++) out[e] = 0.f;
}

// RAW = true (raw_params.cuh): `scales`, `rotations` are un-activated, `shs` / `dL_dsh` are _features_dc and
// its gradient, `shs_rest` / `dL_dsh_rest` are _features_rest and its gradient, `opacity_raw` the
// un-activated opacity; all parameter gradients are written with respect to the raw parameters.
// 128-thread CTAs, at least 6 per SM (<= 80 registers): 0.129 ms; 256-thread CTAs 0.134 ms, forcing 64 registers 0.140 ms
constexpr int GB_BLOCK = 128;
constexpr int GB_MB = 6;
template <bool RAW>
__global__ void __launch_bounds__(GB_BLOCK, GB_MB)
geom_backward_kernel(int P, int D, int M, const float* __restrict__ means3D, const int* __restrict__ radii,
                     const float* __restrict__ shs, const float* __restrict__ shs_rest,
                     const float* __restrict__ opacity_raw, float* __restrict__ dL_dsh_rest,
                     const uint8_t* __restrict__ clamped, bool colors_are_precomp,
                     const float* __restrict__ scales, float scale_modifier, const float* __restrict__ rotations,
                     const float* __restrict__ cov3D_precomp, const float* __restrict__ viewmatrix,
                     const float* __restrict__ projmatrix, const float* __restrict__ campos, float h_x, float h_y,
                     float tanfovx, float tanfovy, const float* __restrict__ gacc, float* __restrict__ dL_dmeans2D,
                     float* __restrict__ dL_dopacity, float* __restrict__ dL_dcolors, float* __restrict__ dL_dmeans3D,
                     float* __restrict__ dL_dcov3D, float* __restrict__ dL_dsh, float* __restrict__ dL_dscales,
                     float* __restrict__ dL_drotations) {
    __shared__ float s_cam[35];
    if (threadIdx.x < 16) s_cam[threadIdx.x] = viewmatrix[threadIdx.x];
    else if (threadIdx.x < 32) s_cam[threadIdx.x] = projmatrix[threadIdx.x - 16];
    else if (threadIdx.x < 35) s_cam[threadIdx.x] = campos[threadIdx.x - 32];
    __syncthreads();
    const float* vm = s_cam;
    const float* proj = s_cam + 16;
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P) return;
    const bool sh_vec = dL_dsh != nullptr && ((M * 3) % 4 == 0) && ((reinterpret_cast<uintptr_t>(dL_dsh) & 15) == 0);
    // 256-bit accesses on the 192-byte SH rows (one 32-byte sector per lane and instruction, gs_math.cuh)
    const bool sh_vec256 = !RAW && sh_vec && M == 16 && ((reinterpret_cast<uintptr_t>(dL_dsh) & 31) == 0);

    if (!(radii[idx] > 0)) {
        store3(dL_dmeans2D + 3 * (size_t)idx, 0.f, 0.f, 0.f);
        dL_dopacity[idx] = 0.f;
        store3(dL_dmeans3D + 3 * (size_t)idx, 0.f, 0.f, 0.f);
        if (dL_dcolors) store3(dL_dcolors + 3 * (size_t)idx, 0.f, 0.f, 0.f);
        if (dL_dcov3D) {
#pragma unroll
            for (int k = 0; k < 6; k++) dL_dcov3D[6 * (size_t)idx + k] = 0.f;
        }
        if (RAW) {
            store3(dL_dsh + 3 * (size_t)idx, 0.f, 0.f, 0.f);
            float z45[RAW_REST];
#pragma unroll
            for (int k = 0; k < RAW_REST; k++) z45[k] = 0.f;
            store_rest_row(dL_dsh_rest, idx, z45);
        } else if (dL_dsh) {
            float* o = dL_dsh + (size_t)idx * M * 3;
            if (sh_vec256) {
                const float z8[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
                for (int i = 0; i < 6; i++) stg256(o + 8 * i, z8);
            } else if (sh_vec) {
                for (int i = 0; i < M * 3 / 4; i++) reinterpret_cast<float4*>(o)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
            } else {
                for (int i = 0; i < M * 3; i++) o[i] = 0.f;
            }
        }
        if (dL_dscales) store3(dL_dscales + 3 * (size_t)idx, 0.f, 0.f, 0.f);
        if (dL_drotations) reinterpret_cast<float4*>(dL_drotations)[idx] = make_float4(0.f, 0.f, 0.f, 0.f);
        return;
    }

    const float4* ga = reinterpret_cast<const float4*>(gacc + (size_t)idx * 12);
    const float4 g0 = ga[0], g1 = ga[1], g2 = ga[2];
    const float g2x = g0.x, g2y = g0.y, g_opacity = g0.z, g_depth = g0.w;
    const float dcx = g1.x, dcy = g1.y, dcz = g1.z;  // dL/dconic .x .y .w  (backward.cu:165)
    const float gc0 = g2.x, gc1 = g2.y, gc2 = g2.z;  // dL/dcolour

    const float mx = means3D[3 * (size_t)idx], my = means3D[3 * (size_t)idx + 1], mz = means3D[3 * (size_t)idx + 2];

    // ---- covariance: recompute what the forward used ----
    float c3[6];
    float4 q = make_float4(0.f, 0.f, 0.f, 0.f), q_raw = q;
    float q_norm = 1.f;
    float sc0 = 0.f, sc1 = 0.f, sc2 = 0.f;
    if (cov3D_precomp != nullptr) {
#pragma unroll
        for (int k = 0; k < 6; k++) c3[k] = cov3D_precomp[6 * (size_t)idx + k];
    } else {
        q = reinterpret_cast<const float4*>(rotations)[idx];
        sc0 = scales[3 * (size_t)idx]; sc1 = scales[3 * (size_t)idx + 1]; sc2 = scales[3 * (size_t)idx + 2];
        if (RAW) {
            q_raw = q;
            q = act_rotation(q, q_norm);
            sc0 = act_scale(sc0); sc1 = act_scale(sc1); sc2 = act_scale(sc2);
        }
        cov3d_from_scale_rot(sc0, sc1, sc2, scale_modifier, q, c3);
    }

    // ---- computeCov2DCUDA, backward.cu:144-274 ----
    const Ewa e = ewa_setup(mx, my, mz, vm, h_x, h_y, tanfovx, tanfovy);
    const float x_grad_mul = (e.txtz < -e.limx || e.txtz > e.limx) ? 0.f : 1.f;
    const float y_grad_mul = (e.tytz < -e.limy || e.tytz > e.limy) ? 0.f : 1.f;
    float a, b, c;
    ewa_cov2d(e, c3, a, b, c);
    const float denom = a * c - b * b;
    float dL_da = 0.f, dL_db = 0.f, dL_dc = 0.f;
    const float denom2inv = 1.0f / ((denom * denom) + 0.0000001f);
    const float T00 = e.T00, T01 = e.T01, T02 = e.T02, T10 = e.T10, T11 = e.T11, T12 = e.T12;
    float dcov[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (denom2inv != 0.f) {
        // conic = adj(C)/det with adj(C) = K = [[c,-b],[-b,a]], so a perturbation dC moves it by -conic dC conic and
        // dL/dC = -(K G K)/det^2, G = [[dcx, dcy],[dcy, dcz]] (the pair accumulates the off-diagonal once for both
        // symmetric entries); 1/det^2 carries the reference's 1e-7 regulariser (backward.cu:203).
        const float k00 = c, k01 = -b, k11 = a;
        const float h00 = dcx * k00 + dcy * k01, h01 = dcx * k01 + dcy * k11;  // H = G K
        const float h10 = dcy * k00 + dcz * k01, h11 = dcy * k01 + dcz * k11;
        dL_da = -denom2inv * (k00 * h00 + k01 * h10);                          // -(K H)_00
        dL_dc = -denom2inv * (k01 * h01 + k11 * h11);                          // -(K H)_11
        dL_db = -2.f * denom2inv * (k00 * h01 + k01 * h11);                    // both off-diagonal entries
        // C = T Sigma T^T restricted to rows t0, t1 of T:  dL/dSigma = t0 p^T + t1 r^T with
        // p = dL_da t0 + (dL_db/2) t1,  r = (dL_db/2) t0 + dL_dc t1  (symmetric); the packed gradient holds the
        // diagonal once and each off-diagonal pair summed.
        const float t0[3] = {T00, T01, T02}, t1[3] = {T10, T11, T12};
        const float hb = 0.5f * dL_db;
        float pv[3], rv[3];
#pragma unroll
        for (int j = 0; j < 3; j++) {
            pv[j] = dL_da * t0[j] + hb * t1[j];
            rv[j] = hb * t0[j] + dL_dc * t1[j];
        }
        dcov[0] = t0[0] * pv[0] + t1[0] * rv[0];
        dcov[3] = t0[1] * pv[1] + t1[1] * rv[1];
        dcov[5] = t0[2] * pv[2] + t1[2] * rv[2];
        dcov[1] = 2.f * (t0[0] * pv[1] + t1[0] * rv[1]);
        dcov[2] = 2.f * (t0[0] * pv[2] + t1[0] * rv[2]);
        dcov[4] = 2.f * (t0[1] * pv[2] + t1[1] * rv[2]);
    }
    // rows of Vrk * T[0] and Vrk * T[1]
    const float vt0x = T00 * c3[0] + T01 * c3[1] + T02 * c3[2];
    const float vt0y = T00 * c3[1] + T01 * c3[3] + T02 * c3[4];
    const float vt0z = T00 * c3[2] + T01 * c3[4] + T02 * c3[5];
    const float vt1x = T10 * c3[0] + T11 * c3[1] + T12 * c3[2];
    const float vt1y = T10 * c3[1] + T11 * c3[3] + T12 * c3[4];
    const float vt1z = T10 * c3[2] + T11 * c3[4] + T12 * c3[5];
    const float dL_dT00 = 2 * vt0x * dL_da + vt1x * dL_db;
    const float dL_dT01 = 2 * vt0y * dL_da + vt1y * dL_db;
    const float dL_dT02 = 2 * vt0z * dL_da + vt1z * dL_db;
    const float dL_dT10 = 2 * vt1x * dL_dc + vt0x * dL_db;
    const float dL_dT11 = 2 * vt1y * dL_dc + vt0y * dL_db;
    const float dL_dT12 = 2 * vt1z * dL_dc + vt0z * dL_db;
    const float dL_dJ00 = vm[0] * dL_dT00 + vm[4] * dL_dT01 + vm[8] * dL_dT02;
    const float dL_dJ02 = vm[2] * dL_dT00 + vm[6] * dL_dT01 + vm[10] * dL_dT02;
    const float dL_dJ11 = vm[1] * dL_dT10 + vm[5] * dL_dT11 + vm[9] * dL_dT12;
    const float dL_dJ12 = vm[2] * dL_dT10 + vm[6] * dL_dT11 + vm[10] * dL_dT12;
    const float tz = 1.f / e.tz, tz2 = tz * tz, tz3 = tz2 * tz;
    const float dL_dtx = x_grad_mul * -h_x * tz2 * dL_dJ02;
    const float dL_dty = y_grad_mul * -h_y * tz2 * dL_dJ12;
    const float dL_dtz = -h_x * tz2 * dL_dJ00 - h_y * tz2 * dL_dJ11 + (2 * h_x * e.tx) * tz3 * dL_dJ02 +
                         (2 * h_y * e.ty) * tz3 * dL_dJ12;
    float dmx = vm[0] * dL_dtx + vm[1] * dL_dty + vm[2] * dL_dtz;  // backward.cu:268-273
    float dmy = vm[4] * dL_dtx + vm[5] * dL_dty + vm[6] * dL_dtz;
    float dmz = vm[8] * dL_dtx + vm[9] * dL_dty + vm[10] * dL_dtz;

    // ---- preprocessCUDA backward, backward.cu:372-403 ----
    const float m_hw = proj[3] * mx + proj[7] * my + proj[11] * mz + proj[15];
    const float m_w = 1.0f / (m_hw + 0.0000001f);
    const float mul1 = (proj[0] * mx + proj[4] * my + proj[8] * mz + proj[12]) * m_w * m_w;
    const float mul2 = (proj[1] * mx + proj[5] * my + proj[9] * mz + proj[13]) * m_w * m_w;
    dmx += (proj[0] * m_w - proj[3] * mul1) * g2x + (proj[1] * m_w - proj[3] * mul2) * g2y;
    dmy += (proj[4] * m_w - proj[7] * mul1) * g2x + (proj[5] * m_w - proj[7] * mul2) * g2y;
    dmz += (proj[8] * m_w - proj[11] * mul1) * g2x + (proj[9] * m_w - proj[11] * mul2) * g2y;
    const float mul3 = vm[2] * mx + vm[6] * my + vm[10] * mz + vm[14];
    dmx += (vm[2] - vm[3] * mul3) * g_depth;
    dmy += (vm[6] - vm[7] * mul3) * g_depth;
    dmz += (vm[10] - vm[11] * mul3) * g_depth;

    // ---- SH backward, backward.cu:20-139 ----
    if (shs != nullptr) {
        const float ox = mx - s_cam[32], oy = my - s_cam[33], oz = mz - s_cam[34];
        const float len = sqrtf(ox * ox + oy * oy + oz * oz);
        const float x = ox / len, y = oy / len, z = oz / len;
        const uint32_t cl = clamped[idx];
        const float r0 = (cl & 1) ? 0.f : gc0, r1 = (cl & 2) ? 0.f : gc1, r2 = (cl & 4) ? 0.f : gc2;
        const float* sh = shs + (size_t)idx * M * 3;
        float w[16];
        float ddx = 0.f, ddy = 0.f, ddz = 0.f;  // dL/ddir
        // s[k] = sh_k . dL/dRGB for the active coefficients.  Fast path (the HT layout, M == 16, 16-byte
        // aligned): twelve 128-bit loads, dots formed on the fly so no 48-float copy lives in registers.
        float sd[16];
#pragma unroll
        for (int k = 0; k < 16; k++) sd[k] = 0.f;
        const int nact = (D + 1) * (D + 1);
        const bool fast = !RAW && (M == 16) && ((reinterpret_cast<uintptr_t>(shs) & 15) == 0);
        if (RAW) {
            float e[RAW_REST];
#pragma unroll
            for (int k = 0; k < RAW_REST; k++) e[k] = 0.f;
            if (D > 0) load_rest_row(shs_rest, idx, 3 * nact - 3, e);
            sd[0] = __ldg(shs + 3 * (size_t)idx) * r0 + __ldg(shs + 3 * (size_t)idx + 1) * r1 + __ldg(shs + 3 * (size_t)idx + 2) * r2;
#pragma unroll
            for (int k = 1; k < 16; k++) sd[k] = e[3 * k - 3] * r0 + e[3 * k - 2] * r1 + e[3 * k - 1] * r2;
        } else if (fast && ((reinterpret_cast<uintptr_t>(shs) & 31) == 0)) {
#pragma unroll
            for (int i = 0; i < 6; i++) {
                if (8 * i < 3 * nact) {
                    float e[8];
                    ldg256_nc(sh + 8 * i, e);
#pragma unroll
                    for (int t = 0; t < 8; t++) {
                        const int el = 8 * i + t, k = el / 3, c = el - 3 * k;
                        sd[k] = fmaf(e[t], c == 0 ? r0 : (c == 1 ? r1 : r2), sd[k]);
                    }
                }
            }
        } else if (fast) {
            const float4* sh4 = reinterpret_cast<const float4*>(sh);
#pragma unroll
            for (int i = 0; i < 12; i++) {
                if (4 * i < 3 * nact) {
                    const float4 v = __ldg(sh4 + i);
                    const float e[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
                    for (int t = 0; t < 4; t++) {
                        const int el = 4 * i + t, k = el / 3, c = el - 3 * k;
                        sd[k] = fmaf(e[t], c == 0 ? r0 : (c == 1 ? r1 : r2), sd[k]);
                    }
                }
            }
        } else {
#pragma unroll
            for (int k = 0; k < 16; k++)
                if (k < nact) sd[k] = sh[3 * k] * r0 + sh[3 * k + 1] * r1 + sh[3 * k + 2] * r2;
        }
        // Basis values w[k] (what dL/dsh_k is proportional to) and the gradient with respect to the view direction.
        // The direction gradient is sum_k t_k * grad B_k(x,y,z) with t_k = C_k (sh_k . dL/dRGB); the B_k are the
        // degree-1..3 polynomials of the real SH basis, so the sum is collected monomial by monomial of grad B_k
        // instead of basis function by basis function.  t_k stays 0 above the active degree.
        float t[16];
#pragma unroll
        for (int k = 0; k < 16; k++) t[k] = 0.f;
        w[0] = HTGS_SH_C0;
        int ncoef = 1;
        if (D > 0) {
            ncoef = 4;
            w[1] = -HTGS_SH_C1 * y; w[2] = HTGS_SH_C1 * z; w[3] = -HTGS_SH_C1 * x;
            t[1] = -HTGS_SH_C1 * sd[1]; t[2] = HTGS_SH_C1 * sd[2]; t[3] = -HTGS_SH_C1 * sd[3];
            const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
            const float xx_m_yy = xx - yy, zz4_m_r = 4.f * zz - xx - yy;  // recurring factors of the degree-2/3 basis
            if (D > 1) {
                ncoef = 9;
                w[4] = HTGS_SH_C2_0 * xy; w[5] = HTGS_SH_C2_1 * yz; w[6] = HTGS_SH_C2_2 * (2.f * zz - xx - yy);
                w[7] = HTGS_SH_C2_3 * xz; w[8] = HTGS_SH_C2_4 * xx_m_yy;
                t[4] = HTGS_SH_C2_0 * sd[4]; t[5] = HTGS_SH_C2_1 * sd[5]; t[6] = HTGS_SH_C2_2 * sd[6];
                t[7] = HTGS_SH_C2_3 * sd[7]; t[8] = HTGS_SH_C2_4 * sd[8];
            }
            if (D > 2) {
                ncoef = 16;
                w[9] = HTGS_SH_C3_0 * y * (3.f * xx - yy);
                w[10] = HTGS_SH_C3_1 * xy * z;
                w[11] = HTGS_SH_C3_2 * y * zz4_m_r;
                w[12] = HTGS_SH_C3_3 * z * (2.f * zz - 3.f * (xx + yy));
                w[13] = HTGS_SH_C3_4 * x * zz4_m_r;
                w[14] = HTGS_SH_C3_5 * z * xx_m_yy;
                w[15] = HTGS_SH_C3_6 * x * (xx - 3.f * yy);
                t[9] = HTGS_SH_C3_0 * sd[9]; t[10] = HTGS_SH_C3_1 * sd[10]; t[11] = HTGS_SH_C3_2 * sd[11];
                t[12] = HTGS_SH_C3_3 * sd[12]; t[13] = HTGS_SH_C3_4 * sd[13]; t[14] = HTGS_SH_C3_5 * sd[14];
                t[15] = HTGS_SH_C3_6 * sd[15];
            }
            // d/dx: linear part, then the monomials 1, x, y, z, xy, yz, xz, and the two quadratic forms
            ddx = t[3] + y * t[4] + z * t[7] + 2.f * x * (t[8] - t[6])
                + xy * (6.f * t[9] - 2.f * t[11]) + yz * t[10] + xz * (2.f * t[14] - 6.f * t[12])
                + t[13] * (zz4_m_r - 2.f * xx) + 3.f * t[15] * xx_m_yy;
            ddy = t[1] + x * t[4] + z * t[5] - 2.f * y * (t[6] + t[8])
                + 3.f * t[9] * xx_m_yy + xz * t[10] + t[11] * (zz4_m_r - 2.f * yy)
                - yz * (6.f * t[12] + 2.f * t[14]) - xy * (2.f * t[13] + 6.f * t[15]);
            ddz = t[2] + y * t[5] + 4.f * z * t[6] + x * t[7]
                + xy * t[10] + 8.f * (yz * t[11] + xz * t[13]) + 3.f * t[12] * (2.f * zz - xx - yy) + t[14] * xx_m_yy;
        }
        float* o = dL_dsh + (size_t)idx * M * 3;
        if (RAW) {
#pragma unroll
            for (int k = 1; k < 16; k++)
                if (k >= ncoef) w[k] = 0.f;
            store3(dL_dsh + 3 * (size_t)idx, w[0] * r0, w[0] * r1, w[0] * r2);
            float e[RAW_REST];
#pragma unroll
            for (int k = 0; k < RAW_REST; k++) {
                const int c = k % 3;
                e[k] = w[1 + k / 3] * (c == 0 ? r0 : (c == 1 ? r1 : r2));
            }
            store_rest_row(dL_dsh_rest, idx, e);
        } else if (dL_dsh == nullptr) {
            // compact SH mode (htgs_backward_compact): the expanded 192 B/Gaussian gradient is not written here at all; the
            // clamp-masked colour gradient it is the outer product of has already left through rgb_extract_kernel
        } else if (fast && sh_vec256) {
            // six 256-bit stores; coefficient and channel of every element are compile-time
#pragma unroll
            for (int k = 1; k < 16; k++)
                if (k >= ncoef) w[k] = 0.f;
#pragma unroll
            for (int i = 0; i < 6; i++) {
                float e[8];
#pragma unroll
                for (int t = 0; t < 8; t++) {
                    const int el = 8 * i + t, k = el / 3, c = el - 3 * k;
                    e[t] = w[k] * (c == 0 ? r0 : (c == 1 ? r1 : r2));
                }
                stg256(o + 8 * i, e);
            }
        } else if (fast && sh_vec) {
            // twelve 128-bit stores; coefficient and channel of every element are compile-time
#pragma unroll
            for (int k = 1; k < 16; k++)
                if (k >= ncoef) w[k] = 0.f;
            float4* o4 = reinterpret_cast<float4*>(o);
#pragma unroll
            for (int i = 0; i < 12; i++) {
                float e[4];
#pragma unroll
                for (int t = 0; t < 4; t++) {
                    const int el = 4 * i + t, k = el / 3, c = el - 3 * k;
                    e[t] = w[k] * (c == 0 ? r0 : (c == 1 ? r1 : r2));
                }
                o4[i] = make_float4(e[0], e[1], e[2], e[3]);
            }
        } else
            write_sh_grad(o, w, ncoef, M, r0, r1, r2);
        // dnormvdv, auxiliary.h:107-117
        const float sum2 = ox * ox + oy * oy + oz * oz;
        const float invsum32 = 1.0f / sqrtf(sum2 * sum2 * sum2);
        dmx += ((+sum2 - ox * ox) * ddx - oy * ox * ddy - oz * ox * ddz) * invsum32;
        dmy += (-ox * oy * ddx + (sum2 - oy * oy) * ddy - oz * oy * ddz) * invsum32;
        dmz += (-ox * oz * ddx - oy * oz * ddy + (sum2 - oz * oz) * ddz) * invsum32;
    }

    // ---- scale / rotation backward (what backward.cu:278-341 computes) ----
    // Sigma = L L^T with L = Rm diag(s), Rm the rotation matrix of the (un-normalised) quaternion and
    // s = scale_modifier * scale.  With dS the symmetric matrix form of the packed dL/dSigma (off-diagonals halved),
    //   dL/dL = 2 dS L,   dL/ds_j = sum_i (dL/dL)_ij Rm_ij,   H = dL/dRm = (dL/dL) diag(s).
    // Rm = (1 - 2|v|^2) I + 2 v v^T + 2 r [v]x  for q = (r, v), hence
    //   dL/dr = 2 v . a,   dL/dv = 2 r a + 2 (H + H^T) v - 4 tr(H) v,   a = axial vector of H - H^T.
    // Like the reference, dL/dscale is the derivative with respect to s (the scale_modifier factor is not applied).
    if (scales != nullptr) {
        const float r = q.x, x = q.y, y = q.z, z = q.w;
        const float Rm[3][3] = {{1.f - 2.f * (y * y + z * z), 2.f * (x * y - r * z), 2.f * (x * z + r * y)},
                                {2.f * (x * y + r * z), 1.f - 2.f * (x * x + z * z), 2.f * (y * z - r * x)},
                                {2.f * (x * z - r * y), 2.f * (y * z + r * x), 1.f - 2.f * (x * x + y * y)}};
        const float s[3] = {scale_modifier * sc0, scale_modifier * sc1, scale_modifier * sc2};
        const float dS[3][3] = {{dcov[0], 0.5f * dcov[1], 0.5f * dcov[2]},
                                {0.5f * dcov[1], dcov[3], 0.5f * dcov[4]},
                                {0.5f * dcov[2], 0.5f * dcov[4], dcov[5]}};
        float Hm[3][3];
        float dsc[3] = {0.f, 0.f, 0.f};
#pragma unroll
        for (int i = 0; i < 3; i++)
#pragma unroll
            for (int j = 0; j < 3; j++) {
                const float gl = 2.f * s[j] * (dS[i][0] * Rm[0][j] + dS[i][1] * Rm[1][j] + dS[i][2] * Rm[2][j]);  // (dL/dL)_ij
                dsc[j] += gl * Rm[i][j];
                Hm[i][j] = gl * s[j];
            }
        const float ax = Hm[2][1] - Hm[1][2], ay = Hm[0][2] - Hm[2][0], az = Hm[1][0] - Hm[0][1];
        const float tr4 = 4.f * (Hm[0][0] + Hm[1][1] + Hm[2][2]);
        const float hx = 2.f * Hm[0][0] * x + (Hm[0][1] + Hm[1][0]) * y + (Hm[0][2] + Hm[2][0]) * z;  // (H + H^T) v
        const float hy = (Hm[0][1] + Hm[1][0]) * x + 2.f * Hm[1][1] * y + (Hm[1][2] + Hm[2][1]) * z;
        const float hz = (Hm[0][2] + Hm[2][0]) * x + (Hm[1][2] + Hm[2][1]) * y + 2.f * Hm[2][2] * z;
        float4 dq;
        dq.x = 2.f * (x * ax + y * ay + z * az);
        dq.y = 2.f * (r * ax + hx) - tr4 * x;
        dq.z = 2.f * (r * ay + hy) - tr4 * y;
        dq.w = 2.f * (r * az + hz) - tr4 * z;
        if (RAW) {
            // exp backward; F.normalize backward: (g - qhat (qhat . g)) / n
            dsc[0] *= sc0; dsc[1] *= sc1; dsc[2] *= sc2;
            const float dot = q.x * dq.x + q.y * dq.y + q.z * dq.z + q.w * dq.w;
            const float inv_n = 1.0f / q_norm;
            dq = make_float4((dq.x - q.x * dot) * inv_n, (dq.y - q.y * dot) * inv_n, (dq.z - q.z * dot) * inv_n,
                             (dq.w - q.w * dot) * inv_n);
            (void)q_raw;
        }
        store3(dL_dscales + 3 * (size_t)idx, dsc[0], dsc[1], dsc[2]);
        reinterpret_cast<float4*>(dL_drotations)[idx] = dq;
    }

    store3(dL_dmeans2D + 3 * (size_t)idx, g2x, g2y, 0.f);
    if (RAW) {
        const float sg = act_opacity(opacity_raw[idx]);
        dL_dopacity[idx] = g_opacity * sg * (1.0f - sg);  // sigmoid backward
    } else
        dL_dopacity[idx] = g_opacity;
    store3(dL_dmeans3D + 3 * (size_t)idx, dmx, dmy, dmz);
    if (dL_dcolors) {
        if (colors_are_precomp) store3(dL_dcolors + 3 * (size_t)idx, gc0, gc1, gc2);
        else store3(dL_dcolors + 3 * (size_t)idx, gc0, gc1, gc2);
    }
    if (dL_dcov3D) {
#pragma unroll
        for (int k = 0; k < 6; k++) dL_dcov3D[6 * (size_t)idx + k] = dcov[k];
    }
}

int launch_geom_backward(int P, int D, int M, const float* means3D, const int* radii, const float* shs,
                         const float* shs_rest, const float* opacity_raw, float* dL_dsh_rest, const float* colors_precomp, const float* scales, float scale_modifier,
                         const float* rotations, const float* cov3D_precomp, const float* viewmatrix,
                         const float* projmatrix, const float* campos, int W, int H, float tanfovx, float tanfovy,
                         const GeomView& g, const float* gacc, float* dL_dmeans2D, float* dL_dopacity,
                         float* dL_dcolors, float* dL_dmeans3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscales,
                         float* dL_drotations, cudaStream_t st) {
    const float h_y = H / (2.0f * tanfovy), h_x = W / (2.0f * tanfovx);  // rasterizer_impl.cu:387-388
    if (shs_rest != nullptr)  // raw-parameter entry point
        geom_backward_kernel<true><<<(P + GB_BLOCK - 1) / GB_BLOCK, GB_BLOCK, 0, st>>>(
            P, D, M, means3D, radii, shs, shs_rest, opacity_raw, dL_dsh_rest, g.clamped, false, scales, scale_modifier,
            rotations, cov3D_precomp, viewmatrix, projmatrix, campos, h_x, h_y, tanfovx, tanfovy, gacc, dL_dmeans2D,
            dL_dopacity, dL_dcolors, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations);
    else
        geom_backward_kernel<false><<<(P + GB_BLOCK - 1) / GB_BLOCK, GB_BLOCK, 0, st>>>(
            P, D, M, means3D, radii, shs, nullptr, nullptr, nullptr, g.clamped, colors_precomp != nullptr, scales,
            scale_modifier, rotations, cov3D_precomp, viewmatrix, projmatrix, campos, h_x, h_y, tanfovx, tanfovy, gacc,
            dL_dmeans2D, dL_dopacity, dL_dcolors, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations);
    HTGS_LAUNCH_CHECK("geom_backward_kernel");
    return 0;
}

}  // namespace htgs
