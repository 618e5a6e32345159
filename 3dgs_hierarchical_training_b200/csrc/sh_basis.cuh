// Real spherical-harmonics basis of degree <= 3 in the reference's sign convention
// (cuda_rasterizer/auxiliary.h:22-39 for the constants, forward.cu:20-71 for how colour is formed from them):
// colour_c = 0.5 + sum_k w[k] * sh[k][c].  Shared by the kernels that need d(colour)/d(sh) = w[k].
#pragma once
#include "gs_math.cuh"

namespace htgs {

// w[0..ncoef) for the unit direction (x, y, z); returns ncoef = (D+1)^2.  Entries >= ncoef are left untouched.
__device__ __forceinline__ int sh_basis(int D, float x, float y, float z, float (&w)[16]) {
    w[0] = HTGS_SH_C0;
    if (D < 1) return 1;
    w[1] = -HTGS_SH_C1 * y;
    w[2] = HTGS_SH_C1 * z;
    w[3] = -HTGS_SH_C1 * x;
    if (D < 2) return 4;
    const float xx = x * x, yy = y * y, zz = z * z, xy = x * y, yz = y * z, xz = x * z;
    const float xx_m_yy = xx - yy, zz4_m_r = 4.f * zz - xx - yy;
    w[4] = HTGS_SH_C2_0 * xy;
    w[5] = HTGS_SH_C2_1 * yz;
    w[6] = HTGS_SH_C2_2 * (2.f * zz - xx - yy);
    w[7] = HTGS_SH_C2_3 * xz;
    w[8] = HTGS_SH_C2_4 * xx_m_yy;
    if (D < 3) return 9;
    w[9] = HTGS_SH_C3_0 * y * (3.f * xx - yy);
    w[10] = HTGS_SH_C3_1 * xy * z;
    w[11] = HTGS_SH_C3_2 * y * zz4_m_r;
    w[12] = HTGS_SH_C3_3 * z * (2.f * zz - 3.f * (xx + yy));
    w[13] = HTGS_SH_C3_4 * x * zz4_m_r;
    w[14] = HTGS_SH_C3_5 * z * xx_m_yy;
    w[15] = HTGS_SH_C3_6 * x * (xx - 3.f * yy);
    return 16;
}

}  // namespace htgs
