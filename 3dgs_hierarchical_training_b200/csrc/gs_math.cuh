// Per-Gaussian projection math with an explicit, contraction-proof operation
// order.  radii, tile rectangles and the depth bits of the sort keys must be
// bit-identical to the reference, whose rounding is fixed by how nvcc AND ptxas
// contract its glm expressions (SASS of cuda_rasterizer/forward.cu:74-256,
// auxiliary.h:41-77 for sm_100a at -O3, default -fmad; the PTX alone is not enough:
// mul.f32/add.f32 without .rn are fused further by ptxas).  Writing every operation as an _rn intrinsic stops
// the compiler from fusing or re-associating anything here.
#pragma once
#include "common.cuh"

namespace htgs {

#define FM(a, b, c) __fmaf_rn((a), (b), (c))
#define MU(a, b) __fmul_rn((a), (b))
#define AD(a, b) __fadd_rn((a), (b))
#define SB(a, b) __fsub_rn((a), (b))
#define DV(a, b) __fdiv_rn((a), (b))

// 256-bit global accesses (sm_100: LDG.E.ENL2.256 / STG.E.ENL2.256).  The 192-byte SH rows are read and written by one
// lane per Gaussian; with 128-bit accesses every instruction touches half of a 32-byte sector per lane, so the memory
// system sees twice the sector requests (and, for stores, twice the partial-sector writes) the data needs.  One lane-
// access of 32 bytes is exactly one sector.  Addresses must be 32-byte aligned.
__device__ __forceinline__ void ldg256_nc(const float* p, float* v) {
    asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7]) : "l"(p));
}
__device__ __forceinline__ void stg256(float* p, const float* v) {
    asm volatile("st.global.v8.f32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "f"(v[0]), "f"(v[1]), "f"(v[2]), "f"(v[3]), "f"(v[4]),
                 "f"(v[5]), "f"(v[6]), "f"(v[7]) : "memory");
}

// row r of the row-vector 4x4 transform: m[12+r] + fma(z, m[8+r], fma(x, m[r], y*m[4+r]))
// (auxiliary.h:58-77)
__device__ __forceinline__ float xform_row(const float* __restrict__ m, int r, float x, float y, float z) {
    return AD(m[12 + r], FM(z, m[8 + r], FM(x, m[r], MU(y, m[4 + r]))));
}

// auxiliary.h:41-44 -- evaluated in fp64 because of the double literals.
__device__ __forceinline__ float ndc2pix(float v, int S) {
    double t = __fma_rn(__dadd_rn((double)v, 1.0), (double)S, -1.0);
    return (float)__dmul_rn(t, 0.5);
}

// auxiliary.h:46-56
__device__ __forceinline__ void get_rect(float px, float py, int radius, int gx, int gy, uint2& rmin, uint2& rmax) {
    const float r = (float)radius;
    int x0 = (int)MU(SB(px, r), 0.0625f);
    int y0 = (int)MU(SB(py, r), 0.0625f);
    int x1 = (int)MU(AD(AD(AD(px, r), 16.0f), -1.0f), 0.0625f);
    int y1 = (int)MU(AD(AD(AD(py, r), 16.0f), -1.0f), 0.0625f);
    rmin.x = min((unsigned)gx, (unsigned)max(x0, 0));
    rmin.y = min((unsigned)gy, (unsigned)max(y0, 0));
    rmax.x = min((unsigned)gx, (unsigned)max(x1, 0));
    rmax.y = min((unsigned)gy, (unsigned)max(y1, 0));
}

// forward.cu:118-152.  q = (r,x,y,z) is used un-normalised (forward.cu:127).
__device__ __forceinline__ void cov3d_from_scale_rot(float s0, float s1, float s2, float mod, float4 q, float* cov) {
    const float sx = MU(mod, s0), sy = MU(mod, s1), sz = MU(mod, s2);
    const float r = q.x, x = q.y, y = q.z, z = q.w;
    // ptxas contracts the reference's mul/add pairs here: one product of each pair is rounded
    // (xz, rx, rz), the other is fused (SASS of preprocessCUDA, sm_100a).
    const float yy = MU(y, y), zz = MU(z, z);
    const float xz = MU(x, z), rx = MU(r, x), rz = MU(r, z);
    const float yy_zz = AD(yy, zz);
    const float xx_zz = FM(x, x, zz);
    const float xx_yy = FM(x, x, yy);
    const float d0 = FM(x, y, -rz), d1 = FM(r, y, xz), d2 = FM(x, y, rz), d3 = FM(y, z, -rx), d4 = FM(-r, y, xz),
                d5 = FM(y, z, rx);
    const float R00 = SB(1.0f, AD(yy_zz, yy_zz));
    const float R01 = AD(d0, d0);
    const float R02 = AD(d1, d1);
    const float R10 = AD(d2, d2);
    const float R11 = SB(1.0f, AD(xx_zz, xx_zz));
    const float R12 = AD(d3, d3);
    const float R20 = AD(d4, d4);
    const float R21 = AD(d5, d5);
    const float R22 = SB(1.0f, AD(xx_yy, xx_yy));
    const float a0 = MU(sx, R00), a1 = MU(sy, R01), a2 = MU(sz, R02);
    const float b0 = MU(sx, R10), b1 = MU(sy, R11), b2 = MU(sz, R12);
    const float c0 = MU(sx, R20), c1 = MU(sy, R21), c2 = MU(sz, R22);
    cov[0] = FM(a2, a2, FM(a0, a0, MU(a1, a1)));
    cov[1] = FM(b2, a2, FM(b0, a0, MU(b1, a1)));
    cov[2] = FM(c2, a2, FM(c0, a0, MU(c1, a1)));
    cov[3] = FM(b2, b2, FM(b0, b0, MU(b1, b1)));
    cov[4] = FM(c2, b2, FM(c0, b0, MU(c1, b1)));
    cov[5] = FM(c2, c2, FM(c0, c0, MU(c1, c1)));
}

struct Ewa {
    float tx, ty, tz;    // view-space point with the 1.3*tanfov clamp applied to x,y
    float txtz, tytz;    // unclamped ratios
    float limx, limy;
    float T00, T01, T02, T10, T11, T12;  // glm T[0][0..2], T[1][0..2] of T = W*J
};

// forward.cu:80-99 (and its twin backward.cu:166-192)
__device__ __forceinline__ Ewa ewa_setup(float px, float py, float pz, const float* __restrict__ vm, float fx, float fy,
                                         float tanfovx, float tanfovy) {
    Ewa e;
    const float tx = xform_row(vm, 0, px, py, pz);
    const float ty = xform_row(vm, 1, px, py, pz);
    const float tz = xform_row(vm, 2, px, py, pz);
    e.limx = MU(tanfovx, 1.3f);
    e.limy = MU(tanfovy, 1.3f);
    e.txtz = DV(tx, tz);
    e.tytz = DV(ty, tz);
    const float cx = fminf(e.limx, fmaxf(-e.limx, e.txtz));
    const float cy = fminf(e.limy, fmaxf(-e.limy, e.tytz));
    e.tx = MU(cx, tz);
    e.ty = MU(cy, tz);
    e.tz = tz;
    const float tz2 = MU(tz, tz);
    const float J00 = DV(fx, tz);
    const float J02 = DV(MU(fx, MU(cx, -tz)), tz2);
    const float J11 = DV(fy, tz);
    const float J12 = DV(MU(fy, MU(cy, -tz)), tz2);
    e.T00 = FM(vm[2], J02, MU(vm[0], J00));
    e.T01 = FM(vm[6], J02, MU(vm[4], J00));
    e.T02 = FM(J02, vm[10], MU(vm[8], J00));
    e.T10 = FM(vm[2], J12, MU(J11, vm[1]));
    e.T11 = FM(vm[6], J12, MU(J11, vm[5]));
    e.T12 = FM(J12, vm[10], MU(J11, vm[9]));
    return e;
}

// forward.cu:101-112: cov = T^T Vrk^T T, then the 0.3 px low-pass on the diagonal.
__device__ __forceinline__ void ewa_cov2d(const Ewa& e, const float* c, float& a, float& b, float& cc) {
    const float u0 = FM(e.T02, c[2], FM(e.T00, c[0], MU(e.T01, c[1])));
    const float v0 = FM(e.T12, c[2], FM(e.T10, c[0], MU(e.T11, c[1])));
    const float u1 = FM(e.T02, c[4], FM(e.T00, c[1], MU(e.T01, c[3])));
    const float v1 = FM(e.T12, c[4], FM(e.T10, c[1], MU(e.T11, c[3])));
    const float u2 = FM(e.T02, c[5], FM(e.T00, c[2], MU(e.T01, c[4])));
    const float v2 = FM(e.T12, c[5], FM(e.T10, c[2], MU(e.T11, c[4])));
    const float c00 = FM(e.T02, u2, FM(e.T00, u0, MU(e.T01, u1)));
    const float c01 = FM(e.T02, v2, FM(e.T00, v0, MU(e.T01, v1)));
    const float c11 = FM(e.T12, v2, FM(e.T10, v0, MU(e.T11, v1)));
    a = AD(c00, 0.3f);
    b = c01;
    cc = AD(c11, 0.3f);
}

// The Gaussian falloff exponent of one (pixel, Gaussian) pair, in the exact
// order nvcc gives forward.cu:339 / backward.cu:525:
//   power = fma(fma(dx, dx*cx, dy*(dy*cz)), -0.5f, -(dy*(dx*cy)))
// (the PTX shows mul/sub; ptxas fuses them -- SASS: FFMA R14, R14, -0.5, -R29)
__device__ __forceinline__ float pair_power(float dx, float dy, float cx, float cy, float cz) {
    const float q = FM(dx, MU(dx, cx), MU(dy, MU(dy, cz)));
    return FM(q, -0.5f, -MU(dy, MU(dx, cy)));
}

// SH basis constants (auxiliary.h:22-39)
#define HTGS_SH_C0 0.28209479177387814f
#define HTGS_SH_C1 0.4886025119029199f
#define HTGS_SH_C2_0 1.0925484305920792f
#define HTGS_SH_C2_1 -1.0925484305920792f
#define HTGS_SH_C2_2 0.31539156525252005f
#define HTGS_SH_C2_3 -1.0925484305920792f
#define HTGS_SH_C2_4 0.5462742152960396f
#define HTGS_SH_C3_0 -0.5900435899266435f
#define HTGS_SH_C3_1 2.890611442640554f
#define HTGS_SH_C3_2 -0.4570457994644658f
#define HTGS_SH_C3_3 0.3731763325901154f
#define HTGS_SH_C3_4 -0.4570457994644658f
#define HTGS_SH_C3_5 1.445305721320277f
#define HTGS_SH_C3_6 -0.5900435899266435f

}  // namespace htgs
