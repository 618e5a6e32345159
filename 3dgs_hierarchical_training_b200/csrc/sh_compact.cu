// View-parallel training without shipping the expanded SH gradient (no reference counterpart).
//
// 81 % of the 236 B/Gaussian gradient that a view-parallel step all-reduces is dL/dsh, and per view that tensor is
// rank one: dL/dsh[i][k][c] = basis_k(direction of Gaussian i seen from the view's camera) * r[i][c], with r the
// clamp-masked colour gradient (3 floats).  So a rank only has to make r and its camera position known to the others
// (12 B/Gaussian/view instead of 192 B/Gaussian), and every rank rebuilds and sums the outer products itself, in the
// same fixed view order -- replicas stay bit-identical, the all-reduce shrinks to the remaining 11 floats/Gaussian.
//
//   rgb_extract_kernel   after the blend backward: r = dL/dcolour from the accumulator rows, zeroed where the forward
//                        clamped the channel (forward.cu:63-70), written to this view's slot -- either a plain store
//                        (peers pull it later / NCCL all-gathers it) or multimem.st through an NVSwitch multicast
//                        address, which lands it in every rank's copy of the slot at once (push).
//   sh_expand_kernel     dL/dsh[i] = sum over views v of basis(mean_i - campos_v) (x) r_v[i], views in slot order; slots
//                        are read through plain pointers (local, or a peer's mapped memory: NVLink peer loads).
//
// Slot layout (floats): [0 .. 3*P4) r of Gaussians 0..P4-1 (P4 = P rounded up to 4), [3*P4 .. 3*P4+4) camera position + pad.
#include "../../include/htgs.h"
#include "sh_basis.cuh"

namespace htgs {

constexpr int SHC_MAX_VIEWS = 16;
struct SlotPtrs {
    const float* p[SHC_MAX_VIEWS];
};

__device__ __forceinline__ void store4(float* a, float4 v, bool mc) {
    if (mc)
        asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(a), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
    else
        *reinterpret_cast<float4*>(a) = v;
}

// one thread per 4 Gaussians: 4 accumulator rows in, three 128-bit stores out
__global__ void __launch_bounds__(256)
rgb_extract_kernel(int P, const float* __restrict__ gacc, const uint8_t* __restrict__ clamped, const float* __restrict__ campos,
                   float* __restrict__ slot, bool multicast) {
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    const int P4 = (P + 3) & ~3;
    if (q == 0) store4(slot + 3 * (size_t)P4, make_float4(campos[0], campos[1], campos[2], 0.f), multicast);
    const int i0 = 4 * q;
    if (i0 >= P) return;
    float r[12];
#pragma unroll
    for (int j = 0; j < 4; j++) {
        const int i = i0 + j;
        float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
        uint32_t cl = 0;
        if (i < P) {
            g = reinterpret_cast<const float4*>(gacc + (size_t)i * 12)[2];
            cl = clamped[i];
        }
        r[3 * j] = (cl & 1) ? 0.f : g.x;
        r[3 * j + 1] = (cl & 2) ? 0.f : g.y;
        r[3 * j + 2] = (cl & 4) ? 0.f : g.z;
    }
    float* o = slot + 3 * (size_t)i0;
    store4(o, make_float4(r[0], r[1], r[2], r[3]), multicast);
    store4(o + 4, make_float4(r[4], r[5], r[6], r[7]), multicast);
    store4(o + 8, make_float4(r[8], r[9], r[10], r[11]), multicast);
}

__global__ void __launch_bounds__(128)
sh_expand_kernel(int P, int D, int M, int n_views, const float* __restrict__ means3D, const __grid_constant__ SlotPtrs slots,
                 float* __restrict__ dL_dsh) {
    __shared__ float s_cam[SHC_MAX_VIEWS][4];
    const int P4 = (P + 3) & ~3;
    if (threadIdx.x < n_views * 4) s_cam[threadIdx.x >> 2][threadIdx.x & 3] = slots.p[threadIdx.x >> 2][3 * (size_t)P4 + (threadIdx.x & 3)];
    __syncthreads();
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= P) return;
    const float mx = means3D[3 * (size_t)idx], my = means3D[3 * (size_t)idx + 1], mz = means3D[3 * (size_t)idx + 2];
    float acc[48];
#pragma unroll
    for (int e = 0; e < 48; e++) acc[e] = 0.f;
    for (int v = 0; v < n_views; v++) {
        const float* rp = slots.p[v] + 3 * (size_t)idx;
        const float r0 = rp[0], r1 = rp[1], r2 = rp[2];
        if (r0 == 0.f && r1 == 0.f && r2 == 0.f) continue;  // culled or unseen in this view
        const float ox = mx - s_cam[v][0], oy = my - s_cam[v][1], oz = mz - s_cam[v][2];
        const float len = sqrtf(ox * ox + oy * oy + oz * oz);  // the direction the forward used: (mean - campos) / length
        float w[16];
#pragma unroll
        for (int k = 0; k < 16; k++) w[k] = 0.f;
        sh_basis(D, ox / len, oy / len, oz / len, w);
#pragma unroll
        for (int k = 0; k < 16; k++) {
            acc[3 * k] = fmaf(w[k], r0, acc[3 * k]);
            acc[3 * k + 1] = fmaf(w[k], r1, acc[3 * k + 1]);
            acc[3 * k + 2] = fmaf(w[k], r2, acc[3 * k + 2]);
        }
    }
    float* o = dL_dsh + (size_t)idx * M * 3;
    if (M == 16 && (reinterpret_cast<uintptr_t>(dL_dsh) & 15) == 0) {
#pragma unroll
        for (int i = 0; i < 12; i++) reinterpret_cast<float4*>(o)[i] = make_float4(acc[4 * i], acc[4 * i + 1], acc[4 * i + 2], acc[4 * i + 3]);
    } else {
#pragma unroll
        for (int e = 0; e < 48; e++)
            if (e < 3 * M) o[e] = acc[e];
        for (int e = 48; e < 3 * M; e++) o[e] = 0.f;
    }
}

int launch_rgb_extract(int P, const float* gacc, const uint8_t* clamped, const float* campos, float* slot, bool multicast,
                       cudaStream_t st) {
    const int quads = (P + 3) / 4;
    rgb_extract_kernel<<<(quads + 255) / 256, 256, 0, st>>>(P, gacc, clamped, campos, slot, multicast);
    HTGS_LAUNCH_CHECK("rgb_extract_kernel");
    return 0;
}

}  // namespace htgs

using namespace htgs;

extern "C" size_t htgs_sh_slot_floats(int P) { return 3 * (size_t)((P + 3) & ~3) + 4; }

extern "C" int htgs_sh_expand(int P, int D, int M, int n_views, const float* means3D, const float* const* slots_host,
                              float* dL_dsh, void* stream) {
    const char* who = "htgs_sh_expand";
    if (P <= 0) return 0;
    if (n_views < 1 || n_views > SHC_MAX_VIEWS) { set_error("%s: n_views=%d outside 1..%d", who, n_views, SHC_MAX_VIEWS); return 1; }
    if (D < 0 || D > 3 || M < (D + 1) * (D + 1)) { set_error("%s: bad SH degree %d / M %d", who, D, M); return 1; }
    if (!means3D || !slots_host || !dL_dsh) { set_error("%s: null pointer", who); return 1; }
    SlotPtrs sp;
    for (int v = 0; v < SHC_MAX_VIEWS; v++) sp.p[v] = nullptr;
    for (int v = 0; v < n_views; v++) {
        if (!slots_host[v] || (reinterpret_cast<size_t>(slots_host[v]) & 15)) { set_error("%s: slot %d null or not 16-byte aligned", who, v); return 1; }
        sp.p[v] = slots_host[v];
    }
    sh_expand_kernel<<<(P + 127) / 128, 128, 0, reinterpret_cast<cudaStream_t>(stream)>>>(P, D, M, n_views, means3D, sp, dL_dsh);
    HTGS_LAUNCH_CHECK("sh_expand_kernel");
    return 0;
}
