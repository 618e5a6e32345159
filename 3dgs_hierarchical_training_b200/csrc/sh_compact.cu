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
#include <mutex>

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

// 256 threads per 1024 Gaussians.  Each thread gathers r of 4 Gaussians (4 accumulator rows in) into shared memory; the CTA
// then writes its 3072 floats as 768 consecutive 128-bit stores, lane after lane -- full 128-byte lines per warp, which is
// what the NVLink multicast path wants (the first version let every thread store its own three float4: a 48-byte lane
// stride, a third of every line per instruction, and on 8 GPUs the push cost as much as the all-reduce it replaced).
constexpr int RGB_CTA_GAUSS = 1024;
__global__ void __launch_bounds__(256)
rgb_extract_kernel(int P, const float* __restrict__ gacc, const uint8_t* __restrict__ clamped, const float* __restrict__ campos,
                   float* __restrict__ slot, bool multicast) {
    __shared__ __align__(16) float s_r[3 * RGB_CTA_GAUSS];
    const int P4 = (P + 3) & ~3;
    if (blockIdx.x == 0 && threadIdx.x == 0) store4(slot + 3 * (size_t)P4, make_float4(campos[0], campos[1], campos[2], 0.f), multicast);
    const int base = blockIdx.x * RGB_CTA_GAUSS;
#pragma unroll
    for (int j = 0; j < 4; j++) {
        const int li = j * 256 + threadIdx.x;  // consecutive lanes read consecutive accumulator rows
        const int i = base + li;
        float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
        uint32_t cl = 0;
        if (i < P) {
            g = reinterpret_cast<const float4*>(gacc + (size_t)i * 12)[2];
            cl = clamped[i];
        }
        s_r[3 * li] = (cl & 1) ? 0.f : g.x;
        s_r[3 * li + 1] = (cl & 2) ? 0.f : g.y;
        s_r[3 * li + 2] = (cl & 4) ? 0.f : g.z;
    }
    __syncthreads();
    const int n4 = (min(RGB_CTA_GAUSS, P4 - base) * 3) / 4;  // float4 this CTA owns (P4 is a multiple of 4, so is the count)
    float* o = slot + 3 * (size_t)base;
    for (int q = threadIdx.x; q < n4; q += 256) store4(o + 4 * q, reinterpret_cast<const float4*>(s_r)[q], multicast);
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
    if (M == 16 && (reinterpret_cast<uintptr_t>(dL_dsh) & 31) == 0) {  // one 32-byte sector per lane and store (gs_math.cuh)
#pragma unroll
        for (int i = 0; i < 6; i++) stg256(o + 8 * i, acc + 8 * i);
    } else if (M == 16 && (reinterpret_cast<uintptr_t>(dL_dsh) & 15) == 0) {
#pragma unroll
        for (int i = 0; i < 12; i++) reinterpret_cast<float4*>(o)[i] = make_float4(acc[4 * i], acc[4 * i + 1], acc[4 * i + 2], acc[4 * i + 3]);
    } else {
#pragma unroll
        for (int e = 0; e < 48; e++)
            if (e < 3 * M) o[e] = acc[e];
        for (int e = 48; e < 3 * M; e++) o[e] = 0.f;
    }
}

// The extraction (and, with a multicast slot, the push over NVLink) needs only the blend backward's accumulator, so it
// runs on a side stream NEXT TO the geometry backward: forked from the caller's stream after the blend backward, joined
// back before the entry point returns control of the stream.  One side stream and one event pair per device, created
// on first use under a lock (process-global state, listed in include/htgs.h); compact-mode backwards on one device must
// therefore come from one stream at a time.
namespace {
struct SideStream {
    cudaStream_t stream = nullptr;
    cudaEvent_t fork = nullptr, join = nullptr;
};
SideStream g_side[64];
std::mutex g_side_mutex;

int side_for_current_device(SideStream** out) {
    int dev = -1;
    HTGS_CUDA_CHECK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) { set_error("compact SH mode: device ordinal %d not supported (max 63)", dev); return 1; }
    std::lock_guard<std::mutex> lock(g_side_mutex);
    SideStream& s = g_side[dev];
    if (!s.stream) {
        cudaStream_t st = nullptr;
        cudaEvent_t a = nullptr, b = nullptr;
        HTGS_CUDA_CHECK(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));
        HTGS_CUDA_CHECK(cudaEventCreateWithFlags(&a, cudaEventDisableTiming));
        HTGS_CUDA_CHECK(cudaEventCreateWithFlags(&b, cudaEventDisableTiming));
        s.fork = a; s.join = b; s.stream = st;
    }
    *out = &s;
    return 0;
}
}  // namespace

int launch_rgb_extract(int P, const float* gacc, const uint8_t* clamped, const float* campos, float* slot, bool multicast,
                       cudaStream_t st) {
    SideStream* side = nullptr;
    if (side_for_current_device(&side)) return 1;
    HTGS_CUDA_CHECK(cudaEventRecord(side->fork, st));
    HTGS_CUDA_CHECK(cudaStreamWaitEvent(side->stream, side->fork, 0));
    rgb_extract_kernel<<<(P + RGB_CTA_GAUSS - 1) / RGB_CTA_GAUSS, 256, 0, side->stream>>>(P, gacc, clamped, campos, slot, multicast);
    HTGS_LAUNCH_CHECK("rgb_extract_kernel");
    HTGS_CUDA_CHECK(cudaEventRecord(side->join, side->stream));
    return 0;
}

// makes the caller's stream wait for the side-stream work of the last launch_rgb_extract on this device
int join_rgb_extract(cudaStream_t st) {
    SideStream* side = nullptr;
    if (side_for_current_device(&side)) return 1;
    HTGS_CUDA_CHECK(cudaStreamWaitEvent(st, side->join, 0));
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
