// Gradient all-reduce of the view-parallel step as one kernel over NVLink peer memory.
//
// Every rank holds the flat 59-float/Gaussian gradient in a buffer that all peers have mapped (torch symmetric memory
// does the allocation, the handle exchange and the cross-GPU barriers; this file is only the data path).  Two-shot:
// rank r reduces slice r -- it reads that slice from every peer with 128-bit peer loads, sums in the fixed order
// 0..N-1, and writes the result back into every peer's buffer with 128-bit peer stores.  Each element is reduced by
// exactly one rank and broadcast, so all replicas end with bit-identical gradients.  Per GPU the links carry
// (N-1)/N of the buffer in each direction, both directions busy at once; NCCL's ring/NVLS all-reduce moved the same
// 236 MB in 0.49 ms at 2 GPUs and 0.64 ms at 8, far from the 900 GB/s per direction the links offer.
#include "../../include/htgs.h"
#include "common.cuh"
#include <algorithm>
#include <atomic>

namespace htgs {

constexpr int AR_MAX_WORLD = 16;
constexpr int AR_THREADS = 512;  // 256-1024 threads, 4-8 float4 per thread, 1-8 CTAs per SM: all 0.36 ms at 2 GPUs (link-bound)
constexpr int AR_UNROLL = 4;
constexpr int AR_G = 2;

struct PeerPtrs {
    float4* p[AR_MAX_WORLD];
};

__device__ __forceinline__ float4 ld_sys(const float4* a) {
    float4 v;
    asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(a) : "memory");
    return v;
}
__device__ __forceinline__ void st_sys(float4* a, float4 v) {
    asm volatile("st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(a), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}

// NVSwitch multicast (NVLS) form of the same two-shot: `mc` is a multicast address that maps the buffer of every
// rank.  multimem.ld_reduce makes the switch fetch the element from all ranks and return their sum; multimem.st makes
// it write the result to all ranks.  Per GPU the links then carry about one buffer in each direction whatever the
// world size (the peer-load form carries 2(N-1)/N buffers), which is what pays from 4 GPUs up.
__global__ void __launch_bounds__(AR_THREADS)
allreduce_multimem_kernel(float4* __restrict__ mc, int rank, int world, long long n4) {
    const long long per = (n4 + world - 1) / world;
    const long long lo = (long long)rank * per;
    const long long hi = lo + per < n4 ? lo + per : n4;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i0 = lo + (long long)blockIdx.x * blockDim.x + threadIdx.x; i0 < hi; i0 += stride * AR_UNROLL) {
        float4 acc[AR_UNROLL];
#pragma unroll
        for (int u = 0; u < AR_UNROLL; u++) {
            const long long i = i0 + u * stride;
            if (i < hi)
                asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                             : "=f"(acc[u].x), "=f"(acc[u].y), "=f"(acc[u].z), "=f"(acc[u].w) : "l"(mc + i) : "memory");
        }
#pragma unroll
        for (int u = 0; u < AR_UNROLL; u++) {
            const long long i = i0 + u * stride;
            if (i < hi)
                asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(mc + i), "f"(acc[u].x), "f"(acc[u].y),
                             "f"(acc[u].z), "f"(acc[u].w) : "memory");
        }
    }
}

template <int WORLD>
__global__ void __launch_bounds__(AR_THREADS)
allreduce_two_shot_kernel(const __grid_constant__ PeerPtrs bufs, int rank, long long n4) {
    const long long per = (n4 + WORLD - 1) / WORLD;
    const long long lo = (long long)rank * per;
    const long long hi = lo + per < n4 ? lo + per : n4;
    const long long stride = (long long)gridDim.x * AR_THREADS;
    for (long long i0 = lo + (long long)blockIdx.x * AR_THREADS + threadIdx.x; i0 < hi; i0 += stride * AR_UNROLL) {
        float4 acc[AR_UNROLL];
#pragma unroll
        for (int u = 0; u < AR_UNROLL; u++) {
            const long long i = i0 + u * stride;
            if (i < hi) acc[u] = ld_sys(bufs.p[0] + i);
        }
#pragma unroll
        for (int p = 1; p < WORLD; p++) {
            float4 t[AR_UNROLL];
#pragma unroll
            for (int u = 0; u < AR_UNROLL; u++) {
                const long long i = i0 + u * stride;
                if (i < hi) t[u] = ld_sys(bufs.p[p] + i);
            }
#pragma unroll
            for (int u = 0; u < AR_UNROLL; u++) {
                acc[u].x += t[u].x; acc[u].y += t[u].y; acc[u].z += t[u].z; acc[u].w += t[u].w;
            }
        }
#pragma unroll
        for (int p = 0; p < WORLD; p++) {
#pragma unroll
            for (int u = 0; u < AR_UNROLL; u++) {
                const long long i = i0 + u * stride;
                if (i < hi) st_sys(bufs.p[p] + i, acc[u]);
            }
        }
    }
}

// launch shape of the multicast kernel; changed only through htgs_allreduce_set_launch_shape (tuning aid)
static std::atomic<int> g_mc_ctas_per_sm{AR_G}, g_mc_threads{AR_THREADS};

}  // namespace htgs

using namespace htgs;

extern "C" int htgs_allreduce_set_launch_shape(int ctas_per_sm, int threads) {
    if (ctas_per_sm < 1 || ctas_per_sm > 16 || threads < 32 || threads > AR_THREADS || (threads & 31)) {
        set_error("htgs_allreduce_set_launch_shape: need 1..16 CTAs per SM and 32..%d threads (multiple of 32)", AR_THREADS);
        return 1;
    }
    g_mc_ctas_per_sm.store(ctas_per_sm, std::memory_order_relaxed);
    g_mc_threads.store(threads, std::memory_order_relaxed);
    return 0;
}

extern "C" int htgs_allreduce_p2p(const void* const* peer_buffers, int rank, int world, long long numel, void* stream) {
    const char* who = "htgs_allreduce_p2p";
    if (world < 1 || world > AR_MAX_WORLD || rank < 0 || rank >= world) { set_error("%s: bad rank %d / world %d (max %d)", who, rank, world, AR_MAX_WORLD); return 1; }
    if (numel <= 0 || world == 1) return 0;
    if (numel % 4) { set_error("%s: numel must be a multiple of 4 (pad the buffer)", who); return 1; }
    if (!peer_buffers) { set_error("%s: null pointer table", who); return 1; }
    PeerPtrs b;
    for (int p = 0; p < world; p++) {
        if (!peer_buffers[p] || (reinterpret_cast<size_t>(peer_buffers[p]) & 15)) { set_error("%s: peer buffer %d null or not 16-byte aligned", who, p); return 1; }
        b.p[p] = reinterpret_cast<float4*>(const_cast<void*>(peer_buffers[p]));
    }
    for (int p = world; p < AR_MAX_WORLD; p++) b.p[p] = nullptr;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    const long long n4 = numel / 4;
    const int grid = 148 * AR_G;
    switch (world) {
        case 2: allreduce_two_shot_kernel<2><<<grid, AR_THREADS, 0, st>>>(b, rank, n4); break;
        case 3: allreduce_two_shot_kernel<3><<<grid, AR_THREADS, 0, st>>>(b, rank, n4); break;
        case 4: allreduce_two_shot_kernel<4><<<grid, AR_THREADS, 0, st>>>(b, rank, n4); break;
        case 5: allreduce_two_shot_kernel<5><<<grid, AR_THREADS, 0, st>>>(b, rank, n4); break;
        case 6: allreduce_two_shot_kernel<6><<<grid, AR_THREADS, 0, st>>>(b, rank, n4); break;
        case 7: allreduce_two_shot_kernel<7><<<grid, AR_THREADS, 0, st>>>(b, rank, n4); break;
        case 8: allreduce_two_shot_kernel<8><<<grid, AR_THREADS, 0, st>>>(b, rank, n4); break;
        default: set_error("%s: world size %d has no instantiation (2..8)", who, world); return 1;
    }
    HTGS_LAUNCH_CHECK("allreduce_two_shot_kernel");
    return 0;
}

extern "C" int htgs_allreduce_multimem(void* multicast_buffer, int rank, int world, long long numel, void* stream) {
    const char* who = "htgs_allreduce_multimem";
    if (world < 1 || rank < 0 || rank >= world) { set_error("%s: bad rank %d / world %d", who, rank, world); return 1; }
    if (numel <= 0 || world == 1) return 0;
    if (numel % 4) { set_error("%s: numel must be a multiple of 4 (pad the buffer)", who); return 1; }
    if (!multicast_buffer || (reinterpret_cast<size_t>(multicast_buffer) & 15)) { set_error("%s: multicast address null or not 16-byte aligned", who); return 1; }
    const int grid_mul = g_mc_ctas_per_sm.load(std::memory_order_relaxed), threads = g_mc_threads.load(std::memory_order_relaxed);
    allreduce_multimem_kernel<<<148 * grid_mul, threads, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
        reinterpret_cast<float4*>(multicast_buffer), rank, world, numel / 4);
    HTGS_LAUNCH_CHECK("allreduce_multimem_kernel");
    return 0;
}
