// Micro-probe (development aid): latency of match.any.sync as a function of how many distinct values the warp holds.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__global__ void probe(uint32_t* out, int mode, int iters) {
    const int lane = threadIdx.x & 31;
    uint32_t v;
    switch (mode) {
        case 0: v = 7; break;                                   // 1 distinct
        case 1: v = lane & 3; break;                            // 4 distinct
        case 2: v = lane & 15; break;                           // 16 distinct
        case 3: v = (lane * 2654435761u) >> 24; break;          // ~30 distinct, scattered values
        case 4: v = 100 + lane; break;                          // 32 distinct, consecutive values
        default: v = (lane * 37u + 11u) & 255u; break;          // 32 distinct, scattered
    }
    uint32_t acc = 0;
    long long t0 = clock64();
    for (int i = 0; i < iters; i++) {
        const uint32_t peers = __match_any_sync(0xffffffffu, v);
        acc += __popc(peers);
        v = (v + (acc & 0u));  // keep a dependency, value unchanged
    }
    long long t1 = clock64();
    if (threadIdx.x == 0) { out[2 * mode] = (uint32_t)((t1 - t0) / iters); out[2 * mode + 1] = acc; }
}
int main() {
    uint32_t* d; cudaMalloc(&d, 64);
    uint32_t h[16];
    for (int m = 0; m < 6; m++) probe<<<1, 32>>>(d, m, 4096);
    cudaMemcpy(h, d, 48, cudaMemcpyDeviceToHost);
    const char* names[6] = {"1 distinct", "4 distinct", "16 distinct", "~30 distinct scattered", "32 distinct consecutive", "32 distinct scattered"};
    for (int m = 0; m < 6; m++) printf("%-26s %u cycles per dependent match.any\n", names[m], h[2 * m]);
    return 0;
}
