// red_probe.cu -- micro-benchmark of global float reductions on the box ("peak = a
// measured micro-benchmark: red.global.add.f32 to uniformly random addresses in a 10*P-float array").
// Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a tools/micro/red_probe.cu -o tools/micro/red_probe
// Prints one JSON document.  Patterns:
//   scalar_random      red.global.add.f32, every lane its own random address in a 10*P-float array (P = 1M)
//   scalar_warp_same   red.global.add.f32, the 32 lanes of a warp hit the SAME random address (the reference backward:
//                      all pixels of a warp add to the same Gaussian's slot, backward.cu:555-598)
//   v4_random          red.global.add.v4.f32, every lane its own random 16-byte row of a [P][12] array (this library's
//                      flush: three such rows per (tile, Gaussian) instance)
//   v4_rows3           three consecutive v4 rows of one random Gaussian per lane (exactly the flush pattern)
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>

__device__ __forceinline__ uint32_t mix(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
    return x;
}

template <int MODE>
__global__ void __launch_bounds__(256) red_kernel(float* buf, uint32_t P, int iters) {
    const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t s = mix(tid * 2654435761u + 12345u);
    for (int i = 0; i < iters; i++) {
        s = mix(s + 0x9e3779b9u);
        if (MODE == 0) {
            float* a = buf + (s % (10u * P));
            asm volatile("red.global.add.f32 [%0], %1;" ::"l"(a), "f"(1.0f) : "memory");
        } else if (MODE == 1) {
            const uint32_t sw = __shfl_sync(0xffffffffu, s, 0);
            float* a = buf + (sw % (10u * P));
            asm volatile("red.global.add.f32 [%0], %1;" ::"l"(a), "f"(1.0f) : "memory");
        } else if (MODE == 2) {
            float* a = buf + (size_t)(s % (3u * P)) * 4;
            asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(a), "f"(1.0f), "f"(1.0f), "f"(1.0f), "f"(1.0f) : "memory");
        } else {
            float* a = buf + (size_t)(s % P) * 12;
#pragma unroll
            for (int r = 0; r < 3; r++)
                asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(a + 4 * r), "f"(1.0f), "f"(1.0f), "f"(1.0f), "f"(1.0f) : "memory");
        }
    }
}

template <int MODE>
static double run(float* buf, uint32_t P, int grid, int iters) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    red_kernel<MODE><<<grid, 256>>>(buf, P, iters);  // warm-up
    cudaDeviceSynchronize();
    float best = 1e30f;
    for (int rep = 0; rep < 5; rep++) {
        cudaEventRecord(e0);
        red_kernel<MODE><<<grid, 256>>>(buf, P, iters);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    return best;
}

int main() {
    const uint32_t P = 1000000;
    float* buf = nullptr;
    if (cudaMalloc(&buf, (size_t)P * 12 * sizeof(float)) != cudaSuccess) { printf("{\"error\": \"cudaMalloc\"}\n"); return 1; }
    cudaMemset(buf, 0, (size_t)P * 12 * sizeof(float));
    cudaDeviceProp prop;
    cudaGetDeviceProperties(&prop, 0);
    const int grid = prop.multiProcessorCount * 8, iters = 256;
    const double lanes = (double)grid * 256 * iters;
    const double t0 = run<0>(buf, P, grid, iters), t1 = run<1>(buf, P, grid, iters), t2 = run<2>(buf, P, grid, iters),
                 t3 = run<3>(buf, P, grid, iters / 4);
    printf("{\"device\": \"%s\", \"sms\": %d, \"array\": \"P = 1e6: 10*P floats (scalar), [P][12] floats (v4)\",\n", prop.name, prop.multiProcessorCount);
    printf(" \"scalar_random\": {\"ms\": %.4f, \"Gred_per_s\": %.2f, \"note\": \"red.global.add.f32, one random address per lane\"},\n", t0, lanes / t0 / 1e6);
    printf(" \"scalar_warp_same\": {\"ms\": %.4f, \"Gred_per_s\": %.2f, \"Gwarp_red_per_s\": %.3f, \"note\": \"32 lanes of a warp on one address (reference backward pattern)\"},\n", t1, lanes / t1 / 1e6, lanes / 32 / t1 / 1e6);
    printf(" \"v4_random\": {\"ms\": %.4f, \"Gred_v4_per_s\": %.2f, \"GB_per_s_payload\": %.1f, \"note\": \"red.global.add.v4.f32, one random 16-B row per lane\"},\n", t2, lanes / t2 / 1e6, lanes * 16 / t2 / 1e6);
    const double lanes3 = (double)grid * 256 * (iters / 4) * 3;
    printf(" \"v4_rows3\": {\"ms\": %.4f, \"Gred_v4_per_s\": %.2f, \"GB_per_s_payload\": %.1f, \"note\": \"three consecutive v4 rows of one random Gaussian per lane (this library's per-instance flush)\"}}\n", t3, lanes3 / t3 / 1e6, lanes3 * 16 / t3 / 1e6);
    cudaFree(buf);
    return 0;
}
