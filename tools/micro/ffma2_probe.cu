// Micro-probe (development aid): does fma.rn.f32x2 (FFMA2) free issue slots on sm_100a?
// Times (a) scalar FFMA chains, (b) the same FLOPs as FFMA2, (c)/(d) each interleaved 1:1 with integer LOP3/IADD work.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ uint64_t pk(float a, float b) { uint64_t r; asm("mov.b64 %0, {%1,%2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ void up(uint64_t r, float& a, float& b) { asm("mov.b64 {%0,%1}, %2;" : "=f"(a), "=f"(b) : "l"(r)); }
__device__ __forceinline__ uint64_t fma2(uint64_t a, uint64_t b, uint64_t c) { uint64_t d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }

template <int MODE>
__global__ void probe(float* out, int iters, float x, float y, uint32_t z) {
    float a[8];
    uint64_t p[4];
    uint32_t u[4] = {threadIdx.x, threadIdx.x * 3u, z, z ^ 7u};
    for (int i = 0; i < 8; i++) a[i] = threadIdx.x * 0.001f + i;
    for (int i = 0; i < 4; i++) p[i] = pk(a[2 * i], a[2 * i + 1]);
    const uint64_t xx = pk(x, x), yy = pk(y, y);
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int r = 0; r < 8; r++) {
            if (MODE == 0 || MODE == 2) {
#pragma unroll
                for (int i = 0; i < 8; i++) a[i] = fmaf(a[i], x, y);
            } else {
#pragma unroll
                for (int i = 0; i < 4; i++) p[i] = fma2(p[i], xx, yy);
            }
            if (MODE >= 2) {
#pragma unroll
                for (int i = 0; i < 4; i++) { u[i] = (u[i] ^ (u[(i + 1) & 3] + z)) + 0x9e3779b9u; u[i] = (u[i] << 3) | (u[i] >> 29); }
            }
        }
    }
    float s = 0.f;
    if (MODE == 1 || MODE == 3) for (int i = 0; i < 4; i++) { float q, w; up(p[i], q, w); s += q + w; }
    else for (int i = 0; i < 8; i++) s += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s + (float)(u[0] + u[1] + u[2] + u[3]);
}

template <int MODE> float run(float* d, int iters) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    probe<MODE><<<148 * 8, 256>>>(d, iters, 1.0001f, 0.5f, 5u);
    cudaEventRecord(e0);
    probe<MODE><<<148 * 8, 256>>>(d, iters, 1.0001f, 0.5f, 5u);
    cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1); return ms;
}
int main() {
    float* d; cudaMalloc(&d, 148 * 8 * 256 * 4);
    const int iters = 4096;
    const double fma = 148.0 * 8 * 256 * (double)iters * 64;
    float t0 = run<0>(d, iters), t1 = run<1>(d, iters), t2 = run<2>(d, iters), t3 = run<3>(d, iters);
    printf("FFMA only        %.3f ms  %.1f TFLOP/s\n", t0, 2 * fma / t0 / 1e9);
    printf("FFMA2 only       %.3f ms  %.1f TFLOP/s\n", t1, 2 * fma / t1 / 1e9);
    printf("FFMA  + int ALU  %.3f ms\n", t2);
    printf("FFMA2 + int ALU  %.3f ms\n", t3);
    return 0;
}
