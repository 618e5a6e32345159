// distCUDA2: mean squared distance to the 3 nearest other points.
// Replaces SimpleKNN::knn (simple-knn/simple_knn.cu:186-222) and its kernels
// coord2Morton (:64-71), boxMinMax (:79-118), boxMeanDist (:148-184).
//
// Same exact-search idea (Morton order, AABBs of runs of sorted points, prune by
// box distance) but: no cudaMalloc / thrust vectors / blocking copies (the
// caller passes one workspace; the scene AABB stays on the device), the Morton
// sort is this library's onesweep radix sort, and the boxes are two-level
// (256-point leaves under 4096-point parents) so a query tests P/4096 parents
// instead of P/1024 boxes and scans 256-point leaves instead of 1024-point ones.
// The result depends only on the three smallest distances, so it is independent
// of traversal order and equals the reference bit for bit.
#include "common.cuh"
#include <cfloat>

namespace htgs {

constexpr int KNN_LEAF = 256;
constexpr int KNN_FAN = 16;  // leaves per parent

struct Box { float3 lo, hi; };

struct KnnWs {
    float* partial;      // [nblk][6]
    float* aabb;         // [6]
    uint64_t* codes[2];  // [P]
    uint32_t* order[2];  // [P]
    Box* leaves;         // [nleaf]
    Box* parents;        // [nparent]
    void* sort_scratch;
    size_t bytes;
};

static KnnWs knn_carve(void* base, int P) {
    KnnWs w;
    char* p = reinterpret_cast<char*>(base);
    auto take = [&](size_t bytes) { p = reinterpret_cast<char*>(align_up(reinterpret_cast<size_t>(p), 256)); char* r = p; p += bytes; return r; };
    const size_t n = (size_t)P;
    const size_t nblk = (n + 1023) / 1024;
    const size_t nleaf = (n + KNN_LEAF - 1) / KNN_LEAF, npar = (nleaf + KNN_FAN - 1) / KNN_FAN;
    w.partial = reinterpret_cast<float*>(take(nblk * 6 * sizeof(float)));
    w.aabb = reinterpret_cast<float*>(take(8 * sizeof(float)));
    for (int i = 0; i < 2; i++) w.codes[i] = reinterpret_cast<uint64_t*>(take(n * sizeof(uint64_t)));
    for (int i = 0; i < 2; i++) w.order[i] = reinterpret_cast<uint32_t*>(take(n * sizeof(uint32_t)));
    w.leaves = reinterpret_cast<Box*>(take(nleaf * sizeof(Box)));
    w.parents = reinterpret_cast<Box*>(take(npar * sizeof(Box)));
    w.sort_scratch = take(sort_scratch_bytes((int64_t)n));
    w.bytes = (size_t)(p - reinterpret_cast<char*>(base)) + 256;
    return w;
}

// Scene AABB with the reference's init value {0,0,0} (simple_knn.cu:192-200):
// the box always contains the origin.
__global__ void __launch_bounds__(1024) aabb_partial_kernel(int P, const float* __restrict__ pts, float* __restrict__ partial) {
    __shared__ float s[32][6];
    const int i = blockIdx.x * 1024 + threadIdx.x;
    float v[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (i < P) {
        const float x = pts[3 * (size_t)i], y = pts[3 * (size_t)i + 1], z = pts[3 * (size_t)i + 2];
        v[0] = fminf(0.f, x); v[1] = fminf(0.f, y); v[2] = fminf(0.f, z);
        v[3] = fmaxf(0.f, x); v[4] = fmaxf(0.f, y); v[5] = fmaxf(0.f, z);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
        for (int k = 0; k < 3; k++) {
            v[k] = fminf(v[k], __shfl_xor_sync(0xffffffffu, v[k], o));
            v[3 + k] = fmaxf(v[3 + k], __shfl_xor_sync(0xffffffffu, v[3 + k], o));
        }
    }
    if ((threadIdx.x & 31) == 0)
        for (int k = 0; k < 6; k++) s[threadIdx.x >> 5][k] = v[k];
    __syncthreads();
    if (threadIdx.x < 6) {
        float r = s[0][threadIdx.x];
        for (int w = 1; w < 32; w++) r = threadIdx.x < 3 ? fminf(r, s[w][threadIdx.x]) : fmaxf(r, s[w][threadIdx.x]);
        partial[blockIdx.x * 6 + threadIdx.x] = r;
    }
}

__global__ void aabb_final_kernel(int nblk, const float* __restrict__ partial, float* __restrict__ aabb) {
    const int k = threadIdx.x;
    if (k >= 6) return;
    float r = 0.f;
    for (int b = 0; b < nblk; b++) r = k < 3 ? fminf(r, partial[b * 6 + k]) : fmaxf(r, partial[b * 6 + k]);
    aabb[k] = r;
}

__device__ __forceinline__ uint32_t prep_morton(uint32_t x) {  // simple_knn.cu:46-53
    x = (x | (x << 16)) & 0x030000FF;
    x = (x | (x << 8)) & 0x0300F00F;
    x = (x | (x << 4)) & 0x030C30C3;
    x = (x | (x << 2)) & 0x09249249;
    return x;
}

__global__ void morton_kernel(int P, const float* __restrict__ pts, const float* __restrict__ aabb,
                              uint64_t* __restrict__ codes, uint32_t* __restrict__ order) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const float x = pts[3 * (size_t)i], y = pts[3 * (size_t)i + 1], z = pts[3 * (size_t)i + 2];
    const float sx = ((x - aabb[0]) / (aabb[3] - aabb[0])) * 1023.f;
    const float sy = ((y - aabb[1]) / (aabb[4] - aabb[1])) * 1023.f;
    const float sz = ((z - aabb[2]) / (aabb[5] - aabb[2])) * 1023.f;
    const uint32_t c = prep_morton((uint32_t)sx) | (prep_morton((uint32_t)sy) << 1) | (prep_morton((uint32_t)sz) << 2);
    codes[i] = (uint64_t)c;
    order[i] = (uint32_t)i;
}

__global__ void __launch_bounds__(KNN_LEAF) leaf_box_kernel(int P, const float* __restrict__ pts,
                                                            const uint32_t* __restrict__ order, Box* __restrict__ leaves) {
    __shared__ float s[KNN_LEAF / 32][6];
    const int i = blockIdx.x * KNN_LEAF + threadIdx.x;
    float v[6] = {FLT_MAX, FLT_MAX, FLT_MAX, -FLT_MAX, -FLT_MAX, -FLT_MAX};
    if (i < P) {
        const uint32_t g = order[i];
        v[0] = v[3] = pts[3 * (size_t)g]; v[1] = v[4] = pts[3 * (size_t)g + 1]; v[2] = v[5] = pts[3 * (size_t)g + 2];
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
        for (int k = 0; k < 3; k++) {
            v[k] = fminf(v[k], __shfl_xor_sync(0xffffffffu, v[k], o));
            v[3 + k] = fmaxf(v[3 + k], __shfl_xor_sync(0xffffffffu, v[3 + k], o));
        }
    }
    if ((threadIdx.x & 31) == 0)
        for (int k = 0; k < 6; k++) s[threadIdx.x >> 5][k] = v[k];
    __syncthreads();
    if (threadIdx.x == 0) {
        Box b;
        float r[6];
        for (int k = 0; k < 6; k++) {
            r[k] = s[0][k];
            for (int w = 1; w < KNN_LEAF / 32; w++) r[k] = k < 3 ? fminf(r[k], s[w][k]) : fmaxf(r[k], s[w][k]);
        }
        b.lo = make_float3(r[0], r[1], r[2]);
        b.hi = make_float3(r[3], r[4], r[5]);
        leaves[blockIdx.x] = b;
    }
}

__global__ void parent_box_kernel(int nleaf, int npar, const Box* __restrict__ leaves, Box* __restrict__ parents) {
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= npar) return;
    Box b = leaves[p * KNN_FAN];
    for (int k = 1; k < KNN_FAN && p * KNN_FAN + k < nleaf; k++) {
        const Box c = leaves[p * KNN_FAN + k];
        b.lo.x = fminf(b.lo.x, c.lo.x); b.lo.y = fminf(b.lo.y, c.lo.y); b.lo.z = fminf(b.lo.z, c.lo.z);
        b.hi.x = fmaxf(b.hi.x, c.hi.x); b.hi.y = fmaxf(b.hi.y, c.hi.y); b.hi.z = fmaxf(b.hi.z, c.hi.z);
    }
    parents[p] = b;
}

__device__ __forceinline__ float box_dist2(const Box& b, const float3& p) {  // simple_knn.cu:120-130
    float dx = 0.f, dy = 0.f, dz = 0.f;
    if (p.x < b.lo.x || p.x > b.hi.x) dx = fminf(fabsf(p.x - b.lo.x), fabsf(p.x - b.hi.x));
    if (p.y < b.lo.y || p.y > b.hi.y) dy = fminf(fabsf(p.y - b.lo.y), fabsf(p.y - b.hi.y));
    if (p.z < b.lo.z || p.z > b.hi.z) dz = fminf(fabsf(p.z - b.lo.z), fabsf(p.z - b.hi.z));
    return dx * dx + dy * dy + dz * dz;
}

__device__ __forceinline__ void update3(const float3& ref, const float3& pt, float* best) {  // simple_knn.cu:132-146
    const float dx = __fsub_rn(pt.x, ref.x), dy = __fsub_rn(pt.y, ref.y), dz = __fsub_rn(pt.z, ref.z);
    float d = __fmaf_rn(dz, dz, __fmaf_rn(dx, dx, __fmul_rn(dy, dy)));
#pragma unroll
    for (int j = 0; j < 3; j++)
        if (best[j] > d) { const float t = best[j]; best[j] = d; d = t; }
}

__device__ __forceinline__ float3 load_pt(const float* __restrict__ pts, uint32_t g) {
    return make_float3(__ldg(pts + 3 * (size_t)g), __ldg(pts + 3 * (size_t)g + 1), __ldg(pts + 3 * (size_t)g + 2));
}

__global__ void __launch_bounds__(256)
knn_mean_dist_kernel(int P, int nleaf, int npar, const float* __restrict__ pts, const uint32_t* __restrict__ order,
                     const Box* __restrict__ leaves, const Box* __restrict__ parents, float* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const uint32_t self = order[i];
    const float3 ref = load_pt(pts, self);
    float best[3] = {FLT_MAX, FLT_MAX, FLT_MAX};
    // seed the pruning bound with the Morton neighbours (simple_knn.cu:157-162)
    for (int j = max(0, i - 3); j <= min(P - 1, i + 3); j++) {
        if (j == i) continue;
        update3(ref, load_pt(pts, order[j]), best);
    }
    const float reject = best[2];
    best[0] = best[1] = best[2] = FLT_MAX;
    for (int pb = 0; pb < npar; pb++) {
        const float dp = box_dist2(parents[pb], ref);
        if (dp > reject || dp > best[2]) continue;
        const int l1 = min(nleaf, (pb + 1) * KNN_FAN);
        for (int lb = pb * KNN_FAN; lb < l1; lb++) {
            const float dl = box_dist2(leaves[lb], ref);
            if (dl > reject || dl > best[2]) continue;
            const int e = min(P, (lb + 1) * KNN_LEAF);
            for (int j = lb * KNN_LEAF; j < e; j++) {
                if (j == i) continue;
                update3(ref, load_pt(pts, order[j]), best);
            }
        }
    }
    out[self] = __fdiv_rn(__fadd_rn(__fadd_rn(best[0], best[1]), best[2]), 3.0f);
}

size_t knn_workspace_bytes(int P) { return knn_carve(nullptr, P < 0 ? 0 : P).bytes; }

int knn_dist2(int P, const float* points, float* out, void* workspace, size_t workspace_bytes, cudaStream_t st) {
    if (P <= 0) return 0;
    if (!points || !out || !workspace) { set_error("htgs_dist2: null pointer"); return 1; }
    if (reinterpret_cast<size_t>(workspace) & 255) { set_error("htgs_dist2: workspace must be 256-byte aligned"); return 1; }
    KnnWs w = knn_carve(workspace, P);
    if (workspace_bytes < w.bytes) { set_error("htgs_dist2: workspace too small (%zu < %zu)", workspace_bytes, w.bytes); return 1; }
    const int nblk = (P + 1023) / 1024;
    const int nleaf = (P + KNN_LEAF - 1) / KNN_LEAF, npar = (nleaf + KNN_FAN - 1) / KNN_FAN;
    aabb_partial_kernel<<<nblk, 1024, 0, st>>>(P, points, w.partial);
    HTGS_LAUNCH_CHECK("aabb_partial_kernel");
    aabb_final_kernel<<<1, 32, 0, st>>>(nblk, w.partial, w.aabb);
    HTGS_LAUNCH_CHECK("aabb_final_kernel");
    morton_kernel<<<(P + 255) / 256, 256, 0, st>>>(P, points, w.aabb, w.codes[0], w.order[0]);
    HTGS_LAUNCH_CHECK("morton_kernel");
    const int end_bit = 30;
    if (sort_pairs(P, end_bit, w.codes[0], w.order[0], w.codes[1], w.order[1], w.sort_scratch, st)) return 1;
    const uint32_t* order = (sort_num_passes(end_bit) & 1) ? w.order[1] : w.order[0];
    leaf_box_kernel<<<nleaf, KNN_LEAF, 0, st>>>(P, points, order, w.leaves);
    HTGS_LAUNCH_CHECK("leaf_box_kernel");
    parent_box_kernel<<<(npar + 127) / 128, 128, 0, st>>>(nleaf, npar, w.leaves, w.parents);
    HTGS_LAUNCH_CHECK("parent_box_kernel");
    knn_mean_dist_kernel<<<(P + 255) / 256, 256, 0, st>>>(P, nleaf, npar, points, order, w.leaves, w.parents, out);
    HTGS_LAUNCH_CHECK("knn_mean_dist_kernel");
    return 0;
}

}  // namespace htgs

extern "C" {
size_t htgs_dist2_workspace_bytes(int P) { return htgs::knn_workspace_bytes(P); }
int htgs_dist2(int P, const float* points, float* out, void* workspace, size_t workspace_bytes, void* stream) {
    return htgs::knn_dist2(P, points, out, workspace, workspace_bytes, reinterpret_cast<cudaStream_t>(stream));
}
}
