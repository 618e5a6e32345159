// "Next" row 4 of SURVEY.md section 8(f), first half: the row surgery behind prune_points and
// densify_and_clone (scene/gaussian_model_ht.py:548-616, :656-672).
//
// The reference rebuilds every parameter, both Adam moments of every parameter and the three statistics
// tensors with one boolean-mask indexing op each -- 21 x (nonzero + host sync + gather) for a prune, and
// 18 x torch.cat plus 6 masked gathers for a clone.  Here the mask is scanned once (offsets[i] = number of
// selected rows before row i, total copied to pinned host memory), and ONE launch moves the rows of all
// tensors: blockIdx.y picks the tensor, threads walk its 32-bit words so reads are fully coalesced and
// writes are coalesced within every run of selected rows.  Row order is preserved, so the result is
// bit-identical to torch's `t[mask]` / `torch.cat((t, t[mask]))`.
#include "../../include/htgs.h"
#include "common.cuh"

namespace htgs {

constexpr int ROW_BLOCK = 256;

struct RowLaunch {
    htgs_row_tensor t[HTGS_ROWS_MAX_TENSORS];
    int n;
};

__global__ void __launch_bounds__(ROW_BLOCK) mask_count_kernel(int P, const uint8_t* __restrict__ mask, uint32_t* __restrict__ block_sums) {
    const int i = blockIdx.x * ROW_BLOCK + threadIdx.x;
    const int c = __syncthreads_count(i < P && mask[i] != 0);
    if (threadIdx.x == 0) block_sums[blockIdx.x] = (uint32_t)c;
}

// exclusive scan of the block totals in place; block_sums[nb] = total, also to *total_out
__global__ void __launch_bounds__(1024) mask_scan_blocks_kernel(int nb, uint32_t* __restrict__ block_sums, uint32_t* __restrict__ total_out) {
    __shared__ uint32_t s_warp[32];
    __shared__ uint32_t s_carry;
    if (threadIdx.x == 0) s_carry = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int base = 0; base < nb; base += 1024) {
        const int i = base + threadIdx.x;
        const uint32_t v = i < nb ? block_sums[i] : 0;
        uint32_t inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t n = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += n;
        }
        if (lane == 31) s_warp[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            uint32_t w = s_warp[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t n = __shfl_up_sync(0xffffffffu, w, o);
                if (lane >= o) w += n;
            }
            s_warp[lane] = w;
        }
        __syncthreads();
        const uint32_t warp_off = warp == 0 ? 0 : s_warp[warp - 1];
        const uint32_t carry = s_carry;
        if (i < nb) block_sums[i] = carry + warp_off + inc - v;
        __syncthreads();
        if (threadIdx.x == 1023) s_carry = carry + warp_off + inc;
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        block_sums[nb] = s_carry;
        *total_out = s_carry;
    }
}

__global__ void __launch_bounds__(ROW_BLOCK) mask_offsets_kernel(int P, const uint8_t* __restrict__ mask,
                                                                 const uint32_t* __restrict__ block_sums,
                                                                 uint32_t* __restrict__ offsets) {
    __shared__ uint32_t s_warp[ROW_BLOCK / 32];
    const int i = blockIdx.x * ROW_BLOCK + threadIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const bool sel = i < P && mask[i] != 0;
    const uint32_t b = __ballot_sync(0xffffffffu, sel);
    if (lane == 0) s_warp[warp] = __popc(b);
    __syncthreads();
    uint32_t before = block_sums[blockIdx.x];
#pragma unroll
    for (int w = 0; w < ROW_BLOCK / 32; w++)
        if (w < warp) before += s_warp[w];
    if (i < P) offsets[i] = before + __popc(b & ((1u << lane) - 1u));
}

// APPEND = false: dst[offsets[r]] = src[r] for selected rows r                    (torch: src[mask])
// APPEND = true : dst[r] = src[r] for all r; dst[P + offsets[r]] = src[r] or 0     (torch: cat(src, src[mask] | zeros))
template <bool APPEND>
__global__ void __launch_bounds__(ROW_BLOCK) move_rows_kernel(int P, const uint8_t* __restrict__ mask,
                                                              const uint32_t* __restrict__ offsets,
                                                              const __grid_constant__ RowLaunch L) {
    const htgs_row_tensor T = L.t[blockIdx.y];
    const uint32_t* __restrict__ src = reinterpret_cast<const uint32_t*>(T.src);
    uint32_t* __restrict__ dst = reinterpret_cast<uint32_t*>(T.dst);
    const long long w = T.row_words, total = (long long)P * w;
    if (total < (1ll << 31)) {  // 32-bit index arithmetic (the 64-bit division below costs ~100 instructions per word)
        const uint32_t w32 = (uint32_t)w, n32 = (uint32_t)total, step = gridDim.x * ROW_BLOCK;
        const uint32_t app = APPEND ? (uint32_t)P : 0u;
        for (uint32_t e = blockIdx.x * ROW_BLOCK + threadIdx.x; e < n32; e += step) {
            const uint32_t r = e / w32;
            const uint32_t v = src[e];
            if (APPEND) dst[e] = v;
            if (mask[r]) dst[(size_t)(app + offsets[r]) * w32 + (e - r * w32)] = (APPEND && T.zero_new) ? 0u : v;
            if (e + step < e) break;  // wrap guard
        }
        return;
    }
    for (long long e = (long long)blockIdx.x * ROW_BLOCK + threadIdx.x; e < total; e += (long long)gridDim.x * ROW_BLOCK) {
        const long long r = e / w;
        const uint32_t v = src[e];
        if (APPEND) dst[e] = v;
        if (mask[r]) {
            const long long c = e - r * w;
            const long long d = ((APPEND ? (long long)P : 0ll) + offsets[r]) * w + c;
            dst[d] = (APPEND && T.zero_new) ? 0u : v;
        }
    }
}

// densify_and_split's tensor work (scene/gaussian_model_ht.py:630-654): the selected rows are replaced by N new
// rows each, appended after the survivors:  dst = cat(src[~mask], repeat(src[mask], N))  -- what
// densification_postfix followed by prune_points(cat(mask, zeros)) leaves.  offsets[] is the scan of `mask`
// (count = selected rows), so a surviving row r lands at r - offsets[r].  zero_new: the N*count new rows are zeros
// (Adam moments); skip_new: they are left for the caller to fill (positions / scales, which are computed).
__global__ void __launch_bounds__(ROW_BLOCK) split_rows_kernel(int P, int N, uint32_t count, const uint8_t* __restrict__ mask,
                                                               const uint32_t* __restrict__ offsets,
                                                               const __grid_constant__ RowLaunch L) {
    const htgs_row_tensor T = L.t[blockIdx.y];
    const uint32_t* __restrict__ src = reinterpret_cast<const uint32_t*>(T.src);
    uint32_t* __restrict__ dst = reinterpret_cast<uint32_t*>(T.dst);
    const long long w = T.row_words, total = (long long)P * w;
    const long long K = (long long)P - count;  // survivors
    for (long long e = (long long)blockIdx.x * ROW_BLOCK + threadIdx.x; e < total; e += (long long)gridDim.x * ROW_BLOCK) {
        const long long r = total < (1ll << 31) ? (long long)((uint32_t)e / (uint32_t)w) : e / w;
        const long long c = e - r * w;
        const uint32_t v = src[e];
        const uint32_t o = offsets[r];
        if (!mask[r]) {
            dst[(r - o) * w + c] = v;
        } else if (T.zero_new != 2) {
            for (int n = 0; n < N; n++) dst[(K + (long long)n * count + o) * w + c] = T.zero_new ? 0u : v;
        }
    }
}

static int check_tensors(const char* who, int n, const htgs_row_tensor* t) {
    if (n < 0 || n > HTGS_ROWS_MAX_TENSORS || (n > 0 && !t)) { set_error("%s: need 0..%d tensors", who, HTGS_ROWS_MAX_TENSORS); return 1; }
    for (int k = 0; k < n; k++)
        if (!t[k].src || !t[k].dst || t[k].row_words <= 0) { set_error("%s: tensor %d: null pointer or row_words <= 0", who, k); return 1; }
    return 0;
}

template <bool APPEND>
static int move_rows(const char* who, int P, const uint8_t* mask, const void* scratch, int n, const htgs_row_tensor* t, cudaStream_t st) {
    if (P <= 0 || n == 0) return 0;
    if (!mask || !scratch) { set_error("%s: null pointer", who); return 1; }
    if (check_tensors(who, n, t)) return 1;
    RowLaunch L;
    L.n = n;
    long long max_words = 0;
    for (int k = 0; k < n; k++) {
        L.t[k] = t[k];
        const long long wds = (long long)P * t[k].row_words;
        if (wds > max_words) max_words = wds;
    }
    const int nb = (P + ROW_BLOCK - 1) / ROW_BLOCK;
    const uint32_t* offsets = reinterpret_cast<const uint32_t*>(reinterpret_cast<const char*>(scratch) + align_up((size_t)(nb + 1) * 4, 256));
    const long long want = (max_words + ROW_BLOCK - 1) / ROW_BLOCK;
    const dim3 grid((unsigned)(want < 148 * 16 ? (want < 1 ? 1 : want) : 148 * 16), (unsigned)n);
    move_rows_kernel<APPEND><<<grid, ROW_BLOCK, 0, st>>>(P, mask, offsets, L);
    HTGS_LAUNCH_CHECK("move_rows_kernel");
    return 0;
}

}  // namespace htgs

using namespace htgs;

extern "C" size_t htgs_row_scan_scratch_bytes(int P) {
    const size_t p = P > 0 ? (size_t)P : 0, nb = (p + ROW_BLOCK - 1) / ROW_BLOCK;
    return align_up((nb + 1) * 4, 256) + align_up(p * 4, 256) + 256;
}

extern "C" int htgs_row_scan(int P, const uint8_t* mask, void* scratch, uint32_t* count_pinned_host, void* stream) {
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    if (!count_pinned_host) { set_error("htgs_row_scan: null count pointer"); return 1; }
    if (P <= 0) { *count_pinned_host = 0; return 0; }
    if (!mask || !scratch) { set_error("htgs_row_scan: null pointer"); return 1; }
    if (reinterpret_cast<size_t>(scratch) & 255) { set_error("htgs_row_scan: scratch must be 256-byte aligned"); return 1; }
    const int nb = (P + ROW_BLOCK - 1) / ROW_BLOCK;
    uint32_t* block_sums = reinterpret_cast<uint32_t*>(scratch);
    uint32_t* offsets = reinterpret_cast<uint32_t*>(reinterpret_cast<char*>(scratch) + align_up((size_t)(nb + 1) * 4, 256));
    uint32_t* total = reinterpret_cast<uint32_t*>(reinterpret_cast<char*>(offsets) + align_up((size_t)P * 4, 256));
    mask_count_kernel<<<nb, ROW_BLOCK, 0, st>>>(P, mask, block_sums);
    HTGS_LAUNCH_CHECK("mask_count_kernel");
    mask_scan_blocks_kernel<<<1, 1024, 0, st>>>(nb, block_sums, total);
    HTGS_LAUNCH_CHECK("mask_scan_blocks_kernel");
    mask_offsets_kernel<<<nb, ROW_BLOCK, 0, st>>>(P, mask, block_sums, offsets);
    HTGS_LAUNCH_CHECK("mask_offsets_kernel");
    HTGS_CUDA_CHECK(cudaMemcpyAsync(count_pinned_host, total, sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
    return 0;
}

extern "C" int htgs_rows_compact(int P, const uint8_t* mask, const void* scratch, int n_tensors, const htgs_row_tensor* tensors,
                                 void* stream) {
    return move_rows<false>("htgs_rows_compact", P, mask, scratch, n_tensors, tensors, reinterpret_cast<cudaStream_t>(stream));
}

extern "C" int htgs_rows_append(int P, const uint8_t* mask, const void* scratch, int n_tensors, const htgs_row_tensor* tensors,
                                void* stream) {
    return move_rows<true>("htgs_rows_append", P, mask, scratch, n_tensors, tensors, reinterpret_cast<cudaStream_t>(stream));
}

extern "C" int htgs_rows_split(int P, int N, uint32_t count, const uint8_t* mask, const void* scratch, int n_tensors,
                               const htgs_row_tensor* tensors, void* stream) {
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    const char* who = "htgs_rows_split";
    if (P <= 0 || n_tensors == 0) return 0;
    if (N < 1) { set_error("%s: N must be >= 1", who); return 1; }
    if (count > (uint32_t)P) { set_error("%s: count exceeds P", who); return 1; }
    if (!mask || !scratch) { set_error("%s: null pointer", who); return 1; }
    if (check_tensors(who, n_tensors, tensors)) return 1;
    RowLaunch L;
    L.n = n_tensors;
    long long max_words = 0;
    for (int k = 0; k < n_tensors; k++) {
        L.t[k] = tensors[k];
        const long long wds = (long long)P * tensors[k].row_words;
        if (wds > max_words) max_words = wds;
    }
    const int nb = (P + ROW_BLOCK - 1) / ROW_BLOCK;
    const uint32_t* offsets = reinterpret_cast<const uint32_t*>(reinterpret_cast<const char*>(scratch) + align_up((size_t)(nb + 1) * 4, 256));
    const long long want = (max_words + ROW_BLOCK - 1) / ROW_BLOCK;
    const dim3 grid((unsigned)(want < 148 * 16 ? (want < 1 ? 1 : want) : 148 * 16), (unsigned)n_tensors);
    split_rows_kernel<<<grid, ROW_BLOCK, 0, st>>>(P, N, count, mask, offsets, L);
    HTGS_LAUNCH_CHECK("split_rows_kernel");
    return 0;
}
