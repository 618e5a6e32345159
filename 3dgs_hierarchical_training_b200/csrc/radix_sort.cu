// L2: hand-written onesweep LSD radix sort of (uint64 key, uint32 value) pairs.
// Replaces cub::DeviceRadixSort::SortPairs(keys, values, R, 0, 32+bit)
// (cuda_rasterizer/rasterizer_impl.cu:304-309).  Stable, 8-bit digits,
// ceil(end_bit/8) scatter passes after ONE up-front histogram pass:
//   sort_hist_kernel      all per-pass digit histograms in a single read of the keys
//   sort_scan_hist_kernel exclusive scan of each 256-bin histogram
//   onesweep_kernel       per pass: rank inside a 4096-key tile with warp
//                         match-any multisplit, chained decoupled look-back of the
//                         per-digit tile prefixes, scatter through shared memory so
//                         each digit's run leaves the SM as one contiguous store.
// Algorithmic HBM traffic: R*(8 + passes*24) bytes (HBM-bound, integer work).
#include "common.cuh"
#include <algorithm>
#include <cstdio>

namespace htgs {

constexpr int SORT_THREADS = 256;
constexpr int SORT_WARPS = SORT_THREADS / 32;
#ifndef HTGS_SORT_IPT
#define HTGS_SORT_IPT 16
#endif
#ifndef HTGS_SORT_MINB
#define HTGS_SORT_MINB 3
#endif
constexpr int SORT_IPT = HTGS_SORT_IPT;
constexpr int SORT_TILE = SORT_THREADS * SORT_IPT;  // 4096 keys per CTA
constexpr int SORT_RADIX = 256;
constexpr int SORT_MAX_PASSES = 8;
constexpr int SORT_LOOK = 8;
#ifndef HTGS_HIST_GRID
#define HTGS_HIST_GRID 4
#endif  // predecessor status words fetched per look-back round trip

constexpr uint32_t FLAG_AGG = 1u << 30;
constexpr uint32_t FLAG_PREFIX = 2u << 30;
constexpr uint32_t FLAG_MASK = 3u << 30;
constexpr uint32_t VALUE_MASK = ~FLAG_MASK;

struct SortScratch {
    uint32_t* hist;     // [SORT_MAX_PASSES][256]
    uint32_t* tickets;  // [SORT_MAX_PASSES]
    uint32_t* status;   // [passes][tiles][256]
};

static inline int64_t sort_tiles(int64_t n) { return (n + SORT_TILE - 1) / SORT_TILE; }

static inline size_t header_bytes() { return align_up((SORT_MAX_PASSES * SORT_RADIX + SORT_MAX_PASSES + 8) * sizeof(uint32_t), 256); }

size_t sort_scratch_bytes(int64_t n) {
    return header_bytes() + align_up((size_t)SORT_MAX_PASSES * (size_t)sort_tiles(n) * SORT_RADIX * sizeof(uint32_t), 256) + 256;
}

int sort_num_passes(int end_bit) { return (end_bit + 7) / 8; }

static SortScratch carve(void* scratch, int64_t n) {
    SortScratch s;
    char* p = reinterpret_cast<char*>(scratch);
    s.hist = reinterpret_cast<uint32_t*>(p);
    s.tickets = s.hist + SORT_MAX_PASSES * SORT_RADIX;
    s.status = reinterpret_cast<uint32_t*>(p + header_bytes());
    (void)n;
    return s;
}

__global__ void __launch_bounds__(SORT_THREADS)
sort_hist_kernel(const uint64_t* __restrict__ keys, int64_t n, const uint32_t* __restrict__ n_dev, int end_bit, int passes,
                 uint32_t* __restrict__ hist) {
    __shared__ uint32_t s_hist[SORT_MAX_PASSES * SORT_RADIX];
    if (n_dev) n = min(n, (int64_t)*n_dev);  // device-side count (n is then the capacity)
    for (int i = threadIdx.x; i < passes * SORT_RADIX; i += SORT_THREADS) s_hist[i] = 0;
    __syncthreads();
    const int64_t stride = (int64_t)gridDim.x * SORT_THREADS;
    constexpr int HU = 4;  // keys in flight per thread
    for (int64_t i0 = (int64_t)blockIdx.x * SORT_THREADS + threadIdx.x; i0 < n; i0 += stride * HU) {
        uint64_t k[HU];
#pragma unroll
        for (int u = 0; u < HU; u++) k[u] = (i0 + u * stride < n) ? keys[i0 + u * stride] : 0ull;
#pragma unroll
        for (int u = 0; u < HU; u++) {
            if (i0 + u * stride >= n) break;
            // (warp-aggregating equal low words with match.any before the atomics was measured: 0.229 -> 0.240 ms for the sort)
            for (int p = 0; p < passes; p++) {
                const int nb = min(8, end_bit - 8 * p);
                const uint32_t d = (uint32_t)(k[u] >> (8 * p)) & ((1u << nb) - 1u);
                atomicAdd(&s_hist[p * SORT_RADIX + d], 1u);
            }
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < passes * SORT_RADIX; i += SORT_THREADS) {
        const uint32_t c = s_hist[i];
        if (c) atomicAdd(&hist[i], c);
    }
}

// 256-wide exclusive scan across the CTA (one value per thread).
__device__ __forceinline__ uint32_t block_excl_scan(uint32_t v, uint32_t* s_tmp /*[SORT_WARPS]*/) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) s_tmp[warp] = inc;
    __syncthreads();
    uint32_t off = 0;
#pragma unroll
    for (int w = 0; w < SORT_WARPS; w++)
        if (w < warp) off += s_tmp[w];
    __syncthreads();
    return off + inc - v;
}

__global__ void __launch_bounds__(SORT_RADIX) sort_scan_hist_kernel(uint32_t* __restrict__ hist) {
    __shared__ uint32_t s_tmp[SORT_WARPS];
    uint32_t* h = hist + blockIdx.x * SORT_RADIX;
    const uint32_t v = h[threadIdx.x];
    const uint32_t e = block_excl_scan(v, s_tmp);
    h[threadIdx.x] = e;
}

__global__ void __launch_bounds__(SORT_THREADS, HTGS_SORT_MINB)
onesweep_kernel(const uint64_t* __restrict__ kin, const uint32_t* __restrict__ vin, uint64_t* __restrict__ kout,
                uint32_t* __restrict__ vout, int64_t n, const uint32_t* __restrict__ n_dev, int shift, int nbits,
                const uint32_t* __restrict__ digit_start, uint32_t* status, uint32_t* ticket) {
    __shared__ uint64_t s_keys[SORT_TILE];
    __shared__ uint8_t s_digit[SORT_TILE];
    __shared__ uint32_t s_whist[SORT_WARPS][SORT_RADIX];
    __shared__ uint32_t s_local_start[SORT_RADIX];
    __shared__ uint32_t s_goff[SORT_RADIX];
    __shared__ uint32_t s_tmp[SORT_WARPS];
    __shared__ uint32_t s_tile;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_tile = atomicAdd(ticket, 1u);
#pragma unroll
    for (int w = 0; w < SORT_WARPS; w++) s_whist[w][tid] = 0;
    // per-warp digit mask words of the tile-id passes: they borrow the front of the key staging buffer, which is not
    // written before the ranking is over (8 warps x 256 words = 8 KB of its 32 KB)
    for (int i = tid; i < SORT_WARPS * SORT_RADIX; i += SORT_THREADS) reinterpret_cast<uint32_t*>(s_keys)[i] = 0;
    __syncthreads();
    const uint32_t wmask_base = (uint32_t)__cvta_generic_to_shared(s_keys) + (uint32_t)warp * (SORT_RADIX * 4u);
    const uint32_t tile = s_tile;
    const int64_t base = (int64_t)tile * SORT_TILE;
    if (n_dev) n = min(n, (int64_t)*n_dev);  // grid was sized for the capacity; surplus tiles leave
    if (base >= n) return;
    const int valid = (int)min((int64_t)SORT_TILE, n - base);
    const uint32_t mask = (1u << nbits) - 1u;
    const uint32_t lanemask_lt = (1u << lane) - 1u;

    uint64_t k[SORT_IPT];
    const int64_t my0 = base + warp * (32 * SORT_IPT) + lane;
#pragma unroll
    for (int i = 0; i < SORT_IPT; i++) {
        const int64_t idx = my0 + i * 32;
        k[i] = idx < n ? kin[idx] : ~0ull;
    }
    // Rank inside the warp's 512-key slice (stable: item order = i-major, lane-minor).  ncu (r02) showed the warps of this
    // kernel stalled on the latency of match.any and of the dependent shared-memory read-modify-write behind it (issue
    // slots 25-40 % busy), so the three steps run as three waves with 8 independent items in flight each:
    //   A  peers = match.any(digit)                                   8 independent MATCH
    //   B  the lowest peer adds the group size to the warp's digit counter with ONE shared-memory atomic and keeps the old
    //      value; atomics of one warp to one address retire in issue order, so item i's prefix counts items < i exactly
    //   C  the old value is broadcast from the leader; rank = old + number of peers in lower lanes
    //
    // Step A has two forms.  match.any costs about one cycle of a shared unit per DISTINCT value in the warp
    // (tools/micro/match_probe.cu: 5 -> 26 cycles for 1 -> 32 distinct).  In the depth-byte passes neighbouring keys are
    // the tile instances of one Gaussian (same depth, so the same digit: ~8 distinct values per warp) and it is the cheap
    // form; in the tile-id passes (shift >= 32) the 32 digits of a warp are nearly all different, which made those two
    // passes 46-56 us against 30-36 us.  There the peers are found through shared memory instead: every lane ORs its lane
    // bit into a per-digit mask word, reads the word back and clears it -- three shared-memory instructions that are
    // conflict-free exactly when the digits differ, issued back to back (a warp's shared-memory instructions execute in
    // program order, so the read sees every lane's bit and the clear comes after every lane's read).
    uint16_t pos[SORT_IPT];
#ifdef HTGS_SORT_SMEM_ALL
    const bool via_smem = true;
#else
    const bool via_smem = shift >= 32;
#endif
#pragma unroll
    for (int h = 0; h < SORT_IPT; h += 8) {
        uint32_t peers[8], prev[8];
        if (via_smem) {
#pragma unroll
            for (int i = 0; i < 8; i++) {
                const uint32_t a = wmask_base + (((uint32_t)(k[h + i] >> shift) & mask) << 2);
                asm volatile("red.shared.or.b32 [%0], %1;" ::"r"(a), "r"(1u << lane) : "memory");
                asm volatile("ld.volatile.shared.u32 %0, [%1];" : "=r"(peers[i]) : "r"(a) : "memory");
                asm volatile("st.volatile.shared.u32 [%0], %1;" ::"r"(a), "r"(0u) : "memory");
            }
        } else {
#pragma unroll
            for (int i = 0; i < 8; i++) peers[i] = __match_any_sync(0xffffffffu, (uint32_t)(k[h + i] >> shift) & mask);
        }
#pragma unroll
        for (int i = 0; i < 8; i++) {
            const uint32_t d = (uint32_t)(k[h + i] >> shift) & mask;
            prev[i] = 0;
            if ((peers[i] & lanemask_lt) == 0) prev[i] = atomicAdd(&s_whist[warp][d], (uint32_t)__popc(peers[i]));  // lowest peer
        }
#pragma unroll
        for (int i = 0; i < 8; i++) {
            const uint32_t p0 = __shfl_sync(0xffffffffu, prev[i], __ffs(peers[i]) - 1);
            pos[h + i] = (uint16_t)(p0 + __popc(peers[i] & lanemask_lt));
        }
    }
    __syncthreads();

    // thread d owns digit d: exclusive scan over warps, tile total, look-back
    uint32_t run = 0;
#pragma unroll
    for (int w = 0; w < SORT_WARPS; w++) {
        const uint32_t c = s_whist[w][tid];
        s_whist[w][tid] = run;
        run += c;
    }
    uint32_t count = run;  // keys of this tile with digit `tid`, padding included
    if ((uint32_t)tid == mask) count -= (uint32_t)(SORT_TILE - valid);  // padding keys are all-ones
    uint32_t* my_status = status + (size_t)tile * SORT_RADIX + tid;
    *reinterpret_cast<volatile uint32_t*>(my_status) = (tile == 0 ? FLAG_PREFIX : FLAG_AGG) | count;

    const uint32_t local_start = block_excl_scan(run, s_tmp);
    s_local_start[tid] = local_start;

    // Chained look-back, SORT_LOOK predecessors per round trip: the status words are independent loads, so
    // a hop over SORT_LOOK tiles costs one L2 latency instead of SORT_LOOK (chains are ~10 hops long when a whole
    // wave of tiles starts together).
    uint32_t excl = 0;
    if (tile > 0) {
        int64_t t = (int64_t)tile - 1;
        uint32_t spins = 0;
        bool found = false;
        while (!found) {
            uint32_t sv[SORT_LOOK];
#pragma unroll
            for (int i = 0; i < SORT_LOOK; i++)
                sv[i] = (t - i >= 0) ? *reinterpret_cast<volatile uint32_t*>(status + (size_t)(t - i) * SORT_RADIX + tid)
                                     : (uint32_t)(2u << 30);  // before tile 0: an empty inclusive prefix (FLAG_PREFIX | 0)
            int used = 0;
#pragma unroll
            for (int i = 0; i < SORT_LOOK; i++) {
                if (found || used < i) continue;  // stop at the first word that is not ready yet
                const uint32_t f = sv[i] & FLAG_MASK;
                if (f == 0) continue;
                excl += sv[i] & VALUE_MASK;
                used = i + 1;
                if (f == FLAG_PREFIX) found = true;
            }
            if (found) break;
            if (used == 0) {
                if (++spins > (1u << 24)) {  // a predecessor never published: stop the grid loudly, never hang the GPU
                    printf("htgs: radix sort look-back timed out (tile %u, digit %d)\n", tile, tid);
                    __trap();  // the launch fails; the caller sees a CUDA error at its next synchronisation
                }
                __nanosleep(32);
            } else {
                spins = 0;
            }
            t -= used;
        }
        *reinterpret_cast<volatile uint32_t*>(my_status) = FLAG_PREFIX | ((excl + count) & VALUE_MASK);
    }
    s_goff[tid] = digit_start[tid] + excl - local_start;
    __syncthreads();

    // tile-local sorted order in shared memory.  Registers are the occupancy limit of this kernel, so
    // nothing but the keys and their 16-bit positions is kept live: the digit of every sorted slot is
    // parked in shared memory (1 B each) so the value pass can find its destination again, and the
    // values are only loaded once the keys have left, re-using the key buffer.
#pragma unroll
    for (int i = 0; i < SORT_IPT; i++) {
        const uint32_t d = (uint32_t)(k[i] >> shift) & mask;
        const uint32_t p = s_local_start[d] + s_whist[warp][d] + pos[i];
        pos[i] = (uint16_t)p;
        s_keys[p] = k[i];
        s_digit[p] = (uint8_t)d;
    }
    __syncthreads();
#pragma unroll
    for (int m = 0; m < SORT_IPT; m++) {
        const int j = tid + m * SORT_THREADS;
        if (j < valid) kout[s_goff[s_digit[j]] + (uint32_t)j] = s_keys[j];
    }
    uint32_t v[SORT_IPT];
#pragma unroll
    for (int i = 0; i < SORT_IPT; i++) {
        const int64_t idx = my0 + i * 32;
        v[i] = idx < n ? vin[idx] : 0u;
    }
    __syncthreads();
    uint32_t* s_vals = reinterpret_cast<uint32_t*>(s_keys);
#pragma unroll
    for (int i = 0; i < SORT_IPT; i++) s_vals[pos[i]] = v[i];
    __syncthreads();
#pragma unroll
    for (int m = 0; m < SORT_IPT; m++) {
        const int j = tid + m * SORT_THREADS;
        if (j < valid) vout[s_goff[s_digit[j]] + (uint32_t)j] = s_vals[j];
    }
}

// Sorts on key bits [0, end_bit).  With n_dev != nullptr the element count is read on the device
// (min(n, *n_dev)); n is then only the capacity the grid and the scratch were sized for, which lets
// the caller enqueue the sort before the count has reached the host.  After the call the sorted pairs are in
// (k_out, v_out) when the pass count is odd and in (k_in, v_in) when it is
// even -- callers pick the starting buffer with sort_num_passes() so the result
// lands where they want it.
int sort_pairs(int64_t n, int end_bit, uint64_t* k_in, uint32_t* v_in, uint64_t* k_out, uint32_t* v_out,
               void* scratch, cudaStream_t st, const uint32_t* n_dev) {
    if (n <= 0) return 0;
    if (n >= (int64_t(1) << 30)) {
        set_error("sort_pairs: n=%lld exceeds the 2^30 limit of the look-back words", (long long)n);
        return 1;
    }
    if (end_bit < 1 || end_bit > 64) {
        set_error("sort_pairs: bad end_bit %d", end_bit);
        return 1;
    }
    const int passes = sort_num_passes(end_bit);
    const int64_t tiles = sort_tiles(n);
    SortScratch s = carve(scratch, n);
    const size_t zero_bytes = header_bytes() + (size_t)passes * tiles * SORT_RADIX * sizeof(uint32_t);
    HTGS_CUDA_CHECK(cudaMemsetAsync(scratch, 0, zero_bytes, st));
    const int hist_blocks = (int)std::min<int64_t>(tiles, 148 * HTGS_HIST_GRID);
    sort_hist_kernel<<<hist_blocks, SORT_THREADS, 0, st>>>(k_in, n, n_dev, end_bit, passes, s.hist);
    HTGS_LAUNCH_CHECK("sort_hist_kernel");
    sort_scan_hist_kernel<<<passes, SORT_RADIX, 0, st>>>(s.hist);
    HTGS_LAUNCH_CHECK("sort_scan_hist_kernel");
    uint64_t* ka = k_in; uint32_t* va = v_in; uint64_t* kb = k_out; uint32_t* vb = v_out;
    for (int p = 0; p < passes; p++) {
        const int nbits = end_bit - 8 * p < 8 ? end_bit - 8 * p : 8;
        onesweep_kernel<<<(unsigned)tiles, SORT_THREADS, 0, st>>>(ka, va, kb, vb, n, n_dev, 8 * p, nbits, s.hist + p * SORT_RADIX,
                                                                   s.status + (size_t)p * tiles * SORT_RADIX,
                                                                   s.tickets + p);
        HTGS_LAUNCH_CHECK("onesweep_kernel");
        uint64_t* tk = ka; ka = kb; kb = tk;
        uint32_t* tv = va; va = vb; vb = tv;
    }
    return 0;
}

}  // namespace htgs
