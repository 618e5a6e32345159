// K2 + K3: (Gaussian, tile) key generation and per-tile ranges.
// Replaces duplicateWithKeys (cuda_rasterizer/rasterizer_impl.cu:70-111), the
// second half of the inclusive scan (:278), cudaMemset + identifyTileRanges
// (:311-318, :116-138).
//
// make_keys is instance-parallel: a block owns 256 consecutive Gaussians, scans
// their tile counts in shared memory, then every thread emits consecutive
// instances (binary search of the owning Gaussian), so the 12 B/instance of
// key+value stores are fully coalesced instead of one divergent rect loop per
// thread.  Emission order is the reference's (Gaussian index, then rect row,
// then column), which the stable sort turns into the same final order.
#include "gs_math.cuh"

namespace htgs {

__global__ void __launch_bounds__(HTGS_PRE_BLOCK)
make_keys_kernel(int P, int gx, int gy, const int* __restrict__ radii, const float2* __restrict__ means2D,
                 const float4* __restrict__ feat, const uint32_t* __restrict__ tiles_touched,
                 const uint32_t* __restrict__ block_offsets, uint32_t* __restrict__ point_offsets,
                 uint64_t* __restrict__ keys, uint32_t* __restrict__ values, uint32_t capacity) {
    __shared__ uint32_t s_off[HTGS_PRE_BLOCK + 1];  // exclusive offsets inside the block
    __shared__ uint32_t s_warp[HTGS_PRE_BLOCK / 32];
    __shared__ uint32_t s_rect[HTGS_PRE_BLOCK];     // rmin.x | rmin.y<<12 | width<<24 .. packed below
    __shared__ uint32_t s_rectw[HTGS_PRE_BLOCK];
    __shared__ uint32_t s_depth[HTGS_PRE_BLOCK];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int idx = blockIdx.x * HTGS_PRE_BLOCK + tid;
    uint32_t cnt = 0;
    if (idx < P) {
        cnt = tiles_touched[idx];
        if (cnt > 0) {
            uint2 rmin, rmax;
            get_rect(means2D[idx].x, means2D[idx].y, radii[idx], gx, gy, rmin, rmax);
            s_rect[tid] = rmin.y * (uint32_t)gx + rmin.x;  // tile id of the rect's first tile
            s_rectw[tid] = rmax.x - rmin.x;
            s_depth[tid] = __float_as_uint(feat[idx].w);
        }
    }
    uint32_t inc = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t n = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += n;
    }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    uint32_t warp_off = 0;
#pragma unroll
    for (int w = 0; w < HTGS_PRE_BLOCK / 32; w++)
        if (w < warp) warp_off += s_warp[w];
    const uint32_t base = block_offsets[blockIdx.x];
    s_off[tid] = warp_off + inc - cnt;
    if (tid == HTGS_PRE_BLOCK - 1) s_off[HTGS_PRE_BLOCK] = warp_off + inc;
    if (idx < P) point_offsets[idx] = base + warp_off + inc;  // inclusive, as the reference stores it
    __syncthreads();

    const uint32_t total = s_off[HTGS_PRE_BLOCK];
    for (uint32_t o = tid; o < total; o += HTGS_PRE_BLOCK) {
        // largest g with s_off[g] <= o  (entries with cnt==0 share offsets with
        // their successor and are skipped by taking the LAST such g)
        int lo = 0, hi = HTGS_PRE_BLOCK;  // invariant: s_off[lo] <= o < s_off[hi]
#pragma unroll
        for (int it = 0; it < 8; it++) {
            const int mid = (lo + hi) >> 1;
            if (s_off[mid] <= o) lo = mid; else hi = mid;
        }
        const uint32_t k = o - s_off[lo];
        const uint32_t w = s_rectw[lo];
        const uint32_t row = k / w, col = k - row * w;
        const uint32_t tile = s_rect[lo] + row * (uint32_t)gx + col;
        if (base + o < capacity) {  // only relevant when the buffers were sized from a predicted count
            keys[(size_t)base + o] = ((uint64_t)tile << 32) | s_depth[lo];
            values[(size_t)base + o] = (uint32_t)(blockIdx.x * HTGS_PRE_BLOCK + lo);
        }
    }
}

// One WARP per tile: lower bound of (tile << 32) in the sorted keys gives the range start, lower bound of
// ((tile+1) << 32) its end.  Tiles with no instances get (0,0) like the reference's memset.
// The kernel is a chain of dependent L2 round trips, so the search is 32-ary: every step the 32 lanes probe 32 evenly
// spaced keys of the remaining interval for both bounds at once and a ballot picks the sub-interval -- 5 steps for
// 3.4 M keys instead of the 22 of a binary search (round 1 ran two 4-ary searches per THREAD, 11 steps).  Measured:
// 12.7 -> 8.6 us at config 5 (1024 tiles), unchanged 12.8 us at config 3 (8160 tiles, where the probes of neighbouring
// tiles stop sharing cache lines early).
__global__ void __launch_bounds__(256)
tile_ranges_kernel(int64_t R, const uint32_t* __restrict__ R_dev, int tiles, const uint64_t* __restrict__ keys,
                   uint2* __restrict__ ranges) {
    const int t = (int)((blockIdx.x * (unsigned)blockDim.x + threadIdx.x) >> 5);
    const int lane = threadIdx.x & 31;
    if (t >= tiles) return;
    if (R_dev) R = min(R, (int64_t)*R_dev);
    const uint64_t target[2] = {(uint64_t)t << 32, (uint64_t)(t + 1) << 32};
    int64_t lo[2] = {0, 0}, hi[2] = {R, R};  // the first index with key >= target lies in [lo, hi]
    while (lo[0] < hi[0] || lo[1] < hi[1]) {
        int64_t at[2], step[2];
        bool ge[2];
#pragma unroll
        for (int s = 0; s < 2; s++) {
            step[s] = (hi[s] - lo[s] + 31) >> 5;  // >= 1 while the interval is not empty
            at[s] = lo[s] + (int64_t)lane * step[s];
            ge[s] = true;                         // probes past the interval count as "key >= target"
        }
        uint64_t k[2] = {0, 0};
#pragma unroll
        for (int s = 0; s < 2; s++)
            if (lo[s] < hi[s] && at[s] < hi[s]) k[s] = __ldg(keys + at[s]);
#pragma unroll
        for (int s = 0; s < 2; s++) {
            if (lo[s] < hi[s]) {
                if (at[s] < hi[s]) ge[s] = k[s] >= target[s];
                const uint32_t m = __ballot_sync(0xffffffffu, ge[s]);
                const int f = m ? __ffs(m) - 1 : 32;  // first probe with key >= target (32: none)
                // keys at probes < f are below the target, the key at probe f is not: the bound lies in (probe f-1, probe f]
                const int64_t nlo = f == 0 ? lo[s] : lo[s] + (int64_t)(f - 1) * step[s] + 1;
                const int64_t nhi = f == 32 ? hi[s] : min(hi[s], lo[s] + (int64_t)f * step[s]);
                lo[s] = nlo;
                hi[s] = nhi;
            } else {
                (void)__ballot_sync(0xffffffffu, false);
            }
        }
    }
    if (lane == 0) ranges[t] = lo[0] == lo[1] ? make_uint2(0, 0) : make_uint2((uint32_t)lo[0], (uint32_t)lo[1]);
}

int launch_make_keys(int P, int W, int H, const int* radii, const GeomView& g, uint64_t* keys, uint32_t* values,
                     int64_t capacity, cudaStream_t st) {
    const int gx = (W + HTGS_TILE - 1) / HTGS_TILE, gy = (H + HTGS_TILE - 1) / HTGS_TILE;
    const int nb = (P + HTGS_PRE_BLOCK - 1) / HTGS_PRE_BLOCK;
    make_keys_kernel<<<nb, HTGS_PRE_BLOCK, 0, st>>>(P, gx, gy, radii, g.means2D, g.feat, g.tiles_touched, g.block_sums,
                                                    g.point_offsets, keys, values, (uint32_t)capacity);
    HTGS_LAUNCH_CHECK("make_keys_kernel");
    return 0;
}

int launch_tile_ranges(int64_t R, const uint32_t* R_dev, int W, int H, const uint64_t* keys_sorted, uint2* ranges,
                       cudaStream_t st) {
    const int gx = (W + HTGS_TILE - 1) / HTGS_TILE, gy = (H + HTGS_TILE - 1) / HTGS_TILE;
    const int tiles = gx * gy;
    tile_ranges_kernel<<<(tiles + 7) / 8, 256, 0, st>>>(R, R_dev, tiles, keys_sorted, ranges);
    HTGS_LAUNCH_CHECK("tile_ranges_kernel");
    return 0;
}

// Number of bits needed to write n (at least 1): the tile-id part of the sort key occupies bits [32, 32 + this),
// the same bit count the reference derives for its CUB call (rasterizer_impl.cu:301).
uint32_t higher_msb(uint32_t n) { return n ? 32u - (uint32_t)__builtin_clz(n) : 1u; }

}  // namespace htgs
