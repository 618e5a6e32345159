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

// One WARP per tile BOUNDARY: b(t) = first index whose key carries a tile id >= t, for t = 0 .. tiles.  Tile t owns
// [b(t), b(t+1)), so warp t writes ranges[t].x and ranges[t-1].y; tiles with no instances get (0,0) like the
// reference's memset (the warp sees that from the keys on either side of its boundary).
// The search is 32-ary: every step the 32 lanes probe 32 evenly spaced keys of the remaining interval and a ballot
// picks the sub-interval -- 5-7 steps for 3.4 M keys instead of the 22 of a binary search.  Only the high word of a
// key (the tile id) is read and compared, all index arithmetic is 32-bit.  ncu (config 3): the two-bounds-per-warp
// 64-bit version of this kernel executed 665 instructions per warp and was ISSUE-bound at 11.3 us, not latency-bound.
__global__ void __launch_bounds__(256)
tile_ranges_kernel(uint32_t R, const uint32_t* __restrict__ R_dev, int tiles, const uint64_t* __restrict__ keys,
                   uint2* __restrict__ ranges) {
    const int t = (int)((blockIdx.x * (unsigned)blockDim.x + threadIdx.x) >> 5);
    const uint32_t lane = threadIdx.x & 31;
    if (t > tiles) return;
    if (R_dev) R = min(R, *R_dev);
    const uint32_t* tile_of = reinterpret_cast<const uint32_t*>(keys) + 1;  // tile id of key i: tile_of[2 * i]
    uint32_t lo = 0, hi = R;  // the boundary lies in [lo, hi]
    while (lo < hi) {
        const uint32_t step = (hi - lo + 31) >> 5;  // >= 1
        const uint32_t at = lo + lane * step;  // R < 2^30 (checked at the C ABI): no overflow, 2 * at fits too
        bool ge = true;                             // probes past the interval count as "tile id >= t"
        if (at < hi) ge = __ldg(tile_of + 2 * at) >= (uint32_t)t;
        const uint32_t m = __ballot_sync(0xffffffffu, ge);
        const uint32_t f = m ? (uint32_t)__ffs(m) - 1 : 32;  // first such probe (32: none)
        // probes below f are under the boundary, probe f is not: the boundary lies in (probe f-1, probe f]
        const uint32_t nlo = f == 0 ? lo : lo + (f - 1) * step + 1;
        const uint32_t nhi = f == 32 ? hi : min(hi, lo + f * step);
        lo = nlo;
        hi = nhi;
    }
    uint32_t* out = reinterpret_cast<uint32_t*>(ranges);
    if (lane == 0 && t < tiles) out[2 * t] = (lo < R && __ldg(tile_of + 2 * lo) == (uint32_t)t) ? lo : 0;
    if (lane == 1 && t > 0) out[2 * t - 1] = (lo > 0 && __ldg(tile_of + 2 * (lo - 1)) == (uint32_t)(t - 1)) ? lo : 0;
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
    tile_ranges_kernel<<<(tiles + 1 + 7) / 8, 256, 0, st>>>((uint32_t)R, R_dev, tiles, keys_sorted, ranges);
    HTGS_LAUNCH_CHECK("tile_ranges_kernel");
    return 0;
}

// Number of bits needed to write n (at least 1): the tile-id part of the sort key occupies bits [32, 32 + this),
// the same bit count the reference derives for its CUB call (rasterizer_impl.cu:301).
uint32_t higher_msb(uint32_t n) { return n ? 32u - (uint32_t)__builtin_clz(n) : 1u; }

}  // namespace htgs
