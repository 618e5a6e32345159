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

// One thread per tile: lower bound of (tile << 32) in the sorted keys gives the
// range start, lower bound of ((tile+1) << 32) its end.  Tiles with no
// instances get (0,0) like the reference's memset.
__global__ void tile_ranges_kernel(int64_t R, const uint32_t* __restrict__ R_dev, int tiles,
                                   const uint64_t* __restrict__ keys, uint2* __restrict__ ranges) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= tiles) return;
    if (R_dev) R = min(R, (int64_t)*R_dev);
    // both lower bounds advance in the same loop, three probes per step each (4-ary search): the kernel is a chain of
    // dependent L2 round trips, so it pays to keep six independent loads in flight and halve the number of steps
    uint32_t bounds[2];
    const uint64_t target0 = (uint64_t)t << 32, target1 = (uint64_t)(t + 1) << 32;
    int64_t lo0 = 0, hi0 = R, lo1 = 0, hi1 = R;  // first index with key >= target lies in [lo, hi]
    while (lo0 < hi0 || lo1 < hi1) {
        const int64_t q0 = (hi0 - lo0) >> 2, q1 = (hi1 - lo1) >> 2;
        const int64_t a0 = lo0 + q0, b0 = lo0 + ((hi0 - lo0) >> 1), c0 = hi0 - 1 - q0;
        const int64_t a1 = lo1 + q1, b1 = lo1 + ((hi1 - lo1) >> 1), c1 = hi1 - 1 - q1;
        const bool go0 = lo0 < hi0, go1 = lo1 < hi1;
        uint64_t ka0 = 0, kb0 = 0, kc0 = 0, ka1 = 0, kb1 = 0, kc1 = 0;
        if (go0) { ka0 = __ldg(keys + a0); kb0 = __ldg(keys + b0); kc0 = __ldg(keys + c0); }
        if (go1) { ka1 = __ldg(keys + a1); kb1 = __ldg(keys + b1); kc1 = __ldg(keys + c1); }
        if (go0) {  // a0 <= b0 <= c0 < hi0
            if (ka0 >= target0) hi0 = a0;
            else if (kb0 >= target0) { lo0 = a0 + 1; hi0 = b0; }
            else if (kc0 >= target0) { lo0 = b0 + 1; hi0 = c0; }
            else lo0 = c0 + 1;
        }
        if (go1) {
            if (ka1 >= target1) hi1 = a1;
            else if (kb1 >= target1) { lo1 = a1 + 1; hi1 = b1; }
            else if (kc1 >= target1) { lo1 = b1 + 1; hi1 = c1; }
            else lo1 = c1 + 1;
        }
    }
    bounds[0] = (uint32_t)lo0;
    bounds[1] = (uint32_t)lo1;
    ranges[t] = bounds[0] == bounds[1] ? make_uint2(0, 0) : make_uint2(bounds[0], bounds[1]);
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
    tile_ranges_kernel<<<(tiles + 127) / 128, 128, 0, st>>>(R, R_dev, tiles, keys_sorted, ranges);
    HTGS_LAUNCH_CHECK("tile_ranges_kernel");
    return 0;
}

// Number of bits needed to write n (at least 1): the tile-id part of the sort key occupies bits [32, 32 + this),
// the same bit count the reference derives for its CUB call (rasterizer_impl.cu:301).
uint32_t higher_msb(uint32_t n) { return n ? 32u - (uint32_t)__builtin_clz(n) : 1u; }

}  // namespace htgs
