/*
 * htgs.h -- C ABI of the B200-native differentiable Gaussian rasterizer.
 *
 * This is the drop-in boundary: a plain `extern "C"` shared library
 * (3dgs_hierarchical_training_b200/libhtgs.so) with raw device pointers and
 * sizes, no torch types.  The Python plugin mirror
 * (3dgs_hierarchical_training_b200/diff_gaussian_rasterization, simple_knn)
 * binds it with ctypes; INTEGRATION.md shows the binding a maintainer of the
 * reference would add.
 *
 * Reference interfaces replaced (paths relative to /root/reference/submodules):
 *   diff-gaussian-rasterization/rasterize_points.h:18-70   (torch glue)
 *   diff-gaussian-rasterization/cuda_rasterizer/rasterizer.h:20-92
 *   simple-knn/spatial.h:13 , simple-knn/simple_knn.h:15-19
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless the name ends in _host;
 *   - `stream` is a cudaStream_t passed as void*; all work is enqueued on it;
 *   - functions return 0 on success, non-zero on error; htgs_last_error()
 *     returns the message of the calling thread's last failure;
 *   - matrices use the reference's row-vector convention (flat m[4*c+r]);
 *   - optional inputs are NULL exactly where the reference passes an empty
 *     tensor (data_ptr()==0);
 *   - work buffers are caller-owned; their sizes come from the *_bytes() queries.
 *     The rasterizer, sort, knn, optimizer and row-surgery entry points keep no
 *     state between calls and are re-entrant per (device, stream).  The
 *     exceptions, each documented at its declaration: the built-in scratch
 *     array of htgs_sort_pairs_static (shared per device), the SSIM window
 *     constants of htgs_l1_ssim_* (uploaded once per device under a lock, first
 *     call outside stream capture), the side stream + event pair of
 *     htgs_backward_compact (one per device, created on first use under a
 *     lock; compact-mode backwards on a device come from one stream at a
 *     time), the launch-shape override of
 *     htgs_allreduce_set_launch_shape, and the htgs_profile_* stage timers
 *     (process-wide measurement aid, not thread-safe).
 */
#ifndef HTGS_H_INCLUDED
#define HTGS_H_INCLUDED

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HTGS_ABI_VERSION 1

int htgs_abi_version(void);
const char* htgs_last_error(void);

/* ---- work-buffer sizes (replace required<GeometryState/ImageState/BinningState>,
 *      cuda_rasterizer/rasterizer_impl.h:66-72) ---- */
size_t htgs_geom_bytes(int P);
size_t htgs_image_bytes(int W, int H);
size_t htgs_binning_bytes(int64_t R, int W, int H);
/* scratch for the backward pass: per-Gaussian gradient accumulators */
size_t htgs_backward_scratch_bytes(int P);

/* Byte offsets of the named fields inside the opaque buffers, for parity tests
 * and debugging.  Returns the number of fields written (<= cap). */
enum {
    HTGS_GEOM_MEANS2D = 0,     /* float2[P]  pixel centre                         */
    HTGS_GEOM_CONIC_OPACITY,   /* float4[P]  conic.xyz, opacity                   */
    HTGS_GEOM_FEAT,            /* float4[P]  r,g,b (SH colour or precomp), depth  */
    HTGS_GEOM_CLAMPED,         /* uint8[P]   bit c set: channel c clamped at 0    */
    HTGS_GEOM_TILES_TOUCHED,   /* uint32[P]                                       */
    HTGS_GEOM_POINT_OFFSETS,   /* uint32[P]  inclusive prefix sum                 */
    HTGS_GEOM_BLOCK_SUMS,      /* uint32[ceil(P/256)+1] internal                  */
    HTGS_GEOM_NUM_RENDERED,    /* uint32[1]  R                                    */
    HTGS_GEOM_NFIELDS
};
enum {
    HTGS_BIN_KEYS_SORTED = 0,  /* uint64[R] */
    HTGS_BIN_VALUES_SORTED,    /* uint32[R]  == point_list; physically first (offset 0) */
    HTGS_BIN_KEYS_ALT,         /* uint64[R]  ping-pong partner */
    HTGS_BIN_VALUES_ALT,       /* uint32[R] */
    HTGS_BIN_SORT_SCRATCH,     /* histograms + look-back state */
    HTGS_BIN_NFIELDS
};
enum {
    HTGS_IMG_RANGES = 0,       /* uint2[tiles] */
    HTGS_IMG_N_CONTRIB,        /* uint32[W*H]  */
    HTGS_IMG_NFIELDS
};
int htgs_geom_layout(int P, size_t* offsets, int cap);
int htgs_binning_layout(int64_t R, int W, int H, size_t* offsets, int cap);
int htgs_image_layout(int W, int H, size_t* offsets, int cap);

/* ---- forward, phase 1: per-Gaussian preprocess + tile-count prefix sum ----
 * Replaces FORWARD::preprocess + cub::DeviceScan::InclusiveSum + the blocking
 * 4-byte copy (cuda_rasterizer/rasterizer_impl.cu:249-282, forward.cu:155-256).
 * Writes radii[P] (0 for culled) and the geometry buffer, then waits for the
 * stream and stores R = number of (Gaussian, tile) instances in
 * *num_rendered_host so the caller can size the binning buffer. */
int htgs_forward_preprocess(int P, int D, int M,
                            const float* means3D, const float* scales, float scale_modifier,
                            const float* rotations, const float* opacities, const float* shs,
                            const float* cov3D_precomp, const float* colors_precomp,
                            const float* viewmatrix, const float* projmatrix, const float* campos,
                            int W, int H, float tanfovx, float tanfovy, int prefiltered,
                            void* geom_buf, int* radii, int64_t* num_rendered_host, void* stream);

/* ---- forward, phase 2: key generation, radix sort, tile ranges, blend ----
 * Replaces duplicateWithKeys + cub::DeviceRadixSort::SortPairs + memset +
 * identifyTileRanges + FORWARD::render (rasterizer_impl.cu:284-336,
 * forward.cu:261-381).  out_color[3,H,W], out_depth[H,W], out_alpha[H,W]. */
int htgs_forward_render(int P, int64_t R, int W, int H, const float* background,
                        const int* radii, void* geom_buf, void* binning_buf, void* image_buf,
                        float* out_color, float* out_depth, float* out_alpha, void* stream);

/* ---- forward without the host round trip ----
 * The reference stalls the host on a 4-byte copy of R in every forward (rasterizer_impl.cu:282)
 * because R sizes the binning buffer.  These two entry points let the caller size that buffer
 * from a PREDICTION (e.g. the previous frame's R plus a margin) and enqueue the whole forward
 * without waiting:
 *   htgs_forward_preprocess_async   same as htgs_forward_preprocess but no synchronisation; R is
 *                                   copied asynchronously into *num_rendered_pinned_host (must be
 *                                   pinned host memory; valid once the stream reached this point,
 *                                   e.g. after an event recorded right after the call);
 *   htgs_forward_render_capacity    same as htgs_forward_render with a binning buffer sized by
 *                                   htgs_binning_bytes(capacity, W, H); the kernels take the true R
 *                                   from device memory.  If R turns out larger than capacity nothing
 *                                   is written out of bounds but the outputs are invalid: the caller
 *                                   checks R <= capacity afterwards and, if not, reruns
 *                                   htgs_forward_render with an exact-size buffer.
 * The backward reads only the sorted Gaussian list, which sits at offset 0 of the binning buffer
 * for every capacity, so htgs_backward needs no knowledge of the capacity that was used. */
int htgs_forward_preprocess_async(int P, int D, int M,
                                  const float* means3D, const float* scales, float scale_modifier,
                                  const float* rotations, const float* opacities, const float* shs,
                                  const float* cov3D_precomp, const float* colors_precomp,
                                  const float* viewmatrix, const float* projmatrix, const float* campos,
                                  int W, int H, float tanfovx, float tanfovy, int prefiltered,
                                  void* geom_buf, int* radii, uint32_t* num_rendered_pinned_host, void* stream);
int htgs_forward_render_capacity(int P, int64_t capacity, int W, int H, const float* background,
                                 const int* radii, void* geom_buf, void* binning_buf, void* image_buf,
                                 float* out_color, float* out_depth, float* out_alpha, void* stream);

/* ---- backward ----
 * Replaces Rasterizer::backward (rasterizer_impl.cu:343-447; backward.cu).
 * Every non-NULL dL_* output is fully written (zeros for culled Gaussians), so
 * the caller may allocate them uninitialised (the reference zero-fills ten
 * tensors per call, rasterize_points.cu:158-167).
 *   dL_dmeans2D[P,3] dL_dopacity[P] dL_dmeans3D[P,3]            always
 *   dL_dcolors[P,3]   when colors_precomp != NULL (else may be NULL)
 *   dL_dsh[P,M,3]     when shs != NULL
 *   dL_dscales[P,3], dL_drotations[P,4]  when scales != NULL
 *   dL_dcov3D[P,6]    when cov3D_precomp != NULL (else may be NULL)
 * R: pass the instance count the forward's binning buffer was SIZED with -- the R of htgs_forward_render, or the
 * capacity of htgs_forward_render_capacity.  The forward blend leaves the per-instance reach masks in that buffer and
 * records the count it carved the buffer with; when the backward's count matches it reuses them, otherwise (e.g. the
 * exact R after a capacity-sized forward) it recomputes them: same result, slightly slower.                        */
int htgs_backward(int P, int D, int M, int64_t R, int W, int H, const float* background,
                  const float* means3D, const float* shs, const float* colors_precomp,
                  const float* alphas, const float* scales, float scale_modifier,
                  const float* rotations, const float* cov3D_precomp,
                  const float* viewmatrix, const float* projmatrix, const float* campos,
                  float tanfovx, float tanfovy, const int* radii,
                  const void* geom_buf, const void* binning_buf, const void* image_buf,
                  const float* dL_dout_color, const float* dL_dout_depth, const float* dL_dout_alpha,
                  void* scratch,
                  float* dL_dmeans2D, float* dL_dopacity, float* dL_dcolors, float* dL_dmeans3D,
                  float* dL_dcov3D, float* dL_dsh, float* dL_dscales, float* dL_drotations,
                  void* stream);

/* Replaces Rasterizer::markVisible (rasterizer_impl.cu:141-153): present[i] =
 * view-space z > 0.2.  `present` is uint8/bool[P]. */
int htgs_mark_visible(int P, const float* means3D, const float* viewmatrix, const float* projmatrix,
                      uint8_t* present, void* stream);

/* Stand-alone stable LSD radix sort of (uint64 key, uint32 value) pairs on key
 * bits [0, end_bit), the contract of cub::DeviceRadixSort::SortPairs as used
 * at rasterizer_impl.cu:304-309.  keys_in/values_in are clobbered.  scratch
 * comes from htgs_sort_scratch_bytes(n). */
size_t htgs_sort_scratch_bytes(int64_t n);
int htgs_sort_pairs(int64_t n, int end_bit, uint64_t* keys_in, uint32_t* values_in,
                    uint64_t* keys_out, uint32_t* values_out, void* scratch, void* stream);

/* htgs_sort_pairs_static has no counterpart in the reference.  It is htgs_sort_pairs for callers
 * that may not allocate device memory at all -- a plugin living inside a host application's
 * pre-reserved arena, a serving process with a fixed and accounted pool, a latency-critical loop
 * that cannot afford the device-wide synchronisation of cudaMalloc/cudaFree.  They still pass
 * their own input and output arrays, but the scratch comes from a fixed array inside the library
 * image (HTGS_SORT_STATIC_SCRATCH_BYTES, held by every GPU that loads the library until exit), so:
 *   - at most HTGS_SORT_STATIC_MAX_PAIRS pairs; a larger n is refused with an error message;
 *   - nothing is allocated or freed, by this call or by anything it uses;
 *   - all calls on a device share that one array.  Each call clears the part it needs first, so
 *     what the previous call left behind is harmless, but calls whose GPU work overlaps (other
 *     streams, other host threads) corrupt each other: keep them ordered;
 *   - the first call on each device looks up the array's device address; issue that first call
 *     outside any stream capture.  The address is cached per device for the life of the process
 *     (do not use this entry point on a device after cudaDeviceReset).
 * Same contract as htgs_sort_pairs otherwise (stable, key bits [0,end_bit), inputs clobbered). */
#define HTGS_SORT_STATIC_MAX_PAIRS (1LL << 20)
#define HTGS_SORT_STATIC_SCRATCH_BYTES 2105856
int htgs_sort_pairs_static(int64_t n, int end_bit, uint64_t* keys_in, uint32_t* values_in,
                           uint64_t* keys_out, uint32_t* values_out, void* stream);

/* ---- measurement aids (no counterpart in the reference, which has no timers or counters) ----
 * htgs_profile_enable(1) makes the forward/backward entry points record a CUDA event pair on the
 * caller's stream around each stage; htgs_profile_read() waits for them, returns the summed
 * milliseconds and launch counts per stage (order: preprocess, scan, make_keys, sort, tile_ranges,
 * blend_fwd, zero_acc, blend_bwd, geom_bwd) and clears the log.  Process-wide, not thread-safe.
 * htgs_blend_stats() re-walks the tile lists of the last forward and writes HTGS_BLEND_NSTATS counters
 * to stats (device, uint64 each): [0] pair tests of the reference's per-pixel loop, [1] blended pairs,
 * [2] pair tests this library executes after reach-mask culling, [3..9] number of (warp, Gaussian)
 * visits in which 0, 1, 2-4, 5-8, 9-16, 17-24, 25-32 of the warp's 32 pixels blended the Gaussian,
 * [10] (Gaussian, tile) instances staged before the tile saturated whose Gaussian cannot reach any pixel of the tile. */
#define HTGS_BLEND_NSTATS 11
#define HTGS_NUM_STAGES 9
void htgs_profile_enable(int on);
int htgs_profile_read(double* ms_sum, int* counts, int cap);
int htgs_blend_stats(int P, int64_t R, int W, int H, const void* geom_buf, const void* binning_buf,
                     const void* image_buf, unsigned long long* stats, void* stream);

/* ---- per-step tail after the backward (opt-in, outside the plugin surface) ----
 * htgs_densification_stats replaces the three boolean-mask updates of
 * trainer/ht3dgs_trainer.py:143-148 + scene/gaussian_model_ht.py:718-721 (each of which syncs the host):
 * for radii[i] > 0:  max_radii2D[i] = max(max_radii2D[i], radii[i]);
 *                    xyz_gradient_accum[i] += ||grad_means2D[i, 0:2]||;  denom[i] += 1.
 * htgs_adam_step is one fused launch for all parameter groups of the torch.optim.Adam the model builds at
 * scene/gaussian_model_ht.py:275-289 (betas, eps shared; lr per group; no weight decay / amsgrad), `step`
 * counting from 1.  Group structs are read on the host; pointers inside are device pointers. */
#define HTGS_ADAM_MAX_GROUPS 8
typedef struct htgs_adam_group {
    float* param;
    const float* grad;
    float* exp_avg;
    float* exp_avg_sq;
    long long numel;
    float lr;
} htgs_adam_group;
int htgs_densification_stats(int P, const int* radii, const float* grad_means2D, float* max_radii2D,
                             float* xyz_gradient_accum, float* denom, void* stream);
int htgs_adam_step(int n_groups, const htgs_adam_group* groups, float beta1, float beta2, float eps,
                   long long step, void* stream);

/* ---- gradient all-reduce over NVLink peer memory (opt-in, no reference counterpart) ----
 * peer_buffers: HOST array of `world` device pointers to the same-sized float32 buffer on every rank, entry `rank`
 * being the caller's own (all mapped into the caller's address space, e.g. by torch symmetric memory or cudaIpc).
 * Two-shot: this rank sums slice `rank` of all buffers in the fixed order 0..world-1 and writes the result into that
 * slice of EVERY buffer; run on all ranks it leaves the element-wise sum everywhere, bit-identical on all ranks.
 * The caller provides the cross-GPU ordering: every rank's buffer must be complete before any rank's kernel starts,
 * and every rank's kernel must have finished before any rank reads the result (a barrier before and after).
 * numel must be a multiple of 4; world <= 8. */
int htgs_allreduce_p2p(const void* const* peer_buffers, int rank, int world, long long numel, void* stream);
/* Same contract, but through one NVSwitch MULTICAST address that maps the buffer of every rank (cuMulticast* /
 * torch symmetric memory `multicast_ptr`): the switch sums the ranks' elements on the load (multimem.ld_reduce) and
 * replicates the store (multimem.st), so each GPU moves about one buffer per direction at any world size.  The
 * summation order inside the switch is the hardware's; all ranks still receive the same bits. */
int htgs_allreduce_multimem(void* multicast_buffer, int rank, int world, long long numel, void* stream);
/* ---- view-parallel steps without shipping the expanded SH gradient (opt-in, no reference counterpart) ----
 * Per view dL/dsh[i][k][c] = basis_k(direction of Gaussian i from the camera) * r[i][c] with r the clamp-masked colour
 * gradient, so ranks exchange r (12 B/Gaussian/view) instead of 192 B/Gaussian and each rebuilds the sum itself.
 *   htgs_sh_slot_floats(P)   floats in one view slot: r of P Gaussians (rows padded to a multiple of 4) + camera position.
 *   htgs_backward_compact    htgs_backward on the SH path, except that dL_dsh is not produced: r and campos are written to
 *                            rgb_slot (16-byte aligned, htgs_sh_slot_floats(P) floats) right after the blend backward.  With
 *                            rgb_slot_is_multicast != 0 the slot address is an NVSwitch multicast address and the writes are
 *                            multimem.st: the slot arrives in every rank's buffer (the caller provides the cross-GPU barrier
 *                            before anyone reads, as for htgs_allreduce_multimem).
 *   htgs_sh_expand           dL_dsh[P,M,3] = sum over the n_views slots (HOST array of device pointers, local or peer-mapped,
 *                            <= 16) of basis (x) r, in slot order -- identical bits on every rank that passes the same order. */
size_t htgs_sh_slot_floats(int P);
int htgs_backward_compact(int P, int D, int M, int64_t R, int W, int H, const float* background, const float* means3D,
                          const float* shs, const float* alphas, const float* scales, float scale_modifier,
                          const float* rotations, const float* cov3D_precomp, const float* viewmatrix, const float* projmatrix,
                          const float* campos, float tanfovx, float tanfovy, const int* radii, const void* geom_buf,
                          const void* binning_buf, const void* image_buf, const float* dL_dout_color, const float* dL_dout_depth,
                          const float* dL_dout_alpha, void* scratch, float* dL_dmeans2D, float* dL_dopacity, float* dL_dmeans3D,
                          float* dL_dcov3D, float* dL_dscales, float* dL_drotations, float* rgb_slot, int rgb_slot_is_multicast,
                          void* stream);
int htgs_sh_expand(int P, int D, int M, int n_views, const float* means3D, const float* const* slots_host, float* dL_dsh,
                   void* stream);

/* Tuning aid for htgs_allreduce_multimem: CTAs per SM (1..16, default 2) and threads per CTA (32..512, default 512)
 * of its kernel, process-wide.  Measured flat on 4 and 8 B200s (0.57-0.60 ms for 236 MB at every shape). */
int htgs_allreduce_set_launch_shape(int ctas_per_sm, int threads);

/* ---- row surgery for prune / clone (first half; opt-in) ----
 * Replaces the per-tensor boolean-mask indexing of _prune_optimizer / prune_points and the per-tensor
 * torch.cat of cat_tensors_to_optimizer / densification_postfix in the clone case
 * (scene/gaussian_model_ht.py:548-616, :656-672): scan the mask once, then move the rows of up to
 * HTGS_ROWS_MAX_TENSORS float32 tensors in one launch.
 *   htgs_row_scan      offsets[i] = number of non-zero mask entries before i (kept in scratch); the total
 *                      is copied asynchronously to *count_pinned_host (pinned host memory, valid once the
 *                      stream reached this point) so the caller can size the outputs.
 *   htgs_rows_compact  dst[offsets[i]] = src[i] where mask[i] != 0          == torch src[mask]
 *   htgs_rows_append   dst[0:P] = src; dst[P + offsets[i]] = src[i] (or 0 when zero_new) where mask[i] != 0
 *                                                                            == torch.cat((src, src[mask])) / cat((src, zeros))
 *   htgs_rows_split    dst = cat(src[mask == 0], repeat(src[mask != 0], N)): what densification_postfix followed by
 *                      prune_points(cat(mask, zeros)) leaves in densify_and_split (:630-654).  count = the total
 *                      htgs_row_scan reported.  zero_new = 1: the N*count new rows are zeros (Adam moments);
 *                      zero_new = 2: they are not written (the caller fills computed positions / scales).
 * mask: uint8/bool [P]; every tensor is [P, row_words] 32-bit words, contiguous; dst sized by the caller. */
#define HTGS_ROWS_MAX_TENSORS 24
typedef struct {
    const void* src;
    void* dst;
    int row_words; /* 32-bit words per row */
    int zero_new;  /* append / split: 1 = new rows are zero (Adam moments of new Gaussians); split only: 2 = new rows untouched */
} htgs_row_tensor;
size_t htgs_row_scan_scratch_bytes(int P);
int htgs_row_scan(int P, const uint8_t* mask, void* scratch, uint32_t* count_pinned_host, void* stream);
int htgs_rows_compact(int P, const uint8_t* mask, const void* scratch, int n_tensors, const htgs_row_tensor* tensors, void* stream);
int htgs_rows_append(int P, const uint8_t* mask, const void* scratch, int n_tensors, const htgs_row_tensor* tensors, void* stream);
int htgs_rows_split(int P, int N, uint32_t count, const uint8_t* mask, const void* scratch, int n_tensors,
                    const htgs_row_tensor* tensors, void* stream);

/* ---- merge-time operations of the hierarchical trainer (second half; opt-in) ----
 * htgs_importance_accumulate  one view of calc_importance (trainer/ht3dgs_trainer.py:1428-1462): for loss = image.sum()
 *     (dL_dout_color = that loss's gradient with respect to the rendered colour, [3,H,W]; depth and alpha receive none)
 *     importance[i][k][c] += |dL/dsh[i][k][c]|, importance float32 [P, M, 3] kept by the caller across the views of a
 *     segment (zero it first; divide by the pixel count afterwards as the reference does).  Runs the blend backward of
 *     the LAST forward (same geometry / binning / image buffers and alphas as htgs_backward) and forms the outer product
 *     of the SH basis with the clamp-masked colour gradient directly: no geometry backward, no gradient tensors, none of
 *     the reference's three abs() hooks.  scratch from htgs_backward_scratch_bytes(P).
 * htgs_importance_select      the bottom-k choice of merge_two_3DGS (:232-238, :246-251): mask[i] = 1 for the k Gaussians
 *     with the smallest max_j importance[i][j] (n_per values per Gaussian), 0 elsewhere.  Ties are broken towards
 *     the lower index (stable radix sort of (amax bits, index) with this library's own sort; torch.topk leaves the
 *     choice among equal values unspecified).  workspace from htgs_importance_select_workspace_bytes(P), 256-B aligned.
 * htgs_transform_points       out[i] = (T [x_i 1])[:3] / (T [x_i 1])[3] for a row-major 4x4 T in DEVICE memory: the
 *     source model's positions carried into the destination frame (:254-257). */
int htgs_importance_accumulate(int P, int D, int M, int64_t R, int W, int H, const float* background,
                               const float* means3D, const float* campos, const float* alphas, const int* radii,
                               const void* geom_buf, const void* binning_buf, const void* image_buf,
                               const float* dL_dout_color, void* scratch, float* importance, void* stream);
size_t htgs_importance_select_workspace_bytes(int P);
int htgs_importance_select(int P, int n_per, const float* importance, int64_t k, void* workspace, uint8_t* mask,
                           void* stream);
int htgs_transform_points(int P, const float* xyz, const float* transform, float* out, void* stream);

/* ---- un-activated, un-concatenated parameters (opt-in) ----
 * The HT model stores _scaling, _rotation, _opacity, _features_dc [P,1,3] and _features_rest [P,15,3]
 * and every render call first materialises exp / F.normalize / sigmoid / torch.cat of them
 * (scene/gaussian_model_ht.py:127-188, call site :834-866); autograd then walks back through the same
 * four ops.  These two entry points take the stored parameters as they are, apply the activations in
 * registers inside K1 / K7 and return gradients with respect to the STORED parameters, so neither the
 * activated tensors nor the 192 B/Gaussian concatenated SH tensor (nor the split of its gradient) ever
 * exist.  M is fixed at 16 (degree-3 storage); D <= 3 is the active degree.  features_rest and
 * dL_dfeatures_rest must be 16-byte aligned.  Pair with htgs_forward_render / _capacity unchanged.
 *   forward : identical outputs to htgs_forward_preprocess_async fed with the activated tensors.
 *   backward: dL_dscaling_raw = dL_dscales * exp(_scaling); dL_drotation_raw = normalize-backward of
 *             dL_drotations; dL_dopacity_raw = dL_dopacity * s(1-s); dL_dfeatures_dc / _rest = the two
 *             slices of dL_dsh. */
int htgs_forward_preprocess_raw_async(int P, int D, const float* xyz, const float* scaling_raw, float scale_modifier,
                                      const float* rotation_raw, const float* opacity_raw,
                                      const float* features_dc, const float* features_rest,
                                      const float* viewmatrix, const float* projmatrix, const float* campos,
                                      int W, int H, float tanfovx, float tanfovy, int prefiltered,
                                      void* geom_buf, int* radii, uint32_t* num_rendered_pinned_host, void* stream);
int htgs_backward_raw(int P, int D, int64_t R, int W, int H, const float* background, const float* xyz,
                      const float* features_dc, const float* features_rest, const float* alphas,
                      const float* scaling_raw, float scale_modifier, const float* rotation_raw,
                      const float* opacity_raw, const float* viewmatrix, const float* projmatrix,
                      const float* campos, float tanfovx, float tanfovy, const int* radii,
                      const void* geom_buf, const void* binning_buf, const void* image_buf,
                      const float* dL_dout_color, const float* dL_dout_depth, const float* dL_dout_alpha,
                      void* scratch, float* dL_dmeans2D, float* dL_dxyz, float* dL_dfeatures_dc,
                      float* dL_dfeatures_rest, float* dL_dopacity_raw, float* dL_dscaling_raw,
                      float* dL_drotation_raw, void* stream);

/* ---- image loss (opt-in) ----
 * loss = (1 - lambda) * mean|pred - gt| + lambda * (1 - SSIM(pred, gt)), the training loss of
 * trainer/losses.py:98-136 (SSIM: 11x11 sigma-1.5 window, zero padding, per channel, :147-200), for
 * pred/gt float32 [C,H,W].  forward writes out3 = {loss, l1, 1-ssim} (device) and, when dmaps != NULL,
 * the three derivative maps [3][C][H][W] the backward consumes; backward writes dL_dpred[C,H,W] =
 * grad_loss * dloss/dpred (grad_loss: device scalar, NULL = 1).  scratch from htgs_l1_ssim_scratch_bytes.
 * The first call on each device uploads the 11 window taps to constant memory with a blocking copy (under a
 * lock): make it outside any stream capture. */
size_t htgs_l1_ssim_scratch_bytes(int C, int H, int W);
int htgs_l1_ssim_forward(int C, int H, int W, const float* pred, const float* gt, float lambda_dssim,
                         float* dmaps, void* scratch, float* out3, void* stream);
int htgs_l1_ssim_backward(int C, int H, int W, const float* pred, const float* gt, const float* dmaps,
                          float lambda_dssim, const float* grad_loss, float* dL_dpred, void* stream);

/* ---- simple-knn ----
 * Replaces distCUDA2 / SimpleKNN::knn (simple-knn/spatial.cu:15-26,
 * simple_knn.cu:186-222): out[i] = mean of the squared distances from point i
 * to its 3 nearest other points.  No allocation, no host sync. */
size_t htgs_dist2_workspace_bytes(int P);
int htgs_dist2(int P, const float* points, float* out, void* workspace, size_t workspace_bytes, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* HTGS_H_INCLUDED */
