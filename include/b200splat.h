/*
 * b200splat.h — C ABI of the B200-native (sm_100a) Gaussian-splatting rasterizer.
 *
 * This is the drop-in boundary for ONE hot path of ShayManor/3d-gaussian-splatting: the call
 *     gsplat.rasterization(means, quats, scales, opacities, colors=shs, viewmats[1,4,4], Ks[1,3,3],
 *                          width, height, packed=False, render_mode, sh_degree, backgrounds[1,3])
 * issued at reference  src/backends/gsplat_backend.py:83-97 and :168-182, its autograd backward
 * (triggered at src/training/trainer.py:505-511), and the densification statistic of
 * src/training/trainer.py:786-791.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless its name ends in `_host`;
 *   - no allocation, no exceptions, no torch types; caller owns every buffer;
 *   - all work is enqueued on `stream` (a cudaStream_t passed as void*); only functions whose
 *     documentation says "synchronises" block the host;
 *   - return 0 on success, B200_EARG (-1) bad argument, B200_ESCRATCH (-2) scratch too small,
 *     B200_ECUDA (-3) CUDA error (text in b200_last_error()).
 *   - quaternions (w,x,y,z); viewmat row-major world->camera, OpenCV axes; images row-major HWC;
 *     tile = 16x16 px; tile_id = ty*tile_w + tx; one camera per call (reference passes batch dim 1,
 *     gsplat_backend.py:174-175).
 *
 * Splat record ("splats", [N,12] fp32, 48 B, 16-B aligned) written by the projection stage and
 * gathered by the blend stages:
 *     [0] mean2d.x  [1] mean2d.y  [2] conic.A  [3] conic.B
 *     [4] conic.C   [5] opacity   [6] red      [7] green
 *     [8] blue      [9] depth     [10] 0       [11] 0
 * The gradient record ("v_splats", same shape) holds the cotangent of each slot.
 */
#ifndef B200SPLAT_H
#define B200SPLAT_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200_OK 0
#define B200_EARG (-1)
#define B200_ESCRATCH (-2)
#define B200_ECUDA (-3)

#define B200_TILE 16
#define B200_SPLAT_FLOATS 12

/* channel sets of the blend stages (reference render_mode parsing, gsplat_backend.py:190-201) */
#define B200_CH_RGB 0   /* "RGB"            -> 3 channels           */
#define B200_CH_D 1     /* "D" / "ED"       -> 1 channel (depth)    */
#define B200_CH_RGBD 2  /* "RGB+D"/"RGB+ED" -> 4 channels (rgb, d)  */

/* Host camera block: 16 viewmat floats (row-major 4x4 world->camera),
 * then fx, fy, cx, cy (pixels; K[0,0],K[1,1],K[0,2],K[1,2] of gsplat_backend.py:10,175),
 * then near_plane, far_plane, eps2d, radius_clip (gsplat defaults 0.01, 1e10, 0.3, 0.0). */
#define B200_CAM_FLOATS 24
/* Packed DEVICE camera block (opaque, 16-B aligned) written by b200_cam_pack. */
#define B200_CAM_DEV_FLOATS 32

int b200_version(void);
const char* b200_last_error(void);

/* ---- stage 0: camera block built ON THE DEVICE ------------------------------------------------
 * The reference hands the backend a device-resident pose (viewpoint['world_view_transform'],
 * trainer.py:459-466) and device intrinsics (self.K, gsplat_backend.py:10,175); gsplat consumes both on
 * the device.  b200_cam_pack does the same: viewmat[16] and K[9] are DEVICE pointers, no host sync.
 * cam_dev[B200_CAM_DEV_FLOATS] is then passed to the projection stages (forward and backward). */
int b200_cam_pack(const float* viewmat, const float* K, float near_plane, float far_plane, float eps2d,
                  float radius_clip, int width, int height, float* cam_dev, void* stream);

/* ---- stage 1: EWA projection + SH colour + per-Gaussian tile count (fused) -------------------
 * Replaces gsplat fully_fused_projection + spherical_harmonics + clamp_min(c+0.5,0) + the counting
 * half of isect_tiles, as invoked through gsplat_backend.py:168-182.
 *   means[N,3] quats[N,4] scales[N,3] (post-exp) opacities[N] (post-sigmoid) shs[N,sh_coeffs,3]
 *   sh_degree in 0..3 with (sh_degree+1)^2 <= sh_coeffs; sh_degree < 0 => `shs` is [N,3] plain colours.
 * Outputs: radii[N,2] int32, splats[N,12], tiles_per_gauss[N] int32,
 *          block_sums[b200_project_blocks(N)] int64 (sum of tiles_per_gauss per 256-Gaussian block; may be NULL
 *          when depth_keys is given — the split-key path recounts in depth order, b200_bin_count_sorted);
 *          optional (both or neither NULL) depth_keys[N] uint32 = fp32 depth bits (0xFFFFFFFF if culled)
 *          and gauss_ids[N] = 0..N-1, the input of the split-key sort below.
 *          optional tile_grid[(tile_h+1)*(tile_w+1)] int32 (zeroed by this call): +1/-1 at the corners of
 *          every visible Gaussian's tile rectangle, the input of b200_tile_prep.
 * Camera: exactly one of cam_host (B200_CAM_FLOATS host floats) and cam_dev (b200_cam_pack output) is
 * non-NULL; width/height must equal the ones the block was packed with. */
int64_t b200_project_blocks(int64_t N);
int b200_project_sh_fwd(const float* means, const float* quats, const float* scales,
                        const float* opacities, const float* shs, int sh_coeffs, int sh_degree,
                        int64_t N, const float* cam_host, const float* cam_dev, int width, int height,
                        int32_t* radii, float* splats, int32_t* tiles_per_gauss,
                        int64_t* block_sums, uint32_t* depth_keys, int32_t* gauss_ids,
                        int32_t* tile_grid, uint8_t* visibility, float* grad_records, void* tile_rects,
                        void* stream);
/* Three optional auxiliary outputs (NULL = skip):
 *   visibility[N,2]  uint8 = radii > 0, the `visibility_filter` of gsplat_backend.py:206-208 (no compare launch);
 *   grad_records[N,12] fp32 (16-B aligned) = zeros: the cotangent records b200_raster_bwd accumulates into (no fill launch);
 *   tile_rects[N] 2 x uint32 = (minx | maxx << 16, miny | maxy << 16), the Gaussian's tile rectangle of SURVEY.md A.3,
 *     (0, 0) when it touches no tile — what b200_bin_coarse gathers behind the depth permutation (8 B instead of
 *     radii + means2d). */

/* ---- stage 2a: exclusive scan of block sums; writes block_offsets[nblocks] and total[1] -------
 * (the cumsum half of gsplat isect_tiles).  Single CTA; nblocks = b200_project_blocks(N). */
int b200_bin_scan(const int64_t* block_sums, int64_t nblocks, int64_t* block_offsets,
                  int64_t* total, void* stream);

/* ---- stage 2b: key emission ---------------------------------------------------------------
 * isect_ids[M] int64 = (tile_id << 32) | depth_bits, flatten_ids[M] int32 = Gaussian index, in
 * gsplat's emission order (Gaussian-major, then tile row-major).  cum_tiles (optional, may be NULL)
 * receives the inclusive int64 cumsum [N].  M is the value b200_bin_scan wrote to total[0]. */
int b200_bin_emit(const float* splats, const int32_t* radii, const int32_t* tiles_per_gauss,
                  const int64_t* block_offsets, int64_t N, int width, int height, int64_t M,
                  int64_t* isect_ids, int32_t* flatten_ids, int64_t* cum_tiles, void* stream);

/* ---- stage 2c: stable LSD radix sort of (uint64 key, uint32 value) pairs over key bits
 * [0, end_bit) (replaces cub::DeviceRadixSort::SortPairs inside gsplat isect_tiles(sort=True)).
 * keys_a/vals_a hold the input; *result_in_b tells which buffer pair holds the sorted output.
 * Scratch: b200_sort_scratch_bytes(M). */
size_t b200_sort_scratch_bytes(int64_t M);
int b200_sort_pairs(uint64_t* keys_a, uint32_t* vals_a, uint64_t* keys_b, uint32_t* vals_b,
                    int64_t M, int end_bit, void* scratch, size_t scratch_bytes,
                    int* result_in_b_host, void* stream);

int b200_sort_pairs_k32(uint32_t* keys_a, uint32_t* vals_a, uint32_t* keys_b, uint32_t* vals_b,
                        int64_t M, int end_bit, void* scratch, size_t scratch_bytes,
                        int* result_in_b_host, void* stream);
int b200_sort_pairs_k16(uint16_t* keys_a, uint32_t* vals_a, uint16_t* keys_b, uint32_t* vals_b,
                        int64_t M, int end_bit, void* scratch, size_t scratch_bytes,
                        int* result_in_b_host, void* stream);

/* Allocation-free variant of b200_sort_pairs_k32.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, serving loops, harnesses
 * that forbid it).  Same contract and same result as b200_sort_pairs_k32, but the scratch lives in a
 * file-scope device array inside the library (2 105 600 B, held by every GPU that loads the library until
 * exit); nothing is allocated or freed by the call.  M > B200_SORT_STATIC_MAX_KEYS is rejected (B200_EARG).
 * Every call shares the array; each call clears what it uses on `stream` first, so calls ordered on the
 * GPU are independent, but calls that OVERLAP on the GPU (different streams without ordering) race. */
#define B200_SORT_STATIC_MAX_KEYS 1048576
int b200_sort_pairs_k32_static(uint32_t* keys_a, uint32_t* vals_a, uint32_t* keys_b, uint32_t* vals_b,
                               int64_t M, int end_bit, int* result_in_b_host, void* stream);

/* ---- stage 2b'/2c': split-key form of the same sort (the fast path) ----------------------------
 * An LSD radix sort of (tile<<32 | depth) keys visits the depth digits first and the tile digits
 * last.  Depth digits do not depend on the tile, so they are sorted over the N Gaussians BEFORE the
 * duplication (b200_sort_pairs_k32 on depth_keys/gauss_ids -> perm), the duplicates are emitted in
 * that order with 16-bit tile keys, and only the tile digits are sorted over the M intersections
 * (b200_sort_pairs_k16).  Output order is identical to sorting the 64-bit keys (ties: ascending
 * Gaussian index); b200_isect_ids rebuilds the sorted 64-bit keys when a caller wants them.
 * Requires tile_w*tile_h <= 65536. */
int b200_bin_count_sorted(const int32_t* tiles_per_gauss, const int32_t* perm, int64_t N,
                          int64_t* block_sums, void* stream);
int b200_bin_emit_sorted(const float* splats, const int32_t* radii, const int32_t* tiles_per_gauss,
                         const int32_t* perm, const int64_t* block_offsets, int64_t N, int width,
                         int height, int64_t M, uint16_t* tile_keys, int32_t* flatten_ids, void* stream);
/* ---- two-level binning (default fast path) ------------------------------------------------------------------------
 * Same result as gsplat isect_tiles(sort=True) + isect_offset_encode (gsplat_backend.py:168-182) — flatten_ids and
 * isect_offsets bit for bit — without sorting the M intersections.  Inputs: projection outputs, `perm` = Gaussian ids
 * in depth order (b200_sort_pairs_k32 of the depth keys), isect_offsets and M from b200_tile_prep.
 *   1. (Gaussian, supertile) pairs, supertile = 4x4 tiles, are emitted in depth order as
 *      (supertile << 16 | 16-bit mask of the covered tiles, id); output offsets by a decoupled look-back scan inside the
 *      kernel, per-supertile counts in CTA-private histograms (M' pairs, M' ~ M/3..M/5, M' <= M);
 *   2. one small kernel turns the counts into supertile ranges and radix digit bases;
 *   3. the pairs are sorted stably on the supertile bits (one or two 8-bit onesweep passes over M', count on the device);
 *   4. one CTA per supertile, one warp per tile, streams the supertile's list once and compacts the ids covering its
 *      tile into flatten_ids[isect_offsets[tile] ...] in list order (= depth order, ties by ascending Gaussian index).
 * tile_rects: the packed per-Gaussian rectangles written by b200_project_sh_fwd, or NULL (then they are recomputed from
 * radii / means2d).  keys_a/vals_a/keys_b/vals_b: M uint32 each (only M' used); scratch: b200_bin_coarse_scratch_bytes(N, M).
 * Needs b200_bin_coarse_supported(width, height): <= 65 536 tiles and <= 4096 supertiles. */
int b200_bin_coarse_supported(int width, int height);
size_t b200_bin_coarse_scratch_bytes(int64_t N, int64_t M_cap);
int b200_bin_coarse(const float* splats, const int32_t* radii, const void* tile_rects, const int32_t* perm,
                    int64_t N, int width, int height, const int32_t* isect_offsets, int64_t M,
                    uint32_t* keys_a, uint32_t* vals_a, uint32_t* keys_b, uint32_t* vals_b,
                    int32_t* flatten_ids, void* scratch, size_t scratch_bytes, void* stream);
/* gsplat's sorted 64-bit isect_ids rebuilt from (flatten_ids, isect_offsets) for callers that ask (parity tests, meta). */
int b200_isect_ids_from_offsets(const int32_t* flatten_ids, const int32_t* isect_offsets, int n_tiles,
                                const float* splats, int64_t M, int64_t* isect_ids, void* stream);

/* Per-tile ranges and radix digit bases WITHOUT a pass over the intersections: the 2-D prefix sum of
 * tile_grid (written by b200_project_sh_fwd) is the intersection count of every tile, so
 *   offsets[n_tiles]  = row-major exclusive scan          (== b200_tile_ranges_k16 of the sorted keys),
 *   total[0]          = M,
 *   digit_base[2*256] = exclusive digit offsets of the two 8-bit passes of the tile-key sort
 * come from one single-CTA kernel; b200_sort_pairs_k16_prepared then skips its histogram pass.
 * Needs (tile_w+1)*(tile_h+1) <= b200_tile_prep_max_cells() (the grid is staged in shared memory). */
/* The corner-difference grid built OUTSIDE the projection kernel (call b200_project_sh_fwd with tile_grid == NULL): a few
 * persistent CTAs recompute every visible Gaussian's tile rectangle from the stored means2d / radii (same fp32
 * operations), accumulate the +1/-1 corners in a shared-memory copy of the grid and flush it once.  tile_grid is zeroed
 * by the call.  Runs on a side stream under the depth sort (rasterization.py). */
int b200_tile_count(const float* splats, const int32_t* radii, int64_t N, int width, int height, int32_t* tile_grid,
                    void* stream);
size_t b200_tile_prep_max_cells(void);
int b200_tile_prep(const int32_t* tile_grid, int width, int height, int32_t* offsets, int64_t* total,
                   uint32_t* digit_base, void* stream);
int b200_sort_pairs_k16_prepared(uint16_t* keys_a, uint32_t* vals_a, uint16_t* keys_b, uint32_t* vals_b,
                                 int64_t M, int end_bit, const uint32_t* digit_base, void* scratch,
                                 size_t scratch_bytes, int* result_in_b_host, void* stream);
int b200_tile_ranges_k16(const uint16_t* sorted_tile_keys, int64_t M, int n_tiles, int32_t* offsets,
                         void* stream);
int b200_isect_ids(const uint16_t* sorted_tile_keys, const int32_t* flatten_ids, const float* splats,
                   int64_t M, int64_t* isect_ids, void* stream);

/* ---- stage 2d: per-tile ranges (gsplat isect_offset_encode) ---------------------------------
 * offsets[n_tiles] int32 = first sorted index whose tile id >= t; all M's for trailing empty tiles. */
int b200_tile_ranges(const int64_t* sorted_isect_ids, int64_t M, int n_tiles, int32_t* offsets,
                     void* stream);

/* ---- stage 3: forward blend (gsplat rasterize_to_pixels forward) ----------------------------
 * render[H,W,CH] alpha[H,W] last_ids[H,W] (index into the sorted list of the last composited
 * Gaussian, -1 if none).  background: DEVICE pointer to n_background <= CH floats (the reference's
 * bg_color tensor, trainer.py:471-478); channels beyond n_background composite over 0. */
int b200_raster_fwd(const float* splats, const int32_t* flatten_ids, const int32_t* offsets,
                    int64_t M, int width, int height, int channels, const float* background,
                    int n_background, float* render, float* alpha, int32_t* last_ids, void* stream);

/* ---- stage 4a: backward blend (gsplat rasterize_to_pixels backward) -------------------------
 * v_render[H,W,CH], v_alpha[H,W] or NULL.  ACCUMULATES into v_splats[N,12] (caller zeroes). */
int b200_raster_bwd(const float* splats, const int32_t* flatten_ids, const int32_t* offsets,
                    int64_t M, int width, int height, int channels, const float* background,
                    int n_background, const float* alpha, const int32_t* last_ids, const float* v_render,
                    const float* v_alpha, float* v_splats, void* stream);

/* ---- stage 4b: projection + SH backward ----------------------------------------------------
 * Chains v_splats through conic/mean2d/depth/colour to the five input tensors.  Writes (does not
 * accumulate) v_means[N,3] v_quats[N,4] v_scales[N,3] v_opacities[N] v_shs[N,sh_coeffs,3]. */
int b200_project_sh_bwd(const float* means, const float* quats, const float* scales,
                        const float* opacities, const float* shs, int sh_coeffs, int sh_degree,
                        int64_t N, const float* cam_host, const float* cam_dev, int width, int height,
                        const int32_t* radii, const float* v_splats,
                        float* v_means, float* v_quats, float* v_scales, float* v_opacities,
                        float* v_shs, void* stream);

/* ---- stages 1 / 4b on the RAW parameter store ----------------------------------
 * Same as b200_project_sh_fwd / _bwd, but the inputs are the reference's raw parameters
 * (src/gaussians/gaussian_model.py:65-76): xyz[N,3], rotation[N,4] (un-normalised), scaling[N,3] (log),
 * opacity[N] (logit), features[N,sh_coeffs,3] (features_dc ++ features_rest, stored contiguously by the
 * parameter store).  The getters' activations (gaussian_model.py:90-108: exp, sigmoid, F.normalize, cat)
 * are applied in registers, so no activated copy of the store is materialised, and the backward stage
 * returns gradients w.r.t. the RAW parameters.  accumulate != 0: the gradients are ADDED into the
 * caller's buffers (the flat gradient store shared by the views of a step and by the NCCL all-reduce);
 * culled Gaussians and SH coefficients above sh_degree are then left untouched. */
int b200_project_sh_fwd_raw(const float* xyz, const float* rotation, const float* scaling,
                            const float* opacity, const float* features, int sh_coeffs, int sh_degree,
                            int64_t N, const float* cam_host, const float* cam_dev, int width, int height,
                            int32_t* radii, float* splats, int32_t* tiles_per_gauss,
                            int64_t* block_sums, uint32_t* depth_keys, int32_t* gauss_ids,
                            int32_t* tile_grid, uint8_t* visibility, float* grad_records, void* tile_rects,
                            void* stream);
int b200_project_sh_bwd_raw(const float* xyz, const float* rotation, const float* scaling,
                            const float* opacity, const float* features, int sh_coeffs, int sh_degree,
                            int64_t N, const float* cam_host, const float* cam_dev, int width, int height,
                            const int32_t* radii, const float* v_splats,
                            float* v_xyz, float* v_rotation, float* v_scaling, float* v_opacity,
                            float* v_features, int accumulate, void* stream);

/* ---- stage 5: densification statistics (reference trainer.py:786-791) -------------------------
 * accum[N,3] += ||xyz_grad[n,:]||_2 (broadcast over the 3 columns); count[N] += 1. */
int b200_densify_stats(const float* xyz_grad, int64_t N, float* accum, float* count, void* stream);

/* ---- f3: the reference's per-view photometric loss, forward + gradient fused ----
 * trainer.py:485-492: x = clamp(render,0,1) (straight-through), loss = (1-lambda) mean|x-gt| +
 * lambda (1 - SSIM_3x3box(x, gt)) (_compute_ssim, trainer.py:354-371).  render/gt are [H,W,3].
 * sums[8]: [0..2] receive sum|x-gt|, the sum of window SSIMs and sum (x-gt)^2 over the n = 3*H*W elements,
 * [3] is an arrival counter (both cleared by the call), [4..7] = (loss, l1, ssim, mse) finished by the last
 * CTA, so that the step needs no scalar kernels; v_render (optional) receives grad_scale * d loss / d render. */
int b200_l1_ssim_fwd_bwd(const float* render, const float* gt, int width, int height,
                         float lambda_dssim, float grad_scale, float* v_render, float* sums, void* stream);

/* ---- f4: the optimiser step over the flat parameter store -----------------------
 * torch.optim.Adam semantics (trainer.py:331-340: six groups, eps 1e-15; no amsgrad / weight decay) in one
 * pass over params / grads / exp_avg / exp_avg_sq of identical flat layout (n_floats, multiple of 4, 16-B
 * aligned).  seg_host: n_seg (<= 8) rows of B200_ADAM_SEG_DOUBLES doubles {begin, end, lr, lr_alt, period, split,
 * clamp_max}: elements [begin,end) use lr, except where period > 0 and (i - begin) % period >= split, which use
 * lr_alt (the features_dc / features_rest rates inside the [N,16,3] SH block: period 48, split 3); the updated
 * parameter is min(p, clamp_max) — +inf for none — which is the reference's per-iteration cap on the log-scales
 * (scaling.data.clamp_(max=scale_ceiling), trainer.py:854-859) folded into the same pass.  step = Adam's t >= 1. */
#define B200_ADAM_SEG_DOUBLES 7
int b200_adam_step(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, int64_t n_floats,
                   const double* seg_host, int n_seg, int step, float beta1, float beta2, float eps,
                   void* stream);
/* Same pass with stage 5 riding along (trainer.py:786-791 takes the statistic right after optimizer.step(), from the
 * same, untouched xyz.grad): accum[n_gauss,3] += ||xyz_grad||_2, count[n_gauss] += 1 for every Gaussian, by the first
 * n_gauss threads of the launch — replaces a separate b200_densify_stats launch.  A segment's `begin` may be negative
 * (a rank stepping only a slice of the flat buffer passes offset pointers and shifted segments). */
int b200_adam_step_stats(float* params, const float* grads, float* exp_avg, float* exp_avg_sq, int64_t n_floats,
                         const double* seg_host, int n_seg, int step, float beta1, float beta2, float eps,
                         const float* xyz_grad, int64_t n_gauss, float* accum, float* count, void* stream);

/* ---- f2: population change of the flat store in one gather -----------------------
 * Data movement of gaussian_model.py:110-366 (clone / split / prune with the Adam moments spliced,
 * _replace_param_in_optimizer :9-50): destination row j of the new store is source row src_row[j]; its
 * moments are copied when keep_state[j] != 0 (or keep_state == NULL) and zeroed otherwise.  layout_*_host:
 * five int64 block offsets {xyz, scaling, rotation, opacity, features} of the flat buffers, in floats.
 * The three moment pointers are all NULL or all set. */
int b200_store_gather(const float* params_src, const float* exp_avg_src, const float* exp_avg_sq_src,
                      const int64_t* layout_src_host, float* params_dst, float* exp_avg_dst,
                      float* exp_avg_sq_dst, const int64_t* layout_dst_host, const int64_t* src_row,
                      const uint8_t* keep_state, int64_t n_dst, void* stream);

/* ---- multi-GPU: the gradient all-reduce as this library's kernel over NVSwitch multicast memory -----------------------
 * The reference has no distributed code; BASELINE.json north_star asks for the parameter gradients to be
 * reduced over NVLink.  mc_ptr is the MULTICAST address of a flat fp32 buffer replicated in symmetric memory on all
 * `world` ranks (host plumbing: torch.distributed._symmetric_memory).  Rank `rank` reads its 1/world slice once with
 * multimem.ld_reduce.add.v4.f32 (the switch sums the replicas), scales it and writes it to every replica with
 * multimem.st.v4.f32.  In place; n_floats % 4 == 0; the caller brackets the call with cross-rank barriers on the same
 * stream.  max_ctas <= 0: 148. */
int b200_allreduce_multimem(float* mc_ptr, int64_t n_floats, int rank, int world, float scale, int max_ctas, void* stream);
/* The slice [begin, end) (floats, multiples of 4) of an n_floats buffer that rank `rank` of `world` reduces / updates. */
void b200_multimem_slice(int64_t n_floats, int rank, int world, int64_t* begin_float, int64_t* end_float);
/* Reduce-scatter of the gradient + Adam on the rank's slice + all-gather of the PARAMETERS in one kernel (the optimiser of
 * trainer.py:331-340,505-511 for the view-sharded step): for every 16-byte vector of this rank's slice the switch-summed
 * gradient (multimem.ld_reduce over mc_grads, times grad_scale) drives b200_adam_step's arithmetic on the LOCAL parameter
 * and moments (`params` = this rank's replica, exp_avg / exp_avg_sq: only the slice is touched — sharded moments), and the
 * updated parameter is written to every replica (multimem.st over mc_params).  Inside [keep_grad_begin, keep_grad_end)
 * (the position block) the reduced gradient is also written to every replica of the gradient buffer, so that the
 * densification statistic (trainer.py:786-791) can be taken locally afterwards.  seg_host as b200_adam_step.  The caller
 * brackets the call with cross-rank barriers. */
int b200_allreduce_adam_multimem(float* mc_grads, float* mc_params, const float* params, float* exp_avg,
                                 float* exp_avg_sq, int64_t n_floats, int rank, int world, float grad_scale,
                                 const double* seg_host, int n_seg, int step, float beta1, float beta2, float eps,
                                 int64_t keep_grad_begin, int64_t keep_grad_end, int max_ctas, void* stream);
/* Every replica of this rank's slice of the symmetric buffer <- the rank's local copy. */
int b200_share_slice_multimem(float* mc_ptr, const float* local, int64_t n_floats, int rank, int world, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200SPLAT_H */
