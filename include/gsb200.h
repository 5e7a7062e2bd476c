/*
 * gsb200.h — C-ABI of libgsb200.so, the B200 (sm_100a) Gaussian-splatting training hot path.
 *
 * The reference (uc-vision/splat-trainer) is pure Python and has no FFI of its own: every
 * kernel on its hot path lives in the un-vendored dependency `taichi-splatting >= 0.31.0`
 * (reference pyproject.toml:14).  The entry points below are therefore what a binding for
 * that dependency's *call sites* needs; each one cites the reference call site it serves.
 * SURVEY.md §8(b) is the contract this header implements.
 *
 * Conventions
 *   - every pointer is a raw DEVICE pointer unless the name starts with `h_`;
 *   - `stream` is a cudaStream_t passed as void*; all work is enqueued on it, nothing
 *     synchronises the host unless the comment says so;
 *   - the library allocates nothing persistent and keeps no global state: the caller
 *     (PyTorch's caching allocator) owns every buffer incl. workspaces (size query first);
 *   - every function returns 0 on success or a negative gs_status; the text of the last
 *     error on the calling thread is returned by gs_last_error();
 *   - all floating point is fp32; quaternions are xyzw (reference scene/io.py:45,104);
 *   - images are H×W×C row-major, pixel centres at (+0.5,+0.5);
 *   - "rows": the tile mapper / rasteriser work on M rows of projected splats.  In the
 *     fused scene path M = N (row i = Gaussian i, culled rows carry opacity 0); in the
 *     operator path (project_to_image → render_projected) M = V compacted rows.
 */
#ifndef GSB200_H
#define GSB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GSB200_VERSION 100          /* 0.1.0 */
#define GS_TILE 16                  /* tile edge in pixels (only supported value) */
#define GS_G2D_STRIDE 6             /* floats per projected row: mean.x mean.y conic.a conic.b conic.c opacity */

typedef enum gs_status {
  GS_OK = 0,
  GS_ERR_INVALID = -1,              /* bad argument                                   */
  GS_ERR_CUDA = -2,                 /* a CUDA call / launch failed (see gs_last_error) */
  GS_ERR_WORKSPACE = -3,            /* workspace too small                            */
  GS_ERR_UNSUPPORTED = -4           /* e.g. channel count not instantiated            */
} gs_status;

/* CameraParams of the reference (trainer/trainer.py:291-301, camera_table/camera_table.py:167-175). */
typedef struct gs_camera {
  float T_camera_world[16];         /* row-major 4×4, x_cam = R x_world + t            */
  float fx, fy, cx, cy;             /* CameraParams.projection                        */
  int32_t width, height;            /* CameraParams.image_size = (W, H)               */
  float near_plane, far_plane;      /* config/trainer/default.yaml:83-84              */
} gs_camera;

/* RasterConfig: the four names the reference passes (trainer/trainer.py:308-310,
 * scene/mlp_scene.py:373) plus the constants SURVEY Appendix A lists. */
typedef struct gs_raster_config {
  int32_t tile_size;                /* 16                                             */
  float alpha_threshold;            /* 1/255: contributions below are skipped         */
  float clamp_max_alpha;            /* 0.99                                           */
  float saturate_threshold;         /* 0.9999: stop a pixel once T < 1 - this         */
  float blur_cov;                   /* 0.3 px² added to the 2-D covariance diagonal   */
  int32_t antialias;                /* opacity *= sqrt(det Σ / det(Σ+blur))           */
  int32_t compute_visibility;       /* forward accumulates Σ α·T per row              */
  int32_t compute_point_heuristic;  /* backward writes prune_cost / split_score       */
  int32_t render_median_depth;      /* forward writes depth of first splat with T<½   */
  float frustum_clamp;              /* 1.3: |x/z| clamp (× half-FOV tangent) in the Jacobian */
} gs_raster_config;

/* Regularisation gradients fused into gs_project_bwd (section 7 has the standalone pass they equal). */
typedef struct gs_reg_fuse {
  float w_scale, w_opacity, w_aspect; /* weights of the scale / opacity / aspect terms                     */
  const float *sums;                  /* device float[4] written by gs_reg_loss(compute_sums=1) this image   */
  const float *depth;                 /* device (n) camera depth plane of gs_project_fwd                     */
} gs_reg_fuse;

void gs_default_raster_config(gs_raster_config *cfg);
int gs_version(void);
const char *gs_last_error(void);
/* Diagnostics: number of kernels this library has launched in this process (what bench.py reports as gpu_launches). */
unsigned long long gs_launch_count(void);

/* ---------------------------------------------------------------------------------------
 * (1) 3D→2D projection + SH colour.   Serves taichi_splatting.project_to_image
 *     (reference scene/mlp_scene.py:415, :375) and evaluate_sh_at (scene/transfer_sh.py:49).
 *     Row i of the outputs belongs to Gaussian i.  Culled Gaussians (outside near/far,
 *     degenerate covariance, opacity <= alpha_threshold, or no tile touched) get
 *     g2d = {0,0,1,0,1,0}, depth = +inf, rgb = 0.
 *     sh: (n,3,K) with K=(deg+1)^2, basis order/sign = rsh_cart_3 (scene/mlp/rsh.py:110-157),
 *     colour = 0.5 + Σ_k Y_k(dir) sh[c][k] (util/misc.py:41-49). sh may be NULL (rgb untouched).
 *     acc (nullable): the raster-backward accumulator (n × acc_stride); rows of in-view
 *     Gaussians are zeroed here so no separate memset is needed.
 *     tile_info (nullable, gs_tile_info_bytes(n, W, H) bytes, 256-B aligned): what the cull learns about each row's
 *     tile footprint (16-B record with the exact hit mask, tile count, per-tile totals); gs_tile_map(tile_info_in =
 *     tile_info) then skips its own footprint pass.  Opaque to the caller.
 *     records (nullable, (n, GS_REC_FLOATS), needs sh): the rasteriser's staged splat record of every in-view row
 *     (gs_raster_aux.records), written here so no later pass re-reads g2d / rgb to build it.
 * ------------------------------------------------------------------------------------- */
int gs_project_fwd(const gs_camera *h_cam, const gs_raster_config *h_cfg, int64_t n,
                   const float *position, const float *log_scaling, const float *rotation,
                   const float *alpha_logit, const float *sh, int sh_degree,
                   float *g2d, float *depth, float *rgb,
                   float *acc, int acc_stride, void *tile_info, float *records, void *stream);
size_t gs_tile_info_bytes(int64_t n, int width, int height);

/* Backward of gs_project_fwd.  d_g2d (row stride d_g2d_stride floats, 6 used),
 * d_rgb (stride d_rgb_stride, 3 used), d_depth (nullable, n) are dL/d(outputs).
 * row_vis (nullable, stride row_vis_stride): rows with row_vis <= 0 are skipped — a splat no pixel
 * blended has an exactly-zero image gradient, so the fused path neither reads its SH row nor
 * touches its gradient row (keeps the gradient buffer all-zero outside the visible set).
 * Gradients are ADDED into g_* when accumulate != 0 (batch = sum, reference
 * trainer/trainer.py:500-514), else written for in-view rows only (culled rows untouched).
 * reg (nullable, host struct): the gradients of the image's regularisation terms (gs_reg_loss's d_opacity, d_depth and
 * d_log_scaling, reference scene/mlp_scene.py:246-288) are computed per row here instead of by two extra sweeps;
 * needs row_vis (the weights) and reg->sums from gs_reg_loss(compute_sums=1). */
int gs_project_bwd(const gs_camera *h_cam, const gs_raster_config *h_cfg, int64_t n,
                   const float *position, const float *log_scaling, const float *rotation,
                   const float *alpha_logit, const float *sh, int sh_degree,
                   const float *g2d,
                   const float *d_g2d, int d_g2d_stride,
                   const float *d_rgb, int d_rgb_stride,
                   const float *d_depth, const float *row_vis, int row_vis_stride, const gs_reg_fuse *reg,
                   float *g_position, float *g_log_scaling, float *g_rotation,
                   float *g_alpha_logit, float *g_sh, int accumulate, void *stream);

/* evaluate_sh_at(sh (n,3,K), positions (n,3), indexes (v), cam_pos (3)) -> (v,3)
 * (reference scene/transfer_sh.py:49).  indexes are int64 (torch default); NULL = arange. */
int gs_sh_eval_fwd(int64_t v, const int64_t *indexes, const float *sh, int sh_degree,
                   const float *position, const float *h_cam_pos, float *rgb, void *stream);
int gs_sh_eval_bwd(int64_t v, const int64_t *indexes, const float *sh, int sh_degree,
                   const float *position, const float *h_cam_pos, const float *d_rgb,
                   float *g_sh, float *g_position, void *stream);

/* ---------------------------------------------------------------------------------------
 * (2) Tile-overlap mapping: (tile, depth) key sort + tile ranges.  Serves the mapper inside
 *     taichi_splatting.render_projected (reference scene/mlp_scene.py:418, :377).
 *     The result equals a stable sort of all (row, tile) overlaps by the 64-bit key
 *     (tile_id << 32 | float_bits(depth)), ties broken by row index.  It is computed without
 *     sorting any O-sized list: rows are sorted by depth (32-bit keys), binned per 8x8-tile
 *     super-tile (one short sort of ~O/6 pairs) and expanded in place into the per-tile lists,
 *     whose ranges come from per-tile totals.
 *       overlap_to_point[o]  row index of the o-th sorted overlap        (max_overlaps)
 *       tile_ranges[2t,2t+1] [begin,end) of tile t in overlap_to_point   (tiles_x*tiles_y*2)
 *       counters[0] = total overlaps O (may exceed max_overlaps: then counters[1] = 1, ranges are
 *                     clamped to max_overlaps and the lists are truncated — re-run with more room)
 *       counters[2] = min(O, max_overlaps), counters[3] = rows touching at least one tile
 *     tile_info_in (nullable): the tile info gs_project_fwd wrote for the same m rows / image size
 *     (must come from the same g2d/depth); NULL = computed here.
 * ------------------------------------------------------------------------------------- */
size_t gs_tile_map_workspace_bytes(int64_t m, int64_t max_overlaps, int width, int height);
int gs_tile_map(const gs_raster_config *h_cfg, int64_t m, const float *g2d, const float *depth,
                const void *tile_info_in, int width, int height, int64_t max_overlaps,
                void *workspace, size_t workspace_bytes,
                int32_t *overlap_to_point, int32_t *tile_ranges, int32_t *counters,
                void *stream);
/* Count only: tiles touched per row (m) and their total (counters[0]); used by the
 * drop-in path to size overlap buffers exactly (the reference reads the total back too). */
int gs_tile_count(const gs_raster_config *h_cfg, int64_t m, const float *g2d, const float *depth,
                  int width, int height, int32_t *counts, int32_t *counters, void *stream);

/* Stable LSD radix sort of (key,value) u32 pairs on bits [begin_bit,end_bit).  n may be given
 * on the host (n_dev == NULL) or as a device scalar bounded by n (grid sizing uses n).
 * keys_out/vals_out receive the result; keys_in/vals_in are clobbered (ping-pong). */
size_t gs_radix_sort_workspace_bytes(int64_t n);
int gs_radix_sort_pairs(uint32_t *keys_in, uint32_t *vals_in, uint32_t *keys_out, uint32_t *vals_out,
                        int64_t n, const int32_t *n_dev, int begin_bit, int end_bit,
                        void *workspace, size_t workspace_bytes, void *stream);

/* ---------------------------------------------------------------------------------------
 * (3) Forward rasterisation: front-to-back alpha compositing per 16×16 tile, splat records
 *     staged in shared memory, block-level early termination.  (render_projected,
 *     reference scene/mlp_scene.py:418; flags trainer/trainer.py:308-310,319.)
 *       image (H,W,C), final_T (H,W) transmittance left, n_contrib (H,W) 1+index (within
 *       the tile list) of the last blended splat, visibility (m, nullable) += Σ α·T,
 *       median_depth (H,W, nullable) depth of the splat that takes T below ½ (0 if none).
 *     channels ∈ {1,3,4}.
 * ------------------------------------------------------------------------------------- */
/* Optional fast-path inputs of the rasteriser (host struct, every member nullable; NULL aux = plain path).
 *   records    (m, GS_REC_FLOATS) staged splat records written by gs_raster_records (or by gs_project_fwd's `records`
 *              output): {mean.x, mean.y, A', B' | C', opacity, pcut, row id | r, g, b, y half-extent} with the conic
 *              pre-scaled for exp2 and pcut = log2(thr/opacity).  With records (channels == 3 only) a tile's batch of
 *              64 splats is gathered by asynchronous 16-byte copies (cp.async, completion on an mbarrier), double
 *              buffered: batch k+1 lands while batch k is composited.
 *   tile_order (T) tile ids, heaviest first (gs_tile_order) + ticket: the kernels then run as persistent CTAs that pull
 *              tiles from this queue (longest-processing-time-first), so one long tile cannot form the tail.
 *   tile_work  (T) forward: OUT, the tile's largest n_contrib (= the backward's trip count for the tile).
 *   counters   (8 x uint64, zeroed by the caller) forward only: [0] (warp, splat) candidate iterations, [1] iterations in
 *              which at least one pixel blended, [2] (pixel, splat) pairs blended, [3] splat records staged.
 *              Instrumented build of the kernel: slower, never used by the training path. */
#define GS_REC_FLOATS 12
typedef struct gs_raster_aux {
  const float *records;
  const int32_t *tile_order;
  uint32_t *ticket;
  int32_t *tile_work;
  unsigned long long *counters;
} gs_raster_aux;

int gs_raster_fwd(const gs_raster_config *h_cfg, int channels, int64_t m,
                  const float *g2d, const float *features, const float *depth,
                  const int32_t *overlap_to_point, const int32_t *tile_ranges,
                  int width, int height,
                  float *image, float *final_T, int32_t *n_contrib,
                  float *visibility, float *median_depth, const gs_raster_aux *aux, void *stream);

/* records[i] for every row with opacity > 0 (3 feature channels).  Rows with opacity 0 are skipped (never listed). */
int gs_raster_records(const gs_raster_config *h_cfg, int64_t m, const float *g2d, const float *features,
                      float *records, void *stream);

/* Processing order of the T tiles, heaviest first, by a single-CTA counting sort on the work estimate:
 *   work_stride == 2: work = tile_ranges (T,2), estimate = end - begin (forward);
 *   work_stride == 1: work = tile_work (T) written by gs_raster_fwd (backward).
 * Also zeroes *ticket (the queue head the persistent kernels advance). */
int gs_tile_order(int tiles, const int32_t *work, int work_stride, int32_t *order, uint32_t *ticket, void *stream);

/* (4) Backward rasterisation: back-to-front re-composite, per-thread 2×2 pixels, warp
 *     transpose-reduction, one vector of atomics per (warp, splat).  acc row layout
 *     (acc_stride = gs_raster_acc_stride(channels) floats, must be zeroed by the caller or by
 *     gs_project_fwd):  [0..5] dL/dg2d, [6..6+C) dL/dfeature, [6+C] prune_cost,
 *     [7+C] split_score, [8+C] visibility (Σ α·T, same value the forward flag produces).
 *     vis_plane (nullable, m): if given, the visibility sums go to this contiguous plane instead of
 *     acc[.][8+C], so later passes find the rows a render touched by streaming 4 B per row. */
int gs_raster_acc_stride(int channels);
int gs_raster_bwd(const gs_raster_config *h_cfg, int channels, int64_t m,
                  const float *g2d, const float *features,
                  const int32_t *overlap_to_point, const int32_t *tile_ranges,
                  int width, int height,
                  const float *final_T, const int32_t *n_contrib, const float *d_image,
                  float *acc, int acc_stride, float *vis_plane, const gs_raster_aux *aux, void *stream);

/* L1 + MSE photometric loss against a target, fused with the reference's post-activation
 * clamp(image,0,1) (scene/color_model.py:154-160; trainer/trainer.py:465-476):
 *   loss = w_l1·mean|c−t| + w_mse·mean(c−t)², c = clamp(pred,0,1)
 * d_pred receives dloss/dpred (0 where pred is outside [0,1]); sums[0] += Σ|c−t|,
 * sums[1] += Σ(c−t)² (caller zeroes; divide by n for the means).  target is float [0,1] or,
 * if target_u8 != NULL, uint8 scaled by 1/255 (trainer/trainer.py:548-553). */
int gs_loss_l1_mse(int64_t n, const float *pred, const float *target, const uint8_t *target_u8,
                   float w_l1, float w_mse, float *d_pred, float *sums, void *stream);

/* ---------------------------------------------------------------------------------------
 * (5) Optimiser, statistics, densification.
 * ------------------------------------------------------------------------------------- */
typedef struct gs_optim_config {
  int32_t mode;                     /* 0 = Adam, 1 = LaProp (VisibilityAwareLaProp, scene/mlp_scene.py:58-60) */
  float beta1, beta2, eps;
  float vis_beta, vis_smooth;       /* config/scene/mlp.yaml:25-29                      */
  float grad_clip;                  /* <= 0 disables                                    */
  int32_t bias_correction;
} gs_optim_config;

enum { GS_GROUP_SCALAR = 0, GS_GROUP_VECTOR = 1, GS_GROUP_LOCAL_VECTOR = 2 };  /* config/scene/mlp.yaml:8-14 */
enum { GS_POST_NONE = 0, GS_POST_NORMALIZE = 1, GS_POST_CLAMP = 2 };          /* scene/mlp_scene.py:236-237 */

/* Per-point visibility state, once per optimiser step and before the group steps:
 * for rows with visible[i] > 0: step[i] += 1; running_vis[i] = lerp(vis_beta). */
int gs_optim_vis_update(const gs_optim_config *h_cfg, int64_t n, const float *visible,
                        float *running_vis, float *step, void *stream);

/* `indexes = points.visible.nonzero()` (reference scene/mlp_scene.py:217) without the host sync: ascending row
 * indexes with visible > 0 into indexes[0..count), count -> count_dev[0]. */
size_t gs_visible_indexes_workspace_bytes(int64_t n);
int gs_visible_indexes(int64_t n, const float *visible, int32_t *indexes, int32_t *count_dev,
                       void *workspace, size_t workspace_bytes, void *stream);

/* gs_visible_indexes for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc).  No counterpart in the reference.  Same outputs
 * as gs_visible_indexes, no workspace argument: a single-pass chained-scan compaction whose scratch (one ticket and one
 * status word per 2048-row tile, 33 KB) is a file-scope device array inside the library, held by every GPU that loads
 * it until the process exits.  The call performs no cudaMalloc / cudaFree / allocating library call: one
 * cudaMemsetAsync of the scratch it uses (the array keeps the last call's contents) and one kernel, on `stream`.
 * Limit: 0 <= n <= GS_VISIBLE_INDEXES_STATIC_MAX_ROWS, beyond which it returns GS_ERR_INVALID with a message.
 * Every call on a device shares the one array: calls that overlap on the GPU race — serialise them on one stream. */
#define GS_VISIBLE_INDEXES_STATIC_MAX_ROWS ((int64_t)1 << 24)
int64_t gs_visible_indexes_static_max_rows(void);
int gs_visible_indexes_static(int64_t n, const float *visible, int32_t *indexes, int32_t *count_dev, void *stream);

/* gs_optim_vis_update + gs_visible_indexes fused into one pass over `visible` (one kernel, no host sync). */
size_t gs_optim_begin_workspace_bytes(int64_t n);
int gs_optim_begin(const gs_optim_config *h_cfg, int64_t n, const float *visible, float *running_vis, float *step,
                   int32_t *indexes, int32_t *count_dev, void *workspace, size_t workspace_bytes, void *stream);

/* One parameter group (ParameterClass.step(visibility, indexes, basis), reference
 * scene/mlp_scene.py:214-230).  Rows with visible[i] <= 0 are skipped (the reference's
 * `indexes = visible.nonzero()` done in-kernel, no host sync).  exp_avg is n×dim;
 * exp_avg_sq is n×dim for SCALAR, n for VECTOR / LOCAL_VECTOR.  LOCAL_VECTOR (dim 3)
 * expresses gradient and update in the basis R(q)·diag(max(exp(s),1e-4)) (point_basis,
 * gaussians/split.py:15-19) built from basis_log_scaling / basis_rotation.
 * post: GS_POST_NORMALIZE renormalises each row, GS_POST_CLAMP clamps to [post_lo,post_hi].
 * zero_grad != 0 clears the consumed gradient rows (reference zero_grad, :155-161).
 * indexes/count_dev (both or neither): the ascending list of rows with visible > 0 and its device-side
 * length from gs_visible_indexes — the kernel then touches only those rows (grid-stride over the list). */
int gs_optim_step(const gs_optim_config *h_cfg, int kind, int64_t n,
                  const int32_t *indexes, const int32_t *count_dev, int dim, float lr,
                  float *param, float *grad, float *exp_avg, float *exp_avg_sq,
                  const float *visible, const float *running_vis, const float *step,
                  const float *basis_log_scaling, const float *basis_rotation,
                  int post, float post_lo, float post_hi, int zero_grad, void *stream);

/* Several parameter groups in ONE launch (the reference steps its groups one after the other,
 * scene/mlp_scene.py:214-230; groups are independent of each other except that `position` reads log_scaling /
 * rotation as its basis BEFORE their update, :219).  The groups of a call run concurrently, so a call is refused
 * (GS_ERR_INVALID) when one of its groups updates a plane another one reads as basis — put that group in a later call.
 * Results equal gs_optim_step called once per group, bit for bit. */
#define GS_OPTIM_MAX_GROUPS 8
typedef struct gs_optim_group {
  int32_t kind, dim;                       /* GS_GROUP_*, elements per row                   */
  float lr;
  float *param, *grad, *exp_avg, *exp_avg_sq;
  const float *basis_ls, *basis_q;         /* LOCAL_VECTOR only                              */
  int32_t post;                            /* GS_POST_*                                      */
  float post_lo, post_hi;
} gs_optim_group;
int gs_optim_step_groups(const gs_optim_config *h_cfg, int64_t n, const int32_t *indexes, const int32_t *count_dev,
                         int n_groups, const gs_optim_group *h_groups, const float *visible, const float *running_vis,
                         const float *step, int zero_grad, void *stream);

/* PointState.add_rendering (reference controller/point_state.py:34-50) + MLPScene.add_rendering
 * (scene/mlp_scene.py:242-244) for one rendering whose rows are Gaussians (fused path, idx=NULL)
 * or listed by idx (m rows -> Gaussian idx[i]).  In-view rows: opacity (g2d[5]) > 0.
 * screen scale = larger std-dev in px of the 2-D Gaussian (from the conic).
 *   max_scale_px = max(.), points_in_view += (vis>0), visibility += vis, scene_visible += vis,
 *   split_score = exp_lerp(split_alpha, old, new), prune_cost = exp_lerp(prune_alpha, old, new)
 *   with exp_lerp(t,a,b) = m + log(lerp(e^(a-m), e^(b-m), t)), m = max(a,b) (util/misc.py:56-59).
 * heur points at (prune_cost, split_score) pairs with row stride heur_stride; vis has stride vis_stride.
 * clear_acc / clear_vis (nullable): after a visible row (vis > 0) is consumed its clear_stride accumulator floats and
 * its vis entry are zeroed, so the raster-backward accumulator is all-zero again for the next image without a memset. */
int gs_accumulate_stats(int64_t m, const int64_t *idx, const float *g2d,
                        const float *vis, int vis_stride, const float *heur, int heur_stride,
                        float split_alpha, float prune_alpha,
                        float *prune_cost, float *split_score, float *max_scale_px,
                        int16_t *points_in_view, float *visibility, float *scene_visible,
                        float *clear_acc, int clear_stride, float *clear_vis, void *stream);

/* take_n (reference controller/target_controller.py:150-160): mask of the k smallest
 * (descending=0) or largest values; ties at the threshold broken by lowest index, so exactly
 * min(k,n) entries are set.  NaNs order last.  Radix select, no full sort. */
size_t gs_select_topk_workspace_bytes(int64_t n);
int gs_select_topk_mask(const float *values, int64_t n, int64_t k, int descending,
                        uint8_t *mask, void *workspace, size_t workspace_bytes, void *stream);

/* Stream compaction of rows (points[keep_mask], reference scene/mlp_scene.py:308): step 1 turns a
 * byte mask into destination slots (slots[i] = rank among kept, or -1) and the kept count
 * (count_dev[0]); step 2 moves any number of row arrays with those slots. */
size_t gs_compact_workspace_bytes(int64_t n);
int gs_compact_slots(const uint8_t *keep_mask, int64_t n, int32_t *slots, int32_t *count_dev,
                     void *workspace, size_t workspace_bytes, void *stream);
int gs_compact_rows(const void *src, void *dst, int64_t n, int row_bytes, const int32_t *slots,
                    void *stream);

/* split_gaussians_uniform(points[split_idx], k=2, random_axis=True) (reference
 * gaussians/split.py:87-113) written to rows [out_offset, out_offset+2S) of the destination
 * arrays: children at ∓sep along the sampled axis in the R·S basis, log_scaling[axis] +=
 * log(1/sqrt 2).  axis ~ Multinomial(∝ max(exp(s),eps)) drawn by inverse CDF from uniform[s]
 * (caller-supplied U[0,1) so every rank and the oracle draw the same axis); uniform == NULL
 * selects argmax (random_axis=False).  The caller copies the remaining attributes
 * (rotation, alpha_logit, feature) with gs_gather_rows(repeat=2). */
int gs_split_positions(int64_t s, const int64_t *split_idx, const float *position,
                       const float *log_scaling, const float *rotation, const float *uniform,
                       float sep, float *out_position, float *out_log_scaling, int64_t out_offset,
                       void *stream);
/* dst[out_offset + i*repeat + r] = src[idx[i]] for r < repeat (repeat_interleave, gaussians/split.py:42-44) */
int gs_gather_rows(const void *src, void *dst, int64_t s, const int64_t *idx, int row_bytes,
                   int repeat, int64_t out_offset, void *stream);

/* ---------------------------------------------------------------------------------------
 * (6) Camera-batch data parallelism (new capability — the reference loops over the batch on one GPU,
 *     trainer/trainer.py:500-514).  Every image's contribution (gradient rows of the splats a pixel blended,
 *     their visibility and heuristics, and which rows were in view) is PACKED into one record; every rank then
 *     FOLDS the records of all images of the step in image order, which reproduces the reference's sequential
 *     `.grad` accumulation and PointState.add_rendering order (controller/point_state.py:34-50) on every replica.
 *     Records may be local or peer-GPU (NVLink-mapped) memory.
 * ------------------------------------------------------------------------------------- */
#define GS_EXCHANGE_MAX_GROUPS 8
#define GS_EXCHANGE_MAX_IMAGES 64
#define GS_EXCHANGE_MAX_PEERS 16
typedef struct gs_exchange_groups {
  int32_t count;                          /* parameter groups, in record-plane order            */
  int32_t dim[GS_EXCHANGE_MAX_GROUPS];    /* floats per row of each group                       */
  float *grad[GS_EXCHANGE_MAX_GROUPS];    /* n × dim dense gradient plane of each group         */
} gs_exchange_groups;

/* bytes of one record for n Gaussians, `cap` visible rows and row_floats = Σ dim floats per row */
size_t gs_exchange_record_bytes(int64_t n, int64_t cap, int row_floats);
size_t gs_exchange_pack_workspace_bytes(int64_t n);
/* After one image's backward (gradient planes hold THIS image's rows only): updates the rank's own max_scale_px
 * (order independent; MAX-reduced across ranks before a densification), writes the record, moves the gradient rows
 * of the visible splats into it and zeroes them at the source, and clears the consumed accumulator rows like
 * gs_accumulate_stats.  record[0] = visible count (int32), record[1] = overflow flag (count > cap: grow and redo). */
int gs_exchange_pack(int64_t n, const float *g2d, const float *vis, int vis_stride, const float *heur,
                     int heur_stride, float *max_scale_px, float *clear_acc, int clear_stride,
                     float *clear_vis, const gs_exchange_groups *groups, void *record, int64_t cap,
                     void *workspace, size_t workspace_bytes, void *stream);
/* records: HOST array of n_images device (or peer) pointers, in image order.  Adds every record's rows into the
 * gradient planes and their visibility into scene_visible for rows [grad_begin, grad_end) (32-row aligned; all rows for the
 * replicated optimiser, this rank's slice for the row-sharded one), and folds
 * the controller statistics of rows [stats_begin, stats_end) (4-row aligned; the whole scene on one GPU, this rank's
 * slice under DP — the statistics are only read at a densification, where the slices are all-gathered):
 * split_score / prune_cost by exp_lerp over in-view rows (0 for rows no pixel blended), points_in_view += 1,
 * visibility += vis for blended rows. */
int gs_exchange_fold(int64_t n, int n_images, const void *const *records, int64_t cap,
                     const gs_exchange_groups *groups, float split_alpha, float prune_alpha, float *prune_cost,
                     float *split_score, int16_t *points_in_view, float *visibility, int64_t stats_begin,
                     int64_t stats_end, int64_t grad_begin, int64_t grad_end, float *scene_visible, void *stream);
/* Row-sharded optimiser under DP (reduce by row slice, step 1/W of the rows, all-gather the updated
 * rows): gs_exchange_fold with [grad_begin, grad_end) = this rank's 32-row-aligned slice sums only those gradient rows
 * (and their `visible`); after the optimiser step on the slice, gs_exchange_publish writes the parameter rows it
 * touched — indexes[0 .. count[0]) relative to row_begin, the list gs_optim_begin produced — into every peer's copy of
 * the parameter planes (peer_bases[q] = peer q's mapping of the buffer that starts at local_base here; param_groups:
 * .grad[g] = this rank's PARAMETER plane of group g).  A barrier across ranks must follow before the planes are read. */
int gs_exchange_publish(const int32_t *indexes, const int32_t *count, int64_t max_rows, int64_t row_begin,
                        const gs_exchange_groups *param_groups, const void *local_base, int n_peers,
                        void *const *peer_bases, void *stream);

/* ---------------------------------------------------------------------------------------
 * (7) Per-step controller / scene passes over every splat.
 * ------------------------------------------------------------------------------------- */
/* MCMCController.step, noise branch (reference controller/mcmc_controller.py:91-100): rows with
 * points_in_view > min_views move by point_basis(log_scaling, rotation) · (randn · soft_lt(sigmoid(alpha_logit),
 * opacity_threshold/2, margin 16) · noise_level)   (soft_lt util/misc.py:20-37, sample_gaussians
 * gaussians/split.py:31-36).  randn: caller-supplied N(0,1) draws (n×3), so every rank and the oracle add the same
 * noise.  The prune/respawn branch (:76-89) is mask algebra + gs_select_topk_mask + the densification kernels. */
int gs_mcmc_noise(int64_t n, const float *alpha_logit, const float *log_scaling, const float *rotation,
                  const int16_t *points_in_view, int min_views, float opacity_threshold, float noise_level,
                  const float *randn, float *position, void *stream);

/* MLPScene.compute_reg / reg_loss (reference scene/mlp_scene.py:246-288; saturate util/misc.py:68-69) over the rows a
 * pixel blended (vis > 0, weight w = vis, m = their count; rows are Gaussians, opacity = g2d[5], depth = z_cam):
 *   scale = mean(w·Σ exp(s)²/depth²)   opacity = mean(w·(1−e^(−4·op))²·Σ exp(s)²/depth²)   aspect = mean(w·max exp(s)/min exp(s))
 * compute_sums != 0: sums[0..2] = Σ of the three terms, sums[3] = m (device floats; the loss is Σ weight_k·sums[k]/m).
 * Any gradient pointer non-NULL: gradients of Σ weight_k·reg_k using sums[3] (from this or an earlier call on the
 * stream) are ADDED to d_opacity[i·stride] (the raster accumulator's opacity column), WRITTEN to d_depth[i] (read by
 * gs_project_bwd for blended rows only) and ADDED to d_log_scaling (n×3).  The specular term belongs to the colour
 * MLP and is not part of this path. */
int gs_reg_loss(int64_t n, float w_scale, float w_opacity, float w_aspect, const float *g2d, const float *depth,
                const float *log_scaling, const float *vis, int vis_stride, float *sums, int compute_sums,
                float *d_opacity, int d_opacity_stride, float *d_depth, float *d_log_scaling, void *stream);

/* ---------------------------------------------------------------------------------------
 * (8) Multi-scale SSIM loss: Trainer.compute_ssim_loss, reference trainer/trainer.py:450-462 with
 *     self.ssim = fused_ssim(padding="valid") (:112; fused-ssim's published algorithm: 11x11 Gaussian window sigma 1.5,
 *     zero padded, C1 = 0.01^2, C2 = 0.03^2, map cropped by 5 px).  loss = weight * mean_l (1 - ssim_l), level l >= 1 on
 *     2x2 average-pooled images.  pred: raw render H x W x C channels-last, clamped to [0,1] on load like the L1/MSE loss
 *     (scene/color_model.py:154-160); target fp32 or uint8.  ssim_sums[l] = Σ of the SSIM map over the valid crop of
 *     level l (device; ssim_l = ssim_sums[l] / ((H_l-10)(W_l-10)C)).  d_pred (nullable): weight * d loss / d pred is
 *     ADDED (run after gs_loss_l1_mse, which writes it). */
size_t gs_ssim_workspace_bytes(int height, int width, int channels, int levels);
int gs_ssim_loss(int height, int width, int channels, int levels, const float *pred, const float *target,
                 const uint8_t *target_u8, float weight, float *d_pred, float *ssim_sums, void *workspace,
                 size_t workspace_bytes, void *stream);
/* The whole image loss of Trainer.compute_losses in one call (reference trainer/trainer.py:465-498: l1 / mse terms +
 * compute_ssim_loss): the results of gs_loss_l1_mse followed by gs_ssim_loss, in 2 * levels launches: the L1/MSE sums are
 * taken by the level-0 SSIM forward on the pixels it already holds, every level's forward also writes the 2x2-pooled
 * inputs of the next level, and the level-0 backward adds the L1/MSE gradient when it writes d_pred — once, no separate
 * sweep of the image.  d_pred is WRITTEN; l1_mse_sums[2] = Σ|d|, Σd² and ssim_sums[levels] are zeroed here. */
int gs_photometric_loss(int height, int width, int channels, int levels, const float *pred, const float *target,
                        const uint8_t *target_u8, float w_l1, float w_mse, float w_ssim, float *d_pred,
                        float *l1_mse_sums, float *ssim_sums, void *workspace, size_t workspace_bytes, void *stream);

/* ---------------------------------------------------------------------------------------
 * (9) Colour MLP path of MLPScene: MLPScene.eval_colors -> ColorModel.forward, reference
 *     scene/mlp_scene.py:352-368, scene/color_model.py:124-149, scene/mlp/torch_mlp.py:52-66,89-135.
 *     feature = LayerNorm([point feature, GLO feature]) (no affine); diffuse = act(base MLP(feature)); specular =
 *     act(directional MLP(feature * a + b), intensity_bias -2) with (a, b) = Linear(real SH of normalize(position -
 *     cam_pos)) (scene/mlp/rsh.py convention); act(o) = sigmoid(o[1:4]) * exp(o[0] + bias) (luminance_activation :74-79).
 *     Precision follows the reference's torch.autocast(float16): fp16 inputs / weights / outputs on every Linear and
 *     on the GLU, fp32 accumulation (tensor cores), fp32 LayerNorm, SH and final activation.
 *     weights[10]: device fp32 arrays shaped like the reference's nn.Linear parameters (row-major (out, in)):
 *     base_w1 (2H,F) base_b1 (2H) base_w2 (4,H) base_b2 (4) dir_wsh (2F,S) dir_bsh (2F) dir_w1 (2H,F) dir_b1 (2H)
 *     dir_w2 (4,H) dir_b2 (4); F = point_features + glo_features, S = (sh_degree+1)^2.
 *     Instantiated: point_features 16, glo_features 32, hidden 32 (one hidden layer), sh_degree 5
 *     (config/scene/mlp.yaml:33-43); anything else returns GS_ERR_UNSUPPORTED. */
typedef struct gs_color_mlp {
  int32_t point_features, glo_features, hidden, sh_degree;
  const float *weights[10];
} gs_color_mlp;

/* rgb[i] = diffuse + specular_weight * specular (Colors.total, scene/color_model.py:84-85) for the rows with
 * in_view[i*in_view_stride] > 0 (NULL: every row; pass the opacity column of g2d); other rows are left untouched.
 * specular (n,3), nullable: the specular term alone (RenderedPoints.attributes.specular, read by reg_loss).
 * glo: device (glo_features) — the image's GLO embedding row, or zeros for a non-training view (mlp_scene.py:356-360). */
int gs_color_mlp_fwd(const gs_color_mlp *h_model, int64_t n, const float *in_view, int in_view_stride,
                     const float *feature, const float *position, const float *h_cam_pos, const float *glo,
                     float specular_weight, float *rgb, float *specular, void *stream);

/* Backward of gs_color_mlp_fwd for the rows with row_w[i*row_w_stride] > 0 (the rows a pixel blended: the raster
 * backward's visibility plane).  d_rgb (stride >= 3, e.g. columns 6..8 of the raster accumulator): d loss / d rgb.
 * The forward is recomputed (nothing is saved between the two kernels).  w_specular_reg / reg_sums (nullable): the
 * specular regulariser of reg_loss (reference scene/mlp_scene.py:262,267: weight * mean_i(w_i * sum_c |specular_ic|)
 * over the blended rows, count = reg_sums[3] from gs_reg_loss) is added to d specular.
 * d_feature (n, point_features) is WRITTEN for the processed rows (ADDED when accumulate != 0; other rows untouched);
 * d_glo (glo_features) and h_grads->weights[10] (shaped like the model's weights) are ACCUMULATED (batch = sum, like
 * .grad).  fp16 operands of the backward carry an internal loss scale (the reference's autocast backward has none). */
int gs_color_mlp_bwd(const gs_color_mlp *h_model, const gs_color_mlp *h_grads, int64_t n, const float *row_w,
                     int row_w_stride, const float *feature, const float *position, const float *h_cam_pos,
                     const float *glo, float specular_weight, const float *d_rgb, int d_rgb_stride,
                     float w_specular_reg, const float *reg_sums, float *d_feature, float *d_glo, int accumulate,
                     void *stream);

/* ---------------------------------------------------------------------------------------
 * (10) Point clustering for the view-similarity features: assign_clusters / kmeans_iter, reference
 *      visibility/cluster.py:139-171 (pykeops argmin of squared distances), used by PointClusters.cluster (:22) on the
 *      splat positions at every evaluation (trainer/trainer.py:327-370).  1 <= d <= 8, k*d*4 <= 48 KB.
 *      gs_assign_clusters: labels[i] = argmin_j |x_i - c_j|^2 (ties: lowest j; a NaN centroid is never nearest).
 *      gs_kmeans: `iters` rounds of (assign, centroid = mean of its points; an emptied cluster becomes NaN = 0/0 like
 *      the reference).  centroids (k,d): initial values in, final values out; labels: the LAST assignment.
 *      No host synchronisation; labels equal the CPU oracle's exactly for equal centroids, centroid sums are fp32
 *      reductions in arbitrary order (as the reference's scatter_add_ on CUDA). */
size_t gs_kmeans_workspace_bytes(int k, int d);
int gs_assign_clusters(int64_t n, int d, int k, const float *x, const float *centroids, int32_t *labels, void *stream);
int gs_kmeans(int64_t n, int d, int k, int iters, const float *x, float *centroids, int32_t *labels, void *workspace,
              size_t workspace_bytes, void *stream);

/* Per-view cluster features (PointClusters.view_features, reference visibility/cluster.py:40-55, called per training view
 * from trainer/trainer.py:337-340,356-359): features[labels[row]] += vis over the entries with vis > threshold.
 * indexes == NULL: vis is the (n) visibility plane of one render (row = entry); else vis / indexes are the (v) compacted
 * (idx, visibility) lists of query_visibility.  features (num_clusters) is zeroed here. */
int gs_view_features(int64_t n, const float *vis, const int64_t *indexes, int64_t v, const int32_t *labels, float threshold,
                     int num_clusters, float *features, void *stream);
/* point_visibility (reference visibility/query_points.py:62-92, used to crop the initial cloud, trainer/init.py): counts[i] =
 * number of cameras in which point i projects inside the image with near < depth < far.  cameras: (num_cameras, 20) floats
 * = image_t_world (4x4 row-major, K·T_camera_world) | width, height, near, far. */
int gs_point_visibility(int64_t n, const float *points, int num_cameras, const float *cameras, int32_t *counts, void *stream);

/* estimate_scale (reference gaussians/loading.py:45-61, pykeops Kmin over all N x N distances): scales[i] = mean
 * distance from points[i] to its num_neighbors (1..8) nearest neighbours, computed as the reference does (the
 * num_neighbors + 1 smallest squared distances to ALL points, the smallest — self — dropped, sqrt, mean).  Uniform grid +
 * growing shell search; results equal the brute-force oracle bit for bit.  One host read (the bounding box): an
 * initialisation-time call (initial log_scaling, gaussians/loading.py:23-34). */
size_t gs_knn_scale_workspace_bytes(int64_t n);
int gs_knn_scale(int64_t n, const float *points, int num_neighbors, float *scales, void *workspace,
                 size_t workspace_bytes, void *stream);

/* torch.optim.Adam's update on a dense fp32 array in one launch (the colour MLP's weights, reference
 * scene/color_model.py:162-170, betas (0.9, 0.999); a GLO row an image touched, :37-44, betas (0.8, 0.95)):
 * m = b1 m + (1-b1) g; v = b2 v + (1-b2) g^2; p -= lr/(1-b1^step) * m / (sqrt(v/(1-b2^step)) + eps); step >= 1.
 * zero_grad != 0 clears grad. */
int gs_adam_step(int64_t n, float *param, float *grad, float *exp_avg, float *exp_avg_sq, float lr, float beta1,
                 float beta2, float eps, int step, int zero_grad, void *stream);

/* The same update on the rows r of a (rows, dim) table with touched[r] > 0 (the GLO embeddings an image used since the
 * last step; SparseAdam semantics: other rows keep parameters AND moments).  steps (rows) int32 and touched (rows) float
 * live on the device: steps[r] is incremented, grad row and touched[r] are cleared.  No host read-back; under camera-batch
 * DP `touched` is summed over the ranks together with the gradients. */
int gs_adam_rows(int rows, int dim, float *param, float *grad, float *exp_avg, float *exp_avg_sq, int32_t *steps,
                 float *touched, float lr, float beta1, float beta2, float eps, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* GSB200_H */
