/* brush_b200.h -- C ABI of the B200-native differentiable splat pipeline.
 *
 * Drop-in boundary for ONE path of wartron/brush: brush_render::Backend::render_splats
 * (crates/brush-render/src/lib.rs:66-86), its autodiff backward
 * (crates/brush-render/src/render.rs:465-626) and the two primitives it is built on
 * (brush-sort/src/lib.rs:32, brush-prefix-sum/src/lib.rs:17).  Plain pointers and sizes only;
 * every pointer is a DEVICE pointer unless its name ends in `_host`.  `stream` is a
 * cudaStream_t passed as void* so a Rust/ctypes binding needs no CUDA headers.
 *
 * All calls enqueue work on `stream` and return; none synchronises unless stated.
 * Return value: 0 = ok, <0 = error (see BRUSH_E_*); brush_last_error() gives the text.
 * A ctx is bound to one device, owns only scratch memory, and is not thread-safe.
 */
#ifndef BRUSH_B200_H
#define BRUSH_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BRUSH_OK 0
#define BRUSH_E_INVALID (-1)   /* bad argument (DimCheck / sh_degree_from_coeffs panics in the reference) */
#define BRUSH_E_CAPACITY (-2)  /* intersection buffer too small (the reference truncates silently) */
#define BRUSH_E_CUDA (-3)      /* CUDA runtime error */

#define BRUSH_TILE_WIDTH 16u   /* helpers.wgsl:1 */
#define BRUSH_PROJECTED_FLOATS 9u /* helpers.wgsl:33-43 ProjectedSplat */

/* == shaders::helpers::RenderUniforms, crates/brush-render/src/shaders/helpers.wgsl:7-30 (112 bytes).
 * Filled by the caller as render.rs:102-116 does; tile_bounds, sh_degree, num_visible and total_splats
 * are (re)computed by brush_render_forward.  viewmat is column-major world->camera. */
typedef struct BrushUniforms {
    float viewmat[16];
    float focal[2];
    uint32_t img_size[2];
    uint32_t tile_bounds[2];
    float pixel_center[2];
    uint32_t sh_degree;
    uint32_t num_visible;
    uint32_t total_splats;
    uint32_t padding;
} BrushUniforms;

/* == brush_render::RenderAux, crates/brush-render/src/lib.rs:20-33.  Caller-allocated device buffers:
 *   projected_splats        f32[N,9]      rows >= num_visible are unspecified
 *   uniforms_buffer         i32[28]       the 112-byte RenderUniforms; word 25 = num_visible
 *   num_intersections       i32[1]
 *   num_visible             i32[1]
 *   final_index             i32[H,W]      (not written when raster_u32)
 *   cum_tiles_hit           i32[N]        inclusive scan; entries >= num_visible equal num_intersections
 *   tile_bins               i32[Ty,Tx,2]
 *   compact_gid_from_isect  i32[isect_capacity]
 *   global_from_compact_gid i32[N]        entries >= num_visible are 0
 *   tile_id_from_isect      i32[isect_capacity] or NULL (sorted tile ids; kept for tests, not in RenderAux)
 * isect_capacity replaces the reference's hard-coded min(N*tiles, 128*65535) (render.rs:204-206). */
typedef struct BrushAux {
    float* projected_splats;
    int32_t* uniforms_buffer;
    int32_t* num_intersections;
    int32_t* num_visible;
    int32_t* final_index;
    int32_t* cum_tiles_hit;
    int32_t* tile_bins;
    int32_t* compact_gid_from_isect;
    int32_t* global_from_compact_gid;
    int32_t* tile_id_from_isect;
    uint64_t isect_capacity;
} BrushAux;

typedef struct brush_ctx brush_ctx;

/* Stage indices for brush_stage_times_ms (== the reference's tracing spans, render.rs:127-267,515-577). */
enum {
    BRUSH_STAGE_PROJECT_SPLATS = 0,
    BRUSH_STAGE_DEPTH_SORT,
    BRUSH_STAGE_PROJECT_VISIBLE,
    BRUSH_STAGE_PREFIX_SUM,
    BRUSH_STAGE_MAP_INTERSECTS,
    BRUSH_STAGE_TILE_SORT,
    BRUSH_STAGE_TILE_BINS,
    BRUSH_STAGE_RASTERIZE,
    BRUSH_STAGE_RASTERIZE_BWD,
    BRUSH_STAGE_GATHER_PROJECT_BWD,
    BRUSH_STAGE_COUNT
};

int brush_ctx_create(int device, brush_ctx** out);
void brush_ctx_destroy(brush_ctx* ctx);
const char* brush_last_error(brush_ctx* ctx);
const char* brush_version(void);

/* Tile-sort strategy: 0 = two-stage (depth sort, then stable sort by tile id: the reference's structure,
 * render.rs:151-156,225-237), 1 = single fused 64-bit (tile<<32 | depth_bits) key sort.  Same output. */
int brush_ctx_set_sort_mode(brush_ctx* ctx, int fused64);
/* View-parallel overlap (no reference counterpart).  `event` (a cudaEvent_t, or NULL to cancel) marks
 * the point at which sh_coeffs becomes readable: the NEXT brush_render_forward on this ctx reads sh_coeffs only
 * behind it (cudaStreamWaitEvent on its stream) and runs everything that does not need colours -- projection, both
 * sorts, the scan, the intersection lists, the tile bins -- before that.  The SH evaluation then runs as its own
 * kernel just before the rasterizer: same arithmetic and bits as the fused kernel.  One-shot: consumed by that
 * forward.  brush_ctx_set_split_colors(1) makes every forward take the deferred-colour route (measurements). */
int brush_ctx_set_sh_ready_event(brush_ctx* ctx, void* event);
int brush_ctx_set_split_colors(brush_ctx* ctx, int enabled);
/* Per-stage CUDA-event timing (the sync-span equivalent, crates/sync-span/src/lib.rs:26-41). Off by default. */
int brush_ctx_set_timing(brush_ctx* ctx, int enabled);
/* Synchronises `stream`, then fills out[0..n) with the last call's stage times in ms. */
int brush_stage_times_ms(brush_ctx* ctx, float* out_host, int n, void* stream);
/* Diagnostics for the roofline report (no reference counterpart; SURVEY.md 6 / 8d "measured on the box").
 * brush_measure_peaks runs four micro-kernels on `stream`, synchronises, and fills out4_host with the device's
 * measured FP32 TFLOP/s (FMA = 2 flop) for scalar FFMA with constant-bank operands [0], scalar FFMA with three register
 * operands [1] and packed FFMA2 (fma.rn.f32x2) [2], and 1e12 MUFU.EX2 per second [3].  While pair counting is on, the two rasterizers count the (pixel,
 * splat) pairs they evaluate after culling; brush_ctx_read_pair_counts synchronises, writes {forward, backward}
 * totals since the last read into out_host[2] and clears them.  Off by default; never on in a timed region. */
int brush_measure_peaks(brush_ctx* ctx, float* out4_host, void* stream);
int brush_ctx_set_pair_counting(brush_ctx* ctx, int enabled);
int brush_ctx_read_pair_counts(brush_ctx* ctx, uint64_t* out_host, void* stream);
/* Synchronises `stream` and returns the sticky device status: BRUSH_OK or BRUSH_E_CAPACITY; clears it. */
int brush_ctx_check(brush_ctx* ctx, void* stream);

/* == render_forward, crates/brush-render/src/render.rs:55-323.
 * means f32[N,3], log_scales f32[N,3], quats f32[N,4] (normalised, w first), sh_coeffs f32[N,K,3],
 * raw_opacity f32[N]; K in {1,4,9,16,25}.  out_img: f32[H,W,4], or u32[H,W] packed RGBA8 if raster_u32. */
int brush_render_forward(brush_ctx* ctx, const BrushUniforms* uniforms_host, uint32_t N, uint32_t K,
                         const float* means, const float* log_scales, const float* quats, const float* sh_coeffs,
                         const float* raw_opacity, int raster_u32, void* out_img, const BrushAux* aux, void* stream);

/* == RenderBackwards::backward, crates/brush-render/src/render.rs:468-625.
 * Inputs: the checkpointed parameters, the forward's out_img and aux, and v_out_img f32[H,W,4].
 * Outputs (dense, fully written; zero for non-visible splats): v_means f32[N,3], v_xy f32[N,2] (the xy_dummy
 * gradient), v_log_scales f32[N,3], v_quats f32[N,4], v_sh_coeffs f32[N,K,3], v_raw_opacity f32[N]. */
int brush_render_backward(brush_ctx* ctx, uint32_t N, uint32_t K, const float* means, const float* log_scales,
                          const float* quats, const float* raw_opacity, const float* out_img,
                          const float* v_out_img, const BrushAux* aux, float* v_means, float* v_xy,
                          float* v_log_scales, float* v_quats, float* v_sh_coeffs, float* v_raw_opacity,
                          void* stream);

/* == brush_sort::radix_argsort, brush-sort/src/lib.rs:32-37.  Stable ascending key-value sort of the first
 * *n_dev entries (device-side count) on the low 4*ceil(bits/4) key bits; entries >= *n_dev of the outputs are
 * unspecified.  max_n = allocated length of all four arrays. */
int brush_radix_argsort_u32(brush_ctx* ctx, const uint32_t* keys, const uint32_t* values, const int32_t* n_dev,
                            uint32_t max_n, uint32_t bits, uint32_t* keys_out, uint32_t* values_out, void* stream);
/* 64-bit-key variant used by the fused (tile | depth) sort. */
int brush_radix_argsort_u64(brush_ctx* ctx, const uint64_t* keys, const uint32_t* values, const int32_t* n_dev,
                            uint32_t max_n, uint32_t bits, uint64_t* keys_out, uint32_t* values_out, void* stream);

/* == brush_prefix_sum::prefix_sum, brush-prefix-sum/src/lib.rs:17.  Inclusive u32 scan of n elements. */
int brush_prefix_sum_u32(brush_ctx* ctx, const uint32_t* in, uint32_t n, uint32_t* out, void* stream);

/* NO COUNTERPART IN THE REFERENCE -- added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The same inclusive u32
 * scan as brush_prefix_sum_u32, with the kernel's scratch (tile ticket + look-back status words) in a file-scope
 * `static __device__` array inside the library instead of a ctx: it takes no ctx and calls no cudaMalloc, cudaFree or
 * allocating routine.  It runs on the CURRENT device, on `stream`; `in` / `out` are the caller's device buffers
 * (distinct).  Limit: n <= BRUSH_NOALLOC_SCAN_MAX_N, larger sizes return BRUSH_E_INVALID.  The error text goes
 * to the optional caller-owned host buffer err_host[err_len].  Every GPU that loads the library holds the array
 * (32 KB + 64 B) until exit.  All calls on a device share it and it keeps the last call's contents (each call clears
 * what it reads, on `stream`): calls that overlap on the GPU (different streams, no ordering between them) race;
 * calls on one stream, or ordered by events, do not. */
#define BRUSH_NOALLOC_SCAN_MAX_N 16777216u
int brush_prefix_sum_u32_noalloc(const uint32_t* in, uint32_t n, uint32_t* out, void* stream, char* err_host,
                                 uint32_t err_len);

/* Host-staged convenience for FFI callers that hold HOST buffers (bench e2e leg, smoke tests): copies the five
 * parameter arrays H2D, renders, copies out_img (f32[H,W,4]) and the two counters back, synchronises. */
int brush_render_forward_host(brush_ctx* ctx, const BrushUniforms* uniforms_host, uint32_t N, uint32_t K,
                              const float* means_host, const float* log_scales_host, const float* quats_host,
                              const float* sh_coeffs_host, const float* raw_opacity_host, uint64_t isect_capacity,
                              float* out_img_host, uint32_t* num_visible_host, uint32_t* num_intersections_host);

/* ---- train-step kernels either side of the path ---- */

/* == (pred_compare - gt).abs().mean() and its gradient, crates/brush-train/src/train.rs:255.
 * pred f32[H,W,4]; gt f32[H,W,C], C = 3 (compare RGB) or 4 (compare RGBA, train.rs:249-253).
 * Writes v_out_img f32[H,W,4] = d loss / d pred and loss_out f32[1] = the mean. */
int brush_l1_loss_backward(brush_ctx* ctx, const float* pred, const float* gt, uint32_t height, uint32_t width,
                           uint32_t gt_channels, float* v_out_img, float* loss_out, void* stream);

/* == the full image loss of SplatTrainer::step and its gradient, crates/brush-train/src/train.rs:249-268 with
 * crates/brush-train/src/ssim.rs:7-112:  loss = (1 - w) * mean|pred_compare - gt| - w * ssim(pred_rgb, gt_rgb),
 * SSIM with the reference's 11-tap sigma-1.5 window, zero padding 6 (maps are (H+2) x (W+2)), variances clamped
 * at 0.  ssim_window must be 11 (the reference default, TrainConfig::ssim_window_size); ssim_weight = 0 gives
 * brush_l1_loss_backward.  Writes v_out_img f32[H,W,4] and loss_out f32[1]. */
int brush_image_loss_backward(brush_ctx* ctx, const float* pred, const float* gt, uint32_t height, uint32_t width,
                              uint32_t gt_channels, float ssim_weight, uint32_t ssim_window, float* v_out_img,
                              float* loss_out, void* stream);

/* == the five optim.step calls of SplatTrainer::step, crates/brush-train/src/train.rs:318-359, as ONE launch.
 * Burn Adam: m = b1 m + (1-b1) g; v = b2 v + (1-b2) g^2; p -= lr * (m/(1-b1^t)) / (sqrt(v/(1-b2^t)) + eps).
 * sh_coeffs_per_channel > 1 marks the SH group [N,K,3]: coefficients k > 0 use lr * sh_rest_lr_scale
 * (the 1/lr_coeffs_sh_scale lerp of train.rs:334-348).  grad_scale multiplies every gradient (1/world for DP). */
typedef struct BrushAdamGroup {
    float* param;
    const float* grad;
    float* m;
    float* v;
    uint64_t count;
    float lr;
    uint32_t sh_coeffs_per_channel;
    float sh_rest_lr_scale;
} BrushAdamGroup;
int brush_adam_step(brush_ctx* ctx, const BrushAdamGroup* groups_host, int n_groups, float beta1, float beta2,
                    float eps, uint32_t step, float grad_scale, void* stream);

/* == rasterize_backwards + gather_grads/project_backwards (render.rs:494-594), the autodiff VJP of the rotation
 * normalisation (gaussian_splats.rs:174-175) and the five optimizer steps of train.rs:318-359, for the
 * single-device step with one view per batch: same rasterizer launch as brush_render_backward, then ONE kernel
 * computes each splat's gradients in registers and applies brush_adam_step's update in place, so the dense
 * gradients never touch HBM.  `rotation` is the raw (un-normalised) parameter, `quats` the normalised copy the
 * forward rendered with; v_xy [N,2] (may be NULL) receives the xy gradient; grad_norm_accum / visible_accum (f32[N],
 * both or neither) receive the refine statistics of train.rs:284-316 in the same pass (+= |v_xy * (w/2, h/2)| *
 * grad_scale, += 1 for visible splats).  All pointers are device pointers;
 * parameters and moments are updated in place.  Results equal brush_render_backward -> brush_quat_normalize_vjp
 * -> brush_adam_step on the same inputs. */
typedef struct BrushTrainState {
    float *means, *log_scales, *rotation, *sh_coeffs, *raw_opacity;          /* parameters */
    float *m_means, *v_means, *m_log_scales, *v_log_scales, *m_rotation, *v_rotation;
    float *m_sh_coeffs, *v_sh_coeffs, *m_raw_opacity, *v_raw_opacity;        /* Adam moments, zero at start */
    float lr_mean, lr_scale, lr_rotation, lr_coeffs_dc, lr_opac;
    float sh_rest_lr_scale;                                                    /* 1 / lr_coeffs_sh_scale */
} BrushTrainState;
int brush_render_backward_adam(brush_ctx* ctx, uint32_t N, uint32_t K, const BrushTrainState* state,
                               const float* quats, const float* out_img, const float* v_out_img,
                               const BrushAux* aux, float* v_xy, float* grad_norm_accum, float* visible_accum,
                               float beta1, float beta2, float eps, uint32_t step, float grad_scale, void* stream);

#define BRUSH_DP_MAX_WORLD 8
/* ---- view-parallel step without a gradient exchange (no reference counterpart) ----
 * Every rank renders its own view and publishes only what rasterize_backwards leaves per splat: the nine sums of
 * render.rs:505-507 (48-byte rows) -- 5x less than the dense gradients.  Then every rank runs
 * gather_grads/project_backwards for EVERY view on those rows (peer loads over NVLink), sums the gradients in view
 * order, and applies the Adam update of brush_render_backward_adam to its own full replica.  Replicas stay
 * bit-identical with no parameter broadcast.
 *   rows[i]:  rank i's row block for THIS step as mapped in this process: f32[N*12 + 32], peer-mapped on every
 *             rank; callers alternate between two blocks by step parity (that is what makes one barrier enough).
 *   flags[i]: rank i's flag block, u32[8], zeroed once at allocation, peer-mapped.
 *   epoch:    the same strictly increasing value (>= 1) on every rank for every step.
 * brush_render_backward_rows = rasterize_backwards into rows[rank] + the view's visibility flags, uniforms and
 * loss term (device f32, may be NULL), then signals every rank.  brush_views_adam_step waits for all ranks'
 * signals and does the rest; grad_norm_accum / visible_accum (f32[N], both or neither) receive the refine
 * statistics of train.rs:284-316 summed over the step's views (+= norm * grad_scale, += visible count);
 * loss_out (device f32, may be NULL) = grad_scale * sum of the ranks' loss terms. */
typedef struct BrushViewExchange {
    int32_t world, rank;
    float* rows[BRUSH_DP_MAX_WORLD];
    uint32_t* flags[BRUSH_DP_MAX_WORLD];
    uint32_t epoch;
} BrushViewExchange;
int brush_render_backward_rows(brush_ctx* ctx, uint32_t N, uint32_t K, const float* out_img, const float* v_out_img,
                               const BrushAux* aux, const float* loss_term, const BrushViewExchange* x,
                               void* stream);
int brush_views_adam_step(brush_ctx* ctx, uint32_t N, uint32_t K, const BrushTrainState* state, const float* quats,
                          const BrushViewExchange* x, float* grad_norm_accum, float* visible_accum, float* loss_out,
                          float beta1, float beta2, float eps, uint32_t step, float grad_scale, void* stream);

/* ---- hybrid view-parallel step (no reference counterpart) ----
 * The SH group is 3K of the 13 + 3K gradient floats of a splat, but per view it is basis(viewdir) (x) v_rgb
 * (gather_grads.wgsl:165-232): three floats of information.  brush_render_backward_hybrid = brush_render_backward
 * without the dense SH gradient: the five small gradients are written densely (for brush_dp_adam_step), and the view's
 * colour rows (v_r, v_g, v_b, visible) go to x->rows[x->rank] (f32[4 N + 32], peer-mapped, two blocks alternating
 * by step parity like the `views` exchange; the trailer holds the view's uniforms and loss term), followed by the
 * signal to every rank.  brush_sh_views_adam_step waits for all ranks' signals, rebuilds
 * sum_v basis(viewdir_v) (x) v_rgb_v for ALL splats in view order from the colour rows (peer loads) and applies the
 * SH part of brush_adam_step to this rank's full SH replica -- bit-identical on every rank.  `means` must be the
 * means the step's forwards rendered with, so call it BEFORE the small groups are updated.  Inbound NVLink bytes
 * per splat and step: 16 (world - 1) + 44 instead of 237 (dense, world 8) or 48 (world - 1) (`views`). */
int brush_render_backward_hybrid(brush_ctx* ctx, uint32_t N, uint32_t K, const float* means, const float* log_scales,
                                 const float* quats, const float* raw_opacity, const float* out_img,
                                 const float* v_out_img, const BrushAux* aux, float* v_means, float* v_xy,
                                 float* v_log_scales, float* v_quats, float* v_raw_opacity, const float* loss_term,
                                 const BrushViewExchange* x, void* stream);
int brush_sh_views_adam_step(brush_ctx* ctx, uint32_t N, uint32_t K, const float* means, float* sh_coeffs,
                             float* m_sh_coeffs, float* v_sh_coeffs, float lr_coeffs_dc, float sh_rest_lr_scale,
                             const BrushViewExchange* x, float* loss_out, float beta1, float beta2, float eps,
                             uint32_t step, float grad_scale, void* stream);

/* == image_to_tensor (crates/brush-dataset/src/lib.rs:11-24: u8 pixels -> f32 / 255) applied on the device after the
 * upload, so the host keeps one byte per channel and the per-step upload is 4x smaller.  n = number of channel
 * values (H*W*C); src 4-byte aligned, dst 16-byte aligned. */
int brush_image_u8_to_f32(brush_ctx* ctx, const uint8_t* src, uint64_t n, float* dst, void* stream);

/* == the rotation normalisation Splats::render applies before the operator and its VJP (Burn elementwise ops +
 * autodiff in the reference, crates/brush-render/src/gaussian_splats.rs:174-175).  quats f32[n,4]. */
int brush_quat_normalize(brush_ctx* ctx, const float* quats, uint32_t n, float* out, void* stream);
int brush_quat_normalize_vjp(brush_ctx* ctx, const float* quats, uint32_t n, float* v_inout, void* stream);

/* == the xy-gradient refine statistics of one view, crates/brush-train/src/train.rs:284-316:
 * grad_norm[g] (+)= |v_xy[g] * (w/2, h/2)|, visible[g] (+)= 1 for the view's visible splats; accumulate = 0
 * overwrites both (first view of a batch). */
int brush_xy_grad_stats(brush_ctx* ctx, const float* v_xy, const BrushAux* aux, uint32_t n, uint32_t width,
                        uint32_t height, int accumulate, float* grad_norm, float* visible, void* stream);

/* ---- view-parallel optimizer step over NVLink peer memory (no reference counterpart: the
 * reference asserts a batch of one view on one device, crates/brush-train/src/train.rs:216-219) ----
 * One launch per rank replaces "all-reduce the dense gradients, then brush_adam_step on every rank": rank r owns
 * the r-th 1/world of every segment, sums that slice of all ranks' `src` buffers through peer loads (or one
 * multimem.ld_reduce when multicast pointers are given), applies the Adam update of brush_adam_step to it (moments
 * m, v are touched only on the owned slice) or just the sum (adam = 0: refine statistics), and stores the result
 * into every rank's `dst` buffer.  All buffers named here except m, v, loss_out must be peer-mapped on every rank
 * (cudaIpc / CUDA VMM / torch symmetric memory); src[i], dst[i], flags[i] are rank i's buffers as mapped in THIS
 * process.  flags[i]: uint32[2 * BRUSH_DP_MAX_WORLD] per lane (see `lane`), zeroed once at allocation.  epoch: the same strictly
 * increasing value (>= 1) on every rank for every call.  Every rank must make the matching call. */
typedef struct BrushDpSegment {
    uint64_t src_offset;   /* floats, multiple of 4 */
    uint64_t dst_offset;   /* floats, multiple of 4 */
    uint64_t count;        /* floats, multiple of 4 */
    int32_t adam;
    float lr;
    uint32_t sh_coeffs_per_channel;
    float sh_rest_lr_scale;
} BrushDpSegment;
typedef struct BrushDpStep {
    int32_t world, rank;
    const float* src[BRUSH_DP_MAX_WORLD];
    float* dst[BRUSH_DP_MAX_WORLD];
    const float* src_multicast;   /* NVSwitch multicast mapping of src (sum = one multimem.ld_reduce), or NULL */
    float* dst_multicast;         /* NVSwitch multicast mapping of dst (broadcast = one multimem.st), or NULL */
    uint32_t* flags[BRUSH_DP_MAX_WORLD];
    float* m;                     /* local Adam moments, indexed like dst */
    float* v;
    uint64_t loss_offset;         /* float index inside src of the rank's loss term */
    float* loss_out;              /* local f32[1] = grad_scale * sum of the ranks' loss terms, or NULL */
    uint32_t epoch;
    /* Two exchanges may be in flight at once on different streams (e.g. the small parameter groups on the step's
     * stream and the SH group on a side stream that the next forward overlaps): each uses its own `lane` (0 or 1) =
     * its own flag words inside flags[] (which must then hold 4 * BRUSH_DP_MAX_WORLD words) and arrival counter.
     * Epochs are per lane.  ctas_per_sm = 0 picks the default grid (6 CTAs per SM); an overlapped exchange should
     * leave room for the kernels it runs beside (2). */
    int32_t lane;
    int32_t ctas_per_sm;
} BrushDpStep;
int brush_dp_adam_step(brush_ctx* ctx, const BrushDpStep* step_host, const BrushDpSegment* segments_host,
                       int n_segments, float beta1, float beta2, float eps, uint32_t adam_step, float grad_scale,
                       void* stream);
/* The slice [begin, end) of a segment of `count` floats that `rank` owns (host helper, no GPU needed). */
void brush_dp_owned_range(uint64_t count, int world, int rank, uint64_t* begin, uint64_t* end);

/* ---- densify / prune: the arithmetic of SplatTrainer::refine_splats, crates/brush-train/src/train.rs:395-553
 * (with prune_points :560-596, concat_splats :598-621, quaternion_vec_multiply :140-177, reset_opacity :205-209).
 * Two calls, because the caller has to allocate the output between them (the reference awaits argwhere twice):
 *   brush_refine_plan   classifies every splat (clone: small & high 2-D gradient; split: large & high gradient),
 *                       applies the two prune tests (opacity < cull_alpha_thresh, max scale > cull_scale_thresh) to
 *                       every row the refine yields, and returns the counts -- it synchronises the stream.
 *   brush_refine_apply  writes the new splat set: surviving post-step splats, then surviving clones (PRE-step values),
 *                       then surviving split copies (pre-step mean + R(q) (sample * scale), scale / 1.6), each in
 *                       index order -- the order of the reference's concat + two order-preserving prunes.
 * brush_refine_apply must follow brush_refine_plan on the same ctx with the same n and buffers.  samples:
 * f32[num_split, 3], the N(0, 0.5) draws of train.rs:497-499 in split-candidate order (an input, so that every rank
 * of a view-parallel run uses the same ones).  reset_alpha != 0: every output opacity = inverse_sigmoid(
 * reset_alpha_value) (train.rs:557-560). */
typedef struct BrushSplatParams {   /* device pointers */
    float* means;         /* [n,3] */
    float* rotation;      /* [n,4] */
    float* sh_coeffs;     /* [n,K,3] */
    float* raw_opacity;   /* [n] */
    float* log_scales;    /* [n,3] */
} BrushSplatParams;
typedef struct BrushRefineConfig {  /* TrainConfig fields, train.rs:48-71 */
    float densify_grad_thresh;
    float densify_size_thresh;
    float cull_alpha_thresh;
    float cull_scale_thresh;
    int32_t reset_alpha;
    float reset_alpha_value;
} BrushRefineConfig;
typedef struct BrushRefineCounts {  /* == RefineStats (train.rs:112-118) + the size of the new set */
    uint32_t num_split, num_cloned, alpha_pruned, scale_pruned, num_out, reserved[3];
} BrushRefineCounts;
int brush_refine_plan(brush_ctx* ctx, uint32_t n, const float* grad_2d_accum, const float* xy_grad_counts,
                      const BrushSplatParams* pre_step, const BrushSplatParams* post_step,
                      const BrushRefineConfig* config, BrushRefineCounts* counts_host, void* stream);
int brush_refine_apply(brush_ctx* ctx, uint32_t n, uint32_t K, const BrushSplatParams* pre_step,
                       const BrushSplatParams* post_step, const float* samples, const BrushRefineConfig* config,
                       const BrushSplatParams* out, void* stream);

/* Number of kernels this library has launched in this process (all contexts). */
uint64_t brush_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* BRUSH_B200_H */
