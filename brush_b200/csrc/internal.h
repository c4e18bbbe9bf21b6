// Internal launch interface between the translation units of libbrush_b200.so.
// Each function enqueues kernels on `st` and returns; no synchronisation.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>

#include "brush_math.cuh"

namespace brush {

// Word offsets inside the device copy of RenderUniforms (i32[28]).
constexpr int UNI_NUM_VISIBLE_WORD = 25;  // render.rs:145-149
// Row stride (floats) of the rasterize_backwards accumulator v_projected[N][VP_STRIDE]:
// (v_x, v_y, v_cxx, v_cxy | v_cyy, v_r, v_g, v_b | v_a, -, -, -); 48-byte rows keep red.global.add.v4.f32 aligned.
constexpr uint32_t VP_STRIDE = 12;

// Device control block, zeroed by one memset at the start of every forward.
struct Control {
    uint32_t ticket[8];        // dynamic tile tickets (one per look-back kernel in the chain)
    uint32_t status;           // sticky error bits (1 = intersection capacity overflow)
    uint32_t num_visible;
    uint32_t num_intersections;
    uint32_t pad[5];
};

// ---- project.cu (-fmad=false) ----
void launch_project_forward(const Cam& cam, uint32_t N, const float* means, const float* log_scales,
                            const float* quats, uint32_t* presort_gid, uint32_t* depth_bits,
                            uint64_t* lb_status, uint32_t* ticket, int32_t* uniforms_dev, int32_t* num_visible_out,
                            uint8_t* visible_mask, cudaStream_t st);
void launch_project_visible(const Cam& cam, uint32_t N, uint32_t K, const float* means, const float* log_scales,
                            const float* quats, const float* coeffs, const float* raw_opac,
                            const uint32_t* global_from_compact, const int32_t* num_visible_dev, float* projected,
                            uint32_t* num_tiles_hit, uint64_t* tile_masks, uint32_t* tile_boxes, bool with_colors,
                            cudaStream_t st);
// The colour columns of `projected` when launch_project_visible ran with with_colors = false.
void launch_project_colors(const Cam& cam, uint32_t N, uint32_t K, const float* means, const float* coeffs,
                           const uint32_t* global_from_compact, const int32_t* num_visible_dev, float* projected,
                           cudaStream_t st);
void launch_map_intersects(const Cam& cam, uint32_t N, const float* projected, const uint32_t* cum_tiles_hit,
                           const int32_t* num_visible_dev, uint64_t capacity, uint32_t* tile_id_from_isect,
                           uint32_t* compact_gid_from_isect, const uint32_t* depth_sorted_bits,
                           uint64_t* key64_from_isect, uint32_t* status_word, const uint64_t* tile_masks,
                           const uint32_t* tile_boxes, cudaStream_t st);

// ---- scan.cu ----
// Inclusive scan of in[0..n) -> out. If total_out != nullptr, also writes the grand total (min(total,cap)) there
// and sets bit 0 of *status_word when total > cap.
void launch_inclusive_scan(const uint32_t* in, uint32_t n, uint32_t* out, uint64_t* lb_status, uint32_t* ticket,
                           int32_t* total_out, uint64_t cap, uint32_t* status_word, cudaStream_t st);
size_t scan_status_words(uint32_t n);

// ---- sort.cu ----
struct SortScratch {
    void* keys_tmp;        // max_n keys (u32 or u64)
    uint32_t* vals_tmp;    // max_n
    uint32_t* hist;        // passes * 256
    uint32_t* lb;          // passes * tiles * 256 status words
    uint32_t* tickets;     // passes
    bool pre_zeroed;       // hist, lb and tickets were zero-filled by the caller (one launch for a whole forward)
};
size_t sort_tiles(uint32_t max_n, int key_bytes);
// Zeroes what the sort needs (hist, lb, tickets) itself.  Result lands in keys_out/vals_out.
void launch_radix_sort_u32(const uint32_t* keys, const uint32_t* vals, const int32_t* n_dev, uint32_t max_n,
                           uint32_t bits, uint32_t* keys_out, uint32_t* vals_out, const SortScratch& s,
                           cudaStream_t st);
void launch_radix_sort_u64(const uint64_t* keys, const uint32_t* vals, const int32_t* n_dev, uint32_t max_n,
                           uint32_t bits, uint64_t* keys_out, uint32_t* vals_out, const SortScratch& s,
                           cudaStream_t st);

// ---- raster.cu ----
void launch_tile_bins(const uint32_t* sorted_tile_ids, const uint64_t* sorted_keys64, const int32_t* n_isect_dev,
                      uint64_t capacity, uint32_t num_tiles, uint32_t* tile_bins, cudaStream_t st);
void launch_rasterize(const Cam& cam, const uint32_t* compact_gid_from_isect, const uint32_t* tile_bins,
                      const float* projected, int raster_u32, void* out_img, uint32_t* final_index,
                      unsigned long long* pair_counter /* diagnostic, normally null */, cudaStream_t st);
void launch_rasterize_backward(const Cam& cam, const uint32_t* compact_gid_from_isect, const uint32_t* tile_bins,
                               const float* projected, const uint32_t* global_from_compact,
                               const uint32_t* final_index, const float* out_img, const float* v_out_img,
                               float* v_projected /* [N,VP_STRIDE] by GLOBAL id, zeroed */,
                               unsigned long long* pair_counter /* diagnostic, normally null */, cudaStream_t st);

// ---- peaks.cu: FP32 / MUFU micro-kernels, the measured roofline denominators of the rasterizers ----
int measure_peaks(float* out4, float* sink, cudaStream_t st);

// ---- backward.cu ----
void launch_gather_project_backward(const Cam& cam, uint32_t N, uint32_t K, const float* means,
                                    const float* log_scales, const float* quats, const float* raw_opac,
                                    const float* v_projected, float* v_means, float* v_xy, float* v_log_scales,
                                    float* v_quats, float* v_coeffs, float* v_raw_opac,
                                    float* colour_rows /* hybrid step: (v_r, v_g, v_b, visible) per splat instead of v_coeffs; else null */,
                                    const uint8_t* visible_mask, cudaStream_t st);

// ---- programmatic dependent launch (PDL) ----
// A training step is ~22 short kernels on one stream; with plain launches each boundary costs a full drain of the
// GPU, the launch latency and the ramp-up of the next grid (13 boundaries between the two rasterizers alone).
// Every kernel of the step starts with PDL_ENTER(): griddepcontrol.wait (all memory of the preceding grid is
// visible; a no-op for a plain launch) followed by griddepcontrol.launch_dependents (the next grid on the stream
// may be scheduled as soon as every CTA of this one has got this far), and is launched through launch_k with the
// programmatic-stream-serialization attribute: the next kernel's CTAs become resident while this kernel's last
// wave drains and fall through their wait the moment it retires.  Nothing before PDL_ENTER() touches global
// memory, so read-after-write and write-after-read order between kernels is exactly the stream order.
// BRUSH_PDL=0 in the environment turns the attribute off (A/B measurements).
bool pdl_enabled();
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
#define PDL_ENTER()              \
    do {                         \
        pdl_wait();              \
        pdl_launch_dependents(); \
    } while (0)

template <typename... KArgs, typename... Args>
static inline cudaError_t launch_k(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                                   Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// The one Adam update every optimizer kernel uses (adam_kernel, dp_adam_kernel, backward_adam_kernel), so the three
// step paths give the same bits on the same gradients.  == AdamState::step of Burn 0.16 as brush-train calls it
// (train.rs:318-359): bias-corrected moments, p -= lr * m_hat / (sqrt(v_hat) + eps).  div.approx / sqrt.approx
// (<= 2 ulp / 1 ulp): the IEEE sequences take their slow path whenever v is tiny -- every splat a view does not
// see -- and WGSL's own `/` and `sqrt` are not correctly rounded either (2.5 ulp).
__device__ __forceinline__ float adam_update_value(float p, float g, float& m, float& v, float lr, float beta1,
                                                   float beta2, float eps, float inv_bc1, float inv_bc2,
                                                   float grad_scale) {
    // explicit roundings: the compiler may not contract these differently in different kernels
    g = __fmul_rn(g, grad_scale);
    m = __fmaf_rn(beta1, m, __fmul_rn(1.0f - beta1, g));
    v = __fmaf_rn(beta2, v, __fmul_rn(__fmul_rn(1.0f - beta2, g), g));
    float root;
    asm("sqrt.approx.f32 %0, %1;" : "=f"(root) : "f"(__fmul_rn(v, inv_bc2)));
    return __fmaf_rn(-lr, __fdividef(__fmul_rn(m, inv_bc1), __fadd_rn(root, eps)), p);
}
#endif

// Fused single-device backward tail (backward.cu): parameters are updated in place.
struct FusedAdamLaunch {
    float *means, *log_scales, *rotation, *sh, *raw_opac;
    float *m_means, *v_means, *m_scales, *v_scales, *m_rot, *v_rot, *m_sh, *v_sh, *m_opac, *v_opac;
    float lr_mean, lr_scale, lr_rot, lr_sh_dc, lr_sh_rest, lr_opac;
    float beta1, beta2, eps, inv_bc1, inv_bc2, grad_scale;
};
// visible_mask: u8[N] written by project_forward; grad_norm_accum / visible_accum (both or neither): the refine
// statistics of train.rs:284-316 are accumulated in the same pass.  v_xy may be null.
void launch_backward_adam(const Cam& cam, uint32_t N, uint32_t K, const float* norm_quats, const float* v_projected,
                          const FusedAdamLaunch& A, float* v_xy, const uint8_t* visible_mask, float* grad_norm_accum,
                          float* visible_accum, cudaStream_t st);
void launch_visible_mask_from_list(const uint32_t* global_from_compact, const int32_t* num_visible_dev, uint32_t N,
                                   uint8_t* visible_mask, cudaStream_t st);

// View-parallel step without a gradient exchange (backward.cu): every rank publishes the nine per-splat sums of ITS
// view (48 B per splat instead of 252 B of dense gradients), then redoes the per-splat backward for every view and
// applies the same Adam update to its full replica.  rows: f32[N * VP_STRIDE + VIEW_TRAILER]; a row's word 9 is the
// view's visibility flag; the trailer holds the view's RenderUniforms (28 words) and its loss term (word 28).
constexpr int DP_MAX_WORLD = BRUSH_DP_MAX_WORLD;
constexpr int VIEW_TRAILER = 32;
constexpr uint32_t RGB_STRIDE = 4;  // hybrid step: colour rows (v_r, v_g, v_b, visible), same trailer behind them
struct ViewsLaunch {
    int world, rank;
    const float* rows[DP_MAX_WORLD];
    uint32_t* flags[DP_MAX_WORLD];
    uint32_t epoch;
    float* loss_out;
    float* grad_norm_accum;
    float* visible_accum;
};
void launch_publish_view(float* rows, uint32_t N, const uint8_t* visible_mask, const void* uniforms_dev,
                         const float* loss_dev, const ViewsLaunch& X, cudaStream_t st);
int launch_views_adam(const ViewsLaunch& X, uint32_t N, uint32_t K, const float* norm_quats, const FusedAdamLaunch& A,
                      cudaStream_t st);
// Hybrid step: publish the trailer of this rank's colour rows + signal; then rebuild the SH gradient of every view from
// the colour rows and apply Adam to the local SH replica.
void launch_publish_colour_rows(float* rows, uint32_t N, const void* uniforms_dev, const float* loss_dev,
                                const ViewsLaunch& X, cudaStream_t st);
int launch_sh_views_adam(const ViewsLaunch& X, uint32_t N, uint32_t K, const float* means, float* sh, float* m_sh,
                         float* v_sh, float lr_sh_dc, float lr_sh_rest, float beta1, float beta2, float eps,
                         float inv_bc1, float inv_bc2, float grad_scale, cudaStream_t st);

// ---- optim.cu ----
// Zero-fill by a kernel on `st` (16-byte granular when aligned).  Used instead of cudaMemsetAsync everywhere in the
// step: driver memsets and small copies may be routed through a copy engine, where they queue behind a concurrent
// host->device upload of the next training view on another stream and stall the compute stream.
cudaError_t zero_async(void* p, size_t bytes, cudaStream_t st);
// Up to six 4-byte-aligned regions in ONE launch, plus (optionally) up to 32 words copied from the launch
// parameters to words_dst: a forward's RenderUniforms ride along with its zero-fills (render.rs:102-124).
constexpr int ZERO_MAX_SEGS = 6;
struct ZeroSegs {
    void* p[ZERO_MAX_SEGS];
    size_t bytes[ZERO_MAX_SEGS];
    uint32_t* words_dst;
    uint32_t n_words;
    uint32_t words[32];
};
cudaError_t zero_multi_async(const ZeroSegs& z, cudaStream_t st);
extern std::atomic<uint64_t> g_launch_count;  // kernels launched by this library (bench.py's gpu_launches)
struct AdamSeg {
    float* param;
    const float* grad;
    float* m;
    float* v;
    uint64_t count;
    float lr;
    uint32_t sh_k;
    float sh_rest_scale;
    uint32_t pad;
};
struct AdamLaunch {
    AdamSeg seg[8];
    int n_groups;
    float beta1, beta2, eps, inv_bc1, inv_bc2, grad_scale;
};
void launch_l1_loss_backward(const float* pred, const float* gt, uint32_t P, uint32_t C, float* v_out,
                             float* loss_out, cudaStream_t st);
void launch_adam(const AdamLaunch& L, cudaStream_t st);
// ---- ssim.cu ---- scratch: sums = 64 zeroable bytes, maps = f32[9 * (H+2) * (W+2)]
void launch_image_loss_backward(const float* pred, const float* gt, uint32_t H, uint32_t W, uint32_t C,
                                float ssim_weight, float* maps, float* sums, float* v_out, float* loss_out,
                                cudaStream_t st);
void launch_image_u8_to_f32(const unsigned char* src, size_t n, float* dst, cudaStream_t st);
void launch_quat_normalize(const float* q, uint32_t n, float* out, cudaStream_t st);
void launch_quat_normalize_vjp(const float* q, uint32_t n, float* v_inout, cudaStream_t st);
void launch_xy_grad_stats(const float* v_xy, const uint8_t* visible_mask, uint32_t n, float half_w, float half_h,
                          int accumulate, float* grad_norm, float* visible, cudaStream_t st);


// ---- refine.cu: densify / prune (train.rs:395-553) ----
struct RefineLaunchCfg {
    float grad_thresh, log_size_thresh;
    double logit_alpha;
    float log_scale_cull;
    int reset_alpha;
    float reset_alpha_raw;
};
size_t refine_scratch_bytes(uint32_t N);
// totals_dev: u32[6] = rows kept of (post-step splats, clones, split copies), split candidates, rows failing the
// opacity test, rows failing the scale test after passing the opacity test
void launch_refine_plan(uint32_t N, const float* grad_accum, const float* grad_counts, const float* pre_log_scales,
                        const float* pre_opac, const float* post_log_scales, const float* post_opac,
                        const RefineLaunchCfg& c, void* scratch, uint32_t** totals_dev, cudaStream_t st);
// pre / post / out: means, rotation, sh, raw_opacity, log_scales
void launch_refine_apply(uint32_t N, uint32_t K, const float* const pre[5], const float* const post[5],
                         const float* samples, const RefineLaunchCfg& c, void* scratch, float* const out[5],
                         cudaStream_t st);

// ---- dp_adam.cu: view-parallel optimizer step over NVLink peer memory ----
struct DpSeg {
    uint64_t src_off, dst_off, count;  // floats; multiples of 4
    int adam;                          // 1: Adam on the summed gradient; 0: plain sum copied to dst
    float lr;
    uint32_t sh_k;
    float sh_rest_scale;
};
struct DpLaunch {
    int world, rank, n_segs;
    const float* src[DP_MAX_WORLD];
    float* dst[DP_MAX_WORLD];
    const float* src_mc;
    float* dst_mc;
    uint32_t* flags[DP_MAX_WORLD];
    float* m;
    float* v;
    uint64_t loss_off;
    float* loss_out;
    uint32_t* arrive;
    uint32_t epoch;
    int lane;         // flag words [lane * 2 * DP_MAX_WORLD, ...) of every rank's flag block
    int ctas_per_sm;  // 0 = default
    float beta1, beta2, eps, inv_bc1, inv_bc2, grad_scale;
    DpSeg seg[8];
};
void launch_dp_adam(const DpLaunch& L, cudaStream_t st);

}  // namespace brush
