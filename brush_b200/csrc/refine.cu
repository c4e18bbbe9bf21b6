// Densification / pruning of the splat set on the device: the arithmetic of SplatTrainer::refine_splats
//   == crates/brush-train/src/train.rs:395-553 (decisions, clone / split values), :560-596 (prune_points),
//      :598-621 (concat_splats), :140-177 (quaternion_vec_multiply), :205-209 (reset_opacity).
// The reference builds index tensors (argwhere), selects, concatenates and masks with a dozen Burn ops and two
// host round trips; here it is three launches:
//   1. refine_classify: per splat, the clone / split decision and the survival (prune tests) of each of the up to
//      three rows it yields -- the post-step splat itself, its clone (pre-step values), its split copy (smaller,
//      offset) -- as one flag byte, plus per-block counts of the four streams;
//   2. refine_scan: exclusive scan of the block counts (one CTA), totals for the caller;
//   3. refine_apply: every splat writes its surviving rows at their final positions.
// The output order is the reference's: surviving post-step splats in index order, then surviving clones in
// index order, then surviving split copies in index order (concat, then two order-preserving prunes).
// Reference quirks kept (see brush_b200/train.py:refine_splats): the shrink / offset of the split splats
// THEMSELVES lands on clones that are dropped -- only the appended copies carry them; alpha_pruned counts all rows
// failing the opacity test, scale_pruned is cumulative.
#include "internal.h"

namespace brush {

constexpr uint32_t RF_THREADS = 256;
constexpr uint8_t RF_KEEP_BASE = 1, RF_CLONE = 2, RF_KEEP_CLONE = 4, RF_SPLIT = 8, RF_KEEP_SPLIT = 16;
// per-block counters: rows kept of the three streams, split candidates (rank of the N(0, 0.5) draw), and the two
// prune statistics
constexpr int RF_COUNTERS = 6;  // keep_base, keep_clone, keep_split, split, alpha_fail, scale_fail_after_alpha

struct RefineCfg {
    float grad_thresh;       // densify_grad_thresh
    float log_size_thresh;   // log(densify_size_thresh): max(exp(ls)) < t  <=>  max(ls) < log t
    double logit_alpha;      // logit(cull_alpha_thresh): sigmoid(x) < t  <=>  x < logit t
    float log_scale_cull;    // log(cull_scale_thresh)
    int reset_alpha;
    float reset_alpha_raw;   // inverse_sigmoid(reset_alpha_value)
};

__device__ __forceinline__ float max3(float a, float b, float c) { return fmaxf(a, fmaxf(b, c)); }

__device__ __forceinline__ uint32_t block_exclusive_scan(uint32_t v, uint32_t* s_warp, uint32_t* total) {
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    uint32_t incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= (uint32_t)o) incl += t;
    }
    __syncthreads();  // s_warp may still be read from a previous call
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    uint32_t base = 0, tot = 0;
#pragma unroll
    for (uint32_t w = 0; w < RF_THREADS / 32; w++) {
        const uint32_t t = s_warp[w];
        if (w < warp) base += t;
        tot += t;
    }
    *total = tot;
    return base + incl - v;
}

// The split copy's log-scales as the reference forms them: log(exp(ls) / 1.6) (train.rs:481-489).
__device__ __forceinline__ float split_log_scale(float ls) { return logf(expf(ls) / 1.6f); }

__global__ void __launch_bounds__(RF_THREADS)
refine_classify_kernel(uint32_t N, const float* __restrict__ grad_accum, const float* __restrict__ grad_counts,
                       const float* __restrict__ pre_log_scales, const float* __restrict__ pre_opac,
                       const float* __restrict__ post_log_scales, const float* __restrict__ post_opac,
                       const RefineCfg cfg, uint8_t* __restrict__ flags, uint32_t* __restrict__ block_counts,
                       uint32_t n_blocks) {
    __shared__ uint32_t s_cnt[RF_COUNTERS];
    if (threadIdx.x < RF_COUNTERS) s_cnt[threadIdx.x] = 0;
    __syncthreads();
    const uint32_t i = blockIdx.x * RF_THREADS + threadIdx.x;
    uint32_t f = 0, alpha_fail = 0, scale_fail = 0;
    if (i < N) {
        const float grad = grad_accum[i] / fmaxf(grad_counts[i], 1.0f);  // train.rs:406
        const bool big = grad >= cfg.grad_thresh;                        // :408
        const float l0 = post_log_scales[3 * (size_t)i], l1 = post_log_scales[3 * (size_t)i + 1],
                    l2 = post_log_scales[3 * (size_t)i + 2];
        const float lmax = max3(l0, l1, l2);
        const bool small = lmax < cfg.log_size_thresh;                   // :409-413
        // prune tests (train.rs:541-553), one per row this splat yields
        const bool a_ok = !((double)post_opac[i] < cfg.logit_alpha);
        const bool s_ok = !(lmax > cfg.log_scale_cull);
        alpha_fail += !a_ok;
        scale_fail += a_ok && !s_ok;
        if (a_ok && s_ok) f |= RF_KEEP_BASE;
        if (big && small) {  // clone: pre-step values (:431-455)
            f |= RF_CLONE;
            const bool ca = !((double)pre_opac[i] < cfg.logit_alpha);
            const bool cs = !(max3(pre_log_scales[3 * (size_t)i], pre_log_scales[3 * (size_t)i + 1],
                                   pre_log_scales[3 * (size_t)i + 2]) > cfg.log_scale_cull);
            alpha_fail += !ca;
            scale_fail += ca && !cs;
            if (ca && cs) f |= RF_KEEP_CLONE;
        }
        if (big && !small) {  // split copy: post-step opacity, scales / 1.6 (:460-515)
            f |= RF_SPLIT;
            const bool cs = !(max3(split_log_scale(l0), split_log_scale(l1), split_log_scale(l2)) > cfg.log_scale_cull);
            alpha_fail += !a_ok;
            scale_fail += a_ok && !cs;
            if (a_ok && cs) f |= RF_KEEP_SPLIT;
        }
        flags[i] = (uint8_t)f;
    }
    const uint32_t v[RF_COUNTERS] = {(f & RF_KEEP_BASE) ? 1u : 0u, (f & RF_KEEP_CLONE) ? 1u : 0u,
                                     (f & RF_KEEP_SPLIT) ? 1u : 0u, (f & RF_SPLIT) ? 1u : 0u, alpha_fail, scale_fail};
#pragma unroll
    for (int c = 0; c < RF_COUNTERS; c++) {
        const uint32_t s = __reduce_add_sync(0xffffffffu, v[c]);
        if ((threadIdx.x & 31u) == 0 && s) atomicAdd(&s_cnt[c], s);
    }
    __syncthreads();
    if (threadIdx.x < RF_COUNTERS) block_counts[threadIdx.x * n_blocks + blockIdx.x] = s_cnt[threadIdx.x];
}

// One CTA: exclusive scan of each counter over the blocks (in place), totals -> totals[RF_COUNTERS].
__global__ void __launch_bounds__(RF_THREADS)
refine_scan_kernel(uint32_t* __restrict__ block_counts, uint32_t n_blocks, uint32_t* __restrict__ totals) {
    __shared__ uint32_t s_warp[RF_THREADS / 32];
    for (int c = 0; c < RF_COUNTERS; c++) {
        uint32_t* row = block_counts + (size_t)c * n_blocks;
        uint32_t carry = 0;
        for (uint32_t base = 0; base < n_blocks; base += RF_THREADS) {
            const uint32_t i = base + threadIdx.x;
            const uint32_t v = i < n_blocks ? row[i] : 0u;
            uint32_t tot;
            const uint32_t ex = block_exclusive_scan(v, s_warp, &tot);
            if (i < n_blocks) row[i] = carry + ex;
            carry += tot;
        }
        if (threadIdx.x == 0) totals[c] = carry;
    }
}

struct SplatSet {  // means [n,3], rotation [n,4], sh [n,K,3], raw_opacity [n], log_scales [n,3]
    float *means, *rotation, *sh, *opac, *log_scales;
};

__device__ __forceinline__ void copy_row(float* __restrict__ dst, const float* __restrict__ src, uint32_t n) {
    if ((n & 3u) == 0 && ((((uintptr_t)dst) | ((uintptr_t)src)) & 15u) == 0) {
        for (uint32_t j = 0; j < n / 4; j++) reinterpret_cast<float4*>(dst)[j] = reinterpret_cast<const float4*>(src)[j];
    } else {
        for (uint32_t j = 0; j < n; j++) dst[j] = src[j];
    }
}

__global__ void __launch_bounds__(RF_THREADS)
refine_apply_kernel(uint32_t N, uint32_t K, const SplatSet pre, const SplatSet post, const float* __restrict__ samples,
                    const RefineCfg cfg, const uint8_t* __restrict__ flags, const uint32_t* __restrict__ block_offsets,
                    uint32_t n_blocks, const uint32_t* __restrict__ totals, const SplatSet out) {
    __shared__ uint32_t s_warp[RF_THREADS / 32];
    const uint32_t i = blockIdx.x * RF_THREADS + threadIdx.x;
    const uint32_t f = i < N ? flags[i] : 0u;
    uint32_t tot;
    const uint32_t p_base = block_offsets[0 * n_blocks + blockIdx.x] +
                            block_exclusive_scan((f & RF_KEEP_BASE) ? 1u : 0u, s_warp, &tot);
    const uint32_t p_clone = totals[0] + block_offsets[1 * n_blocks + blockIdx.x] +
                             block_exclusive_scan((f & RF_KEEP_CLONE) ? 1u : 0u, s_warp, &tot);
    const uint32_t p_split = totals[0] + totals[1] + block_offsets[2 * n_blocks + blockIdx.x] +
                             block_exclusive_scan((f & RF_KEEP_SPLIT) ? 1u : 0u, s_warp, &tot);
    const uint32_t split_rank = block_offsets[3 * n_blocks + blockIdx.x] +
                                block_exclusive_scan((f & RF_SPLIT) ? 1u : 0u, s_warp, &tot);
    if (i >= N) return;
    const uint32_t row_sh = 3u * K;
    auto opac_of = [&](float v) { return cfg.reset_alpha ? cfg.reset_alpha_raw : v; };  // train.rs:557-560
    if (f & RF_KEEP_BASE) {
        copy_row(out.means + 3 * (size_t)p_base, post.means + 3 * (size_t)i, 3);
        copy_row(out.rotation + 4 * (size_t)p_base, post.rotation + 4 * (size_t)i, 4);
        copy_row(out.sh + (size_t)row_sh * p_base, post.sh + (size_t)row_sh * i, row_sh);
        out.opac[p_base] = opac_of(post.opac[i]);
        copy_row(out.log_scales + 3 * (size_t)p_base, post.log_scales + 3 * (size_t)i, 3);
    }
    if (f & RF_KEEP_CLONE) {
        copy_row(out.means + 3 * (size_t)p_clone, pre.means + 3 * (size_t)i, 3);
        copy_row(out.rotation + 4 * (size_t)p_clone, pre.rotation + 4 * (size_t)i, 4);
        copy_row(out.sh + (size_t)row_sh * p_clone, pre.sh + (size_t)row_sh * i, row_sh);
        out.opac[p_clone] = opac_of(pre.opac[i]);
        copy_row(out.log_scales + 3 * (size_t)p_clone, pre.log_scales + 3 * (size_t)i, 3);
    }
    if (f & RF_KEEP_SPLIT) {
        // mean = pre-step mean + R(q_post) * (sample * scale_post)  (train.rs:491-514, quaternion_vec_multiply :140-177)
        const float qw = post.rotation[4 * (size_t)i], qx = post.rotation[4 * (size_t)i + 1],
                    qy = post.rotation[4 * (size_t)i + 2], qz = post.rotation[4 * (size_t)i + 3];
        const float l0 = post.log_scales[3 * (size_t)i], l1 = post.log_scales[3 * (size_t)i + 1],
                    l2 = post.log_scales[3 * (size_t)i + 2];
        const float vx = samples[3 * (size_t)split_rank] * expf(l0), vy = samples[3 * (size_t)split_rank + 1] * expf(l1),
                    vz = samples[3 * (size_t)split_rank + 2] * expf(l2);
        const float t1 = qw * vx + qy * vz - qz * vy;
        const float t2 = qw * vy - qx * vz + qz * vx;
        const float t3 = qw * vz + qx * vy - qy * vx;
        const float t4 = qx * vx + qy * vy + qz * vz;
        float* m = out.means + 3 * (size_t)p_split;
        m[0] = pre.means[3 * (size_t)i] + (vx + (qw * t1 + qx * t4 - qy * t3 + qz * t2) * 2.0f);
        m[1] = pre.means[3 * (size_t)i + 1] + (vy + (qw * t2 - qx * t3 + qy * t4 + qz * t1) * 2.0f);
        m[2] = pre.means[3 * (size_t)i + 2] + (vz + (qw * t3 + qx * t2 - qy * t1 + qz * t4) * 2.0f);
        float* r = out.rotation + 4 * (size_t)p_split;
        r[0] = qw; r[1] = qx; r[2] = qy; r[3] = qz;
        copy_row(out.sh + (size_t)row_sh * p_split, post.sh + (size_t)row_sh * i, row_sh);
        out.opac[p_split] = opac_of(post.opac[i]);
        float* s = out.log_scales + 3 * (size_t)p_split;
        s[0] = split_log_scale(l0); s[1] = split_log_scale(l1); s[2] = split_log_scale(l2);
    }
}

uint32_t refine_blocks(uint32_t N) { return (N + RF_THREADS - 1) / RF_THREADS; }
size_t refine_scratch_bytes(uint32_t N) {  // flags u8[N] (padded), block counters u32[RF_COUNTERS][blocks], totals u32[8]
    return (((size_t)N + 15) & ~(size_t)15) + (size_t)RF_COUNTERS * refine_blocks(N) * 4 + 32;
}
static void refine_layout(void* scratch, uint32_t N, uint8_t** flags, uint32_t** counts, uint32_t** totals) {
    *flags = (uint8_t*)scratch;
    *counts = (uint32_t*)((char*)scratch + (((size_t)N + 15) & ~(size_t)15));
    *totals = *counts + (size_t)RF_COUNTERS * refine_blocks(N);
}

void launch_refine_plan(uint32_t N, const float* grad_accum, const float* grad_counts, const float* pre_log_scales,
                        const float* pre_opac, const float* post_log_scales, const float* post_opac,
                        const RefineLaunchCfg& c, void* scratch, uint32_t** totals_dev, cudaStream_t st) {
    uint8_t* flags;
    uint32_t *counts, *totals;
    refine_layout(scratch, N, &flags, &counts, &totals);
    const RefineCfg cfg{c.grad_thresh, c.log_size_thresh, c.logit_alpha, c.log_scale_cull, c.reset_alpha, c.reset_alpha_raw};
    const uint32_t nb = refine_blocks(N);
    g_launch_count += 2;
    refine_classify_kernel<<<nb, RF_THREADS, 0, st>>>(N, grad_accum, grad_counts, pre_log_scales, pre_opac,
                                                      post_log_scales, post_opac, cfg, flags, counts, nb);
    refine_scan_kernel<<<1, RF_THREADS, 0, st>>>(counts, nb, totals);
    *totals_dev = totals;
}

void launch_refine_apply(uint32_t N, uint32_t K, const float* const pre[5], const float* const post[5],
                         const float* samples, const RefineLaunchCfg& c, void* scratch, float* const out[5],
                         cudaStream_t st) {
    uint8_t* flags;
    uint32_t *counts, *totals;
    refine_layout(scratch, N, &flags, &counts, &totals);
    const RefineCfg cfg{c.grad_thresh, c.log_size_thresh, c.logit_alpha, c.log_scale_cull, c.reset_alpha, c.reset_alpha_raw};
    auto set = [](const float* const p[5]) {
        return SplatSet{(float*)p[0], (float*)p[1], (float*)p[2], (float*)p[3], (float*)p[4]};
    };
    const uint32_t nb = refine_blocks(N);
    g_launch_count++;
    refine_apply_kernel<<<nb, RF_THREADS, 0, st>>>(N, K, set(pre), set(post), samples, cfg, flags, counts, nb, totals,
                                                   set((const float* const*)out));
}

}  // namespace brush
