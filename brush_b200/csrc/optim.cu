// Train-step kernels either side of the render path: the L1 image loss that produces
// v_out_img directly, and a fused multi-tensor Adam over the five parameter groups.
//   loss  == (pred_compare - gt).abs().mean()                  crates/brush-train/src/train.rs:255
//   Adam  == AdamConfig::new().with_epsilon(1e-15) (Burn 0.16) crates/brush-train/src/train.rs:184,318-359
#include <cstdlib>

#include "internal.h"

namespace brush {

std::atomic<uint64_t> g_launch_count{0};

bool pdl_enabled() {
    static const bool on = [] {
        const char* e = getenv("BRUSH_PDL");
        return !(e && e[0] == '0');
    }();
    return on;
}

// pred f32[P,4]; gt f32[P,C] (C = 3: RGB compare, C = 4: RGBA compare).  v_out[P,4] = d mean|pred-gt| / d pred.
// loss_out[0] += sum |pred-gt| / (P*C)   (must be zeroed by the caller; one atomic per CTA)
__global__ void __launch_bounds__(256)
l1_loss_backward_kernel(const float4* __restrict__ pred, const float* __restrict__ gt, uint32_t P, uint32_t Cc,
                        float4* __restrict__ v_out, float* __restrict__ loss_out) {
    __shared__ float s_part[8];
    const float inv = 1.0f / ((float)P * (float)Cc);
    float acc = 0.f;
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < P; i += gridDim.x * blockDim.x) {
        const float4 p = pred[i];
        float d[4] = {0.f, 0.f, 0.f, 0.f};
        if (Cc == 4) {
            const float4 g = __ldg(reinterpret_cast<const float4*>(gt) + i);
            d[0] = p.x - g.x; d[1] = p.y - g.y; d[2] = p.z - g.z; d[3] = p.w - g.w;
        } else {
            d[0] = p.x - __ldg(gt + 3 * (size_t)i);
            d[1] = p.y - __ldg(gt + 3 * (size_t)i + 1);
            d[2] = p.z - __ldg(gt + 3 * (size_t)i + 2);
        }
        float4 v;
        v.x = d[0] > 0.f ? inv : (d[0] < 0.f ? -inv : 0.f);
        v.y = d[1] > 0.f ? inv : (d[1] < 0.f ? -inv : 0.f);
        v.z = d[2] > 0.f ? inv : (d[2] < 0.f ? -inv : 0.f);
        v.w = Cc == 4 ? (d[3] > 0.f ? inv : (d[3] < 0.f ? -inv : 0.f)) : 0.f;
        v_out[i] = v;
        acc += fabsf(d[0]) + fabsf(d[1]) + fabsf(d[2]) + fabsf(d[3]);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x < 8) {
        float v = s_part[threadIdx.x];
#pragma unroll
        for (int o = 4; o > 0; o >>= 1) v += __shfl_xor_sync(0xffu, v, o);
        if (threadIdx.x == 0) atomicAdd(loss_out, v * inv);
    }
}

void launch_l1_loss_backward(const float* pred, const float* gt, uint32_t P, uint32_t C, float* v_out,
                             float* loss_out, cudaStream_t st) {
    if (P == 0) return;
    uint32_t blocks = (P + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    g_launch_count++;
    l1_loss_backward_kernel<<<blocks, 256, 0, st>>>(reinterpret_cast<const float4*>(pred), gt, P, C,
                                                    reinterpret_cast<float4*>(v_out), loss_out);
}

// One launch for all parameter groups.  Element e of group s: lr = s.lr, except SH coefficients k > 0 which
// use lr * sh_rest_scale (the reference's 1/lr_coeffs_sh_scale lerp, train.rs:334-348; Adam's update is linear
// in lr so the lerp is exactly a smaller lr).
__device__ __forceinline__ void adam_one(float& p, float g, float& m, float& v, float lr, const AdamLaunch& L) {
    p = adam_update_value(p, g, m, v, lr, L.beta1, L.beta2, L.eps, L.inv_bc1, L.inv_bc2, L.grad_scale);
}

__global__ void __launch_bounds__(256)
adam_kernel(const __grid_constant__ AdamLaunch L) {
    const uint64_t gid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (int s = 0; s < L.n_groups; s++) {
        const AdamSeg& g = L.seg[s];
        const uint32_t row = g.sh_k > 1 ? 3u * g.sh_k : 0u;  // SH group: elements 3..row-1 of each row use the "rest" lr
        const float lr_rest = g.lr * g.sh_rest_scale;
        const bool vec = (((uintptr_t)g.param | (uintptr_t)g.grad | (uintptr_t)g.m | (uintptr_t)g.v) & 15u) == 0;
        const uint64_t n4 = vec ? g.count / 4 : 0;
        // 16-byte path: 7 x 16 B of traffic per thread per iteration, fully coalesced
        for (uint64_t j = gid; j < n4; j += stride) {
            float4 p = reinterpret_cast<float4*>(g.param)[j];
            const float4 gr = __ldcs(reinterpret_cast<const float4*>(g.grad) + j);  // read once: stream past L2
            float4 m = reinterpret_cast<float4*>(g.m)[j];
            float4 v = reinterpret_cast<float4*>(g.v)[j];
            float lr0 = g.lr, lr1 = g.lr, lr2 = g.lr, lr3 = g.lr;
            if (row) {  // row >= 12, so r + i wraps at most once
                const uint32_t r = (uint32_t)((4 * j) % row);
                auto lr_of = [&](uint32_t e) { return (e >= row ? e - row : e) >= 3u ? lr_rest : g.lr; };
                lr0 = lr_of(r); lr1 = lr_of(r + 1u); lr2 = lr_of(r + 2u); lr3 = lr_of(r + 3u);
            }
            adam_one(p.x, gr.x, m.x, v.x, lr0, L);
            adam_one(p.y, gr.y, m.y, v.y, lr1, L);
            adam_one(p.z, gr.z, m.z, v.z, lr2, L);
            adam_one(p.w, gr.w, m.w, v.w, lr3, L);
            reinterpret_cast<float4*>(g.param)[j] = p;
            reinterpret_cast<float4*>(g.m)[j] = m;
            reinterpret_cast<float4*>(g.v)[j] = v;
        }
        for (uint64_t e = 4 * n4 + gid; e < g.count; e += stride) {
            float p = g.param[e], m = g.m[e], v = g.v[e];
            const float lr = (row && (e % row) >= 3u) ? lr_rest : g.lr;
            adam_one(p, g.grad[e], m, v, lr, L);
            g.param[e] = p; g.m[e] = m; g.v[e] = v;
        }
    }
}

void launch_adam(const AdamLaunch& L, cudaStream_t st) {
    uint64_t mx = 0;
    for (int s = 0; s < L.n_groups; s++) mx = L.seg[s].count > mx ? L.seg[s].count : mx;
    if (mx == 0) return;
    uint64_t blocks = (mx / 4 + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    if (blocks == 0) blocks = 1;
    g_launch_count++;
    adam_kernel<<<(uint32_t)blocks, 256, 0, st>>>(L);
}


// ---- rotation normalisation either side of the operator (gaussian_splats.rs:174-175; the reference leaves
// both directions to Burn's elementwise ops + autodiff) ----
__global__ void __launch_bounds__(256)
quat_normalize_kernel(const float4* __restrict__ q, uint32_t n, float4* __restrict__ out) {
    PDL_ENTER();
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 v = q[i];
    const float inv = 1.0f / sqrtf(v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w);
    out[i] = make_float4(v.x * inv, v.y * inv, v.z * inv, v.w * inv);
}
// v_q = (v_qhat - qhat * dot(qhat, v_qhat)) / |q|, in place on the gradient
__global__ void __launch_bounds__(256)
quat_normalize_vjp_kernel(const float4* __restrict__ q, uint32_t n, float4* __restrict__ v_inout) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 a = q[i];
    const float inv = 1.0f / sqrtf(a.x * a.x + a.y * a.y + a.z * a.z + a.w * a.w);
    const float4 h = make_float4(a.x * inv, a.y * inv, a.z * inv, a.w * inv);
    const float4 g = v_inout[i];
    const float d = h.x * g.x + h.y * g.y + h.z * g.z + h.w * g.w;
    v_inout[i] = make_float4((g.x - h.x * d) * inv, (g.y - h.y * d) * inv, (g.z - h.z * d) * inv,
                             (g.w - h.w * d) * inv);
}
// Refine statistics of one view (train.rs:284-316): grad_norm[g] (+)= |v_xy[g] * (w/2, h/2)|,
// visible[g] (+)= 1 for the view's visible splats.  accumulate = 0 overwrites (first view of a batch).
__global__ void __launch_bounds__(256)
xy_grad_stats_kernel(const float2* __restrict__ v_xy, const uint8_t* __restrict__ visible_mask, uint32_t n,
                     float half_w, float half_h, int accumulate, float* __restrict__ grad_norm,
                     float* __restrict__ visible) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float2 g = v_xy[i];
    const float gx = g.x * half_w, gy = g.y * half_h;
    const float nrm = sqrtf(gx * gx + gy * gy);
    const float vis = (float)visible_mask[i];
    grad_norm[i] = accumulate ? grad_norm[i] + nrm : nrm;
    visible[i] = accumulate ? visible[i] + vis : vis;
}

// == image_to_tensor's u8 -> f32 / 255 (crates/brush-dataset/src/lib.rs), done after the upload instead of before
// it: the host keeps (and PCIe carries) one byte per channel.  Four channels' worth per thread.
__global__ void __launch_bounds__(256)
image_u8_to_f32_kernel(const uchar4* __restrict__ src, size_t n4, float4* __restrict__ dst,
                       const unsigned char* __restrict__ src1, float* __restrict__ dst1, uint32_t ntail) {
    const size_t stride = (size_t)gridDim.x * 256u;
    for (size_t i = (size_t)blockIdx.x * 256u + threadIdx.x; i < n4; i += stride) {
        const uchar4 v = src[i];
        dst[i] = make_float4((float)v.x / 255.0f, (float)v.y / 255.0f, (float)v.z / 255.0f, (float)v.w / 255.0f);
    }
    if (blockIdx.x == 0 && threadIdx.x < ntail) dst1[threadIdx.x] = (float)src1[threadIdx.x] / 255.0f;
}
void launch_image_u8_to_f32(const unsigned char* src, size_t n, float* dst, cudaStream_t st) {
    if (!n) return;
    const size_t n4 = n / 4;
    size_t blocks = (n4 + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    if (blocks == 0) blocks = 1;
    g_launch_count++;
    image_u8_to_f32_kernel<<<(uint32_t)blocks, 256, 0, st>>>((const uchar4*)src, n4, (float4*)dst, src + 4 * n4,
                                                             dst + 4 * n4, (uint32_t)(n - 4 * n4));
}

void launch_quat_normalize(const float* q, uint32_t n, float* out, cudaStream_t st) {
    if (!n) return;
    g_launch_count++;
    launch_k(quat_normalize_kernel, (n + 255) / 256, 256, 0, st, (const float4*)q, n, (float4*)out);
}
void launch_quat_normalize_vjp(const float* q, uint32_t n, float* v_inout, cudaStream_t st) {
    if (!n) return;
    g_launch_count++;
    quat_normalize_vjp_kernel<<<(n + 255) / 256, 256, 0, st>>>((const float4*)q, n, (float4*)v_inout);
}
void launch_xy_grad_stats(const float* v_xy, const uint8_t* visible_mask, uint32_t n, float half_w, float half_h,
                          int accumulate, float* grad_norm, float* visible, cudaStream_t st) {
    if (!n) return;
    g_launch_count++;
    xy_grad_stats_kernel<<<(n + 255) / 256, 256, 0, st>>>((const float2*)v_xy, visible_mask, n, half_w, half_h,
                                                          accumulate, grad_norm, visible);
}

__global__ void __launch_bounds__(256)
zero_fill_kernel(uint4* __restrict__ p16, size_t n16, unsigned char* __restrict__ tail, uint32_t ntail) {
    PDL_ENTER();
    const size_t stride = (size_t)gridDim.x * 256u;
    for (size_t i = (size_t)blockIdx.x * 256u + threadIdx.x; i < n16; i += stride) p16[i] = make_uint4(0u, 0u, 0u, 0u);
    if (blockIdx.x == 0 && threadIdx.x < ntail) tail[threadIdx.x] = 0;
}
cudaError_t zero_async(void* p, size_t bytes, cudaStream_t st) {
    if (!bytes) return cudaSuccess;
    unsigned char* b = (unsigned char*)p;
    if (((uintptr_t)b & 15u) || bytes < 16) {  // unaligned or tiny: byte-wise through the tail path when it fits
        if (bytes <= 256) {
            g_launch_count++;
            launch_k(zero_fill_kernel, 1, 256, 0, st, (uint4*)nullptr, (size_t)0, b, (uint32_t)bytes);
            return cudaGetLastError();
        }
        return cudaMemsetAsync(p, 0, bytes, st);
    }
    const size_t n16 = bytes / 16;
    size_t blocks = (n16 + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    g_launch_count++;
    launch_k(zero_fill_kernel, (uint32_t)blocks, 256, 0, st, (uint4*)b, n16, b + n16 * 16, (uint32_t)(bytes - n16 * 16));
    return cudaGetLastError();
}

__global__ void __launch_bounds__(256)
zero_multi_kernel(const __grid_constant__ ZeroSegs z) {
    PDL_ENTER();
    const size_t stride = (size_t)gridDim.x * 256u, t0 = (size_t)blockIdx.x * 256u + threadIdx.x;
    if (z.words_dst && t0 < z.n_words) z.words_dst[t0] = z.words[t0];
#pragma unroll
    for (int s = 0; s < ZERO_MAX_SEGS; s++) {
        if (!z.bytes[s]) continue;
        const uintptr_t a = (uintptr_t)z.p[s];
        const size_t head = (16 - (a & 15u)) & 15u;  // bytes before the first 16-byte boundary (multiple of 4)
        const size_t hb = head < z.bytes[s] ? head : z.bytes[s];
        uint32_t* w = (uint32_t*)z.p[s];
        if (t0 < hb / 4) w[t0] = 0u;
        const size_t rest = z.bytes[s] - hb, n16 = rest / 16;
        uint4* p16 = (uint4*)((char*)z.p[s] + hb);
        for (size_t i = t0; i < n16; i += stride) p16[i] = make_uint4(0u, 0u, 0u, 0u);
        uint32_t* tail = (uint32_t*)(p16 + n16);
        if (t0 < (rest - n16 * 16) / 4) tail[t0] = 0u;
    }
}
cudaError_t zero_multi_async(const ZeroSegs& z, cudaStream_t st) {
    size_t mx = 0;
    for (int s = 0; s < ZERO_MAX_SEGS; s++) mx = z.bytes[s] > mx ? z.bytes[s] : mx;
    if (!mx && !z.words_dst) return cudaSuccess;
    size_t blocks = (mx / 16 + 255) / 256;
    blocks = blocks < 1 ? 1 : (blocks > 148 * 8 ? 148 * 8 : blocks);
    g_launch_count++;
    launch_k(zero_multi_kernel, (uint32_t)blocks, 256, 0, st, z);
    return cudaGetLastError();
}

}  // namespace brush
