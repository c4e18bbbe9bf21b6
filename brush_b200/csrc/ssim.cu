// Image loss of the train step, forward and backward in two launches:
//     loss = (1 - w) * mean|pred_compare - gt|  -  w * ssim(pred_rgb, gt_rgb)
//   == crates/brush-train/src/train.rs:249-268 (loss) and crates/brush-train/src/ssim.rs:7-112 (SSIM);
//   the reference differentiates it with Burn autodiff (conv2d backward x 5 + elementwise chain).
// SSIM as the reference defines it: 11-tap Gaussian window (sigma 1.5, normalised 1-D, outer product), grouped
// conv2d with ZERO padding of ceil(11/2) = 6 on every side, so the statistic maps are (H+2) x (W+2); variances
// clamped at 0; C1 = 0.01^2, C2 = 0.03^2; mean over the 3 x (H+2) x (W+2) map.
//
// Kernel A (ssim_stats): per 16x16 tile of the map and per channel, separable window sums of x, y, x^2, y^2, xy in
//   shared memory -> S and its three partials dS/dmu_x, dS/dE[x^2], dS/dE[xy] (scratch maps), sum of S.
// Kernel B (loss_backward): d ssim / d x[p] = sum_q win(q, p) * (dmu(q) + 2 x[p] dExx(q) + y[p] dExy(q)), again
//   separable in shared memory, fused with the L1 term; writes v_out_img and the loss value.
#include "internal.h"

namespace brush {

constexpr int SS_WIN = 11;
constexpr int SS_PAD = 6;          // ssim.rs:47 window_size.div_ceil(2)
constexpr int SS_T = 16;           // tile
constexpr int SS_IN = SS_T + SS_WIN - 1;  // 26

struct SsimWin { float w[SS_WIN]; };

static SsimWin make_window() {  // ssim.rs:7-15
    SsimWin g;
    float vals[SS_WIN], sum = 0.f;
    for (int x = 0; x < SS_WIN; x++) {
        float d = (float)x - (float)(SS_WIN / 2);
        vals[x] = expf(-(d * d) / (2.0f * 1.5f * 1.5f));
        sum += vals[x];
    }
    for (int x = 0; x < SS_WIN; x++) g.w[x] = vals[x] / sum;
    return g;
}

// pred f32[H,W,4]; gt f32[H,W,C].  maps: f32[3 channels][3 partials][(H+2)*(W+2)]
__global__ void __launch_bounds__(256)
ssim_stats_kernel(const float* __restrict__ pred, const float* __restrict__ gt, int H, int W, int C,
                  const __grid_constant__ SsimWin win, float* __restrict__ maps, float* __restrict__ sums) {
    __shared__ float s_x[SS_IN][SS_IN + 1], s_y[SS_IN][SS_IN + 1];
    __shared__ float s_h[5][SS_IN][SS_T + 1];
    __shared__ float s_red[8];
    const int MH = H + 2, MW = W + 2;
    const int q0x = blockIdx.x * SS_T, q0y = blockIdx.y * SS_T;
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    const float c1 = 0.01f * 0.01f, c2 = 0.03f * 0.03f;
    float s_acc = 0.f;
    for (int c = 0; c < 3; c++) {
        __syncthreads();
        // input pixel of map position q and tap k: q + k - PAD
        for (int i = threadIdx.x; i < SS_IN * SS_IN; i += 256) {
            const int ly = i / SS_IN, lx = i - ly * SS_IN;
            const int py = q0y + ly - SS_PAD, px = q0x + lx - SS_PAD;
            float xv = 0.f, yv = 0.f;
            if (py >= 0 && py < H && px >= 0 && px < W) {
                const size_t p = (size_t)py * W + px;
                xv = __ldg(pred + p * 4 + c);
                yv = __ldg(gt + p * C + c);
            }
            s_x[ly][lx] = xv;
            s_y[ly][lx] = yv;
        }
        __syncthreads();
        for (int i = threadIdx.x; i < SS_IN * SS_T; i += 256) {  // horizontal pass
            const int ly = i >> 4, ox = i & 15;
            float a = 0.f, b = 0.f, aa = 0.f, bb = 0.f, ab = 0.f;
#pragma unroll
            for (int k = 0; k < SS_WIN; k++) {
                const float xv = s_x[ly][ox + k], yv = s_y[ly][ox + k], w = win.w[k];
                const float wx = w * xv, wy = w * yv;
                a += wx; b += wy; aa = fmaf(wx, xv, aa); bb = fmaf(wy, yv, bb); ab = fmaf(wx, yv, ab);
            }
            s_h[0][ly][ox] = a; s_h[1][ly][ox] = b; s_h[2][ly][ox] = aa; s_h[3][ly][ox] = bb; s_h[4][ly][ox] = ab;
        }
        __syncthreads();
        float mu_x = 0.f, mu_y = 0.f, exx = 0.f, eyy = 0.f, exy = 0.f;
#pragma unroll
        for (int k = 0; k < SS_WIN; k++) {  // vertical pass
            const float w = win.w[k];
            mu_x = fmaf(w, s_h[0][ty + k][tx], mu_x); mu_y = fmaf(w, s_h[1][ty + k][tx], mu_y);
            exx = fmaf(w, s_h[2][ty + k][tx], exx); eyy = fmaf(w, s_h[3][ty + k][tx], eyy);
            exy = fmaf(w, s_h[4][ty + k][tx], exy);
        }
        const int qx = q0x + tx, qy = q0y + ty;
        if (qx < MW && qy < MH) {
            // ssim.rs:63-109
            const float mu_xx = mu_x * mu_x, mu_yy = mu_y * mu_y, mu_xy = mu_x * mu_y;
            const float vx = exx - mu_xx, vy = eyy - mu_yy;
            const float gate = vx > 0.f ? 1.f : 0.f;  // clamp_min(0) of sigma_xx: gradient passes where unclamped
            const float sxx = fmaxf(vx, 0.f), syy = fmaxf(vy, 0.f), sxy = exy - mu_xy;
            const float A = 2.f * mu_xy + c1, B = 2.f * sxy + c2, Cc = mu_xx + mu_yy + c1, D = sxx + syy + c2;
            const float inv_cd = 1.0f / (Cc * D);
            const float S = A * B * inv_cd;
            s_acc += S;
            // partials of S w.r.t. mu_x (total, through A, B, C, D), E[x^2] and E[xy]
            const float dmu = 2.f * mu_y * (B - A) * inv_cd - S * 2.f * mu_x / Cc + S * 2.f * mu_x * gate / D;
            const float dexx = -S * gate / D;
            const float dexy = 2.f * A * inv_cd;
            const size_t q = (size_t)qy * MW + qx, plane = (size_t)MH * MW;
            maps[(size_t)(c * 3 + 0) * plane + q] = dmu;
            maps[(size_t)(c * 3 + 1) * plane + q] = dexx;
            maps[(size_t)(c * 3 + 2) * plane + q] = dexy;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s_acc += __shfl_xor_sync(0xffffffffu, s_acc, o);
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = s_acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        float t = 0.f;
        for (int i = 0; i < 8; i++) t += s_red[i];
        atomicAdd(sums + 0, t);
    }
}

// v_out f32[H,W,4] = d loss / d pred;  loss_out = (1-w) * L1 - w * ssim
__global__ void __launch_bounds__(256)
image_loss_backward_kernel(const float* __restrict__ pred, const float* __restrict__ gt, int H, int W, int C,
                           float ssim_weight, const __grid_constant__ SsimWin win, const float* __restrict__ maps,
                           float* __restrict__ sums, uint32_t* __restrict__ ticket, float* __restrict__ v_out,
                           float* __restrict__ loss_out) {
    __shared__ float s_m[3][SS_IN][SS_IN + 1];
    __shared__ float s_h[3][SS_IN][SS_T + 1];
    __shared__ float s_red[8];
    const int MH = H + 2, MW = W + 2;
    const int p0x = blockIdx.x * SS_T, p0y = blockIdx.y * SS_T;
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    const int px = p0x + tx, py = p0y + ty;
    const bool inside = px < W && py < H;
    const size_t p = (size_t)py * W + px, plane = (size_t)MH * MW;
    float4 x4 = make_float4(0.f, 0.f, 0.f, 0.f);
    float y4[4] = {0.f, 0.f, 0.f, 0.f};
    if (inside) {
        x4 = __ldg(reinterpret_cast<const float4*>(pred) + p);
        for (int c = 0; c < C; c++) y4[c] = __ldg(gt + p * C + c);
    }
    const float xs[3] = {x4.x, x4.y, x4.z};
    float g_ssim[3] = {0.f, 0.f, 0.f};
    if (ssim_weight > 0.f) {
        for (int c = 0; c < 3; c++) {
            __syncthreads();
            // pixel p feeds map positions q in [p-4, p+6]; tap t = q - (p - 4) carries window weight w[10 - t]
            for (int i = threadIdx.x; i < SS_IN * SS_IN; i += 256) {
                const int ly = i / SS_IN, lx = i - ly * SS_IN;
                const int qy = p0y + ly - (SS_WIN - 1 - SS_PAD), qx = p0x + lx - (SS_WIN - 1 - SS_PAD);
                const bool ok = qy >= 0 && qy < MH && qx >= 0 && qx < MW;
                const size_t q = ok ? (size_t)qy * MW + qx : 0;
#pragma unroll
                for (int m = 0; m < 3; m++) s_m[m][ly][lx] = ok ? __ldg(maps + (size_t)(c * 3 + m) * plane + q) : 0.f;
            }
            __syncthreads();
            for (int i = threadIdx.x; i < SS_IN * SS_T; i += 256) {
                const int ly = i >> 4, ox = i & 15;
                float a = 0.f, b = 0.f, d = 0.f;
#pragma unroll
                for (int t = 0; t < SS_WIN; t++) {
                    const float w = win.w[SS_WIN - 1 - t];
                    a = fmaf(w, s_m[0][ly][ox + t], a); b = fmaf(w, s_m[1][ly][ox + t], b); d = fmaf(w, s_m[2][ly][ox + t], d);
                }
                s_h[0][ly][ox] = a; s_h[1][ly][ox] = b; s_h[2][ly][ox] = d;
            }
            __syncthreads();
            float a = 0.f, b = 0.f, d = 0.f;
#pragma unroll
            for (int t = 0; t < SS_WIN; t++) {
                const float w = win.w[SS_WIN - 1 - t];
                a = fmaf(w, s_h[0][ty + t][tx], a); b = fmaf(w, s_h[1][ty + t][tx], b); d = fmaf(w, s_h[2][ty + t][tx], d);
            }
            g_ssim[c] = a + 2.f * xs[c] * b + y4[c] * d;
        }
    }
    // L1 term (train.rs:249-255): compare RGB, or RGBA when the ground truth has alpha
    const float inv_l1 = 1.0f / ((float)H * (float)W * (float)C);
    const float k_l1 = (1.0f - ssim_weight) * inv_l1;
    const float k_ss = -ssim_weight / (3.0f * (float)MH * (float)MW);
    float l1 = 0.f;
    if (inside) {
        const float d0 = x4.x - y4[0], d1 = x4.y - y4[1], d2 = x4.z - y4[2], d3 = C == 4 ? x4.w - y4[3] : 0.f;
        auto sgn = [](float d) { return d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f); };
        float4 v;
        v.x = k_l1 * sgn(d0) + k_ss * g_ssim[0];
        v.y = k_l1 * sgn(d1) + k_ss * g_ssim[1];
        v.z = k_l1 * sgn(d2) + k_ss * g_ssim[2];
        v.w = k_l1 * sgn(d3);
        reinterpret_cast<float4*>(v_out)[p] = v;
        l1 = fabsf(d0) + fabsf(d1) + fabsf(d2) + fabsf(d3);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) l1 += __shfl_xor_sync(0xffffffffu, l1, o);
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = l1;
    __syncthreads();
    if (threadIdx.x == 0) {
        float t = 0.f;
        for (int i = 0; i < 8; i++) t += s_red[i];
        atomicAdd(sums + 1, t);
        __threadfence();
        if (atomicAdd(ticket, 1u) == gridDim.x * gridDim.y - 1u) {  // last block: combine the two means
            __threadfence();
            const float ssim_mean = *(volatile float*)(sums + 0) / (3.0f * (float)MH * (float)MW);
            const float l1_mean = *(volatile float*)(sums + 1) * inv_l1;
            *loss_out = (1.0f - ssim_weight) * l1_mean - ssim_weight * ssim_mean;
            sums[2] = ssim_mean;
            sums[3] = l1_mean;
        }
    }
}

// scratch: maps f32[9 * (H+2)*(W+2)], sums f32[4] + ticket u32 (zeroed here)
void launch_image_loss_backward(const float* pred, const float* gt, uint32_t H, uint32_t W, uint32_t C,
                                float ssim_weight, float* maps, float* sums, float* v_out, float* loss_out,
                                cudaStream_t st) {
    static const SsimWin win = make_window();
    cudaMemsetAsync(sums, 0, 32, st);
    if (ssim_weight > 0.f) {
        dim3 grid((W + 2 + SS_T - 1) / SS_T, (H + 2 + SS_T - 1) / SS_T);
        g_launch_count++;
        ssim_stats_kernel<<<grid, 256, 0, st>>>(pred, gt, (int)H, (int)W, (int)C, win, maps, sums);
    }
    dim3 grid((W + SS_T - 1) / SS_T, (H + SS_T - 1) / SS_T);
    g_launch_count++;
    image_loss_backward_kernel<<<grid, 256, 0, st>>>(pred, gt, (int)H, (int)W, (int)C, ssim_weight, win, maps, sums,
                                                      reinterpret_cast<uint32_t*>(sums + 4), v_out, loss_out);
}

}  // namespace brush
