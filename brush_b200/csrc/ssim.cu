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
constexpr int SS_TX = 32, SS_TY = 16;            // output tile (x, y) of one 256-thread CTA
constexpr int SS_INX = SS_TX + SS_WIN - 1;       // 42 input columns
constexpr int SS_INY = SS_TY + SS_WIN - 1;       // 26 input rows
constexpr int SS_HITEMS = SS_INY * (SS_TX / 4);  // horizontal pass: one thread = 4 adjacent outputs of one row

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
    __shared__ float s_x[SS_INY][SS_INX + 1], s_y[SS_INY][SS_INX + 1];
    __shared__ float s_h[5][SS_INY][SS_TX + 1];
    __shared__ float s_red[8];
    const int MH = H + 2, MW = W + 2;
    const int q0x = blockIdx.x * SS_TX, q0y = blockIdx.y * SS_TY;
    const int tx = threadIdx.x & 31, typ = threadIdx.x >> 5;  // vertical pass: outputs (tx, 2 typ) and (tx, 2 typ + 1)
    const float c1 = 0.01f * 0.01f, c2 = 0.03f * 0.03f;
    float s_acc = 0.f;
    for (int c = 0; c < 3; c++) {
        __syncthreads();
        // input pixel of map position q and tap k: q + k - PAD
        for (int i = threadIdx.x; i < SS_INY * SS_INX; i += 256) {
            const int ly = i / SS_INX, lx = i - ly * SS_INX;
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
        // horizontal pass, register-tiled: 14 loads feed 4 outputs x 11 taps (the warp's addresses are
        // conflict-free: row stride 43 words, group stride 4)
        if (threadIdx.x < SS_HITEMS) {
            const int ly = threadIdx.x >> 3, ox0 = (threadIdx.x & 7) * 4;
            float xr[SS_WIN + 3], yr[SS_WIN + 3];
#pragma unroll
            for (int j = 0; j < SS_WIN + 3; j++) { xr[j] = s_x[ly][ox0 + j]; yr[j] = s_y[ly][ox0 + j]; }
#pragma unroll
            for (int o = 0; o < 4; o++) {
                float a = 0.f, b = 0.f, aa = 0.f, bb = 0.f, ab = 0.f;
#pragma unroll
                for (int k = 0; k < SS_WIN; k++) {
                    const float xv = xr[o + k], yv = yr[o + k], w = win.w[k];
                    const float wx = w * xv, wy = w * yv;
                    a += wx; b += wy; aa = fmaf(wx, xv, aa); bb = fmaf(wy, yv, bb); ab = fmaf(wx, yv, ab);
                }
                s_h[0][ly][ox0 + o] = a; s_h[1][ly][ox0 + o] = b; s_h[2][ly][ox0 + o] = aa;
                s_h[3][ly][ox0 + o] = bb; s_h[4][ly][ox0 + o] = ab;
            }
        }
        __syncthreads();
        // vertical pass: 12 loads per quantity feed 2 outputs
        float acc[2][5];
#pragma unroll
        for (int o = 0; o < 2; o++)
#pragma unroll
            for (int m = 0; m < 5; m++) acc[o][m] = 0.f;
#pragma unroll
        for (int m = 0; m < 5; m++) {
            float col[SS_WIN + 1];
#pragma unroll
            for (int j = 0; j < SS_WIN + 1; j++) col[j] = s_h[m][2 * typ + j][tx];
#pragma unroll
            for (int k = 0; k < SS_WIN; k++) {
                acc[0][m] = fmaf(win.w[k], col[k], acc[0][m]);
                acc[1][m] = fmaf(win.w[k], col[k + 1], acc[1][m]);
            }
        }
#pragma unroll
        for (int o = 0; o < 2; o++) {
        const float mu_x = acc[o][0], mu_y = acc[o][1], exx = acc[o][2], eyy = acc[o][3], exy = acc[o][4];
        const int ty = 2 * typ + o;
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
    __shared__ float s_m[3][SS_INY][SS_INX + 1];
    __shared__ float s_h[3][SS_INY][SS_TX + 1];
    __shared__ float s_red[8];
    const int MH = H + 2, MW = W + 2;
    const int p0x = blockIdx.x * SS_TX, p0y = blockIdx.y * SS_TY;
    const int tx = threadIdx.x & 31, typ = threadIdx.x >> 5;  // pixels (tx, 2 typ) and (tx, 2 typ + 1) of the tile
    const int px = p0x + tx;
    const size_t plane = (size_t)MH * MW;
    bool inside[2];
    float4 x4[2];
    float y4[2][4];
    float g_ssim[2][3];
#pragma unroll
    for (int o = 0; o < 2; o++) {
        const int py = p0y + 2 * typ + o;
        inside[o] = px < W && py < H;
        x4[o] = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int c = 0; c < 4; c++) y4[o][c] = 0.f;
#pragma unroll
        for (int c = 0; c < 3; c++) g_ssim[o][c] = 0.f;
        if (inside[o]) {
            const size_t p = (size_t)py * W + px;
            x4[o] = __ldg(reinterpret_cast<const float4*>(pred) + p);
            y4[o][0] = __ldg(gt + p * C); y4[o][1] = __ldg(gt + p * C + 1); y4[o][2] = __ldg(gt + p * C + 2);
            if (C == 4) y4[o][3] = __ldg(gt + p * C + 3);
        }
    }
    if (ssim_weight > 0.f) {
#pragma unroll
        for (int c = 0; c < 3; c++) {
            __syncthreads();
            // pixel p feeds map positions q in [p-4, p+6]; tap t = q - (p - 4) carries window weight w[10 - t]
            for (int i = threadIdx.x; i < SS_INY * SS_INX; i += 256) {
                const int ly = i / SS_INX, lx = i - ly * SS_INX;
                const int qy = p0y + ly - (SS_WIN - 1 - SS_PAD), qx = p0x + lx - (SS_WIN - 1 - SS_PAD);
                const bool ok = qy >= 0 && qy < MH && qx >= 0 && qx < MW;
                const size_t q = ok ? (size_t)qy * MW + qx : 0;
#pragma unroll
                for (int m = 0; m < 3; m++) s_m[m][ly][lx] = ok ? __ldg(maps + (size_t)(c * 3 + m) * plane + q) : 0.f;
            }
            __syncthreads();
            if (threadIdx.x < SS_HITEMS) {  // horizontal pass, 4 outputs per thread
                const int ly = threadIdx.x >> 3, ox0 = (threadIdx.x & 7) * 4;
#pragma unroll
                for (int m = 0; m < 3; m++) {
                    float r[SS_WIN + 3];
#pragma unroll
                    for (int j = 0; j < SS_WIN + 3; j++) r[j] = s_m[m][ly][ox0 + j];
#pragma unroll
                    for (int o = 0; o < 4; o++) {
                        float a = 0.f;
#pragma unroll
                        for (int t = 0; t < SS_WIN; t++) a = fmaf(win.w[SS_WIN - 1 - t], r[o + t], a);
                        s_h[m][ly][ox0 + o] = a;
                    }
                }
            }
            __syncthreads();
            float acc[2][3];
#pragma unroll
            for (int m = 0; m < 3; m++) {  // vertical pass, 2 outputs per thread
                float col[SS_WIN + 1];
#pragma unroll
                for (int j = 0; j < SS_WIN + 1; j++) col[j] = s_h[m][2 * typ + j][tx];
                acc[0][m] = acc[1][m] = 0.f;
#pragma unroll
                for (int t = 0; t < SS_WIN; t++) {
                    acc[0][m] = fmaf(win.w[SS_WIN - 1 - t], col[t], acc[0][m]);
                    acc[1][m] = fmaf(win.w[SS_WIN - 1 - t], col[t + 1], acc[1][m]);
                }
            }
#pragma unroll
            for (int o = 0; o < 2; o++) {
                const float xc = c == 0 ? x4[o].x : (c == 1 ? x4[o].y : x4[o].z);
                g_ssim[o][c] = acc[o][0] + 2.f * xc * acc[o][1] + y4[o][c] * acc[o][2];
            }
        }
    }
    // L1 term (train.rs:249-255): compare RGB, or RGBA when the ground truth has alpha
    const float inv_l1 = 1.0f / ((float)H * (float)W * (float)C);
    const float k_l1 = (1.0f - ssim_weight) * inv_l1;
    const float k_ss = -ssim_weight / (3.0f * (float)MH * (float)MW);
    float l1 = 0.f;
#pragma unroll
    for (int o = 0; o < 2; o++) {
        if (!inside[o]) continue;
        const size_t p = (size_t)(p0y + 2 * typ + o) * W + px;
        const float d0 = x4[o].x - y4[o][0], d1 = x4[o].y - y4[o][1], d2 = x4[o].z - y4[o][2];
        const float d3 = C == 4 ? x4[o].w - y4[o][3] : 0.f;
        auto sgn = [](float d) { return d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f); };
        float4 v;
        v.x = k_l1 * sgn(d0) + k_ss * g_ssim[o][0];
        v.y = k_l1 * sgn(d1) + k_ss * g_ssim[o][1];
        v.z = k_l1 * sgn(d2) + k_ss * g_ssim[o][2];
        v.w = k_l1 * sgn(d3);
        reinterpret_cast<float4*>(v_out)[p] = v;
        l1 += fabsf(d0) + fabsf(d1) + fabsf(d2) + fabsf(d3);
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
    zero_async(sums, 32, st);
    if (ssim_weight > 0.f) {
        dim3 grid((W + 2 + SS_TX - 1) / SS_TX, (H + 2 + SS_TY - 1) / SS_TY);
        g_launch_count++;
        ssim_stats_kernel<<<grid, 256, 0, st>>>(pred, gt, (int)H, (int)W, (int)C, win, maps, sums);
    }
    dim3 grid((W + SS_TX - 1) / SS_TX, (H + SS_TY - 1) / SS_TY);
    g_launch_count++;
    image_loss_backward_kernel<<<grid, 256, 0, st>>>(pred, gt, (int)H, (int)W, (int)C, ssim_weight, win, maps, sums,
                                                      reinterpret_cast<uint32_t*>(sums + 4), v_out, loss_out);
}

}  // namespace brush
