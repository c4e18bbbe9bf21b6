// Image loss of the train step, forward and backward in two launches:
//     loss = (1 - w) * mean|pred_compare - gt|  -  w * ssim(pred_rgb, gt_rgb)
//   == crates/brush-train/src/train.rs:249-268 (loss) and crates/brush-train/src/ssim.rs:7-112 (SSIM);
//   the reference differentiates it with Burn autodiff (conv2d backward x 5 + elementwise chain).
// SSIM as the reference defines it: 11-tap Gaussian window (sigma 1.5, normalised 1-D, outer product), grouped
// conv2d with ZERO padding of ceil(11/2) = 6 on every side, so the statistic maps are (H+2) x (W+2); variances
// clamped at 0; C1 = 0.01^2, C2 = 0.03^2; mean over the 3 x (H+2) x (W+2) map.
//
// Both kernels are row-streaming separable convolutions with ONE WARP per work item = (strip of 32 columns,
// segment of R rows, colour channel), no CTA-level synchronisation in the loop:
//   * the warp walks down its segment one input row at a time; the row (32 + 10 halo values per quantity) goes
//     through a per-warp shared-memory line, and every lane forms the 11-tap horizontal sums of ITS column;
//   * the vertical pass lives in registers: 11 partial sums per quantity are in flight per lane, one output row
//     is started and one completed per input row (the loop is unrolled over the 11 phases so that every slot is a
//     fixed register);
//   * R is chosen by the host so that all work items are resident at once (a single wave, no tail).
// About 2.4x fewer instructions per pixel than a tile-based separable pass (halo re-loads, index arithmetic and
// the shared-memory round trip of the intermediate all go away): 0.205 -> see DESIGN.md.
//
// Kernel A (ssim_stats): window sums of x, y, x^2, y^2, xy -> S and its three partials dS/dmu_x, dS/dE[x^2],
//   dS/dE[xy] (scratch maps), sum of S.
// Kernel B (loss_backward): d ssim / d x[p] = sum_q win(q, p) * (dmu(q) + 2 x[p] dExx(q) + y[p] dExy(q)),
//   fused with the L1 term; writes v_out_img and the loss value.
#include "internal.h"

namespace brush {

constexpr int SS_WIN = 11;
constexpr int SS_PAD = 6;           // ssim.rs:47 window_size.div_ceil(2)
constexpr int SS_STRIP = 32;        // columns per work item (one per lane)
constexpr int SS_LINE = 48;         // floats per shared-memory line (32 + 10 halo, padded)
constexpr int SS_WARPS = 4;         // work items per CTA
constexpr int SS_DEPTH = 4;         // input rows in flight per warp (cp.async ring of shared-memory lines)

// 4-byte asynchronous copy global -> shared; `ok == false` writes a zero without touching `src` (zero padding of
// the convolution and rows / columns outside the image).
// (A copy of size 0 performs no global access, so `src` may lie outside the tensor then.)
__device__ __forceinline__ void cp_async_f32(uint32_t dst, const float* src, bool ok) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(dst), "l"(src), "r"(ok ? 4 : 0) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
// Keeps a loop-invariant value in its register: the compiler otherwise re-derives shared-memory addresses from
// the special registers in every row.
__device__ __forceinline__ uint32_t pinned(uint32_t v) {
    asm volatile("mov.b32 %0, %0;" : "+r"(v));
    return v;
}
__device__ __forceinline__ float rcp_fast(float x) {
    float y;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

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

struct SsimGrid { int strips, segs, rows; };  // work items = strips * segs * 3 channels; rows per segment

// One step of the in-register vertical pass.  Q quantities, 11 slots each; at input row i (phase PH = i % 11) the
// slot started j rows ago takes tap j of the new horizontal sums h[]; slot PH is (re)started with tap 0.  After
// the call slot (PH + 1) % 11 -- started 10 rows ago -- is complete.
template <int Q, int PH, bool FLIP>
__device__ __forceinline__ void vertical_step(float (&acc)[Q][SS_WIN], const float (&h)[Q], const SsimWin& win) {
#pragma unroll
    for (int j = 0; j < SS_WIN; j++) {
        const int slot = (PH - j + SS_WIN) % SS_WIN;
        const float w = win.w[FLIP ? SS_WIN - 1 - j : j];
#pragma unroll
        for (int m = 0; m < Q; m++) acc[m][slot] = j == 0 ? w * h[m] : fmaf(w, h[m], acc[m][slot]);
    }
}

// ---- kernel A -------------------------------------------------------------------------------------------------
struct StatsRow {  // state of one ssim_stats work item
    const float* xp; const float* yp;  // the next input row to be REQUESTED, at this lane's first column
    float* mp;                         // the next map row to be written, plane 0, at this lane's column
    size_t plane; int x_step, y_step, y_halo, MW;
    int py, H;                         // image row of the next input row to be requested
    bool col_ok, col2_ok, out_ok, halo;
    uint32_t line0;                    // shared-window address of this lane's word in the warp's ring: SS_DEPTH x (x line, y line)
    uint32_t fill, use;                // ring slots of the next request / the next row to be consumed
    int lane;
    float s_acc;
};

// Requests one more input row (42 values of x and of y) into the next ring slot.  Rows are requested SS_DEPTH - 1
// ahead of their use: at 16 warps per SM one row of math does not cover a DRAM round trip.
__device__ __forceinline__ void stats_fetch(StatsRow& S) {
    const bool row_ok = (unsigned)S.py < (unsigned)S.H;
    const bool a = row_ok && S.col_ok, b = row_ok && S.col2_ok;
    const uint32_t dst = S.line0 + S.fill * (2u * SS_LINE * 4u);
    cp_async_f32(dst, S.xp, a);
    cp_async_f32(dst + SS_LINE * 4u, S.yp, a);
    if (S.halo) {
        cp_async_f32(dst + 128u, S.xp + 128, b);
        cp_async_f32(dst + SS_LINE * 4u + 128u, S.yp + S.y_halo, b);
    }
    cp_async_commit();
    S.xp += S.x_step; S.yp += S.y_step; S.py++;
    S.fill = (S.fill + 1u) & (SS_DEPTH - 1);
}

template <int PH>
__device__ __forceinline__ void stats_row(StatsRow& S, float (&acc)[5][SS_WIN], const SsimWin& win, int i, int n_in) {
    if (i >= n_in) return;
    stats_fetch(S);                 // row i + SS_DEPTH - 1 (rows past the segment are requested too: guarded, unused)
    cp_async_wait<SS_DEPTH - 1>();  // row i has landed
    __syncwarp();
    const uint32_t sx = S.line0 + S.use * (2u * SS_LINE * 4u), sy = sx + SS_LINE * 4u;
    S.use = (S.use + 1u) & (SS_DEPTH - 1);
    float h[5] = {0.f, 0.f, 0.f, 0.f, 0.f};  // horizontal sums of x, y, x^2, y^2, xy at this lane's column
#pragma unroll
    for (int k = 0; k < SS_WIN; k++) {
        float xv, yv;
        asm volatile("ld.shared.f32 %0, [%1];" : "=f"(xv) : "r"(sx + 4u * k));
        asm volatile("ld.shared.f32 %0, [%1];" : "=f"(yv) : "r"(sy + 4u * k));
        const float w = win.w[k];
        const float wx = w * xv, wy = w * yv;
        h[0] += wx; h[1] += wy; h[2] = fmaf(wx, xv, h[2]); h[3] = fmaf(wy, yv, h[3]); h[4] = fmaf(wx, yv, h[4]);
    }
    __syncwarp();  // every lane has read the slot before the next request overwrites it
    vertical_step<5, PH, false>(acc, h, win);
    if (i >= SS_WIN - 1) {  // map row q = q0 + i - 10 is complete in slot (PH + 1) % 11
        constexpr int slot = (PH + 1) % SS_WIN;
        if (S.out_ok) {
            // ssim.rs:63-109
            const float c1 = 0.01f * 0.01f, c2 = 0.03f * 0.03f;
            const float mu_x = acc[0][slot], mu_y = acc[1][slot], exx = acc[2][slot], eyy = acc[3][slot],
                        exy = acc[4][slot];
            const float mu_xx = mu_x * mu_x, mu_yy = mu_y * mu_y, mu_xy = mu_x * mu_y;
            const float vx = exx - mu_xx, vy = eyy - mu_yy;
            const float gate = vx > 0.f ? 1.f : 0.f;  // clamp_min(0) of sigma_xx: gradient passes where unclamped
            const float sxx = fmaxf(vx, 0.f), syy = fmaxf(vy, 0.f), sxy = exy - mu_xy;
            const float A = 2.f * mu_xy + c1, B = 2.f * sxy + c2, Cc = mu_xx + mu_yy + c1, D = sxx + syy + c2;
            // C >= 1e-4 and D >= 9e-4: two approximate reciprocals (1 ulp) instead of four IEEE divisions
            const float inv_c = rcp_fast(Cc), inv_d = rcp_fast(D);
            const float inv_cd = inv_c * inv_d;
            const float Sv = A * B * inv_cd;
            S.s_acc += Sv;
            // partials of S w.r.t. mu_x (total, through A, B, C, D), E[x^2] and E[xy]
            const float sd = Sv * gate * inv_d;
            const float dmu = 2.f * (mu_y * (B - A) * inv_cd + mu_x * (sd - Sv * inv_c));
            const float dexx = -sd;
            const float dexy = 2.f * A * inv_cd;
            S.mp[0] = dmu;
            S.mp[S.plane] = dexx;
            S.mp[2 * S.plane] = dexy;
        }
        S.mp += S.MW;
    }
}

// pred f32[H,W,4]; gt f32[H,W,C].  maps: f32[3 channels][3 partials][(H+2)*(W+2)]
__global__ void __launch_bounds__(32 * SS_WARPS)
ssim_stats_kernel(const float* __restrict__ pred, const float* __restrict__ gt, int H, int W, int C,
                  const __grid_constant__ SsimWin win, const SsimGrid G, float* __restrict__ maps,
                  float* __restrict__ sums) {
    __shared__ __align__(16) float s_line[SS_WARPS][SS_DEPTH][2][SS_LINE];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int item = blockIdx.x * SS_WARPS + warp;
    PDL_ENTER();
    if (item >= G.strips * G.segs * 3) return;  // whole warps only; no CTA-wide barrier below
    StatsRow S;
    const int MW = W + 2, MH = H + 2;
    const int c = item % 3;
    const int c0 = ((item / 3) % G.strips) * SS_STRIP;
    const int q0 = (item / (3 * G.strips)) * G.rows;
    const int q_end = min(q0 + G.rows, MH);
    const int px = c0 - SS_PAD + lane;  // this lane's first input column; lanes < 10 also fetch column px + 32
    S.H = H; S.MW = MW; S.plane = (size_t)MH * MW;
    S.py = q0 - SS_PAD;
    S.col_ok = px >= 0 && px < W;
    S.col2_ok = lane < SS_WIN - 1 && px + 32 >= 0 && px + 32 < W;
    S.out_ok = c0 + lane < MW;
    S.x_step = 4 * W; S.y_step = C * W; S.y_halo = 32 * C;
    S.xp = pred + ((ptrdiff_t)S.py * W + px) * 4 + c;
    S.yp = gt + ((ptrdiff_t)S.py * W + px) * C + c;
    S.mp = maps + (size_t)(c * 3) * S.plane + (size_t)q0 * MW + (c0 + lane);
    S.line0 = pinned((uint32_t)__cvta_generic_to_shared(&s_line[warp][0][0][0]) + 4u * lane);
    S.fill = S.use = 0u;
    S.halo = lane < SS_WIN - 1;
    S.lane = lane;
    S.s_acc = 0.f;
    const int n_in = q_end - q0 + SS_WIN - 1;  // input rows q0 - 6 .. q_end - 1 + 4
    float acc[5][SS_WIN];
#pragma unroll
    for (int d = 0; d < SS_DEPTH - 1; d++) stats_fetch(S);
    for (int i = 0; i < n_in; i += SS_WIN) {
        stats_row<0>(S, acc, win, i, n_in); stats_row<1>(S, acc, win, i + 1, n_in);
        stats_row<2>(S, acc, win, i + 2, n_in); stats_row<3>(S, acc, win, i + 3, n_in);
        stats_row<4>(S, acc, win, i + 4, n_in); stats_row<5>(S, acc, win, i + 5, n_in);
        stats_row<6>(S, acc, win, i + 6, n_in); stats_row<7>(S, acc, win, i + 7, n_in);
        stats_row<8>(S, acc, win, i + 8, n_in); stats_row<9>(S, acc, win, i + 9, n_in);
        stats_row<10>(S, acc, win, i + 10, n_in);
    }
    cp_async_wait<0>();
    float s_acc = S.s_acc;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s_acc += __shfl_xor_sync(0xffffffffu, s_acc, o);
    if (lane == 0) atomicAdd(sums + 0, s_acc);
}

// ---- kernel B -------------------------------------------------------------------------------------------------
constexpr uint32_t SB_SLOT = 5u * SS_LINE * 4u;  // ring slot of kernel B: three map lines, then x and y of the pixel row
struct BackRow {  // state of one loss_backward work item
    const float* mq;    // the next map row to be REQUESTED, plane 0, at this lane's first column
    const float* xq; const float* yq;  // pred / gt of the pixel row that completes with that map row, channel c
    int x_step, y_step, py, H; bool pix_ok;
    size_t plane;
    int MH, MW, qy, c;
    bool col_ok, col2_ok, halo;
    uint32_t line0;     // shared-window address of this lane's word in the warp's ring: SS_DEPTH x three lines
    uint32_t fill, use;
    int lane;
};

__device__ __forceinline__ void back_fetch(BackRow& B) {
    const bool row_ok = (unsigned)B.qy < (unsigned)B.MH;
    const bool a = row_ok && B.col_ok, b = row_ok && B.col2_ok;
    const uint32_t dst = B.line0 + B.fill * SB_SLOT;
#pragma unroll
    for (int m = 0; m < 3; m++) {
        cp_async_f32(dst + m * (SS_LINE * 4u), B.mq + (size_t)m * B.plane, a);
        if (B.halo) cp_async_f32(dst + m * (SS_LINE * 4u) + 128u, B.mq + (size_t)m * B.plane + 32, b);
    }
    // the pixel row this map row completes (10 rows up) travels with it: its x and y are needed by the epilogue,
    // and a dependent global load there would stall the warp for a full memory round trip per row
    const bool c = B.pix_ok && (unsigned)B.py < (unsigned)B.H;
    cp_async_f32(dst + 3u * (SS_LINE * 4u), B.xq, c);
    cp_async_f32(dst + 4u * (SS_LINE * 4u), B.yq, c);
    cp_async_commit();
    B.xq += B.x_step; B.yq += B.y_step; B.py++;
    B.mq += B.MW; B.qy++;
    B.fill = (B.fill + 1u) & (SS_DEPTH - 1);
}

struct BackOut {
    const float* xq; const float* yq; float* vq;  // pred / gt / v_out at the next pixel row to complete, channel c
    int x_step, y_step, py, H, C;  // (xq / yq: only the alpha channel of 4-channel targets is read through them)
    bool col_ok;
    float k_l1, k_ss, l1;
};

template <int PH>
__device__ __forceinline__ void back_row(BackRow& B, BackOut& O, float (&acc)[3][SS_WIN], const SsimWin& win, int i,
                                         int n_in) {
    if (i >= n_in) return;
    back_fetch(B);                  // row i + SS_DEPTH - 1 (rows past the segment are requested too: guarded, unused)
    cp_async_wait<SS_DEPTH - 1>();  // row i has landed
    __syncwarp();
    const uint32_t sm = B.line0 + B.use * SB_SLOT;
    B.use = (B.use + 1u) & (SS_DEPTH - 1);
    float x, y;  // this lane's pixel of the row that completes now (valid from row 10 on)
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(x) : "r"(sm + 3u * (SS_LINE * 4u)));
    asm volatile("ld.shared.f32 %0, [%1];" : "=f"(y) : "r"(sm + 4u * (SS_LINE * 4u)));
    // tap t = q - (p - 4) carries window weight w[10 - t]
    float h[3] = {0.f, 0.f, 0.f};
#pragma unroll
    for (int t = 0; t < SS_WIN; t++) {
        const float w = win.w[SS_WIN - 1 - t];
#pragma unroll
        for (int m = 0; m < 3; m++) {
            float v;
            asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(sm + 4u * (m * SS_LINE + t)));
            h[m] = fmaf(w, v, h[m]);
        }
    }
    __syncwarp();
    vertical_step<3, PH, true>(acc, h, win);
    if (i >= SS_WIN - 1) {  // pixel row p0 + i - 10 is complete in slot (PH + 1) % 11
        constexpr int slot = (PH + 1) % SS_WIN;
        if (O.col_ok && O.py < O.H) {

            const float g = acc[0][slot] + 2.f * x * acc[1][slot] + y * acc[2][slot];
            const float d = x - y;
            const float sg = d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f);
            O.vq[0] = O.k_l1 * sg + O.k_ss * g;
            O.l1 += fabsf(d);
            if (B.c == 2) {  // the alpha component rides with the last colour channel (train.rs:249-255)
                float va = 0.f;
                if (O.C == 4) {
                    const float d3 = __ldg(O.xq + 1) - __ldg(O.yq + 1);
                    va = O.k_l1 * (d3 > 0.f ? 1.f : (d3 < 0.f ? -1.f : 0.f));
                    O.l1 += fabsf(d3);
                }
                O.vq[1] = va;
            }
        }
        O.xq += O.x_step; O.yq += O.y_step; O.vq += O.x_step; O.py++;
    }
}

// v_out f32[H,W,4] = d loss / d pred;  loss_out = (1-w) * L1 - w * ssim.   ssim_weight > 0 (the L1-only loss is
// l1_loss_backward_kernel, optim.cu)
__global__ void __launch_bounds__(32 * SS_WARPS)
image_loss_backward_kernel(const float* __restrict__ pred, const float* __restrict__ gt, int H, int W, int C,
                           float ssim_weight, const __grid_constant__ SsimWin win, const SsimGrid G,
                           const float* __restrict__ maps, float* __restrict__ sums, uint32_t* __restrict__ ticket,
                           float* __restrict__ v_out, float* __restrict__ loss_out) {
    __shared__ __align__(16) float s_line[SS_WARPS][SS_DEPTH][5][SS_LINE];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    PDL_ENTER();
    const int item = blockIdx.x * SS_WARPS + warp;
    const int MH = H + 2, MW = W + 2;
    const float inv_l1 = 1.0f / ((float)H * (float)W * (float)C);
    float l1 = 0.f;
    if (item < G.strips * G.segs * 3) {
        BackRow B;
        const int c = item % 3;
        const int c0 = ((item / 3) % G.strips) * SS_STRIP;
        const int p0 = (item / (3 * G.strips)) * G.rows;
        const int qx = c0 - (SS_WIN - 1 - SS_PAD) + lane;  // pixel p feeds map positions q in [p-4, p+6]
        B.plane = (size_t)MH * MW; B.MH = MH; B.MW = MW; B.c = c;
        B.qy = p0 - (SS_WIN - 1 - SS_PAD);
        B.col_ok = qx >= 0 && qx < MW;
        B.col2_ok = lane < SS_WIN - 1 && qx + 32 >= 0 && qx + 32 < MW;
        B.mq = maps + (size_t)(c * 3) * B.plane + ((ptrdiff_t)B.qy * MW + qx);
        B.line0 = pinned((uint32_t)__cvta_generic_to_shared(&s_line[warp][0][0][0]) + 4u * lane);
        B.fill = B.use = 0u;
        B.halo = lane < SS_WIN - 1;
        B.lane = lane;
        B.H = H; B.py = p0 - (SS_WIN - 1); B.pix_ok = c0 + lane < W;
        B.x_step = 4 * W; B.y_step = C * W;
        B.xq = pred + ((ptrdiff_t)B.py * W + c0 + lane) * 4 + c;
        B.yq = gt + ((ptrdiff_t)B.py * W + c0 + lane) * C + c;
        BackOut O;
        O.H = H; O.C = C; O.py = p0; O.col_ok = c0 + lane < W;
        O.x_step = 4 * W; O.y_step = C * W;
        O.xq = pred + ((size_t)p0 * W + c0 + lane) * 4 + c;
        O.yq = gt + ((size_t)p0 * W + c0 + lane) * C + c;
        O.vq = v_out + ((size_t)p0 * W + c0 + lane) * 4 + c;
        O.k_l1 = (1.0f - ssim_weight) * inv_l1;
        O.k_ss = -ssim_weight / (3.0f * (float)MH * (float)MW);
        O.l1 = 0.f;
        const int p_end = min(p0 + G.rows, H);
        const int n_in = p_end - p0 + SS_WIN - 1;  // map rows p0 - 4 .. p_end - 1 + 6
        float acc[3][SS_WIN];
        if (p_end > p0) {
#pragma unroll
            for (int d = 0; d < SS_DEPTH - 1; d++) back_fetch(B);
            for (int i = 0; i < n_in; i += SS_WIN) {
                back_row<0>(B, O, acc, win, i, n_in); back_row<1>(B, O, acc, win, i + 1, n_in);
                back_row<2>(B, O, acc, win, i + 2, n_in); back_row<3>(B, O, acc, win, i + 3, n_in);
                back_row<4>(B, O, acc, win, i + 4, n_in); back_row<5>(B, O, acc, win, i + 5, n_in);
                back_row<6>(B, O, acc, win, i + 6, n_in); back_row<7>(B, O, acc, win, i + 7, n_in);
                back_row<8>(B, O, acc, win, i + 8, n_in); back_row<9>(B, O, acc, win, i + 9, n_in);
                back_row<10>(B, O, acc, win, i + 10, n_in);
            }
        }
        cp_async_wait<0>();
        l1 = O.l1;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) l1 += __shfl_xor_sync(0xffffffffu, l1, o);
    if (lane == 0 && l1 != 0.f) atomicAdd(sums + 1, l1);
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        if (atomicAdd(ticket, 1u) == gridDim.x - 1u) {  // last block: combine the two means
            __threadfence();
            const float ssim_mean = *(volatile float*)(sums + 0) / (3.0f * (float)MH * (float)MW);
            const float l1_mean = *(volatile float*)(sums + 1) * inv_l1;
            *loss_out = (1.0f - ssim_weight) * l1_mean - ssim_weight * ssim_mean;
            sums[2] = ssim_mean;
            sums[3] = l1_mean;
        }
    }
}

// Rows per segment such that every work item of the launch is resident at once (one wave): the per-item cost is
// (rows + 10) input rows, so fewer, longer segments also waste less on the vertical halo.
static SsimGrid plan_grid(int width, int height, const void* kernel) {
    static int sm_count = 0;
    if (!sm_count) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev);
    }
    int ctas_per_sm = 0;
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&ctas_per_sm, kernel, 32 * SS_WARPS, 0);
    const int slots = max(1, sm_count * max(1, ctas_per_sm) * SS_WARPS);
    SsimGrid G;
    G.strips = (width + SS_STRIP - 1) / SS_STRIP;
    const int max_segs = max(1, slots / (G.strips * 3));
    G.rows = max(16, (height + max_segs - 1) / max_segs);
    G.segs = (height + G.rows - 1) / G.rows;
    return G;
}

// scratch: maps f32[9 * (H+2)*(W+2)], sums f32[4] + ticket u32 (zeroed here)
void launch_image_loss_backward(const float* pred, const float* gt, uint32_t H, uint32_t W, uint32_t C,
                                float ssim_weight, float* maps, float* sums, float* v_out, float* loss_out,
                                cudaStream_t st) {
    static const SsimWin win = make_window();
    if (!(ssim_weight > 0.f)) {  // plain L1 (ssim_weight = 0): one pass, no window
        zero_async(loss_out, 4, st);
        launch_l1_loss_backward(pred, gt, H * W, C, v_out, loss_out, st);
        return;
    }
    zero_async(sums, 32, st);
    const SsimGrid GA = plan_grid((int)W + 2, (int)H + 2, (const void*)ssim_stats_kernel);
    const int items_a = GA.strips * GA.segs * 3;
    g_launch_count++;
    launch_k(ssim_stats_kernel, (items_a + SS_WARPS - 1) / SS_WARPS, 32 * SS_WARPS, 0, st, pred, gt, (int)H, (int)W, (int)C,
             win, GA, maps, sums);
    const SsimGrid GB = plan_grid((int)W, (int)H, (const void*)image_loss_backward_kernel);
    const int items_b = GB.strips * GB.segs * 3;
    g_launch_count++;
    launch_k(image_loss_backward_kernel, (items_b + SS_WARPS - 1) / SS_WARPS, 32 * SS_WARPS, 0, st, pred, gt, (int)H, (int)W,
             (int)C, ssim_weight, win, GB, (const float*)maps, sums, reinterpret_cast<uint32_t*>(sums + 4), v_out, loss_out);
}

}  // namespace brush
