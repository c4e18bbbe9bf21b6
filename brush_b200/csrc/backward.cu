// Fused per-splat backward: gather_grads + project_backwards in one coalesced pass over all splats.
//   == crates/brush-render/src/shaders/gather_grads.wgsl:165-232       (SH VJP, opacity VJP, xy scatter)
//   == crates/brush-render/src/shaders/project_backwards.wgsl:75-227   (conic/cov2d/J/W/R/S VJP chain)
// The reference runs these as two kernels over three intermediate buffers; fusing them reads the
// rasterizer's nine accumulated sums (v_projected[g][0..9), rows by global splat id, stride VP_STRIDE) once and
// writes every row of each dense gradient once, so the reference's zero-fill of the outputs disappears too.
#include "internal.h"
#include "peer.cuh"

namespace brush {

BD V3 project_pix_vjp(const Cam& c, V3 p, V2 v_xy) {  // project_backwards.wgsl:19-23
    float rw = 1.0f / (p.z + 1e-6f);
    float vx = c.fx * v_xy.x, vy = c.fy * v_xy.y;
    return {vx * rw, vy * rw, (-(vx * p.x + vy * p.y) * rw) * rw};
}

BD float el(const M3& m, int col, int row) {
    const V3& c = m.c[col];
    return row == 0 ? c.x : (row == 1 ? c.y : c.z);
}

BD V4 quat_to_rotmat_vjp(V4 q, const M3& R) {  // project_backwards.wgsl:25-57
    float w = q.x, x = q.y, y = q.z, z = q.w;
    return {2.0f * (x * (el(R, 1, 2) - el(R, 2, 1)) + y * (el(R, 2, 0) - el(R, 0, 2)) + z * (el(R, 0, 1) - el(R, 1, 0))),
            2.0f * (-2.f * x * (el(R, 1, 1) + el(R, 2, 2)) + y * (el(R, 0, 1) + el(R, 1, 0)) +
                    z * (el(R, 0, 2) + el(R, 2, 0)) + w * (el(R, 1, 2) - el(R, 2, 1))),
            2.0f * (x * (el(R, 0, 1) + el(R, 1, 0)) - 2.f * y * (el(R, 0, 0) + el(R, 2, 2)) +
                    z * (el(R, 1, 2) + el(R, 2, 1)) + w * (el(R, 2, 0) - el(R, 0, 2))),
            2.0f * (x * (el(R, 0, 2) + el(R, 2, 0)) + y * (el(R, 1, 2) + el(R, 2, 1)) -
                    2.f * z * (el(R, 0, 0) + el(R, 1, 1)) + w * (el(R, 0, 1) - el(R, 1, 0)))};
}

BD V3 cov2d_to_conic_vjp(V3 conic, V3 v_conic) {  // project_backwards.wgsl:59-72
    M2 X = {{{conic.x, conic.y}, {conic.y, conic.z}}};
    M2 G = {{{v_conic.x, v_conic.y * 0.5f}, {v_conic.y * 0.5f, v_conic.z}}};
    M2 S = mul(mul(X, G), X);
    return {-S.c[0].x, -(S.c[1].x + S.c[0].y), -S.c[1].y};
}

// project_backwards.wgsl:83-226: VJP of (mean, scale = exp(log_scale), quat) -> (xy, conic) for one splat.
BD void project_backward_chain(const Cam& cam, V3 mean, V3 scale, V4 quat, V2 v_xy, V3 v_conic, V3& v_mean_out,
                               V3& v_log_scale_out, V4& v_quat_out) {
    const M3& W = cam.W;
    const V3 p_view = mul(W, mean) + cam.t;
    V3 v_mean = mul(transpose(W), project_pix_vjp(cam, p_view, v_xy));
    const V3 cov2d = calc_cov2d(cam, p_view, scale, quat);
    const V3 conic = cov_to_conic(cov2d);
    const V3 v_cov2d = cov2d_to_conic_vjp(conic, v_conic);
    const float rz = 1.0f / p_view.z;
    const float rz2 = rz * rz;
    const M3 J = {{{cam.fx * rz, 0.f, 0.f}, {0.f, cam.fy * rz, 0.f},
                   {-cam.fx * p_view.x * rz2, -cam.fy * p_view.y * rz2, 0.f}}};  // no frustum clamp here (quirk)
    const M3 R = quat_to_rotmat(quat);
    const M3 S = {{{scale.x, 0.f, 0.f}, {0.f, scale.y, 0.f}, {0.f, 0.f, scale.z}}};
    const M3 M = mul(R, S);
    const M3 V = mul(M, transpose(M));
    const M3 v_cov = {{{v_cov2d.x, 0.5f * v_cov2d.y, 0.f}, {0.5f * v_cov2d.y, v_cov2d.z, 0.f}, {0.f, 0.f, 0.f}}};
    const M3 T = mul(J, W);
    const M3 v_V = mul(mul(transpose(T), v_cov), T);
    const M3 v_T = add(mul(mul(v_cov, T), transpose(V)), mul(mul(transpose(v_cov), T), V));
    const float v_cov3d0 = v_V.c[0].x;
    const float v_cov3d1 = v_V.c[0].y + v_V.c[1].x;
    const float v_cov3d2 = v_V.c[0].z + v_V.c[2].x;
    const float v_cov3d3 = v_V.c[1].y;
    const float v_cov3d4 = v_V.c[1].z + v_V.c[2].y;
    const float v_cov3d5 = v_V.c[2].z;
    const M3 v_J = mul(v_T, transpose(W));
    const float rz3 = rz2 * rz;
    const V3 v_t = {-cam.fx * rz2 * v_J.c[2].x, -cam.fy * rz2 * v_J.c[2].y,
                    -cam.fx * rz2 * v_J.c[0].x + 2.0f * cam.fx * p_view.x * rz3 * v_J.c[2].x -
                        cam.fy * rz2 * v_J.c[1].y + 2.0f * cam.fy * p_view.y * rz3 * v_J.c[2].y};
    v_mean = v_mean + V3{dot3(v_t, W.c[0]), dot3(v_t, W.c[1]), dot3(v_t, W.c[2])};
    const M3 v_V_symm = {{{v_cov3d0, 0.5f * v_cov3d1, 0.5f * v_cov3d2},
                          {0.5f * v_cov3d1, v_cov3d3, 0.5f * v_cov3d4},
                          {0.5f * v_cov3d2, 0.5f * v_cov3d4, v_cov3d5}}};
    const M3 v_M = mul(scale_all(v_V_symm, 2.0f), M);
    const V3 v_scale = {dot3(R.c[0], v_M.c[0]), dot3(R.c[1], v_M.c[1]), dot3(R.c[2], v_M.c[2])};
    const M3 v_R = mul(v_M, S);
    const V4 v_quat = quat_to_rotmat_vjp(quat, v_R);

    v_mean_out = v_mean;
    v_log_scale_out = {v_scale.x * scale.x, v_scale.y * scale.y, v_scale.z * scale.z};
    v_quat_out = v_quat;
}

// SH_DENSE = false (the hybrid view-parallel step): the SH gradient is not formed here.  The splat's colour gradient
// v_rgb -- all the SH gradient of one view is made of: v_coeffs[g,k,:] = basis_k(viewdir) * v_rgb, gather_grads.wgsl:
// 165-232 -- goes to colour_rows[g] = (v_r, v_g, v_b, visible) instead, 16 bytes instead of 12 K, for
// sh_views_adam_kernel on every rank to expand.
template <int K, bool SH_DENSE>
__global__ void __launch_bounds__(256)
gather_project_backward_kernel(const __grid_constant__ Cam cam, const float* __restrict__ means,
                               const float* __restrict__ log_scales, const float* __restrict__ quats,
                               const float* __restrict__ raw_opac, const float* __restrict__ v_projected,
                               float* __restrict__ v_means, float* __restrict__ v_xy_out,
                               float* __restrict__ v_log_scales, float* __restrict__ v_quats,
                               float* __restrict__ v_coeffs, float* __restrict__ v_raw_opac,
                               float4* __restrict__ colour_rows, const uint8_t* __restrict__ visible_mask) {
    // One thread per splat in GLOBAL id order: the rasterizer accumulated its nine sums into row g, so every
    // read and write here is coalesced and every output row is written exactly once (dense gradients, zero
    // for splats no pixel saw: render.rs:598-624), with no separate zero-fill pass.
    const uint32_t g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= cam.total_splats) return;
    const float4* vp = reinterpret_cast<const float4*>(v_projected + (size_t)g * VP_STRIDE);
    const float4 vp0 = vp[0], vp1 = vp[1];
    const V2 v_xy = {vp0.x, vp0.y};
    const V3 v_conic = {vp0.z, vp0.w, vp1.x};
    const V3 v_rgb = {vp1.y, vp1.z, vp1.w};
    const float v_alpha = reinterpret_cast<const float*>(vp)[8];
    float* out_sh = SH_DENSE ? v_coeffs + (size_t)g * (3 * K) : nullptr;
    if (!SH_DENSE) colour_rows[g] = make_float4(v_rgb.x, v_rgb.y, v_rgb.z, (float)visible_mask[g]);
    // Every output is linear in the nine sums, so an all-zero row (culled splat, or no contributing pixel) gives
    // exact zeros; skipping the chain also keeps 0 * inf out of the result for splats behind the camera.
    if (vp0.x == 0.f && vp0.y == 0.f && vp0.z == 0.f && vp0.w == 0.f && vp1.x == 0.f && vp1.y == 0.f &&
        vp1.z == 0.f && vp1.w == 0.f && v_alpha == 0.f) {
        if (!SH_DENSE) {
        } else if ((3 * K) % 4 == 0) {
#pragma unroll
            for (int j = 0; j < (3 * K) / 4; j++) reinterpret_cast<float4*>(out_sh)[j] = make_float4(0.f, 0.f, 0.f, 0.f);
        } else {
#pragma unroll
            for (int j = 0; j < 3 * K; j++) out_sh[j] = 0.f;
        }
        v_raw_opac[g] = 0.f;
        reinterpret_cast<float2*>(v_xy_out)[g] = make_float2(0.f, 0.f);
        v_means[3 * (size_t)g] = 0.f; v_means[3 * (size_t)g + 1] = 0.f; v_means[3 * (size_t)g + 2] = 0.f;
        v_log_scales[3 * (size_t)g] = 0.f; v_log_scales[3 * (size_t)g + 1] = 0.f; v_log_scales[3 * (size_t)g + 2] = 0.f;
        reinterpret_cast<float4*>(v_quats)[g] = make_float4(0.f, 0.f, 0.f, 0.f);
        return;
    }

    const V3 mean = {__ldg(means + 3 * (size_t)g), __ldg(means + 3 * (size_t)g + 1), __ldg(means + 3 * (size_t)g + 2)};
    const V3 scale = {__expf(__ldg(log_scales + 3 * (size_t)g)), __expf(__ldg(log_scales + 3 * (size_t)g + 1)),
                      __expf(__ldg(log_scales + 3 * (size_t)g + 2))};
    const float4 q4 = __ldg(reinterpret_cast<const float4*>(quats) + g);
    const V4 quat = {q4.x, q4.y, q4.z, q4.w};

    // ---- gather_grads.wgsl:175-231 ----
    {
        if (SH_DENSE) {
        V3 d = mean - cam.t;
        float inv = rsqrtf(dot3(d, d));
        V3 viewdir = d * inv;
        float b[K];
        sh_basis<K>(cam.sh_degree, viewdir, b);
        float* out = out_sh;
        if ((3 * K) % 4 == 0) {
            float o[3 * K];
#pragma unroll
            for (int k = 0; k < K; k++) { o[3 * k] = b[k] * v_rgb.x; o[3 * k + 1] = b[k] * v_rgb.y; o[3 * k + 2] = b[k] * v_rgb.z; }
#pragma unroll
            for (int j = 0; j < (3 * K) / 4; j++)
                reinterpret_cast<float4*>(out)[j] = make_float4(o[4 * j], o[4 * j + 1], o[4 * j + 2], o[4 * j + 3]);
        } else {
#pragma unroll
            for (int k = 0; k < K; k++) { out[3 * k] = b[k] * v_rgb.x; out[3 * k + 1] = b[k] * v_rgb.y; out[3 * k + 2] = b[k] * v_rgb.z; }
        }
        }
        float sgm = 1.0f / (1.0f + __expf(-__ldg(raw_opac + g)));
        v_raw_opac[g] = v_alpha * (sgm * (1.0f - sgm));
        reinterpret_cast<float2*>(v_xy_out)[g] = make_float2(v_xy.x, v_xy.y);
    }

    V3 v_mean, v_log_scale;
    V4 v_quat;
    project_backward_chain(cam, mean, scale, quat, v_xy, v_conic, v_mean, v_log_scale, v_quat);

    v_means[3 * (size_t)g] = v_mean.x; v_means[3 * (size_t)g + 1] = v_mean.y; v_means[3 * (size_t)g + 2] = v_mean.z;
    v_log_scales[3 * (size_t)g] = v_log_scale.x;
    v_log_scales[3 * (size_t)g + 1] = v_log_scale.y;
    v_log_scales[3 * (size_t)g + 2] = v_log_scale.z;
    reinterpret_cast<float4*>(v_quats)[g] = make_float4(v_quat.x, v_quat.y, v_quat.z, v_quat.w);
}

void launch_gather_project_backward(const Cam& cam, uint32_t N, uint32_t K, const float* means,
                                    const float* log_scales, const float* quats, const float* raw_opac,
                                    const float* v_projected, float* v_means, float* v_xy, float* v_log_scales,
                                    float* v_quats, float* v_coeffs, float* v_raw_opac, float* colour_rows,
                                    const uint8_t* visible_mask, cudaStream_t st) {
    if (N == 0) return;
    dim3 grid((N + 255) / 256), block(256);
    g_launch_count++;
#define GB_ARGS cam, means, log_scales, quats, raw_opac, v_projected, v_means, \
                v_xy, v_log_scales, v_quats, v_coeffs, v_raw_opac, (float4*)colour_rows, visible_mask
    if (colour_rows) {  // hybrid step: no dense SH gradient (K plays no part)
        gather_project_backward_kernel<1, false><<<grid, block, 0, st>>>(GB_ARGS);
        return;
    }
    switch (K) {
        case 1: gather_project_backward_kernel<1, true><<<grid, block, 0, st>>>(GB_ARGS); break;
        case 4: gather_project_backward_kernel<4, true><<<grid, block, 0, st>>>(GB_ARGS); break;
        case 9: gather_project_backward_kernel<9, true><<<grid, block, 0, st>>>(GB_ARGS); break;
        case 16: gather_project_backward_kernel<16, true><<<grid, block, 0, st>>>(GB_ARGS); break;
        case 25: gather_project_backward_kernel<25, true><<<grid, block, 0, st>>>(GB_ARGS); break;
    }
#undef GB_ARGS
}

// ---------------------------------------------------------------------------------------------
// Single-device fast path: the per-splat backward, the VJP of the rotation normalisation and the Adam update of
// all five parameter groups in ONE pass, so the dense gradients (244 B per splat at SH degree 3) are never written
// to or re-read from HBM.  == gather_grads + project_backwards (above) -> autodiff of
// gaussian_splats.rs:174-175 -> the five optim.step calls of train.rs:318-359, same formulas and operation order
// as gather_project_backward_kernel + quat_normalize_vjp_kernel + adam_kernel.
//   phase 1 (thread = splat): nine sums -> v_mean, v_log_scale, v_quat (+ normalisation VJP), v_opac; Adam on
//       those 11 parameters; SH basis and v_rgb of the splat go to shared memory; v_xy is written out for the
//       refine statistics.
//   phase 2 (block-cooperative): the block's 256 SH rows are contiguous, so parameter and moments stream through
//       as coalesced float4s; element e of row s has gradient basis[s][e / 3] * v_rgb[s][e % 3].
BD void adam_update(float& p, float g, float& m, float& v, float lr, const FusedAdamLaunch& A) {
    p = adam_update_value(p, g, m, v, lr, A.beta1, A.beta2, A.eps, A.inv_bc1, A.inv_bc2, A.grad_scale);
}

// The block's 256 SH rows, [256][3K] floats, are contiguous and 16-byte aligned (256 * 3K is a multiple of 4):
// parameter and moments stream through as coalesced float4s; grad(srow, k, channel) supplies the gradient of
// coefficient k / channel of the block's srow-th splat.
template <int K, typename GradFn>
BD void sh_rows_adam(const FusedAdamLaunch& A, uint32_t N, GradFn grad) {
    const size_t total = (size_t)N * (3 * K);
    const size_t base = (size_t)blockIdx.x * 256u * (3 * K);
    constexpr int F4 = 256 * 3 * K / 4;
    constexpr int U = 4;  // float4 triples in flight per thread: the pass is pure streaming, latency-bound otherwise
    for (int i0 = threadIdx.x; i0 < F4; i0 += 256 * U) {
        float4 p[U], m[U], v[U];
        bool full[U];
#pragma unroll
        for (int u = 0; u < U; u++) {
            const int i = i0 + 256 * u;
            const size_t e0 = base + 4 * (size_t)i;
            full[u] = i < F4 && e0 + 3 < total;
            if (full[u]) {
                p[u] = *reinterpret_cast<const float4*>(A.sh + e0);
                m[u] = *reinterpret_cast<const float4*>(A.m_sh + e0);
                v[u] = *reinterpret_cast<const float4*>(A.v_sh + e0);
            }
        }
#pragma unroll
        for (int u = 0; u < U; u++) {
            const int i = i0 + 256 * u;
            const size_t e0 = base + 4 * (size_t)i;
            if (i >= F4 || e0 >= total) continue;
            float gsh[4], lr[4];
            int srow = (4 * i) / (3 * K), r = 4 * i - srow * (3 * K), kk = r / 3, ch = r - 3 * kk;
#pragma unroll
            for (int c = 0; c < 4; c++) {
                gsh[c] = grad(srow, kk, ch);
                lr[c] = r >= 3 ? A.lr_sh_rest : A.lr_sh_dc;
                r++;
                if (++ch == 3) { ch = 0; kk++; }
                if (r == 3 * K) { r = 0; kk = 0; srow++; }
            }
            if (full[u]) {
                adam_update(p[u].x, gsh[0], m[u].x, v[u].x, lr[0], A);
                adam_update(p[u].y, gsh[1], m[u].y, v[u].y, lr[1], A);
                adam_update(p[u].z, gsh[2], m[u].z, v[u].z, lr[2], A);
                adam_update(p[u].w, gsh[3], m[u].w, v[u].w, lr[3], A);
                *reinterpret_cast<float4*>(A.sh + e0) = p[u];
                *reinterpret_cast<float4*>(A.m_sh + e0) = m[u];
                *reinterpret_cast<float4*>(A.v_sh + e0) = v[u];
            } else {  // the last, partial float4 of the whole array
                for (int c = 0; c < 4 && e0 + c < total; c++) {
                    float ps = A.sh[e0 + c], ms = A.m_sh[e0 + c], vs = A.v_sh[e0 + c];
                    adam_update(ps, gsh[c], ms, vs, lr[c], A);
                    A.sh[e0 + c] = ps; A.m_sh[e0 + c] = ms; A.v_sh[e0 + c] = vs;
                }
            }
        }
    }
}

// Adam on the 11 per-splat parameters of splat g (order of train.rs:320-356: means, opacity, [sh], rotation, scale)
BD void small_params_adam(const FusedAdamLaunch& A, uint32_t g, float* mean, float* ls, float4 q_raw, float opac,
                          V3 v_mean, V3 v_ls, V4 v_q, float v_opac) {
    const float gm[3] = {v_mean.x, v_mean.y, v_mean.z}, gs[3] = {v_ls.x, v_ls.y, v_ls.z};
#pragma unroll
    for (int c = 0; c < 3; c++) {
        const size_t e = 3 * (size_t)g + c;
        float m = A.m_means[e], v = A.v_means[e];
        adam_update(mean[c], gm[c], m, v, A.lr_mean, A);
        A.means[e] = mean[c]; A.m_means[e] = m; A.v_means[e] = v;
        m = A.m_scales[e]; v = A.v_scales[e];
        adam_update(ls[c], gs[c], m, v, A.lr_scale, A);
        A.log_scales[e] = ls[c]; A.m_scales[e] = m; A.v_scales[e] = v;
    }
    float m = A.m_opac[g], v = A.v_opac[g];
    adam_update(opac, v_opac, m, v, A.lr_opac, A);
    A.raw_opac[g] = opac; A.m_opac[g] = m; A.v_opac[g] = v;
    float4 m4 = reinterpret_cast<float4*>(A.m_rot)[g], v4 = reinterpret_cast<float4*>(A.v_rot)[g];
    adam_update(q_raw.x, v_q.x, m4.x, v4.x, A.lr_rot, A);
    adam_update(q_raw.y, v_q.y, m4.y, v4.y, A.lr_rot, A);
    adam_update(q_raw.z, v_q.z, m4.z, v4.z, A.lr_rot, A);
    adam_update(q_raw.w, v_q.w, m4.w, v4.w, A.lr_rot, A);
    reinterpret_cast<float4*>(A.rotation)[g] = q_raw;
    reinterpret_cast<float4*>(A.m_rot)[g] = m4;
    reinterpret_cast<float4*>(A.v_rot)[g] = v4;
}

// VJP of q / |q| (== quat_normalize_vjp_kernel)
BD V4 quat_normalize_vjp(float4 q_raw, V4 v_q) {
    const float inv = 1.0f / sqrtf(q_raw.x * q_raw.x + q_raw.y * q_raw.y + q_raw.z * q_raw.z + q_raw.w * q_raw.w);
    const float hx = q_raw.x * inv, hy = q_raw.y * inv, hz = q_raw.z * inv, hw = q_raw.w * inv;
    const float d = hx * v_q.x + hy * v_q.y + hz * v_q.z + hw * v_q.w;
    return {(v_q.x - hx * d) * inv, (v_q.y - hy * d) * inv, (v_q.z - hz * d) * inv, (v_q.w - hw * d) * inv};
}

#ifndef BA_MINB
#define BA_MINB 2
#endif
template <int K>
__global__ void __launch_bounds__(256, BA_MINB)
backward_adam_kernel(const __grid_constant__ Cam cam, const float* __restrict__ norm_quats,
                     const float* __restrict__ v_projected, const __grid_constant__ FusedAdamLaunch A,
                     float* __restrict__ v_xy_out, const uint8_t* __restrict__ visible_mask,
                     float* __restrict__ grad_norm_accum, float* __restrict__ visible_accum) {
    constexpr int ROW = K + 3;
    __shared__ float s_sh[256][ROW + 1];  // basis[K], v_rgb[3]
    const uint32_t N = cam.total_splats;
    const uint32_t g = blockIdx.x * 256u + threadIdx.x;
    PDL_ENTER();
    if (g < N) {
        const float4* vp = reinterpret_cast<const float4*>(v_projected + (size_t)g * VP_STRIDE);
        const float4 vp0 = vp[0], vp1 = vp[1];
        const float v_alpha = reinterpret_cast<const float*>(vp)[8];
        const bool zero = vp0.x == 0.f && vp0.y == 0.f && vp0.z == 0.f && vp0.w == 0.f && vp1.x == 0.f &&
                          vp1.y == 0.f && vp1.z == 0.f && vp1.w == 0.f && v_alpha == 0.f;
        float mean[3] = {A.means[3 * (size_t)g], A.means[3 * (size_t)g + 1], A.means[3 * (size_t)g + 2]};
        float ls[3] = {A.log_scales[3 * (size_t)g], A.log_scales[3 * (size_t)g + 1], A.log_scales[3 * (size_t)g + 2]};
        float4 q_raw = reinterpret_cast<const float4*>(A.rotation)[g];
        float opac = A.raw_opac[g];
        V3 v_mean = {0.f, 0.f, 0.f}, v_ls = {0.f, 0.f, 0.f};
        V4 v_q = {0.f, 0.f, 0.f, 0.f};
        float v_opac = 0.f;
        float b[K];
#pragma unroll
        for (int k = 0; k < K; k++) b[k] = 0.f;
        if (!zero) {  // all outputs are linear in the nine sums: an all-zero row gives exact zeros
            const V3 mean_v = {mean[0], mean[1], mean[2]};
            const V3 scale = {__expf(ls[0]), __expf(ls[1]), __expf(ls[2])};
            const float4 q4 = __ldg(reinterpret_cast<const float4*>(norm_quats) + g);
            project_backward_chain(cam, mean_v, scale, V4{q4.x, q4.y, q4.z, q4.w}, V2{vp0.x, vp0.y},
                                   V3{vp0.z, vp0.w, vp1.x}, v_mean, v_ls, v_q);
            v_q = quat_normalize_vjp(q_raw, v_q);
            const float sgm = 1.0f / (1.0f + __expf(-opac));
            v_opac = v_alpha * (sgm * (1.0f - sgm));
            const V3 dv = mean_v - cam.t;
            sh_basis<K>(cam.sh_degree, dv * rsqrtf(dot3(dv, dv)), b);
        }
        if (v_xy_out) reinterpret_cast<float2*>(v_xy_out)[g] = make_float2(vp0.x, vp0.y);
        if (grad_norm_accum) {  // refine statistics (== xy_grad_stats_kernel + visible_flags_kernel + two adds)
            const float gx = vp0.x * (0.5f * (float)cam.img_w), gy = vp0.y * (0.5f * (float)cam.img_h);
            grad_norm_accum[g] += sqrtf(gx * gx + gy * gy) * A.grad_scale;
            visible_accum[g] += (float)visible_mask[g];
        }
#pragma unroll
        for (int k = 0; k < K; k++) s_sh[threadIdx.x][k] = b[k];
        s_sh[threadIdx.x][K] = vp1.y; s_sh[threadIdx.x][K + 1] = vp1.z; s_sh[threadIdx.x][K + 2] = vp1.w;
        small_params_adam(A, g, mean, ls, q_raw, opac, v_mean, v_ls, v_q, v_opac);
    }
    __syncthreads();
    sh_rows_adam<K>(A, N, [&](int srow, int kk, int ch) { return s_sh[srow][kk] * s_sh[srow][K + ch]; });
}

void launch_backward_adam(const Cam& cam, uint32_t N, uint32_t K, const float* norm_quats, const float* v_projected,
                          const FusedAdamLaunch& A, float* v_xy, const uint8_t* visible_mask, float* grad_norm_accum,
                          float* visible_accum, cudaStream_t st) {
    if (N == 0) return;
    dim3 grid((N + 255) / 256), block(256);
    g_launch_count++;
    switch (K) {
        case 1: launch_k(backward_adam_kernel<1>, grid, block, 0, st, cam, norm_quats, v_projected, A, v_xy, visible_mask,
                         grad_norm_accum, visible_accum); break;
        case 4: launch_k(backward_adam_kernel<4>, grid, block, 0, st, cam, norm_quats, v_projected, A, v_xy, visible_mask,
                         grad_norm_accum, visible_accum); break;
        case 9: launch_k(backward_adam_kernel<9>, grid, block, 0, st, cam, norm_quats, v_projected, A, v_xy, visible_mask,
                         grad_norm_accum, visible_accum); break;
        case 16: launch_k(backward_adam_kernel<16>, grid, block, 0, st, cam, norm_quats, v_projected, A, v_xy, visible_mask,
                         grad_norm_accum, visible_accum); break;
        case 25: launch_k(backward_adam_kernel<25>, grid, block, 0, st, cam, norm_quats, v_projected, A, v_xy, visible_mask,
                         grad_norm_accum, visible_accum); break;
    }
}

// ---------------------------------------------------------------------------------------------
// View-parallel step without exchanging dense gradients.  The per-splat backward is nonlinear in the
// camera but cheap, and its input -- the nine sums rasterize_backwards leaves per splat -- is 5x smaller than its
// output.  So every rank reads every view's rows (48 B per splat and view, over NVLink for the peers' views),
// redoes project_backwards + the SH VJP per view in registers, sums in a fixed view order, and applies Adam to
// its own full replica: replicas stay bit-identical without any parameter broadcast, and the only cross-GPU
// traffic of a training step is 48 * N * (world - 1) bytes per rank.
//   publish_view_kernel: marks the view's visible splats in word 9 of their rows, copies the view's uniforms and
//       loss term behind the rows;  arrive_kernel: release-stores this step's epoch into every rank's flag block.
//   views_adam_kernel: acquire-waits for all ranks' epochs, then per block of 256 splats: for each view, the
//       block's 12 KB of rows is fetched coalesced with cp.async into a ring in shared memory (two views in flight
//       while one is processed), and each thread runs the chain for its splat; SH gradients accumulate in shared memory;
//       Adam as in backward_adam_kernel.  Row buffers alternate between two halves by step parity, so no second
//       barrier is needed: a rank can overwrite a half only after every peer has arrived at the step in between.
__global__ void __launch_bounds__(256)
publish_view_kernel(float* __restrict__ rows, uint32_t N, const uint8_t* __restrict__ visible_mask,
                    const uint32_t* __restrict__ uniforms, const float* __restrict__ loss) {
    const uint32_t i = blockIdx.x * 256u + threadIdx.x;
    if (i < N) rows[(size_t)i * VP_STRIDE + 9] = (float)visible_mask[i];  // the rows are still in L2
    float* trailer = rows + (size_t)N * VP_STRIDE;
    if (i < 28) reinterpret_cast<uint32_t*>(trailer)[i] = uniforms[i];
    if (i == 28) trailer[28] = loss ? *loss : 0.0f;
}

// Fallback when the mask of the forward that produced `aux` is no longer held by the ctx: rebuild it from the
// compact list (zero-filled first).
__global__ void __launch_bounds__(256)
visible_mask_from_list_kernel(const uint32_t* __restrict__ global_from_compact,
                              const int32_t* __restrict__ num_visible_dev, uint8_t* __restrict__ visible_mask) {
    const uint32_t c = blockIdx.x * 256u + threadIdx.x;
    if (c < (uint32_t)*num_visible_dev) visible_mask[global_from_compact[c]] = 1;
}
void launch_visible_mask_from_list(const uint32_t* global_from_compact, const int32_t* num_visible_dev, uint32_t N,
                                   uint8_t* visible_mask, cudaStream_t st) {
    if (!N) return;
    zero_async(visible_mask, N, st);
    g_launch_count++;
    visible_mask_from_list_kernel<<<(N + 255) / 256, 256, 0, st>>>(global_from_compact, num_visible_dev, visible_mask);
}

struct ArriveLaunch { uint32_t* flags[DP_MAX_WORLD]; int world, rank; uint32_t epoch; };
__global__ void arrive_kernel(const __grid_constant__ ArriveLaunch L) {
    if (threadIdx.x < (uint32_t)L.world) {
        __threadfence_system();
        st_release_sys(L.flags[threadIdx.x] + L.rank, L.epoch);
    }
}

void launch_publish_view(float* rows, uint32_t N, const uint8_t* visible_mask, const void* uniforms_dev,
                         const float* loss_dev, const ViewsLaunch& X, cudaStream_t st) {
    const uint32_t threads = N > 32 ? N : 32;
    g_launch_count += 2;
    publish_view_kernel<<<(threads + 255) / 256, 256, 0, st>>>(rows, N, visible_mask,
                                                               (const uint32_t*)uniforms_dev, loss_dev);
    ArriveLaunch L{};
    for (int r = 0; r < X.world; r++) L.flags[r] = X.flags[r];
    L.world = X.world; L.rank = X.rank; L.epoch = X.epoch;
    arrive_kernel<<<1, 32, 0, st>>>(L);
}

constexpr int VIEW_STAGES = 3;
template <int K>
__global__ void __launch_bounds__(256, 2)
views_adam_kernel(const __grid_constant__ ViewsLaunch X, uint32_t N, const float* __restrict__ norm_quats,
                  const __grid_constant__ FusedAdamLaunch A) {
    constexpr int S = (3 * K) | 1;  // odd row stride: thread-per-row accumulation is bank-conflict free
    extern __shared__ float4 views_smem[];
    float* s_g = reinterpret_cast<float*>(views_smem);            // [256][S] SH gradients, summed over views
    float4* s_rows = reinterpret_cast<float4*>(s_g + 256 * S);    // [VIEW_STAGES][256 * 3] rows of this block, per view
    __shared__ Cam s_cam[DP_MAX_WORLD];
    const int world = X.world, tid = threadIdx.x;

    if (tid < world) {  // every rank has published this step's rows
        const uint32_t* f = X.flags[X.rank] + tid;
        while ((int32_t)(ld_acquire_sys(f) - X.epoch) < 0) {}
    }
    __syncthreads();
    if (tid < world) {
        BrushUniforms u;
        const float* trailer = X.rows[tid] + (size_t)N * VP_STRIDE;
        float4* dst = reinterpret_cast<float4*>(&u);
#pragma unroll
        for (int j = 0; j < 7; j++) dst[j] = ld_peer_v4(trailer + 4 * j);
        s_cam[tid] = make_cam(u);
    }
    if (blockIdx.x == 0 && tid == 0 && X.loss_out) {
        float l = 0.f;
        for (int w = 0; w < world; w++) l += ld_peer_v4(X.rows[w] + (size_t)N * VP_STRIDE + 28).x;
        *X.loss_out = l * A.grad_scale;
    }

    const uint32_t first = blockIdx.x * 256u, g = first + tid;
    const int nf4 = 3 * (int)min(256u, N - first);
    // VIEW_STAGES - 1 views' rows in flight per block, global -> shared without passing through registers
    auto fetch = [&](int w) {
        if (w < world) {
            const float* src = X.rows[w] + (size_t)first * VP_STRIDE;
            const uint32_t dst = (uint32_t)__cvta_generic_to_shared(s_rows + (w % VIEW_STAGES) * 768);
#pragma unroll
            for (int j = 0; j < 3; j++)
                if (tid + 256 * j < nf4)
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + 16u * (tid + 256 * j)),
                                 "l"(src + 4 * (tid + 256 * j)) : "memory");
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
#pragma unroll
    for (int w = 0; w < VIEW_STAGES - 1; w++) fetch(w);

    float mean[3] = {0.f, 0.f, 0.f}, ls[3] = {0.f, 0.f, 0.f}, opac = 0.f;
    float4 q_raw = make_float4(0.f, 0.f, 0.f, 1.f), q4 = q_raw;
    if (g < N) {
#pragma unroll
        for (int c = 0; c < 3; c++) { mean[c] = A.means[3 * (size_t)g + c]; ls[c] = A.log_scales[3 * (size_t)g + c]; }
        q_raw = reinterpret_cast<const float4*>(A.rotation)[g];
        q4 = __ldg(reinterpret_cast<const float4*>(norm_quats) + g);
        opac = A.raw_opac[g];
    }
    const V3 mean_v = {mean[0], mean[1], mean[2]};
    const V3 scale = {__expf(ls[0]), __expf(ls[1]), __expf(ls[2])};
#pragma unroll 4
    for (int e = 0; e < 3 * K; e++) s_g[tid * S + e] = 0.f;
    V3 v_mean = {0.f, 0.f, 0.f}, v_ls = {0.f, 0.f, 0.f};
    V4 v_q = {0.f, 0.f, 0.f, 0.f};
    float v_alpha_sum = 0.f, norm_sum = 0.f, vis_sum = 0.f;

    for (int w = 0; w < world; w++) {
        asm volatile("cp.async.wait_group %0;" ::"n"(VIEW_STAGES - 2) : "memory");
        __syncthreads();  // view w has landed for every thread; view w - 1 is consumed (first pass: s_cam is complete)
        fetch(w + VIEW_STAGES - 1);  // into the slot view w - 1 occupied
        if (g < N) {
            const float4* row = s_rows + (w % VIEW_STAGES) * 768 + 3 * tid;
            const float4 r0 = row[0], r1 = row[1], r2 = row[2];
            const Cam& cam = s_cam[w];
            vis_sum += r2.y;
            const bool zero = r0.x == 0.f && r0.y == 0.f && r0.z == 0.f && r0.w == 0.f && r1.x == 0.f &&
                              r1.y == 0.f && r1.z == 0.f && r1.w == 0.f && r2.x == 0.f;
            if (!zero) {
                V3 vm, vl;
                V4 vq;
                project_backward_chain(cam, mean_v, scale, V4{q4.x, q4.y, q4.z, q4.w}, V2{r0.x, r0.y},
                                       V3{r0.z, r0.w, r1.x}, vm, vl, vq);
                v_mean = v_mean + vm;
                v_ls = v_ls + vl;
                v_q = {v_q.x + vq.x, v_q.y + vq.y, v_q.z + vq.z, v_q.w + vq.w};
                v_alpha_sum += r2.x;
                const float gx = r0.x * (0.5f * (float)cam.img_w), gy = r0.y * (0.5f * (float)cam.img_h);
                norm_sum += sqrtf(gx * gx + gy * gy);
                float b[K];
                const V3 dv = mean_v - cam.t;
                sh_basis<K>(cam.sh_degree, dv * rsqrtf(dot3(dv, dv)), b);
#pragma unroll
                for (int k = 0; k < K; k++) {
                    s_g[tid * S + 3 * k] += b[k] * r1.y;
                    s_g[tid * S + 3 * k + 1] += b[k] * r1.z;
                    s_g[tid * S + 3 * k + 2] += b[k] * r1.w;
                }
            }
        }
    }
    if (g < N) {
        v_q = quat_normalize_vjp(q_raw, v_q);
        const float sgm = 1.0f / (1.0f + __expf(-opac));
        const float v_opac = v_alpha_sum * (sgm * (1.0f - sgm));
        small_params_adam(A, g, mean, ls, q_raw, opac, v_mean, v_ls, v_q, v_opac);
        if (X.grad_norm_accum) {  // refine statistics, train.rs:284-316, summed over the step's views
            X.grad_norm_accum[g] += norm_sum * A.grad_scale;
            X.visible_accum[g] += vis_sum;
        }
    }
    __syncthreads();
    sh_rows_adam<K>(A, N, [&](int srow, int kk, int ch) { return s_g[srow * S + 3 * kk + ch]; });
}

template <int K>
static int launch_views_adam_k(const ViewsLaunch& X, uint32_t N, const float* norm_quats, const FusedAdamLaunch& A,
                               cudaStream_t st) {
    const size_t smem = (size_t)256 * ((3 * K) | 1) * 4 + (size_t)VIEW_STAGES * 256 * 3 * 16;
    cudaError_t e = cudaFuncSetAttribute(views_adam_kernel<K>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    views_adam_kernel<K><<<(N + 255) / 256, 256, smem, st>>>(X, N, norm_quats, A);
    return 0;
}

int launch_views_adam(const ViewsLaunch& X, uint32_t N, uint32_t K, const float* norm_quats, const FusedAdamLaunch& A,
                      cudaStream_t st) {
    if (N == 0) return 0;
    g_launch_count++;
    switch (K) {
        case 1: return launch_views_adam_k<1>(X, N, norm_quats, A, st);
        case 4: return launch_views_adam_k<4>(X, N, norm_quats, A, st);
        case 9: return launch_views_adam_k<9>(X, N, norm_quats, A, st);
        case 16: return launch_views_adam_k<16>(X, N, norm_quats, A, st);
        case 25: return launch_views_adam_k<25>(X, N, norm_quats, A, st);
    }
    return -1;
}

// ---------------------------------------------------------------------------------------------
// Hybrid view-parallel step: the SH group is 3K of the 13 + 3K gradient floats of a splat (81 % at SH
// degree 3), but per view it is the outer product basis(viewdir) (x) v_rgb -- three floats of information.  So the
// SH gradient is never exchanged: every rank publishes its view's colour rows (v_r, v_g, v_b, visible: 16 B per splat,
// written by gather_project_backward_kernel<K, false>), reads every other view's rows over NVLink, rebuilds
// sum_v basis(viewdir_v) (x) v_rgb_v for ALL splats in a fixed view order and applies Adam to its own full SH replica
// (bit-identical on every rank, no parameter broadcast).  The eleven remaining gradient floats go through
// dp_adam_kernel as dense sums.  Inbound NVLink bytes per splat and step: 16 (world - 1) + 44, against 237 for the
// dense exchange at world 8 and 48 (world - 1) for the `views` exchange.
__global__ void publish_trailer_kernel(float* __restrict__ trailer, const uint32_t* __restrict__ uniforms,
                                       const float* __restrict__ loss) {
    const uint32_t i = threadIdx.x;
    if (i < 28) reinterpret_cast<uint32_t*>(trailer)[i] = uniforms[i];
    if (i == 28) trailer[28] = loss ? *loss : 0.0f;
}

void launch_publish_colour_rows(float* rows, uint32_t N, const void* uniforms_dev, const float* loss_dev,
                                const ViewsLaunch& X, cudaStream_t st) {
    g_launch_count += 2;
    publish_trailer_kernel<<<1, 32, 0, st>>>(rows + (size_t)N * RGB_STRIDE, (const uint32_t*)uniforms_dev, loss_dev);
    ArriveLaunch L{};
    for (int r = 0; r < X.world; r++) L.flags[r] = X.flags[r];
    L.world = X.world; L.rank = X.rank; L.epoch = X.epoch;
    arrive_kernel<<<1, 32, 0, st>>>(L);
}

struct ShViewsLaunch {
    const float* means;  // the means the step's forwards rendered with (BEFORE this step's update)
    float *sh, *m_sh, *v_sh;
    float lr_sh_dc, lr_sh_rest, beta1, beta2, eps, inv_bc1, inv_bc2, grad_scale;
};

template <int K>
__global__ void __launch_bounds__(256, 2)
sh_views_adam_kernel(const __grid_constant__ ViewsLaunch X, uint32_t N, const __grid_constant__ ShViewsLaunch P) {
    constexpr int S = (3 * K) | 1;
    extern __shared__ float4 shv_smem[];
    float* s_g = reinterpret_cast<float*>(shv_smem);             // [256][S] SH gradients, summed over views
    float4* s_rows = reinterpret_cast<float4*>(s_g + 256 * S);   // [world][256] colour rows of this block, per view
    __shared__ Cam s_cam[DP_MAX_WORLD];
    const int world = X.world, tid = threadIdx.x;
    if (tid < world) {
        const uint32_t* f = X.flags[X.rank] + tid;
        while ((int32_t)(ld_acquire_sys(f) - X.epoch) < 0) __nanosleep(64);
    }
    __syncthreads();
    if (tid < world) {
        BrushUniforms u;
        const float* trailer = X.rows[tid] + (size_t)N * RGB_STRIDE;
        float4* dst = reinterpret_cast<float4*>(&u);
#pragma unroll
        for (int j = 0; j < 7; j++) dst[j] = ld_peer_v4(trailer + 4 * j);
        s_cam[tid] = make_cam(u);
    }
    if (blockIdx.x == 0 && tid == 0 && X.loss_out) {
        float l = 0.f;
        for (int w = 0; w < world; w++) l += ld_peer_v4(X.rows[w] + (size_t)N * RGB_STRIDE + 28).x;
        *X.loss_out = l * P.grad_scale;
    }
    const uint32_t first = blockIdx.x * 256u, g = first + tid;
    const int count = (int)min(256u, N - first);
    // Every view's colour rows of the block (4 KB each) are requested at once: with a ring of three the block paid an
    // NVLink round trip per view behind a barrier, one after the other (0.06 ms per view and step at 1 M splats).
    for (int w = 0; w < world; w++) {
        if (tid < count) {
            const float* src = X.rows[w] + ((size_t)first + tid) * RGB_STRIDE;
            const uint32_t dst = (uint32_t)__cvta_generic_to_shared(s_rows + w * 256 + tid);
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
        }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    V3 mean_v = {0.f, 0.f, 0.f};
    if (g < N) mean_v = {P.means[3 * (size_t)g], P.means[3 * (size_t)g + 1], P.means[3 * (size_t)g + 2]};
    // The sum over views lives in registers (it was a read-modify-write of 3K shared-memory words per view); the
    // view order is fixed, so every rank forms the same sum.
    float acc[3 * K];
#pragma unroll
    for (int e = 0; e < 3 * K; e++) acc[e] = 0.f;
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();  // the rows have landed for every thread, and s_cam is complete
    if (g < N) {
        for (int w = 0; w < world; w++) {
            const float4 r = s_rows[w * 256 + tid];
            if (!(r.x == 0.f && r.y == 0.f && r.z == 0.f)) {
                const Cam& cam = s_cam[w];
                float b[K];
                const V3 dv = mean_v - cam.t;
                sh_basis<K>(cam.sh_degree, dv * rsqrtf(dot3(dv, dv)), b);
#pragma unroll
                for (int k = 0; k < K; k++) {
                    acc[3 * k] += b[k] * r.x;
                    acc[3 * k + 1] += b[k] * r.y;
                    acc[3 * k + 2] += b[k] * r.z;
                }
            }
        }
    }
#pragma unroll
    for (int e = 0; e < 3 * K; e++) s_g[tid * S + e] = acc[e];
    __syncthreads();
    FusedAdamLaunch A{};
    A.sh = P.sh; A.m_sh = P.m_sh; A.v_sh = P.v_sh;
    A.lr_sh_dc = P.lr_sh_dc; A.lr_sh_rest = P.lr_sh_rest;
    A.beta1 = P.beta1; A.beta2 = P.beta2; A.eps = P.eps; A.inv_bc1 = P.inv_bc1; A.inv_bc2 = P.inv_bc2;
    A.grad_scale = P.grad_scale;
    sh_rows_adam<K>(A, N, [&](int srow, int kk, int ch) { return s_g[srow * S + 3 * kk + ch]; });
}

template <int K>
static int launch_sh_views_adam_k(const ViewsLaunch& X, uint32_t N, const ShViewsLaunch& P, cudaStream_t st) {
    const size_t smem = (size_t)256 * ((3 * K) | 1) * 4 + (size_t)X.world * 256 * 16;
    cudaError_t e = cudaFuncSetAttribute(sh_views_adam_kernel<K>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    sh_views_adam_kernel<K><<<(N + 255) / 256, 256, smem, st>>>(X, N, P);
    return 0;
}

int launch_sh_views_adam(const ViewsLaunch& X, uint32_t N, uint32_t K, const float* means, float* sh, float* m_sh,
                         float* v_sh, float lr_sh_dc, float lr_sh_rest, float beta1, float beta2, float eps,
                         float inv_bc1, float inv_bc2, float grad_scale, cudaStream_t st) {
    if (N == 0) return 0;
    g_launch_count++;
    const ShViewsLaunch P{means, sh, m_sh, v_sh, lr_sh_dc, lr_sh_rest, beta1, beta2, eps, inv_bc1, inv_bc2, grad_scale};
    switch (K) {
        case 1: return launch_sh_views_adam_k<1>(X, N, P, st);
        case 4: return launch_sh_views_adam_k<4>(X, N, P, st);
        case 9: return launch_sh_views_adam_k<9>(X, N, P, st);
        case 16: return launch_sh_views_adam_k<16>(X, N, P, st);
        case 25: return launch_sh_views_adam_k<25>(X, N, P, st);
    }
    return -1;
}

}  // namespace brush
