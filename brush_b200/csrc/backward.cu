// Fused per-splat backward: gather_grads + project_backwards in one coalesced pass over all splats.
//   == crates/brush-render/src/shaders/gather_grads.wgsl:165-232       (SH VJP, opacity VJP, xy scatter)
//   == crates/brush-render/src/shaders/project_backwards.wgsl:75-227   (conic/cov2d/J/W/R/S VJP chain)
// The reference runs these as two kernels over three intermediate buffers; fusing them reads the
// rasterizer's nine accumulated sums (v_projected[g][0..9), rows by global splat id, stride VP_STRIDE) once and
// writes every row of each dense gradient once, so the reference's zero-fill of the outputs disappears too.
#include "internal.h"

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

template <int K>
__global__ void __launch_bounds__(256)
gather_project_backward_kernel(const __grid_constant__ Cam cam, const float* __restrict__ means,
                               const float* __restrict__ log_scales, const float* __restrict__ quats,
                               const float* __restrict__ raw_opac, const float* __restrict__ v_projected,
                               float* __restrict__ v_means, float* __restrict__ v_xy_out,
                               float* __restrict__ v_log_scales, float* __restrict__ v_quats,
                               float* __restrict__ v_coeffs, float* __restrict__ v_raw_opac) {
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
    float* out_sh = v_coeffs + (size_t)g * (3 * K);
    // Every output is linear in the nine sums, so an all-zero row (culled splat, or no contributing pixel) gives
    // exact zeros; skipping the chain also keeps 0 * inf out of the result for splats behind the camera.
    if (vp0.x == 0.f && vp0.y == 0.f && vp0.z == 0.f && vp0.w == 0.f && vp1.x == 0.f && vp1.y == 0.f &&
        vp1.z == 0.f && vp1.w == 0.f && v_alpha == 0.f) {
        if ((3 * K) % 4 == 0) {
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
        float sgm = 1.0f / (1.0f + __expf(-__ldg(raw_opac + g)));
        v_raw_opac[g] = v_alpha * (sgm * (1.0f - sgm));
        reinterpret_cast<float2*>(v_xy_out)[g] = make_float2(v_xy.x, v_xy.y);
    }

    // ---- project_backwards.wgsl:83-226 ----
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

    v_means[3 * (size_t)g] = v_mean.x; v_means[3 * (size_t)g + 1] = v_mean.y; v_means[3 * (size_t)g + 2] = v_mean.z;
    v_log_scales[3 * (size_t)g] = v_scale.x * scale.x;
    v_log_scales[3 * (size_t)g + 1] = v_scale.y * scale.y;
    v_log_scales[3 * (size_t)g + 2] = v_scale.z * scale.z;
    reinterpret_cast<float4*>(v_quats)[g] = make_float4(v_quat.x, v_quat.y, v_quat.z, v_quat.w);
}

void launch_gather_project_backward(const Cam& cam, uint32_t N, uint32_t K, const float* means,
                                    const float* log_scales, const float* quats, const float* raw_opac,
                                    const float* v_projected, float* v_means, float* v_xy, float* v_log_scales,
                                    float* v_quats, float* v_coeffs, float* v_raw_opac, cudaStream_t st) {
    if (N == 0) return;
    dim3 grid((N + 255) / 256), block(256);
    g_launch_count++;
#define GB_ARGS cam, means, log_scales, quats, raw_opac, v_projected, v_means, \
                v_xy, v_log_scales, v_quats, v_coeffs, v_raw_opac
    switch (K) {
        case 1: gather_project_backward_kernel<1><<<grid, block, 0, st>>>(GB_ARGS); break;
        case 4: gather_project_backward_kernel<4><<<grid, block, 0, st>>>(GB_ARGS); break;
        case 9: gather_project_backward_kernel<9><<<grid, block, 0, st>>>(GB_ARGS); break;
        case 16: gather_project_backward_kernel<16><<<grid, block, 0, st>>>(GB_ARGS); break;
        case 25: gather_project_backward_kernel<25><<<grid, block, 0, st>>>(GB_ARGS); break;
    }
#undef GB_ARGS
}

}  // namespace brush
