/* C caller of the libbrush_b200.so ABI (include/brush_b200.h), written the way the Rust FFI crate calls it
 * (rust/brush-b200-sys, rust/brush-render-b200/render_b200.rs): raw cudaMalloc buffers, nothing from Python or torch.
 * Links only libbrush_b200.so + the CUDA runtime.
 *
 *   test_abi <dir> N K W H
 *
 * reads   <dir>/{means,log_scales,quats,sh_coeffs,raw_opacity}.f32 and target.f32 [H,W,3]  (raw little-endian f32;
 *         quats normalised); the upstream gradient is that of loss = mean((out_rgb - target)^2), render.rs:786-789
 * runs    brush_ctx_create -> brush_render_forward -> brush_render_backward -> brush_ctx_check, then
 *         brush_render_forward_host on the same inputs (host buffers), and brush_prefix_sum_u32 /
 *         brush_radix_argsort_u32 on small vectors (brush-prefix-sum/src/lib.rs:110-175, brush-sort/src/lib.rs:164-265)
 * writes  <dir>/out_{img,img_host,projected,gfc,v_means,v_xy,v_log_scales,v_quats,v_sh,v_opac}.bin and counts.txt
 * The Python test (tests/test_abi_c.py) compares them with the reference's golden vectors; exit code != 0 on any
 * ABI error.  Camera = test_reference's (render.rs:734-746): position (0,0,-8), identity, fov 90 degrees. */
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "brush_b200.h"

#define CK(call)                                                                        \
    do {                                                                                \
        cudaError_t e__ = (call);                                                       \
        if (e__ != cudaSuccess) {                                                       \
            fprintf(stderr, "%s: %s\n", #call, cudaGetErrorString(e__));                \
            return 2;                                                                   \
        }                                                                               \
    } while (0)
#define BR(call)                                                                        \
    do {                                                                                \
        int r__ = (call);                                                               \
        if (r__ != BRUSH_OK) {                                                          \
            fprintf(stderr, "%s -> %d: %s\n", #call, r__, brush_last_error(ctx));       \
            return 3;                                                                   \
        }                                                                               \
    } while (0)

static float* read_f32(const char* dir, const char* name, size_t n) {
    char path[1024];
    snprintf(path, sizeof path, "%s/%s.f32", dir, name);
    FILE* f = fopen(path, "rb");
    if (!f) { perror(path); exit(4); }
    float* p = (float*)malloc(n * sizeof(float) + 16);
    if (fread(p, sizeof(float), n, f) != n) { fprintf(stderr, "%s: short read\n", path); exit(4); }
    fclose(f);
    return p;
}
static int write_dev(const char* dir, const char* name, const void* dev, size_t bytes) {
    void* h = malloc(bytes + 16);
    if (cudaMemcpy(h, dev, bytes, cudaMemcpyDeviceToHost) != cudaSuccess) return 1;
    char path[1024];
    snprintf(path, sizeof path, "%s/%s.bin", dir, name);
    FILE* f = fopen(path, "wb");
    if (!f) { perror(path); return 1; }
    fwrite(h, 1, bytes, f);
    fclose(f);
    free(h);
    return 0;
}
static void* dev_upload(const void* host, size_t bytes) {
    void* d = NULL;
    if (cudaMalloc(&d, bytes ? bytes : 16) != cudaSuccess) exit(5);
    if (bytes && cudaMemcpy(d, host, bytes, cudaMemcpyHostToDevice) != cudaSuccess) exit(5);
    return d;
}
static void* dev_alloc_poisoned(size_t bytes) {  /* the reference noise-fills "empty" buffers under test */
    void* d = NULL;
    if (cudaMalloc(&d, bytes ? bytes : 16) != cudaSuccess) exit(5);
    cudaMemset(d, 0xCF, bytes ? bytes : 16);
    return d;
}

int main(int argc, char** argv) {
    if (argc != 6) { fprintf(stderr, "usage: %s dir N K W H\n", argv[0]); return 1; }
    const char* dir = argv[1];
    const uint32_t N = (uint32_t)atoi(argv[2]), K = (uint32_t)atoi(argv[3]);
    const uint32_t W = (uint32_t)atoi(argv[4]), H = (uint32_t)atoi(argv[5]);
    printf("%s\n", brush_version());

    float* means = read_f32(dir, "means", (size_t)N * 3);
    float* log_scales = read_f32(dir, "log_scales", (size_t)N * 3);
    float* quats = read_f32(dir, "quats", (size_t)N * 4);
    float* sh = read_f32(dir, "sh_coeffs", (size_t)N * K * 3);
    float* opac = read_f32(dir, "raw_opacity", (size_t)N);
    float* target = read_f32(dir, "target", (size_t)W * H * 3);
    float* v_out = (float*)calloc((size_t)W * H * 4, sizeof(float));

    brush_ctx* ctx = NULL;
    if (brush_ctx_create(0, &ctx) != BRUSH_OK || !ctx) { fprintf(stderr, "brush_ctx_create failed\n"); return 3; }
    cudaStream_t st;
    CK(cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking));

    /* render.rs:102-116 with the test camera: viewmat = translate(0,0,8) column-major; focal = 0.5 W / tan(45 deg) */
    BrushUniforms u;
    memset(&u, 0, sizeof u);
    u.viewmat[0] = u.viewmat[5] = u.viewmat[10] = u.viewmat[15] = 1.0f;
    u.viewmat[14] = 8.0f;
    u.focal[0] = u.focal[1] = 0.5f * (float)W;
    u.img_size[0] = W; u.img_size[1] = H;
    u.pixel_center[0] = 0.5f * (float)W; u.pixel_center[1] = 0.5f * (float)H;

    const uint32_t tx = (W + BRUSH_TILE_WIDTH - 1) / BRUSH_TILE_WIDTH, ty = (H + BRUSH_TILE_WIDTH - 1) / BRUSH_TILE_WIDTH;
    uint64_t cap = (uint64_t)N * tx * ty;          /* render.rs:204-206 */
    if (cap > 128ull * 65535ull) cap = 128ull * 65535ull;

    float* d_means = (float*)dev_upload(means, (size_t)N * 12);
    float* d_ls = (float*)dev_upload(log_scales, (size_t)N * 12);
    float* d_quats = (float*)dev_upload(quats, (size_t)N * 16);
    float* d_sh = (float*)dev_upload(sh, (size_t)N * K * 12);
    float* d_opac = (float*)dev_upload(opac, (size_t)N * 4);

    BrushAux aux;
    memset(&aux, 0, sizeof aux);
    aux.projected_splats = (float*)dev_alloc_poisoned((size_t)N * BRUSH_PROJECTED_FLOATS * 4);
    aux.uniforms_buffer = (int32_t*)dev_alloc_poisoned(28 * 4);
    aux.num_intersections = (int32_t*)dev_alloc_poisoned(4);
    aux.num_visible = (int32_t*)dev_alloc_poisoned(4);
    aux.final_index = (int32_t*)dev_alloc_poisoned((size_t)W * H * 4);
    aux.cum_tiles_hit = (int32_t*)dev_alloc_poisoned((size_t)N * 4);
    aux.tile_bins = (int32_t*)dev_alloc_poisoned((size_t)tx * ty * 8);
    aux.compact_gid_from_isect = (int32_t*)dev_alloc_poisoned((size_t)cap * 4);
    aux.global_from_compact_gid = (int32_t*)dev_alloc_poisoned((size_t)N * 4);
    aux.tile_id_from_isect = NULL;
    aux.isect_capacity = cap;
    float* d_img = (float*)dev_alloc_poisoned((size_t)W * H * 16);

    BR(brush_ctx_set_timing(ctx, 1));
    BR(brush_render_forward(ctx, &u, N, K, d_means, d_ls, d_quats, d_sh, d_opac, 0, d_img, &aux, st));

    /* d loss / d out for loss = mean((out_rgb - target)^2) over H*W*3 values; alpha gets no gradient */
    float* img_h = (float*)malloc((size_t)W * H * 16);
    CK(cudaMemcpyAsync(img_h, d_img, (size_t)W * H * 16, cudaMemcpyDeviceToHost, st));
    CK(cudaStreamSynchronize(st));
    for (size_t p = 0; p < (size_t)W * H; p++)
        for (int c = 0; c < 3; c++) v_out[4 * p + c] = 2.0f * (img_h[4 * p + c] - target[3 * p + c]) / (float)((size_t)W * H * 3);
    float* d_vout = (float*)dev_upload(v_out, (size_t)W * H * 16);

    float* g_means = (float*)dev_alloc_poisoned((size_t)N * 12);
    float* g_xy = (float*)dev_alloc_poisoned((size_t)N * 8);
    float* g_ls = (float*)dev_alloc_poisoned((size_t)N * 12);
    float* g_quats = (float*)dev_alloc_poisoned((size_t)N * 16);
    float* g_sh = (float*)dev_alloc_poisoned((size_t)N * K * 12);
    float* g_opac = (float*)dev_alloc_poisoned((size_t)N * 4);
    BR(brush_render_backward(ctx, N, K, d_means, d_ls, d_quats, d_opac, d_img, d_vout, &aux, g_means, g_xy, g_ls,
                             g_quats, g_sh, g_opac, st));
    BR(brush_ctx_check(ctx, st));  /* synchronises; BRUSH_E_CAPACITY if the lists were truncated */
    float stage_ms[BRUSH_STAGE_COUNT];
    BR(brush_stage_times_ms(ctx, stage_ms, BRUSH_STAGE_COUNT, st));

    int32_t nv = -1, ni = -1;
    CK(cudaMemcpy(&nv, aux.num_visible, 4, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(&ni, aux.num_intersections, 4, cudaMemcpyDeviceToHost));
    int32_t uni[28];
    CK(cudaMemcpy(uni, aux.uniforms_buffer, sizeof uni, cudaMemcpyDeviceToHost));
    if (uni[25] != nv || (uint32_t)uni[26] != N) { fprintf(stderr, "uniforms word 25/26 mismatch\n"); return 6; }

    int bad = 0;
    bad |= write_dev(dir, "out_img", d_img, (size_t)W * H * 16);
    bad |= write_dev(dir, "out_projected", aux.projected_splats, (size_t)N * 36);
    bad |= write_dev(dir, "out_gfc", aux.global_from_compact_gid, (size_t)N * 4);
    bad |= write_dev(dir, "out_v_means", g_means, (size_t)N * 12);
    bad |= write_dev(dir, "out_v_xy", g_xy, (size_t)N * 8);
    bad |= write_dev(dir, "out_v_log_scales", g_ls, (size_t)N * 12);
    bad |= write_dev(dir, "out_v_quats", g_quats, (size_t)N * 16);
    bad |= write_dev(dir, "out_v_sh", g_sh, (size_t)N * K * 12);
    bad |= write_dev(dir, "out_v_opac", g_opac, (size_t)N * 4);
    if (bad) return 7;

    /* the host-staged convenience entry: same render from HOST buffers */
    float* img_host = (float*)malloc((size_t)W * H * 16);
    uint32_t nv_h = 0, ni_h = 0;
    BR(brush_render_forward_host(ctx, &u, N, K, means, log_scales, quats, sh, opac, cap, img_host, &nv_h, &ni_h));
    {
        char path[1024];
        snprintf(path, sizeof path, "%s/out_img_host.bin", dir);
        FILE* f = fopen(path, "wb");
        fwrite(img_host, 1, (size_t)W * H * 16, f);
        fclose(f);
    }
    if ((int32_t)nv_h != nv || (int32_t)ni_h != ni) { fprintf(stderr, "host-staged counts differ\n"); return 8; }

    /* the two primitives: inclusive scan (brush-prefix-sum/src/lib.rs:17) and stable argsort (brush-sort/src/lib.rs:32) */
    enum { M = 5000 };
    uint32_t h_in[M], h_out[M], h_keys[M], h_vals[M], h_ko[M], h_vo[M];
    for (uint32_t i = 0; i < M; i++) { h_in[i] = i % 7u; h_keys[i] = (i * 2654435761u) >> 20; h_vals[i] = i; }
    uint32_t *d_in = (uint32_t*)dev_upload(h_in, sizeof h_in), *d_out = (uint32_t*)dev_alloc_poisoned(sizeof h_out);
    BR(brush_prefix_sum_u32(ctx, d_in, M, d_out, st));
    uint32_t *d_k = (uint32_t*)dev_upload(h_keys, sizeof h_keys), *d_v = (uint32_t*)dev_upload(h_vals, sizeof h_vals);
    uint32_t *d_ko = (uint32_t*)dev_alloc_poisoned(sizeof h_ko), *d_vo = (uint32_t*)dev_alloc_poisoned(sizeof h_vo);
    int32_t n_sort = M - 100;
    int32_t* d_n = (int32_t*)dev_upload(&n_sort, 4);
    BR(brush_radix_argsort_u32(ctx, d_k, d_v, d_n, M, 12, d_ko, d_vo, st));
    CK(cudaStreamSynchronize(st));
    CK(cudaMemcpy(h_out, d_out, sizeof h_out, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(h_ko, d_ko, sizeof h_ko, cudaMemcpyDeviceToHost));
    CK(cudaMemcpy(h_vo, d_vo, sizeof h_vo, cudaMemcpyDeviceToHost));
    uint32_t acc = 0;
    for (uint32_t i = 0; i < M; i++) { acc += h_in[i]; if (h_out[i] != acc) { fprintf(stderr, "scan[%u]\n", i); return 9; } }
    for (int32_t i = 1; i < n_sort; i++) {
        if (h_ko[i - 1] > h_ko[i] || (h_ko[i - 1] == h_ko[i] && h_vo[i - 1] > h_vo[i])) { fprintf(stderr, "sort[%d]\n", i); return 9; }
        if (h_keys[h_vo[i]] != h_ko[i]) { fprintf(stderr, "sort payload[%d]\n", i); return 9; }
    }

    /* error behaviour: K = 5 is the reference's "Invalid nr. of sh bases" panic (render.rs:51) */
    int rc = brush_render_forward(ctx, &u, N, 5, d_means, d_ls, d_quats, d_sh, d_opac, 0, d_img, &aux, st);
    if (rc != BRUSH_E_INVALID || !strstr(brush_last_error(ctx), "Invalid nr. of sh bases")) { fprintf(stderr, "K=5 not rejected\n"); return 10; }

    char path[1024];
    snprintf(path, sizeof path, "%s/counts.txt", dir);
    FILE* f = fopen(path, "w");
    fprintf(f, "%d %d %llu\n", nv, ni, (unsigned long long)brush_launch_count());
    for (int i = 0; i < BRUSH_STAGE_COUNT; i++) fprintf(f, "%.6f\n", stage_ms[i]);
    fclose(f);
    brush_ctx_destroy(ctx);
    printf("ok: num_visible=%d num_intersections=%d\n", nv, ni);
    return 0;
}
