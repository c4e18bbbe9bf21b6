// TEST INFRASTRUCTURE ONLY -- CPU oracle for the brush differentiable splat path.
//
// A scalar C++ restatement of the reference's WGSL shaders and their host
// sequencing.  Only tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline / --impl reference legs may build, load or call it; the product
// library (brush_b200/csrc) never links or includes anything from oracle/.
//
// Parity status: PINNED for projection (xy, conic), SH degree 3, rasterize
// forward/backward and project_backwards by the reference's own golden vectors
// crates/brush-render/test_cases/{tiny_case,basic_case}.safetensors at the
// tolerances of crates/brush-render/src/render.rs:815-830 (see
// tests/test_oracle_golden.py).  UNPINNED (oracle is the authority): SH degree
// 4, the packed-u32 output, >256 splats per tile, depth-tie order,
// intersection-capacity behaviour.
//
// All citations are relative to /root/reference/.
// Build: see oracle/Makefile.  Two targets from this one file:
//   libbrush_oracle.so               the PARITY oracle: -O2 -ffp-contract=off, no fast-math, f64 accumulation of the
//                                    rasterize_backwards sums (independent of the thread count).
//   _timed/libbrush_oracle_timed.so  the TIMED CPU baseline (-DORACLE_TIMED -O3 -march=native, built on the machine
//                                    that runs bench.py): same algorithm and results to rounding, but the
//                                    rasterize_backwards sums go to thread-private f32 accumulators (no atomics) and
//                                    the LSD sorts take 8-bit digits (half the passes, identical output).
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <ctime>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif
#include "oracle_math.h"

using namespace orc;

// crates/brush-render/src/shaders/helpers.wgsl:7-30 (112 bytes).
struct Uniforms {
    float viewmat[16];  // column-major mat4
    float focal[2];
    uint32_t img_size[2];
    uint32_t tile_bounds[2];
    float pixel_center[2];
    uint32_t sh_degree;
    uint32_t num_visible;
    uint32_t total_splats;
    uint32_t padding;
};
static_assert(sizeof(Uniforms) == 112, "RenderUniforms layout");

static const uint32_t TILE_WIDTH = 16;                      // helpers.wgsl:1
static const uint32_t TILE_SIZE = TILE_WIDTH * TILE_WIDTH;  // helpers.wgsl:3
static const float COV_BLUR = 0.3f;                         // helpers.wgsl:166

static inline vec3 ld3(const float* p, size_t i) { return {p[3 * i], p[3 * i + 1], p[3 * i + 2]}; }
static inline vec4 ld4(const float* p, size_t i) { return {p[4 * i], p[4 * i + 1], p[4 * i + 2], p[4 * i + 3]}; }
static inline mat3 view_W(const Uniforms& u) {
    const float* m = u.viewmat;
    return {{{m[0], m[1], m[2]}, {m[4], m[5], m[6]}, {m[8], m[9], m[10]}}};
}
static inline vec3 view_t(const Uniforms& u) { return {u.viewmat[12], u.viewmat[13], u.viewmat[14]}; }
static inline vec3 exp3(vec3 v) { return {det_expf(v.x), det_expf(v.y), det_expf(v.z)}; }

struct bbox4 { uint32_t minx, miny, maxx, maxy; };

// helpers.wgsl:55-62
static bbox4 get_bbox(vec2 center, vec2 dims, const uint32_t bounds[2]) {
    int32_t bx = (int32_t)bounds[0], by = (int32_t)bounds[1];
    bbox4 b;
    b.minx = (uint32_t)clampi(sat_f2i(center.x - dims.x), 0, bx);
    b.miny = (uint32_t)clampi(sat_f2i(center.y - dims.y), 0, by);
    b.maxx = (uint32_t)clampi(sat_f2i((center.x + dims.x) + 1.0f), 0, bx);
    b.maxy = (uint32_t)clampi(sat_f2i((center.y + dims.y) + 1.0f), 0, by);
    return b;
}
// helpers.wgsl:64-71
static bbox4 get_tile_bbox(vec2 pix_center, uint32_t pix_radius, const uint32_t tile_bounds[2]) {
    vec2 tile_center = {pix_center.x / (float)TILE_WIDTH, pix_center.y / (float)TILE_WIDTH};
    float tile_radius = (float)pix_radius / (float)TILE_WIDTH;
    return get_bbox(tile_center, {tile_radius, tile_radius}, tile_bounds);
}
// helpers.wgsl:74-109  (quat stored w,x,y,z)
static mat3 quat_to_rotmat(vec4 q) {
    float w = q.x, x = q.y, y = q.z, z = q.w;
    float x2 = x * x, y2 = y * y, z2 = z * z;
    float xy = x * y, xz = x * z, yz = y * z;
    float wx = w * x, wy = w * y, wz = w * z;
    return {{{1.0f - 2.0f * (y2 + z2), 2.0f * (xy + wz), 2.0f * (xz - wy)},
             {2.0f * (xy - wz), 1.0f - 2.0f * (x2 + z2), 2.0f * (yz + wx)},
             {2.0f * (xz + wy), 2.0f * (yz - wx), 1.0f - 2.0f * (x2 + y2)}}};
}
// helpers.wgsl:119-122
static vec2 project_pix(const float fxfy[2], vec3 p_view, const float pp[2]) {
    float px = p_view.x / p_view.z, py = p_view.y / p_view.z;
    return {px * fxfy[0] + pp[0], py * fxfy[1] + pp[1]};
}
// helpers.wgsl:124-158
static vec3 calc_cov2d(const Uniforms& u, vec3 p_view, vec3 scale, vec4 quat) {
    float fx = u.focal[0], fy = u.focal[1];
    float iw = (float)u.img_size[0], ih = (float)u.img_size[1];
    float tan_fov_x = (0.5f * iw) / fx, tan_fov_y = (0.5f * ih) / fy;
    float lims_pos_x = (iw - u.pixel_center[0]) / fx + 0.3f * tan_fov_x;
    float lims_pos_y = (ih - u.pixel_center[1]) / fy + 0.3f * tan_fov_y;
    float lims_neg_x = u.pixel_center[0] / fx + 0.3f * tan_fov_x;
    float lims_neg_y = u.pixel_center[1] / fy + 0.3f * tan_fov_y;
    float rz = 1.0f / p_view.z;
    float rz2 = rz * rz;
    float tx = p_view.z * clampf(p_view.x * rz, -lims_neg_x, lims_pos_x);
    float ty = p_view.z * clampf(p_view.y * rz, -lims_neg_y, lims_pos_y);
    mat3 M = quat_to_rotmat(quat);
    M.c[0] = M.c[0] * scale.x;
    M.c[1] = M.c[1] * scale.y;
    M.c[2] = M.c[2] * scale.z;
    mat3x2 J = {{{fx * rz, 0.0f}, {0.0f, fy * rz}, {((-fx) * tx) * rz2, ((-fy) * ty) * rz2}}};
    mat3 W = view_W(u);
    mat3 V = mul(M, transpose(M));
    mat3x2 T = mul(J, W);
    mat3x2 TV = mul(T, V);
    // cov = TV * transpose(T): cov[c] = TV * (T[0][c], T[1][c], T[2][c])
    vec2 cov0 = mul(TV, vec3{T.c[0].x, T.c[1].x, T.c[2].x});
    vec2 cov1 = mul(TV, vec3{T.c[0].y, T.c[1].y, T.c[2].y});
    float c00 = cov0.x + COV_BLUR;
    float c11 = cov1.y + COV_BLUR;
    float c01 = cov0.y;
    return {c00, c01, c11};
}
// helpers.wgsl:160-164
static vec3 cov_to_conic(vec3 cov2d) {
    float det = cov2d.x * cov2d.z - cov2d.y * cov2d.y;
    float inv_det = 1.0f / det;
    return {cov2d.z * inv_det, (-cov2d.y) * inv_det, cov2d.x * inv_det};
}
// helpers.wgsl:192-201 (the opac argument is unused there)
static uint32_t radius_from_conic(vec3 conic) {
    float det = 1.0f / (conic.x * conic.z - conic.y * conic.y);
    float cx = conic.z * det, cz = conic.x * det;
    float b = 0.5f * (cx + cz);
    float s = sqrtf(fmaxw(0.1f, b * b - det));
    float v1 = b + s;
    float v2 = b - s;
    float radius = 3.0f * sqrtf(fmaxw(0.0f, fmaxw(v1, v2)));
    return sat_f2u(ceilf(radius));
}
// helpers.wgsl:220-236
static bool check_edge(vec2 p1, vec2 p2, vec2 center, const mat2& conic) {
    vec2 edge = p2 - p1;
    vec2 f = p1 - center;
    float a = dot(rowmul(edge, conic), edge);
    float b = 2.0f * dot(rowmul(f, conic), edge);
    float c = dot(rowmul(f, conic), f) - 1.0f;
    float disc = b * b - (4.0f * a) * c;
    if (disc < 0.0f) return false;
    float sq = sqrtf(disc);
    float t1 = ((-b) - sq) / (2.0f * a);
    float t2 = ((-b) + sq) / (2.0f * a);
    return (t1 >= 0.0f && t1 <= 1.0f) || (t2 >= 0.0f && t2 <= 1.0f);
}
static inline float signw(float v) { return v > 0.0f ? 1.0f : (v < 0.0f ? -1.0f : 0.0f); }
// helpers.wgsl:238-262
static bool ellipse_intersects_aabb(vec2 box_pos, vec2 box_extent, vec2 center, const mat2& conic) {
    vec2 d = center - box_pos;
    if (fabsf(d.x) <= box_extent.x && fabsf(d.y) <= box_extent.y) return true;
    vec2 cs = {signw(d.x), signw(d.y)};
    vec2 nearest = {box_pos.x + cs.x * box_extent.x, box_pos.y + cs.y * box_extent.y};
    vec2 cp = nearest - center;
    if (dot(rowmul(cp, conic), cp) <= 1.0f) return true;
    vec2 e1 = {nearest.x - (cs.x * 2.0f) * box_extent.x, nearest.y - 0.0f};
    vec2 e2 = {nearest.x - 0.0f, nearest.y - (cs.y * 2.0f) * box_extent.y};
    return check_edge(nearest, e1, center, conic) || check_edge(nearest, e2, center, conic);
}
// helpers.wgsl:264-279
static bool can_be_visible(uint32_t tx, uint32_t ty, vec2 xy, vec3 conic, float opac) {
    float sigma = det_logf(opac * 255.0f);
    if (sigma <= 0.0f) return false;
    float s2 = 2.0f * sigma;
    vec3 cs = {conic.x / s2, conic.y / s2, conic.z / s2};
    vec2 ext = {(float)TILE_WIDTH / 2.0f, (float)TILE_WIDTH / 2.0f};
    vec2 tc = {(float)(tx * TILE_WIDTH) + ext.x, (float)(ty * TILE_WIDTH) + ext.y};
    mat2 m = {{{cs.x, cs.y}, {cs.y, cs.z}}};
    return ellipse_intersects_aabb(tc, ext, xy, m);
}
static inline float sigmoidf(float x) { return 1.0f / (1.0f + det_expf(-x)); }  // project_visible.wgsl:149-151
static inline vec3 normalize3(vec3 v) {
    float l = sqrtf(dot(v, v));
    return {v.x / l, v.y / l, v.z / l};
}

// ---------------------------------------------------------------------------
// Spherical harmonics.  project_visible.wgsl:51-147 (forward) and
// gather_grads.wgsl:17-112 (VJP w.r.t. coefficients; no gradient to viewdir).
// ---------------------------------------------------------------------------
static vec3 sh_coeffs_to_color(uint32_t degree, vec3 d, const float* sh /* [K][3] */) {
    auto C = [&](int k) { return vec3{sh[3 * k], sh[3 * k + 1], sh[3 * k + 2]}; };
    vec3 colors = 0.2820947917738781f * C(0);
    if (degree == 0) return colors;
    float x = d.x, y = d.y, z = d.z;
    float fTmp0A = 0.48860251190292f;
    colors = colors + fTmp0A * ((((-y) * C(1)) + (z * C(2))) - (x * C(3)));
    if (degree == 1) return colors;
    float z2 = z * z;
    float fTmp0B = -1.092548430592079f * z;
    float fTmp1A = 0.5462742152960395f;
    float fC1 = x * x - y * y;
    float fS1 = 2.f * x * y;
    float pSH6 = (0.9461746957575601f * z2 - 0.3153915652525201f);
    float pSH7 = fTmp0B * x;
    float pSH5 = fTmp0B * y;
    float pSH8 = fTmp1A * fC1;
    float pSH4 = fTmp1A * fS1;
    colors = colors + ((((pSH4 * C(4) + pSH5 * C(5)) + pSH6 * C(6)) + pSH7 * C(7)) + pSH8 * C(8));
    if (degree == 2) return colors;
    float fTmp0C = -2.285228997322329f * z2 + 0.4570457994644658f;
    float fTmp1B = 1.445305721320277f * z;
    float fTmp2A = -0.5900435899266435f;
    float fC2 = x * fC1 - y * fS1;
    float fS2 = x * fS1 + y * fC1;
    float pSH12 = z * (1.865881662950577f * z2 - 1.119528997770346f);
    float pSH13 = fTmp0C * x;
    float pSH11 = fTmp0C * y;
    float pSH14 = fTmp1B * fC1;
    float pSH10 = fTmp1B * fS1;
    float pSH15 = fTmp2A * fC2;
    float pSH9 = fTmp2A * fS2;
    colors = colors + ((((((pSH9 * C(9) + pSH10 * C(10)) + pSH11 * C(11)) + pSH12 * C(12)) + pSH13 * C(13)) +
                        pSH14 * C(14)) + pSH15 * C(15));
    if (degree == 3) return colors;
    float fTmp0D = z * (-4.683325804901025f * z2 + 2.007139630671868f);
    float fTmp1C = 3.31161143515146f * z2 - 0.47308734787878f;
    float fTmp2B = -1.770130769779931f * z;
    float fTmp3A = 0.6258357354491763f;
    float fC3 = x * fC2 - y * fS2;
    float fS3 = x * fS2 + y * fC2;
    float pSH20 = (1.984313483298443f * z * pSH12 - 1.006230589874905f * pSH6);
    float pSH21 = fTmp0D * x;
    float pSH19 = fTmp0D * y;
    float pSH22 = fTmp1C * fC1;
    float pSH18 = fTmp1C * fS1;
    float pSH23 = fTmp2B * fC2;
    float pSH17 = fTmp2B * fS2;
    float pSH24 = fTmp3A * fC3;
    float pSH16 = fTmp3A * fS3;
    colors = colors + ((((((((pSH16 * C(16) + pSH17 * C(17)) + pSH18 * C(18)) + pSH19 * C(19)) + pSH20 * C(20)) +
                          pSH21 * C(21)) + pSH22 * C(22)) + pSH23 * C(23)) + pSH24 * C(24));
    return colors;
}

// basis[k] such that v_coeffs[k] = basis[k] * v_color (gather_grads.wgsl:17-112).
static void sh_basis(uint32_t degree, vec3 d, float* b /* [25] */) {
    b[0] = 0.2820947917738781f;
    if (degree == 0) return;
    float x = d.x, y = d.y, z = d.z;
    float fTmp0A = 0.48860251190292f;
    b[1] = (-fTmp0A) * y;
    b[2] = fTmp0A * z;
    b[3] = (-fTmp0A) * x;
    if (degree == 1) return;
    float z2 = z * z;
    float fTmp0B = -1.092548430592079f * z;
    float fTmp1A = 0.5462742152960395f;
    float fC1 = x * x - y * y;
    float fS1 = 2.f * x * y;
    b[6] = (0.9461746957575601f * z2 - 0.3153915652525201f);
    b[7] = fTmp0B * x;
    b[5] = fTmp0B * y;
    b[8] = fTmp1A * fC1;
    b[4] = fTmp1A * fS1;
    if (degree == 2) return;
    float fTmp0C = -2.285228997322329f * z2 + 0.4570457994644658f;
    float fTmp1B = 1.445305721320277f * z;
    float fTmp2A = -0.5900435899266435f;
    float fC2 = x * fC1 - y * fS1;
    float fS2 = x * fS1 + y * fC1;
    b[12] = z * (1.865881662950577f * z2 - 1.119528997770346f);
    b[13] = fTmp0C * x;
    b[11] = fTmp0C * y;
    b[14] = fTmp1B * fC1;
    b[10] = fTmp1B * fS1;
    b[15] = fTmp2A * fC2;
    b[9] = fTmp2A * fS2;
    if (degree == 3) return;
    float fTmp0D = z * (-4.683325804901025f * z2 + 2.007139630671868f);
    float fTmp1C = 3.31161143515146f * z2 - 0.47308734787878f;
    float fTmp2B = -1.770130769779931f * z;
    float fTmp3A = 0.6258357354491763f;
    float fC3 = x * fC2 - y * fS2;
    float fS3 = x * fS2 + y * fC2;
    b[20] = (1.984313483298443f * z * b[12] + -1.006230589874905f * b[6]);
    b[21] = fTmp0D * x;
    b[19] = fTmp0D * y;
    b[22] = fTmp1C * fC1;
    b[18] = fTmp1C * fS1;
    b[23] = fTmp2B * fC2;
    b[17] = fTmp2B * fS2;
    b[24] = fTmp3A * fC3;
    b[16] = fTmp3A * fS3;
}

static int threads_env() {
    const char* e = getenv("BRUSH_ORACLE_THREADS");
    int t = e ? atoi(e) : 1;
    return t < 1 ? 1 : t;
}
#ifdef _OPENMP
#define OMP_FOR _Pragma("omp parallel for schedule(dynamic, 64) num_threads(g_threads)")
#else
#define OMP_FOR
#endif
static int g_threads = 1;
// Wall-clock seconds of each stage of the last oracle_render_forward / oracle_render_backward call, in the order of
// the reference's tracing spans (render.rs:127-267,515-577): ProjectSplats, DepthSort, ProjectVisibile, PrefixSum,
// MapGaussiansToIntersect, Tile sort, GetTileBinEdges, Rasterize, RasterizeBackwards, GatherGrads, ProjectBackwards.
static double g_stage_s[11];
static inline double now_s() {
#ifdef _OPENMP
    return omp_get_wtime();
#else
    return (double)clock() / CLOCKS_PER_SEC;
#endif
}
struct StageClock {
    int i;
    double t0;
    explicit StageClock(int i_) : i(i_), t0(now_s()) {}
    ~StageClock() { g_stage_s[i] = now_s() - t0; }
};

extern "C" {
void oracle_stage_seconds(double* out11) {
    for (int i = 0; i < 11; i++) out11[i] = g_stage_s[i];
}

// Number of worker threads for the timed CPU baseline. 1 (default) = the strictly
// serial parity oracle.
void oracle_set_threads(int t) { g_threads = t < 1 ? 1 : t; }
int oracle_get_threads() { return g_threads; }
int oracle_max_threads() {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}
float oracle_det_expf(float x) { return det_expf(x); }
float oracle_det_logf(float x) { return det_logf(x); }

// ---------------------------------------------------------------------------
// project_forward.wgsl:15-68.  Compaction order: the reference uses a global
// atomicAdd (nondeterministic); the canonical order here is ascending gid.
// ---------------------------------------------------------------------------
int oracle_project_forward(const Uniforms* up, const float* means, const float* log_scales, const float* quats,
                           uint32_t* global_from_presort_gid, float* depths, uint32_t* num_visible) {
    const Uniforms& u = *up;
    uint32_t N = u.total_splats;
    std::vector<float> z(N);
    std::vector<uint8_t> vis(N);
    mat3 W = view_W(u);
    vec3 t = view_t(u);
    OMP_FOR
    for (int64_t g = 0; g < (int64_t)N; g++) {
        vis[g] = 0;
        vec3 mean = ld3(means, g);
        vec3 p_view = mul(W, mean) + t;
        if (p_view.z <= 0.01f) continue;
        vec3 scale = exp3(ld3(log_scales, g));
        vec4 quat = ld4(quats, g);
        vec3 cov2d = calc_cov2d(u, p_view, scale, quat);
        float det = cov2d.x * cov2d.z - cov2d.y * cov2d.y;
        if (det == 0.0f) continue;
        vec3 conic = cov_to_conic(cov2d);
        vec2 xy = project_pix(u.focal, p_view, u.pixel_center);
        uint32_t radius = radius_from_conic(conic);
        bbox4 b = get_tile_bbox(xy, radius, u.tile_bounds);
        if ((b.maxx - b.minx) == 0u || (b.maxy - b.miny) == 0u) continue;
        vis[g] = 1;
        z[g] = p_view.z;
    }
    uint32_t w = 0;
    for (uint32_t g = 0; g < N; g++) {
        if (vis[g]) {
            global_from_presort_gid[w] = g;
            depths[w] = z[g];
            w++;
        }
    }
    *num_visible = w;
    return 0;
}

// ---------------------------------------------------------------------------
// brush-sort/src/lib.rs:32-147 + shaders: LSD radix sort, 4 bits per pass,
// ceil(bits/4) passes, stable, only the first n entries take part.
// ---------------------------------------------------------------------------
int oracle_radix_argsort_u32(const uint32_t* keys, const uint32_t* vals, uint32_t n, uint32_t bits,
                             uint32_t* keys_out, uint32_t* vals_out) {
    if (bits > 32) return -1;
    std::vector<uint32_t> k0(keys, keys + n), v0(vals, vals + n), k1(n), v1(n);
#ifdef ORACLE_TIMED
    // timed build: the same stable LSD sort over the same 4*ceil(bits/4) key bits, 8 bits per pass
    const uint32_t rb = ((bits + 3) / 4) * 4;
    for (uint32_t shift = 0; shift < rb; shift += 8) {
        const uint32_t nb = rb - shift < 8 ? rb - shift : 8, mask = (1u << nb) - 1u;
        size_t cnt[257] = {0};
        for (uint32_t i = 0; i < n; i++) cnt[((uint32_t)(k0[i] >> shift) & mask) + 1]++;
        for (int b = 0; b < 256; b++) cnt[b + 1] += cnt[b];
        for (uint32_t i = 0; i < n; i++) {
            size_t d = cnt[(uint32_t)(k0[i] >> shift) & mask]++;
            k1[d] = k0[i];
            v1[d] = v0[i];
        }
        k0.swap(k1);
        v0.swap(v1);
    }
#else
    uint32_t passes = (bits + 3) / 4;
    for (uint32_t pass = 0; pass < passes; pass++) {
        uint32_t shift = pass * 4;
        size_t cnt[17] = {0};
        for (uint32_t i = 0; i < n; i++) cnt[((k0[i] >> shift) & 0xF) + 1]++;
        for (int b = 0; b < 16; b++) cnt[b + 1] += cnt[b];
        for (uint32_t i = 0; i < n; i++) {
            size_t d = cnt[(k0[i] >> shift) & 0xF]++;
            k1[d] = k0[i];
            v1[d] = v0[i];
        }
        k0.swap(k1);
        v0.swap(v1);
    }
#endif
    std::copy(k0.begin(), k0.end(), keys_out);
    std::copy(v0.begin(), v0.end(), vals_out);
    return 0;
}

// 64-bit-key variant (north-star fused (tile<<32 | depth_bits) key): same LSD algorithm.
int oracle_radix_argsort_u64(const uint64_t* keys, const uint32_t* vals, uint32_t n, uint32_t bits,
                             uint64_t* keys_out, uint32_t* vals_out) {
    if (bits > 64) return -1;
    std::vector<uint64_t> k0(keys, keys + n), k1(n);
    std::vector<uint32_t> v0(vals, vals + n), v1(n);
#ifdef ORACLE_TIMED
    // timed build: the same stable LSD sort over the same 4*ceil(bits/4) key bits, 8 bits per pass
    const uint32_t rb = ((bits + 3) / 4) * 4;
    for (uint32_t shift = 0; shift < rb; shift += 8) {
        const uint32_t nb = rb - shift < 8 ? rb - shift : 8, mask = (1u << nb) - 1u;
        size_t cnt[257] = {0};
        for (uint32_t i = 0; i < n; i++) cnt[((uint32_t)(k0[i] >> shift) & mask) + 1]++;
        for (int b = 0; b < 256; b++) cnt[b + 1] += cnt[b];
        for (uint32_t i = 0; i < n; i++) {
            size_t d = cnt[(uint32_t)(k0[i] >> shift) & mask]++;
            k1[d] = k0[i];
            v1[d] = v0[i];
        }
        k0.swap(k1);
        v0.swap(v1);
    }
#else
    uint32_t passes = (bits + 3) / 4;
    for (uint32_t pass = 0; pass < passes; pass++) {
        uint32_t shift = pass * 4;
        size_t cnt[17] = {0};
        for (uint32_t i = 0; i < n; i++) cnt[((k0[i] >> shift) & 0xF) + 1]++;
        for (int b = 0; b < 16; b++) cnt[b + 1] += cnt[b];
        for (uint32_t i = 0; i < n; i++) {
            size_t d = cnt[(k0[i] >> shift) & 0xF]++;
            k1[d] = k0[i];
            v1[d] = v0[i];
        }
        k0.swap(k1);
        v0.swap(v1);
    }
#endif
    std::copy(k0.begin(), k0.end(), keys_out);
    std::copy(v0.begin(), v0.end(), vals_out);
    return 0;
}

// brush-prefix-sum/src/lib.rs:17-102: inclusive u32 scan (wrapping).
int oracle_prefix_sum_u32(const uint32_t* in, uint32_t n, uint32_t* out) {
    uint32_t acc = 0;
    for (uint32_t i = 0; i < n; i++) {
        acc += in[i];
        out[i] = acc;
    }
    return 0;
}

// ---------------------------------------------------------------------------
// project_visible.wgsl:163-258.  projected[c] = (x, y, cxx, cxy, cyy, r, g, b, opac)
// ---------------------------------------------------------------------------
int oracle_project_visible(const Uniforms* up, uint32_t K, const float* means, const float* log_scales,
                           const float* quats, const float* coeffs, const float* raw_opacities,
                           const uint32_t* global_from_compact_gid, uint32_t num_visible, float* projected,
                           uint32_t* num_tiles_hit) {
    const Uniforms& u = *up;
    mat3 W = view_W(u);
    vec3 t = view_t(u);
    OMP_FOR
    for (int64_t c = 0; c < (int64_t)num_visible; c++) {
        uint32_t g = global_from_compact_gid[c];
        vec3 mean = ld3(means, g);
        vec3 scale = exp3(ld3(log_scales, g));
        vec4 quat = ld4(quats, g);
        float opac = sigmoidf(raw_opacities[g]);
        vec3 p_view = mul(W, mean) + t;
        vec3 cov2d = calc_cov2d(u, p_view, scale, quat);
        vec3 conic = cov_to_conic(cov2d);
        vec2 xy = project_pix(u.focal, p_view, u.pixel_center);
        vec3 viewdir = normalize3(mean - t);  // project_visible.wgsl:232-233: "camera_pos" = viewmat[3].xyz
        vec3 color = sh_coeffs_to_color(u.sh_degree, viewdir, coeffs + (size_t)g * K * 3);
        color = {color.x + 0.5f, color.y + 0.5f, color.z + 0.5f};
        uint32_t radius = radius_from_conic(conic);
        bbox4 b = get_tile_bbox(xy, radius, u.tile_bounds);
        uint32_t area = 0;
        for (uint32_t ty = b.miny; ty < b.maxy; ty++)
            for (uint32_t tx = b.minx; tx < b.maxx; tx++)
                if (can_be_visible(tx, ty, xy, conic, opac)) area++;
        float* p = projected + (size_t)c * 9;
        p[0] = xy.x; p[1] = xy.y;
        p[2] = conic.x; p[3] = conic.y; p[4] = conic.z;
        p[5] = color.x; p[6] = color.y; p[7] = color.z; p[8] = opac;
        num_tiles_hit[c] = area;
    }
    return 0;
}

// map_gaussian_to_intersects.wgsl:10-48.  `capacity` = arrayLength(tile_id_from_isect).
int oracle_map_gaussian_to_intersects(const Uniforms* up, uint32_t num_visible, const float* projected,
                                      const uint32_t* cum_tiles_hit, uint64_t capacity, uint32_t* tile_id_from_isect,
                                      uint32_t* compact_gid_from_isect) {
    const Uniforms& u = *up;
    OMP_FOR
    for (int64_t c = 0; c < (int64_t)num_visible; c++) {
        const float* p = projected + (size_t)c * 9;
        vec2 xy = {p[0], p[1]};
        vec3 conic = {p[2], p[3], p[4]};
        float opac = p[8];
        uint32_t radius = radius_from_conic(conic);
        bbox4 b = get_tile_bbox(xy, radius, u.tile_bounds);
        uint64_t isect_id = c > 0 ? cum_tiles_hit[c - 1] : 0u;
        for (uint32_t ty = b.miny; ty < b.maxy; ty++)
            for (uint32_t tx = b.minx; tx < b.maxx; tx++)
                if (can_be_visible(tx, ty, xy, conic, opac) && isect_id < capacity) {
                    tile_id_from_isect[isect_id] = tx + ty * u.tile_bounds[0];
                    compact_gid_from_isect[isect_id] = (uint32_t)c;
                    isect_id++;
                }
    }
    return 0;
}

// get_tile_bin_edges.wgsl:15-42.  tile_bins must be pre-zeroed (render.rs:241-244).
int oracle_get_tile_bin_edges(const uint32_t* sorted_tile_ids, uint32_t num_intersections, uint32_t* tile_bins) {
    for (uint32_t i = 0; i < num_intersections; i++) {
        uint32_t cur = sorted_tile_ids[i];
        if (i == num_intersections - 1) tile_bins[cur * 2 + 1] = num_intersections;
        if (i == 0) {
            tile_bins[cur * 2 + 0] = 0u;
        } else {
            uint32_t prev = sorted_tile_ids[i - 1];
            if (prev != cur) {
                tile_bins[prev * 2 + 1] = i;
                tile_bins[cur * 2 + 0] = i;
            }
        }
    }
    return 0;
}

// ---------------------------------------------------------------------------
// rasterize.wgsl:20-115.  out_img: f32[H,W,4] or packed u32[H,W] (raster_u32).
// The batching through workgroup memory does not change per-pixel results, so
// each pixel simply walks its tile's range front to back.
// ---------------------------------------------------------------------------
int oracle_rasterize(const Uniforms* up, const uint32_t* compact_gid_from_isect, const uint32_t* tile_bins,
                     const float* projected, int raster_u32, void* out_img, uint32_t* final_index) {
    const Uniforms& u = *up;
    uint32_t Wd = u.img_size[0], Hd = u.img_size[1];
    int64_t ntiles = (int64_t)u.tile_bounds[0] * u.tile_bounds[1];
#ifdef _OPENMP
#pragma omp parallel for schedule(dynamic, 2) num_threads(g_threads)
#endif
    for (int64_t tile_id = 0; tile_id < ntiles; tile_id++) {
        uint32_t tx = tile_id % u.tile_bounds[0], ty = tile_id / u.tile_bounds[0];
        uint32_t r0 = tile_bins[tile_id * 2], r1 = tile_bins[tile_id * 2 + 1];
        for (uint32_t ly = 0; ly < TILE_WIDTH; ly++)
            for (uint32_t lx = 0; lx < TILE_WIDTH; lx++) {
                uint32_t px = tx * TILE_WIDTH + lx, py = ty * TILE_WIDTH + ly;
                if (!(px < Wd && py < Hd)) continue;
                float pcx = (float)px + 0.5f, pcy = (float)py + 0.5f;
                float T = 1.0f;
                vec3 pix = {0, 0, 0};
                uint32_t final_idx = 0;
                for (uint32_t i = r0; i < r1; i++) {
                    const float* p = projected + (size_t)compact_gid_from_isect[i] * 9;
                    float dx = p[0] - pcx, dy = p[1] - pcy;
                    float sigma = 0.5f * (p[2] * dx * dx + p[4] * dy * dy) + p[3] * dx * dy;
                    float vis = expf(-sigma);
                    float alpha = fminf(0.999f, p[8] * vis);
                    if (sigma >= 0.0f && alpha >= 1.0f / 255.0f) {
                        float next_T = T * (1.0f - alpha);
                        if (next_T <= 1e-4f) break;
                        float fac = alpha * T;
                        pix = {pix.x + p[5] * fac, pix.y + p[6] * fac, pix.z + p[7] * fac};
                        T = next_T;
                        final_idx = i;
                    }
                }
                size_t pix_id = (size_t)px + (size_t)py * Wd;
                float fc[4] = {pix.x, pix.y, pix.z, 1.0f - T};
                if (raster_u32) {
                    uint32_t cu[4];
                    for (int k = 0; k < 4; k++) cu[k] = sat_f2u(clampf(fc[k] * 255.0f, 0.0f, 255.0f));
                    ((uint32_t*)out_img)[pix_id] = cu[0] | (cu[1] << 8) | (cu[2] << 16) | (cu[3] << 24);
                } else {
                    float* o = (float*)out_img + pix_id * 4;
                    o[0] = fc[0]; o[1] = fc[1]; o[2] = fc[2]; o[3] = fc[3];
                    final_index[pix_id] = final_idx;
                }
            }
    }
    return 0;
}

// ---------------------------------------------------------------------------
// rasterize_backwards.wgsl:140-304.  Per-(tile, splat) sums are formed in f32 in
// pixel order (the reference's subgroupAdd) and accumulated per compact id in
// f64 (the reference's f32 atomics have run-dependent order; f64 keeps the
// oracle independent of thread count).  Outputs are indexed by COMPACT gid:
// v_xy[V,2], v_conic[V,3], v_colors[V,4] and must be zeroed by the caller.
// ---------------------------------------------------------------------------
int oracle_rasterize_backwards(const Uniforms* up, const uint32_t* compact_gid_from_isect, const uint32_t* tile_bins,
                               const float* projected, const uint32_t* final_index, const float* output,
                               const float* v_output, uint32_t num_points, float* v_xy, float* v_conic,
                               float* v_colors) {
    const Uniforms& u = *up;
    uint32_t Wd = u.img_size[0], Hd = u.img_size[1];
    int64_t ntiles = (int64_t)u.tile_bounds[0] * u.tile_bounds[1];
#ifdef ORACLE_TIMED
    // timed build: one f32 accumulator array per thread (no atomics), summed per splat afterwards
    const int nthr = g_threads;
    const size_t acc_rows = u.num_visible < num_points ? u.num_visible : num_points;  // rows are compact ids < V
    const size_t acc_len = acc_rows * 9;
    std::vector<std::vector<float>> acc_thr(nthr);
#else
    std::vector<double> acc((size_t)num_points * 9, 0.0);
#endif
#ifdef _OPENMP
#pragma omp parallel num_threads(g_threads)
#endif
    {
#ifdef ORACLE_TIMED
#ifdef _OPENMP
    std::vector<float>& acc = acc_thr[omp_get_thread_num()];
#else
    std::vector<float>& acc = acc_thr[0];
#endif
    acc.assign(acc_len, 0.0f);
#endif
#ifdef _OPENMP
#pragma omp for schedule(dynamic, 2)
#endif
    for (int64_t tile_id = 0; tile_id < ntiles; tile_id++) {
        uint32_t tx = tile_id % u.tile_bounds[0], ty = tile_id / u.tile_bounds[0];
        uint32_t r0 = tile_bins[tile_id * 2], r1 = tile_bins[tile_id * 2 + 1];
        if (r1 <= r0) continue;
        // per-pixel running state
        float T[TILE_SIZE], Tf[TILE_SIZE], buf[TILE_SIZE][3], vo[TILE_SIZE][4];
        uint32_t fin[TILE_SIZE];
        bool inside[TILE_SIZE];
        for (uint32_t l = 0; l < TILE_SIZE; l++) {
            uint32_t px = tx * TILE_WIDTH + l % TILE_WIDTH, py = ty * TILE_WIDTH + l / TILE_WIDTH;
            inside[l] = px < Wd && py < Hd;
            buf[l][0] = buf[l][1] = buf[l][2] = 0.0f;
            vo[l][0] = vo[l][1] = vo[l][2] = vo[l][3] = 0.0f;
            fin[l] = 0; T[l] = Tf[l] = 1.0f;
            if (inside[l]) {
                size_t pix_id = (size_t)px + (size_t)py * Wd;
                Tf[l] = 1.0f - output[pix_id * 4 + 3];
                T[l] = Tf[l];
                fin[l] = final_index[pix_id];
                for (int k = 0; k < 4; k++) vo[l][k] = v_output[pix_id * 4 + k];
            }
        }
        for (uint32_t i = r1; i-- > r0;) {  // back to front
            uint32_t cg = compact_gid_from_isect[i];
            const float* p = projected + (size_t)cg * 9;
            float s[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
            bool any = false;
            for (uint32_t l = 0; l < TILE_SIZE; l++) {
                if (!(inside[l] && i <= fin[l])) continue;
                float pcx = (float)(tx * TILE_WIDTH + l % TILE_WIDTH) + 0.5f;
                float pcy = (float)(ty * TILE_WIDTH + l / TILE_WIDTH) + 0.5f;
                float dx = p[0] - pcx, dy = p[1] - pcy;
                float sigma = 0.5f * (p[2] * dx * dx + p[4] * dy * dy) + p[3] * dx * dy;
                float vis = expf(-sigma);
                float alpha = fminf(0.99f, p[8] * vis);
                if (!(sigma >= 0.0f && alpha >= 1.0f / 255.0f)) continue;
                any = true;
                float ra = 1.0f / (1.0f - alpha);
                T[l] *= ra;
                float fac = alpha * T[l];
                float v_alpha = (p[5] * T[l] - buf[l][0] * ra) * vo[l][0] + (p[6] * T[l] - buf[l][1] * ra) * vo[l][1] +
                                (p[7] * T[l] - buf[l][2] * ra) * vo[l][2];
                v_alpha += Tf[l] * ra * vo[l][3];
                buf[l][0] += p[5] * fac; buf[l][1] += p[6] * fac; buf[l][2] += p[7] * fac;
                float v_sigma = -p[8] * vis * v_alpha;
                s[0] += v_sigma * (p[2] * dx + p[3] * dy);
                s[1] += v_sigma * (p[3] * dx + p[4] * dy);
                s[2] += 0.5f * v_sigma * dx * dx;
                s[3] += v_sigma * dx * dy;
                s[4] += 0.5f * v_sigma * dy * dy;
                s[5] += fac * vo[l][0];
                s[6] += fac * vo[l][1];
                s[7] += fac * vo[l][2];
                s[8] += vis * v_alpha;
            }
            if (any) {
#ifdef ORACLE_TIMED
                float* a = &acc[(size_t)cg * 9];
                for (int k = 0; k < 9; k++) a[k] += s[k];
#else
                double* a = &acc[(size_t)cg * 9];
                for (int k = 0; k < 9; k++) {
#ifdef _OPENMP
#pragma omp atomic
#endif
                    a[k] += (double)s[k];
                }
#endif
            }
        }
    }
    }  // omp parallel
#ifdef ORACLE_TIMED
    OMP_FOR
    for (int64_t c = 0; c < (int64_t)acc_rows; c++) {
        float a[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
        for (int t = 0; t < nthr; t++)
            if (!acc_thr[t].empty())
                for (int k = 0; k < 9; k++) a[k] += acc_thr[t][(size_t)c * 9 + k];
        v_xy[c * 2] += a[0]; v_xy[c * 2 + 1] += a[1];
        v_conic[c * 3] += a[2]; v_conic[c * 3 + 1] += a[3]; v_conic[c * 3 + 2] += a[4];
        for (int k = 0; k < 4; k++) v_colors[c * 4 + k] += a[5 + k];
    }
#else
    for (size_t c = 0; c < num_points; c++) {
        const double* a = &acc[c * 9];
        v_xy[c * 2] += (float)a[0]; v_xy[c * 2 + 1] += (float)a[1];
        v_conic[c * 3] += (float)a[2]; v_conic[c * 3 + 1] += (float)a[3]; v_conic[c * 3 + 2] += (float)a[4];
        for (int k = 0; k < 4; k++) v_colors[c * 4 + k] += (float)a[5 + k];
    }
#endif
    return 0;
}

// ---------------------------------------------------------------------------
// gather_grads.wgsl:165-232.  v_coeffs[N,K,3], v_opacs[N], v_xy_global[N,2]
// are dense and pre-zeroed (render.rs:539-547).
// ---------------------------------------------------------------------------
int oracle_gather_grads(const Uniforms* up, uint32_t K, const uint32_t* global_from_compact_gid, uint32_t num_visible,
                        const float* raw_opacities, const float* means, const float* v_colors,
                        const float* v_xy_local, float* v_coeffs, float* v_opacs, float* v_xy_global) {
    const Uniforms& u = *up;
    vec3 t = view_t(u);
    OMP_FOR
    for (int64_t c = 0; c < (int64_t)num_visible; c++) {
        uint32_t g = global_from_compact_gid[c];
        vec3 vc = {v_colors[c * 4], v_colors[c * 4 + 1], v_colors[c * 4 + 2]};
        float v_alpha = v_colors[c * 4 + 3];
        vec3 viewdir = normalize3(ld3(means, g) - t);
        float b[25];
        sh_basis(u.sh_degree, viewdir, b);
        float* out = v_coeffs + (size_t)g * K * 3;
        for (uint32_t k = 0; k < K; k++) {
            out[3 * k] = b[k] * vc.x;
            out[3 * k + 1] = b[k] * vc.y;
            out[3 * k + 2] = b[k] * vc.z;
        }
        float s = sigmoidf(raw_opacities[g]);
        v_opacs[g] = v_alpha * (s * (1.0f - s));
        v_xy_global[g * 2] = v_xy_local[c * 2];
        v_xy_global[g * 2 + 1] = v_xy_local[c * 2 + 1];
    }
    return 0;
}

// project_backwards.wgsl:19-23
static vec3 project_pix_vjp(const float fxfy[2], vec3 p_view, vec2 v_xy) {
    float rw = 1.0f / (p_view.z + 1e-6f);
    float vpx = fxfy[0] * v_xy.x, vpy = fxfy[1] * v_xy.y;
    return {vpx * rw, vpy * rw, (-(vpx * p_view.x + vpy * p_view.y) * rw) * rw};
}
// project_backwards.wgsl:25-57; v_R.c[i] = column i, .x/.y/.z = rows 0/1/2 (v_R[i][j] = c[i].{j})
static vec4 quat_to_rotmat_vjp(vec4 q, const mat3& R) {
    float w = q.x, x = q.y, y = q.z, z = q.w;
    auto r = [&](int i, int j) { const vec3& c = R.c[i]; return j == 0 ? c.x : (j == 1 ? c.y : c.z); };
    return {2.0f * (x * (r(1, 2) - r(2, 1)) + y * (r(2, 0) - r(0, 2)) + z * (r(0, 1) - r(1, 0))),
            2.0f * (-2.f * x * (r(1, 1) + r(2, 2)) + y * (r(0, 1) + r(1, 0)) + z * (r(0, 2) + r(2, 0)) +
                    w * (r(1, 2) - r(2, 1))),
            2.0f * (x * (r(0, 1) + r(1, 0)) - 2.f * y * (r(0, 0) + r(2, 2)) + z * (r(1, 2) + r(2, 1)) +
                    w * (r(2, 0) - r(0, 2))),
            2.0f * (x * (r(0, 2) + r(2, 0)) + y * (r(1, 2) + r(2, 1)) - 2.f * z * (r(0, 0) + r(1, 1)) +
                    w * (r(0, 1) - r(1, 0)))};
}
// project_backwards.wgsl:59-72
static vec3 cov2d_to_conic_vjp(vec3 conic, vec3 v_conic) {
    mat2 X = {{{conic.x, conic.y}, {conic.y, conic.z}}};
    mat2 G = {{{v_conic.x, v_conic.y / 2.0f}, {v_conic.y / 2.0f, v_conic.z}}};
    mat2 S = mul(mul(X, G), X);
    return {-S.c[0].x, -(S.c[1].x + S.c[0].y), -S.c[1].y};
}

// ---------------------------------------------------------------------------
// project_backwards.wgsl:75-227.  v_xys / v_conics are indexed by COMPACT gid;
// v_means[N,3], v_scales[N,3], v_quats[N,4] are dense and pre-zeroed.
// ---------------------------------------------------------------------------
int oracle_project_backwards(const Uniforms* up, const float* means, const float* log_scales, const float* quats,
                             const uint32_t* global_from_compact_gid, uint32_t num_visible, const float* v_xys,
                             const float* v_conics, float* v_means, float* v_scales, float* v_quats) {
    const Uniforms& u = *up;
    mat3 W = view_W(u);
    vec3 tr = view_t(u);
    float fx = u.focal[0], fy = u.focal[1];
    OMP_FOR
    for (int64_t c = 0; c < (int64_t)num_visible; c++) {
        vec3 v_conic = ld3(v_conics, c);
        vec2 v_xy = {v_xys[c * 2], v_xys[c * 2 + 1]};
        uint32_t g = global_from_compact_gid[c];
        vec3 mean = ld3(means, g);
        vec3 scale = exp3(ld3(log_scales, g));
        vec4 quat = ld4(quats, g);
        vec3 p_view = mul(W, mean) + tr;
        vec3 v_mean = mul(transpose(W), project_pix_vjp(u.focal, p_view, v_xy));
        vec3 cov2d = calc_cov2d(u, p_view, scale, quat);
        vec3 conic = cov_to_conic(cov2d);
        vec3 v_cov2d = cov2d_to_conic_vjp(conic, v_conic);
        float rz = 1.0f / p_view.z;
        float rz2 = rz * rz;
        mat3 J = {{{fx * rz, 0.0f, 0.0f},
                   {0.0f, fy * rz, 0.0f},
                   {(-fx) * p_view.x * rz2, (-fy) * p_view.y * rz2, 0.0f}}};
        mat3 R = quat_to_rotmat(quat);
        mat3 S = {{{scale.x, 0, 0}, {0, scale.y, 0}, {0, 0, scale.z}}};
        mat3 M = mul(R, S);
        mat3 V = mul(M, transpose(M));
        mat3 v_cov = {{{v_cov2d.x, 0.5f * v_cov2d.y, 0.0f}, {0.5f * v_cov2d.y, v_cov2d.z, 0.0f}, {0.0f, 0.0f, 0.0f}}};
        mat3 T = mul(J, W);
        mat3 Tt = transpose(T);
        mat3 Vt = transpose(V);
        mat3 v_V = mul(mul(Tt, v_cov), T);
        mat3 v_T = add(mul(mul(v_cov, T), Vt), mul(mul(transpose(v_cov), T), V));
        float v_cov3d0 = v_V.c[0].x;
        float v_cov3d1 = v_V.c[0].y + v_V.c[1].x;
        float v_cov3d2 = v_V.c[0].z + v_V.c[2].x;
        float v_cov3d3 = v_V.c[1].y;
        float v_cov3d4 = v_V.c[1].z + v_V.c[2].y;
        float v_cov3d5 = v_V.c[2].z;
        mat3 v_J = mul(v_T, transpose(W));
        float rz3 = rz2 * rz;
        vec3 v_t = {(-fx) * rz2 * v_J.c[2].x, (-fy) * rz2 * v_J.c[2].y,
                    (-fx) * rz2 * v_J.c[0].x + 2.0f * fx * p_view.x * rz3 * v_J.c[2].x - fy * rz2 * v_J.c[1].y +
                        2.0f * fy * p_view.y * rz3 * v_J.c[2].y};
        v_mean = v_mean + vec3{dot(v_t, W.c[0]), dot(v_t, W.c[1]), dot(v_t, W.c[2])};
        mat3 v_V_symm = {{{v_cov3d0, 0.5f * v_cov3d1, 0.5f * v_cov3d2},
                          {0.5f * v_cov3d1, v_cov3d3, 0.5f * v_cov3d4},
                          {0.5f * v_cov3d2, 0.5f * v_cov3d4, v_cov3d5}}};
        mat3 v_M = mul(scale_cols(v_V_symm, 2.0f), M);
        vec3 v_scale = {dot(R.c[0], v_M.c[0]), dot(R.c[1], v_M.c[1]), dot(R.c[2], v_M.c[2])};
        mat3 v_R = mul(v_M, S);
        vec4 v_quat = quat_to_rotmat_vjp(quat, v_R);
        v_means[g * 3] = v_mean.x; v_means[g * 3 + 1] = v_mean.y; v_means[g * 3 + 2] = v_mean.z;
        v_scales[g * 3] = v_scale.x * scale.x; v_scales[g * 3 + 1] = v_scale.y * scale.y;
        v_scales[g * 3 + 2] = v_scale.z * scale.z;
        v_quats[g * 4] = v_quat.x; v_quats[g * 4 + 1] = v_quat.y; v_quats[g * 4 + 2] = v_quat.z;
        v_quats[g * 4 + 3] = v_quat.w;
    }
    return 0;
}

// ---------------------------------------------------------------------------
// render_forward: crates/brush-render/src/render.rs:55-323 sequencing.
// All output arrays are caller-allocated with the reference's shapes:
//   projected f32[N,9], uniforms_out (112 B, num_visible filled in),
//   num_intersections u32[1], num_visible u32[1], final_index u32[H,W],
//   cum_tiles_hit u32[N], tile_bins u32[Ty,Tx,2],
//   compact_gid_from_isect u32[capacity], global_from_compact_gid u32[N],
//   tile_id_from_isect u32[capacity] (sorted; not part of RenderAux, kept for tests).
// Divergence from render.rs:204-206: capacity is an argument (the reference hard
// codes min(N*T, 128*65535)); entries beyond it are dropped exactly as
// map_gaussian_to_intersects.wgsl:40 does, and -2 is returned instead of silence.
// ---------------------------------------------------------------------------
int oracle_render_forward(const Uniforms* uin, uint32_t N, uint32_t K, const float* means, const float* log_scales,
                          const float* quats, const float* coeffs, const float* raw_opacities, int raster_u32,
                          void* out_img, float* projected, Uniforms* uniforms_out, uint32_t* num_intersections,
                          uint32_t* num_visible_out, uint32_t* final_index, uint32_t* cum_tiles_hit,
                          uint32_t* tile_bins, uint32_t* compact_gid_from_isect, uint32_t* global_from_compact_gid,
                          uint32_t* tile_id_from_isect, uint64_t capacity) {
    Uniforms u = *uin;
    u.total_splats = N;
    u.num_visible = 0;
    if (!(K == 1 || K == 4 || K == 9 || K == 16 || K == 25)) return -1;  // render.rs:44-53
    const uint32_t deg_of[26] = {0, 0, 0, 0, 1, 0, 0, 0, 0, 2, 0, 0, 0, 0, 0, 0, 3, 0, 0, 0, 0, 0, 0, 0, 0, 4};
    u.sh_degree = deg_of[K];
    u.tile_bounds[0] = (u.img_size[0] + TILE_WIDTH - 1) / TILE_WIDTH;  // render.rs:82-85
    u.tile_bounds[1] = (u.img_size[1] + TILE_WIDTH - 1) / TILE_WIDTH;

    // render.rs:123-156: ProjectSplats + DepthSort
    std::vector<uint32_t> presort(N, 0u), depth_bits(N), sorted_depth(N);
    std::vector<float> depths(N);
    uint32_t V = 0;
    {
        StageClock c(0);
        oracle_project_forward(&u, means, log_scales, quats, presort.data(), depths.data(), &V);
    }
    u.num_visible = V;
    {
        StageClock c(1);
        for (uint32_t i = 0; i < V; i++) depth_bits[i] = f2bits(depths[i]);
        std::fill(global_from_compact_gid, global_from_compact_gid + N, 0u);
        oracle_radix_argsort_u32(depth_bits.data(), presort.data(), V, 32, sorted_depth.data(), global_from_compact_gid);
    }

    // render.rs:161-184: ProjectVisible (num_tiles_hit zeroed)
    std::vector<uint32_t> num_tiles_hit(N, 0u);
    {
        StageClock c(2);
        oracle_project_visible(&u, K, means, log_scales, quats, coeffs, raw_opacities, global_from_compact_gid, V,
                               projected, num_tiles_hit.data());
    }
    // render.rs:186-192: PrefixSum over all N
    {
        StageClock c(3);
        oracle_prefix_sum_u32(num_tiles_hit.data(), N, cum_tiles_hit);
    }
    uint32_t I = N > 0 ? cum_tiles_hit[N - 1] : 0u;
    int status = 0;
    if ((uint64_t)I > capacity) {
        status = -2;
        I = (uint32_t)capacity;
    }
    // render.rs:211-237: MapGaussiansToIntersect + tile sort
    uint32_t num_tiles = u.tile_bounds[0] * u.tile_bounds[1];
    // only the first I entries are ever written or read (the reference allocates `max_intersects` uninitialised)
    std::vector<uint32_t> tid(I ? I : 1), cg(I ? I : 1);
    {
        StageClock c(4);
        oracle_map_gaussian_to_intersects(&u, V, projected, cum_tiles_hit, I, tid.data(), cg.data());
    }
    uint32_t bits = 32 - (num_tiles ? __builtin_clz(num_tiles) : 32);
    {
        StageClock c(5);
        oracle_radix_argsort_u32(tid.data(), cg.data(), I, bits, tile_id_from_isect, compact_gid_from_isect);
    }
    // render.rs:239-262: GetTileBinEdges
    {
        StageClock c(6);
        std::fill(tile_bins, tile_bins + (size_t)num_tiles * 2, 0u);
        oracle_get_tile_bin_edges(tile_id_from_isect, I, tile_bins);
    }
    // render.rs:267-307: Rasterize
    {
        StageClock c(7);
        oracle_rasterize(&u, compact_gid_from_isect, tile_bins, projected, raster_u32, out_img, final_index);
    }
    *uniforms_out = u;
    *num_visible_out = V;
    *num_intersections = I;
    return status;
}

// ---------------------------------------------------------------------------
// RenderBackwards::backward: render.rs:468-625.  Gradients are dense and zeroed
// here (render.rs:505-507,539-540,547,573-575): v_means[N,3], v_xy[N,2]
// (the xy_dummy gradient), v_log_scales[N,3], v_quats[N,4], v_coeffs[N,K,3],
// v_raw_opac[N].
// ---------------------------------------------------------------------------
int oracle_render_backward(const Uniforms* up, uint32_t N, uint32_t K, const float* means, const float* log_scales,
                           const float* quats, const float* raw_opacities, const float* out_img,
                           const float* v_out_img, const float* projected, const uint32_t* final_index,
                           const uint32_t* tile_bins, const uint32_t* compact_gid_from_isect,
                           const uint32_t* global_from_compact_gid, float* v_means, float* v_xy,
                           float* v_log_scales, float* v_quats, float* v_coeffs, float* v_raw_opac) {
    const Uniforms& u = *up;
    uint32_t V = u.num_visible;
    std::vector<float> v_xy_local((size_t)N * 2, 0.f), v_conics((size_t)N * 3, 0.f), v_colors((size_t)N * 4, 0.f);
    {
        StageClock c(8);
        oracle_rasterize_backwards(&u, compact_gid_from_isect, tile_bins, projected, final_index, out_img, v_out_img, N,
                                   v_xy_local.data(), v_conics.data(), v_colors.data());
    }
    {
        StageClock c(9);
        std::fill(v_coeffs, v_coeffs + (size_t)N * K * 3, 0.f);
        std::fill(v_raw_opac, v_raw_opac + N, 0.f);
        std::fill(v_xy, v_xy + (size_t)N * 2, 0.f);
        oracle_gather_grads(&u, K, global_from_compact_gid, V, raw_opacities, means, v_colors.data(),
                            v_xy_local.data(), v_coeffs, v_raw_opac, v_xy);
    }
    {
        StageClock c(10);
        std::fill(v_means, v_means + (size_t)N * 3, 0.f);
        std::fill(v_log_scales, v_log_scales + (size_t)N * 3, 0.f);
        std::fill(v_quats, v_quats + (size_t)N * 4, 0.f);
        oracle_project_backwards(&u, means, log_scales, quats, global_from_compact_gid, V, v_xy_local.data(),
                                 v_conics.data(), v_means, v_log_scales, v_quats);
    }
    return 0;
}

}  // extern "C"
