// Device math for the splat path.  The projection / culling functions produce the INTEGER outputs of
// the pipeline (visibility, radius, tile bbox, tile hit tests, depth bits), which must be bit-exact
// against the CPU oracle.  Translation units that include this header for those functions are
// compiled with -fmad=false (see build.py), so every * and + is a separately rounded IEEE binary32
// operation; the canonical evaluation order is:
//   dot(a,b) = ((a.x*b.x)+(a.y*b.y))+(a.z*b.z);  M*v = ((M[0]*v.x)+(M[1]*v.y))+(M[2]*v.z)
//   exp/log  = det_expf/det_logf (explicit fmaf only), sqrt and / correctly rounded,
//   f32->int conversions saturating with NaN -> 0 (cvt.rzi semantics).
// Reference semantics: crates/brush-render/src/shaders/helpers.wgsl (cited per function).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/brush_b200.h"

namespace brush {

constexpr uint32_t TILE_WIDTH = BRUSH_TILE_WIDTH;
constexpr uint32_t TILE_SIZE = TILE_WIDTH * TILE_WIDTH;
constexpr float COV_BLUR = 0.3f;  // helpers.wgsl:166

struct V2 { float x, y; };
struct V3 { float x, y, z; };
struct V4 { float x, y, z, w; };
struct M3 { V3 c[3]; };    // columns
struct M32 { V2 c[3]; };   // 3 columns x 2 rows
struct M2 { V2 c[2]; };

#define BD __device__ __forceinline__

BD V2 operator+(V2 a, V2 b) { return {a.x + b.x, a.y + b.y}; }
BD V2 operator-(V2 a, V2 b) { return {a.x - b.x, a.y - b.y}; }
BD V2 operator*(V2 a, float s) { return {a.x * s, a.y * s}; }
BD V3 operator+(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
BD V3 operator-(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
BD V3 operator*(V3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
BD float dot2(V2 a, V2 b) { return (a.x * b.x) + (a.y * b.y); }
BD float dot3(V3 a, V3 b) { return ((a.x * b.x) + (a.y * b.y)) + (a.z * b.z); }
BD V3 mul(const M3& m, V3 v) { return ((m.c[0] * v.x) + (m.c[1] * v.y)) + (m.c[2] * v.z); }
BD M3 mul(const M3& a, const M3& b) { return {{mul(a, b.c[0]), mul(a, b.c[1]), mul(a, b.c[2])}}; }
BD M3 transpose(const M3& m) {
    return {{{m.c[0].x, m.c[1].x, m.c[2].x}, {m.c[0].y, m.c[1].y, m.c[2].y}, {m.c[0].z, m.c[1].z, m.c[2].z}}};
}
BD M3 add(const M3& a, const M3& b) { return {{a.c[0] + b.c[0], a.c[1] + b.c[1], a.c[2] + b.c[2]}}; }
BD M3 scale_all(const M3& m, float s) { return {{m.c[0] * s, m.c[1] * s, m.c[2] * s}}; }
BD V2 mul(const M32& m, V3 v) { return ((m.c[0] * v.x) + (m.c[1] * v.y)) + (m.c[2] * v.z); }
BD M32 mul(const M32& a, const M3& b) { return {{mul(a, b.c[0]), mul(a, b.c[1]), mul(a, b.c[2])}}; }
BD V2 rowmul(V2 v, const M2& m) { return {dot2(v, m.c[0]), dot2(v, m.c[1])}; }
BD M2 mul(const M2& a, const M2& b) {
    return {{(a.c[0] * b.c[0].x) + (a.c[1] * b.c[0].y), (a.c[0] * b.c[1].x) + (a.c[1] * b.c[1].y)}};
}

// ---- deterministic transcendentals (same algorithm as the oracle's spec, IEEE-only operations) ----
BD float det_expf(float x) {
    if (x != x) return x;
    if (x > 88.72283935546875f) return __int_as_float(0x7f800000);
    if (x < -103.97208404541015625f) return 0.0f;
    float t = __fmul_rn(x, 1.44269504088896341f);
    float kf = rintf(t);
    float r = __fmaf_rn(kf, -0.693359375f, x);
    r = __fmaf_rn(kf, 2.12194440e-4f, r);
    float p = 1.9875691500e-4f;
    p = __fmaf_rn(p, r, 1.3981999507e-3f);
    p = __fmaf_rn(p, r, 8.3334519073e-3f);
    p = __fmaf_rn(p, r, 4.1665795894e-2f);
    p = __fmaf_rn(p, r, 1.6666665459e-1f);
    p = __fmaf_rn(p, r, 5.0000001201e-1f);
    float r2 = __fmul_rn(r, r);
    float y = __fmaf_rn(p, r2, r);
    y = __fadd_rn(y, 1.0f);
    int k = (int)kf;
    int k1 = k / 2;
    int k2 = k - k1;
    float s1 = __int_as_float((k1 + 127) << 23);
    float s2 = __int_as_float((k2 + 127) << 23);
    return __fmul_rn(__fmul_rn(y, s1), s2);
}

BD float det_logf(float x) {
    if (x != x) return x;
    if (x < 0.0f) return __int_as_float(0x7fc00000);
    if (x == 0.0f) return __int_as_float(0xff800000);
    if (x == __int_as_float(0x7f800000)) return x;
    int e = 0;
    uint32_t ix = __float_as_uint(x);
    if (ix < 0x00800000u) {
        x = __fmul_rn(x, 8388608.0f);
        ix = __float_as_uint(x);
        e = -23;
    }
    e += (int)(ix >> 23) - 127;
    float m = __uint_as_float((ix & 0x007fffffu) | 0x3f800000u);
    if (m > 1.41421356237f) {
        m = __fmul_rn(m, 0.5f);
        e += 1;
    }
    float f = __fadd_rn(m, -1.0f);
    float z = __fmul_rn(f, f);
    float p = 7.0376836292e-2f;
    p = __fmaf_rn(p, f, -1.1514610310e-1f);
    p = __fmaf_rn(p, f, 1.1676998740e-1f);
    p = __fmaf_rn(p, f, -1.2420140846e-1f);
    p = __fmaf_rn(p, f, 1.4249322787e-1f);
    p = __fmaf_rn(p, f, -1.6668057665e-1f);
    p = __fmaf_rn(p, f, 2.0000714765e-1f);
    p = __fmaf_rn(p, f, -2.4999993993e-1f);
    p = __fmaf_rn(p, f, 3.3333331174e-1f);
    float y = __fmul_rn(__fmul_rn(f, z), p);
    float fe = (float)e;
    y = __fmaf_rn(fe, -2.12194440e-4f, y);
    y = __fmaf_rn(z, -0.5f, y);
    float r = __fadd_rn(f, y);
    r = __fmaf_rn(fe, 0.693359375f, r);
    return r;
}

BD float div_rn(float a, float b) { return __fdiv_rn(a, b); }
BD float sqrt_rn(float a) { return __fsqrt_rn(a); }
BD int clampi(int v, int lo, int hi) { return min(max(v, lo), hi); }
BD float clampf(float v, float lo, float hi) { return fminf(fmaxf(v, lo), hi); }

// ---- per-call constants derived from RenderUniforms (helpers.wgsl:7-30) ----
struct Cam {
    M3 W;      // viewmat[0..2].xyz
    V3 t;      // viewmat[3].xyz
    float fx, fy, cx, cy;
    uint32_t img_w, img_h, tiles_x, tiles_y;
    uint32_t sh_degree, total_splats;
};

__host__ __device__ inline Cam make_cam(const BrushUniforms& u) {
    Cam c;
    const float* m = u.viewmat;
    c.W = {{{m[0], m[1], m[2]}, {m[4], m[5], m[6]}, {m[8], m[9], m[10]}}};
    c.t = {m[12], m[13], m[14]};
    c.fx = u.focal[0]; c.fy = u.focal[1];
    c.cx = u.pixel_center[0]; c.cy = u.pixel_center[1];
    c.img_w = u.img_size[0]; c.img_h = u.img_size[1];
    c.tiles_x = u.tile_bounds[0]; c.tiles_y = u.tile_bounds[1];
    c.sh_degree = u.sh_degree; c.total_splats = u.total_splats;
    return c;
}

struct TileBox { uint32_t minx, miny, maxx, maxy; };

// helpers.wgsl:55-71
BD TileBox get_tile_bbox(V2 pix_center, uint32_t pix_radius, uint32_t tiles_x, uint32_t tiles_y) {
    float tcx = div_rn(pix_center.x, (float)TILE_WIDTH), tcy = div_rn(pix_center.y, (float)TILE_WIDTH);
    float tr = div_rn((float)pix_radius, (float)TILE_WIDTH);
    TileBox b;
    b.minx = (uint32_t)clampi(__float2int_rz(tcx - tr), 0, (int)tiles_x);
    b.miny = (uint32_t)clampi(__float2int_rz(tcy - tr), 0, (int)tiles_y);
    b.maxx = (uint32_t)clampi(__float2int_rz((tcx + tr) + 1.0f), 0, (int)tiles_x);
    b.maxy = (uint32_t)clampi(__float2int_rz((tcy + tr) + 1.0f), 0, (int)tiles_y);
    return b;
}

// helpers.wgsl:74-109 (quat = w,x,y,z)
BD M3 quat_to_rotmat(V4 q) {
    float w = q.x, x = q.y, y = q.z, z = q.w;
    float x2 = x * x, y2 = y * y, z2 = z * z;
    float xy = x * y, xz = x * z, yz = y * z;
    float wx = w * x, wy = w * y, wz = w * z;
    return {{{1.0f - 2.0f * (y2 + z2), 2.0f * (xy + wz), 2.0f * (xz - wy)},
             {2.0f * (xy - wz), 1.0f - 2.0f * (x2 + z2), 2.0f * (yz + wx)},
             {2.0f * (xz + wy), 2.0f * (yz - wx), 1.0f - 2.0f * (x2 + y2)}}};
}

// helpers.wgsl:119-122
BD V2 project_pix(const Cam& c, V3 p) {
    float px = div_rn(p.x, p.z), py = div_rn(p.y, p.z);
    return {px * c.fx + c.cx, py * c.fy + c.cy};
}

// helpers.wgsl:124-158
BD V3 calc_cov2d(const Cam& c, V3 p_view, V3 scale, V4 quat) {
    float iw = (float)c.img_w, ih = (float)c.img_h;
    float tan_fov_x = div_rn(0.5f * iw, c.fx), tan_fov_y = div_rn(0.5f * ih, c.fy);
    float lims_pos_x = div_rn(iw - c.cx, c.fx) + 0.3f * tan_fov_x;
    float lims_pos_y = div_rn(ih - c.cy, c.fy) + 0.3f * tan_fov_y;
    float lims_neg_x = div_rn(c.cx, c.fx) + 0.3f * tan_fov_x;
    float lims_neg_y = div_rn(c.cy, c.fy) + 0.3f * tan_fov_y;
    float rz = div_rn(1.0f, p_view.z);
    float rz2 = rz * rz;
    float tx = p_view.z * clampf(p_view.x * rz, -lims_neg_x, lims_pos_x);
    float ty = p_view.z * clampf(p_view.y * rz, -lims_neg_y, lims_pos_y);
    M3 M = quat_to_rotmat(quat);
    M.c[0] = M.c[0] * scale.x;
    M.c[1] = M.c[1] * scale.y;
    M.c[2] = M.c[2] * scale.z;
    M32 J = {{{c.fx * rz, 0.0f}, {0.0f, c.fy * rz}, {((-c.fx) * tx) * rz2, ((-c.fy) * ty) * rz2}}};
    M3 V = mul(M, transpose(M));
    M32 T = mul(J, c.W);
    M32 TV = mul(T, V);
    V2 cov0 = mul(TV, V3{T.c[0].x, T.c[1].x, T.c[2].x});
    V2 cov1 = mul(TV, V3{T.c[0].y, T.c[1].y, T.c[2].y});
    return {cov0.x + COV_BLUR, cov0.y, cov1.y + COV_BLUR};
}

// helpers.wgsl:160-164
BD V3 cov_to_conic(V3 cov2d) {
    float det = cov2d.x * cov2d.z - cov2d.y * cov2d.y;
    float inv_det = div_rn(1.0f, det);
    return {cov2d.z * inv_det, (-cov2d.y) * inv_det, cov2d.x * inv_det};
}

// helpers.wgsl:192-201
BD uint32_t radius_from_conic(V3 conic) {
    float det = div_rn(1.0f, conic.x * conic.z - conic.y * conic.y);
    float cx = conic.z * det, cz = conic.x * det;
    float b = 0.5f * (cx + cz);
    float s = sqrt_rn(fmaxf(0.1f, b * b - det));
    float v1 = b + s;
    float v2 = b - s;
    float radius = 3.0f * sqrt_rn(fmaxf(0.0f, fmaxf(v1, v2)));
    return __float2uint_rz(ceilf(radius));
}

// helpers.wgsl:220-236
BD bool check_edge(V2 p1, V2 p2, V2 center, const M2& conic) {
    V2 edge = p2 - p1;
    V2 f = p1 - center;
    float a = dot2(rowmul(edge, conic), edge);
    float b = 2.0f * dot2(rowmul(f, conic), edge);
    float c = dot2(rowmul(f, conic), f) - 1.0f;
    float disc = b * b - (4.0f * a) * c;
    if (disc < 0.0f) return false;
    float sq = sqrt_rn(disc);
    float t1 = div_rn((-b) - sq, 2.0f * a);
    float t2 = div_rn((-b) + sq, 2.0f * a);
    return (t1 >= 0.0f && t1 <= 1.0f) || (t2 >= 0.0f && t2 <= 1.0f);
}
BD float signw(float v) { return v > 0.0f ? 1.0f : (v < 0.0f ? -1.0f : 0.0f); }

// helpers.wgsl:238-262
BD bool ellipse_intersects_aabb(V2 box_pos, V2 ext, V2 center, const M2& conic) {
    V2 d = center - box_pos;
    if (fabsf(d.x) <= ext.x && fabsf(d.y) <= ext.y) return true;
    V2 cs = {signw(d.x), signw(d.y)};
    V2 nearest = {box_pos.x + cs.x * ext.x, box_pos.y + cs.y * ext.y};
    V2 cp = nearest - center;
    if (dot2(rowmul(cp, conic), cp) <= 1.0f) return true;
    V2 e1 = {nearest.x - (cs.x * 2.0f) * ext.x, nearest.y - 0.0f};
    V2 e2 = {nearest.x - 0.0f, nearest.y - (cs.y * 2.0f) * ext.y};
    return check_edge(nearest, e1, center, conic) || check_edge(nearest, e2, center, conic);
}

// helpers.wgsl:264-279, split so that the per-splat part (log, conic scaling) is computed once.
struct VisTest {
    M2 conic_scaled;
    V2 xy;
    bool any;
};
BD VisTest make_vis_test(V2 xy, V3 conic, float opac) {
    VisTest v;
    float sigma = det_logf(opac * 255.0f);
    v.any = !(sigma <= 0.0f);
    float s2 = 2.0f * sigma;
    float a = div_rn(conic.x, s2), b = div_rn(conic.y, s2), c = div_rn(conic.z, s2);
    v.conic_scaled = {{{a, b}, {b, c}}};
    v.xy = xy;
    return v;
}
BD bool can_be_visible(const VisTest& v, uint32_t tx, uint32_t ty) {
    if (!v.any) return false;
    V2 ext = {8.0f, 8.0f};
    V2 tc = {(float)(tx * TILE_WIDTH) + ext.x, (float)(ty * TILE_WIDTH) + ext.y};
    return ellipse_intersects_aabb(tc, ext, v.xy, v.conic_scaled);
}

// The same predicate, decided cheaply wherever that is safe.  can_be_visible asks whether the ellipse
// {p : (p-c)^T C (p-c) <= 1} meets the tile (helpers.wgsl:238-262: centre inside the box, nearest corner inside the
// ellipse, or the boundary crossing one of the two edges next to that corner -- two IEEE sqrt and four IEEE divisions
// per candidate).  Away from its decision boundary that equals "min of the quadratic form over the box < 1", and the
// minimum of a convex form whose centre lies outside the box is on one of the two edges facing the centre, where it is
// a clamped 1-D parabola: ~45 instructions, no division.  Whenever that minimum is within a margin of 1 (two orders
// above the rounding error of either evaluation), or the reference's sign() logic degenerates (centre within 1e-3 px of
// a tile centre line), or the conic is not positive definite, the reference's exact sequence decides -- so the result
// is the reference's bit for bit, and only one candidate in ~1e5 pays for it.
BD bool can_be_visible_fast(const VisTest& v, uint32_t tx, uint32_t ty) {
    if (!v.any) return false;
    const float a = v.conic_scaled.c[0].x, b = v.conic_scaled.c[0].y, c = v.conic_scaled.c[1].y;
    const float bx = (float)(tx * TILE_WIDTH) + 8.0f, by = (float)(ty * TILE_WIDTH) + 8.0f;
    const float dx = v.xy.x - bx, dy = v.xy.y - by;  // the reference's d (exactly)
    const float adx = fabsf(dx), ady = fabsf(dy);
    if (adx <= 8.0f && ady <= 8.0f) return true;     // the reference's first test, same operations
    bool exact = !(adx > 1e-3f) || !(ady > 1e-3f) || !(a > 0.0f) || !(c > 0.0f) || !(__fmaf_rn(a, c, -b * b) > 0.0f);
    // the box in ellipse-centred coordinates, and its two edges that face the centre
    const float x0 = -dx - 8.0f, x1 = -dx + 8.0f, y0 = -dy - 8.0f, y1 = -dy + 8.0f;
    const float xe = dx > 0.0f ? x1 : x0, ye = dy > 0.0f ? y1 : y0;
    const float ys = fminf(fmaxf(__fdividef(-b * xe, c), y0), y1);  // minimiser on the edge px = xe
    const float xs = fminf(fmaxf(__fdividef(-b * ye, a), x0), x1);  // minimiser on the edge py = ye
    const float q1 = __fmaf_rn(__fmaf_rn(2.0f * b, xe, c * ys), ys, a * xe * xe);
    const float q2 = __fmaf_rn(__fmaf_rn(2.0f * b, ye, a * xs), xs, c * ye * ye);
    const float qmin = fminf(q1, q2);
    const float X = fmaxf(fabsf(x0), fabsf(x1)), Y = fmaxf(fabsf(y0), fabsf(y1));
    const float mag = __fmaf_rn(a * X, X, __fmaf_rn(c * Y, Y, 2.0f * fabsf(b) * X * Y));
    exact = exact || !(fabsf(qmin - 1.0f) > __fmaf_rn(8e-6f, mag, 8e-6f));  // also catches NaN / inf
    if (exact) return can_be_visible(v, tx, ty);
    return qmin < 1.0f;
}

BD float sigmoid_det(float x) { return div_rn(1.0f, 1.0f + det_expf(-x)); }  // project_visible.wgsl:149-151
BD V3 normalize3(V3 v) {
    float l = sqrt_rn(dot3(v, v));
    return {div_rn(v.x, l), div_rn(v.y, l), div_rn(v.z, l)};
}
BD V3 exp3(V3 v) { return {det_expf(v.x), det_expf(v.y), det_expf(v.z)}; }

// SH basis values b[k] (k < (deg+1)^2) at unit direction d.  project_visible.wgsl:51-147 evaluates
// colour = sum_k b[k]*coeff[k] (+0.5); gather_grads.wgsl:17-112 uses the same b[k] for the VJP.
template <int MAXK>
BD void sh_basis(uint32_t degree, V3 d, float* b) {
    b[0] = 0.2820947917738781f;
    if (degree == 0) return;
    float x = d.x, y = d.y, z = d.z;
    const float fTmp0A = 0.48860251190292f;
    b[1] = (-fTmp0A) * y;
    b[2] = fTmp0A * z;
    b[3] = (-fTmp0A) * x;
    if (degree == 1) return;
    float z2 = z * z;
    float fTmp0B = -1.092548430592079f * z;
    const float fTmp1A = 0.5462742152960395f;
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
    const float fTmp2A = -0.5900435899266435f;
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
    if (MAXK > 16) {
        float fTmp0D = z * (-4.683325804901025f * z2 + 2.007139630671868f);
        float fTmp1C = 3.31161143515146f * z2 - 0.47308734787878f;
        float fTmp2B = -1.770130769779931f * z;
        const float fTmp3A = 0.6258357354491763f;
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
}

}  // namespace brush
