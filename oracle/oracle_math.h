// TEST INFRASTRUCTURE ONLY -- CPU oracle for the brush splat pipeline.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
// legs may build, load or call anything under oracle/.  The product
// (brush_b200/) never includes, links or executes this code.
//
// Scalar float32 helpers that restate WGSL built-in semantics used by the
// reference shaders (crates/brush-render/src/shaders/*.wgsl).  Everything here
// is compiled with -ffp-contract=off: every * and + below is a separately
// rounded IEEE-754 binary32 operation; the only fused operations are the
// explicit fmaf() calls inside det_expf / det_logf.
//
// Canonical evaluation order (WGSL leaves it to the driver, so the reference
// itself is not bit-defined; this is the order both the oracle and the CUDA
// kernels use so integer outputs can be compared bit-exactly):
//   dot(a,b)      = ((a.x*b.x) + (a.y*b.y)) + (a.z*b.z)
//   M * v         = ((M[0]*v.x) + (M[1]*v.y)) + (M[2]*v.z)       (columns M[c])
//   A * B         = columns A * B[c]
//   v * M (row)   = (dot(v,M[0]), dot(v,M[1]), ...)
//   exp / log     = det_expf / det_logf below (<= ~1 ulp, IEEE-only ops)
//   sqrt, /       = correctly rounded IEEE
//   f32 -> i32/u32 conversions saturate, NaN -> 0
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>

namespace orc {

static inline uint32_t f2bits(float f) { uint32_t u; std::memcpy(&u, &f, 4); return u; }
static inline float bits2f(uint32_t u) { float f; std::memcpy(&f, &u, 4); return f; }

// exp(x), float-only, Cody-Waite reduction + degree-5 polynomial (Cephes expf
// coefficients), two-step power-of-two scaling so denormal results round once.
static inline float det_expf(float x) {
    if (x != x) return x;
    if (x > 88.72283935546875f) return INFINITY;
    if (x < -103.97208404541015625f) return 0.0f;
    float t = x * 1.44269504088896341f;
    float kf = rintf(t);  // round-to-nearest-even
    float r = fmaf(kf, -0.693359375f, x);
    r = fmaf(kf, 2.12194440e-4f, r);
    float p = 1.9875691500e-4f;
    p = fmaf(p, r, 1.3981999507e-3f);
    p = fmaf(p, r, 8.3334519073e-3f);
    p = fmaf(p, r, 4.1665795894e-2f);
    p = fmaf(p, r, 1.6666665459e-1f);
    p = fmaf(p, r, 5.0000001201e-1f);
    float r2 = r * r;
    float y = fmaf(p, r2, r);
    y = y + 1.0f;
    int k = (int)kf;
    int k1 = k / 2;  // trunc toward zero
    int k2 = k - k1;
    float s1 = bits2f((uint32_t)(k1 + 127) << 23);
    float s2 = bits2f((uint32_t)(k2 + 127) << 23);
    return (y * s1) * s2;
}

// log(x) for x > 0, float-only (Cephes logf polynomial).
static inline float det_logf(float x) {
    if (x != x) return x;
    if (x < 0.0f) return NAN;
    if (x == 0.0f) return -INFINITY;
    if (x == INFINITY) return x;
    int e = 0;
    uint32_t ix = f2bits(x);
    if (ix < 0x00800000u) {  // denormal: scale by 2^23
        x = x * 8388608.0f;
        ix = f2bits(x);
        e = -23;
    }
    e += (int)(ix >> 23) - 127;
    float m = bits2f((ix & 0x007fffffu) | 0x3f800000u);  // [1,2)
    if (m > 1.41421356237f) {
        m = m * 0.5f;
        e += 1;
    }
    float f = m - 1.0f;
    float z = f * f;
    float p = 7.0376836292e-2f;
    p = fmaf(p, f, -1.1514610310e-1f);
    p = fmaf(p, f, 1.1676998740e-1f);
    p = fmaf(p, f, -1.2420140846e-1f);
    p = fmaf(p, f, 1.4249322787e-1f);
    p = fmaf(p, f, -1.6668057665e-1f);
    p = fmaf(p, f, 2.0000714765e-1f);
    p = fmaf(p, f, -2.4999993993e-1f);
    p = fmaf(p, f, 3.3333331174e-1f);
    float y = (f * z) * p;
    float fe = (float)e;
    y = fmaf(fe, -2.12194440e-4f, y);
    y = fmaf(z, -0.5f, y);
    float r = f + y;
    r = fmaf(fe, 0.693359375f, r);
    return r;
}

static inline int32_t sat_f2i(float f) {  // WGSL i32(f32): truncate, saturate
    if (f != f) return 0;
    if (f >= 2147483648.0f) return INT32_MAX;
    if (f <= -2147483648.0f) return INT32_MIN;
    return (int32_t)f;
}
static inline uint32_t sat_f2u(float f) {  // WGSL u32(f32)
    if (f != f) return 0u;
    if (f >= 4294967296.0f) return UINT32_MAX;
    if (f <= 0.0f) return 0u;
    return (uint32_t)f;
}
static inline int32_t clampi(int32_t v, int32_t lo, int32_t hi) { return v < lo ? lo : (v > hi ? hi : v); }
// WGSL min/max/clamp on floats; written so NaN handling matches fminf/fmaxf on the GPU side.
static inline float fminw(float a, float b) { return fminf(a, b); }
static inline float fmaxw(float a, float b) { return fmaxf(a, b); }
static inline float clampf(float v, float lo, float hi) { return fminw(fmaxw(v, lo), hi); }

struct vec2 { float x, y; };
struct vec3 { float x, y, z; };
struct vec4 { float x, y, z, w; };
struct mat3 { vec3 c[3]; };   // column-major, c[i] = column i (WGSL mat3x3f)
struct mat3x2 { vec2 c[3]; }; // 3 columns of 2 rows (WGSL mat3x2f)
struct mat2 { vec2 c[2]; };

static inline vec2 operator+(vec2 a, vec2 b) { return {a.x + b.x, a.y + b.y}; }
static inline vec2 operator-(vec2 a, vec2 b) { return {a.x - b.x, a.y - b.y}; }
static inline vec2 operator*(vec2 a, float s) { return {a.x * s, a.y * s}; }
static inline vec3 operator+(vec3 a, vec3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
static inline vec3 operator-(vec3 a, vec3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
static inline vec3 operator*(vec3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
static inline vec3 operator*(float s, vec3 a) { return {s * a.x, s * a.y, s * a.z}; }
static inline float dot(vec2 a, vec2 b) { return (a.x * b.x) + (a.y * b.y); }
static inline float dot(vec3 a, vec3 b) { return ((a.x * b.x) + (a.y * b.y)) + (a.z * b.z); }

static inline vec3 mul(const mat3& m, vec3 v) { return ((m.c[0] * v.x) + (m.c[1] * v.y)) + (m.c[2] * v.z); }
static inline mat3 mul(const mat3& a, const mat3& b) { return {{mul(a, b.c[0]), mul(a, b.c[1]), mul(a, b.c[2])}}; }
static inline mat3 transpose(const mat3& m) {
    return {{{m.c[0].x, m.c[1].x, m.c[2].x}, {m.c[0].y, m.c[1].y, m.c[2].y}, {m.c[0].z, m.c[1].z, m.c[2].z}}};
}
static inline mat3 scale_cols(const mat3& m, float s) { return {{m.c[0] * s, m.c[1] * s, m.c[2] * s}}; }
static inline mat3 add(const mat3& a, const mat3& b) { return {{a.c[0] + b.c[0], a.c[1] + b.c[1], a.c[2] + b.c[2]}}; }
static inline vec2 mul(const mat3x2& m, vec3 v) { return ((m.c[0] * v.x) + (m.c[1] * v.y)) + (m.c[2] * v.z); }
static inline mat3x2 mul(const mat3x2& a, const mat3& b) { return {{mul(a, b.c[0]), mul(a, b.c[1]), mul(a, b.c[2])}}; }
static inline vec2 rowmul(vec2 v, const mat2& m) { return {dot(v, m.c[0]), dot(v, m.c[1])}; }
static inline mat2 mul(const mat2& a, const mat2& b) {
    return {{(a.c[0] * b.c[0].x) + (a.c[1] * b.c[0].y), (a.c[0] * b.c[1].x) + (a.c[1] * b.c[1].y)}};
}

}  // namespace orc
