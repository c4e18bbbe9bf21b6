// Tile binning and the per-tile alpha compositor, forward and backward.
//   tile_bins          == crates/brush-render/src/shaders/get_tile_bin_edges.wgsl:15-42
//   rasterize          == crates/brush-render/src/shaders/rasterize.wgsl:20-115
//   rasterize_backward == crates/brush-render/src/shaders/rasterize_backwards.wgsl:140-304
// FP32/MUFU-bound stages: compiled with FMA contraction on and ex2.approx for exp(-sigma); parity is
// the 1e-5 absolute image tolerance / 1e-3 relative gradient tolerance, not bit equality.
#include <cstdlib>
#include <type_traits>

#include "internal.h"

namespace brush {

// ---------------------------------------------------------------------------------------------
// Four consecutive intersections per thread (one 16-byte load + the element before them): the first version walked one
// element per thread per grid-stride iteration, each a dependent pair of loads behind the *n_dev load (13 us for 12 MB).
__global__ void __launch_bounds__(256)
tile_bins_kernel(const uint32_t* __restrict__ tile_ids, const uint64_t* __restrict__ keys64,
                 const int32_t* __restrict__ n_dev, uint64_t capacity, uint32_t* __restrict__ tile_bins) {
    PDL_ENTER();
    const uint32_t n = (uint32_t)min((uint64_t)max(*n_dev, 0), capacity);
    auto tile_of = [&](uint32_t i) { return keys64 ? (uint32_t)(keys64[i] >> 32) : tile_ids[i]; };
    const bool vec = !keys64 && (reinterpret_cast<uintptr_t>(tile_ids) & 15u) == 0;
    for (uint32_t i0 = (blockIdx.x * blockDim.x + threadIdx.x) * 4u; i0 < n; i0 += gridDim.x * blockDim.x * 4u) {
        uint32_t t[4];
        if (vec && i0 + 4u <= n) {
            const uint4 q = __ldg(reinterpret_cast<const uint4*>(tile_ids + i0));
            t[0] = q.x; t[1] = q.y; t[2] = q.z; t[3] = q.w;
        } else {
#pragma unroll
            for (uint32_t j = 0; j < 4u; j++) t[j] = i0 + j < n ? tile_of(i0 + j) : 0u;
        }
        uint32_t prev = i0 > 0 ? tile_of(i0 - 1u) : 0u;
#pragma unroll
        for (uint32_t j = 0; j < 4u; j++) {
            const uint32_t i = i0 + j;
            if (i >= n) break;
            const uint32_t cur = t[j];
            if (i == n - 1) tile_bins[cur * 2 + 1] = n;
            if (i == 0) {
                tile_bins[cur * 2] = 0u;
            } else if (prev != cur) {
                tile_bins[prev * 2 + 1] = i;
                tile_bins[cur * 2] = i;
            }
            prev = cur;
        }
    }
}

void launch_tile_bins(const uint32_t* sorted_tile_ids, const uint64_t* sorted_keys64, const int32_t* n_isect_dev,
                      uint64_t capacity, uint32_t num_tiles, uint32_t* tile_bins, cudaStream_t st) {
    // tile_bins was zero-filled by the forward's single init launch (int_zeros, render.rs:241-244)
    (void)num_tiles;
    if (capacity == 0) return;
    uint64_t blocks = (capacity + 1023) / 1024;
    if (blocks > 148 * 8) blocks = 148 * 8;
    g_launch_count++;
    launch_k(tile_bins_kernel, (uint32_t)blocks, 256, 0, st, sorted_tile_ids, sorted_keys64, n_isect_dev, capacity,
             tile_bins);
}

// ---------------------------------------------------------------------------------------------
// One CTA per 16x16 tile, 8 warps, each warp owns an 8x4 pixel block.  The tile's depth-sorted splats are
// staged through shared memory in batches of 256 (rasterize.wgsl:57-71).  B200-first differences from the
// reference loop (which walks every splat of the tile for every pixel, ~35 instructions per pair):
//  * at staging time each splat is put in "log2 form": q(d) = L + A dx^2 + B dx dy + C dy^2 with
//    L = log2(opac), (A,B,C) = -log2(e) * (cxx/2, cxy, cyy/2), so alpha = 2^q, and the visibility test
//    "sigma >= 0 && alpha >= 1/255" becomes "q - L <= 0 && q >= log2(1/255)" -- no exp for rejected pairs;
//  * the staging thread also decides, once per splat, which of the tile's eight 8x4 pixel blocks (= warps) the
//    splat's alpha >= 1/255 ellipse can reach (bounding box, then an exact ellipse-vs-rectangle test) and stores
//    the 8-bit mask; every warp turns its bit into a depth-ordered survivor list by ballot and only walks that
//    (about 40% of the tile's list at 1M splats / 1080p);
//  * survivors are processed four at a time: their geometry is independent of the compositing state, so the four
//    dependency chains overlap; the order-dependent update is predicated, not branched;
//  * whole-warp and whole-tile early exit once every pixel's transmittance is spent.
// Shared-memory layout of one CTA (byte offsets from the CTA's shared window; accessed with explicit
// ld.shared so the inner loops carry one 32-bit address register and no generic->shared conversions).
constexpr uint32_t SM_SLOTS = TILE_SIZE + 1;          // slot 256 is a dummy splat that no pixel ever sees (list padding)
constexpr uint32_t SM_G0 = 0;                         // float4[257]: x, y, A, B
// The shared-memory data pipe is what the backward compositor runs into (ncu: l1tex data-pipe wavefronts 85 % of peak,
// short-scoreboard the top stall) and the second limiter of the forward (62 %); a warp-uniform LDS.128 costs two
// wavefronts, an LDS.64 one.  So each kernel lays a splat out such that its inner loop loads no word it does not use:
//   forward  (BWD = false): G1 = C, L, opacity, tag          COL = r, g, b, mask      loop: G0.xyzw, G1.xy, COL.xyzw (5 wavefronts)
//   backward (BWD = true):  G1 = C, L, opacity, r            COL = g, b, mask, tag    loop: G0.xyzw, G1.xyzw, COL.xy (5, was 6)
constexpr uint32_t SM_G1 = SM_G0 + 16 * SM_SLOTS;     // float4[257]
constexpr uint32_t SM_COL = SM_G1 + 16 * SM_SLOTS;    // float4[257]
constexpr uint32_t SM_LIST = SM_COL + 16 * SM_SLOTS;  // u16[8][260]: per warp, LOW HALVES of the shared-window addresses of the survivors' G0 entries
constexpr uint32_t LIST_PAD = 4;                      // lists are padded with the dummy slot to a multiple of 4
constexpr uint32_t LIST_SLACK = 4;                    // entries past the padding that may be read (never used) by a load-ahead
constexpr uint32_t LIST_STRIDE = 2 * (TILE_SIZE + LIST_PAD + LIST_SLACK);
constexpr uint32_t SM_BYTES = SM_LIST + LIST_STRIDE * (TILE_SIZE / 32);
// The forward has shared memory to spare and keeps full 32-bit entries (no PRMT per entry).
constexpr uint32_t LIST_STRIDE32 = 4 * (TILE_SIZE + LIST_PAD);
constexpr uint32_t SM_BYTES_FWD = SM_LIST + LIST_STRIDE32 * (TILE_SIZE / 32);

// Packed forward: per warp, a buffer of 16 PAIRS of survivors in pair-interleaved form (see composite_forward_packed).
constexpr uint32_t PB_PAIR = 80;                      // bytes per pair: 5 x 16
constexpr uint32_t PB_WARP = 16 * PB_PAIR;            // 32 survivors per chunk
constexpr uint32_t SM_PB = SM_BYTES_FWD;              // behind the forward's lists
constexpr uint32_t SM_BYTES_FWD_PACKED = SM_PB + (TILE_SIZE / 32) * PB_WARP;

constexpr uint32_t NO_CONTRIBUTOR = 0xffffffffu;      // "no splat of this batch contributed" (never a shared-window address)

constexpr float LOG2E = 1.4426950408889634f;
constexpr float LN2 = 0.6931471805599453f;
constexpr float LOG2_INV255 = -7.994353436858858f;  // log2(1/255)

__device__ __forceinline__ float ex2_approx(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float rcp_approx(float x) {
    float y;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float4 lds128(uint32_t a) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a));
    return v;
}
// Loads of data that does not change while a batch is processed (the staged splats, the survivor lists).  Not
// volatile, so the compiler may hoist them over the stores of the pair tile and overlap them with the math of
// the previous group; `token` is a value that only exists after the barrier / __syncwarp that published the
// data (it differs from batch to batch), which keeps the loads below that point and stops values of different
// batches from being merged.
__device__ __forceinline__ float4 lds128_ro(uint32_t a, uint32_t token) {
    float4 v;
    asm("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4]; // %5" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(a), "r"(token));
    return v;
}
__device__ __forceinline__ uint2 lds64u_ro(uint32_t a, uint32_t token) {
    uint2 v;
    asm("ld.shared.v2.u32 {%0, %1}, [%2]; // %3" : "=r"(v.x), "=r"(v.y) : "r"(a), "r"(token));
    return v;
}
__device__ __forceinline__ float2 lds64(uint32_t a) {
    float2 v;
    asm volatile("ld.shared.v2.f32 {%0, %1}, [%2];" : "=f"(v.x), "=f"(v.y) : "r"(a));
    return v;
}
__device__ __forceinline__ float2 lds64_ro(uint32_t a, uint32_t token) {
    float2 v;
    asm("ld.shared.v2.f32 {%0, %1}, [%2]; // %3" : "=f"(v.x), "=f"(v.y) : "r"(a), "r"(token));
    return v;
}
__device__ __forceinline__ uint32_t lds_u32(uint32_t a) {
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a));
    return v;
}
__device__ __forceinline__ uint4 lds128u(uint32_t a) {
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a));
    return v;
}
__device__ __forceinline__ void sts128(uint32_t a, float4 v) {
    asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(a), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ void sts_u32(uint32_t a, uint32_t v) {
    asm volatile("st.shared.u32 [%0], %1;" ::"r"(a), "r"(v) : "memory");
}
__device__ __forceinline__ void sts_u16(uint32_t a, uint32_t v) {
    asm volatile("st.shared.u16 [%0], %1;" ::"r"(a), "h"((unsigned short)v) : "memory");
}
__device__ __forceinline__ uint32_t lds_u16(uint32_t a) {
    unsigned short v;
    asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(a));
    return v;
}
// Four consecutive entries of a survivor list -> full shared-window addresses.  The lists hold 16-bit halves
// (half the shared memory of u32 entries: what lets a fourth backward CTA fit on an SM); the staged splats all
// lie within one 64 KB window of the CTA's shared memory (checked once per CTA), so one PRMT per entry puts
// the common upper half back.
__device__ __forceinline__ void list_load4(uint32_t addr, uint32_t sbase, uint32_t (&aa)[4], uint32_t token) {
    const uint2 l = lds64u_ro(addr, token);
    aa[0] = __byte_perm(l.x, sbase, 0x7610);
    aa[1] = __byte_perm(l.x, sbase, 0x7632);
    aa[2] = __byte_perm(l.y, sbase, 0x7610);
    aa[3] = __byte_perm(l.y, sbase, 0x7632);
}
__device__ __forceinline__ void check_list_window(uint32_t sbase) {
    if (((sbase ^ (sbase + SM_LIST)) >> 16) != 0u) __trap();
}
// The dummy slot: L = -inf, so q = -inf fails the visibility test for every pixel.
__device__ __forceinline__ void init_dummy_slot(uint32_t sbase) {
    sts128(sbase + SM_G0 + TILE_SIZE * 16, make_float4(0.f, 0.f, 0.f, 0.f));
    sts128(sbase + SM_G1 + TILE_SIZE * 16, make_float4(0.f, __int_as_float(0xff800000), 0.f, 0.f));
    sts128(sbase + SM_COL + TILE_SIZE * 16, make_float4(0.f, 0.f, 0.f, 0.f));
}

// Does the splat's alpha >= 1/255 region reach the (continuous) rectangle [rx0, rx1] x [ry0, ry1] of pixel centres?
// In log2 form the pair test is sp(d) >= log2(1/255) - L with sp a concave quadratic in d = centre - pixel, so the
// question is whether max sp over the rectangle clears that bound: yes if the centre is inside, else the maximum
// lies on one of the four edges, where sp is a 1-D parabola.  The bound carries a slack well above the rounding
// error of sp, so this can only keep a splat the per-pixel test then rejects, never drop one it would accept.
struct CullForm {
    float x, y, A, B, C, h1, h2, L;
};
__device__ __forceinline__ bool ellipse_reaches_block(const CullForm& f, float rx0, float rx1, float ry0, float ry1) {
    const float dx0 = f.x - rx1, dx1 = f.x - rx0, dy0 = f.y - ry1, dy1 = f.y - ry0;  // d ranges over [d0, d1]
    if (dx0 <= 0.0f && dx1 >= 0.0f && dy0 <= 0.0f && dy1 >= 0.0f) return true;
    const float ya = fminf(fmaxf(f.h1 * dx0, dy0), dy1), yb = fminf(fmaxf(f.h1 * dx1, dy0), dy1);
    const float xa = fminf(fmaxf(f.h2 * dy0, dx0), dx1), xb = fminf(fmaxf(f.h2 * dy1, dx0), dx1);
    const float s0 = fmaf(f.C * ya, ya, fmaf(f.B, ya, f.A * dx0) * dx0);
    const float s1 = fmaf(f.C * yb, yb, fmaf(f.B, yb, f.A * dx1) * dx1);
    const float s2 = fmaf(f.C * dy0, dy0, fmaf(f.B, dy0, f.A * xa) * xa);
    const float s3 = fmaf(f.C * dy1, dy1, fmaf(f.B, dy1, f.A * xb) * xb);
    const float smax = fmaxf(fmaxf(s0, s1), fmaxf(s2, s3));
    const float dxm = fmaxf(fabsf(dx0), fabsf(dx1)), dym = fmaxf(fabsf(dy0), fabsf(dy1));
    const float mag = fabsf(f.A) * dxm * dxm + fabsf(f.C) * dym * dym + fabsf(f.B) * dxm * dym;
    const float bound = (LOG2_INV255 - f.L) - (0.01f + 1e-5f * mag);
    return !(smax < bound);  // NaN (degenerate conic) keeps the splat
}

// Put one splat of the tile's list into shared memory in "log2 form" (see the kernel comment), together with the
// 8-bit mask of the tile's 8x4 pixel blocks (= warps) its alpha >= 1/255 region can reach: the cull is done ONCE
// per splat by the staging thread, and every warp's list pass is then a bit test.
// `tag` travels in G1.w: the backward uses it as the row of its gradient accumulator.
// (tile_px, tile_py) = pixel coordinates of the tile's first pixel.
// Returns whether the splat is "plain": opacity <= PLAIN_MAX_OPAC (so neither compositor's alpha clamp, 0.999 forward /
// 0.99 backward, can bind: alpha <= opacity) and a conic that is positive definite with a margin (so sigma >= 0 for
// every pixel, in exact arithmetic and under rounding).  A batch of plain splats is composited by the loop variant
// without the `sigma < 0` test and the clamp: two instructions per pair fewer, the same bits.
constexpr float PLAIN_MAX_OPAC = 0.989f;
template <bool BWD>
__device__ __forceinline__ bool stage_splat(uint32_t sbase, uint32_t slot, const float* __restrict__ projected,
                                            uint32_t cg, uint32_t tag, float tile_px, float tile_py) {
    const float* p = projected + (size_t)cg * 9;
    const float x = __ldg(p), y = __ldg(p + 1), cxx = __ldg(p + 2), cxy = __ldg(p + 3), cyy = __ldg(p + 4);
    const float opac = __ldg(p + 8);
    CullForm f;
    f.x = x; f.y = y;
    f.A = -0.5f * LOG2E * cxx; f.B = -LOG2E * cxy; f.C = -0.5f * LOG2E * cyy;
    f.L = __log2f(opac);
    f.h1 = -f.B * rcp_approx(2.0f * f.C); f.h2 = -f.B * rcp_approx(2.0f * f.A);
    sts128(sbase + SM_G0 + slot * 16, make_float4(x, y, f.A, f.B));
    const float col_r = __ldg(p + 5);
    sts128(sbase + SM_G1 + slot * 16, make_float4(f.C, f.L, opac, BWD ? col_r : __uint_as_float(tag)));
    // first stage: bounding box of the region.  sigma(d) <= t = ln(255 opac)  =>  |dx| <= sqrt(2 t cyy / det),
    // |dy| <= sqrt(2 t cxx / det); slightly inflated so that rounding can only keep a splat, never drop one.
    const float t = __logf(255.0f * opac) + 1e-3f;
    const float det = cxx * cyy - cxy * cxy;
    const float k = 2.0f * t * rcp_approx(det);
    float ex = sqrtf(k * cyy) * 1.001f + 1e-2f;
    float ey = sqrtf(k * cxx) * 1.001f + 1e-2f;
    if (!(det > 0.0f) || !(ex < 1e30f) || !(ey < 1e30f)) ex = ey = 1e30f;  // degenerate: never cull
    const float bx0 = x - ex, bx1 = x + ex, by0 = y - ey, by1 = y + ey;
    uint32_t mask = 0;
#pragma unroll
    for (uint32_t b = 0; b < TILE_SIZE / 32u; b++) {  // block b = warp b: 8 wide, 4 tall
        const float rx0 = tile_px + (float)((b & 1u) * 8u) + 0.5f, ry0 = tile_py + (float)((b >> 1) * 4u) + 0.5f;
        const float rx1 = rx0 + 7.0f, ry1 = ry0 + 3.0f;
        bool hit = bx0 <= rx1 && bx1 >= rx0 && by0 <= ry1 && by1 >= ry0;
        if (hit) hit = ellipse_reaches_block(f, rx0, rx1, ry0, ry1);
        mask |= (hit ? 1u : 0u) << b;
    }
    if (BWD) sts128(sbase + SM_COL + slot * 16, make_float4(__ldg(p + 6), __ldg(p + 7), __uint_as_float(mask), __uint_as_float(tag)));
    else sts128(sbase + SM_COL + slot * 16, make_float4(col_r, __ldg(p + 6), __ldg(p + 7), __uint_as_float(mask)));
    // |cxy| <= 0.99 sqrt(cxx cyy): sigma >= 0.01 * (cxx dx^2 + cyy dy^2) / 2, five orders above its rounding error
    return opac <= PLAIN_MAX_OPAC && cxx > 0.0f && cyy > 0.0f && cxy * cxy <= 0.98f * (cxx * cyy);
}

// List pass of one warp over the staged batch: one lane per splat tests its block's bit; the shared-window
// addresses of the survivors' G0 entries are appended, in depth order, to the warp's list, which is then padded
// with the dummy slot to a multiple of LIST_PAD.  Returns the list length (without padding).
// Only slots in [t_lo, t_max) take part.
template <bool LIST16>  // LIST16 <=> the backward kernel (mask in COL.z there, COL.w in the forward)
__device__ __forceinline__ uint32_t cull_to_list(uint32_t sbase, uint32_t list_addr, uint32_t lane, uint32_t warp,
                                                 uint32_t t_lo, uint32_t t_max) {
    uint32_t n_list = 0;
    for (uint32_t c0 = t_lo & ~31u; c0 < t_max; c0 += 32) {
        const uint32_t mine = c0 + lane;
        bool hit = false;
        if (mine >= t_lo && mine < t_max) hit = (lds_u32(sbase + SM_COL + mine * 16u + (LIST16 ? 8u : 12u)) >> warp) & 1u;
        const uint32_t mask = __ballot_sync(0xffffffffu, hit);
        if (hit) {
            const uint32_t pos = n_list + __popc(mask & ((1u << lane) - 1u));
            if (LIST16) sts_u16(list_addr + 2u * pos, sbase + mine * 16u);
            else sts_u32(list_addr + 4u * pos, sbase + mine * 16u);
        }
        n_list += __popc(mask);
    }
    if (lane < LIST_PAD) {
        if (LIST16) sts_u16(list_addr + 2u * (n_list + lane), sbase + TILE_SIZE * 16u);
        else sts_u32(list_addr + 4u * (n_list + lane), sbase + TILE_SIZE * 16u);
    }
    __syncwarp();
    return n_list;
}

// The forward compositing loop of one warp over its survivor list of one batch (see the kernel comment).
// GENERAL = false: every splat of the batch is "plain" (stage_splat), so sigma >= 0 and alpha <= 0.989 hold for every
// pair and the `sp <= 0` compare and the 0.999 clamp are dropped -- same values, two instructions per pair fewer.
template <bool GENERAL, bool COUNT>
__device__ __forceinline__ uint32_t composite_forward(uint32_t list_addr, uint32_t n_list, float pcx, float pcy, float& T,
                                                      float& pr, float& pg, float& pb, uint32_t& groups_walked) {
    uint32_t final_a = NO_CONTRIBUTOR;
    for (uint32_t i = 0; i < n_list; i += LIST_PAD) {
#ifndef BRUSH_COUNT_SLOW
        if (COUNT) groups_walked++;
#endif
        const uint4 a4 = lds128u(list_addr + 4u * i);
        const uint32_t aa[4] = {a4.x, a4.y, a4.z, a4.w};
        const bool live0 = T > 0.0f;
        // a finished pixel sees nothing: its visibility threshold is +inf (one select per group instead of a
        // third compare per pair)
        const float q_min = live0 ? LOG2_INV255 : __int_as_float(0x7f800000);
        bool act[4];
        float alpha[4], Tj[5];
        Tj[0] = T;
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const float4 g0 = lds128(aa[j] + SM_G0);
            const float2 g1 = lds64(aa[j] + SM_G1);  // C, L: the loop needs nothing else of G1
            const float dx = g0.x - pcx, dy = g0.y - pcy;
            const float sp = fmaf(g1.x * dy, dy, fmaf(g0.w, dy, g0.z * dx) * dx);  // -sigma * log2(e)
            const float q = sp + g1.y;                                              // log2(opac * exp(-sigma))
            act[j] = GENERAL ? (sp <= 0.0f && q >= q_min) : (q >= q_min);
            alpha[j] = act[j] ? (GENERAL ? fminf(0.999f, ex2_approx(q)) : ex2_approx(q)) : 0.0f;
            Tj[j + 1] = Tj[j] * (1.0f - alpha[j]);
        }
        if (!__any_sync(0xffffffffu, live0 && Tj[4] <= 1e-4f)) {
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const float4 c = lds128(aa[j] + SM_COL);
                const float fac = alpha[j] * Tj[j];
                pr = fmaf(c.x, fac, pr); pg = fmaf(c.y, fac, pg); pb = fmaf(c.z, fac, pb);
                final_a = act[j] ? aa[j] : final_a;
            }
            T = Tj[4];
            continue;
        }
#ifdef BRUSH_COUNT_SLOW  // experiment build: count the groups that take the stop-test path instead
        if (COUNT) groups_walked++;
#endif
#pragma unroll
        for (int j = 0; j < 4; j++) {  // some pixel finishes in this group (at most 32 times per warp and tile)
            const bool live = act[j] && T > 0.0f;
            const float4 c = lds128(aa[j] + SM_COL);
            const float next_T = T * (1.0f - alpha[j]);
            const bool stop = live && next_T <= 1e-4f;
            const bool contrib = live && !(next_T <= 1e-4f);
            const float fac = contrib ? alpha[j] * T : 0.0f;
            pr = fmaf(c.x, fac, pr); pg = fmaf(c.y, fac, pg); pb = fmaf(c.z, fac, pb);
            T = contrib ? next_T : T;
            T = stop ? -T : T;
            final_a = contrib ? aa[j] : final_a;
        }
        if (__all_sync(0xffffffffu, T <= 0.0f)) break;
    }
    return final_a;
}

// ---- packed FP32 (sm_100 FFMA2 / FMUL2 / FADD2): two IEEE-rn operations per issued instruction ----
typedef unsigned long long f2;
__device__ __forceinline__ f2 pack2(float lo, float hi) { f2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void unpack2(f2 v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ f2 fma2(f2 a, f2 b, f2 c) { f2 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ f2 mul2(f2 a, f2 b) { f2 r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ f2 add2(f2 a, f2 b) { f2 r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ f2 sub2(f2 a, f2 b) { f2 r; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ void lds128_f2(uint32_t a, f2& lo, f2& hi) {
    asm volatile("ld.shared.v2.b64 {%0, %1}, [%2];" : "=l"(lo), "=l"(hi) : "r"(a));
}

// The same compositing loop with the per-pair geometry on packed FP32: a lane still owns one pixel, but it evaluates
// TWO survivors per instruction.  Survivor lists are arbitrary subsets of the staged batch, so each warp first repacks
// its next 32 survivors (one per lane) into a pair-interleaved buffer -- per pair (x1,x2,y1,y2 | A1,A2,B1,B2 |
// C1,C2,L1,L2 | r1,r2,g1,g2 | b1,b2,addr1,addr2) -- whose 16-byte loads are then operand pairs as they stand.
// Operation order per pixel and splat is the scalar loop's (packed operations round each half like the scalar ones),
// so the image is bit-identical to composite_forward's.
template <bool GENERAL, bool COUNT>
__device__ __forceinline__ uint32_t composite_forward_packed(uint32_t list_addr, uint32_t pb, uint32_t dummy_addr, uint32_t lane,
                                                             uint32_t n_list, float pcx, float pcy, float& T, float& pr,
                                                             float& pg, float& pb_, uint32_t& groups_walked) {
    uint32_t final_a = NO_CONTRIBUTOR;
    const f2 PCX = pack2(pcx, pcx), PCY = pack2(pcy, pcy), ONE = pack2(1.0f, 1.0f);
    const uint32_t my = pb + (lane >> 1) * PB_PAIR + (lane & 1u) * 4u;
    for (uint32_t c0 = 0; c0 < n_list; c0 += 32) {
        {   // repack: lane l takes survivor c0 + l (the dummy splat past the end of the list)
            const uint32_t a = c0 + lane < n_list ? lds_u32(list_addr + 4u * (c0 + lane)) : dummy_addr;
            const float4 g0 = lds128(a + SM_G0), g1 = lds128(a + SM_G1), col = lds128(a + SM_COL);
            __syncwarp();  // the previous chunk has been consumed by every lane
            sts_u32(my + 0, __float_as_uint(g0.x)); sts_u32(my + 8, __float_as_uint(g0.y));
            sts_u32(my + 16, __float_as_uint(g0.z)); sts_u32(my + 24, __float_as_uint(g0.w));
            sts_u32(my + 32, __float_as_uint(g1.x)); sts_u32(my + 40, __float_as_uint(g1.y));
            sts_u32(my + 48, __float_as_uint(col.x)); sts_u32(my + 56, __float_as_uint(col.y));
            sts_u32(my + 64, __float_as_uint(col.z)); sts_u32(my + 72, a);
            __syncwarp();
        }
        const uint32_t n_chunk = min(32u, n_list - c0);
        bool done = false;
        for (uint32_t i = 0; i < n_chunk; i += 4) {
            if (COUNT) groups_walked++;
            const bool live0 = T > 0.0f;
            const float q_min = live0 ? LOG2_INV255 : __int_as_float(0x7f800000);
            bool act[4];
            float alpha[4], Tj[5];
            uint32_t aa[4];
            f2 cr[2], cg[2];
            float cb[4];
            Tj[0] = T;
#pragma unroll
            for (int h = 0; h < 2; h++) {  // two pairs per group of four survivors
                const uint32_t pa = pb + ((i >> 1) + h) * PB_PAIR;
                f2 X, Y, A, B, C, L, bb, ad;
                lds128_f2(pa, X, Y);
                lds128_f2(pa + 16, A, B);
                lds128_f2(pa + 32, C, L);
                lds128_f2(pa + 48, cr[h], cg[h]);
                lds128_f2(pa + 64, bb, ad);
                const f2 dx = sub2(X, PCX), dy = sub2(Y, PCY);
                const f2 sp = fma2(mul2(C, dy), dy, mul2(fma2(B, dy, mul2(A, dx)), dx));  // -sigma * log2(e)
                const f2 q = add2(sp, L);
                float sp0, sp1, q0, q1;
                unpack2(sp, sp0, sp1);
                unpack2(q, q0, q1);
                act[2 * h] = GENERAL ? (sp0 <= 0.0f && q0 >= q_min) : (q0 >= q_min);
                act[2 * h + 1] = GENERAL ? (sp1 <= 0.0f && q1 >= q_min) : (q1 >= q_min);
                alpha[2 * h] = act[2 * h] ? (GENERAL ? fminf(0.999f, ex2_approx(q0)) : ex2_approx(q0)) : 0.0f;
                alpha[2 * h + 1] = act[2 * h + 1] ? (GENERAL ? fminf(0.999f, ex2_approx(q1)) : ex2_approx(q1)) : 0.0f;
                float o0, o1;
                unpack2(sub2(ONE, pack2(alpha[2 * h], alpha[2 * h + 1])), o0, o1);
                Tj[2 * h + 1] = Tj[2 * h] * o0;
                Tj[2 * h + 2] = Tj[2 * h + 1] * o1;
                unpack2(bb, cb[2 * h], cb[2 * h + 1]);
                float a0, a1;
                unpack2(ad, a0, a1);
                aa[2 * h] = __float_as_uint(a0); aa[2 * h + 1] = __float_as_uint(a1);
            }
            float crr[4], cgg[4];
            unpack2(cr[0], crr[0], crr[1]); unpack2(cr[1], crr[2], crr[3]);
            unpack2(cg[0], cgg[0], cgg[1]); unpack2(cg[1], cgg[2], cgg[3]);
            if (!__any_sync(0xffffffffu, live0 && Tj[4] <= 1e-4f)) {
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const float fac = alpha[j] * Tj[j];
                    pr = fmaf(crr[j], fac, pr); pg = fmaf(cgg[j], fac, pg); pb_ = fmaf(cb[j], fac, pb_);
                    final_a = act[j] ? aa[j] : final_a;
                }
                T = Tj[4];
                continue;
            }
#pragma unroll
            for (int j = 0; j < 4; j++) {  // some pixel finishes in this group (at most 32 times per warp and tile)
                const bool live = act[j] && T > 0.0f;
                const float next_T = T * (1.0f - alpha[j]);
                const bool stop = live && next_T <= 1e-4f;
                const bool contrib = live && !(next_T <= 1e-4f);
                const float fac = contrib ? alpha[j] * T : 0.0f;
                pr = fmaf(crr[j], fac, pr); pg = fmaf(cgg[j], fac, pg); pb_ = fmaf(cb[j], fac, pb_);
                T = contrib ? next_T : T;
                T = stop ? -T : T;
                final_a = contrib ? aa[j] : final_a;
            }
            if (__all_sync(0xffffffffu, T <= 0.0f)) { done = true; break; }
        }
        if (done) break;
    }
    return final_a;
}

// COUNT: the diagnostic build of the kernel that also counts the (pixel, splat) pairs it evaluates (32 lanes x the
// list entries each warp walks, padding included) into *pair_counter; launched only while brush_ctx_set_pair_counting
// is on (bench.py's "pairs_evaluated", outside every timed region).
template <bool RASTER_U32, bool COUNT, bool PACKED>
__global__ void __launch_bounds__(TILE_SIZE, 5)  // 48 registers: five CTAs per SM (measured best of 4 / 5 / 6)
rasterize_kernel(const __grid_constant__ Cam cam, const uint32_t* __restrict__ compact_gid_from_isect,
                 const uint32_t* __restrict__ tile_bins, const float* __restrict__ projected,
                 void* __restrict__ out_img, uint32_t* __restrict__ final_index,
                 unsigned long long* __restrict__ pair_counter) {
    __shared__ __align__(16) unsigned char smem[PACKED ? SM_BYTES_FWD_PACKED : SM_BYTES_FWD];
    const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(smem);
    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    const uint32_t list_addr = sbase + SM_LIST + warp * LIST_STRIDE32;
    const uint32_t pb_addr = sbase + SM_PB + warp * PB_WARP;
    const uint32_t tile_id = blockIdx.x + blockIdx.y * cam.tiles_x;
    const uint32_t bx0 = blockIdx.x * TILE_WIDTH + (warp & 1u) * 8u, by0 = blockIdx.y * TILE_WIDTH + (warp >> 1) * 4u;
    const uint32_t px = bx0 + (lane & 7u), py = by0 + (lane >> 3);
    const bool inside = px < cam.img_w && py < cam.img_h;
    const float pcx = (float)px + 0.5f, pcy = (float)py + 0.5f;

    PDL_ENTER();
    const uint2 range = __ldg(reinterpret_cast<const uint2*>(tile_bins) + tile_id);
    const uint32_t num_batches = (range.y - range.x + TILE_SIZE - 1) / TILE_SIZE;
    // Transmittance; a pixel that is finished (or outside the image) carries it NEGATED, so "still compositing"
    // is the single test T > 0 and no separate flag has to be kept and combined per splat.  T stays > 1e-4
    // while a pixel is live (rasterize.wgsl stops before going below), so the sign is free.
    float T = inside ? 1.0f : -1.0f;
    float pr = 0.f, pg = 0.f, pb = 0.f;
    uint32_t final_idx = 0;
    uint32_t groups_walked = 0;
    if (tid == 0) init_dummy_slot(sbase);

    for (uint32_t b = 0; b < num_batches; b++) {
        // barrier before the batch buffer is overwritten + whole-tile early exit (the reference keeps
        // finished threads spinning through every batch)
        if (__syncthreads_and(T <= 0.0f)) break;
        const uint32_t batch_start = range.x + b * TILE_SIZE;
        const uint32_t remaining = min(TILE_SIZE, range.y - batch_start);
        bool odd = false;
        if (tid < remaining) {
            const uint32_t cg = __ldg(compact_gid_from_isect + batch_start + tid);
            odd = !stage_splat<false>(sbase, tid, projected, cg, cg, (float)(blockIdx.x * TILE_WIDTH), (float)(blockIdx.y * TILE_WIDTH));
        }
        const int any_odd = __syncthreads_or(odd);
        if (__all_sync(0xffffffffu, T <= 0.0f)) continue;
        const uint32_t n_list = cull_to_list<false>(sbase, list_addr, lane, warp, 0u, remaining);
        // Four survivors per iteration, composited OPTIMISTICALLY: a pixel stops at most once, so the group is
        // first run without the stop logic -- alpha is zeroed for pairs that fail the visibility test or belong to
        // a finished pixel, and the transmittance chain T0..T4 is formed unconditionally.  Only when some pixel
        // of the warp would cross T <= 1e-4 inside the group (T4 tells: T never grows) is the group redone with
        // the per-splat stop test (rasterize.wgsl: stop BEFORE adding the crossing splat).  The common path
        // has no per-splat compare / select on the compositing state: the ALU pipe (compares, selects, min)
        // was this kernel's limiter (ncu: 61% ALU at 78% issue).  Same arithmetic in both paths, so the result
        // does not depend on which one ran.
        uint32_t final_a;
        if (PACKED)
            final_a = any_odd ? composite_forward_packed<true, COUNT>(list_addr, pb_addr, sbase + TILE_SIZE * 16u, lane, n_list, pcx, pcy, T, pr, pg, pb, groups_walked)
                              : composite_forward_packed<false, COUNT>(list_addr, pb_addr, sbase + TILE_SIZE * 16u, lane, n_list, pcx, pcy, T, pr, pg, pb, groups_walked);
        else
            final_a = any_odd ? composite_forward<true, COUNT>(list_addr, n_list, pcx, pcy, T, pr, pg, pb, groups_walked)
                              : composite_forward<false, COUNT>(list_addr, n_list, pcx, pcy, T, pr, pg, pb, groups_walked);
        if (final_a != NO_CONTRIBUTOR) final_idx = batch_start + ((final_a - sbase) >> 4);
    }
    if (COUNT && lane == 0 && groups_walked) atomicAdd(pair_counter, (unsigned long long)groups_walked * (LIST_PAD * 32u));
    if (inside) {
        const size_t pix_id = (size_t)px + (size_t)py * cam.img_w;
        const float a = 1.0f - fabsf(T);
        if (RASTER_U32) {  // rasterize.wgsl:106-109
            uint32_t r8 = __float2uint_rz(fminf(fmaxf(pr * 255.0f, 0.f), 255.f));
            uint32_t g8 = __float2uint_rz(fminf(fmaxf(pg * 255.0f, 0.f), 255.f));
            uint32_t b8 = __float2uint_rz(fminf(fmaxf(pb * 255.0f, 0.f), 255.f));
            uint32_t a8 = __float2uint_rz(fminf(fmaxf(a * 255.0f, 0.f), 255.f));
            reinterpret_cast<uint32_t*>(out_img)[pix_id] = r8 | (g8 << 8) | (b8 << 16) | (a8 << 24);
        } else {
            reinterpret_cast<float4*>(out_img)[pix_id] = make_float4(pr, pg, pb, a);
            final_index[pix_id] = final_idx;
        }
    }
}

void launch_rasterize(const Cam& cam, const uint32_t* compact_gid_from_isect, const uint32_t* tile_bins,
                      const float* projected, int raster_u32, void* out_img, uint32_t* final_index,
                      unsigned long long* pair_counter, cudaStream_t st) {
    if (cam.tiles_x == 0 || cam.tiles_y == 0) return;
    dim3 grid(cam.tiles_x, cam.tiles_y);
    g_launch_count++;
#define RK_ARGS cam, compact_gid_from_isect, tile_bins, projected, out_img, final_index, pair_counter
    // BRUSH_RASTER_PACKED=1: the packed-FP32 compositing loop (A/B).  Same bits, 24 % fewer issued instructions per
    // group -- and the same 0.377 ms at config 4 (FFMA2 holds the FP32 pipe about as long as the two scalar
    // instructions it replaces), so the scalar loop stays the default.
    static const bool packed = [] {
        const char* e = getenv("BRUSH_RASTER_PACKED");
        return e ? atoi(e) != 0 : false;
    }();
    if (packed) {
        if (raster_u32 && pair_counter) launch_k(rasterize_kernel<true, true, true>, grid, TILE_SIZE, 0, st, RK_ARGS);
        else if (raster_u32) launch_k(rasterize_kernel<true, false, true>, grid, TILE_SIZE, 0, st, RK_ARGS);
        else if (pair_counter) launch_k(rasterize_kernel<false, true, true>, grid, TILE_SIZE, 0, st, RK_ARGS);
        else launch_k(rasterize_kernel<false, false, true>, grid, TILE_SIZE, 0, st, RK_ARGS);
    } else {
        if (raster_u32 && pair_counter) launch_k(rasterize_kernel<true, true, false>, grid, TILE_SIZE, 0, st, RK_ARGS);
        else if (raster_u32) launch_k(rasterize_kernel<true, false, false>, grid, TILE_SIZE, 0, st, RK_ARGS);
        else if (pair_counter) launch_k(rasterize_kernel<false, true, false>, grid, TILE_SIZE, 0, st, RK_ARGS);
        else launch_k(rasterize_kernel<false, false, false>, grid, TILE_SIZE, 0, st, RK_ARGS);
    }
#undef RK_ARGS
}

// ---------------------------------------------------------------------------------------------
// Backward.  Per pixel, back to front (rasterize_backwards.wgsl:194-272), same warp-level culling as the
// forward.  The reference sums each of the nine per-splat gradient terms over the subgroup for every
// (subgroup, splat) step and adds them with a CAS loop (rasterize_backwards.wgsl:79-133,274-300).  A warp
// butterfly per step makes this kernel issue-bound (ncu, profiles/r1a: ~107 of ~130 instructions per step
// were reduction and gradient math on half-empty warps), so the work is split in two phases:
//
//  phase 1 (lane = pixel): walk the warp's survivor list in depth order, 16 survivors per round in four groups of
//      four; per (pixel, splat) pair keep only the two scalars that depend on the compositing order,
//      va = vis * v_alpha and fac = alpha * T, in a per-warp shared-memory tile [16 splats][32 pixels].  EVERY
//      survivor takes a row, active pixels or not (measured: 96% have some, the cull is exact per block), so the
//      rows sit at compile-time offsets, the row -> splat map is the list itself and a group carries no vote,
//      branch or address arithmetic (0.775 -> 0.714 ms against rows handed out on demand).
//  phase 2 (lane = half a splat row): accumulates, in registers and without any cross-lane traffic, the
//      moments of va about the splat centre (1, dx, dy, dx^2, dx dy, dy^2) and fac * v_out.rgb over its 16
//      pixels; the nine gradient terms are linear in those sums (v_sigma = -opac * va).  One xor-16 exchange,
//      then two red.global.add.v4.f32 and one scalar red per row -- the native replacement for the reference's
//      CAS loops; a row of zeros is dropped.
constexpr uint32_t BW_ROW = 32u * 8u + 16u;  // one row: 32 x (va, fac) + pad => conflict-free LDS.128 at lane stride
constexpr uint32_t BW_ROWS = 16;             // rows per warp; phase 2 gives each row to two lanes (16 pixels each)
constexpr uint32_t BW_GROUP = 4;             // survivors examined per phase-1 iteration (independent geometry => ILP)
constexpr uint32_t BW_PAIR = SM_BYTES;
constexpr uint32_t BW_PAIR_WARP = BW_ROWS * BW_ROW;
constexpr uint32_t BW_PIX = BW_PAIR + (TILE_SIZE / 32u) * BW_PAIR_WARP;  // per warp float[3][32]: v_out.r / .g / .b of its pixels (512 B stride)
constexpr uint32_t BW_BYTES = BW_PIX + (TILE_SIZE / 32u) * 512u;

__device__ __forceinline__ void sts64(uint32_t a, float x, float y) {
    asm volatile("st.shared.v2.f32 [%0], {%1, %2};" ::"r"(a), "f"(x), "f"(y) : "memory");
}
__device__ __forceinline__ void red_add_v4(float* p, float a, float b, float c, float d) {
    asm volatile("red.relaxed.gpu.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d)
                 : "memory");
}

// Phase 2 for the first k (<= 16) rows of the warp's pair tile: lane l and lane l+16 share row l, the low lane
// sums pixel rows 0-1 of the 8x4 block, the high lane rows 2-3; one xor-16 exchange completes the nine sums.
// (bx, by) = pixel-centre coordinates of the warp's first pixel.
// Row r belongs to entry r of the survivor list at `entries` (16-bit halves, see list_load4).
__device__ __forceinline__ void flush_rows(uint32_t dummy_addr, uint32_t pair_addr, uint32_t pix_addr,
                                           uint32_t entries, uint32_t lane, uint32_t k, float bx, float by,
                                           float* __restrict__ v_projected) {
    __syncwarp();
    const uint32_t r_idx = lane & 15u, half = lane >> 4;
    bool valid = r_idx < k;
    const uint32_t a = !valid ? dummy_addr : __byte_perm(lds_u16(entries + 2u * r_idx), dummy_addr, 0x7610);
    const float4 g0 = lds128(a + SM_G0);  // x, y, A, B
    const float4 g1 = lds128(a + SM_G1);  // C, L, opacity, r
    const uint32_t gid = lds_u32(a + SM_COL + 12u);  // global splat id (the row of its gradient accumulator)
    const uint32_t row = pair_addr + r_idx * BW_ROW + half * 128u;
    const uint32_t pix = pix_addr + half * 64u;  // three planes of 32 floats: v_out.r, .g, .b of the warp's pixels
    const float by_h = by + (float)(2u * half);
    // Moments of va about the splat centre.  Along a pixel row dy is constant, so only the x-moments
    // (1, dx, dx^2) are accumulated per pixel; dy and dy^2 are applied once per row.
    float r = 0.f, g = 0.f, b = 0.f;
    float m0[2], m1[2], m2[2];
    const float dx0 = g0.x - bx;
#pragma unroll
    for (int py = 0; py < 2; py++) {
        float a0 = 0.f, a1 = 0.f, a2 = 0.f;
#pragma unroll
        for (int kx = 0; kx < 8; kx += 4) {
            const int p = py * 8 + kx;
            // four pixels per step: their (va, fac) pairs and ONE 16-byte load per colour plane (a float4 v_out per
            // pixel cost four of these wavefront-bound loads per four pixels and carried an unused alpha)
            const float4 e0 = lds128(row + p * 8), e1 = lds128(row + p * 8 + 16);
            const float4 vr = lds128(pix + p * 4), vg = lds128(pix + 128u + p * 4), vb = lds128(pix + 256u + p * 4);
            const float va_[4] = {e0.x, e0.z, e1.x, e1.z}, fac_[4] = {e0.y, e0.w, e1.y, e1.w};
            const float cr_[4] = {vr.x, vr.y, vr.z, vr.w}, cg_[4] = {vg.x, vg.y, vg.z, vg.w}, cb_[4] = {vb.x, vb.y, vb.z, vb.w};
#pragma unroll
            for (int u = 0; u < 4; u++) {  // same pixel order as before: the sums are the same sums
                const float dx = dx0 - (float)(kx + u);
                const float t1 = va_[u] * dx;
                a0 += va_[u]; a1 = fmaf(va_[u], dx, a1); a2 = fmaf(t1, dx, a2);
                r = fmaf(fac_[u], cr_[u], r); g = fmaf(fac_[u], cg_[u], g); b = fmaf(fac_[u], cb_[u], b);
            }
        }
        m0[py] = a0; m1[py] = a1; m2[py] = a2;
    }
    const float dy0 = g0.y - by_h, dy1 = dy0 - 1.0f;
    float s0 = m0[0] + m0[1], sx = m1[0] + m1[1], sxx = m2[0] + m2[1];
    float sy = fmaf(dy1, m0[1], dy0 * m0[0]), sxy = fmaf(dy1, m1[1], dy0 * m1[0]);
    float syy = fmaf(dy1 * dy1, m0[1], (dy0 * dy0) * m0[0]);
    const uint32_t F = 0xffffffffu;
    s0 += __shfl_xor_sync(F, s0, 16); sx += __shfl_xor_sync(F, sx, 16); sy += __shfl_xor_sync(F, sy, 16);
    sxx += __shfl_xor_sync(F, sxx, 16); sxy += __shfl_xor_sync(F, sxy, 16); syy += __shfl_xor_sync(F, syy, 16);
    r += __shfl_xor_sync(F, r, 16); g += __shfl_xor_sync(F, g, 16); b += __shfl_xor_sync(F, b, 16);
    // rows without a single active pixel (every entry is an exact zero) add nothing; NaNs still propagate
    valid = valid && !(fabsf(s0) + fabsf(sxx) + fabsf(syy) + fabsf(r) + fabsf(g) + fabsf(b) == 0.0f);
    if (valid) {
        // v_sigma = -opac*va; conic = -ln2 * (2A, B, 2C) in log2 form
        const float opac = g1.z, w = LN2 * opac;
        float* dst = v_projected + (size_t)gid * VP_STRIDE;
        if (half == 0) {
            red_add_v4(dst, w * fmaf(2.0f * g0.z, sx, g0.w * sy), w * fmaf(g0.w, sx, 2.0f * g1.x * sy),
                       -0.5f * opac * sxx, -opac * sxy);
        } else {
            red_add_v4(dst + 4, -0.5f * opac * syy, r, g, b);
            atomicAdd(dst + 8, s0);
        }
    }
    __syncwarp();
}

__global__ void __launch_bounds__(TILE_SIZE, 3)  // 80 registers; a fourth CTA at 64 registers measured 4% slower
rasterize_backward_kernel(const __grid_constant__ Cam cam, const uint32_t* __restrict__ compact_gid_from_isect,
                          const uint32_t* __restrict__ tile_bins, const float* __restrict__ projected,
                          const uint32_t* __restrict__ global_from_compact,
                          const uint32_t* __restrict__ final_index, const float* __restrict__ out_img,
                          const float* __restrict__ v_out_img, float* __restrict__ v_projected,
                          unsigned long long* __restrict__ pair_counter) {
    extern __shared__ __align__(16) unsigned char smem[];
    const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(smem);
    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    const uint32_t list_addr = sbase + SM_LIST + warp * LIST_STRIDE;
    const uint32_t pair_addr = sbase + BW_PAIR + warp * BW_PAIR_WARP;
    const uint32_t pix_addr = sbase + BW_PIX + warp * 512u;
    const uint32_t tile_id = blockIdx.x;
    const uint32_t tile_x = tile_id % cam.tiles_x, tile_y = tile_id / cam.tiles_x;
    const uint32_t bx0 = tile_x * TILE_WIDTH + (warp & 1u) * 8u, by0 = tile_y * TILE_WIDTH + (warp >> 1) * 4u;
    const uint32_t px = bx0 + (lane & 7u), py = by0 + (lane >> 3);
    const bool inside = px < cam.img_w && py < cam.img_h;
    const float pcx = (float)px + 0.5f, pcy = (float)py + 0.5f;
    const float rx0 = (float)bx0 + 0.5f, ry0 = (float)by0 + 0.5f;  // pixel centre of the warp's first pixel

    PDL_ENTER();
    const uint2 range = __ldg(reinterpret_cast<const uint2*>(tile_bins) + tile_id);
    if (range.y <= range.x) return;
    if (tid == 0) {
        check_list_window(sbase);
        init_dummy_slot(sbase);
    }

    float T_final = 1.0f;
    float4 v_out = make_float4(0.f, 0.f, 0.f, 0.f);
    uint32_t final_isect = 0;
    if (inside) {
        const size_t pix_id = (size_t)px + (size_t)py * cam.img_w;
        T_final = 1.0f - __ldg(out_img + pix_id * 4 + 3);
        final_isect = __ldg(final_index + pix_id);
        v_out = __ldg(reinterpret_cast<const float4*>(v_out_img) + pix_id);
    }
    sts_u32(pix_addr + lane * 4u, __float_as_uint(v_out.x));  // colour planes for phase 2 (flush_rows)
    sts_u32(pix_addr + 128u + lane * 4u, __float_as_uint(v_out.y));
    sts_u32(pix_addr + 256u + lane * 4u, __float_as_uint(v_out.z));
    // Running state of the pixel (rasterize_backwards.wgsl:194-272 keeps T and the colour `buffer` behind the
    // current splat): v_alpha = ra * (dot(c, v_rgb) * T_before - dot(buffer, v_rgb) + T_final * v_a), so only
    // D = T_final * v_a - dot(buffer, v_rgb) is carried, not the three buffer channels.
    float T = T_final;
    float D = T_final * v_out.w;
    // Splats behind every pixel's last contributor are skipped warp-wide, and batches (and slots of the first
    // batch) behind the last contributor of the WHOLE tile are not even staged: at 4K / 6 M splats most of a tile's
    // list lies behind the point where its pixels saturate (13.6 evaluated pairs per staged intersection before).
    const uint32_t warp_final = __reduce_max_sync(0xffffffffu, inside ? final_isect : 0u);
    const bool warp_any_inside = __any_sync(0xffffffffu, inside);
    __shared__ uint32_t s_warp_final[TILE_SIZE / 32];
    if (lane == 0) s_warp_final[warp] = warp_final;
    __syncthreads();
    uint32_t tile_final = 0;
#pragma unroll
    for (uint32_t w2 = 0; w2 < TILE_SIZE / 32; w2++) tile_final = max(tile_final, s_warp_final[w2]);
    if (tile_final < range.x) return;  // no pixel of the tile has a contributor (final_index is 0 then)
    tile_final = min(tile_final, range.y - 1u);

    const uint32_t num_batches = (range.y - range.x + TILE_SIZE - 1) / TILE_SIZE;
    for (uint32_t b = (range.y - 1u - tile_final) / TILE_SIZE; b < num_batches; b++) {
        __syncthreads();
        const uint32_t batch_end = range.y - b * TILE_SIZE;  // slot t of this batch <-> isect batch_end-1-t
        const uint32_t remaining = min(TILE_SIZE, batch_end - range.x);
        const uint32_t t_lo_tile = batch_end - 1u > tile_final ? batch_end - 1u - tile_final : 0u;
        bool odd = false;
        if (tid >= t_lo_tile && tid < remaining) {  // gradient rows are indexed by GLOBAL splat id, so the per-splat backward is coalesced
            const uint32_t cg = __ldg(compact_gid_from_isect + batch_end - 1u - tid);
            odd = !stage_splat<true>(sbase, tid, projected, cg, __ldg(global_from_compact + cg), (float)(tile_x * TILE_WIDTH),
                               (float)(tile_y * TILE_WIDTH));
        }
        const int any_odd = __syncthreads_or(odd);
        if (!warp_any_inside || batch_end - remaining > warp_final) continue;
        // isect <= final  <=>  slot >= batch_end-1-final; compared as shared-window addresses of the G0 entries
        const uint32_t t_lo_warp = batch_end - 1u > warp_final ? batch_end - 1u - warp_final : 0u;
        const uint32_t a_lo = !inside ? 0xffffffffu
                                      : sbase + (batch_end - 1u > final_isect ? (batch_end - 1u - final_isect) * 16u : 0u);
        const uint32_t n_list = cull_to_list<true>(sbase, list_addr, lane, warp, t_lo_warp, remaining);
        if (pair_counter && lane == 0 && n_list)  // diagnostic (brush_ctx_set_pair_counting), per batch, not per pair
            atomicAdd(pair_counter, (unsigned long long)((n_list + BW_GROUP - 1u) / BW_GROUP) * (BW_GROUP * 32u));
        const uint32_t token = n_list ^ b;  // exists only after cull_to_list's __syncwarp (see lds128_ro)
        // GENERAL = false: every splat of the batch is "plain" (stage_splat): sigma >= 0 and alpha <= 0.989 for every
        // pair, so the `sp <= 0` compare and the clamp are dropped -- same values, two instructions per pair fewer.
        auto walk_list = [&](auto general_tag) {
            constexpr bool GENERAL = decltype(general_tag)::value;
            uint32_t an[BW_GROUP];  // the next group's list entries, loaded one group ahead (the list is padded)
            list_load4(list_addr, sbase, an, token);
            for (uint32_t g = 0; g < n_list; g += BW_ROWS) {
    #pragma unroll
                for (uint32_t u = 0; u < BW_ROWS / BW_GROUP; u++) {
                    if (g + BW_GROUP * u >= n_list) break;
                    const uint32_t aa[BW_GROUP] = {an[0], an[1], an[2], an[3]};
                    list_load4(list_addr + 2u * (g + BW_GROUP * (u + 1u)), sbase, an, token);
                    float alpha[BW_GROUP], ra[BW_GROUP], visra[BW_GROUP], col_r[BW_GROUP];
    #pragma unroll
                    for (int j = 0; j < (int)BW_GROUP; j++) {
                        const float4 g0 = lds128_ro(aa[j] + SM_G0, token);
                        const float4 g1 = lds128_ro(aa[j] + SM_G1, token);
                        const float dx = g0.x - pcx, dy = g0.y - pcy;
                        const float sp = fmaf(g1.x * dy, dy, fmaf(g0.w, dy, g0.z * dx) * dx);
                        const float q = sp + g1.y;
                        // the same predicate, on the same instruction sequence, as the forward
                        const bool act = GENERAL ? (aa[j] >= a_lo && sp <= 0.0f && q >= LOG2_INV255)
                                                 : (aa[j] >= a_lo && q >= LOG2_INV255);
                        const float vis = act ? ex2_approx(sp) : 0.0f;  // inactive: alpha = 0, ra = 1, a row of zeros
                        alpha[j] = GENERAL ? fminf(0.99f, g1.z * vis) : g1.z * vis;  // clamp 0.99 here, 0.999 forward (reference quirk)
                        ra[j] = rcp_approx(1.0f - alpha[j]);
                        visra[j] = vis * ra[j];
                        col_r[j] = g1.w;
                    }
    #pragma unroll
                    for (int j = 0; j < (int)BW_GROUP; j++) {
                        const float2 gb = lds64_ro(aa[j] + SM_COL, token);  // g, b (r came with G1)
                        const float uu = fmaf(gb.y, v_out.z, fmaf(gb.x, v_out.y, col_r[j] * v_out.x));
                        const float va = visra[j] * fmaf(uu, T, D);
                        T *= ra[j];
                        const float fac = alpha[j] * T;
                        D = fmaf(-fac, uu, D);
                        sts64(pair_addr + (BW_GROUP * u + (uint32_t)j) * BW_ROW + lane * 8u, va, fac);
                    }
                }
                flush_rows(sbase + TILE_SIZE * 16u, pair_addr, pix_addr, list_addr + 2u * g, lane, min(BW_ROWS, n_list - g),
                           rx0, ry0, v_projected);
            }
        };
        if (any_odd) walk_list(std::true_type{});
        else walk_list(std::false_type{});
    }
}

void launch_rasterize_backward(const Cam& cam, const uint32_t* compact_gid_from_isect, const uint32_t* tile_bins,
                               const float* projected, const uint32_t* global_from_compact,
                               const uint32_t* final_index, const float* out_img,
                               const float* v_out_img, float* v_projected, unsigned long long* pair_counter,
                               cudaStream_t st) {
    const uint32_t tiles = cam.tiles_x * cam.tiles_y;
    if (tiles == 0) return;
    static std::atomic<uint64_t> configured_devices{0};  // the opt-in to > 48 KB of shared memory is per device
    int dev = 0;
    cudaGetDevice(&dev);
    if (!(configured_devices.load(std::memory_order_acquire) >> dev & 1ull)) {  // idempotent: a race only repeats the call
        cudaFuncSetAttribute(rasterize_backward_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)BW_BYTES);
        configured_devices.fetch_or(1ull << dev, std::memory_order_release);
    }
    g_launch_count++;
    launch_k(rasterize_backward_kernel, tiles, TILE_SIZE, BW_BYTES, st, cam, compact_gid_from_isect, tile_bins, projected,
             global_from_compact, final_index, out_img, v_out_img, v_projected, pair_counter);
}

}  // namespace brush
