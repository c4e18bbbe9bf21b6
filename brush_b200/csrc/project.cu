// Projection stages of the forward chain.  Compiled with -fmad=false: these kernels decide the
// integer outputs (visibility, depth order, tile lists), which are bit-exact against the oracle.
//   project_forward   == crates/brush-render/src/shaders/project_forward.wgsl:15-68
//   project_visible   == crates/brush-render/src/shaders/project_visible.wgsl:163-258
//   map_intersects    == crates/brush-render/src/shaders/map_gaussian_to_intersects.wgsl:10-48
#include "internal.h"
#include "lookback.cuh"

namespace brush {

constexpr int PF_THREADS = 256;

// Stage a block's packed vec3 rows (12-byte stride) through shared memory with 16-byte loads.
// Rows [first, first+256) -> smem[3*i .. 3*i+2].
__device__ __forceinline__ void stage_vec3_rows(const float* __restrict__ src, uint32_t first, uint32_t n_total,
                                                float* smem) {
    const uint32_t tid = threadIdx.x;
    if (first + PF_THREADS <= n_total) {
        // 256 rows * 12 B = 3072 B = 192 float4, block base is 3072-byte aligned.
        if (tid < 192) {
            float4 v = __ldg(reinterpret_cast<const float4*>(src + (size_t)first * 3) + tid);
            reinterpret_cast<float4*>(smem)[tid] = v;
        }
    } else {
        uint32_t rows = n_total - first;
        for (uint32_t i = tid; i < rows * 3; i += PF_THREADS) smem[i] = __ldg(src + (size_t)first * 3 + i);
    }
}

__global__ void __launch_bounds__(PF_THREADS)
project_forward_kernel(const __grid_constant__ Cam cam, const float* __restrict__ means,
                       const float* __restrict__ log_scales, const float* __restrict__ quats,
                       uint32_t* __restrict__ presort_gid, uint32_t* __restrict__ depth_bits, uint64_t* lb_status,
                       uint32_t* ticket, int32_t* uniforms_dev, int32_t* num_visible_out,
                       uint8_t* __restrict__ visible_mask) {
    __shared__ __align__(16) float s_mean[PF_THREADS * 3];
    __shared__ __align__(16) float s_lscale[PF_THREADS * 3];
    __shared__ uint32_t s_warp_tot[PF_THREADS / 32];
    __shared__ uint32_t s_tile, s_base;
    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    const uint32_t N = cam.total_splats;
    PDL_ENTER();
    if (tid == 0) s_tile = atomicAdd(ticket, 1u);
    __syncthreads();
    const uint32_t tile = s_tile;
    const uint32_t first = tile * PF_THREADS;
    stage_vec3_rows(means, first, N, s_mean);
    stage_vec3_rows(log_scales, first, N, s_lscale);
    __syncthreads();

    const uint32_t g = first + tid;
    bool visible = false;
    float depth = 0.0f;
    if (g < N) {
        V3 mean = {s_mean[3 * tid], s_mean[3 * tid + 1], s_mean[3 * tid + 2]};
        V3 p_view = mul(cam.W, mean) + cam.t;
        if (!(p_view.z <= 0.01f)) {
            V3 scale = exp3(V3{s_lscale[3 * tid], s_lscale[3 * tid + 1], s_lscale[3 * tid + 2]});
            float4 q4 = __ldg(reinterpret_cast<const float4*>(quats) + g);
            V3 cov2d = calc_cov2d(cam, p_view, scale, V4{q4.x, q4.y, q4.z, q4.w});
            float det = cov2d.x * cov2d.z - cov2d.y * cov2d.y;
            if (!(det == 0.0f)) {
                V3 conic = cov_to_conic(cov2d);
                V2 xy = project_pix(cam, p_view);
                uint32_t radius = radius_from_conic(conic);
                TileBox b = get_tile_bbox(xy, radius, cam.tiles_x, cam.tiles_y);
                visible = !((b.maxx - b.minx) == 0u || (b.maxy - b.miny) == 0u);
                depth = p_view.z;
            }
        }
    }
    if (g < N) visible_mask[g] = visible ? 1 : 0;  // per-gid visibility for the refine statistics (train.rs:300-314)
    // Deterministic compaction: ballot -> warp offsets -> block scan -> chained look-back across tiles.
    const uint32_t ballot = __ballot_sync(0xffffffffu, visible);
    const uint32_t in_warp = __popc(ballot & ((1u << lane) - 1u));
    if (lane == 0) s_warp_tot[warp] = __popc(ballot);
    __syncthreads();
    if (warp == 0) {
        uint32_t v = lane < PF_THREADS / 32 ? s_warp_tot[lane] : 0u;
        uint32_t incl = v;
#pragma unroll
        for (int o = 1; o < PF_THREADS / 32; o <<= 1) {
            uint32_t n = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= (uint32_t)o) incl += n;
        }
        uint32_t block_total = __shfl_sync(0xffffffffu, incl, PF_THREADS / 32 - 1);
        if (lane < PF_THREADS / 32) s_warp_tot[lane] = incl - v;  // exclusive warp offsets
        uint32_t excl = lookback_exclusive(lb_status, tile, block_total);
        if (lane == 0) {
            s_base = excl;
            if (first + PF_THREADS >= N) {  // last tile: the inclusive total is num_visible
                uniforms_dev[UNI_NUM_VISIBLE_WORD] = (int32_t)(excl + block_total);
                *num_visible_out = (int32_t)(excl + block_total);
            }
        }
    }
    __syncthreads();
    if (visible) {
        uint32_t w = s_base + s_warp_tot[warp] + in_warp;
        presort_gid[w] = g;
        depth_bits[w] = __float_as_uint(depth);  // positive floats sort as u32 (render.rs:153-155)
    }
}

void launch_project_forward(const Cam& cam, uint32_t N, const float* means, const float* log_scales,
                            const float* quats, uint32_t* presort_gid, uint32_t* depth_bits, uint64_t* lb_status,
                            uint32_t* ticket, int32_t* uniforms_dev, int32_t* num_visible_out, uint8_t* visible_mask,
                            cudaStream_t st) {
    if (N == 0) return;
    uint32_t tiles = (N + PF_THREADS - 1) / PF_THREADS;
    g_launch_count++;
    launch_k(project_forward_kernel, tiles, PF_THREADS, 0, st, cam, means, log_scales, quats, presort_gid, depth_bits,
             lb_status, ticket, uniforms_dev, num_visible_out, visible_mask);
}

// ---------------------------------------------------------------------------------------------
template <int K>
__device__ __forceinline__ V3 eval_sh_color(uint32_t degree, V3 viewdir, const float* __restrict__ coeffs,
                                            uint32_t g) {
    float c[3 * K];
    const float* base = coeffs + (size_t)g * (3 * K);
    if ((3 * K) % 4 == 0) {
#pragma unroll
        for (int j = 0; j < (3 * K) / 4; j++) {
            float4 v = __ldg(reinterpret_cast<const float4*>(base) + j);
            c[4 * j] = v.x; c[4 * j + 1] = v.y; c[4 * j + 2] = v.z; c[4 * j + 3] = v.w;
        }
    } else {
#pragma unroll
        for (int j = 0; j < 3 * K; j++) c[j] = __ldg(base + j);
    }
    float b[K];
    sh_basis<K>(degree, viewdir, b);
    V3 col = {0.f, 0.f, 0.f};
#pragma unroll
    for (int k = 0; k < K; k++) {
        col.x += b[k] * c[3 * k];
        col.y += b[k] * c[3 * k + 1];
        col.z += b[k] * c[3 * k + 2];
    }
    return col;
}

// Tile hits of 32 splats, found by the whole warp.  A thread-per-splat walk over its own bounding box leaves most
// lanes idle (boxes differ from 1 to hundreds of tiles; measured 13 of 32 lanes active), so the (splat, tile)
// candidates of the warp's 32 splats are flattened: lane i of each pass tests candidate base + i, whichever splat
// it belongs to (found by binary search over the warp's running box sizes), with the reference's exact test
// (can_be_visible, helpers.wgsl:264-279).  Each owner lane collects the ballot bits of its own candidates: their
// count is num_tiles_hit, and for boxes of up to 64 tiles the bits themselves are kept as a row-major hit mask,
// which map_intersects expands without repeating the test.
constexpr uint32_t BOX_BIG = 0x80000000u;  // tile_boxes flag: box > 64 tiles (or coordinates > 1023): no mask
struct TileHits { uint32_t count; uint32_t box; uint64_t mask; };

BD TileHits warp_tile_hits(const VisTest& vt, const TileBox& b, bool valid, uint32_t lane) {
    const uint32_t FULL = 0xffffffffu;
    const uint32_t w = valid && vt.any ? b.maxx - b.minx : 0u, h = valid && vt.any ? b.maxy - b.miny : 0u;
    const uint32_t n = w * h;
    uint32_t incl = n;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(FULL, incl, o);
        if (lane >= (uint32_t)o) incl += t;
    }
    const uint32_t excl = incl - n, total = __shfl_sync(FULL, incl, 31);
    const uint32_t origin = b.minx | (b.miny << 16);
    TileHits out{0u, 0u, 0ull};
    for (uint32_t base = 0; base < total; base += 32) {
        const uint32_t k = base + lane;
        uint32_t owner = 0;  // number of lanes whose boxes end at or before candidate k
#pragma unroll
        for (int step = 16; step >= 1; step >>= 1) {
            const uint32_t v = __shfl_sync(FULL, incl, (owner + step - 1) & 31u);
            if (v <= k) owner += step;
        }
        owner &= 31u;
        const uint32_t o_excl = __shfl_sync(FULL, excl, owner), o_w = __shfl_sync(FULL, w, owner);
        const uint32_t o_origin = __shfl_sync(FULL, origin, owner);
        VisTest ot;
        ot.any = true;
        ot.conic_scaled.c[0].x = __shfl_sync(FULL, vt.conic_scaled.c[0].x, owner);
        ot.conic_scaled.c[0].y = __shfl_sync(FULL, vt.conic_scaled.c[0].y, owner);
        ot.conic_scaled.c[1].y = __shfl_sync(FULL, vt.conic_scaled.c[1].y, owner);
        ot.conic_scaled.c[1].x = ot.conic_scaled.c[0].y;
        ot.xy.x = __shfl_sync(FULL, vt.xy.x, owner);
        ot.xy.y = __shfl_sync(FULL, vt.xy.y, owner);
        bool pass = false;
        if (k < total) {
            const uint32_t j = k - o_excl;
            // j / o_w for small integers: (j + 0.5) / w is never within rounding distance of an integer
            const uint32_t row = (uint32_t)__float2uint_rz(__fdividef((float)j + 0.5f, (float)o_w));
            const uint32_t col = j - row * o_w;
            pass = can_be_visible_fast(ot, (o_origin & 0xffffu) + col, (o_origin >> 16) + row);
        }
        const uint32_t ballot = __ballot_sync(FULL, pass);
        if (n) {  // this lane's own candidates inside the pass
            const int s0 = max((int)excl - (int)base, 0), e0 = min((int)incl - (int)base, 32);
            if (e0 > s0) {
                const uint32_t len = (uint32_t)(e0 - s0);
                const uint32_t seg = (ballot >> s0) & (len == 32u ? 0xffffffffu : ((1u << len) - 1u));
                out.count += __popc(seg);
                const uint32_t pos = base + (uint32_t)s0 - excl;
                if (pos < 64u) out.mask |= (uint64_t)seg << pos;
            }
        }
    }
    const bool big = n > 64u || b.minx > 1023u || b.miny > 1023u || w > 1023u;
    out.box = big ? BOX_BIG : (b.minx | (b.miny << 10) | (w << 20));
    return out;
}

// The reference walks the bounding box of a CIRCLE of radius ceil(3 sqrt(lambda_max)) (helpers.wgsl:192-201,55-71)
// and tests every tile of it against the alpha >= 1/255 ELLIPSE (can_be_visible).  The candidates outside the
// ellipse's own axis-aligned extent can never pass, so the box handed to the tile test is first clipped to that
// extent: for {d : d^T C d <= 1}, |dx| <= sqrt(c / det C), |dy| <= sqrt(a / det C).  The margin (0.02 px + 0.1 %)
// is two orders above the rounding error of the exact test at 8K coordinates, so only tiles the exact test
// rejects are dropped: the hit lists stay bit-identical (row-major order inside a sub-box is the order inside
// the box), with a quarter fewer candidates on the 1 M scene and more on larger splats.
BD void shrink_to_ellipse(TileBox& b, const VisTest& vt) {
    if (!vt.any) return;
    const float a = vt.conic_scaled.c[0].x, h = vt.conic_scaled.c[0].y, c = vt.conic_scaled.c[1].y;
    const float det = a * c - h * h;
    if (!(det > 0.0f) || !(a > 0.0f) || !(c > 0.0f)) return;
    const float ex = sqrtf(c / det), ey = sqrtf(a / det);
    if (!(ex < 1e7f) || !(ey < 1e7f)) return;
    const float mx = ex * 1.001f + 0.02f, my = ey * 1.001f + 0.02f;
    const float inv = 1.0f / (float)TILE_WIDTH;
    const float lo_x = floorf((vt.xy.x - mx) * inv), hi_x = floorf((vt.xy.x + mx) * inv) + 1.0f;
    const float lo_y = floorf((vt.xy.y - my) * inv), hi_y = floorf((vt.xy.y + my) * inv) + 1.0f;
    const uint32_t minx = max(b.minx, (uint32_t)fmaxf(lo_x, 0.0f)), maxx = min(b.maxx, (uint32_t)fmaxf(hi_x, 0.0f));
    const uint32_t miny = max(b.miny, (uint32_t)fmaxf(lo_y, 0.0f)), maxy = min(b.maxy, (uint32_t)fmaxf(hi_y, 0.0f));
    b.minx = minx; b.miny = miny;
    b.maxx = max(maxx, minx); b.maxy = max(maxy, miny);
}

// COLOR = false: the SH evaluation is left to project_colors_kernel (same arithmetic, same bits), so that everything
// up to the rasterizer runs without reading sh_coeffs -- in the view-parallel step the SH part of the parameter
// exchange is still in flight on another stream while the next forward projects, sorts and bins (api.cu).
template <int K, bool COLOR>
__global__ void __launch_bounds__(256)
project_visible_kernel(const __grid_constant__ Cam cam, const float* __restrict__ means,
                       const float* __restrict__ log_scales, const float* __restrict__ quats,
                       const float* __restrict__ coeffs, const float* __restrict__ raw_opac,
                       const uint32_t* __restrict__ global_from_compact, const int32_t* __restrict__ num_visible_dev,
                       float* __restrict__ projected, uint32_t* __restrict__ num_tiles_hit,
                       uint64_t* __restrict__ tile_masks, uint32_t* __restrict__ tile_boxes) {
    const uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    PDL_ENTER();
    const uint32_t V = (uint32_t)*num_visible_dev;
    const bool valid = c < V;  // V <= total_splats
    if (c < cam.total_splats && !valid) num_tiles_hit[c] = 0u;  // the scan runs over all N (render.rs:164-165,187)
    if (__all_sync(0xffffffffu, !valid)) return;
    VisTest vt{};
    TileBox b{0u, 0u, 0u, 0u};
    if (valid) {
        const uint32_t g = global_from_compact[c];
        V3 mean = {__ldg(means + 3 * (size_t)g), __ldg(means + 3 * (size_t)g + 1), __ldg(means + 3 * (size_t)g + 2)};
        V3 ls = {__ldg(log_scales + 3 * (size_t)g), __ldg(log_scales + 3 * (size_t)g + 1),
                 __ldg(log_scales + 3 * (size_t)g + 2)};
        float4 q4 = __ldg(reinterpret_cast<const float4*>(quats) + g);
        V3 scale = exp3(ls);
        float opac = sigmoid_det(__ldg(raw_opac + g));
        V3 p_view = mul(cam.W, mean) + cam.t;
        V3 cov2d = calc_cov2d(cam, p_view, scale, V4{q4.x, q4.y, q4.z, q4.w});
        V3 conic = cov_to_conic(cov2d);
        V2 xy = project_pix(cam, p_view);
        uint32_t radius = radius_from_conic(conic);
        b = get_tile_bbox(xy, radius, cam.tiles_x, cam.tiles_y);
        vt = make_vis_test(xy, conic, opac);
        shrink_to_ellipse(b, vt);
        float* p = projected + (size_t)c * 9;
        p[0] = xy.x; p[1] = xy.y; p[2] = conic.x; p[3] = conic.y; p[4] = conic.z;
        p[8] = opac;
        if (COLOR) {
            V3 viewdir = normalize3(mean - cam.t);  // project_visible.wgsl:232-233
            V3 color = eval_sh_color<K>(cam.sh_degree, viewdir, coeffs, g);
            p[5] = color.x + 0.5f; p[6] = color.y + 0.5f; p[7] = color.z + 0.5f;
        }
    }
    const TileHits hits = warp_tile_hits(vt, b, valid, threadIdx.x & 31u);
    if (valid) {
        num_tiles_hit[c] = hits.count;
        tile_masks[c] = hits.mask;
        tile_boxes[c] = hits.box;
    }
}

// == the colour half of project_visible.wgsl:163-258 (sh_coeffs_to_color, :51-147) for the deferred-colour forward:
// projected[c][5..8) = SH(degree, normalize(mean - viewmat[3].xyz)) + 0.5 for every compact id c < num_visible.
template <int K>
__global__ void __launch_bounds__(256)
project_colors_kernel(const __grid_constant__ Cam cam, const float* __restrict__ means,
                      const float* __restrict__ coeffs, const uint32_t* __restrict__ global_from_compact,
                      const int32_t* __restrict__ num_visible_dev, float* __restrict__ projected) {
    const uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
    PDL_ENTER();
    if (c >= (uint32_t)*num_visible_dev) return;
    const uint32_t g = global_from_compact[c];
    const V3 mean = {__ldg(means + 3 * (size_t)g), __ldg(means + 3 * (size_t)g + 1), __ldg(means + 3 * (size_t)g + 2)};
    const V3 viewdir = normalize3(mean - cam.t);
    const V3 color = eval_sh_color<K>(cam.sh_degree, viewdir, coeffs, g);
    float* p = projected + (size_t)c * 9;
    p[5] = color.x + 0.5f; p[6] = color.y + 0.5f; p[7] = color.z + 0.5f;
}

void launch_project_visible(const Cam& cam, uint32_t N, uint32_t K, const float* means, const float* log_scales,
                            const float* quats, const float* coeffs, const float* raw_opac,
                            const uint32_t* global_from_compact, const int32_t* num_visible_dev, float* projected,
                            uint32_t* num_tiles_hit, uint64_t* tile_masks, uint32_t* tile_boxes, bool with_colors,
                            cudaStream_t st) {
    if (N == 0) return;
    dim3 grid((N + 255) / 256), block(256);
    g_launch_count++;
#define PV_ARGS cam, means, log_scales, quats, coeffs, raw_opac, global_from_compact, num_visible_dev, projected, \
                num_tiles_hit, tile_masks, tile_boxes
#define PV_CASE(k)                                                                                   \
    case k:                                                                                          \
        if (with_colors) launch_k(project_visible_kernel<k, true>, grid, block, 0, st, PV_ARGS);     \
        else launch_k(project_visible_kernel<1, false>, grid, block, 0, st, PV_ARGS);                \
        break;
    switch (K) {
        PV_CASE(1) PV_CASE(4) PV_CASE(9) PV_CASE(16) PV_CASE(25)
    }
#undef PV_CASE
#undef PV_ARGS
}

void launch_project_colors(const Cam& cam, uint32_t N, uint32_t K, const float* means, const float* coeffs,
                           const uint32_t* global_from_compact, const int32_t* num_visible_dev, float* projected,
                           cudaStream_t st) {
    if (N == 0) return;
    dim3 grid((N + 255) / 256), block(256);
    g_launch_count++;
#define PC_ARGS cam, means, coeffs, global_from_compact, num_visible_dev, projected
    switch (K) {
        case 1: launch_k(project_colors_kernel<1>, grid, block, 0, st, PC_ARGS); break;
        case 4: launch_k(project_colors_kernel<4>, grid, block, 0, st, PC_ARGS); break;
        case 9: launch_k(project_colors_kernel<9>, grid, block, 0, st, PC_ARGS); break;
        case 16: launch_k(project_colors_kernel<16>, grid, block, 0, st, PC_ARGS); break;
        case 25: launch_k(project_colors_kernel<25>, grid, block, 0, st, PC_ARGS); break;
    }
#undef PC_ARGS
}

// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
map_intersects_kernel(const __grid_constant__ Cam cam, const float* __restrict__ projected,
                      const uint32_t* __restrict__ cum_tiles_hit, const int32_t* __restrict__ num_visible_dev,
                      uint64_t capacity, uint32_t* __restrict__ tile_id_from_isect,
                      uint32_t* __restrict__ compact_gid_from_isect, const uint32_t* __restrict__ depth_sorted_bits,
                      uint64_t* __restrict__ key64_from_isect, const uint64_t* __restrict__ tile_masks,
                      const uint32_t* __restrict__ tile_boxes) {
    const uint32_t FULL = 0xffffffffu;
    const uint32_t c = blockIdx.x * blockDim.x + threadIdx.x, lane = threadIdx.x & 31u;
    PDL_ENTER();
    const uint32_t V = (uint32_t)*num_visible_dev;
    const bool valid = c < V;
    if (__all_sync(FULL, !valid)) return;
    const uint64_t first = valid && c > 0 ? cum_tiles_hit[c - 1] : 0u;
    const uint32_t depth_bits = valid && key64_from_isect ? depth_sorted_bits[c] : 0u;
    auto emit = [&](uint64_t isect, uint32_t cg, uint32_t dbits, uint32_t tx, uint32_t ty) {
        if (isect < capacity) {
            const uint32_t tile_id = tx + ty * cam.tiles_x;
            if (key64_from_isect) key64_from_isect[isect] = ((uint64_t)tile_id << 32) | (uint64_t)dbits;
            else tile_id_from_isect[isect] = tile_id;
            compact_gid_from_isect[isect] = cg;
        }
    };
    const uint32_t box = valid ? tile_boxes[c] : 0u;
    const bool big = valid && (box & BOX_BIG);
    {   // Expand project_visible's hit masks (row-major order, as the reference's loop), the WARP's 32 runs together: they
        // are contiguous in the output, so lane i writes output position base + i -- full 128-byte stores -- after
        // finding the splat that owns the position (binary search over the lanes' run starts, five shuffles) and the
        // k-th set bit of its mask.  One thread walking its own run wrote 4 bytes per 32-byte sector and took as long
        // as the warp's longest run (config 5: 0.69 ms for 613 MB).
        const uint64_t start0 = __shfl_sync(FULL, (unsigned long long)first, 0);
        const uint64_t last = valid ? cum_tiles_hit[c] : 0u;
        const uint32_t n_valid = __popc(__ballot_sync(FULL, valid));  // valid lanes are a prefix of the warp
        const uint32_t total = (uint32_t)(__shfl_sync(FULL, (unsigned long long)last, (int)n_valid - 1) - start0);
        const uint32_t rel_first = valid ? (uint32_t)(first - start0) : total;
        const uint64_t m = valid && !big ? tile_masks[c] : 0ull;
        const uint32_t m_lo = (uint32_t)m, m_hi = (uint32_t)(m >> 32);
        for (uint32_t p0 = 0; p0 < total; p0 += 32) {
            const uint32_t pos = p0 + lane;
            uint32_t owner = 0;
#pragma unroll
            for (uint32_t step = 16; step >= 1; step >>= 1) {
                const uint32_t cand = owner + step;
                const uint32_t v = __shfl_sync(FULL, rel_first, (int)cand);
                if (v <= pos) owner = cand;  // the LAST lane whose run starts at or before pos (empty runs share a start)
            }
            const uint32_t o_box = __shfl_sync(FULL, box, (int)owner);
            const uint32_t o_lo = __shfl_sync(FULL, m_lo, (int)owner), o_hi = __shfl_sync(FULL, m_hi, (int)owner);
            const uint32_t o_first = __shfl_sync(FULL, rel_first, (int)owner);
            const uint32_t o_dbits = key64_from_isect ? __shfl_sync(FULL, depth_bits, (int)owner) : 0u;
            if (pos < total && !(o_box & BOX_BIG)) {
                const uint32_t k = pos - o_first, cl = (uint32_t)__popc(o_lo);
                const bool in_hi = k >= cl;
                const uint32_t j = __fns(in_hi ? o_hi : o_lo, 0u, (int)(in_hi ? k - cl : k) + 1) + (in_hi ? 32u : 0u);
                const uint32_t minx = o_box & 1023u, miny = (o_box >> 10) & 1023u, w = o_box >> 20;
                const uint32_t row = (uint32_t)__float2uint_rz(__fdividef((float)j + 0.5f, (float)w));
                emit(start0 + pos, c - lane + owner, o_dbits, minx + j - row * w, miny + row);
            }
        }
    }
    // Boxes beyond the 64-tile mask: the whole warp walks one such splat at a time, 32 tiles per pass, with the
    // reference's test; the ballot gives every hit its position in the splat's row-major run.
    uint32_t pending = __ballot_sync(FULL, big);
    if (!pending) return;
    VisTest vt{};
    TileBox b{0u, 0u, 0u, 0u};
    if (big) {
        const float* p = projected + (size_t)c * 9;
        const V2 xy = {p[0], p[1]};
        const V3 conic = {p[2], p[3], p[4]};
        b = get_tile_bbox(xy, radius_from_conic(conic), cam.tiles_x, cam.tiles_y);
        vt = make_vis_test(xy, conic, p[8]);
    }
    while (pending) {
        const int src = __ffs((int)pending) - 1;
        pending &= pending - 1u;
        VisTest ot;
        ot.any = __shfl_sync(FULL, (int)vt.any, src) != 0;
        ot.conic_scaled.c[0].x = __shfl_sync(FULL, vt.conic_scaled.c[0].x, src);
        ot.conic_scaled.c[0].y = __shfl_sync(FULL, vt.conic_scaled.c[0].y, src);
        ot.conic_scaled.c[1].y = __shfl_sync(FULL, vt.conic_scaled.c[1].y, src);
        ot.conic_scaled.c[1].x = ot.conic_scaled.c[0].y;
        ot.xy.x = __shfl_sync(FULL, vt.xy.x, src);
        ot.xy.y = __shfl_sync(FULL, vt.xy.y, src);
        const uint32_t minx = __shfl_sync(FULL, b.minx, src), miny = __shfl_sync(FULL, b.miny, src);
        const uint32_t bw = __shfl_sync(FULL, b.maxx, src) - minx, bh = __shfl_sync(FULL, b.maxy, src) - miny;
        const uint32_t cg = __shfl_sync(FULL, c, src), dbits = __shfl_sync(FULL, depth_bits, src);
        uint64_t isect = __shfl_sync(FULL, (unsigned long long)first, src);
        const uint32_t total = bw * bh;
        const float fw = (float)bw;
        for (uint32_t base = 0; base < total; base += 32) {
            const uint32_t j = base + lane;
            bool pass = false;
            uint32_t tx = 0, ty = 0;
            if (j < total) {
                const uint32_t row = (uint32_t)__float2uint_rz(__fdividef((float)j + 0.5f, fw));
                tx = minx + j - row * bw;
                ty = miny + row;
                pass = can_be_visible(ot, tx, ty);
            }
            const uint32_t ballot = __ballot_sync(FULL, pass);
            if (pass) emit(isect + __popc(ballot & ((1u << lane) - 1u)), cg, dbits, tx, ty);
            isect += __popc(ballot);
        }
    }
}

void launch_map_intersects(const Cam& cam, uint32_t N, const float* projected, const uint32_t* cum_tiles_hit,
                           const int32_t* num_visible_dev, uint64_t capacity, uint32_t* tile_id_from_isect,
                           uint32_t* compact_gid_from_isect, const uint32_t* depth_sorted_bits,
                           uint64_t* key64_from_isect, uint32_t* /*status_word*/, const uint64_t* tile_masks,
                           const uint32_t* tile_boxes, cudaStream_t st) {
    if (N == 0) return;
    g_launch_count++;
    launch_k(map_intersects_kernel, (N + 255) / 256, 256, 0, st, cam, projected, cum_tiles_hit, num_visible_dev, capacity,
             tile_id_from_isect, compact_gid_from_isect, depth_sorted_bits, key64_from_isect, tile_masks, tile_boxes);
}

}  // namespace brush
