// C ABI of libbrush_b200.so (include/brush_b200.h): context, scratch management and the host-side
// sequencing of the forward chain (== render_forward, crates/brush-render/src/render.rs:55-323) and the
// backward chain (== RenderBackwards::backward, render.rs:468-625).
#include <cmath>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include <nvtx3/nvToolsExt.h>  // header-only; ranges cost nothing unless a tool (nsys, ncu --nvtx) is attached

#include "internal.h"

using namespace brush;

// NVTX range names == the reference's tracing spans (crates/brush-render/src/render.rs:127-267,515-577, collected by
// crates/sync-span/src/lib.rs:26-41), spelling included ("ProjectVisibile", "Tile sort").  The fused per-splat
// backward covers the reference's "GatherGrads" and "ProjectBackwards" spans.
static const char* const kStageSpan[BRUSH_STAGE_COUNT] = {
    "ProjectSplats", "DepthSort", "ProjectVisibile", "PrefixSum", "MapGaussiansToIntersect", "Tile sort",
    "GetTileBinEdges", "Rasterize", "RasterizeBackwards", "GatherGrads+ProjectBackwards"};
struct NvtxRange {
    explicit NvtxRange(const char* name) { nvtxRangePushA(name); }
    ~NvtxRange() { nvtxRangePop(); }
};

struct Buf {
    void* p = nullptr;
    size_t bytes = 0;
};

struct brush_ctx {
    int device = 0;
    std::string err;
    int fused64 = 0;
    int timing = 0;
    void* sh_ready = nullptr;  // cudaEvent_t: the NEXT forward defers its first read of sh_coeffs behind it (one-shot)
    int split_colors = 0;      // always run the deferred-colour forward (A/B measurements)
    int pair_counting = 0;  // diagnostic: the rasterizers count the pairs they evaluate (sticky block, bytes 128..143)
    // scratch, grown on demand (growth synchronises; steady state never allocates)
    Buf control, presort_gid, depth_bits, depth_sorted, num_tiles_hit, keys_tmp, vals_tmp, hist, lb, tickets,
        tile_ids_unsorted, cg_unsorted, tile_ids_sorted, v_projected, sticky, loss_scratch, visible_mask, tile_masks, tile_boxes,
        refine_scratch;
    uint32_t refine_n = 0;  // n of the last brush_refine_plan (brush_refine_apply must match)
    const void* visible_owner = nullptr;  // aux.uniforms_buffer of the forward whose visibility visible_mask holds
    // host-staged path
    Buf h_means, h_lscales, h_quats, h_sh, h_opac, h_img, h_aux[9];
    cudaEvent_t ev[BRUSH_STAGE_COUNT + 1][2];
    bool ev_used[BRUSH_STAGE_COUNT];
    bool ev_ready = false;
    // host copies of the uniforms written by recent forwards, keyed by aux.uniforms_buffer, so the backward
    // needs no device readback (a forward's uniforms are part of the autodiff state, render.rs:436-446)
    std::vector<std::pair<const void*, BrushUniforms>> uni_cache;
};

static int fail(brush_ctx* c, int code, const std::string& msg) {
    if (c) c->err = msg;
    return code;
}
static int cuda_fail(brush_ctx* c, cudaError_t e, const char* where) {
    return fail(c, BRUSH_E_CUDA, std::string(where) + ": " + cudaGetErrorString(e));
}
#define CU(call)                                                     \
    do {                                                             \
        cudaError_t e__ = (call);                                    \
        if (e__ != cudaSuccess) return cuda_fail(ctx, e__, #call);   \
    } while (0)

static int ensure(brush_ctx* ctx, Buf& b, size_t bytes) {
    if (bytes <= b.bytes) return BRUSH_OK;
    if (b.p) CU(cudaFree(b.p));
    b.p = nullptr;
    b.bytes = 0;
    size_t want = bytes + bytes / 8 + 256;
    CU(cudaMalloc(&b.p, want));
    b.bytes = want;
    return BRUSH_OK;
}
#define ENSURE(buf, bytes)                         \
    do {                                           \
        int r__ = ensure(ctx, ctx->buf, (bytes));  \
        if (r__ != BRUSH_OK) return r__;           \
    } while (0)

struct StageTimer {
    brush_ctx* c;
    cudaStream_t st;
    int stage;
    StageTimer(brush_ctx* c_, cudaStream_t s, int stage_) : c(c_), st(s), stage(stage_) {
        nvtxRangePushA(kStageSpan[stage]);
        if (c->timing) {
            cudaEventRecord(c->ev[stage][0], st);
            c->ev_used[stage] = true;
        }
    }
    ~StageTimer() {
        if (c->timing) cudaEventRecord(c->ev[stage][1], st);
        nvtxRangePop();
    }
};

static unsigned long long* pair_counter_of(brush_ctx* c, int which) {
    return c->pair_counting ? reinterpret_cast<unsigned long long*>((char*)c->sticky.p + 128) + which : nullptr;
}

static bool valid_k(uint32_t K) { return K == 1 || K == 4 || K == 9 || K == 16 || K == 25; }
static uint32_t degree_of(uint32_t K) { return K == 1 ? 0 : K == 4 ? 1 : K == 9 ? 2 : K == 16 ? 3 : 4; }
static uint32_t bits_for(uint32_t num_tiles) {  // render.rs:227
    uint32_t b = 0;
    while (b < 32 && (num_tiles >> b)) b++;
    return b;
}

// Layout of the control buffer (zeroed per forward): [Control][lb_project u64 x ..][lb_scan u64 x ..]
struct ControlLayout {
    size_t lb_project_off, lb_scan_off, total;
};
static ControlLayout control_layout(uint32_t N) {
    ControlLayout l;
    l.lb_project_off = sizeof(Control);
    size_t project_tiles = ((size_t)N + 255) / 256 + 1;
    l.lb_scan_off = l.lb_project_off + project_tiles * 8;
    l.total = l.lb_scan_off + scan_status_words(N) * 8;
    return l;
}

// Host copy of the uniforms a forward wrote into aux.uniforms_buffer (part of the autodiff state, render.rs:436-446):
// from the ctx's cache, or read back (synchronises) when the aux comes from another ctx / process.
static int uniforms_of(brush_ctx* ctx, const BrushAux* aux, uint32_t N, uint32_t K, cudaStream_t st, BrushUniforms* out) {
    bool cached = false;
    for (auto& e : ctx->uni_cache)
        if (e.first == aux->uniforms_buffer) { *out = e.second; cached = true; }
    if (!cached) {
        CU(cudaMemcpyAsync(out, aux->uniforms_buffer, sizeof(*out), cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
    }
    if (out->total_splats != N || out->sh_degree != degree_of(K))
        return fail(ctx, BRUSH_E_INVALID, "aux.uniforms_buffer does not match N / K");
    return BRUSH_OK;
}

extern "C" {
#pragma GCC visibility push(default)

const char* brush_version(void) { return "brush_b200 0.1 (sm_100a)"; }

int brush_ctx_create(int device, brush_ctx** out) {
    if (!out) return BRUSH_E_INVALID;
    *out = nullptr;
    cudaError_t e = cudaSetDevice(device);
    if (e != cudaSuccess) return BRUSH_E_CUDA;
    brush_ctx* ctx = new brush_ctx();
    ctx->device = device;
    for (int i = 0; i < BRUSH_STAGE_COUNT; i++) {
        cudaEventCreate(&ctx->ev[i][0]);
        cudaEventCreate(&ctx->ev[i][1]);
        ctx->ev_used[i] = false;
    }
    ctx->ev_ready = true;
    if (cudaMalloc(&ctx->sticky.p, 256) != cudaSuccess || cudaMemset(ctx->sticky.p, 0, 256) != cudaSuccess) {
        delete ctx;
        return BRUSH_E_CUDA;
    }
    ctx->sticky.bytes = 256;
    *out = ctx;
    return BRUSH_OK;
}

void brush_ctx_destroy(brush_ctx* ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    Buf* bufs[] = {&ctx->control, &ctx->presort_gid, &ctx->depth_bits, &ctx->depth_sorted, &ctx->num_tiles_hit,
                   &ctx->keys_tmp, &ctx->vals_tmp, &ctx->hist, &ctx->lb, &ctx->tickets, &ctx->tile_ids_unsorted,
                   &ctx->cg_unsorted, &ctx->tile_ids_sorted, &ctx->v_projected, &ctx->sticky, &ctx->loss_scratch, &ctx->h_means, &ctx->h_lscales,
                   &ctx->h_quats, &ctx->h_sh, &ctx->h_opac, &ctx->h_img, &ctx->visible_mask, &ctx->tile_masks,
                   &ctx->tile_boxes, &ctx->refine_scratch};
    for (Buf* b : bufs)
        if (b->p) cudaFree(b->p);
    for (Buf& b : ctx->h_aux)
        if (b.p) cudaFree(b.p);
    if (ctx->ev_ready)
        for (int i = 0; i < BRUSH_STAGE_COUNT; i++) {
            cudaEventDestroy(ctx->ev[i][0]);
            cudaEventDestroy(ctx->ev[i][1]);
        }
    delete ctx;
}

const char* brush_last_error(brush_ctx* ctx) { return ctx ? ctx->err.c_str() : "null ctx"; }

int brush_ctx_set_sort_mode(brush_ctx* ctx, int fused64) {
    if (!ctx) return BRUSH_E_INVALID;
    ctx->fused64 = fused64 ? 1 : 0;
    return BRUSH_OK;
}
int brush_ctx_set_timing(brush_ctx* ctx, int enabled) {
    if (!ctx) return BRUSH_E_INVALID;
    ctx->timing = enabled ? 1 : 0;
    return BRUSH_OK;
}

int brush_ctx_set_sh_ready_event(brush_ctx* ctx, void* event) {
    if (!ctx) return BRUSH_E_INVALID;
    ctx->sh_ready = event;
    return BRUSH_OK;
}
int brush_ctx_set_split_colors(brush_ctx* ctx, int enabled) {
    if (!ctx) return BRUSH_E_INVALID;
    ctx->split_colors = enabled ? 1 : 0;
    return BRUSH_OK;
}
int brush_ctx_set_pair_counting(brush_ctx* ctx, int enabled) {
    if (!ctx) return BRUSH_E_INVALID;
    ctx->pair_counting = enabled ? 1 : 0;
    return BRUSH_OK;
}
int brush_ctx_read_pair_counts(brush_ctx* ctx, uint64_t* out_host, void* stream) {
    if (!ctx || !out_host) return BRUSH_E_INVALID;
    CU(cudaSetDevice(ctx->device));
    CU(cudaStreamSynchronize((cudaStream_t)stream));
    CU(cudaMemcpy(out_host, (char*)ctx->sticky.p + 128, 16, cudaMemcpyDeviceToHost));
    CU(cudaMemset((char*)ctx->sticky.p + 128, 0, 16));
    return BRUSH_OK;
}
int brush_measure_peaks(brush_ctx* ctx, float* out4_host, void* stream) {
    if (!ctx || !out4_host) return BRUSH_E_INVALID;
    CU(cudaSetDevice(ctx->device));
    int r = measure_peaks(out4_host, (float*)((char*)ctx->sticky.p + 192), (cudaStream_t)stream);
    if (r != 0) return cuda_fail(ctx, (cudaError_t)r, "measure_peaks");
    return BRUSH_OK;
}

int brush_stage_times_ms(brush_ctx* ctx, float* out_host, int n, void* stream) {
    if (!ctx || !out_host) return BRUSH_E_INVALID;
    CU(cudaStreamSynchronize((cudaStream_t)stream));
    for (int i = 0; i < n && i < BRUSH_STAGE_COUNT; i++) {
        out_host[i] = 0.f;
        if (ctx->ev_used[i]) cudaEventElapsedTime(&out_host[i], ctx->ev[i][0], ctx->ev[i][1]);
    }
    return BRUSH_OK;
}

int brush_ctx_check(brush_ctx* ctx, void* stream) {
    if (!ctx) return BRUSH_E_INVALID;
    CU(cudaStreamSynchronize((cudaStream_t)stream));
    uint32_t sticky = 0;
    uint32_t* dev = reinterpret_cast<uint32_t*>(ctx->sticky.p);
    CU(cudaMemcpy(&sticky, dev, 4, cudaMemcpyDeviceToHost));
    if (sticky) {
        CU(cudaMemset(dev, 0, 4));
        if (sticky & 1u)
            return fail(ctx, BRUSH_E_CAPACITY, "intersection count exceeded isect_capacity; lists were truncated");
    }
    return BRUSH_OK;
}

static int sort_scratch(brush_ctx* ctx, uint32_t max_n, int key_bytes, uint32_t bits, SortScratch* s) {
    uint32_t rb = ((bits + 3) / 4) * 4;
    uint32_t passes = (rb + 7) / 8;
    size_t tiles = sort_tiles(max_n, key_bytes);
    ENSURE(keys_tmp, (size_t)max_n * key_bytes);
    ENSURE(vals_tmp, (size_t)max_n * 4);
    ENSURE(hist, 8 * 256 * 4);
    ENSURE(lb, (size_t)(passes ? passes : 1) * tiles * 256 * 4);
    ENSURE(tickets, 64);
    s->keys_tmp = ctx->keys_tmp.p;
    s->vals_tmp = (uint32_t*)ctx->vals_tmp.p;
    s->hist = (uint32_t*)ctx->hist.p;
    s->lb = (uint32_t*)ctx->lb.p;
    s->tickets = (uint32_t*)ctx->tickets.p;
    s->pre_zeroed = false;
    return BRUSH_OK;
}

int brush_render_forward(brush_ctx* ctx, const BrushUniforms* uniforms_host, uint32_t N, uint32_t K,
                         const float* means, const float* log_scales, const float* quats, const float* sh_coeffs,
                         const float* raw_opacity, int raster_u32, void* out_img, const BrushAux* aux, void* stream) {
    if (!ctx) return BRUSH_E_INVALID;
    if (!uniforms_host || !aux || !out_img) return fail(ctx, BRUSH_E_INVALID, "null argument");
    if (!valid_k(K)) return fail(ctx, BRUSH_E_INVALID, "Invalid nr. of sh bases " + std::to_string(K));  // render.rs:51
    if (N > 0 && (!means || !log_scales || !quats || !sh_coeffs || !raw_opacity))
        return fail(ctx, BRUSH_E_INVALID, "null parameter array");
    if (!aux->uniforms_buffer || !aux->num_intersections || !aux->num_visible || !aux->tile_bins ||
        (!raster_u32 && !aux->final_index) ||
        (N > 0 && (!aux->projected_splats || !aux->cum_tiles_hit || !aux->global_from_compact_gid)) ||
        (aux->isect_capacity > 0 && !aux->compact_gid_from_isect))
        return fail(ctx, BRUSH_E_INVALID, "null aux buffer");
    if (N >= (1u << 30) || aux->isect_capacity >= (1ull << 30))
        return fail(ctx, BRUSH_E_INVALID, "N and isect_capacity must be < 2^30");
    cudaStream_t st = (cudaStream_t)stream;
    CU(cudaSetDevice(ctx->device));
    NvtxRange span("render_forward");  // render.rs:71

    BrushUniforms u = *uniforms_host;
    u.tile_bounds[0] = (u.img_size[0] + TILE_WIDTH - 1) / TILE_WIDTH;  // render.rs:82-85
    u.tile_bounds[1] = (u.img_size[1] + TILE_WIDTH - 1) / TILE_WIDTH;
    u.sh_degree = degree_of(K);
    u.total_splats = N;
    u.num_visible = 0;
    u.padding = 0;
    const Cam cam = make_cam(u);
    const uint32_t num_tiles = u.tile_bounds[0] * u.tile_bounds[1];
    const uint64_t cap = aux->isect_capacity;
    const uint32_t cap32 = (uint32_t)cap;
    const int fused = ctx->fused64;

    const ControlLayout cl = control_layout(N);
    ENSURE(control, cl.total);
    ENSURE(presort_gid, (size_t)N * 4);
    ENSURE(depth_bits, (size_t)N * 4);
    ENSURE(depth_sorted, (size_t)N * 4);
    ENSURE(num_tiles_hit, (size_t)N * 4);
    ENSURE(visible_mask, (size_t)N + 16);
    ENSURE(tile_masks, (size_t)N * 8);
    ENSURE(tile_boxes, (size_t)N * 4);
    ENSURE(tile_ids_unsorted, (size_t)cap * (fused ? 8 : 4));
    ENSURE(cg_unsorted, (size_t)cap * 4);
    if (fused || !aux->tile_id_from_isect) ENSURE(tile_ids_sorted, (size_t)cap * (fused ? 8 : 4));
    const uint32_t tile_bits = bits_for(num_tiles);
    // Sort scratch.  The status words of BOTH sorts live behind the control block in one allocation, so one
    // launch zero-fills everything a forward needs zeroed (12 fills in the first version of this path).
    SortScratch ss_depth, ss_tile;
    size_t zero_total = 0;
    {
        uint32_t mx = N > cap32 ? N : cap32;
        int r = sort_scratch(ctx, mx, fused ? 8 : 4, fused ? 32 + tile_bits : 32, &ss_tile);  // keys_tmp / vals_tmp
        if (r != BRUSH_OK) return r;
        ss_depth = ss_tile;
        const size_t tile_passes = ((((fused ? 32 + tile_bits : tile_bits) + 3) / 4) * 4 + 7) / 8;
        const size_t lb_depth = 4 * sort_tiles(N, 4) * 256 * 4;
        const size_t lb_tile = (tile_passes ? tile_passes : 1) * sort_tiles(cap32, fused ? 8 : 4) * 256 * 4;
        size_t off = (cl.total + 255) / 256 * 256;
        const size_t o_hist_d = off; off += 8 * 256 * 4;
        const size_t o_hist_t = off; off += 8 * 256 * 4;
        const size_t o_tick_d = off; off += 64;
        const size_t o_tick_t = off; off += 64;
        const size_t o_lb_d = off; off += lb_depth;
        const size_t o_lb_t = off; off += lb_tile;
        zero_total = off;
        ENSURE(control, zero_total);
        char* base = (char*)ctx->control.p;
        ss_depth.hist = (uint32_t*)(base + o_hist_d); ss_depth.tickets = (uint32_t*)(base + o_tick_d);
        ss_depth.lb = (uint32_t*)(base + o_lb_d); ss_depth.pre_zeroed = true;
        ss_tile.hist = (uint32_t*)(base + o_hist_t); ss_tile.tickets = (uint32_t*)(base + o_tick_t);
        ss_tile.lb = (uint32_t*)(base + o_lb_t); ss_tile.pre_zeroed = true;
    }
    for (int i = 0; i <= BRUSH_STAGE_RASTERIZE; i++) ctx->ev_used[i] = false;

    char* ctl = (char*)ctx->control.p;
    Control* control = (Control*)ctl;
    uint64_t* lb_project = (uint64_t*)(ctl + cl.lb_project_off);
    uint64_t* lb_scan = (uint64_t*)(ctl + cl.lb_scan_off);
    uint32_t* sticky = (uint32_t*)ctx->sticky.p;
    uint32_t* gfc = (uint32_t*)aux->global_from_compact_gid;

    {
        bool found = false;
        for (auto& e : ctx->uni_cache)
            if (e.first == aux->uniforms_buffer) { e.second = u; found = true; }
        if (!found) {
            if (ctx->uni_cache.size() >= 256) ctx->uni_cache.erase(ctx->uni_cache.begin());
            ctx->uni_cache.emplace_back(aux->uniforms_buffer, u);
        }
    }
    {
        ZeroSegs z{};
        z.p[0] = ctl; z.bytes[0] = zero_total;
        z.p[1] = aux->num_visible; z.bytes[1] = 4;
        z.p[2] = aux->num_intersections; z.bytes[2] = 4;
        z.p[3] = gfc; z.bytes[3] = (size_t)N * 4;  // int_zeros, render.rs:124
        z.p[4] = aux->tile_bins; z.bytes[4] = (size_t)num_tiles * 8;  // int_zeros, render.rs:241-244
        z.words_dst = (uint32_t*)aux->uniforms_buffer; z.n_words = 28;  // the uniform upload, render.rs:102-116
        memcpy(z.words, &u, sizeof(u));
        CU(zero_multi_async(z, st));
    }

    {   // render.rs:127-142
        StageTimer t(ctx, st, BRUSH_STAGE_PROJECT_SPLATS);
        launch_project_forward(cam, N, means, log_scales, quats, (uint32_t*)ctx->presort_gid.p,
                               (uint32_t*)ctx->depth_bits.p, lb_project, &control->ticket[0], aux->uniforms_buffer,
                               aux->num_visible, (uint8_t*)ctx->visible_mask.p, st);
        ctx->visible_owner = aux->uniforms_buffer;
    }
    {   // render.rs:151-156
        StageTimer t(ctx, st, BRUSH_STAGE_DEPTH_SORT);
        launch_radix_sort_u32((uint32_t*)ctx->depth_bits.p, (uint32_t*)ctx->presort_gid.p, aux->num_visible, N, 32,
                              (uint32_t*)ctx->depth_sorted.p, gfc, ss_depth, st);
    }
    // Deferred colours: when the caller announced that sh_coeffs is still being written on another stream
    // (brush_ctx_set_sh_ready_event), nothing before the rasterizer reads it; the SH evaluation runs as its own
    // kernel behind the event.  Same arithmetic, same bits as the fused kernel.
    void* const sh_ready = ctx->sh_ready;
    ctx->sh_ready = nullptr;
    const bool defer_colors = sh_ready != nullptr || ctx->split_colors;
    {   // render.rs:168-184
        StageTimer t(ctx, st, BRUSH_STAGE_PROJECT_VISIBLE);
        launch_project_visible(cam, N, K, means, log_scales, quats, sh_coeffs, raw_opacity, gfc, aux->num_visible,
                               aux->projected_splats, (uint32_t*)ctx->num_tiles_hit.p, (uint64_t*)ctx->tile_masks.p,
                               (uint32_t*)ctx->tile_boxes.p, !defer_colors, st);
    }
    {   // render.rs:186-192
        StageTimer t(ctx, st, BRUSH_STAGE_PREFIX_SUM);
        launch_inclusive_scan((uint32_t*)ctx->num_tiles_hit.p, N, (uint32_t*)aux->cum_tiles_hit, lb_scan,
                              &control->ticket[1], aux->num_intersections, cap, sticky, st);
    }
    uint32_t* tid_unsorted = (uint32_t*)ctx->tile_ids_unsorted.p;
    uint32_t* tid_sorted = aux->tile_id_from_isect && !fused ? (uint32_t*)aux->tile_id_from_isect
                                                             : (uint32_t*)ctx->tile_ids_sorted.p;
    {   // render.rs:211-223
        StageTimer t(ctx, st, BRUSH_STAGE_MAP_INTERSECTS);
        launch_map_intersects(cam, N, aux->projected_splats, (uint32_t*)aux->cum_tiles_hit, aux->num_visible, cap,
                              fused ? nullptr : tid_unsorted, (uint32_t*)ctx->cg_unsorted.p,
                              (uint32_t*)ctx->depth_sorted.p, fused ? (uint64_t*)tid_unsorted : nullptr, sticky,
                              (const uint64_t*)ctx->tile_masks.p, (const uint32_t*)ctx->tile_boxes.p, st);
    }
    {   // render.rs:227-237
        StageTimer t(ctx, st, BRUSH_STAGE_TILE_SORT);
        if (cap32) {
            if (fused)
                launch_radix_sort_u64((uint64_t*)tid_unsorted, (uint32_t*)ctx->cg_unsorted.p, aux->num_intersections,
                                      cap32, 32 + tile_bits, (uint64_t*)tid_sorted,
                                      (uint32_t*)aux->compact_gid_from_isect, ss_tile, st);
            else
                launch_radix_sort_u32(tid_unsorted, (uint32_t*)ctx->cg_unsorted.p, aux->num_intersections, cap32,
                                      tile_bits, tid_sorted, (uint32_t*)aux->compact_gid_from_isect, ss_tile, st);
        }
    }
    {   // render.rs:239-262
        StageTimer t(ctx, st, BRUSH_STAGE_TILE_BINS);
        launch_tile_bins(fused ? nullptr : tid_sorted, fused ? (uint64_t*)tid_sorted : nullptr,
                         aux->num_intersections, cap, num_tiles, (uint32_t*)aux->tile_bins, st);
    }
    if (defer_colors) {
        if (sh_ready) CU(cudaStreamWaitEvent(st, (cudaEvent_t)sh_ready, 0));
        NvtxRange span("ProjectColors");
        launch_project_colors(cam, N, K, means, sh_coeffs, gfc, aux->num_visible, aux->projected_splats, st);
    }
    {   // render.rs:267-307
        StageTimer t(ctx, st, BRUSH_STAGE_RASTERIZE);
        launch_rasterize(cam, (uint32_t*)aux->compact_gid_from_isect, (uint32_t*)aux->tile_bins,
                         aux->projected_splats, raster_u32, out_img, (uint32_t*)aux->final_index,
                         pair_counter_of(ctx, 0), st);
    }
    CU(cudaGetLastError());
    return BRUSH_OK;
}

int brush_render_backward(brush_ctx* ctx, uint32_t N, uint32_t K, const float* means, const float* log_scales,
                          const float* quats, const float* raw_opacity, const float* out_img,
                          const float* v_out_img, const BrushAux* aux, float* v_means, float* v_xy,
                          float* v_log_scales, float* v_quats, float* v_sh_coeffs, float* v_raw_opacity,
                          void* stream) {
    if (!ctx) return BRUSH_E_INVALID;
    if (!valid_k(K)) return fail(ctx, BRUSH_E_INVALID, "Invalid nr. of sh bases " + std::to_string(K));
    if (!aux || !out_img || !v_out_img ||
        (N > 0 && (!v_means || !v_xy || !v_log_scales || !v_quats || !v_sh_coeffs || !v_raw_opacity)))
        return fail(ctx, BRUSH_E_INVALID, "null argument");
    cudaStream_t st = (cudaStream_t)stream;
    CU(cudaSetDevice(ctx->device));
    NvtxRange span("render_gaussians backwards");  // render.rs:474
    BrushUniforms u;
    {
        int r = uniforms_of(ctx, aux, N, K, st, &u);
        if (r != BRUSH_OK) return r;
    }
    const Cam cam = make_cam(u);
    ENSURE(v_projected, (size_t)N * VP_STRIDE * 4 + 16);
    ctx->ev_used[BRUSH_STAGE_RASTERIZE_BWD] = ctx->ev_used[BRUSH_STAGE_GATHER_PROJECT_BWD] = false;
    if (N) {
        // The reference zero-fills nine accumulators and six outputs (render.rs:505-507,539-540,547,573-575).  Here
        // only the rasterizer's accumulator is zeroed: the per-splat backward writes every row of every output,
        // zeros included, in one coalesced pass.
        CU(zero_async(ctx->v_projected.p, (size_t)N * VP_STRIDE * 4, st));
    }
    {   // render.rs:515-532
        StageTimer t(ctx, st, BRUSH_STAGE_RASTERIZE_BWD);
        launch_rasterize_backward(cam, (uint32_t*)aux->compact_gid_from_isect, (uint32_t*)aux->tile_bins,
                                  aux->projected_splats, (uint32_t*)aux->global_from_compact_gid,
                                  (uint32_t*)aux->final_index, out_img, v_out_img, (float*)ctx->v_projected.p,
                                  pair_counter_of(ctx, 1), st);
    }
    {   // render.rs:542-594
        StageTimer t(ctx, st, BRUSH_STAGE_GATHER_PROJECT_BWD);
        launch_gather_project_backward(cam, N, K, means, log_scales, quats, raw_opacity,
                                       (float*)ctx->v_projected.p, v_means, v_xy, v_log_scales, v_quats, v_sh_coeffs,
                                       v_raw_opacity, nullptr, nullptr, st);
    }
    CU(cudaGetLastError());
    return BRUSH_OK;
}

// Per-gid visibility of the forward that produced `aux`: project_forward leaves it in the ctx; if another forward
// has run on this ctx since, it is rebuilt from the aux's compact list.
static int visible_mask_for(brush_ctx* ctx, const BrushAux* aux, uint32_t N, cudaStream_t st, const uint8_t** out) {
    ENSURE(visible_mask, (size_t)N + 16);
    if (ctx->visible_owner != aux->uniforms_buffer) {
        launch_visible_mask_from_list((const uint32_t*)aux->global_from_compact_gid, aux->num_visible, N,
                                      (uint8_t*)ctx->visible_mask.p, st);
        ctx->visible_owner = aux->uniforms_buffer;
    }
    *out = (const uint8_t*)ctx->visible_mask.p;
    return BRUSH_OK;
}

static bool train_state_ok(const BrushTrainState& s) {
    return s.means && s.log_scales && s.rotation && s.sh_coeffs && s.raw_opacity && s.m_means && s.v_means &&
           s.m_log_scales && s.v_log_scales && s.m_rotation && s.v_rotation && s.m_sh_coeffs && s.v_sh_coeffs &&
           s.m_raw_opacity && s.v_raw_opacity &&
           !(((uintptr_t)s.rotation | (uintptr_t)s.m_rotation | (uintptr_t)s.v_rotation | (uintptr_t)s.sh_coeffs |
              (uintptr_t)s.m_sh_coeffs | (uintptr_t)s.v_sh_coeffs) & 15u);
}
static FusedAdamLaunch fused_adam_launch(const BrushTrainState& s, float beta1, float beta2, float eps, uint32_t step,
                                         float grad_scale) {
    FusedAdamLaunch A{};
    A.means = s.means; A.log_scales = s.log_scales; A.rotation = s.rotation; A.sh = s.sh_coeffs; A.raw_opac = s.raw_opacity;
    A.m_means = s.m_means; A.v_means = s.v_means; A.m_scales = s.m_log_scales; A.v_scales = s.v_log_scales;
    A.m_rot = s.m_rotation; A.v_rot = s.v_rotation; A.m_sh = s.m_sh_coeffs; A.v_sh = s.v_sh_coeffs;
    A.m_opac = s.m_raw_opacity; A.v_opac = s.v_raw_opacity;
    A.lr_mean = s.lr_mean; A.lr_scale = s.lr_scale; A.lr_rot = s.lr_rotation; A.lr_sh_dc = s.lr_coeffs_dc;
    A.lr_sh_rest = s.lr_coeffs_dc * s.sh_rest_lr_scale; A.lr_opac = s.lr_opac;
    A.beta1 = beta1; A.beta2 = beta2; A.eps = eps; A.grad_scale = grad_scale;
    A.inv_bc1 = (float)(1.0 / (1.0 - std::pow((double)beta1, (double)step)));
    A.inv_bc2 = (float)(1.0 / (1.0 - std::pow((double)beta2, (double)step)));
    return A;
}

static int views_launch(brush_ctx* ctx, const BrushViewExchange* x, ViewsLaunch* X) {
    if (!x || x->world < 1 || x->world > BRUSH_DP_MAX_WORLD || x->rank < 0 || x->rank >= x->world || x->epoch < 1)
        return fail(ctx, BRUSH_E_INVALID, "view exchange: world 1..8, 0 <= rank < world, epoch >= 1");
    *X = ViewsLaunch{};
    X->world = x->world; X->rank = x->rank; X->epoch = x->epoch;
    for (int r = 0; r < x->world; r++) {
        if (!x->rows[r] || !x->flags[r]) return fail(ctx, BRUSH_E_INVALID, "view exchange: null peer buffer");
        if ((uintptr_t)x->rows[r] & 15u) return fail(ctx, BRUSH_E_INVALID, "view exchange: rows must be 16-byte aligned");
        X->rows[r] = x->rows[r]; X->flags[r] = x->flags[r];
    }
    return BRUSH_OK;
}

int brush_render_backward_rows(brush_ctx* ctx, uint32_t N, uint32_t K, const float* out_img, const float* v_out_img,
                               const BrushAux* aux, const float* loss_term, const BrushViewExchange* x,
                               void* stream) {
    if (!ctx) return BRUSH_E_INVALID;
    if (!valid_k(K)) return fail(ctx, BRUSH_E_INVALID, "Invalid nr. of sh bases " + std::to_string(K));
    if (!aux || !out_img || !v_out_img || N == 0) return fail(ctx, BRUSH_E_INVALID, "null argument or N == 0");
    ViewsLaunch X;
    int r = views_launch(ctx, x, &X);
    if (r != BRUSH_OK) return r;
    cudaStream_t st = (cudaStream_t)stream;
    CU(cudaSetDevice(ctx->device));
    BrushUniforms u;
    {
        int r = uniforms_of(ctx, aux, N, K, st, &u);
        if (r != BRUSH_OK) return r;
    }
    const Cam cam = make_cam(u);
    float* rows = x->rows[x->rank];
    ctx->ev_used[BRUSH_STAGE_RASTERIZE_BWD] = ctx->ev_used[BRUSH_STAGE_GATHER_PROJECT_BWD] = false;
    CU(zero_async(rows, (size_t)N * VP_STRIDE * 4, st));
    {
        StageTimer t(ctx, st, BRUSH_STAGE_RASTERIZE_BWD);
        launch_rasterize_backward(cam, (uint32_t*)aux->compact_gid_from_isect, (uint32_t*)aux->tile_bins,
                                  aux->projected_splats, (uint32_t*)aux->global_from_compact_gid,
                                  (uint32_t*)aux->final_index, out_img, v_out_img, rows, pair_counter_of(ctx, 1), st);
    }
    const uint8_t* mask = nullptr;
    r = visible_mask_for(ctx, aux, N, st, &mask);
    if (r != BRUSH_OK) return r;
    launch_publish_view(rows, N, mask, aux->uniforms_buffer, loss_term, X, st);
    CU(cudaGetLastError());
    return BRUSH_OK;
}

int brush_render_backward_hybrid(brush_ctx* ctx, uint32_t N, uint32_t K, const float* means, const float* log_scales,
                                 const float* quats, const float* raw_opacity, const float* out_img,
                                 const float* v_out_img, const BrushAux* aux, float* v_means, float* v_xy,
                                 float* v_log_scales, float* v_quats, float* v_raw_opacity, const float* loss_term,
                                 const BrushViewExchange* x, void* stream) {
    if (!ctx) return BRUSH_E_INVALID;
    if (!valid_k(K)) return fail(ctx, BRUSH_E_INVALID, "Invalid nr. of sh bases " + std::to_string(K));
    if (!aux || !out_img || !v_out_img || N == 0 || !means || !log_scales || !quats || !raw_opacity || !v_means ||
        !v_xy || !v_log_scales || !v_quats || !v_raw_opacity)
        return fail(ctx, BRUSH_E_INVALID, "null argument or N == 0");
    ViewsLaunch X;
    int r = views_launch(ctx, x, &X);
    if (r != BRUSH_OK) return r;
    cudaStream_t st = (cudaStream_t)stream;
    CU(cudaSetDevice(ctx->device));
    NvtxRange span("render_gaussians backwards");
    BrushUniforms u;
    r = uniforms_of(ctx, aux, N, K, st, &u);
    if (r != BRUSH_OK) return r;
    const Cam cam = make_cam(u);
    ENSURE(v_projected, (size_t)N * VP_STRIDE * 4 + 16);
    ctx->ev_used[BRUSH_STAGE_RASTERIZE_BWD] = ctx->ev_used[BRUSH_STAGE_GATHER_PROJECT_BWD] = false;
    CU(zero_async(ctx->v_projected.p, (size_t)N * VP_STRIDE * 4, st));
    {
        StageTimer t(ctx, st, BRUSH_STAGE_RASTERIZE_BWD);
        launch_rasterize_backward(cam, (uint32_t*)aux->compact_gid_from_isect, (uint32_t*)aux->tile_bins,
                                  aux->projected_splats, (uint32_t*)aux->global_from_compact_gid,
                                  (uint32_t*)aux->final_index, out_img, v_out_img, (float*)ctx->v_projected.p,
                                  pair_counter_of(ctx, 1), st);
    }
    const uint8_t* mask = nullptr;
    r = visible_mask_for(ctx, aux, N, st, &mask);
    if (r != BRUSH_OK) return r;
    float* rows = x->rows[x->rank];
    {
        StageTimer t(ctx, st, BRUSH_STAGE_GATHER_PROJECT_BWD);
        launch_gather_project_backward(cam, N, K, means, log_scales, quats, raw_opacity, (float*)ctx->v_projected.p,
                                       v_means, v_xy, v_log_scales, v_quats, nullptr, v_raw_opacity, rows, mask, st);
    }
    launch_publish_colour_rows(rows, N, aux->uniforms_buffer, loss_term, X, st);
    CU(cudaGetLastError());
    return BRUSH_OK;
}

int brush_sh_views_adam_step(brush_ctx* ctx, uint32_t N, uint32_t K, const float* means, float* sh_coeffs,
                             float* m_sh_coeffs, float* v_sh_coeffs, float lr_coeffs_dc, float sh_rest_lr_scale,
                             const BrushViewExchange* x, float* loss_out, float beta1, float beta2, float eps,
                             uint32_t step, float grad_scale, void* stream) {
    if (!ctx) return BRUSH_E_INVALID;
    if (!valid_k(K)) return fail(ctx, BRUSH_E_INVALID, "Invalid nr. of sh bases " + std::to_string(K));
    if (!means || !sh_coeffs || !m_sh_coeffs || !v_sh_coeffs || N == 0 || step < 1 ||
        (((uintptr_t)sh_coeffs | (uintptr_t)m_sh_coeffs | (uintptr_t)v_sh_coeffs) & 15u))
        return fail(ctx, BRUSH_E_INVALID, "null / misaligned argument or N == 0");
    ViewsLaunch X;
    int r = views_launch(ctx, x, &X);
    if (r != BRUSH_OK) return r;
    X.loss_out = loss_out;
    CU(cudaSetDevice(ctx->device));
    const float inv_bc1 = (float)(1.0 / (1.0 - std::pow((double)beta1, (double)step)));
    const float inv_bc2 = (float)(1.0 / (1.0 - std::pow((double)beta2, (double)step)));
    if (launch_sh_views_adam(X, N, K, means, sh_coeffs, m_sh_coeffs, v_sh_coeffs, lr_coeffs_dc,
                             lr_coeffs_dc * sh_rest_lr_scale, beta1, beta2, eps, inv_bc1, inv_bc2, grad_scale,
                             (cudaStream_t)stream) != 0)
        return fail(ctx, BRUSH_E_CUDA, "sh_views_adam launch failed");
    CU(cudaGetLastError());
    return BRUSH_OK;
}

int brush_views_adam_step(brush_ctx* ctx, uint32_t N, uint32_t K, const BrushTrainState* state, const float* quats,
                          const BrushViewExchange* x, float* grad_norm_accum, float* visible_accum, float* loss_out,
                          float beta1, float beta2, float eps, uint32_t step, float grad_scale, void* stream) {
    if (!ctx) return BRUSH_E_INVALID;
    if (!valid_k(K)) return fail(ctx, BRUSH_E_INVALID, "Invalid nr. of sh bases " + std::to_string(K));
    if (!state || !quats || N == 0 || step < 1 || ((uintptr_t)quats & 15u) || !train_state_ok(*state) ||
        (!grad_norm_accum) != (!visible_accum))
        return fail(ctx, BRUSH_E_INVALID, "null / misaligned argument or N == 0");
    ViewsLaunch X;
    int r = views_launch(ctx, x, &X);
    if (r != BRUSH_OK) return r;
    X.loss_out = loss_out; X.grad_norm_accum = grad_norm_accum; X.visible_accum = visible_accum;
    cudaStream_t st = (cudaStream_t)stream;
    CU(cudaSetDevice(ctx->device));
    const FusedAdamLaunch A = fused_adam_launch(*state, beta1, beta2, eps, step, grad_scale);
    {
        StageTimer t(ctx, st, BRUSH_STAGE_GATHER_PROJECT_BWD);
        if (launch_views_adam(X, N, K, quats, A, st) != 0) return fail(ctx, BRUSH_E_CUDA, "views_adam launch failed");
    }
    CU(cudaGetLastError());
    return BRUSH_OK;
}

int brush_render_backward_adam(brush_ctx* ctx, uint32_t N, uint32_t K, const BrushTrainState* state,
                               const float* quats, const float* out_img, const float* v_out_img,
                               const BrushAux* aux, float* v_xy, float* grad_norm_accum, float* visible_accum,
                               float beta1, float beta2, float eps, uint32_t step, float grad_scale, void* stream) {
    if (!ctx) return BRUSH_E_INVALID;
    if (!valid_k(K)) return fail(ctx, BRUSH_E_INVALID, "Invalid nr. of sh bases " + std::to_string(K));
    if (!aux || !out_img || !v_out_img || !state || step < 1) return fail(ctx, BRUSH_E_INVALID, "null argument");
    const BrushTrainState& s = *state;
    if ((!grad_norm_accum) != (!visible_accum)) return fail(ctx, BRUSH_E_INVALID, "statistics: both accumulators or neither");
    if (N > 0 && (!quats || !s.means || !s.log_scales || !s.rotation || !s.sh_coeffs || !s.raw_opacity ||
                  !s.m_means || !s.v_means || !s.m_log_scales || !s.v_log_scales || !s.m_rotation || !s.v_rotation ||
                  !s.m_sh_coeffs || !s.v_sh_coeffs || !s.m_raw_opacity || !s.v_raw_opacity))
        return fail(ctx, BRUSH_E_INVALID, "null argument");
    if (((uintptr_t)s.rotation | (uintptr_t)s.m_rotation | (uintptr_t)s.v_rotation | (uintptr_t)s.sh_coeffs |
         (uintptr_t)s.m_sh_coeffs | (uintptr_t)s.v_sh_coeffs | (uintptr_t)quats) & 15u)
        return fail(ctx, BRUSH_E_INVALID, "rotation / sh_coeffs buffers must be 16-byte aligned");
    cudaStream_t st = (cudaStream_t)stream;
    CU(cudaSetDevice(ctx->device));
    BrushUniforms u;
    {
        int r = uniforms_of(ctx, aux, N, K, st, &u);
        if (r != BRUSH_OK) return r;
    }
    const Cam cam = make_cam(u);
    ENSURE(v_projected, (size_t)N * VP_STRIDE * 4 + 16);
    ctx->ev_used[BRUSH_STAGE_RASTERIZE_BWD] = ctx->ev_used[BRUSH_STAGE_GATHER_PROJECT_BWD] = false;
    if (N) CU(zero_async(ctx->v_projected.p, (size_t)N * VP_STRIDE * 4, st));
    {
        StageTimer t(ctx, st, BRUSH_STAGE_RASTERIZE_BWD);
        launch_rasterize_backward(cam, (uint32_t*)aux->compact_gid_from_isect, (uint32_t*)aux->tile_bins,
                                  aux->projected_splats, (uint32_t*)aux->global_from_compact_gid,
                                  (uint32_t*)aux->final_index, out_img, v_out_img, (float*)ctx->v_projected.p,
                                  pair_counter_of(ctx, 1), st);
    }
    const FusedAdamLaunch A = fused_adam_launch(s, beta1, beta2, eps, step, grad_scale);
    {   // gather_grads + project_backwards + normalisation VJP + Adam in one pass
        StageTimer t(ctx, st, BRUSH_STAGE_GATHER_PROJECT_BWD);
        const uint8_t* mask = nullptr;
        if (grad_norm_accum) {
            int r = visible_mask_for(ctx, aux, N, st, &mask);
            if (r != BRUSH_OK) return r;
        }
        launch_backward_adam(cam, N, K, quats, (float*)ctx->v_projected.p, A, v_xy, mask, grad_norm_accum,
                             visible_accum, st);
    }
    CU(cudaGetLastError());
    return BRUSH_OK;
}

int brush_radix_argsort_u32(brush_ctx* ctx, const uint32_t* keys, const uint32_t* values, const int32_t* n_dev,
                            uint32_t max_n, uint32_t bits, uint32_t* keys_out, uint32_t* values_out, void* stream) {
    if (!ctx) return BRUSH_E_INVALID;
    if (bits > 32) return fail(ctx, BRUSH_E_INVALID, "sorting_bits must be <= 32");  // brush-sort/src/lib.rs:39
    if (max_n >= (1u << 30)) return fail(ctx, BRUSH_E_INVALID, "max_n must be < 2^30");
    if (max_n && (!keys || !values || !n_dev || !keys_out || !values_out))
        return fail(ctx, BRUSH_E_INVALID, "null argument");
    CU(cudaSetDevice(ctx->device));
    SortScratch ss;
    int r = sort_scratch(ctx, max_n, 4, bits, &ss);
    if (r != BRUSH_OK) return r;
    launch_radix_sort_u32(keys, values, n_dev, max_n, bits, keys_out, values_out, ss, (cudaStream_t)stream);
    CU(cudaGetLastError());
    return BRUSH_OK;
}

int brush_radix_argsort_u64(brush_ctx* ctx, const uint64_t* keys, const uint32_t* values, const int32_t* n_dev,
                            uint32_t max_n, uint32_t bits, uint64_t* keys_out, uint32_t* values_out, void* stream) {
    if (!ctx) return BRUSH_E_INVALID;
    if (bits > 64) return fail(ctx, BRUSH_E_INVALID, "sorting_bits must be <= 64");
    if (max_n >= (1u << 30)) return fail(ctx, BRUSH_E_INVALID, "max_n must be < 2^30");
    if (max_n && (!keys || !values || !n_dev || !keys_out || !values_out))
        return fail(ctx, BRUSH_E_INVALID, "null argument");
    CU(cudaSetDevice(ctx->device));
    SortScratch ss;
    int r = sort_scratch(ctx, max_n, 8, bits, &ss);
    if (r != BRUSH_OK) return r;
    launch_radix_sort_u64(keys, values, n_dev, max_n, bits, keys_out, values_out, ss, (cudaStream_t)stream);
    CU(cudaGetLastError());
    return BRUSH_OK;
}

int brush_prefix_sum_u32(brush_ctx* ctx, const uint32_t* in, uint32_t n, uint32_t* out, void* stream) {
    if (!ctx) return BRUSH_E_INVALID;
    if (n && (!in || !out)) return fail(ctx, BRUSH_E_INVALID, "null argument");
    if (n == 0) return BRUSH_OK;
    CU(cudaSetDevice(ctx->device));
    cudaStream_t st = (cudaStream_t)stream;
    size_t bytes = 64 + scan_status_words(n) * 8;
    ENSURE(control, bytes);
    CU(zero_async(ctx->control.p, bytes, st));
    launch_inclusive_scan(in, n, out, (uint64_t*)((char*)ctx->control.p + 64), (uint32_t*)ctx->control.p, nullptr,
                          ~0ull, nullptr, st);
    CU(cudaGetLastError());
    return BRUSH_OK;
}

int brush_render_forward_host(brush_ctx* ctx, const BrushUniforms* uniforms_host, uint32_t N, uint32_t K,
                              const float* means_host, const float* log_scales_host, const float* quats_host,
                              const float* sh_coeffs_host, const float* raw_opacity_host, uint64_t isect_capacity,
                              float* out_img_host, uint32_t* num_visible_host, uint32_t* num_intersections_host) {
    if (!ctx) return BRUSH_E_INVALID;
    if (!uniforms_host || !out_img_host) return fail(ctx, BRUSH_E_INVALID, "null argument");
    if (!valid_k(K)) return fail(ctx, BRUSH_E_INVALID, "Invalid nr. of sh bases " + std::to_string(K));
    CU(cudaSetDevice(ctx->device));
    const uint32_t W = uniforms_host->img_size[0], H = uniforms_host->img_size[1];
    const uint32_t tx = (W + 15) / 16, ty = (H + 15) / 16;
    const size_t n1 = N ? N : 1;
    ENSURE(h_means, n1 * 12);
    ENSURE(h_lscales, n1 * 12);
    ENSURE(h_quats, n1 * 16);
    ENSURE(h_sh, n1 * K * 12);
    ENSURE(h_opac, n1 * 4);
    ENSURE(h_img, (size_t)W * H * 16 + 16);
    const size_t aux_bytes[9] = {n1 * 36, 128, 16, 16, (size_t)W * H * 4 + 16, n1 * 4, (size_t)tx * ty * 8 + 16,
                                 (size_t)(isect_capacity ? isect_capacity : 1) * 4, n1 * 4};
    for (int i = 0; i < 9; i++) {
        int r = ensure(ctx, ctx->h_aux[i], aux_bytes[i]);
        if (r != BRUSH_OK) return r;
    }
    cudaStream_t st = 0;
    if (N) {
        CU(cudaMemcpyAsync(ctx->h_means.p, means_host, (size_t)N * 12, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(ctx->h_lscales.p, log_scales_host, (size_t)N * 12, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(ctx->h_quats.p, quats_host, (size_t)N * 16, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(ctx->h_sh.p, sh_coeffs_host, (size_t)N * K * 12, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(ctx->h_opac.p, raw_opacity_host, (size_t)N * 4, cudaMemcpyHostToDevice, st));
    }
    BrushAux aux;
    aux.projected_splats = (float*)ctx->h_aux[0].p;
    aux.uniforms_buffer = (int32_t*)ctx->h_aux[1].p;
    aux.num_intersections = (int32_t*)ctx->h_aux[2].p;
    aux.num_visible = (int32_t*)ctx->h_aux[3].p;
    aux.final_index = (int32_t*)ctx->h_aux[4].p;
    aux.cum_tiles_hit = (int32_t*)ctx->h_aux[5].p;
    aux.tile_bins = (int32_t*)ctx->h_aux[6].p;
    aux.compact_gid_from_isect = (int32_t*)ctx->h_aux[7].p;
    aux.global_from_compact_gid = (int32_t*)ctx->h_aux[8].p;
    aux.tile_id_from_isect = nullptr;
    aux.isect_capacity = isect_capacity;
    int r = brush_render_forward(ctx, uniforms_host, N, K, (float*)ctx->h_means.p, (float*)ctx->h_lscales.p,
                                 (float*)ctx->h_quats.p, (float*)ctx->h_sh.p, (float*)ctx->h_opac.p, 0,
                                 ctx->h_img.p, &aux, st);
    if (r != BRUSH_OK) return r;
    CU(cudaMemcpyAsync(out_img_host, ctx->h_img.p, (size_t)W * H * 16, cudaMemcpyDeviceToHost, st));
    if (num_visible_host) CU(cudaMemcpyAsync(num_visible_host, aux.num_visible, 4, cudaMemcpyDeviceToHost, st));
    if (num_intersections_host)
        CU(cudaMemcpyAsync(num_intersections_host, aux.num_intersections, 4, cudaMemcpyDeviceToHost, st));
    return brush_ctx_check(ctx, st);
}


static RefineLaunchCfg refine_cfg(const BrushRefineConfig* c) {
    RefineLaunchCfg r;
    r.grad_thresh = c->densify_grad_thresh;
    r.log_size_thresh = logf(c->densify_size_thresh);
    r.logit_alpha = std::log((double)c->cull_alpha_thresh / (1.0 - (double)c->cull_alpha_thresh));
    r.log_scale_cull = logf(c->cull_scale_thresh);
    r.reset_alpha = c->reset_alpha;
    r.reset_alpha_raw = (float)std::log((double)c->reset_alpha_value / (1.0 - (double)c->reset_alpha_value));
    return r;
}
static bool params_ok(const BrushSplatParams* p) {
    return p && p->means && p->rotation && p->sh_coeffs && p->raw_opacity && p->log_scales;
}

int brush_refine_plan(brush_ctx* ctx, uint32_t n, const float* grad_2d_accum, const float* xy_grad_counts,
                      const BrushSplatParams* pre_step, const BrushSplatParams* post_step,
                      const BrushRefineConfig* config, BrushRefineCounts* counts_host, void* stream) {
    if (!ctx) return BRUSH_E_INVALID;
    if (!grad_2d_accum || !xy_grad_counts || !params_ok(pre_step) || !params_ok(post_step) || !config || !counts_host)
        return fail(ctx, BRUSH_E_INVALID, "null argument");
    if (!(config->cull_alpha_thresh > 0.f && config->cull_alpha_thresh < 1.f) || !(config->densify_size_thresh > 0.f) ||
        !(config->cull_scale_thresh > 0.f))
        return fail(ctx, BRUSH_E_INVALID, "refine: thresholds out of range");
    memset(counts_host, 0, sizeof(*counts_host));
    ctx->refine_n = n;
    if (n == 0) return BRUSH_OK;
    CU(cudaSetDevice(ctx->device));
    cudaStream_t st = (cudaStream_t)stream;
    ENSURE(refine_scratch, refine_scratch_bytes(n));
    uint32_t* totals_dev = nullptr;
    launch_refine_plan(n, grad_2d_accum, xy_grad_counts, pre_step->log_scales, pre_step->raw_opacity,
                       post_step->log_scales, post_step->raw_opacity, refine_cfg(config), ctx->refine_scratch.p,
                       &totals_dev, st);
    CU(cudaGetLastError());
    uint32_t t[6];
    CU(cudaMemcpyAsync(t, totals_dev, sizeof(t), cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));  // the caller allocates the new set from these counts (train.rs awaits argwhere)
    counts_host->num_out = t[0] + t[1] + t[2];
    counts_host->num_split = t[3];
    counts_host->alpha_pruned = t[4];
    counts_host->scale_pruned = t[4] + t[5];  // cumulative, train.rs:555
    // every row is either kept or counted by one of the two prune tests: n + clones + splits = kept + pruned
    counts_host->num_cloned = counts_host->num_out + t[4] + t[5] - n - t[3];
    return BRUSH_OK;
}

int brush_refine_apply(brush_ctx* ctx, uint32_t n, uint32_t K, const BrushSplatParams* pre_step,
                       const BrushSplatParams* post_step, const float* samples, const BrushRefineConfig* config,
                       const BrushSplatParams* out, void* stream) {
    if (!ctx) return BRUSH_E_INVALID;
    if (!params_ok(pre_step) || !params_ok(post_step) || !config) return fail(ctx, BRUSH_E_INVALID, "null argument");
    if (!valid_k(K)) return fail(ctx, BRUSH_E_INVALID, "Invalid nr. of sh bases");
    if (n != ctx->refine_n || (n && !ctx->refine_scratch.p))
        return fail(ctx, BRUSH_E_INVALID, "brush_refine_apply must follow brush_refine_plan with the same n");
    if (n == 0) return BRUSH_OK;
    if (!out) return fail(ctx, BRUSH_E_INVALID, "null argument");
    CU(cudaSetDevice(ctx->device));
    const float* pre[5] = {pre_step->means, pre_step->rotation, pre_step->sh_coeffs, pre_step->raw_opacity, pre_step->log_scales};
    const float* post[5] = {post_step->means, post_step->rotation, post_step->sh_coeffs, post_step->raw_opacity, post_step->log_scales};
    float* o[5] = {out->means, out->rotation, out->sh_coeffs, out->raw_opacity, out->log_scales};
    launch_refine_apply(n, K, pre, post, samples, refine_cfg(config), ctx->refine_scratch.p, o, (cudaStream_t)stream);
    CU(cudaGetLastError());
    return BRUSH_OK;
}

uint64_t brush_launch_count(void) { return g_launch_count.load(); }

int brush_image_u8_to_f32(brush_ctx* ctx, const uint8_t* src, uint64_t n, float* dst, void* stream) {
    if (!ctx) return BRUSH_E_INVALID;
    if (n && (!src || !dst)) return fail(ctx, BRUSH_E_INVALID, "null argument");
    if (((uintptr_t)src & 3u) || ((uintptr_t)dst & 15u)) return fail(ctx, BRUSH_E_INVALID, "image buffers must be aligned");
    CU(cudaSetDevice(ctx->device));
    launch_image_u8_to_f32(src, (size_t)n, dst, (cudaStream_t)stream);
    CU(cudaGetLastError());
    return BRUSH_OK;
}
int brush_quat_normalize(brush_ctx* ctx, const float* quats, uint32_t n, float* out, void* stream) {
    if (!ctx) return BRUSH_E_INVALID;
    if (n && (!quats || !out)) return fail(ctx, BRUSH_E_INVALID, "null argument");
    CU(cudaSetDevice(ctx->device));
    launch_quat_normalize(quats, n, out, (cudaStream_t)stream);
    CU(cudaGetLastError());
    return BRUSH_OK;
}
int brush_quat_normalize_vjp(brush_ctx* ctx, const float* quats, uint32_t n, float* v_inout, void* stream) {
    if (!ctx) return BRUSH_E_INVALID;
    if (n && (!quats || !v_inout)) return fail(ctx, BRUSH_E_INVALID, "null argument");
    CU(cudaSetDevice(ctx->device));
    launch_quat_normalize_vjp(quats, n, v_inout, (cudaStream_t)stream);
    CU(cudaGetLastError());
    return BRUSH_OK;
}
int brush_xy_grad_stats(brush_ctx* ctx, const float* v_xy, const BrushAux* aux, uint32_t n, uint32_t width,
                        uint32_t height, int accumulate, float* grad_norm, float* visible, void* stream) {
    if (!ctx) return BRUSH_E_INVALID;
    if (!aux || (n && (!v_xy || !grad_norm || !visible))) return fail(ctx, BRUSH_E_INVALID, "null argument");
    CU(cudaSetDevice(ctx->device));
    const uint8_t* mask = nullptr;
    if (n) {
        int r = visible_mask_for(ctx, aux, n, (cudaStream_t)stream, &mask);
        if (r != BRUSH_OK) return r;
    }
    launch_xy_grad_stats(v_xy, mask, n, 0.5f * width, 0.5f * height, accumulate, grad_norm, visible,
                         (cudaStream_t)stream);
    CU(cudaGetLastError());
    return BRUSH_OK;
}

int brush_l1_loss_backward(brush_ctx* ctx, const float* pred, const float* gt, uint32_t height, uint32_t width,
                           uint32_t gt_channels, float* v_out_img, float* loss_out, void* stream) {
    if (!ctx) return BRUSH_E_INVALID;
    if (!pred || !gt || !v_out_img || !loss_out) return fail(ctx, BRUSH_E_INVALID, "null argument");
    if (gt_channels != 3 && gt_channels != 4) return fail(ctx, BRUSH_E_INVALID, "gt_channels must be 3 or 4");
    CU(cudaSetDevice(ctx->device));
    cudaStream_t st = (cudaStream_t)stream;
    CU(zero_async(loss_out, 4, st));
    launch_l1_loss_backward(pred, gt, height * width, gt_channels, v_out_img, loss_out, st);
    CU(cudaGetLastError());
    return BRUSH_OK;
}

int brush_image_loss_backward(brush_ctx* ctx, const float* pred, const float* gt, uint32_t height, uint32_t width,
                              uint32_t gt_channels, float ssim_weight, uint32_t ssim_window, float* v_out_img,
                              float* loss_out, void* stream) {
    if (!ctx) return BRUSH_E_INVALID;
    if (!pred || !gt || !v_out_img || !loss_out) return fail(ctx, BRUSH_E_INVALID, "null argument");
    if (gt_channels != 3 && gt_channels != 4) return fail(ctx, BRUSH_E_INVALID, "gt_channels must be 3 or 4");
    if (!(ssim_weight >= 0.f && ssim_weight <= 1.f)) return fail(ctx, BRUSH_E_INVALID, "ssim_weight must be in [0, 1]");
    if (ssim_weight > 0.f && ssim_window != 11)
        return fail(ctx, BRUSH_E_INVALID, "ssim_window must be 11 (TrainConfig::ssim_window_size default)");
    if (height == 0 || width == 0) return fail(ctx, BRUSH_E_INVALID, "empty image");
    CU(cudaSetDevice(ctx->device));
    const size_t plane = (size_t)(height + 2) * (width + 2);
    ENSURE(loss_scratch, 64 + 9 * plane * 4);
    float* sums = (float*)ctx->loss_scratch.p;
    launch_image_loss_backward(pred, gt, height, width, gt_channels, ssim_weight, sums + 16, sums, v_out_img,
                               loss_out, (cudaStream_t)stream);
    CU(cudaGetLastError());
    return BRUSH_OK;
}

int brush_adam_step(brush_ctx* ctx, const BrushAdamGroup* groups_host, int n_groups, float beta1, float beta2,
                    float eps, uint32_t step, float grad_scale, void* stream) {
    if (!ctx) return BRUSH_E_INVALID;
    if (!groups_host || n_groups < 1 || n_groups > 8 || step < 1)
        return fail(ctx, BRUSH_E_INVALID, "adam: 1..8 groups, step >= 1");
    CU(cudaSetDevice(ctx->device));
    AdamLaunch L{};
    L.n_groups = n_groups;
    for (int i = 0; i < n_groups; i++) {
        const BrushAdamGroup& g = groups_host[i];
        if (g.count && (!g.param || !g.grad || !g.m || !g.v)) return fail(ctx, BRUSH_E_INVALID, "adam: null group buffer");
        L.seg[i] = AdamSeg{g.param, g.grad, g.m, g.v, g.count, g.lr, g.sh_coeffs_per_channel, g.sh_rest_lr_scale, 0};
    }
    L.beta1 = beta1; L.beta2 = beta2; L.eps = eps; L.grad_scale = grad_scale;
    L.inv_bc1 = (float)(1.0 / (1.0 - std::pow((double)beta1, (double)step)));
    L.inv_bc2 = (float)(1.0 / (1.0 - std::pow((double)beta2, (double)step)));
    launch_adam(L, (cudaStream_t)stream);
    CU(cudaGetLastError());
    return BRUSH_OK;
}

void brush_dp_owned_range(uint64_t count, int world, int rank, uint64_t* begin, uint64_t* end) {
    const uint64_t n4 = count / 4;
    if (begin) *begin = 4 * (n4 * (uint64_t)rank / (uint64_t)(world > 0 ? world : 1));
    if (end) *end = 4 * (n4 * (uint64_t)(rank + 1) / (uint64_t)(world > 0 ? world : 1));
}

int brush_dp_adam_step(brush_ctx* ctx, const BrushDpStep* sp, const BrushDpSegment* segs, int n_segs, float beta1,
                       float beta2, float eps, uint32_t adam_step, float grad_scale, void* stream) {
    if (!ctx) return BRUSH_E_INVALID;
    if (!sp || !segs || n_segs < 1 || n_segs > 8 || adam_step < 1 || sp->epoch < 1)
        return fail(ctx, BRUSH_E_INVALID, "dp_adam: 1..8 segments, adam_step >= 1, epoch >= 1");
    if (sp->world < 1 || sp->world > BRUSH_DP_MAX_WORLD || sp->rank < 0 || sp->rank >= sp->world)
        return fail(ctx, BRUSH_E_INVALID, "dp_adam: world must be 1..8 and 0 <= rank < world");
    CU(cudaSetDevice(ctx->device));
    DpLaunch L{};
    L.world = sp->world; L.rank = sp->rank; L.n_segs = n_segs;
    for (int r = 0; r < sp->world; r++) {
        if (!sp->src[r] || !sp->dst[r] || !sp->flags[r]) return fail(ctx, BRUSH_E_INVALID, "dp_adam: null peer buffer");
        if (((uintptr_t)sp->src[r] | (uintptr_t)sp->dst[r]) & 15u)
            return fail(ctx, BRUSH_E_INVALID, "dp_adam: peer buffers must be 16-byte aligned");
        L.src[r] = sp->src[r]; L.dst[r] = sp->dst[r]; L.flags[r] = sp->flags[r];
    }
    L.src_mc = sp->src_multicast; L.dst_mc = sp->dst_multicast;
    L.m = sp->m; L.v = sp->v;
    L.loss_off = sp->loss_offset; L.loss_out = sp->loss_out;
    if (sp->lane < 0 || sp->lane > 1 || sp->ctas_per_sm < 0 || sp->ctas_per_sm > 16)
        return fail(ctx, BRUSH_E_INVALID, "dp_adam: lane must be 0 or 1, ctas_per_sm 0..16");
    L.arrive = (uint32_t*)ctx->sticky.p + 16 + sp->lane;
    L.epoch = sp->epoch;
    L.lane = sp->lane;
    L.ctas_per_sm = sp->ctas_per_sm;
    for (int i = 0; i < n_segs; i++) {
        const BrushDpSegment& g = segs[i];
        if ((g.src_offset | g.dst_offset | g.count) & 3u)
            return fail(ctx, BRUSH_E_INVALID, "dp_adam: segment offsets and counts must be multiples of 4 floats");
        if (g.adam && (!sp->m || !sp->v || (((uintptr_t)sp->m | (uintptr_t)sp->v) & 15u)))
            return fail(ctx, BRUSH_E_INVALID, "dp_adam: Adam segments need 16-byte aligned m and v");
        L.seg[i] = DpSeg{g.src_offset, g.dst_offset, g.count, g.adam, g.lr, g.sh_coeffs_per_channel, g.sh_rest_lr_scale};
    }
    L.beta1 = beta1; L.beta2 = beta2; L.eps = eps; L.grad_scale = grad_scale;
    L.inv_bc1 = (float)(1.0 / (1.0 - std::pow((double)beta1, (double)adam_step)));
    L.inv_bc2 = (float)(1.0 / (1.0 - std::pow((double)beta2, (double)adam_step)));
    launch_dp_adam(L, (cudaStream_t)stream);
    CU(cudaGetLastError());
    return BRUSH_OK;
}
#pragma GCC visibility pop
}  // extern "C"
