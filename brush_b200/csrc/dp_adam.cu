// View-parallel optimizer step as ONE kernel over NVLink peer memory.
//
// The reference is single-device.  The data-parallel layout BASELINE.json prescribes -- every rank renders its own
// view, dense per-Gaussian gradients are summed, identical Adam everywhere -- is usually an NCCL all-reduce followed
// by a replicated Adam (3 passes over the gradient and 7 over the parameters' worth of HBM on EVERY rank).  Here
// rank r owns 1/world of every parameter group and one launch does, for the owned slice only:
//     g   = sum over ranks of grad_rank[e]          peer loads over NVLink (or ONE multimem.ld_reduce: the
//                                                    NVSwitch adds the ranks' values in flight)
//     m,v = Adam moments (kept only for the slice: optimizer state is sharded, 1/world of the traffic)
//     p  -= lr * mhat / (sqrt(vhat) + eps)          == brush_adam_step, same formula and operation order
//     param_rank'[e] = p for every rank'            peer stores (or ONE multimem.st broadcast)
// so the exchange is fused with the math tile by tile, replicas stay bit-identical (one owner computes each
// element), and the two refine statistics and the loss ride along as plain-sum segments.
// Cross-GPU ordering uses two flag barriers in peer memory (release/acquire at system scope): "my gradients are
// complete" before the first peer load, and "I have finished reading your gradients and writing your parameters"
// before the kernel retires.
#include <cstdlib>

#include "internal.h"
#include "peer.cuh"

namespace brush {

// Wait until every rank has published `epoch` in slot [which][rank'] of THIS rank's flag block.
__device__ __forceinline__ void wait_flags(const uint32_t* my_flags, int which, int world, uint32_t epoch) {
    if (threadIdx.x < (uint32_t)world) {
        const uint32_t* f = my_flags + which * DP_MAX_WORLD + threadIdx.x;
        // back off between polls: an overlapped exchange waits here beside the next forward's kernels, and hundreds of
        // CTAs hammering one L2 line with system-scope loads slow those kernels down (measured: project_forward 2x)
        while ((int32_t)(ld_acquire_sys(f) - epoch) < 0) __nanosleep(256);
    }
    __syncthreads();
}

__device__ __forceinline__ float adam_elem(float p, float g, float& m, float& v, float lr, const DpLaunch& L) {
    return adam_update_value(p, g, m, v, lr, L.beta1, L.beta2, L.eps, L.inv_bc1, L.inv_bc2, L.grad_scale);
}

template <bool MM_LD, bool MM_ST, int U>
__global__ void __launch_bounds__(256)
dp_adam_kernel(const __grid_constant__ DpLaunch L) {
    const int world = L.world, rank = L.rank;
    const int lane_off = L.lane * 2 * DP_MAX_WORLD;
    uint32_t* my_flags = L.flags[rank] + lane_off;
    // ---- barrier 0: every rank's gradient buffer is complete (the producing kernels precede this launch in
    // stream order on their own GPU; the fence + release publishes them at system scope) ----
    if (blockIdx.x == 0 && threadIdx.x < (uint32_t)world) {
        __threadfence_system();
        st_release_sys(L.flags[threadIdx.x] + lane_off + 0 * DP_MAX_WORLD + rank, L.epoch);
    }
    wait_flags(my_flags, 0, world, L.epoch);

    const uint64_t gid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (int s = 0; s < L.n_segs; s++) {
        const DpSeg& sg = L.seg[s];
        const uint32_t row = sg.sh_k > 1 ? 3u * sg.sh_k : 0u;
        const float lr_rest = sg.lr * sg.sh_rest_scale;
        // owned slice [b4, e4) in units of float4 (segment offsets and counts are multiples of 4 floats)
        const uint64_t n4 = sg.count / 4;
        const uint64_t b4 = n4 * (uint64_t)rank / (uint64_t)world, e4 = n4 * (uint64_t)(rank + 1) / (uint64_t)world;
        // U float4 per thread per iteration, all loads issued before any use: NVLink round trips are ~2 us, so the
        // bytes in flight per SM decide the achieved bandwidth
        for (uint64_t j = b4 + gid; j < e4; j += U * stride) {
            float4 g[U], p[U], m[U], v[U];
            bool ok[U];
#pragma unroll
            for (int u = 0; u < U; u++) {
                const uint64_t ju = j + u * stride;
                ok[u] = ju < e4;
                const uint64_t goff = sg.src_off + 4 * ju, poff = sg.dst_off + 4 * ju;
                g[u] = make_float4(0.f, 0.f, 0.f, 0.f);
                if (!ok[u]) continue;
                if (MM_LD) {
                    g[u] = multimem_ld_reduce_v4(L.src_mc + goff);
                } else {
#pragma unroll 8
                    for (int r = 0; r < world; r++) {  // fixed rank order: every replica of the sum is the same sum
                        const float4 t = ld_peer_v4(L.src[r] + goff);
                        g[u].x += t.x; g[u].y += t.y; g[u].z += t.z; g[u].w += t.w;
                    }
                }
                if (sg.adam) {
                    p[u] = *reinterpret_cast<const float4*>(L.dst[rank] + poff);
                    m[u] = *reinterpret_cast<const float4*>(L.m + poff);
                    v[u] = *reinterpret_cast<const float4*>(L.v + poff);
                }
            }
#pragma unroll
            for (int u = 0; u < U; u++) {
                if (!ok[u]) continue;
                const uint64_t ju = j + u * stride;
                const uint64_t poff = sg.dst_off + 4 * ju;
                float4 out;
                if (sg.adam) {
                    float lr0 = sg.lr, lr1 = sg.lr, lr2 = sg.lr, lr3 = sg.lr;
                    if (row) {
                        const uint32_t r0 = (uint32_t)((4 * ju) % row);
                        auto lr_of = [&](uint32_t e) { return (e >= row ? e - row : e) >= 3u ? lr_rest : sg.lr; };
                        lr0 = lr_of(r0); lr1 = lr_of(r0 + 1u); lr2 = lr_of(r0 + 2u); lr3 = lr_of(r0 + 3u);
                    }
                    out.x = adam_elem(p[u].x, g[u].x, m[u].x, v[u].x, lr0, L);
                    out.y = adam_elem(p[u].y, g[u].y, m[u].y, v[u].y, lr1, L);
                    out.z = adam_elem(p[u].z, g[u].z, m[u].z, v[u].z, lr2, L);
                    out.w = adam_elem(p[u].w, g[u].w, m[u].w, v[u].w, lr3, L);
                    *reinterpret_cast<float4*>(L.m + poff) = m[u];
                    *reinterpret_cast<float4*>(L.v + poff) = v[u];
                } else {
                    out = g[u];  // plain-sum segment (refine statistics)
                }
                if (MM_ST) {
                    multimem_st_v4(L.dst_mc + poff, out);
                } else {
#pragma unroll 8
                    for (int r = 0; r < world; r++) st_peer_v4(L.dst[r] + poff, out);
                }
            }
        }
    }
    // loss: every rank sums the ranks' scalars itself (same order => same value everywhere)
    if (blockIdx.x == 0 && threadIdx.x == 0 && L.loss_out) {
        float acc = 0.f;
        for (int r = 0; r < world; r++) {
            float t;
            asm volatile("ld.global.relaxed.sys.f32 %0, [%1];" : "=f"(t) : "l"(L.src[r] + L.loss_off));
            acc += t;
        }
        *L.loss_out = acc * L.grad_scale;
    }
    // ---- barrier 1: this rank has finished reading every peer's gradients and writing every peer's parameters.
    // The last CTA to arrive publishes that, then holds the kernel open until all peers have said the same, so the
    // next kernel on this stream may overwrite the gradients and read the parameters. ----
    __syncthreads();
    __shared__ bool s_last;
    if (threadIdx.x == 0) {
        __threadfence_system();
        s_last = atomicAdd(L.arrive, 1u) == gridDim.x - 1u;
    }
    __syncthreads();
    if (!s_last) return;
    if (threadIdx.x == 0) *L.arrive = 0u;
    if (threadIdx.x < (uint32_t)world) {
        __threadfence_system();
        st_release_sys(L.flags[threadIdx.x] + lane_off + 1 * DP_MAX_WORLD + rank, L.epoch);
    }
    wait_flags(my_flags, 1, world, L.epoch);
}

void launch_dp_adam(const DpLaunch& L, cudaStream_t st) {
    g_launch_count++;
    int sms = 148;
    {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    }
    int blocks = sms * (L.ctas_per_sm > 0 ? L.ctas_per_sm : 6);
    // BRUSH_DP_SIDE_BLOCKS: grid of the side-lane (overlapped SH) launch, for A/B measurements of its SM footprint
    static const int env_side_blocks = [] {
        const char* e = getenv("BRUSH_DP_SIDE_BLOCKS");
        return e ? atoi(e) : 0;
    }();
    if (L.lane == 1 && env_side_blocks > 0) blocks = env_side_blocks;
    const bool mm_ld = L.src_mc != nullptr, mm_st = L.dst_mc != nullptr;
    // float4s in flight per thread.  BRUSH_DP_U=2|4|8 overrides (A/B measurements).
    static const int env_u = [] {
        const char* e = getenv("BRUSH_DP_U");
        return e ? atoi(e) : 0;
    }();
    const int u = env_u ? env_u : 2;
#define DP_LAUNCH(LD, ST)                                                             \
    do {                                                                              \
        if (u >= 8) dp_adam_kernel<LD, ST, 8><<<blocks, 256, 0, st>>>(L);             \
        else if (u >= 4) dp_adam_kernel<LD, ST, 4><<<blocks, 256, 0, st>>>(L);        \
        else dp_adam_kernel<LD, ST, 2><<<blocks, 256, 0, st>>>(L);                    \
    } while (0)
    if (mm_ld && mm_st) DP_LAUNCH(true, true);
    else if (mm_ld) DP_LAUNCH(true, false);
    else if (mm_st) DP_LAUNCH(false, true);
    else DP_LAUNCH(false, false);
#undef DP_LAUNCH
}

}  // namespace brush
