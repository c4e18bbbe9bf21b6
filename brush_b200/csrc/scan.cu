// Inclusive u32 prefix sum == brush_prefix_sum::prefix_sum (brush-prefix-sum/src/lib.rs:17-102, which runs
// 2*levels+1 launches of 512-thread Hillis-Steele blocks).  Here: ONE launch, one read and one write of
// the data (8 B/element, the algorithmic minimum), chained scan with decoupled look-back.
#include "internal.h"
#include "lookback.cuh"

namespace brush {

constexpr int SC_THREADS = 256;
constexpr int SC_ITEMS = 4;
constexpr int SC_TILE = SC_THREADS * SC_ITEMS;

size_t scan_status_words(uint32_t n) { return (size_t)(n + SC_TILE - 1) / SC_TILE + 1; }

__global__ void __launch_bounds__(SC_THREADS)
inclusive_scan_kernel(const uint32_t* __restrict__ in, uint32_t n, uint32_t* __restrict__ out, uint64_t* lb_status,
                      uint32_t* ticket, int32_t* total_out, uint64_t cap, uint32_t* status_word) {
    __shared__ uint32_t s_warp[SC_THREADS / 32];
    __shared__ uint32_t s_tile, s_base;
    const uint32_t tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    PDL_ENTER();
    if (tid == 0) s_tile = atomicAdd(ticket, 1u);
    __syncthreads();
    const uint32_t tile = s_tile;
    const uint32_t base = tile * SC_TILE + tid * SC_ITEMS;
    uint32_t v[SC_ITEMS];
    const bool aligned = ((reinterpret_cast<uintptr_t>(in) | reinterpret_cast<uintptr_t>(out)) & 15u) == 0;
    if (aligned && base + SC_ITEMS <= n) {
        uint4 q = __ldg(reinterpret_cast<const uint4*>(in + base));
        v[0] = q.x; v[1] = q.y; v[2] = q.z; v[3] = q.w;
    } else {
#pragma unroll
        for (int i = 0; i < SC_ITEMS; i++) v[i] = base + i < n ? __ldg(in + base + i) : 0u;
    }
    v[1] += v[0]; v[2] += v[1]; v[3] += v[2];
    uint32_t incl = v[3];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= (uint32_t)o) incl += t;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        uint32_t w = lane < SC_THREADS / 32 ? s_warp[lane] : 0u;
        uint32_t wi = w;
#pragma unroll
        for (int o = 1; o < SC_THREADS / 32; o <<= 1) {
            uint32_t t = __shfl_up_sync(0xffffffffu, wi, o);
            if (lane >= (uint32_t)o) wi += t;
        }
        uint32_t block_total = __shfl_sync(0xffffffffu, wi, SC_THREADS / 32 - 1);
        if (lane < SC_THREADS / 32) s_warp[lane] = wi - w;
        uint32_t excl = lookback_exclusive(lb_status, tile, block_total);
        if (lane == 0) {
            s_base = excl;
            if ((uint64_t)(tile + 1) * SC_TILE >= n && total_out) {
                uint64_t total = (uint64_t)excl + block_total;
                if (total > cap) {
                    if (status_word) atomicOr(status_word, 1u);
                    total = cap;
                }
                *total_out = (int32_t)total;
            }
        }
    }
    __syncthreads();
    const uint32_t off = s_base + s_warp[warp] + (incl - v[3]);
    if (aligned && base + SC_ITEMS <= n) {
        *reinterpret_cast<uint4*>(out + base) = make_uint4(v[0] + off, v[1] + off, v[2] + off, v[3] + off);
    } else {
#pragma unroll
        for (int i = 0; i < SC_ITEMS; i++)
            if (base + i < n) out[base + i] = v[i] + off;
    }
}

void launch_inclusive_scan(const uint32_t* in, uint32_t n, uint32_t* out, uint64_t* lb_status, uint32_t* ticket,
                           int32_t* total_out, uint64_t cap, uint32_t* status_word, cudaStream_t st) {
    if (n == 0) return;
    uint32_t tiles = (n + SC_TILE - 1) / SC_TILE;
    g_launch_count++;
    launch_k(inclusive_scan_kernel, tiles, SC_THREADS, 0, st, in, n, out, lb_status, ticket, total_out, cap, status_word);
}

}  // namespace brush
