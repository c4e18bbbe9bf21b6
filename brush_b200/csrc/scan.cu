// Inclusive u32 prefix sum == brush_prefix_sum::prefix_sum (brush-prefix-sum/src/lib.rs:17-102, which runs
// 2*levels+1 launches of 512-thread Hillis-Steele blocks).  Here: ONE launch, one read and one write of
// the data (8 B/element, the algorithmic minimum), chained scan with decoupled look-back.
#include <cstdio>

#include "internal.h"
#include "lookback.cuh"

namespace brush {

constexpr int SC_THREADS = 256;
constexpr int SC_VEC = 4;                  // uint4 loads
constexpr int SC_CHUNKS = 4;               // uint4s per thread: 16 elements, 4096 per tile
constexpr int SC_ITEMS = SC_VEC * SC_CHUNKS;
constexpr int SC_TILE = SC_THREADS * SC_ITEMS;
// Tile size matters more than it looks: at 1 M elements every tile of a 1024-element tiling is resident at once, no
// predecessor is "inclusive" yet, and the last tile's look-back is 977 / 32 serial rounds of L2 round trips (15 us
// for 8 MB of traffic).  4096-element tiles: 245 tiles, 8 rounds.

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
    // warp-striped: chunk j of lane l of warp w = uint4 number (w * SC_CHUNKS + j) * 32 + l of the tile, so every
    // load and store instruction of a warp covers 512 contiguous bytes
    const uint32_t tile_base = tile * SC_TILE;
    uint32_t v[SC_CHUNKS][SC_VEC];
    const bool aligned = ((reinterpret_cast<uintptr_t>(in) | reinterpret_cast<uintptr_t>(out)) & 15u) == 0;
    const bool full = aligned && (uint64_t)tile_base + SC_TILE <= n;
#pragma unroll
    for (int j = 0; j < SC_CHUNKS; j++) {
        const uint32_t e = tile_base + ((warp * SC_CHUNKS + j) * 32u + lane) * SC_VEC;
        if (full) {
            const uint4 q = __ldg(reinterpret_cast<const uint4*>(in + e));
            v[j][0] = q.x; v[j][1] = q.y; v[j][2] = q.z; v[j][3] = q.w;
        } else {
#pragma unroll
            for (int i = 0; i < SC_VEC; i++) v[j][i] = e + i < n ? __ldg(in + e + i) : 0u;
        }
    }
    // per chunk: serial scan of the four elements, warp scan of the chunk totals; chunks of a warp are consecutive
    uint32_t chunk_off[SC_CHUNKS];
    uint32_t warp_running = 0;
#pragma unroll
    for (int j = 0; j < SC_CHUNKS; j++) {
        v[j][1] += v[j][0]; v[j][2] += v[j][1]; v[j][3] += v[j][2];
        uint32_t incl = v[j][3];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
            if (lane >= (uint32_t)o) incl += t;
        }
        chunk_off[j] = warp_running + incl - v[j][3];
        warp_running += __shfl_sync(0xffffffffu, incl, 31);
    }
    if (lane == 31) s_warp[warp] = warp_running;
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
    const uint32_t base_off = s_base + s_warp[warp];
#pragma unroll
    for (int j = 0; j < SC_CHUNKS; j++) {
        const uint32_t e = tile_base + ((warp * SC_CHUNKS + j) * 32u + lane) * SC_VEC;
        const uint32_t off = base_off + chunk_off[j];
        if (full) {
            *reinterpret_cast<uint4*>(out + e) = make_uint4(v[j][0] + off, v[j][1] + off, v[j][2] + off, v[j][3] + off);
        } else {
#pragma unroll
            for (int i = 0; i < SC_VEC; i++)
                if (e + i < n) out[e + i] = v[j][i] + off;
        }
    }
}

void launch_inclusive_scan(const uint32_t* in, uint32_t n, uint32_t* out, uint64_t* lb_status, uint32_t* ticket,
                           int32_t* total_out, uint64_t cap, uint32_t* status_word, cudaStream_t st) {
    if (n == 0) return;
    uint32_t tiles = (n + SC_TILE - 1) / SC_TILE;
    g_launch_count++;
    launch_k(inclusive_scan_kernel, tiles, SC_THREADS, 0, st, in, n, out, lb_status, ticket, total_out, cap, status_word);
}

// ---- the allocation-free entry point (brush_prefix_sum_u32_noalloc, include/brush_b200.h) ----
// Same kernel, but the look-back status words and the tile ticket live in a file-scope device array instead of the
// ctx's cudaMalloc'ed control block: callers under CUDA graph capture or in harnesses that forbid cudaMalloc get a
// scan that allocates nothing.  Every GPU that loads the library holds the array (32 KB + 64 B) until exit.
// Layout: word 0 = ticket (low 32 bits), words 8.. = one status word per tile (+1 spare, as scan_status_words).
constexpr uint32_t NOALLOC_MAX_N = BRUSH_NOALLOC_SCAN_MAX_N;
constexpr uint32_t NOALLOC_CTRL_WORDS = 8;
constexpr uint32_t NOALLOC_WORDS = NOALLOC_CTRL_WORDS + (NOALLOC_MAX_N + SC_TILE - 1) / SC_TILE + 1;
static __device__ __align__(256) unsigned long long g_noalloc_scan_scratch[NOALLOC_WORDS];

// The array keeps the previous call's tickets and status words: clear the part this call reads, on the stream.
__global__ void __launch_bounds__(256) noalloc_scan_clear_kernel(unsigned long long* scratch, uint32_t words) {
    PDL_ENTER();
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < words; i += gridDim.x * blockDim.x) scratch[i] = 0ull;
}

// In host code `g_noalloc_scan_scratch` is only a handle for the runtime, never a device address: the address is
// looked up once per device, with that device current, and cached.
constexpr int NOALLOC_MAX_DEVICES = 64;
static std::atomic<unsigned long long*> g_noalloc_ptr[NOALLOC_MAX_DEVICES];

static cudaError_t noalloc_scratch_address(unsigned long long** out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= NOALLOC_MAX_DEVICES) return cudaErrorInvalidDevice;
    unsigned long long* p = g_noalloc_ptr[dev].load(std::memory_order_acquire);
    if (!p) {
        e = cudaGetSymbolAddress((void**)&p, g_noalloc_scan_scratch);
        if (e != cudaSuccess) return e;
        g_noalloc_ptr[dev].store(p, std::memory_order_release);  // racing threads store the same address
    }
    *out = p;
    return cudaSuccess;
}

}  // namespace brush

static int noalloc_fail(char* err_host, uint32_t err_len, int code, const char* msg, const char* detail = "") {
    if (err_host && err_len) snprintf(err_host, err_len, "%s%s", msg, detail);
    return code;
}

extern "C" __attribute__((visibility("default"))) int brush_prefix_sum_u32_noalloc(
    const uint32_t* in, uint32_t n, uint32_t* out, void* stream, char* err_host, uint32_t err_len) {
    using namespace brush;
    if (err_host && err_len) err_host[0] = 0;
    if (n && (!in || !out)) return noalloc_fail(err_host, err_len, BRUSH_E_INVALID, "brush_prefix_sum_u32_noalloc: null argument");
    if (n > NOALLOC_MAX_N) {
        char detail[96];
        snprintf(detail, sizeof detail, "n = %u exceeds BRUSH_NOALLOC_SCAN_MAX_N = %u (static scratch)", n, NOALLOC_MAX_N);
        return noalloc_fail(err_host, err_len, BRUSH_E_INVALID, "brush_prefix_sum_u32_noalloc: ", detail);
    }
    if (n == 0) return BRUSH_OK;
    unsigned long long* scratch = nullptr;
    cudaError_t e = noalloc_scratch_address(&scratch);
    if (e != cudaSuccess)
        return noalloc_fail(err_host, err_len, BRUSH_E_CUDA, "brush_prefix_sum_u32_noalloc: cudaGetSymbolAddress: ",
                            cudaGetErrorString(e));
    cudaStream_t st = (cudaStream_t)stream;
    const uint32_t words = NOALLOC_CTRL_WORDS + (uint32_t)scan_status_words(n);
    g_launch_count++;
    e = launch_k(noalloc_scan_clear_kernel, (words + 255) / 256, 256, 0, st, scratch, words);
    if (e == cudaSuccess) {
        launch_inclusive_scan(in, n, out, reinterpret_cast<uint64_t*>(scratch + NOALLOC_CTRL_WORDS),
                              reinterpret_cast<uint32_t*>(scratch), nullptr, ~0ull, nullptr, st);
        e = cudaGetLastError();
    }
    if (e != cudaSuccess)
        return noalloc_fail(err_host, err_len, BRUSH_E_CUDA, "brush_prefix_sum_u32_noalloc: launch: ", cudaGetErrorString(e));
    return BRUSH_OK;
}
