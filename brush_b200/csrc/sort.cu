// Onesweep least-significant-digit radix sort (Adinets & Merrill 2022), key-value, stable, 8-bit digits.
// == brush_sort::radix_argsort (brush-sort/src/lib.rs:32-147), which runs 5 kernels per 4-bit pass
// (count / reduce / scan / scan_add / scatter: 40 launches for 32 bits).  Here: one histogram launch
// for all passes + ONE launch per 8-bit pass; each pass reads and writes every key/value once and
// resolves the cross-tile digit offsets with decoupled look-back instead of a global scan.
//
// The element count lives on the device (*n_dev), as in the reference (indirect dispatch): the grid
// is sized for max_n and tiles past *n_dev exit at once.
// Sort order contract: stable, ascending on the low 4*ceil(bits/4) key bits (exactly the bits the
// reference's ceil(bits/4) passes look at).
#include <cstdlib>

#include "internal.h"

namespace brush {

constexpr int RS_THREADS = 256;
constexpr int RS_WARPS = RS_THREADS / 32;
constexpr int RS_RADIX = 256;
constexpr int RS_MAX_PASSES = 8;
constexpr int LB_WINDOW = 16;

// Lanes holding the same 8-bit digit, by eight ballots: fixed latency, unlike match.any whose cost grows with the
// number of distinct values in the warp (random digits: ~28 distinct of 32 -- measured 22 us per 1M-key pass with
// match.any).  Lanes with valid == false match nothing.
__device__ __forceinline__ uint32_t match_digit(uint32_t d, bool valid) {
    uint32_t peers = __ballot_sync(0xffffffffu, valid);
#pragma unroll
    for (int b = 0; b < 8; b++) {
        const bool bit = (d >> b) & 1u;
        const uint32_t bal = __ballot_sync(0xffffffffu, bit);
        peers &= bit ? bal : ~bal;
    }
    return peers;
}

#ifdef BRUSH_SORT_PROF  // debug build only (scratch/): phase clocks of one tile
__device__ unsigned long long g_sort_prof[16];
#define PROF_MARK(i) do { if (tid == 0 && tile == gridDim.x / 8) g_sort_prof[i] = clock64(); } while (0)
#else
#define PROF_MARK(i) do { } while (0)
#endif

// Keys per thread.  BRUSH_SORT_ITEMS=8|16 overrides the 32-bit-key default (A/B: 8 halves the registers and doubles
// the resident CTAs, at twice the look-back work per key).
template <typename KeyT> struct SortCfg;
template <> struct SortCfg<uint32_t> { static constexpr int ITEMS = 16; };
template <> struct SortCfg<uint64_t> { static constexpr int ITEMS = 8; };
static int items_u32() {
    static const int v = [] {
        const char* e = getenv("BRUSH_SORT_ITEMS");
        const int x = e ? atoi(e) : 0;
        return (x == 8 || x == 16) ? x : SortCfg<uint32_t>::ITEMS;
    }();
    return v;
}

size_t sort_tiles(uint32_t max_n, int key_bytes) {
    uint32_t tile = RS_THREADS * (key_bytes == 8 ? SortCfg<uint64_t>::ITEMS : items_u32());
    return ((size_t)max_n + tile - 1) / tile;
}

struct PassPlan {
    uint32_t passes;
    uint32_t shift[RS_MAX_PASSES];
    uint32_t mask[RS_MAX_PASSES];
};

static PassPlan make_plan(uint32_t bits) {
    PassPlan p{};
    uint32_t rb = ((bits + 3) / 4) * 4;  // bits the reference's 4-bit passes cover
    p.passes = (rb + 7) / 8;
    for (uint32_t i = 0; i < p.passes; i++) {
        uint32_t nb = rb - 8 * i < 8 ? rb - 8 * i : 8;
        p.shift[i] = 8 * i;
        p.mask[i] = (1u << nb) - 1u;
    }
    return p;
}

// ---- histogram of every pass's digit, one read of the keys ----
template <typename KeyT>
__global__ void __launch_bounds__(RS_THREADS)
radix_hist_kernel(const KeyT* __restrict__ keys, const int32_t* __restrict__ n_dev, uint32_t max_n,
                  const __grid_constant__ PassPlan plan, uint32_t* __restrict__ hist) {
    __shared__ uint32_t s_hist[RS_MAX_PASSES * RS_RADIX];
    for (uint32_t i = threadIdx.x; i < plan.passes * RS_RADIX; i += RS_THREADS) s_hist[i] = 0;
    PDL_ENTER();
    const uint32_t n = min((uint32_t)max(*n_dev, 0), max_n);
    __syncthreads();
    const uint32_t lane = threadIdx.x & 31u;
    // warp-uniform loop bound so match_any always sees the whole warp
    // four independent key loads per thread per iteration (the loop is otherwise one L2/HBM round trip per key)
    const uint32_t stride = gridDim.x * RS_THREADS * 4;
    for (uint32_t base = blockIdx.x * RS_THREADS * 4; base < n; base += stride) {
        KeyT k4[4];
        bool valid4[4];
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const uint32_t i = base + u * RS_THREADS + threadIdx.x;
            valid4[u] = i < n;
            k4[u] = valid4[u] ? keys[i] : (KeyT)0;
        }
#pragma unroll
        for (int u = 0; u < 4; u++) {
            const bool valid = valid4[u];
            const KeyT k = k4[u];
            for (uint32_t p = 0; p < plan.passes; p++) {
                const uint32_t d = (uint32_t)(k >> plan.shift[p]) & plan.mask[p];
                // Clustered digits (tile-row bits, depth exponents) would serialise up to 32-way on one shared
                // word, so they are counted warp-aggregated (nine ballots); scattered digits (mantissa bytes, tile
                // columns) go straight to a shared atomic.  The warp picks per digit by looking at how many lanes
                // agree with their neighbour.
                const uint32_t nb = __shfl_down_sync(0xffffffffu, d, 1);
                const uint32_t same = __ballot_sync(0xffffffffu, valid && nb == d);
                if (__popc(same) >= 8) {
                    const uint32_t peers = match_digit(d, valid);
                    if (valid && lane == (uint32_t)(31 - __clz(peers)))
                        atomicAdd(&s_hist[p * RS_RADIX + d], __popc(peers));
                } else if (valid) {
                    atomicAdd(&s_hist[p * RS_RADIX + d], 1u);
                }
            }
        }
    }
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < plan.passes * RS_RADIX; i += RS_THREADS)
        if (s_hist[i]) atomicAdd(&hist[i], s_hist[i]);
}

// ---- one onesweep pass ----
// Two kernels with ONE body (onesweep_body.inc) that differ only in their launch bounds:
//   onesweep_kernel      128 registers, two CTAs per SM -- fastest while a pass is a single wave of tiles (its time is
//                        then ONE tile's latency: the 1 M-key depth sort);
//   onesweep_kernel_occ3 80 registers (no spills), three CTAs per SM -- faster as soon as a pass is many waves, where the
//                        serial phases of a tile overlap only across CTAs (76.7 M keys: 1.94 -> 1.59 ms; 64 registers /
//                        four CTAs spill and gain nothing more).
// (A shared __device__ function changed the register allocation of the first: the body is textually included.)
template <typename KeyT, int ITEMS>
__global__ void __launch_bounds__(RS_THREADS)
onesweep_kernel(const KeyT* __restrict__ keys_in, const uint32_t* __restrict__ vals_in, KeyT* __restrict__ keys_out,
                uint32_t* __restrict__ vals_out, const int32_t* __restrict__ n_dev, uint32_t max_n, uint32_t shift,
                uint32_t mask, const uint32_t* __restrict__ hist, uint32_t* lb, uint32_t* ticket) {
#include "onesweep_body.inc"
}

template <typename KeyT, int ITEMS>
__global__ void __launch_bounds__(RS_THREADS, 3)
onesweep_kernel_occ3(const KeyT* __restrict__ keys_in, const uint32_t* __restrict__ vals_in, KeyT* __restrict__ keys_out,
                uint32_t* __restrict__ vals_out, const int32_t* __restrict__ n_dev, uint32_t max_n, uint32_t shift,
                uint32_t mask, const uint32_t* __restrict__ hist, uint32_t* lb, uint32_t* ticket) {
#include "onesweep_body.inc"
}

template <typename KeyT>
__global__ void copy_first_n_kernel(const KeyT* __restrict__ k, const uint32_t* __restrict__ v,
                                    const int32_t* __restrict__ n_dev, uint32_t max_n, KeyT* __restrict__ ko,
                                    uint32_t* __restrict__ vo) {
    const uint32_t n = min((uint32_t)max(*n_dev, 0), max_n);
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        ko[i] = k[i];
        vo[i] = v[i];
    }
}

template <typename KeyT>
static void radix_sort_impl(const KeyT* keys, const uint32_t* vals, const int32_t* n_dev, uint32_t max_n,
                            uint32_t bits, KeyT* keys_out, uint32_t* vals_out, const SortScratch& s,
                            cudaStream_t st) {
    if (max_n == 0) return;
    const PassPlan plan = make_plan(bits);
    if (plan.passes == 0) {
        g_launch_count++;
        copy_first_n_kernel<KeyT><<<296, 256, 0, st>>>(keys, vals, n_dev, max_n, keys_out, vals_out);
        return;
    }
    const uint32_t tiles = (uint32_t)sort_tiles(max_n, sizeof(KeyT));
    const int items = sizeof(KeyT) == 8 ? SortCfg<uint64_t>::ITEMS : items_u32();
    if (!s.pre_zeroed) {
        ZeroSegs z{};
        z.p[0] = s.hist; z.bytes[0] = (size_t)plan.passes * RS_RADIX * 4;
        z.p[1] = s.lb; z.bytes[1] = (size_t)plan.passes * tiles * RS_RADIX * 4;
        z.p[2] = s.tickets; z.bytes[2] = RS_MAX_PASSES * 4;
        zero_multi_async(z, st);
    }
    uint32_t hist_blocks = (max_n + RS_THREADS * 4 - 1) / (RS_THREADS * 4);
    static const uint32_t hist_cap = [] {
        const char* e = getenv("BRUSH_HIST_BLOCKS");
        const int x = e ? atoi(e) : 0;
        return x > 0 ? (uint32_t)x : 148u * 4u;
    }();
    if (hist_blocks > hist_cap) hist_blocks = hist_cap;  // every block ends with up to passes*256 global atomics
    g_launch_count += 1 + plan.passes;
    launch_k(radix_hist_kernel<KeyT>, hist_blocks, RS_THREADS, 0, st, keys, n_dev, max_n, plan, s.hist);
    const KeyT* src_k = keys;
    const uint32_t* src_v = vals;
    KeyT* tmp_k = reinterpret_cast<KeyT*>(s.keys_tmp);
    for (uint32_t p = 0; p < plan.passes; p++) {
        const bool to_out = ((plan.passes - 1 - p) % 2) == 0;
        KeyT* dst_k = to_out ? keys_out : tmp_k;
        uint32_t* dst_v = to_out ? vals_out : s.vals_tmp;
        if (items == 8 || sizeof(KeyT) == 8)
            launch_k(onesweep_kernel<KeyT, 8>, tiles, RS_THREADS, 0, st, src_k, src_v, dst_k, dst_v, n_dev, max_n,
                     plan.shift[p], plan.mask[p], (const uint32_t*)(s.hist + p * RS_RADIX),
                     s.lb + (size_t)p * tiles * RS_RADIX, s.tickets + p);
        else if constexpr (sizeof(KeyT) == 4) {
            // more tiles than two CTAs per SM hold at once?  The buffer's size decides: the count lives on the device.
            if (tiles > 2u * 148u)
                launch_k(onesweep_kernel_occ3<KeyT, 16>, tiles, RS_THREADS, 0, st, src_k, src_v, dst_k, dst_v, n_dev, max_n,
                         plan.shift[p], plan.mask[p], (const uint32_t*)(s.hist + p * RS_RADIX),
                         s.lb + (size_t)p * tiles * RS_RADIX, s.tickets + p);
            else
                launch_k(onesweep_kernel<KeyT, 16>, tiles, RS_THREADS, 0, st, src_k, src_v, dst_k, dst_v, n_dev, max_n,
                         plan.shift[p], plan.mask[p], (const uint32_t*)(s.hist + p * RS_RADIX),
                         s.lb + (size_t)p * tiles * RS_RADIX, s.tickets + p);
        }
        src_k = dst_k;
        src_v = dst_v;
    }
}

#ifdef BRUSH_SORT_PROF
extern "C" __attribute__((visibility("default"))) void brush_debug_sort_prof(unsigned long long* out) {
    cudaMemcpyFromSymbol(out, g_sort_prof, sizeof(unsigned long long) * 16);
}
#endif

void launch_radix_sort_u32(const uint32_t* keys, const uint32_t* vals, const int32_t* n_dev, uint32_t max_n,
                           uint32_t bits, uint32_t* keys_out, uint32_t* vals_out, const SortScratch& s,
                           cudaStream_t st) {
    radix_sort_impl<uint32_t>(keys, vals, n_dev, max_n, bits, keys_out, vals_out, s, st);
}
void launch_radix_sort_u64(const uint64_t* keys, const uint32_t* vals, const int32_t* n_dev, uint32_t max_n,
                           uint32_t bits, uint64_t* keys_out, uint32_t* vals_out, const SortScratch& s,
                           cudaStream_t st) {
    radix_sort_impl<uint64_t>(keys, vals, n_dev, max_n, bits, keys_out, vals_out, s, st);
}

}  // namespace brush
