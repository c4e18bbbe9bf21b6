// Single-pass chained scan with decoupled look-back (Merrill & Garland) -- the B200 replacement for
// the reference's multi-level prefix sum (brush-prefix-sum/src/lib.rs:17-102) and for the atomic
// compaction counter of project_forward.wgsl:65 (made deterministic: ascending gid order).
//
// One 64-bit status word per tile: [63:32] = flag (0 empty, 1 aggregate, 2 inclusive), [31:0] = value.
// Flag and value travel in one word, so a single volatile 64-bit store/load is the whole protocol.
// Status words must be zero before the launch; tiles take their index from an atomic ticket so a
// tile never waits on one that has not started.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace brush {

#define LB_AGG (1ull << 32)
#define LB_INC (2ull << 32)

__device__ __forceinline__ uint32_t warp_sum_u32(uint32_t v) {
    return __reduce_add_sync(0xffffffffu, v);
}

// Called by all 32 lanes of ONE warp of the tile.  Publishes `aggregate`, walks back over the
// predecessors 32 at a time, publishes the inclusive prefix and returns the exclusive prefix.
__device__ __forceinline__ uint32_t lookback_exclusive(uint64_t* status, uint32_t tile, uint32_t aggregate) {
    const uint32_t lane = threadIdx.x & 31u;
    volatile uint64_t* vs = status;
    if (tile == 0) {
        if (lane == 0) vs[0] = LB_INC | aggregate;
        return 0u;
    }
    if (lane == 0) vs[tile] = LB_AGG | aggregate;
    uint32_t exclusive = 0;
    int look = (int)tile - 1;
    while (true) {
        int idx = look - (int)lane;
        uint64_t s;
        do {
            s = idx >= 0 ? vs[idx] : LB_INC;
        } while (__any_sync(0xffffffffu, (s >> 32) == 0ull));
        uint32_t inc_mask = __ballot_sync(0xffffffffu, (s >> 32) == 2ull);
        uint32_t val = (uint32_t)s;
        if (inc_mask) {
            uint32_t first = __ffs(inc_mask) - 1u;
            exclusive += warp_sum_u32(lane <= first ? val : 0u);
            break;
        }
        exclusive += warp_sum_u32(val);
        look -= 32;
    }
    if (lane == 0) vs[tile] = LB_INC | (uint64_t)(exclusive + aggregate);
    return exclusive;
}

}  // namespace brush
