// scan.cuh -- warp scan + single-pass "decoupled look-back" prefix across tiles, used by the stable
// live-ray compaction (stand-alone op and inside the persistent renderer).
#pragma once
#include <stdint.h>

namespace ngp {

__device__ __forceinline__ uint32_t warp_inclusive_sum(uint32_t v, int lane) {
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t o = __shfl_up_sync(0xffffffffu, v, d);
        if (lane >= d) v += o;
    }
    return v;
}
__device__ __forceinline__ uint32_t warp_sum(uint32_t v) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
    return v;
}

// status word: [63:34] tag  [33:32] flag (0 = nothing, 1 = tile aggregate, 2 = inclusive prefix)  [31:0] value.
// The tag makes stale words from an earlier pass over the same array read as "nothing".
constexpr uint32_t kFlagAggregate = 1u, kFlagInclusive = 2u;

__device__ __forceinline__ unsigned long long pack_status(uint32_t tag, uint32_t flag, uint32_t value) {
    return ((unsigned long long)tag << 34) | ((unsigned long long)flag << 32) | value;
}
__device__ __forceinline__ unsigned long long ld_acquire_u64(const unsigned long long *p) {
    unsigned long long v;
    asm volatile("ld.global.acquire.gpu.b64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_u64(unsigned long long *p, unsigned long long v) {
    asm volatile("st.global.release.gpu.b64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// Called by ONE full warp of the tile.  Publishes this tile's aggregate, walks back over predecessors
// (32 at a time) until an inclusive prefix is found, publishes this tile's inclusive prefix and returns the
// exclusive prefix (sum of aggregates of tiles < tile) on every lane.
// Requires: tiles are numbered by an atomic ticket, so all predecessors have started.
__device__ __forceinline__ uint32_t lookback_exclusive_prefix(unsigned long long *status, uint32_t tile, uint32_t aggregate,
                                                              uint32_t tag, int lane) {
    if (tile == 0) {
        if (lane == 0) st_release_u64(status, pack_status(tag, kFlagInclusive, aggregate));
        return 0;
    }
    if (lane == 0) st_release_u64(status + tile, pack_status(tag, kFlagAggregate, aggregate));
    uint32_t running = 0;
    int look = (int)tile - 1;
    for (;;) {
        const int idx = look - lane;
        unsigned long long w;
        uint32_t flag;
        do {
            w = idx >= 0 ? ld_acquire_u64(status + idx) : pack_status(tag, kFlagInclusive, 0);
            flag = ((uint32_t)(w >> 34) == tag) ? (uint32_t)(w >> 32) & 3u : 0u;
        } while (__any_sync(0xffffffffu, flag == 0));
        const uint32_t incl_mask = __ballot_sync(0xffffffffu, flag == kFlagInclusive);
        if (incl_mask) {
            const int nearest = __ffs(incl_mask) - 1;  // lowest lane = closest predecessor holding a full prefix
            running += warp_sum(lane <= nearest ? (uint32_t)w : 0u);
            break;
        }
        running += warp_sum((uint32_t)w);
        look -= 32;
    }
    if (lane == 0) st_release_u64(status + tile, pack_status(tag, kFlagInclusive, running + aggregate));
    return running;
}

}  // namespace ngp
