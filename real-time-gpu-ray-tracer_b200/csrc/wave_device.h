// wave_device.h — small device helpers shared by the wavefront kernels (kernels_wave.cu) and the persistent path
// kernel (kernels_persist.cu): tile-major pixel ids, shard row mapping, warp-aggregated atomics.
#pragma once

#include "rt_types.h"

#include <cstdint>

constexpr uint32_t SLOT_IDLE = 0xFFFFFFFFu;
constexpr uint32_t SLOT_FRESH = 0xFEFEFEFEu;     // byte pattern of the frame-start memset: the slot has no pixel yet

static __device__ __forceinline__ uint32_t lane_id() { return threadIdx.x & 31u; }

// tile-major pixel id -> local pixel coordinates (8x4 tiles, so that a warp's 32 consecutive ids form one tile)
static __device__ __forceinline__ void pixel_coords(uint32_t pid, uint32_t tiles_x, uint32_t &lx, uint32_t &ly) {
    const uint32_t tile = pid >> 5, in = pid & 31u;
    lx = (tile % tiles_x) * 8u + (in & 7u);
    ly = (tile / tiles_x) * 4u + (in >> 3);
}
static __device__ __forceinline__ uint32_t global_row(const FrameD &fr, uint32_t ly) {
    return ((ly / fr.tile_rows) * fr.shard_world + fr.shard_rank) * fr.tile_rows + (ly % fr.tile_rows);
}

// warp-aggregated append: every lane with `want` gets a distinct index from *counter (one atomic per warp)
static __device__ __forceinline__ uint32_t warp_append(uint32_t *counter, bool want) {
    const uint32_t mask = __ballot_sync(0xffffffffu, want);
    if (mask == 0u) return 0u;
    const uint32_t leader = __ffs(mask) - 1u;
    uint32_t base = 0;
    if (lane_id() == leader) base = atomicAdd(counter, (uint32_t) __popc(mask));
    base = __shfl_sync(0xffffffffu, base, leader);
    return base + (uint32_t) __popc(mask & ((1u << lane_id()) - 1u));
}
