// wave.h — host side of the wavefront path tracer (kernels_wave.cu): slot-streamed path state, ray queue, and
// the wave loop.  Replaces the reference's one-thread-per-pixel megakernel `render` (src/Global/Kernel.cu:105-147).
#pragma once

#include "build.h"
#include "device_util.h"
#include "rt_types.h"

#include <cstdint>
#include <vector>

// Device-resident state of the paths in flight.  A SLOT carries one pixel at a time through all of its samples
// and bounces (the reference's per-pixel XORWOW stream is consumed sequentially, so a pixel is the unit of
// sequential work); when the pixel is finished the slot takes the next unstarted pixel.
struct WaveBuffers {
    float4 *ray_o;        // origin.xyz
    float4 *ray_d;        // direction.xyz (not normalised after a Rough scatter, like the reference)
    float4 *hit;          // t, u, v, bits(leaf ref)
    int *hit_inst;        // instance index, -1 = miss
    float4 *thr;          // path throughput.xyz, w = bits(sample index << 8 | depth)
    float4 *acc;          // radiance sum of the finished samples.xyz, w = bits(pixel id in tile-major order), 0xFFFFFFFF = idle slot
    uint4 *rng_a;         // XORWOW d, v0, v1, v2
    uint2 *rng_b;         // XORWOW v3, v4
    uint32_t *queue;      // slots with a ray to trace this wave
    uint32_t *counts;     // [w] rays queued for wave w of the current batch
    uint32_t *cursors;    // [w] fetch cursor of the trace kernel for wave w
    uint32_t *next_pixel; // next unstarted pixel (tile-major id)
    uint32_t n_slots;
    uint32_t n_pixels;    // group-local tile-major ids in [0, n_pixels): (tiles of this group) * 32
    uint32_t tiles_x;
    uint32_t group, n_groups;   // slot group: 8x4-pixel tile t of the shard belongs to group t % n_groups
};

struct WaveStats {
    uint64_t rays = 0;
    uint32_t waves = 0;
    uint32_t launches = 0;
};

class WaveRenderer {
public:
    // renders one frame of this context's shard into fr; returns after the last wave has been enqueued AND counted
    // (the wave loop reads a 4-byte counter back every `batch` waves to learn when all paths have ended)
    void render(const SceneD &sc, const CameraD &cam, const FrameD &fr, cudaStream_t stream, int sm_count, WaveStats &stats);
    uint32_t slots_override = 0;   // 0 = default
    int mode = 0;                  // 0: binary LBVH traversal, 1: compressed 8-wide
    int refill_min = 8;            // lanes of a warp that must be idle before the warp pulls new rays
    int node_min = 20;             // the node phase of a warp ends when fewer lanes than this stand on an inner node
    bool use_graphs = true;        // replay each group's wave batch as a CUDA graph
    int finish_divisor = 24;       // a group whose last wave had <= n_slots / finish_divisor rays is finished by wave_finish_kernel (0: never)
    int n_groups = 8;              // slot groups whose wave sequences run concurrently on their own streams
    int blocks_per_sm_override = 0; // > 0: cap the resident trace blocks per SM (tuning)
private:
    void reserve(uint32_t n_slots);
    DeviceBuffer<float4> ray_o_, ray_d_, hit_, thr_, acc_;
    DeviceBuffer<int> hit_inst_;
    DeviceBuffer<uint4> rng_a_;
    DeviceBuffer<uint2> rng_b_;
    DeviceBuffer<uint32_t> queue_, counters_;
    PinnedBuffer<uint32_t> h_counts_;
    static constexpr int MAX_GROUPS = 16;
    cudaStream_t g_stream_[MAX_GROUPS] = {};
    cudaEvent_t g_event_[MAX_GROUPS] = {};
    cudaEvent_t fork_event_ = nullptr;
    cudaGraphExec_t g_exec_[MAX_GROUPS] = {};     // the captured wave batch of each group ...
    std::vector<char> g_key_[MAX_GROUPS];         // ... and the kernel arguments it was captured with
public:
    ~WaveRenderer();
private:
    int trace_blocks_per_sm_[4] = {0, 0, 0, 0};
};

// ---- persistent path kernel (kernels_persist.cu) --------------------------------------------------------------
// Path state of the resident threads: one entry per THREAD of the persistent grid (thread id = slot), so every
// access is coalesced and the arrays (104 B x ~115 k threads) stay in L2.
struct PersistBuffers {
    float4 *ray_o, *ray_d;   // world ray of the thread's path segment in flight; ray_o.w = bits(rays of the pixel so far)
    float4 *thr;             // throughput.xyz, w = bits(sample << 16 | depth)
    float4 *acc;             // radiance sum of the finished samples.xyz, w = bits(pixel id, tile-major)
    uint4 *rng_a;            // XORWOW d, v0, v1, v2
    uint2 *rng_b;            // XORWOW v3, v4
    float4 *hit;             // split kernel only (kernels_split.cu): the finished ray's hit t, u, v, bits(leaf ref) ...
    int *hit_inst;           // ... and instance (-1 = miss), handed from the trace lane to the shade warp
    uint32_t *next_pixel;    // next unstarted pixel (index into the hand-out order)
    const uint32_t *tile_order;   // k-th 8x4 tile handed out
    uint32_t *tile_cost;     // rays traced per tile in this frame (orders the next frame)
    uint32_t n_pixels;       // tiles * 32
    uint32_t n_tiles, tiles_x;
    uint32_t n_blas_nodes;   // records in the BLAS node pool (bounds the shared-memory staging variant)
    // ---- frames in flight: ONE launch serves the pixel queues of n_frames consecutive frames back to back, so the lanes
    //      that run out of pixels of frame f start on frame f + 1 instead of idling through f's tail.  Frame f owns plane f
    //      of every per-pixel array (stride frame_px pixels) and copy f of the per-frame TLAS / instance records.
    uint32_t n_frames;       // >= 1
    uint32_t frame_px;       // local pixels per plane: rng / radiance / packed rgba8 / AOV stride between frames
    uint32_t full_px;        // pixels of a full (peer-mapped) frame: stride of the rgba8 ring when fr.rgba8_is_full_frame
    uint32_t tlas_node_stride, tlas_leaf_stride, inst_stride;   // records between the frames' copies (0 when n_frames == 1)
    uint32_t valid_pixels;   // pixels of one frame that lie inside the image (the others are tile padding)
    uint32_t ticket;         // value written to the completion flags (distinguishes this window from the previous one)
    unsigned long long *frame_rays;   // [n_frames] rays traced per frame (null: not wanted)
    uint32_t *frame_done;             // [n_frames] finished pixels per frame (null: no completion tracking)
    volatile uint32_t *host_flags;    // [n_frames] mapped pinned host memory: = ticket once the frame is complete
    volatile uint32_t *peer_flags;    // [n_frames] this rank's row of rank 0's device flag table (tile shards), or null
    const CameraD *frame_cams;        // [n_frames] camera of every frame of the window (rt_set_window_cameras), or null: the one camera
};

// kernels_split.cu
int split_slots_per_block();
int split_blocks_per_sm(bool count);
void launch_persist_split(const SceneD &sc, const CameraD &cam, const FrameD &fr, const PersistBuffers &pb, uint32_t blocks, int node_min, bool count,
                          cudaStream_t stream);

class PersistRenderer {
public:
    // one path-kernel launch per frame (plus the tile ordering); rays are counted on the device (fr.ray_counter[0])
    void render(const SceneD &sc, const CameraD &cam, const FrameD &fr, cudaStream_t stream, int sm_count) { render_window(sc, cam, fr, nullptr, stream, sm_count); }
    // the same launch over a window of frames (win: the frames-in-flight fields of PersistBuffers; null = one frame)
    void render_window(const SceneD &sc, const CameraD &cam, const FrameD &fr, const PersistBuffers *win, cudaStream_t stream, int sm_count);
    int node_min = 20;             // the warp takes a node step while at least this many lanes stand on an inner node
    int shade_min = 8;             // ... and shades finished rays once this many lanes wait for it
    int blocks_per_sm_override = 0;
    int tile_order = 1;            // 1: dear tiles first (cost = rays of the previous frame), 0: always scan order
    bool split = false;            // A/B: the kernel with a dedicated shade warp per CTA (kernels_split.cu) instead of persist8_kernel
    uint32_t launches = 0;         // kernels launched by the last render()
    uint32_t n_blas_nodes = 0;     // set by the caller: size of the BLAS wide-node pool
    bool single_instance = false;  // set by the caller: the scene is one instance -> persist8_kernel<.., .., SINGLE> (no top-level walk)
private:
    DeviceBuffer<float4> ray_o_, ray_d_, thr_, acc_;
    DeviceBuffer<uint4> rng_a_;
    DeviceBuffer<uint2> rng_b_;
    DeviceBuffer<float4> hit_;
    DeviceBuffer<int> hit_inst_;
    int split_blocks_per_sm_[2] = {0, 0};
    DeviceBuffer<uint32_t> counter_, tile_cost_, vals_a_, vals_b_;
    DeviceBuffer<uint64_t> keys_a_, keys_b_;
    RadixSorter sorter_;
    uint64_t history_sig_ = 0;
    int blocks_per_sm_[2] = {0, 0};
};
