// api_ctx.h — the opaque context behind include/rtb200.h, shared by the translation units that implement the C-ABI
// (api.cu: scene commits, builds, the synchronous frame; frames.cu: frames in flight, the frame ring).
#pragma once

#include "../../include/rtb200.h"

#include "build.h"
#include "host_scene.h"
#include "render.h"
#include "vtk_reader.h"
#include "wave.h"

#include <chrono>
#include <cstdlib>
#include <memory>
#include <string>
#include <thread>
#include <vector>

struct FrameWindow;          // frames.cu

// ---- the ring: 2 windows x RT_MAX_FRAMES_IN_FLIGHT full RGBA8 frames + the flag table, on rank 0's device --------------
struct FrameRing {
    uint8_t *base = nullptr;
    size_t frame_bytes = 0;
    uint32_t width = 0, height = 0, world = 0;
    int mode = 0;                      // 0 none, 1 owner (cudaMalloc), 2 IPC mapping, 3 attached to another context of this process
    static size_t flag_words(uint32_t world) { return (size_t) 2 * world * RT_MAX_FRAMES_IN_FLIGHT; }
    static size_t total_bytes(size_t frame_bytes, uint32_t world) {
        return (size_t) 2 * RT_MAX_FRAMES_IN_FLIGHT * frame_bytes + flag_words(world) * sizeof(uint32_t) + 256;
    }
    uint8_t *frame(uint32_t w, uint32_t k) const { return base + ((size_t) w * RT_MAX_FRAMES_IN_FLIGHT + k) * frame_bytes; }
    uint32_t *flag_table() const {
        const size_t off = ((size_t) 2 * RT_MAX_FRAMES_IN_FLIGHT * frame_bytes + 255) / 256 * 256;
        return reinterpret_cast<uint32_t *>(base + off);
    }
    uint32_t *flags(uint32_t w, uint32_t rank) const { return flag_table() + ((size_t) w * world + rank) * RT_MAX_FRAMES_IN_FLIGHT; }
};


struct rt_ctx {
    int device = 0;
    uint32_t flags = 0;
    std::string err;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev[5] = {};
    int sm_count = 148;

    HostScene hs;                            // host half: validated inputs + device-layout records
    rt_update_cb update_cb = nullptr;
    void *update_user = nullptr;

    // device scene
    DeviceBuffer<float> d_tri_verts;
    DeviceBuffer<TriShade> d_tri_shade;
    DeviceBuffer<SphereD> d_spheres;
    DeviceBuffer<QuadD> d_quads;
    DeviceBuffer<float4> d_roughs, d_metals, d_dielectrics;
    DeviceBuffer<unsigned char> d_ext;       // records of the extension primitive type (rt_prims.h), if the build has one
    DeviceBuffer<InstanceD> d_instances;
    DeviceBuffer<InstanceXformIn> d_xform_in;
    DeviceBuffer<float4> d_wcentroid;
    PinnedBuffer<InstanceXformIn> h_xform_in;
    LbvhBuilder blas, tlas;
    // The big per-frame uploads of the animation (triangle vertices and shading records, 52 MB at 629 k triangles) come
    // from HostScene's vectors; page-locking their storage in place makes the copies DMA at PCIe speed instead of being
    // staged through the driver's bounce buffer.  The registration follows the vector's allocation.
    struct PinnedRange {
        void *p = nullptr;
        size_t bytes = 0;
        void release() { if (p) cudaHostUnregister(p); p = nullptr; bytes = 0; }
        void ensure(void *q, size_t b) {
            if (q == p && b == bytes) return;
            release();
            if (q && b >= (4u << 20)) {
                if (cudaHostRegister(q, b, cudaHostRegisterDefault) == cudaSuccess) { p = q; bytes = b; } else cudaGetLastError();
            }
        }
        ~PinnedRange() { release(); }
    };
    PinnedRange pin_verts, pin_shade;
    DeviceBuffer<uint32_t> d_rng_tables;   // XORWOW sequence-jump matrices per hex digit (kernels_render.cu)
    bool rng_tables_ready = false;
    WaveRenderer wave;
    PersistRenderer persist;
    int path_mode = 1;              // 1: persistent path kernel (default), 0: wave loop; RT_FLAG_WAVEFRONT / RTB200_MODE select
    bool geometry_committed = false, materials_committed = false, instances_configured = false, built = false, camera_set = false;
    rt_build_stats build_stats{};
    // rt_build_acceleration_structure ends with the instance update + TLAS build for the current transform arguments.  The
    // FIRST frame rendered after it (the animation renders exactly one frame per rebuild) takes that TLAS instead of
    // building the identical one again; any later frame, a changed transform or an update callback rebuilds as always.
    bool tlas_fresh = false;

    // camera / frame
    uint32_t shard_rank = 0, shard_world = 1, tile_rows = 8, local_rows = 0;
    void *peer_rgba8 = nullptr;
    // frames in flight (frames.cu): two windows alternate; the ring is rank 0's full-frame store of a tile shard
    FrameWindow *win[2] = {nullptr, nullptr};
    FrameRing *ring = nullptr;
    uint32_t next_window = 0;
    std::vector<CameraD> window_cams;        // rt_set_window_cameras: cameras of the frames of the NEXT window (consumed by it)
    DeviceBuffer<uint8_t> d_rgba8;
    DeviceBuffer<float> d_radiance, d_aov_t;
    DeviceBuffer<int> d_aov_inst, d_aov_prim;
    DeviceBuffer<RngState> d_rng;
    DeviceBuffer<unsigned long long> d_counters;   // [0] rays, [1] overflow flag (low word)
    PinnedBuffer<unsigned long long> h_counters;

    // VTK series: the user's own triangles / instances are kept so that every frame can be recomposed as
    // [VTK triangles | user triangles], [user instances | VTK instances] (reference src/Global/Renderer.cu:17,106-116)
    std::shared_ptr<const VtkSeries> vtk;    // shared with the lanes of rt_run_vtk_animation
    // one parsed + converted series frame.  Two of them: the one being committed and the one a host thread reads ahead
    // (rt_prefetch_vtk_frame) while the GPU renders; the vectors keep their capacity from frame to frame.
    struct VtkStage {
        VtkFrame vf;
        VtkCompact compact;                 // device expansion (default): per-point tables + strip ids instead of `tris`
        bool expand_on_device = true;
        std::vector<rt_triangle> tris;
        std::vector<rt_instance_desc> inst;
        std::vector<rt_instance_xform> xf;
        size_t frame = (size_t) -1;
        bool ok = false;
        std::string err;
        float parse_ms = 0.0f;
        std::thread worker;
        void load(const VtkSeries &series, size_t f) {
            const auto t0 = std::chrono::steady_clock::now();
            frame = f;
            err.clear();
            ok = vf.read(series.dir + series.files[f].first, err);
            if (ok) {
                if (expand_on_device) vtk_convert_compact(vf, compact, inst, xf);
                else vtk_convert(vf, tris, inst, xf);
            }
            parse_ms = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
        }
        void join() { if (worker.joinable()) worker.join(); }
    };
    VtkStage vtk_stage[2];
    // strips -> triangle records on the device (launch_expand_strips); RTB200_VTK_EXPAND=host keeps the host conversion
    bool vtk_expand_on_device = true;
    DeviceBuffer<float> d_vtk_points, d_vtk_normals;
    DeviceBuffer<int> d_vtk_ids;
    DeviceBuffer<uint32_t> d_vtk_first_tri, d_vtk_first_id;
    int vtk_ahead = -1;             // stage a read-ahead is running on / has finished in (-1: none)
    std::vector<rt_triangle> user_tris;
    std::vector<rt_instance_desc> user_inst;
    bool composing = false;
    size_t n_user_inst = 0;
    float vtk_parse_ms = 0.0f;
    // rt_run_vtk_animation: the extra renderer contexts (lanes) and the pinned frame each lane delivers; kept between
    // calls so that their device buffers, page-locked ranges and jump tables are allocated once
    std::vector<rt_ctx *> lane_ctx;
    PinnedBuffer<uint8_t> h_frame;
    // Host waits of the frame path.  cudaStreamSynchronize spins; with several lanes per GPU and several ranks per host the
    // spinning threads would take the cores the other lanes' parsing and commits need, so inside the animation loop the
    // wait sleeps on a blocking-sync event instead.
    cudaEvent_t ev_block = nullptr;
    bool blocking_wait = false;
    void wait() {
        if (blocking_wait && ev_block) {
            CUDA_CHECK(cudaEventRecord(ev_block, stream));
            CUDA_CHECK(cudaEventSynchronize(ev_block));
        } else {
            CUDA_CHECK(cudaStreamSynchronize(stream));
        }
    }
};


inline int rt_fail(rt_ctx *ctx, int code, const std::string &msg) {
    if (ctx) ctx->err = msg;
    return code;
}

#define RT_API_BEGIN(ctx)                                   \
    if (!(ctx)) return RT_ERR_ARGUMENT;                     \
    try {                                                   \
        if (cudaSetDevice((ctx)->device) != cudaSuccess) return rt_fail(ctx, RT_ERR_CUDA, "cudaSetDevice failed");
#define RT_API_END(ctx)                                     \
    } catch (const CudaError &e) {                          \
        cudaGetLastError();                                 \
        return rt_fail(ctx, RT_ERR_CUDA, e.what());         \
    } catch (const std::exception &e) {                     \
        return rt_fail(ctx, RT_ERR_ARGUMENT, e.what());     \
    }                                                       \
    return RT_OK;

// a scene of one instance runs the path kernel's SINGLE instantiation (kernels_persist.cu); RTB200_SINGLE_INSTANCE=0: A/B runs
inline bool rt_single_instance(const rt_ctx *ctx) {
    static const bool enabled = !(getenv("RTB200_SINGLE_INSTANCE") && atoi(getenv("RTB200_SINGLE_INSTANCE")) == 0);
    return enabled && ctx->hs.inst_desc.size() == 1;
}

// shared between api.cu and frames.cu
void rt_compute_local_rows(rt_ctx *c);
SceneD rt_scene_view(rt_ctx *ctx);                            // device pointers of the committed + built scene
void rt_release_frames(rt_ctx *ctx);
bool rt_window_in_flight(const rt_ctx *ctx, int window);      // frames.cu                          // frames.cu: frees the windows and the ring (rt_cleanup)
