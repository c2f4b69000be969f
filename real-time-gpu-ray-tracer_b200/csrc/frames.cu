// frames.cu — frames in flight and the multi-GPU frame ring (include/rtb200.h: rt_submit_frames .. rt_run_frames,
// rt_frame_ring_*).
//
// The reference keeps two frames in flight: instances, TLAS and TraverseData are double-buffered and a copy stream
// prepares frame k+1 while the render stream traces frame k (src/Global/Renderer.cu:189-207, 278-317).  Here a WINDOW
// of up to RT_MAX_FRAMES_IN_FLIGHT frames is prepared on the window's stream (per frame: update callback, transform
// arguments H2D, instance matrices, TLAS rebuild, per-pixel XORWOW states) and then ONE launch of the persistent path
// kernel serves the frames' pixel queues back to back (kernels_persist.cu): a lane that finds frame k's queue empty
// takes a pixel of frame k+1, so the sparse tail of a frame — a pixel is sequential work — is filled with the next
// frame's pixels.  The lane that finishes a frame's last pixel raises the frame's flag in mapped host memory (and, for
// tile shards with a frame ring, in rank 0's flag table), so finished frames are read back while the window runs on.
#include "api_ctx.h"

#include <algorithm>
#include <chrono>
#include <cstring>
#include <thread>

struct FrameWindow {
    cudaStream_t stream = nullptr, copy_stream = nullptr;
    cudaEvent_t ev_begin = nullptr, ev_prep = nullptr, ev_end = nullptr, ev_copy = nullptr, ev_fork = nullptr;
    LbvhBuilder tlas;
    DeviceBuffer<InstanceXformIn> d_xform_in;
    PinnedBuffer<InstanceXformIn> h_xform_in;
    DeviceBuffer<InstanceD> d_inst_build;
    DeviceBuffer<float4> d_wcentroid;
    // per-frame copies (frame k at k * stride)
    DeviceBuffer<BvhNode8> tlas_nodes;
    DeviceBuffer<uint32_t> tlas_leaf;
    DeviceBuffer<InstanceD> instances;
    DeviceBuffer<RngState> rng;
    DeviceBuffer<float> radiance, aov_t;
    DeviceBuffer<uint8_t> rgba8;
    DeviceBuffer<int> aov_inst, aov_prim;
    DeviceBuffer<unsigned long long> counters;     // [0..7] as FrameD::ray_counter, [8 + k] rays of frame k
    DeviceBuffer<uint32_t> frame_done;
    PinnedBuffer<unsigned long long> h_counters;
    PinnedBuffer<uint32_t> h_peer_flags;           // rank 0: poll target for the ring's flag table
    DeviceBuffer<CameraD> d_cams;                  // per-frame cameras of the window (rt_set_window_cameras)
    PinnedBuffer<CameraD> h_cams;
    uint32_t *h_flags = nullptr;                   // mapped pinned: frame k complete <=> h_flags[k] == ticket
    PersistRenderer persist;
    // state of the window in flight
    bool in_flight = false;
    uint32_t n_frames = 0, ticket = 0, launches = 0;
    uint64_t first_frame = 0;
    uint32_t frame_px = 0, local_rows = 0, width = 0;
    std::chrono::steady_clock::time_point submitted;
    bool waited[RT_MAX_FRAMES_IN_FLIGHT] = {};
    float latency_ms[RT_MAX_FRAMES_IN_FLIGHT] = {};

    void create() {
        if (stream) return;
        CUDA_CHECK(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
        CUDA_CHECK(cudaStreamCreateWithFlags(&copy_stream, cudaStreamNonBlocking));
        for (cudaEvent_t *e: {&ev_begin, &ev_prep, &ev_end}) CUDA_CHECK(cudaEventCreate(e));
        CUDA_CHECK(cudaEventCreateWithFlags(&ev_copy, cudaEventDisableTiming));
        CUDA_CHECK(cudaEventCreateWithFlags(&ev_fork, cudaEventDisableTiming));
        CUDA_CHECK(cudaHostAlloc(&h_flags, RT_MAX_FRAMES_IN_FLIGHT * sizeof(uint32_t), cudaHostAllocMapped));
        memset(h_flags, 0, RT_MAX_FRAMES_IN_FLIGHT * sizeof(uint32_t));
    }
    ~FrameWindow() {
        if (stream) cudaStreamSynchronize(stream);
        if (copy_stream) cudaStreamSynchronize(copy_stream);
        for (cudaEvent_t e: {ev_begin, ev_prep, ev_end, ev_copy, ev_fork}) if (e) cudaEventDestroy(e);
        if (h_flags) cudaFreeHost(h_flags);
        if (stream) cudaStreamDestroy(stream);
        if (copy_stream) cudaStreamDestroy(copy_stream);
    }
};

bool rt_window_in_flight(const rt_ctx *ctx, int window) { return ctx->win[window] && ctx->win[window]->in_flight; }

void rt_release_frames(rt_ctx *ctx) {
    for (FrameWindow *&w: ctx->win) { delete w; w = nullptr; }
    if (ctx->ring) {
        FrameRing *r = ctx->ring;
        if (r->mode == 1 && r->base) cudaFree(r->base);
        if (r->mode == 2 && r->base) cudaIpcCloseMemHandle(r->base);
        delete r;
        ctx->ring = nullptr;
        ctx->peer_rgba8 = nullptr;
    }
}

namespace {

// Back-pressure of the frame ring.  A rank may not store into window slot w again before rank 0 has let go of the frames the
// slot held: rank 0 publishes, per slot, the ticket of the last window it RELEASED (two words behind the flag table, inside
// the allocation's slack), and the other ranks wait for it before they submit the slot's next window.  Without it a rank
// that runs ahead of a slow consumer on rank 0 (a sink that writes files) would overwrite frames rank 0 has not read yet.
uint32_t *ring_ack(const FrameRing *r, uint32_t window) { return r->flag_table() + FrameRing::flag_words(r->world) + window; }

uint32_t max_shard_rows(const rt_ctx *c) {
    const uint32_t tiles = (c->hs.window.height + c->tile_rows - 1) / c->tile_rows;
    return c->shard_world == 1 ? c->hs.window.height : ((tiles + c->shard_world - 1) / c->shard_world) * c->tile_rows;
}

uint32_t valid_local_rows(const rt_ctx *c) {
    uint32_t rows = 0;
    for (uint32_t ly = 0; ly < c->local_rows; ly++) {
        const uint32_t gy = ((ly / c->tile_rows) * c->shard_world + c->shard_rank) * c->tile_rows + (ly % c->tile_rows);
        if (gy < c->hs.cam.height) rows++;
    }
    return rows;
}

int check_window(rt_ctx *ctx, uint32_t window, const char *who) {
    if (window > 1 || !ctx->win[window] || !ctx->win[window]->in_flight) return rt_fail(ctx, RT_ERR_STATE, std::string(who) + ": no such window in flight");
    return RT_OK;
}

// waits until the host flag of frame k shows the window's ticket; watches the stream so that a failed launch is an
// error return, not a hang
int wait_flag(rt_ctx *ctx, FrameWindow &w, uint32_t k) {
    volatile uint32_t *flag = w.h_flags + k;
    uint32_t spins = 0;
    while (*flag != w.ticket) {
        if (++spins % 2048u == 0u) {
            const cudaError_t q = cudaStreamQuery(w.stream);
            if (q == cudaSuccess) {
                if (*flag == w.ticket) break;
                return rt_fail(ctx, RT_ERR_CUDA, "frames in flight: the window finished without completing the frame");
            }
            if (q != cudaErrorNotReady) { cudaGetLastError(); return rt_fail(ctx, RT_ERR_CUDA, std::string("frames in flight: ") + cudaGetErrorString(q)); }
            if (ctx->blocking_wait) std::this_thread::sleep_for(std::chrono::microseconds(50));
            else std::this_thread::yield();
        }
    }
    std::atomic_thread_fence(std::memory_order_acquire);
    return RT_OK;
}

// rank 0 of a tile shard: every rank's flag of (window, frame k) must carry the ticket
int wait_peers(rt_ctx *ctx, FrameWindow &w, uint32_t window, uint32_t k) {
    FrameRing *r = ctx->ring;
    if (!r || ctx->shard_rank != 0 || r->world <= 1) return RT_OK;
    w.h_peer_flags.reserve(FrameRing::flag_words(r->world));
    const auto t0 = std::chrono::steady_clock::now();
    for (;;) {
        CUDA_CHECK(cudaMemcpyAsync(w.h_peer_flags.get(), r->flags(window, 0), (size_t) r->world * RT_MAX_FRAMES_IN_FLIGHT * sizeof(uint32_t),
                                   cudaMemcpyDeviceToHost, w.copy_stream));
        CUDA_CHECK(cudaStreamSynchronize(w.copy_stream));
        bool all = true;
        for (uint32_t q = 0; q < r->world; q++) all = all && w.h_peer_flags.get()[(size_t) q * RT_MAX_FRAMES_IN_FLIGHT + k] == w.ticket;
        if (all) return RT_OK;
        if (std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() > 120.0)
            return rt_fail(ctx, RT_ERR_STATE, "frame ring: a rank did not deliver its tiles of the frame within 120 s (did every rank submit the same window?)");
        std::this_thread::yield();
    }
}

} // namespace

extern "C" {

int rt_submit_frames(rt_ctx *ctx, uint64_t first_frame, uint32_t n_frames, const uint64_t *seeds, uint64_t seed0, uint32_t *window) {
    RT_API_BEGIN(ctx)
    if (!ctx->built || !ctx->camera_set || !ctx->materials_committed) return rt_fail(ctx, RT_ERR_STATE, "rt_submit_frames: scene not built / camera or materials missing");
    if (n_frames < 1 || n_frames > RT_MAX_FRAMES_IN_FLIGHT || !window) return rt_fail(ctx, RT_ERR_ARGUMENT, "rt_submit_frames: 1 <= n_frames <= RT_MAX_FRAMES_IN_FLIGHT");
    if ((ctx->flags & (RT_FLAG_MEGAKERNEL | RT_FLAG_WAVEFRONT | RT_FLAG_TRAVERSE_BVH2 | RT_FLAG_COUNT_TRAVERSAL)) || ctx->path_mode != 1 || ctx->hs.cam.depth == 0)
        return rt_fail(ctx, RT_ERR_STATE, "rt_submit_frames: frames in flight run on the default path kernel only (no comparison / counting mode, depth >= 1)");
    {
        const std::string e = ctx->hs.check_materials();
        if (!e.empty()) return rt_fail(ctx, RT_ERR_ARGUMENT, "rt_submit_frames: " + e);
    }
    if (ctx->peer_rgba8 && !ctx->ring) return rt_fail(ctx, RT_ERR_STATE, "rt_submit_frames: a raw peer framebuffer holds one frame; use a frame ring");
    uint32_t wi = ctx->next_window & 1u;
    if (ctx->win[wi] && ctx->win[wi]->in_flight) wi ^= 1u;
    if (ctx->win[wi] && ctx->win[wi]->in_flight) return rt_fail(ctx, RT_ERR_STATE, "rt_submit_frames: both windows are in flight (rt_release_window first)");
    if (!ctx->win[wi]) ctx->win[wi] = new FrameWindow();
    FrameWindow &w = *ctx->win[wi];
    w.create();
    HostScene &hs = ctx->hs;
    const int n_inst = (int) hs.inst_desc.size();
    rt_compute_local_rows(ctx);
    const uint32_t W = hs.window.width;
    const size_t px = (size_t) W * max_shard_rows(ctx);
    const uint32_t node_stride = (uint32_t) n_inst + 1u, leaf_stride = (uint32_t) n_inst, inst_stride = (uint32_t) n_inst;
    const bool aov = ctx->flags & RT_FLAG_RECORD_PRIMARY_AOV;
    // allocations (grow-only; a window's buffers are its own, so the other window may be running)
    w.d_xform_in.reserve((size_t) n_frames * n_inst);
    w.h_xform_in.reserve((size_t) RT_MAX_FRAMES_IN_FLIGHT * n_inst);
    w.d_inst_build.reserve(n_inst);
    w.d_wcentroid.reserve(n_inst);
    w.tlas.emit_bvh2 = false;
    w.tlas.sah_rotations = ctx->tlas.sah_rotations;
    w.tlas.prepare_tlas(n_inst);
    w.tlas_nodes.reserve((size_t) n_frames * node_stride);
    w.tlas_leaf.reserve((size_t) n_frames * leaf_stride);
    w.instances.reserve((size_t) n_frames * inst_stride);
    w.rng.reserve(n_frames * px + 1);
    w.radiance.reserve(n_frames * px * 3 + 3);
    w.rgba8.reserve(n_frames * px * 4 + 4);
    if (aov) { w.aov_t.reserve(n_frames * px + 1); w.aov_inst.reserve(n_frames * px + 1); w.aov_prim.reserve(n_frames * px + 1); }
    w.counters.reserve(8 + RT_MAX_FRAMES_IN_FLIGHT);
    w.h_counters.reserve(8 + RT_MAX_FRAMES_IN_FLIGHT);
    w.frame_done.reserve(RT_MAX_FRAMES_IN_FLIGHT);
    const bool cams = !ctx->window_cams.empty();
    if (cams && ctx->window_cams.size() != n_frames) {
        ctx->window_cams.clear();
        return rt_fail(ctx, RT_ERR_ARGUMENT, "rt_submit_frames: rt_set_window_cameras gave another number of cameras than this window has frames");
    }
    if (cams) { w.d_cams.reserve(RT_MAX_FRAMES_IN_FLIGHT); w.h_cams.reserve(RT_MAX_FRAMES_IN_FLIGHT); }
    ctx->d_rng_tables.reserve(RNG_TABLE_WORDS);

    cudaStream_t s = w.stream;
    // everything already queued on the context's stream (commits, builds, a synchronous frame) comes first
    CUDA_CHECK(cudaEventRecord(w.ev_fork, ctx->stream));
    CUDA_CHECK(cudaStreamWaitEvent(s, w.ev_fork, 0));
    if (!ctx->rng_tables_ready) {
        // the jump tables are shared by every stream of the context: build them where everybody is ordered behind
        launch_rng_tables(ctx->d_rng_tables.get(), ctx->rng_tables_ready, ctx->stream);
        CUDA_CHECK(cudaEventRecord(w.ev_fork, ctx->stream));
        CUDA_CHECK(cudaStreamWaitEvent(s, w.ev_fork, 0));
    }
    ctx->tlas_fresh = false;
    if (ctx->ring && ctx->shard_rank != 0 && w.ticket != 0u) {
        // rank 0 must have released this slot's previous window (ring_ack)
        w.h_peer_flags.reserve(FrameRing::flag_words(ctx->ring->world));
        const auto t0 = std::chrono::steady_clock::now();
        for (;;) {
            CUDA_CHECK(cudaMemcpyAsync(w.h_peer_flags.get(), ring_ack(ctx->ring, wi), sizeof(uint32_t), cudaMemcpyDeviceToHost, w.copy_stream));
            CUDA_CHECK(cudaStreamSynchronize(w.copy_stream));
            if (w.h_peer_flags.get()[0] >= w.ticket) break;
            if (std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() > 120.0)
                return rt_fail(ctx, RT_ERR_STATE, "frame ring: rank 0 did not release the previous window of this slot within 120 s");
            std::this_thread::yield();
        }
    }
    w.ticket++;
    if (w.ticket == 0u) w.ticket = 1u;
    w.n_frames = n_frames;
    w.first_frame = first_frame;
    w.frame_px = (uint32_t) px;
    w.local_rows = ctx->local_rows;
    w.width = W;
    w.launches = 0;
    for (uint32_t k = 0; k < RT_MAX_FRAMES_IN_FLIGHT; k++) { w.waited[k] = false; w.latency_ms[k] = 0.0f; }
    w.submitted = std::chrono::steady_clock::now();
    CUDA_CHECK(cudaEventRecord(w.ev_begin, s));
    CUDA_CHECK(cudaMemsetAsync(w.counters.get(), 0, (8 + RT_MAX_FRAMES_IN_FLIGHT) * sizeof(unsigned long long), s));
    CUDA_CHECK(cudaMemsetAsync(w.frame_done.get(), 0, RT_MAX_FRAMES_IN_FLIGHT * sizeof(uint32_t), s));
    if (cams) {
        std::copy(ctx->window_cams.begin(), ctx->window_cams.end(), w.h_cams.get());
        CUDA_CHECK(cudaMemcpyAsync(w.d_cams.get(), w.h_cams.get(), n_frames * sizeof(CameraD), cudaMemcpyHostToDevice, s));
        ctx->window_cams.clear();                     // the cameras belong to this window only
    }

    FrameD fr{};
    FrameRing *ring = ctx->ring;
    fr.rgba8 = ring ? ring->frame(wi, 0) : w.rgba8.get();
    fr.rgba8_is_full_frame = ring ? 1u : 0u;
    fr.radiance = w.radiance.get();
    fr.aov_t = aov ? w.aov_t.get() : nullptr;
    fr.aov_inst = aov ? w.aov_inst.get() : nullptr;
    fr.aov_prim = aov ? w.aov_prim.get() : nullptr;
    fr.rng = w.rng.get();
    fr.ray_counter = w.counters.get();
    fr.overflow_flag = reinterpret_cast<unsigned int *>(w.counters.get() + 1);
    fr.shard_rank = ctx->shard_rank; fr.shard_world = ctx->shard_world; fr.tile_rows = ctx->tile_rows; fr.local_rows = ctx->local_rows;

    // ---- preparation of every frame of the window --------------------------------------------------------------------
    for (uint32_t k = 0; k < n_frames; k++) {
        if (ctx->update_cb) ctx->update_cb(hs.xforms.data(), ctx->n_user_inst, (size_t) (first_frame + k), ctx->update_user);
        InstanceXformIn *hx = w.h_xform_in.get() + (size_t) k * n_inst;
        hs.fill_xform_in(hx, ctx->blas.seg_root_host);
        InstanceXformIn *dx = w.d_xform_in.get() + (size_t) k * n_inst;
        CUDA_CHECK(cudaMemcpyAsync(dx, hx, (size_t) n_inst * sizeof(InstanceXformIn), cudaMemcpyHostToDevice, s));
        launch_instance_update(dx, n_inst, w.d_inst_build.get(), w.tlas.prim_lo(), w.tlas.prim_hi(), w.d_wcentroid.get(), w.tlas.bounds(),
                               ctx->d_tri_verts.get(), s);
        w.tlas.build_tlas(n_inst, tlas_morton_bits((size_t) n_inst), s, ctx->sm_count);
        CUDA_CHECK(cudaMemcpyAsync(w.tlas_nodes.get() + (size_t) k * node_stride, w.tlas.nodes8(), (size_t) node_stride * sizeof(BvhNode8), cudaMemcpyDeviceToDevice, s));
        CUDA_CHECK(cudaMemcpyAsync(w.tlas_leaf.get() + (size_t) k * leaf_stride, w.tlas.leaf_prims8(), (size_t) leaf_stride * sizeof(uint32_t), cudaMemcpyDeviceToDevice, s));
        CUDA_CHECK(cudaMemcpyAsync(w.instances.get() + (size_t) k * inst_stride, w.d_inst_build.get(), (size_t) inst_stride * sizeof(InstanceD), cudaMemcpyDeviceToDevice, s));
        launch_rng_init(w.rng.get() + (size_t) k * px, seeds ? seeds[k] : seed0, hs.cam, fr, s, ctx->d_rng_tables.get(), ctx->rng_tables_ready);
        w.launches += w.tlas.launches + 3 + 1;
    }
    CUDA_CHECK(cudaEventRecord(w.ev_prep, s));

    // ---- one launch for the whole window -------------------------------------------------------------------------
    SceneD sc = rt_scene_view(ctx);
    sc.tlas_nodes8 = w.tlas_nodes.get();
    sc.tlas_leaf_inst = w.tlas_leaf.get();
    sc.instances = w.instances.get();
    PersistBuffers win{};
    win.n_frames = n_frames;
    win.frame_px = (uint32_t) px;
    win.full_px = W * hs.window.height;
    win.tlas_node_stride = node_stride; win.tlas_leaf_stride = leaf_stride; win.inst_stride = inst_stride;
    win.valid_pixels = W * valid_local_rows(ctx);
    win.ticket = w.ticket;
    win.frame_rays = w.counters.get() + 8;
    win.frame_done = w.frame_done.get();
    uint32_t *dev_flags = nullptr;
    CUDA_CHECK(cudaHostGetDevicePointer(reinterpret_cast<void **>(&dev_flags), w.h_flags, 0));
    win.host_flags = dev_flags;
    win.peer_flags = ring ? ring->flags(wi, ctx->shard_rank) : nullptr;
    win.frame_cams = cams ? w.d_cams.get() : nullptr;
    w.persist.node_min = ctx->persist.node_min; w.persist.shade_min = ctx->persist.shade_min;
    w.persist.blocks_per_sm_override = ctx->persist.blocks_per_sm_override; w.persist.tile_order = ctx->persist.tile_order;
    w.persist.split = ctx->persist.split;
    w.persist.n_blas_nodes = (uint32_t) ctx->blas.n_prims();
    w.persist.single_instance = rt_single_instance(ctx);
    if (win.valid_pixels == 0u) {
        // a shard without a single scan-line inside the image: nothing to trace, every frame is complete at once
        for (uint32_t k = 0; k < n_frames; k++) w.h_flags[k] = w.ticket;
        if (win.peer_flags) for (uint32_t k = 0; k < n_frames; k++) CUDA_CHECK(cudaMemcpyAsync((void *) (win.peer_flags + k), &w.h_flags[k], sizeof(uint32_t), cudaMemcpyHostToDevice, s));
    } else {
        w.persist.render_window(sc, hs.cam, fr, &win, s, ctx->sm_count);
        w.launches += w.persist.launches;
    }
    CUDA_CHECK(cudaEventRecord(w.ev_end, s));
    CUDA_CHECK(cudaMemcpyAsync(w.h_counters.get(), w.counters.get(), (8 + RT_MAX_FRAMES_IN_FLIGHT) * sizeof(unsigned long long), cudaMemcpyDeviceToHost, s));
    w.in_flight = true;
    ctx->next_window = wi ^ 1u;
    *window = wi;
    RT_API_END(ctx)
}

int rt_set_window_cameras(rt_ctx *ctx, const rt_camera_input *cameras, uint32_t n_frames) {
    RT_API_BEGIN(ctx)
    ctx->window_cams.clear();
    if (n_frames == 0) return RT_OK;
    if (!ctx->camera_set) return rt_fail(ctx, RT_ERR_STATE, "rt_set_window_cameras: configure the camera first (window size, samples and depth come from it)");
    if (!cameras || n_frames > RT_MAX_FRAMES_IN_FLIGHT) return rt_fail(ctx, RT_ERR_ARGUMENT, "rt_set_window_cameras: 1 <= n_frames <= RT_MAX_FRAMES_IN_FLIGHT cameras");
    std::vector<CameraD> cams(n_frames);
    for (uint32_t k = 0; k < n_frames; k++) {
        const std::string e = HostScene::make_camera(ctx->hs.window, cameras[k], cams[k]);
        if (!e.empty()) return rt_fail(ctx, RT_ERR_ARGUMENT, e);
        if (cams[k].sqrtSamples != ctx->hs.cam.sqrtSamples || cams[k].depth != ctx->hs.cam.depth)
            return rt_fail(ctx, RT_ERR_ARGUMENT, "rt_set_window_cameras: sample_count and ray_trace_depth are those of the configured camera for every frame of a window");
    }
    ctx->window_cams = cams;
    RT_API_END(ctx)
}

int rt_wait_frame(rt_ctx *ctx, uint32_t window, uint32_t k, rt_frame_out *out) {
    RT_API_BEGIN(ctx)
    int rc = check_window(ctx, window, "rt_wait_frame");
    if (rc != RT_OK) return rc;
    FrameWindow &w = *ctx->win[window];
    if (k >= w.n_frames) return rt_fail(ctx, RT_ERR_ARGUMENT, "rt_wait_frame: frame index beyond the window");
    rc = wait_flag(ctx, w, k);
    if (rc != RT_OK) return rc;
    if (!w.waited[k]) {
        w.latency_ms[k] = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - w.submitted).count();
        w.waited[k] = true;
    }
    rc = wait_peers(ctx, w, window, k);
    if (rc != RT_OK) return rc;
    if (out) {
        memset(out, 0, sizeof(*out));
        // the frame's ray count sits in device memory; it is read with the frame (4 + 8 bytes on the copy stream)
        unsigned long long rays = 0, overflow = 0;
        CUDA_CHECK(cudaMemcpyAsync(&w.h_counters.get()[8 + k], w.counters.get() + 8 + k, sizeof(unsigned long long), cudaMemcpyDeviceToHost, w.copy_stream));
        CUDA_CHECK(cudaMemcpyAsync(&w.h_counters.get()[1], w.counters.get() + 1, sizeof(unsigned long long), cudaMemcpyDeviceToHost, w.copy_stream));
        CUDA_CHECK(cudaStreamSynchronize(w.copy_stream));
        rays = w.h_counters.get()[8 + k];
        overflow = w.h_counters.get()[1];
        out->rays = rays;
        out->primary_rays = (uint64_t) w.width * valid_local_rows(ctx) * ctx->hs.cam.sqrtSamples * ctx->hs.cam.sqrtSamples;
        out->render_ms = w.latency_ms[k];
        out->stack_overflow = (uint32_t) overflow;
        out->kernel_launches = w.launches / std::max(1u, w.n_frames);
    }
    RT_API_END(ctx)
}

int rt_read_frame(rt_ctx *ctx, uint32_t window, uint32_t k, int kind, void *host_dst, size_t dst_bytes) {
    RT_API_BEGIN(ctx)
    int rc = check_window(ctx, window, "rt_read_frame");
    if (rc != RT_OK) return rc;
    FrameWindow &w = *ctx->win[window];
    if (k >= w.n_frames) return rt_fail(ctx, RT_ERR_ARGUMENT, "rt_read_frame: frame index beyond the window");
    if (!w.waited[k]) {
        rc = rt_wait_frame(ctx, window, k, nullptr);
        if (rc != RT_OK) return rc;
    }
    const size_t px = (size_t) w.width * w.local_rows, off = (size_t) k * w.frame_px;
    const void *src = nullptr;
    size_t n = 0;
    const bool aov = ctx->flags & RT_FLAG_RECORD_PRIMARY_AOV;
    switch (kind) {
        case RT_BUFFER_RGBA8:
            if (ctx->ring) { src = ctx->ring->frame(window, k); n = ctx->ring->frame_bytes; }
            else { src = w.rgba8.get() + off * 4; n = px * 4; }
            break;
        case RT_BUFFER_RADIANCE: src = w.radiance.get() + off * 3; n = px * 12; break;
        case RT_BUFFER_PRIMARY_T: src = aov ? w.aov_t.get() + off : nullptr; n = px * 4; break;
        case RT_BUFFER_PRIMARY_INSTANCE: src = aov ? w.aov_inst.get() + off : nullptr; n = px * 4; break;
        case RT_BUFFER_PRIMARY_PRIM: src = aov ? w.aov_prim.get() + off : nullptr; n = px * 4; break;
        default: return rt_fail(ctx, RT_ERR_ARGUMENT, "rt_read_frame: unknown buffer kind");
    }
    if (!src) return rt_fail(ctx, RT_ERR_STATE, "primary AOVs need RT_FLAG_RECORD_PRIMARY_AOV");
    if (!host_dst || dst_bytes < n) return rt_fail(ctx, RT_ERR_ARGUMENT, "rt_read_frame: destination too small");
    CUDA_CHECK(cudaMemcpyAsync(host_dst, src, n, cudaMemcpyDeviceToHost, w.copy_stream));
    if (ctx->blocking_wait) {
        CUDA_CHECK(cudaEventRecord(ctx->ev_block, w.copy_stream));
        CUDA_CHECK(cudaEventSynchronize(ctx->ev_block));
    } else {
        CUDA_CHECK(cudaStreamSynchronize(w.copy_stream));
    }
    RT_API_END(ctx)
}

int rt_release_window(rt_ctx *ctx, uint32_t window, rt_window_stats *stats) {
    RT_API_BEGIN(ctx)
    const int rc = check_window(ctx, window, "rt_release_window");
    if (rc != RT_OK) return rc;
    FrameWindow &w = *ctx->win[window];
    w.in_flight = false;                       // whatever happens below, the handle is free again
    CUDA_CHECK(cudaStreamSynchronize(w.stream));
    CUDA_CHECK(cudaStreamSynchronize(w.copy_stream));
    if (stats) {
        memset(stats, 0, sizeof(*stats));
        stats->rays = w.h_counters.get()[0];
        stats->frames = w.n_frames;
        stats->kernel_launches = w.launches;
        stats->stack_overflow = (uint32_t) w.h_counters.get()[1];
        CUDA_CHECK(cudaEventElapsedTime(&stats->device_ms, w.ev_begin, w.ev_end));
        CUDA_CHECK(cudaEventElapsedTime(&stats->prepare_ms, w.ev_begin, w.ev_prep));
    }
    if (ctx->ring && ctx->shard_rank == 0 && ctx->ring->world > 1) {       // the slot's frames may be overwritten now (ring_ack)
        w.h_peer_flags.reserve(FrameRing::flag_words(ctx->ring->world));
        w.h_peer_flags.get()[0] = w.ticket;
        CUDA_CHECK(cudaMemcpyAsync(ring_ack(ctx->ring, window), w.h_peer_flags.get(), sizeof(uint32_t), cudaMemcpyHostToDevice, w.copy_stream));
        CUDA_CHECK(cudaStreamSynchronize(w.copy_stream));
    }
    if (w.h_counters.get()[1] != 0) return rt_fail(ctx, RT_ERR_OVERFLOW, "frames in flight: a traversal stack overflowed: the frames of this window are incomplete");
    RT_API_END(ctx)
}

int rt_run_frames(rt_ctx *ctx, uint64_t first_frame, size_t n_frames, uint64_t seed, uint32_t frames_per_window, rt_frame_sink sink, void *user,
                  rt_loop_stats *stats) {
    RT_API_BEGIN(ctx)
    if (frames_per_window < 1 || frames_per_window > RT_MAX_FRAMES_IN_FLIGHT) return rt_fail(ctx, RT_ERR_ARGUMENT, "rt_run_frames: 1 <= frames_per_window <= RT_MAX_FRAMES_IN_FLIGHT");
    for (uint32_t q = 0; q < 2; q++)
        if (ctx->win[q] && ctx->win[q]->in_flight) return rt_fail(ctx, RT_ERR_STATE, "rt_run_frames: release the windows in flight first");
    rt_loop_stats tot{};
    const auto t0 = std::chrono::steady_clock::now();
    CUDA_CHECK(cudaEventRecord(ctx->ev[0], ctx->stream));          // every window's stream is ordered behind the context's stream
    FrameWindow *last = nullptr;
    struct Pending { uint32_t window, n; uint64_t first; };
    std::vector<Pending> queue;                 // submitted, not yet drained (at most two)
    size_t next = 0;
    int rc = RT_OK;
    bool stopped = false;
    auto submit = [&]() {
        const uint32_t n = (uint32_t) std::min<size_t>(frames_per_window, n_frames - next);
        uint32_t wi = 0;
        rc = rt_submit_frames(ctx, first_frame + next, n, nullptr, seed, &wi);
        if (rc == RT_OK) { queue.push_back({wi, n, first_frame + next}); next += n; }
    };
    while (rc == RT_OK && !stopped && (next < n_frames || !queue.empty())) {
        while (rc == RT_OK && queue.size() < 2 && next < n_frames) submit();
        if (rc != RT_OK || queue.empty()) break;
        const Pending p = queue.front();
        queue.erase(queue.begin());
        for (uint32_t k = 0; k < p.n && rc == RT_OK && !stopped; k++) {
            rt_frame_out out{};
            rc = rt_wait_frame(ctx, p.window, k, &out);
            if (rc != RT_OK) break;
            tot.frames++;
            tot.rays += out.rays;
            tot.max_latency_ms = std::max<double>(tot.max_latency_ms, out.render_ms);
            if (sink) {
                // with a frame ring the pixels of every rank are in rank 0's ring: the other ranks deliver none
                const bool none = ctx->ring && ctx->shard_rank != 0;
                const size_t bytes = none ? 0 : ctx->ring ? ctx->ring->frame_bytes : (size_t) ctx->hs.window.width * ctx->local_rows * 4;
                if (!none) {
                    ctx->h_frame.reserve(bytes + 4);
                    rc = rt_read_frame(ctx, p.window, k, RT_BUFFER_RGBA8, ctx->h_frame.get(), bytes);
                    if (rc != RT_OK) break;
                }
                if (sink(user, (size_t) (p.first + k), none ? nullptr : ctx->h_frame.get(), bytes, &out) != 0) stopped = true;
            }
        }
        rt_window_stats ws{};
        last = ctx->win[p.window];
        const int rr = rt_release_window(ctx, p.window, &ws);
        if (rc == RT_OK) rc = rr;
        tot.device_ms += ws.device_ms;
        tot.kernel_launches += ws.kernel_launches;
    }
    for (const Pending &p: queue) rt_release_window(ctx, p.window, nullptr);      // after an error or a stop
    tot.wall_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    if (last && rc == RT_OK && !stopped) {
        float span = 0.0f;
        CUDA_CHECK(cudaEventSynchronize(ctx->ev[0]));
        CUDA_CHECK(cudaEventElapsedTime(&span, ctx->ev[0], last->ev_end));
        tot.span_ms = span;
    }
    if (stats) *stats = tot;
    if (rc != RT_OK) return rc;
    if (stopped) return rt_fail(ctx, RT_ERR_STATE, "rt_run_frames: stopped by the frame sink");
    RT_API_END(ctx)
}

// ---- the frame ring -------------------------------------------------------------------------------------------------
static int ring_preconditions(rt_ctx *ctx, const char *who) {
    if (!ctx->camera_set) return rt_fail(ctx, RT_ERR_STATE, std::string(who) + ": configure the camera first");
    if (ctx->ring) return rt_fail(ctx, RT_ERR_STATE, std::string(who) + ": this context already has a frame ring");
    for (uint32_t q = 0; q < 2; q++)
        if (ctx->win[q] && ctx->win[q]->in_flight) return rt_fail(ctx, RT_ERR_STATE, std::string(who) + ": not while frames are in flight");
    return RT_OK;
}

static void ring_adopt(rt_ctx *ctx, FrameRing *r) {
    // the ranks of a shard compare tickets through the ring's flag table: every rank counts its windows from the moment
    // the ring exists (and must submit the same sequence of windows from then on)
    for (FrameWindow *w: ctx->win) if (w) w->ticket = 0;
    ctx->next_window = 0;
    ctx->ring = r;
    ctx->peer_rgba8 = r->frame(0, 0);          // the synchronous rt_render_frame stores into slot 0 of window 0
}

int rt_frame_ring_create(rt_ctx *ctx, void *ipc_handle_out) {
    RT_API_BEGIN(ctx)
    const int rc = ring_preconditions(ctx, "rt_frame_ring_create");
    if (rc != RT_OK) return rc;
    FrameRing *r = new FrameRing();
    r->width = ctx->hs.window.width; r->height = ctx->hs.window.height; r->world = ctx->shard_world;
    r->frame_bytes = (size_t) r->width * r->height * 4;
    const size_t bytes = FrameRing::total_bytes(r->frame_bytes, r->world);
    if (cudaMalloc(&r->base, bytes) != cudaSuccess) { cudaGetLastError(); delete r; return rt_fail(ctx, RT_ERR_CUDA, "rt_frame_ring_create: out of device memory"); }
    r->mode = 1;
    CUDA_CHECK(cudaMemsetAsync(r->base, 0, bytes, ctx->stream));
    CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
    if (ipc_handle_out) {
        cudaIpcMemHandle_t h;
        static_assert(sizeof(cudaIpcMemHandle_t) == RT_IPC_HANDLE_BYTES, "CUDA IPC handle size");
        if (cudaIpcGetMemHandle(&h, r->base) != cudaSuccess) {
            const std::string e = cudaGetErrorString(cudaGetLastError());
            cudaFree(r->base);
            delete r;
            return rt_fail(ctx, RT_ERR_CUDA, "rt_frame_ring_create: cudaIpcGetMemHandle: " + e);
        }
        memcpy(ipc_handle_out, &h, sizeof(h));
    }
    ring_adopt(ctx, r);
    RT_API_END(ctx)
}

int rt_frame_ring_open(rt_ctx *ctx, const void *ipc_handle) {
    RT_API_BEGIN(ctx)
    if (!ipc_handle) return rt_fail(ctx, RT_ERR_ARGUMENT, "rt_frame_ring_open: null handle");
    const int rc = ring_preconditions(ctx, "rt_frame_ring_open");
    if (rc != RT_OK) return rc;
    cudaIpcMemHandle_t h;
    memcpy(&h, ipc_handle, sizeof(h));
    void *p = nullptr;
    const cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) { cudaGetLastError(); return rt_fail(ctx, RT_ERR_CUDA, std::string("rt_frame_ring_open: cudaIpcOpenMemHandle: ") + cudaGetErrorString(e)); }
    FrameRing *r = new FrameRing();
    r->base = static_cast<uint8_t *>(p);
    r->width = ctx->hs.window.width; r->height = ctx->hs.window.height; r->world = ctx->shard_world;
    r->frame_bytes = (size_t) r->width * r->height * 4;
    r->mode = 2;
    ring_adopt(ctx, r);
    RT_API_END(ctx)
}

int rt_frame_ring_attach(rt_ctx *ctx, rt_ctx *owner) {
    RT_API_BEGIN(ctx)
    if (!owner || !owner->ring || owner->ring->mode != 1) return rt_fail(ctx, RT_ERR_ARGUMENT, "rt_frame_ring_attach: the owner has no ring of its own");
    const int rc = ring_preconditions(ctx, "rt_frame_ring_attach");
    if (rc != RT_OK) return rc;
    const FrameRing *o = owner->ring;
    if (o->width != ctx->hs.window.width || o->height != ctx->hs.window.height || o->world != ctx->shard_world)
        return rt_fail(ctx, RT_ERR_ARGUMENT, "rt_frame_ring_attach: the ring was made for another frame size or shard count");
    if (owner->device != ctx->device) {
        int can = 0;
        CUDA_CHECK(cudaDeviceCanAccessPeer(&can, ctx->device, owner->device));
        if (!can) return rt_fail(ctx, RT_ERR_CUDA, "rt_frame_ring_attach: no peer access between the two devices");
        const cudaError_t e = cudaDeviceEnablePeerAccess(owner->device, 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); return rt_fail(ctx, RT_ERR_CUDA, std::string("cudaDeviceEnablePeerAccess: ") + cudaGetErrorString(e)); }
        cudaGetLastError();
    }
    FrameRing *r = new FrameRing(*o);
    r->mode = 3;
    ring_adopt(ctx, r);
    RT_API_END(ctx)
}

int rt_frame_ring_close(rt_ctx *ctx) {
    RT_API_BEGIN(ctx)
    for (uint32_t q = 0; q < 2; q++)
        if (ctx->win[q] && ctx->win[q]->in_flight) return rt_fail(ctx, RT_ERR_STATE, "rt_frame_ring_close: not while frames are in flight");
    if (ctx->ring) {
        CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
        FrameRing *r = ctx->ring;
        if (r->mode == 1 && r->base) cudaFree(r->base);
        if (r->mode == 2 && r->base) cudaIpcCloseMemHandle(r->base);
        delete r;
        ctx->ring = nullptr;
        ctx->peer_rgba8 = nullptr;
    }
    RT_API_END(ctx)
}

} // extern "C"
