// api.cu — the C-ABI of include/rtb200.h: an opaque context that owns every host/device resource, mirrors
// the reference's Renderer::* entry points and returns error codes instead of exiting.
#include "../../include/rtb200.h"

#include "api_ctx.h"
#include "image_io.h"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <functional>
#include <mutex>
#include <cstring>
#include <map>
#include <thread>
#include <memory>
#include <string>
#include <tuple>
#include <vector>

void rt_compute_local_rows(rt_ctx *c) {
    const uint32_t H = c->hs.window.height;
    const uint32_t tiles = (H + c->tile_rows - 1) / c->tile_rows;
    const uint32_t local_tiles = c->shard_rank < tiles ? (tiles - c->shard_rank + c->shard_world - 1) / c->shard_world : 0;
    c->local_rows = c->shard_world == 1 ? H : local_tiles * c->tile_rows;
}

SceneD rt_scene_view(rt_ctx *ctx) {
    SceneD sc{};
    sc.tlas_nodes = ctx->tlas.nodes();
    sc.blas_nodes = ctx->blas.nodes();
    sc.instances = ctx->d_instances.get();
    sc.leaf_tris = ctx->blas.leaf_tris();
    sc.leaf_prims = ctx->blas.leaf_prims();
    sc.tri_shade = ctx->d_tri_shade.get();
    sc.spheres = ctx->d_spheres.get();
    sc.quads = ctx->d_quads.get();
    sc.roughs = ctx->d_roughs.get();
    sc.metals = ctx->d_metals.get();
    sc.dielectrics = ctx->d_dielectrics.get();
    sc.tlas_root_ref = ctx->hs.inst_desc.size() > 1 ? 0 : make_leaf_ref(0u, 0u);
    sc.tlas_nodes8 = ctx->tlas.nodes8();
    sc.blas_nodes8 = ctx->blas.nodes8();
    sc.tlas_leaf_inst = ctx->tlas.leaf_prims8();
    sc.leaf_tris8 = ctx->blas.leaf_tris8();
    sc.leaf_prims8 = ctx->blas.leaf_prims8();
    sc.ext_prims = ctx->d_ext.get();
    return sc;
}

namespace {

int fail(rt_ctx *ctx, int code, const std::string &msg) { return rt_fail(ctx, code, msg); }

void compute_local_rows(rt_ctx *c) { rt_compute_local_rows(c); }

void alloc_frame(rt_ctx *c) {
    compute_local_rows(c);
    // planes are allocated for the LARGEST shard (rank 0's), so equal-sized gathers may read every rank's plane
    const uint32_t tiles = (c->hs.window.height + c->tile_rows - 1) / c->tile_rows;
    const uint32_t max_rows = c->shard_world == 1 ? c->hs.window.height : ((tiles + c->shard_world - 1) / c->shard_world) * c->tile_rows;
    const size_t px = (size_t) c->hs.window.width * max_rows;
    c->d_rgba8.reserve(px * 4 + 4);
    c->d_radiance.reserve(px * 3 + 3);
    c->d_rng.reserve(px + 1);
    if (c->flags & RT_FLAG_RECORD_PRIMARY_AOV) { c->d_aov_t.reserve(px + 1); c->d_aov_inst.reserve(px + 1); c->d_aov_prim.reserve(px + 1); }
    c->d_counters.reserve(8);
    c->h_counters.reserve(8);
}

void update_instances_and_tlas(rt_ctx *c) {
    const int n = (int) c->hs.inst_desc.size();
    c->hs.fill_xform_in(c->h_xform_in.get(), c->blas.seg_root_host);
    CUDA_CHECK(cudaMemcpyAsync(c->d_xform_in.get(), c->h_xform_in.get(), (size_t) n * sizeof(InstanceXformIn), cudaMemcpyHostToDevice, c->stream));
    launch_instance_update(c->d_xform_in.get(), n, c->d_instances.get(), c->tlas.prim_lo(), c->tlas.prim_hi(), c->d_wcentroid.get(),
                           c->tlas.bounds(), c->d_tri_verts.get(), c->stream);
    c->tlas.build_tlas(n, tlas_morton_bits((size_t) n), c->stream, c->sm_count);
}

} // namespace

extern "C" {

const char *rt_version(void) { return "rtb200 0.1 (sm_100a)"; }

size_t rt_struct_size(const char *name) {
    if (!name) return 0;
#define RT_SZ(T) if (strcmp(name, #T) == 0) return sizeof(T);
    RT_SZ(rt_vec3) RT_SZ(rt_sphere) RT_SZ(rt_parallelogram) RT_SZ(rt_triangle) RT_SZ(rt_rough) RT_SZ(rt_metal) RT_SZ(rt_dielectric)
    RT_SZ(rt_instance_desc) RT_SZ(rt_instance_xform) RT_SZ(rt_window) RT_SZ(rt_camera_input) RT_SZ(rt_frame_out) RT_SZ(rt_build_stats) RT_SZ(rt_animation_stats) RT_SZ(rt_camera_motion) RT_SZ(rt_window_stats) RT_SZ(rt_loop_stats)
#undef RT_SZ
    return 0;
}

int rt_create(rt_ctx **out, int cuda_device, uint32_t flags) {
    if (!out) return RT_ERR_ARGUMENT;
    *out = nullptr;
    int count = 0;
    if (cudaGetDeviceCount(&count) != cudaSuccess || cuda_device < 0 || cuda_device >= count) { cudaGetLastError(); return RT_ERR_CUDA; }
    if (cudaSetDevice(cuda_device) != cudaSuccess) return RT_ERR_CUDA;
    rt_ctx *c = new rt_ctx();
    c->device = cuda_device;
    c->flags = flags;
    c->hs.flags = flags;
    try {
        CUDA_CHECK(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
        for (auto &e: c->ev) CUDA_CHECK(cudaEventCreate(&e));
        CUDA_CHECK(cudaEventCreateWithFlags(&c->ev_block, cudaEventBlockingSync | cudaEventDisableTiming));
        CUDA_CHECK(cudaDeviceGetAttribute(&c->sm_count, cudaDevAttrMultiProcessorCount, cuda_device));
    } catch (const std::exception &) {
        delete c;
        return RT_ERR_CUDA;
    }
    if (const char *e = getenv("RTB200_SLOTS")) c->wave.slots_override = (uint32_t) strtoul(e, nullptr, 0);
    if (const char *e = getenv("RTB200_REFILL_MIN")) c->wave.refill_min = atoi(e);
    if (const char *e = getenv("RTB200_NODE_MIN")) c->wave.node_min = atoi(e);
    if (const char *e = getenv("RTB200_TRACE_BLOCKS_PER_SM")) c->wave.blocks_per_sm_override = atoi(e);
    if (const char *e = getenv("RTB200_GROUPS")) c->wave.n_groups = atoi(e);
    if (const char *e = getenv("RTB200_GRAPHS")) c->wave.use_graphs = atoi(e) != 0;
    if (const char *e = getenv("RTB200_FINISH_DIVISOR")) c->wave.finish_divisor = atoi(e);
    // Vote thresholds of the path kernel, re-tuned under frames in flight (profiles/r02s_sweep.txt, r02t_sweep.txt): a node step
    // while >= 18 lanes stand on a node, a shade step once >= 10 lanes wait (round 1, one frame at a time: 20 / 8).
    // C2 1903 -> 1928 Mrays/s, 1/8 shard 21.51 -> 21.32 ms, C5 unchanged.  The wave comparison mode keeps its own 20.
    c->persist.node_min = getenv("RTB200_NODE_MIN") ? c->wave.node_min : 18;
    c->persist.shade_min = 10;
    c->persist.blocks_per_sm_override = c->wave.blocks_per_sm_override;
    if (const char *e = getenv("RTB200_SHADE_MIN")) c->persist.shade_min = atoi(e);
    if (const char *e = getenv("RTB200_TILE_ORDER")) c->persist.tile_order = atoi(e);
    if (const char *e = getenv("RTB200_SPLIT")) c->persist.split = atoi(e) != 0;       // A/B: dedicated shade warp per CTA (kernels_split.cu)
    if (const char *e = getenv("RTB200_VTK_EXPAND")) c->vtk_expand_on_device = strcmp(e, "host") != 0;
    for (auto &st: c->vtk_stage) st.expand_on_device = c->vtk_expand_on_device;
    if (flags & RT_FLAG_WAVEFRONT) c->path_mode = 0;
    if (const char *e = getenv("RTB200_MODE")) c->path_mode = strcmp(e, "wave") == 0 ? 0 : 1;     // tuning / A-B runs
    *out = c;
    return RT_OK;
}

int rt_cleanup(rt_ctx *ctx) {
    if (!ctx) return RT_ERR_ARGUMENT;
    cudaSetDevice(ctx->device);
    for (rt_ctx *l: ctx->lane_ctx) rt_cleanup(l);
    ctx->lane_ctx.clear();
    for (auto &st: ctx->vtk_stage) st.join();
    if (ctx->stream) { cudaStreamSynchronize(ctx->stream); }
    rt_release_frames(ctx);
    for (auto &e: ctx->ev) if (e) cudaEventDestroy(e);
    if (ctx->ev_block) cudaEventDestroy(ctx->ev_block);
    cudaStream_t s = ctx->stream;
    delete ctx;   // frees device buffers
    if (s) cudaStreamDestroy(s);
    return RT_OK;
}

const char *rt_last_error(const rt_ctx *ctx) { return ctx ? ctx->err.c_str() : "null context"; }

int rt_commit_geometry(rt_ctx *ctx, const rt_sphere *spheres, size_t sphere_count, const rt_parallelogram *parallelograms,
                       size_t parallelogram_count, const rt_triangle *triangles, size_t triangle_count) {
    RT_API_BEGIN(ctx)
    for (int q = 0; q < 2; q++) if (rt_window_in_flight(ctx, q)) return fail(ctx, RT_ERR_STATE, "rt_commit_geometry: not while frames are in flight (rt_release_window first): the window's kernel is reading the scene");
    // the vectors are about to grow: their old storage must not stay registered when it is freed
    if (triangle_count * 9 > ctx->hs.tri_verts.capacity()) ctx->pin_verts.release();
    if (triangle_count > ctx->hs.tri_shade.capacity()) ctx->pin_shade.release();
    const std::string e = ctx->hs.commit_geometry(spheres, sphere_count, parallelograms, parallelogram_count, triangles, triangle_count);
    if (!e.empty()) return fail(ctx, RT_ERR_ARGUMENT, e);
    if (ctx->composing) {        // animation frames: the same vectors are refilled every frame
        ctx->pin_verts.ensure(ctx->hs.tri_verts.data(), ctx->hs.tri_verts.capacity() * sizeof(ctx->hs.tri_verts[0]));
        ctx->pin_shade.ensure(ctx->hs.tri_shade.data(), ctx->hs.tri_shade.capacity() * sizeof(ctx->hs.tri_shade[0]));
    }
    if (!ctx->composing) ctx->user_tris.assign(triangles, triangles + triangle_count);
    ctx->built = false;
    ctx->d_spheres.upload(ctx->hs.sphere_rec.data(), ctx->hs.sphere_rec.size(), ctx->stream);
    ctx->d_quads.upload(ctx->hs.quad_rec.data(), ctx->hs.quad_rec.size(), ctx->stream);
    ctx->d_tri_verts.upload(ctx->hs.tri_verts.data(), ctx->hs.tri_verts.size(), ctx->stream);
    ctx->d_tri_shade.upload(ctx->hs.tri_shade.data(), ctx->hs.tri_shade.size(), ctx->stream);
    ctx->wait();
    ctx->geometry_committed = true;
    RT_API_END(ctx)
}

int rt_commit_materials(rt_ctx *ctx, const rt_rough *roughs, size_t rough_count, const rt_metal *metals, size_t metal_count) {
    RT_API_BEGIN(ctx)
    for (int q = 0; q < 2; q++) if (rt_window_in_flight(ctx, q)) return fail(ctx, RT_ERR_STATE, "rt_commit_materials: not while frames are in flight (rt_release_window first): the window's kernel is reading the scene");
    const std::string e = ctx->hs.commit_materials(roughs, rough_count, metals, metal_count);
    if (!e.empty()) return fail(ctx, RT_ERR_ARGUMENT, e);
    ctx->d_roughs.upload(ctx->hs.roughs.data(), ctx->hs.roughs.size(), ctx->stream);
    ctx->d_metals.upload(ctx->hs.metals.data(), ctx->hs.metals.size(), ctx->stream);
    CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
    ctx->materials_committed = true;
    RT_API_END(ctx)
}

int rt_commit_dielectrics(rt_ctx *ctx, const rt_dielectric *d, size_t count) {
    RT_API_BEGIN(ctx)
    for (int q = 0; q < 2; q++) if (rt_window_in_flight(ctx, q)) return fail(ctx, RT_ERR_STATE, "rt_commit_dielectrics: not while frames are in flight (rt_release_window first): the window's kernel is reading the scene");
    const std::string e = ctx->hs.commit_dielectrics(d, count);
    if (!e.empty()) return fail(ctx, RT_ERR_ARGUMENT, e);
    ctx->d_dielectrics.upload(ctx->hs.dielectrics.data(), ctx->hs.dielectrics.size(), ctx->stream);
    CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
    RT_API_END(ctx)
}

int rt_commit_extension_primitives(rt_ctx *ctx, const void *records, size_t record_bytes, size_t count) {
    RT_API_BEGIN(ctx)
    for (int q = 0; q < 2; q++) if (rt_window_in_flight(ctx, q)) return fail(ctx, RT_ERR_STATE, "rt_commit_extension_primitives: not while frames are in flight (rt_release_window first): the window's kernel is reading the scene");
    const std::string e = ctx->hs.commit_extension(records, record_bytes, count);
    if (!e.empty()) return fail(ctx, RT_ERR_ARGUMENT, e);
    ctx->built = false;
    ctx->d_ext.upload(ctx->hs.ext_rec.data(), ctx->hs.ext_rec.size(), ctx->stream);
    CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
    RT_API_END(ctx)
}

int rt_primitive_type_count(void) { return RT_PRIMITIVE_TYPE_COUNT; }

int rt_configure_instances(rt_ctx *ctx, const rt_instance_desc *instances, size_t instance_count, const rt_instance_xform *initial_xforms,
                           rt_update_cb update, void *user) {
    RT_API_BEGIN(ctx)
    if (!ctx->geometry_committed) return fail(ctx, RT_ERR_STATE, "rt_configure_instances: commit geometry first");
    const std::string e = ctx->hs.configure_instances(instances, instance_count, initial_xforms);
    if (!e.empty()) return fail(ctx, RT_ERR_ARGUMENT, e);
    if (!ctx->composing) {
        ctx->user_inst.assign(instances, instances + instance_count);
        ctx->n_user_inst = instance_count;
        ctx->update_cb = update;
        ctx->update_user = user;
        if (update) update(ctx->hs.xforms.data(), instance_count, 0, user);   // reference src/Global/Renderer.cu:111
    }
    ctx->instances_configured = true;
    ctx->built = false;
    RT_API_END(ctx)
}

int rt_set_instance_xforms(rt_ctx *ctx, const rt_instance_xform *xforms, size_t instance_count) {
    RT_API_BEGIN(ctx)
    if (!xforms || (instance_count != ctx->hs.xforms.size() && instance_count != ctx->n_user_inst)) return fail(ctx, RT_ERR_ARGUMENT, "rt_set_instance_xforms: count mismatch");
    std::copy(xforms, xforms + instance_count, ctx->hs.xforms.begin());
    ctx->tlas_fresh = false;
    RT_API_END(ctx)
}

int rt_build_acceleration_structure(rt_ctx *ctx) {
    RT_API_BEGIN(ctx)
    for (int q = 0; q < 2; q++) if (rt_window_in_flight(ctx, q)) return fail(ctx, RT_ERR_STATE, "rt_build_acceleration_structure: not while frames are in flight (rt_release_window first): the window's kernel is reading the scene");
    if (!ctx->geometry_committed || !ctx->instances_configured) return fail(ctx, RT_ERR_STATE, "rt_build_acceleration_structure: commit geometry and instances first");
    HostScene &hs = ctx->hs;
    const std::string e = hs.plan_blas();
    if (!e.empty()) return fail(ctx, RT_ERR_ARGUMENT, e);
    const size_t n_inst = hs.inst_desc.size();
    const GeometryD geo{ctx->d_tri_verts.get(), ctx->d_spheres.get(), ctx->d_quads.get(), ctx->d_ext.get()};
    const size_t n_prims = hs.seg_offset.back();
    ctx->blas.emit_bvh2 = ctx->tlas.emit_bvh2 = (ctx->flags & (RT_FLAG_TRAVERSE_BVH2 | RT_FLAG_KEEP_LBVH)) != 0;
    // rotated subtrees no longer cover one contiguous Morton range, which multi-primitive leaves (a tuning experiment) rely on
    ctx->blas.sah_rotations = ctx->tlas.sah_rotations = (ctx->flags & RT_FLAG_SAH_REFINE) && !getenv("RTB200_LEAF_MAX");
    {
        const char *e = getenv("RTB200_PLOC");      // A/B runs
        const bool want = e ? atoi(e) != 0 : (ctx->flags & RT_FLAG_SAH_BUILD) != 0;
        ctx->blas.ploc = want && !ctx->composing && !getenv("RTB200_LEAF_MAX");
        ctx->tlas.ploc = false;
    }
    ctx->blas.wait_event = ctx->tlas.wait_event = ctx->blocking_wait ? ctx->ev_block : nullptr;
    ctx->blas.prepare_blas(hs.seg_offset, hs.seg_type);          // allocations stay outside the timed build
    ctx->d_instances.reserve(n_inst);
    ctx->d_xform_in.reserve(n_inst);
    ctx->d_wcentroid.reserve(n_inst);
    ctx->h_xform_in.reserve(n_inst);
    ctx->tlas.prepare_tlas((int) n_inst);
    CUDA_CHECK(cudaEventRecord(ctx->ev[0], ctx->stream));
    ctx->blas.build_blas(geo, hs.seg_offset, hs.seg_type, hs.seg_first, morton_bits_for(n_prims / std::max<size_t>(1, hs.seg_type.size())),
                         ctx->stream, ctx->sm_count);
    CUDA_CHECK(cudaEventRecord(ctx->ev[1], ctx->stream));
    // TLAS (also rebuilt every frame)
    update_instances_and_tlas(ctx);
    CUDA_CHECK(cudaEventRecord(ctx->ev[2], ctx->stream));
    ctx->wait();
    rt_build_stats &st = ctx->build_stats;
    CUDA_CHECK(cudaEventElapsedTime(&st.blas_ms, ctx->ev[0], ctx->ev[1]));
    CUDA_CHECK(cudaEventElapsedTime(&st.tlas_ms, ctx->ev[1], ctx->ev[2]));
    st.blas_count = hs.seg_type.size();
    st.primitive_count = n_prims;
    st.lbvh_nodes = n_prims > 0 ? n_prims - 1 : 0;
    uint32_t wc[4] = {0, 0, 0, 0}, wt[4] = {0, 0, 0, 0};
    CUDA_CHECK(cudaMemcpy(wc, ctx->blas.wide_counters(), sizeof(wc), cudaMemcpyDeviceToHost));
    CUDA_CHECK(cudaMemcpy(wt, ctx->tlas.wide_counters(), sizeof(wt), cudaMemcpyDeviceToHost));
    st.wide_nodes = wc[0];
    st.node_bytes = (uint64_t) (wc[0] + wt[0]) * sizeof(BvhNode8);
    st.prim_bytes = hs.prim_bytes();
    st.kernel_launches = ctx->blas.launches + ctx->tlas.launches + 2;
    ctx->built = true;
    ctx->tlas_fresh = true;
    RT_API_END(ctx)
}

int rt_get_build_stats(const rt_ctx *ctx, rt_build_stats *out) {
    if (!ctx || !out) return RT_ERR_ARGUMENT;
    *out = ctx->build_stats;
    return RT_OK;
}

int rt_configure_camera(rt_ctx *ctx, const rt_window *window, const rt_camera_input *camera) {
    RT_API_BEGIN(ctx)
    for (int q = 0; q < 2; q++) if (rt_window_in_flight(ctx, q)) return fail(ctx, RT_ERR_STATE, "rt_configure_camera: not while frames are in flight (rt_release_window first): the window's kernel is reading the scene");
    if (!window || !camera) return fail(ctx, RT_ERR_ARGUMENT, "rt_configure_camera: null argument");
    const std::string e = ctx->hs.configure_camera(*window, *camera);
    if (!e.empty()) return fail(ctx, RT_ERR_ARGUMENT, e);
    alloc_frame(ctx);
    ctx->camera_set = true;
    RT_API_END(ctx)
}

int rt_camera_step(const rt_camera_input *current, const rt_camera_motion *motion, rt_camera_input *next, int *moved) {
    if (!current || !motion || !next) return RT_ERR_ARGUMENT;
    const bool m = camera_step(*current, *motion, *next);
    if (moved) *moved = m ? 1 : 0;
    return RT_OK;
}

int rt_set_tile_shard(rt_ctx *ctx, uint32_t rank, uint32_t world, uint32_t tile_rows, void *peer_rgba8) {
    RT_API_BEGIN(ctx)
    if (world == 0 || rank >= world || tile_rows == 0) return fail(ctx, RT_ERR_ARGUMENT, "rt_set_tile_shard: bad rank/world/tile_rows");
    if (ctx->ring) return fail(ctx, RT_ERR_STATE, "rt_set_tile_shard: close the frame ring first (it was made for the current shard)");
    for (int q = 0; q < 2; q++) if (rt_window_in_flight(ctx, q)) return fail(ctx, RT_ERR_STATE, "rt_set_tile_shard: not while frames are in flight");
    ctx->shard_rank = rank; ctx->shard_world = world; ctx->tile_rows = tile_rows; ctx->peer_rgba8 = peer_rgba8;
    if (ctx->camera_set) alloc_frame(ctx);
    RT_API_END(ctx)
}

int rt_render_frame(rt_ctx *ctx, uint64_t frame_index, uint64_t seed, rt_frame_out *out) {
    RT_API_BEGIN(ctx)
    if (!ctx->built || !ctx->camera_set || !ctx->materials_committed) return fail(ctx, RT_ERR_STATE, "rt_render_frame: scene not built / camera or materials missing");
    {
        const std::string e = ctx->hs.check_materials();
        if (!e.empty()) return fail(ctx, RT_ERR_ARGUMENT, "rt_render_frame: " + e);
    }
    cudaStream_t s = ctx->stream;
    if (ctx->update_cb) ctx->update_cb(ctx->hs.xforms.data(), ctx->n_user_inst, (size_t) frame_index, ctx->update_user);
    CUDA_CHECK(cudaEventRecord(ctx->ev[0], s));
    const bool reuse_tlas = ctx->tlas_fresh && !ctx->update_cb;        // built a moment ago for these very transforms (api_ctx.h)
    ctx->tlas_fresh = false;
    if (!reuse_tlas) update_instances_and_tlas(ctx);
    CUDA_CHECK(cudaEventRecord(ctx->ev[1], s));

    const SceneD sc = rt_scene_view(ctx);

    FrameD fr{};
    fr.rgba8 = ctx->peer_rgba8 ? (uint8_t *) ctx->peer_rgba8 : ctx->d_rgba8.get();
    fr.rgba8_is_full_frame = ctx->peer_rgba8 ? 1u : 0u;
    fr.radiance = ctx->d_radiance.get();
    const bool aov = ctx->flags & RT_FLAG_RECORD_PRIMARY_AOV;
    fr.aov_t = aov ? ctx->d_aov_t.get() : nullptr;
    fr.aov_inst = aov ? ctx->d_aov_inst.get() : nullptr;
    fr.aov_prim = aov ? ctx->d_aov_prim.get() : nullptr;
    fr.rng = ctx->d_rng.get();
    fr.ray_counter = ctx->d_counters.get();
    fr.overflow_flag = reinterpret_cast<unsigned int *>(ctx->d_counters.get() + 1);
    fr.count_traversal = (ctx->flags & RT_FLAG_COUNT_TRAVERSAL) ? 1u : 0u;
    fr.traverse_bvh2 = (ctx->flags & RT_FLAG_TRAVERSE_BVH2) ? 1u : 0u;
    fr.shard_rank = ctx->shard_rank; fr.shard_world = ctx->shard_world; fr.tile_rows = ctx->tile_rows; fr.local_rows = ctx->local_rows;

    CUDA_CHECK(cudaMemsetAsync(ctx->d_counters.get(), 0, 8 * sizeof(unsigned long long), s));
    ctx->d_rng_tables.reserve(RNG_TABLE_WORDS);
    launch_rng_init(ctx->d_rng.get(), seed, ctx->hs.cam, fr, s, ctx->d_rng_tables.get(), ctx->rng_tables_ready);
    CUDA_CHECK(cudaEventRecord(ctx->ev[2], s));
    WaveStats ws{};
    // ray_trace_depth == 0: rayColor's loop body never runs (reference src/Global/Kernel.cu:64), every sample returns (1,1,1)
    // without tracing; the one-thread-per-pixel kernel has that loop shape, the queue-driven kernels always trace one ray
    const bool mega = (ctx->flags & RT_FLAG_MEGAKERNEL) || ctx->hs.cam.depth == 0;
    const bool persist = !mega && ctx->path_mode == 1 && !fr.traverse_bvh2;
    if (mega) {
        launch_render(sc, ctx->hs.cam, fr, s);
    } else if (persist) {
        ctx->persist.n_blas_nodes = (uint32_t) ctx->blas.n_prims();      // <= allocated records of the node pool
        ctx->persist.single_instance = rt_single_instance(ctx);
        ctx->persist.render(sc, ctx->hs.cam, fr, s, ctx->sm_count);
        ws.launches = ctx->persist.launches;
    } else {
        ctx->wave.mode = fr.traverse_bvh2 ? 0 : 1;
        ctx->wave.render(sc, ctx->hs.cam, fr, s, ctx->sm_count, ws);
    }
    CUDA_CHECK(cudaEventRecord(ctx->ev[3], s));
    CUDA_CHECK(cudaMemcpyAsync(ctx->h_counters.get(), ctx->d_counters.get(), 8 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, s));
    ctx->wait();
    if (out) {
        memset(out, 0, sizeof(*out));
        out->rays = ws.rays + ctx->h_counters.get()[0];      // rays of the wave queues (host-counted) + rays counted on the device
        out->waves = ws.waves;
        out->stack_overflow = (uint32_t) ctx->h_counters.get()[1];
        out->node_fetches = ctx->h_counters.get()[2];
        out->prim_fetches = ctx->h_counters.get()[3];
        out->instance_fetches = ctx->h_counters.get()[4];
        uint64_t rows = 0;
        for (uint32_t ly = 0; ly < ctx->local_rows; ly++) {
            const uint32_t gy = ((ly / ctx->tile_rows) * ctx->shard_world + ctx->shard_rank) * ctx->tile_rows + (ly % ctx->tile_rows);
            if (gy < ctx->hs.cam.height) rows++;
        }
        out->primary_rays = rows * ctx->hs.cam.width * ctx->hs.cam.sqrtSamples * ctx->hs.cam.sqrtSamples;
        CUDA_CHECK(cudaEventElapsedTime(&out->render_ms, ctx->ev[0], ctx->ev[3]));
        CUDA_CHECK(cudaEventElapsedTime(&out->tlas_ms, ctx->ev[0], ctx->ev[1]));
        CUDA_CHECK(cudaEventElapsedTime(&out->rng_ms, ctx->ev[1], ctx->ev[2]));
        CUDA_CHECK(cudaEventElapsedTime(&out->trace_ms, ctx->ev[2], ctx->ev[3]));
        out->kernel_launches = (reuse_tlas ? 0 : ctx->tlas.launches + 3 /*instance update: bounds init, matrices, world boxes*/) + 1 /*rng*/ + (mega ? 1 : ws.launches);
    }
    // a traversal stack that overflowed dropped subtrees: the frame is wrong, and the caller must hear about it (the
    // reference's 64-entry stacks overflow silently, Q15); `out` is filled all the same
    if (ctx->h_counters.get()[1] != 0) return fail(ctx, RT_ERR_OVERFLOW, "rt_render_frame: a traversal stack overflowed (BVH deeper than the per-thread stack): the frame is incomplete");
    RT_API_END(ctx)
}

static int plane_info(rt_ctx *ctx, int kind, void **ptr, size_t *bytes) {
    const size_t px = (size_t) ctx->hs.window.width * ctx->local_rows;
    switch (kind) {
        case RT_BUFFER_RGBA8:
            // with a frame ring the texels of every rank are stored into rank 0's full frame (slot 0 of window 0 for rt_render_frame)
            if (ctx->ring) { *ptr = ctx->ring->frame(0, 0); *bytes = ctx->ring->frame_bytes; return RT_OK; }
            *ptr = ctx->d_rgba8.get(); *bytes = px * 4; return RT_OK;
        case RT_BUFFER_RADIANCE: *ptr = ctx->d_radiance.get(); *bytes = px * 12; return RT_OK;
        case RT_BUFFER_PRIMARY_T: *ptr = ctx->d_aov_t.get(); *bytes = px * 4; break;
        case RT_BUFFER_PRIMARY_INSTANCE: *ptr = ctx->d_aov_inst.get(); *bytes = px * 4; break;
        case RT_BUFFER_PRIMARY_PRIM: *ptr = ctx->d_aov_prim.get(); *bytes = px * 4; break;
        default: return fail(ctx, RT_ERR_ARGUMENT, "unknown buffer kind");
    }
    if (!(ctx->flags & RT_FLAG_RECORD_PRIMARY_AOV)) return fail(ctx, RT_ERR_STATE, "primary AOVs need RT_FLAG_RECORD_PRIMARY_AOV");
    return RT_OK;
}

int rt_framebuffer_device_ptr(rt_ctx *ctx, int kind, void **dev_ptr, size_t *bytes) {
    RT_API_BEGIN(ctx)
    if (!ctx->camera_set) return fail(ctx, RT_ERR_STATE, "configure the camera first");
    void *p = nullptr;
    size_t n = 0;
    const int rc = plane_info(ctx, kind, &p, &n);
    if (rc != RT_OK) return rc;
    if (dev_ptr) *dev_ptr = p;
    if (bytes) *bytes = n;
    RT_API_END(ctx)
}

int rt_read_framebuffer(rt_ctx *ctx, int kind, void *host_dst, size_t dst_bytes) {
    RT_API_BEGIN(ctx)
    if (!ctx->camera_set) return fail(ctx, RT_ERR_STATE, "configure the camera first");
    void *p = nullptr;
    size_t n = 0;
    const int rc = plane_info(ctx, kind, &p, &n);
    if (rc != RT_OK) return rc;
    if (!host_dst || dst_bytes < n) return fail(ctx, RT_ERR_ARGUMENT, "rt_read_framebuffer: destination too small");
    CUDA_CHECK(cudaMemcpyAsync(host_dst, p, n, cudaMemcpyDeviceToHost, ctx->stream));
    ctx->wait();
    RT_API_END(ctx)
}

int rt_probe_read_bandwidth(rt_ctx *ctx, size_t bytes, int repeats, double *gbs) {
    RT_API_BEGIN(ctx)
    if (!gbs || bytes < 4096 || repeats < 1) return fail(ctx, RT_ERR_ARGUMENT, "rt_probe_read_bandwidth: bad argument");
    *gbs = probe_read_bandwidth(bytes, repeats, ctx->sm_count, ctx->stream);
    RT_API_END(ctx)
}

int rt_write_ppm(rt_ctx *ctx, const char *path) {
    RT_API_BEGIN(ctx)
    if (!ctx->camera_set || !path) return fail(ctx, RT_ERR_STATE, "rt_write_ppm: no frame");
    // the plane rt_read_framebuffer returns: this context's rows, or — with a frame ring — the assembled full frame on rank 0
    void *src = nullptr;
    size_t n_bytes = 0;
    const int rc = plane_info(ctx, RT_BUFFER_RGBA8, &src, &n_bytes);
    if (rc != RT_OK) return rc;
    const uint32_t W = ctx->hs.window.width, H = (uint32_t) (n_bytes / ((size_t) W * 4));
    std::vector<uint8_t> img((size_t) W * H * 4);
    CUDA_CHECK(cudaMemcpyAsync(img.data(), src, img.size(), cudaMemcpyDeviceToHost, ctx->stream));
    CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
    FILE *f = fopen(path, "wb");
    if (!f) return fail(ctx, RT_ERR_IO, std::string("cannot open ") + path);
    const size_t len = strlen(path);
    if (len >= 4 && strcmp(path + len - 4, ".png") == 0) {           // "*.png": 8-bit RGBA PNG, top-down (image_io.h)
        std::vector<uint8_t> png;
        png_encode_rgba8(img.data(), W, H, true, png);
        const bool ok = fwrite(png.data(), 1, png.size(), f) == png.size();
        fclose(f);
        if (!ok) return fail(ctx, RT_ERR_IO, std::string("short write to ") + path);
        return RT_OK;
    }
    fprintf(f, "P6\n%u %u\n255\n", W, H);
    std::vector<uint8_t> row((size_t) W * 3);
    bool ok = true;
    for (uint32_t y = 0; y < H; y++) {   // row 0 of the framebuffer is the bottom scan-line: flip
        const uint8_t *line = &img[(size_t) (H - 1 - y) * W * 4];
        for (uint32_t x = 0; x < W; x++) { row[3 * x] = line[4 * x]; row[3 * x + 1] = line[4 * x + 1]; row[3 * x + 2] = line[4 * x + 2]; }
        ok = ok && fwrite(row.data(), 1, row.size(), f) == row.size();
    }
    ok = (fclose(f) == 0) && ok;
    if (!ok) return fail(ctx, RT_ERR_IO, std::string("short write to ") + path);
    RT_API_END(ctx)
}

int rt_write_image(rt_ctx *ctx, int kind, const char *path) {
    if (kind == RT_BUFFER_RGBA8) return rt_write_ppm(ctx, path);
    RT_API_BEGIN(ctx)
    if (!ctx->camera_set || !path) return fail(ctx, RT_ERR_STATE, "rt_write_image: no frame");
    void *p = nullptr;
    size_t n = 0;
    const int rc = plane_info(ctx, kind, &p, &n);
    if (rc != RT_OK) return rc;
    std::vector<uint8_t> buf(n);
    CUDA_CHECK(cudaMemcpyAsync(buf.data(), p, n, cudaMemcpyDeviceToHost, ctx->stream));
    CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
    FILE *f = fopen(path, "wb");
    if (!f) return fail(ctx, RT_ERR_IO, std::string("cannot open ") + path);
    const uint32_t W = ctx->hs.window.width, H = ctx->local_rows;
    const size_t len = strlen(path);
    if (len >= 4 && strcmp(path + len - 4, ".exr") == 0 && (kind == RT_BUFFER_RADIANCE || kind == RT_BUFFER_PRIMARY_T)) {
        // "*.exr": uncompressed 32-bit float OpenEXR, top-down (image_io.h); radiance -> R, G, B; primary t -> Z
        std::vector<uint8_t> exr;
        exr_encode_float(reinterpret_cast<const float *>(buf.data()), W, H, kind == RT_BUFFER_RADIANCE ? 3 : 1, true, exr);
        const bool ok_exr = fwrite(exr.data(), 1, exr.size(), f) == exr.size();
        fclose(f);
        if (!ok_exr) return fail(ctx, RT_ERR_IO, std::string("short write to ") + path);
        return RT_OK;
    }
    if (kind == RT_BUFFER_RADIANCE) fprintf(f, "PF\n%u %u\n-1.0\n", W, H);
    else if (kind == RT_BUFFER_PRIMARY_T) fprintf(f, "Pf\n%u %u\n-1.0\n", W, H);
    else fprintf(f, "RTAOV1 %u %u %d\n", W, H, kind);
    const bool ok = fwrite(buf.data(), 1, n, f) == n;
    fclose(f);
    if (!ok) return fail(ctx, RT_ERR_IO, std::string("short write to ") + path);
    RT_API_END(ctx)
}

int rt_read_lbvh(rt_ctx *ctx, size_t which, size_t *n_prims, uint64_t *sorted_keys, uint32_t *sorted_prim_ids, int32_t *left, int32_t *right,
                 int32_t *parent, float *node_bounds) {
    RT_API_BEGIN(ctx)
    if (!ctx->built) return fail(ctx, RT_ERR_STATE, "rt_read_lbvh: build first");
    if (which > 1) return fail(ctx, RT_ERR_ARGUMENT, "rt_read_lbvh: which = 0 (BLAS batch) or 1 (TLAS)");
    LbvhBuilder &b = which == 0 ? ctx->blas : ctx->tlas;
    const size_t n = (size_t) b.n_prims();
    if (n_prims) *n_prims = n;
    const size_t m = n > 1 ? n - 1 : 0;
    cudaStream_t s = ctx->stream;
    if (sorted_keys) CUDA_CHECK(cudaMemcpyAsync(sorted_keys, b.sorted_keys(), n * 8, cudaMemcpyDeviceToHost, s));
    if (sorted_prim_ids) CUDA_CHECK(cudaMemcpyAsync(sorted_prim_ids, b.sorted_vals(), n * 4, cudaMemcpyDeviceToHost, s));
    if (left && m) CUDA_CHECK(cudaMemcpyAsync(left, b.left(), m * 4, cudaMemcpyDeviceToHost, s));
    if (right && m) CUDA_CHECK(cudaMemcpyAsync(right, b.right(), m * 4, cudaMemcpyDeviceToHost, s));
    if (parent && m) CUDA_CHECK(cudaMemcpyAsync(parent, b.parent(), m * 4, cudaMemcpyDeviceToHost, s));
    if (node_bounds && m) {
        std::vector<float4> lo(m), hi(m);
        CUDA_CHECK(cudaMemcpyAsync(lo.data(), b.node_lo(), m * 16, cudaMemcpyDeviceToHost, s));
        CUDA_CHECK(cudaMemcpyAsync(hi.data(), b.node_hi(), m * 16, cudaMemcpyDeviceToHost, s));
        CUDA_CHECK(cudaStreamSynchronize(s));
        for (size_t i = 0; i < m; i++) {
            node_bounds[6 * i] = lo[i].x; node_bounds[6 * i + 1] = lo[i].y; node_bounds[6 * i + 2] = lo[i].z;
            node_bounds[6 * i + 3] = hi[i].x; node_bounds[6 * i + 4] = hi[i].y; node_bounds[6 * i + 5] = hi[i].z;
        }
    }
    CUDA_CHECK(cudaStreamSynchronize(s));
    RT_API_END(ctx)
}

int rt_read_wide_bvh(rt_ctx *ctx, size_t which, size_t *n_nodes, size_t *n_leaf_prims, void *nodes, uint32_t *leaf_prims) {
    RT_API_BEGIN(ctx)
    if (!ctx->built) return fail(ctx, RT_ERR_STATE, "rt_read_wide_bvh: build first");
    if (which > 1) return fail(ctx, RT_ERR_ARGUMENT, "rt_read_wide_bvh: which = 0 (BLAS batch) or 1 (TLAS)");
    LbvhBuilder &b = which == 0 ? ctx->blas : ctx->tlas;
    uint32_t wc[4];
    CUDA_CHECK(cudaMemcpyAsync(wc, b.wide_counters(), sizeof(wc), cudaMemcpyDeviceToHost, ctx->stream));
    CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
    if (n_nodes) *n_nodes = wc[0];
    if (n_leaf_prims) *n_leaf_prims = (size_t) b.n_prims();
    if (nodes) CUDA_CHECK(cudaMemcpyAsync(nodes, b.nodes8(), (size_t) wc[0] * sizeof(BvhNode8), cudaMemcpyDeviceToHost, ctx->stream));
    if (leaf_prims) CUDA_CHECK(cudaMemcpyAsync(leaf_prims, b.leaf_prims8(), (size_t) b.n_prims() * 4, cudaMemcpyDeviceToHost, ctx->stream));
    CUDA_CHECK(cudaStreamSynchronize(ctx->stream));
    RT_API_END(ctx)
}

int rt_load_vtk_series(rt_ctx *ctx, const char *series_path, size_t *frame_count) {
    RT_API_BEGIN(ctx)
    if (!series_path) return fail(ctx, RT_ERR_ARGUMENT, "rt_load_vtk_series: null path");
    auto series = std::make_shared<VtkSeries>();
    std::string err;
    if (!series->load(series_path, err)) return fail(ctx, RT_ERR_IO, err);
    if (frame_count) *frame_count = series->files.size();
    ctx->vtk = std::move(series);
    RT_API_END(ctx)
}

int rt_set_vtk_frame(rt_ctx *ctx, size_t frame) {
    RT_API_BEGIN(ctx)
    if (!ctx->vtk) return fail(ctx, RT_ERR_STATE, "rt_set_vtk_frame: load a series first");
    if (!ctx->geometry_committed || !ctx->instances_configured) return fail(ctx, RT_ERR_STATE, "rt_set_vtk_frame: commit geometry and instances first");
    if (frame >= ctx->vtk->files.size()) return fail(ctx, RT_ERR_ARGUMENT, "rt_set_vtk_frame: frame out of range");
    for (int q = 0; q < 2; q++) if (rt_window_in_flight(ctx, q)) return fail(ctx, RT_ERR_STATE, "rt_set_vtk_frame: not while frames are in flight (rt_release_window first): the window's kernel is reading the scene");
    // the frame may have been read ahead by rt_prefetch_vtk_frame; otherwise read it now
    int cur = -1;
    if (ctx->vtk_ahead >= 0) {
        ctx->vtk_stage[ctx->vtk_ahead].join();
        if (ctx->vtk_stage[ctx->vtk_ahead].frame == frame) cur = ctx->vtk_ahead;
        ctx->vtk_ahead = -1;
    }
    if (cur < 0) { cur = 0; ctx->vtk_stage[0].load(*ctx->vtk, frame); }
    rt_ctx::VtkStage &st = ctx->vtk_stage[cur];
    if (!st.ok) return fail(ctx, RT_ERR_IO, st.err);
    ctx->vtk_parse_ms = st.parse_ms;
    std::vector<rt_triangle> &tris = st.tris;
    const std::vector<rt_instance_desc> &inst = st.inst;
    const std::vector<rt_instance_xform> &xf = st.xf;
    st.frame = (size_t) -1;                                     // consumed (tris gets the user's triangles appended below)
    // compose: VTK triangles at the head of the triangle array, user TRIANGLE instances shifted by their count,
    // VTK instances after the user's (reference src/Global/Renderer.cu:17,106-108,114-116)
    const bool on_device = st.expand_on_device;
    const size_t n_vtk_tris = on_device ? st.compact.n_tris : tris.size();
    if (!on_device) tris.insert(tris.end(), ctx->user_tris.begin(), ctx->user_tris.end());
    std::vector<rt_instance_desc> all_inst(ctx->user_inst);
    for (rt_instance_desc &d: all_inst) if (d.type == RT_TRIANGLE) d.first_prim += n_vtk_tris;
    all_inst.insert(all_inst.end(), inst.begin(), inst.end());
    std::vector<rt_instance_xform> all_xf(ctx->hs.xforms.begin(), ctx->hs.xforms.begin() + ctx->n_user_inst);
    all_xf.insert(all_xf.end(), xf.begin(), xf.end());
    const std::vector<rt_sphere> spheres(ctx->hs.spheres);
    const std::vector<rt_parallelogram> quads(ctx->hs.quads);
    struct ComposeGuard { rt_ctx *c; ~ComposeGuard() { c->composing = false; } } compose_guard{ctx};     // also on a thrown CudaError
    ctx->composing = true;
    int rc = RT_OK;
    if (on_device) {
        // spheres and parallelograms stay as committed; the frame's triangles are expanded from the strips on the device
        // and the caller's own triangles (converted on the host, a handful) follow them
        const VtkCompact &vc = st.compact;
        const size_t n_user = ctx->user_tris.size(), n_all = n_vtk_tris + n_user;
        const std::string e = ctx->hs.commit_triangle_tail(n_vtk_tris, ctx->user_tris.data(), n_user);
        if (!e.empty()) { ctx->composing = false; return fail(ctx, RT_ERR_ARGUMENT, e); }
        cudaStream_t s = ctx->stream;
        ctx->d_tri_verts.reserve(n_all * 9);
        ctx->d_tri_shade.reserve(n_all);
        ctx->d_vtk_points.upload(&vc.points.data()->x, vc.points.size() * 3, s);
        ctx->d_vtk_normals.upload(&vc.normals.data()->x, vc.normals.size() * 3, s);
        ctx->d_vtk_ids.upload(vc.ids.data(), vc.ids.size(), s);
        ctx->d_vtk_first_tri.upload(vc.strip_first_tri.data(), vc.strip_first_tri.size(), s);
        ctx->d_vtk_first_id.upload(vc.strip_first_id.data(), vc.strip_first_id.size(), s);
        ctx->hs.note_material(RT_METAL, 0u);                    // every particle triangle is METAL #0 (reference src/Global/VTKReader.cu:178)
        launch_expand_strips(ctx->d_vtk_points.get(), ctx->d_vtk_normals.get(), ctx->d_vtk_ids.get(), ctx->d_vtk_first_tri.get(), ctx->d_vtk_first_id.get(),
                             (int) vc.strip_first_id.size(), (uint32_t) n_vtk_tris, RT_METAL, 0u, ctx->d_tri_verts.get(), ctx->d_tri_shade.get(), s);
        if (n_user) {
            CUDA_CHECK(cudaMemcpyAsync(ctx->d_tri_verts.get() + n_vtk_tris * 9, ctx->hs.tri_verts.data() + n_vtk_tris * 9, n_user * 9 * sizeof(float), cudaMemcpyHostToDevice, s));
            CUDA_CHECK(cudaMemcpyAsync(ctx->d_tri_shade.get() + n_vtk_tris, ctx->hs.tri_shade.data() + n_vtk_tris, n_user * sizeof(TriShade), cudaMemcpyHostToDevice, s));
        }
        ctx->built = false;          // the build that follows on the same stream is ordered after the expansion
    } else {
        rc = rt_commit_geometry(ctx, spheres.data(), spheres.size(), quads.data(), quads.size(), tris.data(), tris.size());
    }
    if (rc == RT_OK) rc = rt_configure_instances(ctx, all_inst.data(), all_inst.size(), all_xf.data(), nullptr, nullptr);
    if (rc == RT_OK) rc = rt_build_acceleration_structure(ctx);
    ctx->composing = false;
    if (rc != RT_OK) return rc;
    if (on_device) ctx->build_stats.kernel_launches += 1;       // expand_strips_kernel
    RT_API_END(ctx)
}

int rt_prefetch_vtk_frame(rt_ctx *ctx, size_t frame) {
    RT_API_BEGIN(ctx)
    if (!ctx->vtk) return fail(ctx, RT_ERR_STATE, "rt_prefetch_vtk_frame: load a series first");
    if (frame >= ctx->vtk->files.size()) return fail(ctx, RT_ERR_ARGUMENT, "rt_prefetch_vtk_frame: frame out of range");
    if (ctx->vtk_ahead >= 0) ctx->vtk_stage[ctx->vtk_ahead].join();       // one read-ahead at a time
    // stage 0 is the one rt_set_vtk_frame reads into when nothing was prefetched; read ahead into stage 1
    rt_ctx::VtkStage *st = &ctx->vtk_stage[1];
    const VtkSeries *series = ctx->vtk.get();
    st->worker = std::thread([st, series, frame]() { st->load(*series, frame); });
    ctx->vtk_ahead = 1;
    RT_API_END(ctx)
}

} // extern "C"

// ---- the animation loop behind the ABI ---------------------------------------------------------------------------
namespace {

// a lane follows the caller's context: same committed scene (the triangles arrive with the lane's own
// rt_set_vtk_frame), camera and shard; own stream, device buffers, read-ahead stages
int sync_lane(rt_ctx *src, rt_ctx *c) {
    HostScene &h = c->hs;
    const HostScene &g = src->hs;
    h.flags = g.flags;
    h.spheres = g.spheres; h.quads = g.quads; h.sphere_rec = g.sphere_rec; h.quad_rec = g.quad_rec;
    h.roughs = g.roughs; h.metals = g.metals; h.dielectrics = g.dielectrics;
    for (int k = 0; k < 3; k++) h.max_mat_index[k] = g.max_mat_index[k];
    h.xforms.assign(g.xforms.begin(), g.xforms.begin() + src->n_user_inst);
    h.window = g.window; h.cam = g.cam;
    c->user_tris = src->user_tris;
    c->user_inst = src->user_inst;
    c->n_user_inst = src->n_user_inst;
    c->update_cb = src->update_cb;
    c->update_user = src->update_user;
    c->vtk = src->vtk;
    c->path_mode = src->path_mode;
    c->shard_rank = src->shard_rank; c->shard_world = src->shard_world; c->tile_rows = src->tile_rows; c->peer_rgba8 = nullptr;
    try {
        c->d_spheres.upload(h.sphere_rec.data(), h.sphere_rec.size(), c->stream);
        c->d_quads.upload(h.quad_rec.data(), h.quad_rec.size(), c->stream);
        c->d_roughs.upload(h.roughs.data(), h.roughs.size(), c->stream);
        c->d_metals.upload(h.metals.data(), h.metals.size(), c->stream);
        c->d_dielectrics.upload(h.dielectrics.data(), h.dielectrics.size(), c->stream);
        CUDA_CHECK(cudaStreamSynchronize(c->stream));
        c->geometry_committed = src->geometry_committed;
        c->materials_committed = src->materials_committed;
        c->instances_configured = src->instances_configured;
        c->camera_set = src->camera_set;
        c->built = false;
        if (c->camera_set) alloc_frame(c);
    } catch (const std::exception &e) {
        src->err = e.what();
        cudaGetLastError();
        return RT_ERR_CUDA;
    }
    return RT_OK;
}

struct AnimationShared {
    std::mutex mu;                       // sink calls and the totals
    rt_animation_stats tot{};
    std::atomic<int> stop{0};            // first error code / sink stop
    std::string err;
    rt_frame_sink sink = nullptr;
    void *user = nullptr;
};

void run_lane(rt_ctx *c, const std::vector<size_t> &frames, uint64_t seed, AnimationShared *sh) {
    if (frames.empty()) return;
    auto bail = [&](int rc, const std::string &msg) {
        std::lock_guard<std::mutex> g(sh->mu);
        int expected = 0;
        if (sh->stop.compare_exchange_strong(expected, rc)) sh->err = msg;
    };
    if (cudaSetDevice(c->device) != cudaSuccess) { bail(RT_ERR_CUDA, "cudaSetDevice failed"); return; }
    size_t plane = 0;
    PinnedBuffer<uint8_t> &host = c->h_frame;
    using clk = std::chrono::steady_clock;
    auto ms_since = [](clk::time_point t) { return std::chrono::duration<double, std::milli>(clk::now() - t).count(); };
    int rc = rt_prefetch_vtk_frame(c, frames[0]);
    for (size_t j = 0; rc == RT_OK && j < frames.size() && sh->stop.load() == 0; j++) {
        clk::time_point t = clk::now();
        rc = rt_set_vtk_frame(c, frames[j]);       // rebuilds of different lanes may overlap freely (the collapse claims its items by ticket)
        if (rc != RT_OK) break;
        const double set_ms = ms_since(t), parse_ms = c->vtk_parse_ms;
        if (j + 1 < frames.size()) rc = rt_prefetch_vtk_frame(c, frames[j + 1]);     // parsed while this frame renders
        if (rc != RT_OK) break;
        rt_frame_out out{};
        t = clk::now();
        rc = rt_render_frame(c, frames[j], seed, &out);
        if (rc != RT_OK) break;
        const double render_ms = ms_since(t);
        plane = (size_t) c->hs.window.width * c->local_rows * 4;
        try {
            host.reserve(plane + 4);
        } catch (const std::exception &e) { c->err = e.what(); rc = RT_ERR_CUDA; break; }
        t = clk::now();
        rc = rt_read_framebuffer(c, RT_BUFFER_RGBA8, host.get(), plane);
        if (rc != RT_OK) break;
        const double read_ms = ms_since(t);
        t = clk::now();
        std::lock_guard<std::mutex> g(sh->mu);
        rt_animation_stats &tot = sh->tot;
        tot.frames++;
        tot.rays += out.rays;
        tot.triangles += c->build_stats.primitive_count;
        tot.kernel_launches += c->build_stats.kernel_launches + out.kernel_launches;
        const double rebuild = (double) c->build_stats.blas_ms + c->build_stats.tlas_ms;
        tot.device_ms += rebuild + out.render_ms;
        if (rebuild > tot.slowest_rebuild_ms) tot.slowest_rebuild_ms = rebuild;
        tot.parse_ms += parse_ms; tot.set_frame_ms += set_ms; tot.render_wall_ms += render_ms; tot.readback_ms += read_ms;
        if (sh->sink && sh->sink(sh->user, frames[j], host.get(), plane, &out) != 0) {
            int expected = 0;
            if (sh->stop.compare_exchange_strong(expected, RT_ERR_STATE)) sh->err = "rt_run_vtk_animation: stopped by the frame sink";
        }
        tot.sink_ms += ms_since(t);
    }
    for (auto &st: c->vtk_stage) st.join();       // a read-ahead may still be running after an early stop
    c->vtk_ahead = -1;
    if (rc != RT_OK) bail(rc, c->err);
}

} // namespace

extern "C" int rt_run_vtk_animation(rt_ctx *ctx, size_t first_frame, size_t frame_stride, size_t max_frames, uint64_t seed, int lanes,
                                    rt_frame_sink sink, void *user, rt_animation_stats *stats) {
    RT_API_BEGIN(ctx)
    if (!ctx->vtk) return fail(ctx, RT_ERR_STATE, "rt_run_vtk_animation: load a series first");
    if (!ctx->geometry_committed || !ctx->instances_configured || !ctx->camera_set || !ctx->materials_committed)
        return fail(ctx, RT_ERR_STATE, "rt_run_vtk_animation: commit geometry, materials, instances and camera first");
    if (frame_stride == 0 || lanes < 1 || lanes > 8) return fail(ctx, RT_ERR_ARGUMENT, "rt_run_vtk_animation: frame_stride >= 1 and 1 <= lanes <= 8");
    if (ctx->peer_rgba8) return fail(ctx, RT_ERR_STATE, "rt_run_vtk_animation: not with a peer framebuffer");
    std::vector<size_t> frames;
    for (size_t f = first_frame; f < ctx->vtk->files.size() && (max_frames == 0 || frames.size() < max_frames); f += frame_stride) frames.push_back(f);
    const int L = (int) std::max<size_t>(1, std::min<size_t>((size_t) lanes, frames.size()));
    while ((int) ctx->lane_ctx.size() < L - 1) {
        rt_ctx *c = nullptr;
        const int rc = rt_create(&c, ctx->device, ctx->flags);
        if (rc != RT_OK) return fail(ctx, rc, "rt_run_vtk_animation: could not create a lane context");
        ctx->lane_ctx.push_back(c);
    }
    std::vector<rt_ctx *> lane_ctx(1, ctx);
    for (int l = 1; l < L; l++) {
        const int rc = sync_lane(ctx, ctx->lane_ctx[l - 1]);
        if (rc != RT_OK) return rc;
        lane_ctx.push_back(ctx->lane_ctx[l - 1]);
    }
    std::vector<std::vector<size_t>> lane_frames(L);
    for (size_t k = 0; k < frames.size(); k++) lane_frames[k % L].push_back(frames[k]);
    std::vector<bool> was_blocking;
    // RTB200_BLOCKING_WAIT=0/1 overrides; default: sleep when several lanes share the host, spin (lower latency) with one
    const char *bw = getenv("RTB200_BLOCKING_WAIT");
    const bool block = bw ? atoi(bw) != 0 : L > 1;
    for (rt_ctx *c: lane_ctx) { was_blocking.push_back(c->blocking_wait); c->blocking_wait = block; }
    AnimationShared sh;
    sh.sink = sink;
    sh.user = user;
    const auto t0 = std::chrono::steady_clock::now();
    std::vector<std::thread> th;
    for (int l = 1; l < L; l++) th.emplace_back(run_lane, lane_ctx[l], std::cref(lane_frames[l]), seed, &sh);
    run_lane(ctx, lane_frames[0], seed, &sh);
    for (std::thread &t: th) t.join();
    sh.tot.wall_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    for (size_t k = 0; k < lane_ctx.size(); k++) lane_ctx[k]->blocking_wait = was_blocking[k];
    cudaSetDevice(ctx->device);
    if (stats) *stats = sh.tot;
    if (sh.stop.load() != 0) return fail(ctx, sh.stop.load(), sh.err);
    RT_API_END(ctx)
}

