/*
 * rtb200.h — C-ABI of the B200-native per-frame ray-tracing hot path.
 *
 * Drop-in boundary for the all-static C++ class `renderer::Renderer` of
 * 3169651074/real-time-gpu-ray-tracer (reference include/Global/Renderer.cuh:109-146).
 * Each entry point names the reference interface it replaces (file:line, relative to the
 * reference root).  Plain pointers and sizes only; no C++/torch types cross this boundary.
 *
 * Threading: a context is single-threaded (one host thread drives one rt_ctx at a time).
 * Errors: every function returns 0 on success or a negative rt_status; nothing exits or throws
 *         (the reference exit()s: include/Global/Global.cuh:242,253; src/Global/Global.cu:14-42).
 */
#ifndef RTB200_H
#define RTB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ---- status codes (reference exit codes kept where they exist) -------------------------- */
enum rt_status {
    RT_OK            = 0,
    RT_ERR_IO        = -1,   /* reference: exit(-1) on file errors, src/Global/Renderer.cu:399-422 */
    RT_ERR_ARGUMENT  = -2,
    RT_ERR_STATE     = -3,   /* call order violated (e.g. render before build)                    */
    RT_ERR_OVERFLOW  = -4,   /* a traversal stack overflowed: the frame was rendered but is incomplete
                                (the reference's 64-entry stacks overflow silently, quirk Q15)      */
    RT_ERR_CUDA      = -200, /* reference: CUDA_ERROR_EXIT_CODE, include/Global/Global.cuh:253    */
    RT_ERR_NCCL      = -300
};

/* ---- enums: values identical to the reference (include/Basic/BasicTypes.cuh:9-16) -------- */
enum rt_primitive_type { RT_SPHERE = 0, RT_PARALLELOGRAM = 1, RT_TRIANGLE = 2 };
enum rt_material_type  { RT_ROUGH = 0, RT_METAL = 1, RT_DIELECTRIC = 2 /* extension, not in the reference */ };

typedef struct rt_vec3 { float x, y, z; } rt_vec3;

/* ---- geometry inputs: the constructor arguments of the reference classes ---------------- */
/* Sphere(materialType, materialIndex, center, radius)    include/Geometry/Sphere.cuh:27-28   */
typedef struct rt_sphere {
    rt_vec3  center;
    float    radius;
    uint32_t material_type;
    uint32_t material_index;
} rt_sphere;

/* Parallelogram(materialType, materialIndex, q, u, v)    include/Geometry/Parallelogram.cuh:26-39
 * (area, normalVector, d are derived inside the library exactly as that constructor does)      */
typedef struct rt_parallelogram {
    rt_vec3  q, u, v;
    uint32_t material_type;
    uint32_t material_index;
} rt_parallelogram;

/* Triangle(materialType, materialIndex, vertexes[, vertexNormals])  include/Geometry/Triangle.cuh:26-46
 * has_normals == 0 selects the 3-vertex constructor (flat normal unit(e1 x e2) on all vertices). */
typedef struct rt_triangle {
    rt_vec3  v[3];
    rt_vec3  n[3];
    uint32_t material_type;
    uint32_t material_index;
    uint32_t has_normals;
    uint32_t _pad;
} rt_triangle;

/* ---- materials (include/Material/Rough.cuh:11-12, include/Material/Metal.cuh:11-13) ------- */
typedef struct rt_rough      { rt_vec3 albedo; } rt_rough;
typedef struct rt_metal      { rt_vec3 albedo; float fuzz; } rt_metal;
typedef struct rt_dielectric { rt_vec3 albedo; float ior; } rt_dielectric;   /* extension */

/* ---- instances ---------------------------------------------------------------------------
 * The reference's instance map carries (PrimitiveType, index) (include/Global/Renderer.cuh:123-126);
 * multi-primitive instances exist only for VTK particles (src/Global/VTKReader.cu:203-209), which
 * pre-fill primitiveCount, boundingBox and centroid.  prim_count <= 1 means "single primitive":
 * count/box/centroid come from the primitive (src/Global/RenderPin.cu:124-140).  For prim_count > 1
 * the caller may pass the local box/centroid (has_bounds = 1, as VTKReader does); otherwise the
 * library uses the merged primitive boxes and the box centre.                                     */
typedef struct rt_instance_desc {
    uint32_t type;         /* rt_primitive_type                                                  */
    uint32_t has_bounds;
    uint64_t first_prim;   /* Instance::primitiveIndex  include/AS/Instance.cuh:29               */
    uint64_t prim_count;   /* Instance::primitiveCount  include/AS/Instance.cuh:31               */
    float    bounds_min[3], bounds_max[3];
    float    centroid[3];
    uint32_t _pad;
} rt_instance_desc;

/* Arguments of Instance::updateTransformArguments (include/AS/Instance.cuh:53, src/AS/Instance.cu:4-17):
 * M = T(shift) * Rx * Ry * Rz * S(scale), rotation in degrees.                                    */
typedef struct rt_instance_xform {
    rt_vec3 shift;
    rt_vec3 rotate_deg;
    rt_vec3 scale;
} rt_instance_xform;

/* Per-frame instance update callback; replaces
 * void (*updateInstances)(Instance*, size_t, size_t)  include/Global/Renderer.cuh:94.
 * It writes transform arguments; matrices, inverse, boxes are derived inside the library.         */
typedef void (*rt_update_cb)(rt_instance_xform *xforms, size_t instance_count, size_t frame, void *user);

/* ---- window + camera (include/Global/Renderer.cuh:46-64) --------------------------------- */
typedef struct rt_window { uint32_t width, height; } rt_window;

typedef struct rt_camera_input {
    rt_vec3  background;
    rt_vec3  center;
    rt_vec3  target;
    float    fov;               /* horizontal, degrees (src/Global/RenderPin.cu:76-78)            */
    rt_vec3  up;
    float    focus_disk_radius;
    float    sample_range;      /* carried, unused by the reference kernel                        */
    uint64_t sample_count;      /* floored to a perfect square (src/Global/RenderPin.cu:93)       */
    uint64_t ray_trace_depth;
} rt_camera_input;

/* ---- scripted camera input (replaces the WASD / mouse controller of the window) -----------------------
 * SDL_OpenGLWindow::calculateNewPosition (src/Global/SDL_OpenGLWindow.cu:182-256) as a pure function of the
 * current camera and one step of input: yaw of the view direction about the camera's V, pitch about its U
 * (Rodrigues rotation, include/Basic/Vec3.cuh:140-160), optional pitch clamp, then W/A/S/D in the horizontal
 * plane, SPACE / LSHIFT along `up`, diagonal moves normalised.  The caller feeds the result to
 * rt_configure_camera before the next rt_render_frame (Renderer.cu:247-255 does exactly that every frame).
 * Transliterated from the reference, same variable flow and branch order (its file needs SDL/GL headers to compile,
 * so it cannot be run here): PARITY UNPINNED.                     */
enum rt_key { RT_KEY_W = 1, RT_KEY_A = 2, RT_KEY_S = 4, RT_KEY_D = 8, RT_KEY_SPACE = 16, RT_KEY_LSHIFT = 32 };
typedef struct rt_camera_motion {
    int32_t  dx, dy;             /* relative mouse motion of this step, pixels                                   */
    uint32_t keys;               /* RT_KEY_* bits held during this step                                         */
    float    mouse_sensitivity;  /* radians per pixel (reference: 0.001, Renderer.cu:224-225)                    */
    float    pitch_limit;        /* radians; the reference passes PI / radians(80) = 2.25 (SDL_OpenGLWindow.cuh:66),
                                    which asin() never exceeds — pass that for identical behaviour             */
    float    move_speed;         /* world units per step (reference start value 2 * 0.05)                        */
} rt_camera_motion;
/* next = current with center/target moved; *moved (may be NULL) = 1 if anything changed.  Host arithmetic only. */
int rt_camera_step(const rt_camera_input *current, const rt_camera_motion *motion, rt_camera_input *next, int *moved);

/* ---- options ----------------------------------------------------------------------------- */
enum rt_flags {
    RT_FLAG_NONE                    = 0,
    /* The reference's Parallelogram::constructBoundingBox is wrong (src/Geometry/Parallelogram.cu:48-50)
     * and culls most of the quad.  Default = bug-compatible (results identical to the reference).  */
    RT_FLAG_FIX_PARALLELOGRAM_BBOX  = 1u << 0,
    /* Record primary-hit AOVs (t, instance, primitive) of each pixel's first sample.              */
    RT_FLAG_RECORD_PRIMARY_AOV      = 1u << 1,
    /* Keep the binary LBVH arrays after the wide collapse so rt_read_lbvh can return them.        */
    RT_FLAG_KEEP_LBVH               = 1u << 2,
    /* Run the instrumented trace kernel: rt_frame_out.{node,prim,instance}_fetches are filled (the roofline's
     * algorithmic-bytes figure).  Slower; never set for timed runs.                                  */
    RT_FLAG_COUNT_TRAVERSAL         = 1u << 3,
    /* Comparison mode: traverse the binary LBVH (64-byte two-box nodes) instead of the compressed 8-wide BVH. */
    RT_FLAG_TRAVERSE_BVH2           = 1u << 4,
    /* Comparison mode: one thread per pixel runs all samples and bounces (the reference's kernel structure,
     * src/Global/Kernel.cu:105-147) instead of the wavefront (shade/generate + persistent trace kernels).          */
    RT_FLAG_MEGAKERNEL              = 1u << 5,
    /* Comparison mode: the wave loop (one shade/generate + one trace launch per bounce of all paths, ray queues in
     * between) instead of the default single persistent launch that fuses both (kernels_persist.cu).                */
    RT_FLAG_WAVEFRONT               = 1u << 6,
    /* Optional SAH refinement of the LBVH: one tree rotation per node, applied bottom-up inside the refit (measured:
     * -1.5 % node visits on config 2, +0.3 ms build per 629 k triangles — off by default).                          */
    RT_FLAG_SAH_REFINE              = 1u << 7,
    /* BLASes built by rt_build_acceleration_structure use PLOC (parallel locally-ordered clustering, Meister & Bittner
     * 2018: bottom-up agglomeration by smallest merged surface area over the Morton order) instead of the Karras LBVH:
     * a tree of near top-down-SAH quality for geometry that is built once, at several times the build time.  The
     * per-frame rebuilds of rt_set_vtk_frame and the TLAS always use the LBVH.                                          */
    RT_FLAG_SAH_BUILD               = 1u << 8
};

typedef struct rt_frame_out {
    uint64_t rays;            /* closest-hit queries traced (= TLAS::hit calls, src/Global/Kernel.cu:68) */
    uint64_t primary_rays;    /* W*H*spp of this context's shard                                         */
    float    render_ms;       /* CUDA-event time, first launch of the frame .. framebuffer ready          */
    float    tlas_ms;         /* CUDA-event time of the per-frame instance update + TLAS rebuild           */
    float    rng_ms;          /* CUDA-event time of the per-pixel XORWOW initialisation                    */
    float    trace_ms;        /* CUDA-event time of the trace/shade kernel                                 */
    uint32_t kernel_launches; /* kernels launched for this frame                                           */
    uint32_t stack_overflow;  /* != 0 if a traversal stack overflowed (reference has no check, Q15)         */
    uint32_t waves;           /* wavefront iterations of this frame (0 in megakernel mode)                  */
    uint32_t _pad;
    uint64_t node_fetches;    /* BVH node records fetched      (only with RT_FLAG_COUNT_TRAVERSAL)                  */
    uint64_t prim_fetches;    /* primitive records fetched     (only with RT_FLAG_COUNT_TRAVERSAL)                  */
    uint64_t instance_fetches;/* instance records fetched      (only with RT_FLAG_COUNT_TRAVERSAL)                  */
} rt_frame_out;

typedef struct rt_build_stats {
    float    blas_ms;          /* triangles resident -> wide BVH ready (CUDA events)              */
    float    tlas_ms;
    uint64_t blas_count;
    uint64_t primitive_count;
    uint64_t lbvh_nodes;       /* binary internal nodes                                            */
    uint64_t wide_nodes;
    uint64_t node_bytes;       /* bytes of traversable node memory                                 */
    uint64_t prim_bytes;       /* bytes of traversable primitive memory                            */
    uint32_t kernel_launches;
    uint32_t _pad;
} rt_build_stats;

enum rt_buffer_kind {
    RT_BUFFER_RGBA8       = 0, /* uchar4, row 0 = bottom scan-line (src/Global/RenderPin.cu:91)   */
    RT_BUFFER_RADIANCE    = 1, /* float3 per pixel, mean over samples before gamma               */
    RT_BUFFER_PRIMARY_T   = 2, /* float per pixel, +inf on miss                                   */
    RT_BUFFER_PRIMARY_INSTANCE = 3, /* int32 per pixel, -1 on miss                                */
    RT_BUFFER_PRIMARY_PRIM     = 4  /* int32 per pixel: (type << 28) | index, -1 on miss          */
};

typedef struct rt_ctx rt_ctx;

/* ---- life cycle -------------------------------------------------------------------------- */
int  rt_create(rt_ctx **out, int cuda_device, uint32_t flags);
/* Renderer::cleanup  include/Global/Renderer.cuh:143-145, src/Global/Renderer.cu:368-392 */
int  rt_cleanup(rt_ctx *ctx);
const char *rt_last_error(const rt_ctx *ctx);
const char *rt_version(void);
/* sizeof() of a struct of this header by name ("rt_triangle", ...), 0 if unknown: lets an FFI binding
 * verify its mirror of the layouts at load time.                                                   */
size_t rt_struct_size(const char *struct_name);

/* ---- scene commits (inputs are copied; caller keeps ownership) ----------------------------- */
/* Renderer::commitGeometryData  include/Global/Renderer.cuh:116, src/Global/Renderer.cu:9-43 */
int rt_commit_geometry(rt_ctx *ctx,
                       const rt_sphere *spheres, size_t sphere_count,
                       const rt_parallelogram *parallelograms, size_t parallelogram_count,
                       const rt_triangle *triangles, size_t triangle_count);
/* Renderer::commitMaterialData  include/Global/Renderer.cuh:119, src/Global/Renderer.cu:45-69 */
int rt_commit_materials(rt_ctx *ctx,
                        const rt_rough *roughs, size_t rough_count,
                        const rt_metal *metals, size_t metal_count);
int rt_commit_dielectrics(rt_ctx *ctx, const rt_dielectric *dielectrics, size_t count); /* extension */
/* ---- adding a geometry type (reference files/AddGeometryType.txt:1-35: ten steps in ten files) -------------------
 * Here a primitive type is one row of the table in csrc/rt_prims.h + one PrimOps<type> specialisation (bounds, hit,
 * surface, local box / centroid).  A build compiled with -DRT_EXT_PRIMITIVE='"header.h"' has a fourth type (value 3 in
 * rt_instance_desc.type); its device records (the header's ExtPrimD, record_bytes = sizeof) are committed here, original
 * order, like the arrays of rt_commit_geometry.  The standard build returns RT_ERR_ARGUMENT.
 * rt_primitive_type_count: 3, or 4 in a build with an extension type.                                              */
int rt_commit_extension_primitives(rt_ctx *ctx, const void *records, size_t record_bytes, size_t count);
int rt_primitive_type_count(void);
/* Renderer::configureInstances  include/Global/Renderer.cuh:123-126, src/Global/Renderer.cu:71-121.
 * initial_xforms may be NULL (identity) — the callback, if given, is invoked with frame 0.      */
int rt_configure_instances(rt_ctx *ctx, const rt_instance_desc *instances, size_t instance_count,
                           const rt_instance_xform *initial_xforms, rt_update_cb update, void *user);
/* Renderer::buildAccelerationStructure  include/Global/Renderer.cuh:129-130, src/Global/Renderer.cu:123-155.
 * Device LBVH build of every BLAS (one per distinct (type, first_prim, prim_count)) + TLAS.       */
int rt_build_acceleration_structure(rt_ctx *ctx);
int rt_get_build_stats(const rt_ctx *ctx, rt_build_stats *out);
/* Renderer::configureCamera  include/Global/Renderer.cuh:133, src/Global/Renderer.cu:157-178 */
int rt_configure_camera(rt_ctx *ctx, const rt_window *window, const rt_camera_input *camera);

/* ---- per frame ---------------------------------------------------------------------------
 * One iteration of Renderer::startRender's loop (src/Global/Renderer.cu:232-338) minus SDL/GL:
 * update callback -> instance transforms -> TLAS rebuild -> render kernel(s) -> framebuffer.
 * `seed` replaces clock64() in curand_init(pixelIndex ^ clock64(), pixelIndex, 0) (src/Global/Kernel.cu:114). */
int rt_render_frame(rt_ctx *ctx, uint64_t frame_index, uint64_t seed, rt_frame_out *out);
/* Set instance transforms directly (no callback); takes effect at the next rt_render_frame.       */
int rt_set_instance_xforms(rt_ctx *ctx, const rt_instance_xform *xforms, size_t instance_count);

/* Headless replacement of the CUDA-GL surface (src/Global/SDL_OpenGLWindow.cu:258-270):
 * copy a framebuffer plane to host memory.  Rows are bottom-up like the reference surface.         */
int rt_read_framebuffer(rt_ctx *ctx, int kind, void *host_dst, size_t dst_bytes);
/* Device pointer of a plane (for NCCL / peer access by the caller); size in bytes via *bytes.      */
int rt_framebuffer_device_ptr(rt_ctx *ctx, int kind, void **dev_ptr, size_t *bytes);
/* Write the RGBA8 plane as a binary PPM (top-down, flipped) — the headless "display"; a path ending in ".png" gets
 * an 8-bit RGBA PNG instead (top-down, uncompressed deflate: no zlib dependency).                    */
int rt_write_ppm(rt_ctx *ctx, const char *path);
/* Dump any plane of this context (rt_buffer_kind) to a file — the headless output stage after the path:
 *   RT_BUFFER_RGBA8    -> binary PPM "P6", or PNG when the path ends in ".png" (as rt_write_ppm)
 *   RT_BUFFER_RADIANCE -> PFM "PF" (float RGB, little-endian, rows bottom-up: the plane's own order)
 *   RT_BUFFER_PRIMARY_T-> PFM "Pf" (one float per pixel, +inf on miss)
 *   either of the two with a path ending in ".exr" -> uncompressed 32-bit float OpenEXR scan-line file, rows top-down
 *                         (channels R, G, B / Z)
 *   RT_BUFFER_PRIMARY_INSTANCE / _PRIM -> "RTAOV1 <width> <height> <kind>\n" + int32 little-endian per pixel, bottom-up */
int rt_write_image(rt_ctx *ctx, int kind, const char *path);

/* ---- multi-GPU image-tile sharding ---------------------------------------------------------
 * This context renders only tile rows t with (t % world) == rank, a tile being `tile_rows` scan-lines.
 * Pixel RNG streams are keyed by the global pixel index, so the union over ranks is bit-identical to
 * a single-GPU frame.  Shard planes are packed: local tile k holds global tile k*world + rank.
 * If peer_rgba8 != NULL the final RGBA8 texels are stored straight into that (peer-mapped, full-size)
 * buffer instead of the local packed plane (fused gather over NVLink).                              */
int rt_set_tile_shard(rt_ctx *ctx, uint32_t rank, uint32_t world, uint32_t tile_rows, void *peer_rgba8);

/* ---- frames in flight (the frame loop behind the ABI) -------------------------------------------------
 * Renderer::startRender keeps two frames in flight: instances, TLAS and TraverseData are double-buffered and a copy
 * stream prepares frame k+1 beside the render stream's frame k (reference src/Global/Renderer.cu:189-207, 278-317).
 * Here a WINDOW of up to RT_MAX_FRAMES_IN_FLIGHT consecutive frames is prepared (update callback, instance matrices,
 * TLAS and per-pixel RNG states of every frame) and then served by ONE launch of the persistent path kernel: the
 * pixel queues of the frames follow one another, so a lane that runs out of pixels of frame k starts on frame k+1
 * instead of idling through k's tail (a pixel is sequential work: at a 1/8 image shard that tail is a third of a frame).
 * Every frame is, bit for bit, the frame rt_render_frame(frame, seed) produces.  Two windows may be in flight
 * (handles 0 and 1): while one drains, the next one's preparation and first pixels already run on a second stream.
 * The window size and the tile shard are those of the context at submit time; so is the camera, unless
 * rt_set_window_cameras gave every frame of the window its own.
 * rt_frame_out of a pipelined frame: rays, primary_rays, stack_overflow (of the window so far) and render_ms = host
 * clock from the window's submission to the frame's completion (its latency); the per-stage times are 0.            */
#ifndef RT_MAX_FRAMES_IN_FLIGHT
#define RT_MAX_FRAMES_IN_FLIGHT 8     /* at most 16: a path carries its frame in four bits */
#endif
/* seeds: one per frame (NULL: `seed0` for every frame).  Returns at once; *window receives the handle.  RT_ERR_STATE
 * when both windows are in flight.                                                                                   */
int rt_submit_frames(rt_ctx *ctx, uint64_t first_frame, uint32_t n_frames, const uint64_t *seeds, uint64_t seed0, uint32_t *window);
/* Cameras of the frames of the NEXT window (n_frames of them, in frame order; consumed by the next rt_submit_frames, which
 * must have that many frames): the reference re-reads its controller every frame (src/Global/Renderer.cu:247-255), so a
 * scripted camera path (rt_camera_step) can run inside a window too.  Window size, sample_count and ray_trace_depth stay
 * those of rt_configure_camera.  Without this call every frame of a window uses the configured camera.                 */
int rt_set_window_cameras(rt_ctx *ctx, const rt_camera_input *cameras, uint32_t n_frames);
/* Blocks until frame k (0-based within the window) is complete on this context's device — and, on rank 0 of a tile shard
 * with a frame ring (below), until every rank's tiles of that frame have landed in the ring.                          */
int rt_wait_frame(rt_ctx *ctx, uint32_t window, uint32_t k, rt_frame_out *out);
/* Copies a plane of frame k to host memory (after rt_wait_frame) on a copy stream, beside the rest of the window.
 * RT_BUFFER_RGBA8 on a context with a frame ring returns the FULL assembled frame (width*height*4 bytes).             */
int rt_read_frame(rt_ctx *ctx, uint32_t window, uint32_t k, int kind, void *host_dst, size_t dst_bytes);
typedef struct rt_window_stats {
    uint64_t rays;            /* closest-hit queries of the whole window                                        */
    uint32_t frames;
    uint32_t kernel_launches; /* kernels launched for the window (preparation of every frame + one path launch) */
    float    device_ms;       /* CUDA events: first preparation launch .. path kernel finished                  */
    float    prepare_ms;      /* ... of which instance updates, TLAS rebuilds and RNG initialisation             */
    uint32_t stack_overflow;
    uint32_t _pad;
} rt_window_stats;
/* Waits for the rest of the window and frees its handle for the next rt_submit_frames.                               */
int rt_release_window(rt_ctx *ctx, uint32_t window, rt_window_stats *stats);
/* The headless loop: frames first_frame .. first_frame + n_frames - 1 in windows of `frames_per_window` (1 ..
 * RT_MAX_FRAMES_IN_FLIGHT), two windows in flight, every finished frame copied to pinned host memory and handed to
 * `sink` in frame order (may be NULL; non-zero return stops the loop with RT_ERR_STATE).  Without a frame ring the sink
 * receives this context's (packed shard) plane; with one, rank 0 receives the full assembled frames and the other
 * ranks are called with (NULL, 0): their pixels are in rank 0's ring.                                                  */
typedef struct rt_loop_stats {
    uint64_t frames, rays, kernel_launches;
    double   wall_ms;         /* host wall clock of the loop                                                      */
    double   device_ms;       /* sum of the windows' device times (windows overlap: >= the wall clock share)       */
    double   span_ms;         /* device clock (CUDA events): start of the loop .. last window finished              */
    double   max_latency_ms;  /* slowest submit -> frame-complete latency                                          */
} rt_loop_stats;
typedef int (*rt_frame_sink)(void *user, size_t frame, const void *rgba8, size_t bytes, const rt_frame_out *out);
int rt_run_frames(rt_ctx *ctx, uint64_t first_frame, size_t n_frames, uint64_t seed, uint32_t frames_per_window,
                  rt_frame_sink sink, void *user, rt_loop_stats *stats);

/* ---- multi-GPU: the frame ring (fused gather) -------------------------------------------------------------
 * Tile-sharded ranks (rt_set_tile_shard) can store their finished texels straight into ONE full-size frame on rank 0
 * instead of gathering packed planes afterwards: the path kernel of every rank writes RGBA8 through a peer mapping over
 * NVLink and raises a per-frame flag in rank 0's memory when its last pixel of the frame is done.  The ring holds
 * 2 windows x RT_MAX_FRAMES_IN_FLIGHT full frames + the flag table.
 *   rank 0            : rt_frame_ring_create  (after rt_set_tile_shard and rt_configure_camera) -> 64-byte CUDA IPC handle
 *   ranks 1..world-1  : rt_frame_ring_open(handle)      (other processes, e.g. torchrun ranks; the caller ships the bytes)
 *   same process      : rt_frame_ring_attach(owner_ctx) (several contexts driven by one process, any devices)
 * With a ring, rt_render_frame stores into frame slot 0 of window 0, and rt_read_framebuffer / rt_framebuffer_device_ptr
 * (RT_BUFFER_RGBA8) refer to that full frame: rank 0 reads it after every rank has finished its call (the callers
 * synchronise).  From the moment the ring exists every rank must submit the same sequence of windows.
 * Errors: RT_ERR_CUDA when the mapping cannot be made (no peer access between the devices), RT_ERR_STATE when the
 * shard / camera are not set, RT_ERR_ARGUMENT for a ring of another frame size.                                        */
#define RT_IPC_HANDLE_BYTES 64
int rt_frame_ring_create(rt_ctx *ctx, void *ipc_handle_out /* RT_IPC_HANDLE_BYTES, may be NULL */);
int rt_frame_ring_open(rt_ctx *ctx, const void *ipc_handle);
int rt_frame_ring_attach(rt_ctx *ctx, rt_ctx *owner);
int rt_frame_ring_close(rt_ctx *ctx);

/* ---- acceleration-structure introspection (parity tests) ----------------------------------- */
/* Binary LBVH as built on the device, for `which` = 0 (the BLAS batch: all BLASes share one sort and one
 * Karras pass, segment id in the key's high bits) or 1 (the TLAS): sorted Morton keys, sorted build-primitive
 * ids, Karras child/parent arrays, node boxes.  Any output pointer may be NULL; *n_prims is always written.
 * Children: value >= 0 -> internal node, value < 0 -> leaf ~value (position in sorted order).      */
int rt_read_lbvh(rt_ctx *ctx, size_t which, size_t *n_prims,
                 uint64_t *sorted_keys, uint32_t *sorted_prim_ids,
                 int32_t *left, int32_t *right, int32_t *parent, float *node_bounds /* 6 per node */);

/* Compressed 8-wide BVH as traversed by the render kernels, `which` as above.  nodes: 80 bytes per node (layout in
 * csrc/rt_types.h BvhNode8), leaf_prims: per leaf-order position the original primitive index (BLAS batch) or the
 * instance index (TLAS).  Roots: node s for BLAS s of the batch, node 0 for the TLAS.  Either pointer may be NULL. */
int rt_read_wide_bvh(rt_ctx *ctx, size_t which, size_t *n_nodes, size_t *n_leaf_prims, void *nodes, uint32_t *leaf_prims);

/* ---- VTK particle-mesh series (Renderer::configureVTKFiles include/Global/Renderer.cuh:113,
 *      src/Global/Renderer.cu:394-443; VTKReader src/Global/VTKReader.cu:16-220) ------------------ */
int rt_load_vtk_series(rt_ctx *ctx, const char *series_path, size_t *frame_count);
int rt_set_vtk_frame(rt_ctx *ctx, size_t frame);
/* Frame-loop overlap (the loop body of Renderer::startRender, src/Global/Renderer.cu:232-338, keeps host preparation of
 * the next frame beside the GPU's current one): start reading + converting series frame `frame` on a host thread and
 * return at once; the next rt_set_vtk_frame(frame) takes the result instead of parsing.  One read-ahead at a time.  */
int rt_prefetch_vtk_frame(rt_ctx *ctx, size_t frame);

/* The whole headless frame loop of the animation behind one call (Renderer::startRender's while-loop,
 * src/Global/Renderer.cu:232-338, with the SDL/GL present replaced by `sink`): renders the series frames
 * first_frame, first_frame + frame_stride, ... (at most max_frames of them; 0 = to the end of the series) — frame
 * sharding over GPUs is first_frame = rank, frame_stride = world.  `lanes` renderer contexts (this one plus lanes-1
 * internal copies of its committed scene, camera and shard, each with its own stream, device buffers and read-ahead
 * thread) run on `lanes` host threads and take the frames round-robin, so one lane's commit + upload overlaps the
 * other's render on the same GPU.  Per frame: rt_set_vtk_frame (frame read ahead), rt_render_frame(frame, seed), RGBA8
 * plane -> pinned host memory, then `sink` (may be NULL) is called — one call at a time, from a lane's thread, frames
 * NOT in order across lanes; a non-zero return stops the loop (RT_ERR_STATE).  Every frame is the one
 * rt_set_vtk_frame + rt_render_frame on a single context would produce.  An rt_update_cb registered with
 * rt_configure_instances is called from the lanes' threads (concurrently for different frames).  The lane contexts are
 * kept until rt_cleanup.                                                                                                  */
typedef struct rt_animation_stats {
    uint64_t frames;              /* frames rendered                                                     */
    uint64_t rays;
    uint64_t triangles;           /* primitives rebuilt, summed over the frames                          */
    uint64_t kernel_launches;
    double   wall_ms;             /* host wall clock of the whole loop                                   */
    double   device_ms;           /* sum over frames of CUDA-event rebuild + render time (lanes overlap) */
    double   slowest_rebuild_ms;  /* max over frames of BLAS batch + TLAS rebuild                        */
    /* where the lanes' host threads spent their time, summed over frames (host wall clock per call) */
    double   parse_ms;            /* read-ahead threads: file -> triangles (overlaps everything else)    */
    double   set_frame_ms;        /* rt_set_vtk_frame: wait for the read-ahead, compose, commit, H2D, rebuild */
    double   render_wall_ms;      /* rt_render_frame                                                     */
    double   readback_ms;         /* RGBA8 plane -> pinned host                                          */
    double   sink_ms;             /* inside the caller's sink (incl. waiting for the other lanes' calls) */
} rt_animation_stats;
int rt_run_vtk_animation(rt_ctx *ctx, size_t first_frame, size_t frame_stride, size_t max_frames, uint64_t seed, int lanes,
                         rt_frame_sink sink, void *user, rt_animation_stats *stats);

/* ---- measurement aid (bench.py roofline denominators; no reference counterpart) ----------------------
 * Streams a device buffer of `bytes` with 128-bit loads from every SM, `repeats` times back to back, and returns
 * the read bandwidth in GB/s (CUDA events).  bytes well below the L2 size (126 MB) measures the L2 -> SM read
 * bandwidth that bounds node/triangle fetches of an L2-resident BVH; bytes well above it measures HBM reads.     */
int rt_probe_read_bandwidth(rt_ctx *ctx, size_t bytes, int repeats, double *gbs);

/* ---- TLAS build that allocates nothing ----------------------------------------------------------------------------
 * This entry point has NO COUNTERPART IN THE REFERENCE; it was added for callers that
 * cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid it).  It is the per-frame
 * TLAS rebuild of the frame path (one block, one launch: csrc/kernels_wide.cu tlas_small_kernel, the kernel rt_render_frame /
 * rt_submit_frames use for up to 1024 instances) with its scratch memory in a file-scope static device array instead of the
 * context's buffers.  It needs no rt_ctx, calls no cudaMalloc / cudaFree / allocating routine, and does not synchronise: two
 * kernel launches and one 4-byte device-to-device copy are enqueued on `cuda_stream` (a cudaStream_t; NULL = default stream)
 * of the CURRENT device, and the call returns.
 *   boxes_dev               in   n_instances x 6 floats (lo x y z, hi x y z): the world boxes the TLAS leaves are to have
 *   nodes_out_dev           out  room for (n_instances + 1) records of 80 bytes (BvhNode8, csrc/rt_types.h), 16-byte aligned;
 *                                node 0 is the root
 *   leaf_instances_out_dev  out  n_instances words: instance index per leaf-order position
 *   n_nodes_out_dev         out  one word: number of wide nodes written (may be NULL)
 * All four are device memory of the current device.  The node contents are those rt_read_wide_bvh(ctx, 1, ...) returns for
 * the same leaf boxes; as there, only their placement depends on the order of the build's atomics.
 * LIMITS: n_instances in 1 .. RT_TLAS_NO_ALLOC_MAX_INSTANCES, else RT_ERR_ARGUMENT with the reason in error_text (error_cap bytes,
 * may be NULL).  The scratch is RT_TLAS_NO_ALLOC_SCRATCH_BYTES (112 KiB) per GPU that loads the library, held until the process
 * exits.  Every call on a device shares that array (it keeps the previous call's contents; the kernels rewrite or reset what
 * they read): calls that overlap on the GPU — two streams, two host threads — RACE and must be serialised by the caller.      */
#define RT_TLAS_NO_ALLOC_MAX_INSTANCES 1024
#define RT_TLAS_NO_ALLOC_SCRATCH_BYTES (112 * 1024)
int rt_build_tlas_no_alloc(const float *boxes_dev, uint32_t n_instances, void *nodes_out_dev, uint32_t *leaf_instances_out_dev,
                           uint32_t *n_nodes_out_dev, void *cuda_stream, char *error_text, size_t error_cap);

#ifdef __cplusplus
}
#endif
#endif /* RTB200_H */
