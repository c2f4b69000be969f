// kernels_persist.cu — the whole frame in ONE persistent launch: every resident thread carries one pixel at a time
// through all of its samples and bounces (the reference's per-pixel XORWOW stream is sequential, so a pixel is the
// unit of sequential work, src/Global/Kernel.cu:107-146) and takes the next unstarted pixel when it is done.
//
// A thread runs the same per-ray state machine as the wavefront trace kernel (trace_closest8 of rt_wide_trace.h,
// step for step) plus one more step, SHADE (hit record, Rough/Metal::scatter, rayColor's bookkeeping, sample
// accumulation, next primary ray — exactly wave_shade_kernel's per-slot work).  The warp only votes on WHICH step
// its lanes take next, so that lanes take the expensive steps together:
//   NODE  : pop the nearest pending child, fetch its 80-byte node, test the 8 child boxes
//   LEAF  : one primitive test (BLAS) or one instance entry (TLAS)
//   SHADE : consume the finished ray's hit, start the path's next ray (or the next sample / pixel)
// Compared with the wave loop there is no ray queue, no per-wave launch and no wave tail: a lane whose ray ends
// waits only until enough lanes of its own warp want to shade.  The per-pixel results are bit-identical (each pixel
// still consumes its own stream in the reference's order); only the schedule differs.
#include "wave.h"

#include "rt_pixel.h"
#include "wave_device.h"

#include <algorithm>
#include <cstdio>
#include <cstdlib>

namespace {

constexpr int PT = 128;
constexpr uint32_t PID_MASK = 0x0FFFFFFFu;      // path state: tile-major pixel id in the low 28 bits, frame of the window above
constexpr uint32_t NG_INNER = 0x00FFFFFFu;      // node group has pending inner hits iff .y > NG_INNER
#ifndef RT_PERSIST_MIN_BLOCKS
#define RT_PERSIST_MIN_BLOCKS 6   /* 80 registers: no spills; measured faster than 8 blocks x 64 registers */
#endif
// Shared-memory staging of the top BVH levels (north_star): the first RT_STAGE_TOP_NODES records of the BLAS node pool
// (the collapse allocates breadth-first, so these are the roots and the levels right below them) are copied into
// shared memory by every CTA and node steps read them from there.  Measured on C2 (profiles/r01h_stage_top_nodes.txt):
// no gain — those records are the ones that already hit in L1 (L1 hit rate 49 % of all sectors) — so it is compiled
// out by default (0); the variant build is `make VARIANT=stage EXTRA=-DRT_STAGE_TOP_NODES=96`.
// Variant (A/B): run the pop / leave-instance transition right after a node or leaf step, with that step's lanes, instead
// of only at the loop top (where it runs at 4-7 lanes: profiles/r02a_persist8_c2_hotlines.txt).
#ifndef RT_POP_IN_PHASE
#define RT_POP_IN_PHASE 0
#endif
#ifndef RT_STAGE_TOP_NODES
#define RT_STAGE_TOP_NODES 0
#endif
enum : uint32_t { ST_DEAD = 0u, ST_TRACE = 1u, ST_SHADE = 2u, ST_FRESH = 3u };
// CAMS: every frame of the window has its own camera pose (pb.frame_cams, global memory, indexed by the path's frame); the
// default instantiation (false) reads the one camera from the kernel parameters and is, instruction for instruction, the
// kernel without this option.
// SINGLE: the scene is ONE instance (configs 2 and 4).  A top level over one box only prunes rays that miss the scene, and entering
// the instance costs every ray a node step on the TLAS root plus a LEAF step of its own; here the ray is taken into the
// instance's space where it starts (in the SHADE step, with the lanes that start rays together) and begins at the BLAS root.  The
// candidates, their order and the acceptance arithmetic are those of the two-level walk, so the frames are the same bit for bit
// (the parity tests run on it: every C2 / C4 test); the TLAS-leaf branch and the leave-instance transition compile away.
template<bool COUNT, bool CAMS = false, bool SINGLE = false>
__global__ void __launch_bounds__(PT, RT_PERSIST_MIN_BLOCKS) persist8_kernel(const __grid_constant__ SceneD sc, const __grid_constant__ CameraD cam,
                                                                            const __grid_constant__ FrameD fr, const __grid_constant__ PersistBuffers pb,
                                                                            int shade_min, int node_min) {
    __shared__ uint2 stack_mem[RT_WIDE_STACK * PT];
    uint2 *const stk = stack_mem + threadIdx.x;               // entry k at stk[k * PT]
#if RT_STAGE_TOP_NODES > 0
    __shared__ uint4 top_nodes[RT_STAGE_TOP_NODES * 5];
    const uint32_t n_top = min((uint32_t) RT_STAGE_TOP_NODES, pb.n_blas_nodes);
    for (uint32_t i = threadIdx.x; i < n_top * 5u; i += PT) top_nodes[i] = __ldg(reinterpret_cast<const uint4 *>(sc.blas_nodes8) + i);
    __syncthreads();
#endif
    const uint32_t tid = blockIdx.x * PT + threadIdx.x;
    const uint32_t slot = tid;                                // path-state entry of this lane
    uint32_t state = ST_FRESH;
    WideRay r{};
    uint2 ng = make_uint2(0u, 0u), tg = make_uint2(0u, 0u);
    int sp = 0, inst = -1, inst_sp = 0, prim_type = 0;
    HitD hit{};
    uint32_t overflow = 0, rays = 0;
    TravStats stats{0, 0, 0};

    for (;;) {
        // ---- cheap transitions: nothing pending in registers -> leave instance / ray finished / pop -----------------
#define RT_ADVANCE()                                                                                                         \
        if (state == ST_TRACE && tg.y == 0u && ng.y <= NG_INNER) {                                                           \
            if (!SINGLE && inst >= 0 && sp == inst_sp && sp != 0) { /* BLAS exhausted, TLAS work left: back to world space */ \
                const float4 o = pb.ray_o[slot], d = pb.ray_d[slot];                                                         \
                inst = -1 - (int) __float_as_uint(d.w);          /* inst < 0: on the TLAS of frame -1 - inst of the window */ \
                r.ray.o = mk3(o.x, o.y, o.z); r.ray.d = mk3(d.x, d.y, d.z);                                                  \
                wide_setup_ray(r);                                                                                           \
            }                                                                                                                \
            if (sp == 0) {                                                                                                   \
                state = ST_SHADE;                                                                                            \
                ng = make_uint2(0u, 0u);                                                                                     \
            } else {                                                                                                         \
                sp--;                                                                                                        \
                const uint2 e = stk[sp * PT];                                                                                \
                if (e.y > NG_INNER) ng = e; else { tg = e; ng = make_uint2(0u, 0u); }                                        \
            }                                                                                                                \
        }
        RT_ADVANCE()
        const bool on_node = state == ST_TRACE && tg.y == 0u && ng.y > NG_INNER;
        const bool on_leaf = state == ST_TRACE && tg.y != 0u;
        const uint32_t node_lanes = __ballot_sync(0xffffffffu, on_node);
        const uint32_t leaf_lanes = __ballot_sync(0xffffffffu, on_leaf);
        const uint32_t shade_lanes = __ballot_sync(0xffffffffu, state >= ST_SHADE);
        if ((node_lanes | leaf_lanes | shade_lanes) == 0u) break;            // every lane ran out of pixels
        const int nn = __popc(node_lanes), ns = __popc(shade_lanes);
        int phase;                                                           // 0 node, 1 leaf, 2 shade
        if (nn >= node_min) phase = 0;
        else if (ns >= shade_min) phase = 2;
        else if (leaf_lanes) phase = 1;
        else if (nn) phase = 0;
        else phase = 2;

        if (phase == 0) {
            // ---- NODE ---------------------------------------------------------------------------------------------
            if (on_node) {
                const uint32_t hits_imask = ng.y;
                const int bit = rt_bfind(hits_imask);
                const uint32_t base = ng.x;
                ng.y &= ~(1u << bit);
                if (ng.y > NG_INNER) {
                    if (sp < RT_WIDE_STACK) { stk[sp * PT] = ng; sp++; } else overflow = 1u;
                }
                const uint32_t sl = ((uint32_t) bit - 24u) ^ (r.oct_inv4 & 0xFFu);
                const uint32_t rel = (uint32_t) __popc(hits_imask & ~(0xFFFFFFFFu << sl));
#if RT_STAGE_TOP_NODES > 0
                BvhNode8 n;
                if (inst >= 0 && base + rel < n_top) {
                    const uint4 *q = top_nodes + (base + rel) * 5u;
                    const uint4 a = q[0];
                    n.n0 = make_float4(__uint_as_float(a.x), __uint_as_float(a.y), __uint_as_float(a.z), __uint_as_float(a.w));
                    n.n1 = q[1]; n.n2 = q[2]; n.n3 = q[3]; n.n4 = q[4];
                } else {
                    n = load_node8(!SINGLE && inst < 0 ? sc.tlas_nodes8 + (size_t) (uint32_t) (-1 - inst) * pb.tlas_node_stride : sc.blas_nodes8, base + rel);
                }
#else
                const BvhNode8 n = load_node8(!SINGLE && inst < 0 ? sc.tlas_nodes8 + (size_t) (uint32_t) (-1 - inst) * pb.tlas_node_stride : sc.blas_nodes8, base + rel);
#endif
                if (COUNT) stats.nodes++;
                const uint32_t hm = node8_intersect(n, r, RT_TMIN, hit.tc);
                ng.x = n.n1.x;
                ng.y = hm & 0xFF000000u;
                if (ng.y) ng.y |= __float_as_uint(n.n0.w) >> 24;
                tg.x = n.n1.y;
                tg.y = hm & 0x00FFFFFFu;
#if RT_POP_IN_PHASE & 1
                RT_ADVANCE()    // a node without hits: pop here, with the lanes of the node step, not alone at the loop top
#endif
            }
        } else if (phase == 1) {
            // ---- LEAF: one primitive test or one instance entry per lane --------------------------------------------
            if (on_leaf) {
                const int idx = rt_bfind(tg.y);
                tg.y &= ~(1u << idx);
                if (SINGLE || inst >= 0) {
                    if (COUNT) stats.prims++;
                    leaf_test8(sc, (uint32_t) prim_type, tg.x + (uint32_t) idx, r.ray, RT_TMIN, SINGLE ? 0 : inst, hit);
#if RT_POP_IN_PHASE & 2
                    RT_ADVANCE()
#endif
                } else {
                    // TLAS leaf: enter one instance, keep the rest of the group and the node group for later
                    if (tg.y) { if (sp < RT_WIDE_STACK) { stk[sp * PT] = tg; sp++; } else overflow = 1u; }
                    if (ng.y > NG_INNER) { if (sp < RT_WIDE_STACK) { stk[sp * PT] = ng; sp++; } else overflow = 1u; }
                    const uint32_t fw = (uint32_t) (-1 - inst);      // frame of the window this ray belongs to
                    inst = (int) __ldg(&sc.tlas_leaf_inst[(size_t) fw * pb.tlas_leaf_stride + tg.x + (uint32_t) idx]);
                    if (COUNT) stats.insts++;
                    float inv[12];
                    int root8;
                    load_instance_wide(sc, inst, inv, root8, prim_type, fw * pb.inst_stride);
                    const f3 wo = r.ray.o, wd = r.ray.d;
                    r.ray.o = xf_point(inv, wo);                     // Instance::hit: origin as point, direction as vector,
                    r.ray.d = xf_vector(inv, wd);                    // not re-normalised (reference src/AS/Instance.cu:26-28)
                    wide_setup_ray(r);
                    inst_sp = sp;
                    ng = make_uint2((uint32_t) root8, 0x80000000u);
                    tg = make_uint2(0u, 0u);
                }
            }
        } else {
            // ---- SHADE: the per-slot work of wave_shade_kernel, for the lanes whose ray has finished ------------------
            const bool shading = state >= ST_SHADE;
            const bool fresh = state == ST_FRESH;
            uint32_t pid = SLOT_IDLE;
            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
            RngState rng{};
            RayD ray{};
            f3 thr = mk3(1.f, 1.f, 1.f);
            uint32_t sample = 0, depth = 0, pix_rays = 0;
            // pid = tile-major pixel id | frame of the window << 28 (PID_MASK / FW below)
            bool need_pixel = shading && fresh;
            bool need_primary = false, acc_dirty = false;
            if (shading && !fresh) {
                acc = pb.acc[slot];
                pid = __float_as_uint(acc.w);
                const float4 o = pb.ray_o[slot], d = pb.ray_d[slot], t4 = pb.thr[slot];
                const uint4 ra = pb.rng_a[slot];
                const uint2 rb = pb.rng_b[slot];
                rng.d = ra.x; rng.v0 = ra.y; rng.v1 = ra.z; rng.v2 = ra.w; rng.v3 = rb.x; rng.v4 = rb.y;
                ray.o = mk3(o.x, o.y, o.z); ray.d = mk3(d.x, d.y, d.z);
                thr = mk3(t4.x, t4.y, t4.z);
                const uint32_t sd = __float_as_uint(t4.w);
                sample = sd >> 16; depth = sd & 0xFFFFu;
                pix_rays = __float_as_uint(o.w);                     // rays of this pixel so far
                const uint32_t fw = pid >> 28;
                if (fr.aov_t && sample == 0u && depth == 0u) {       // primary hit of the pixel's first sample
                    uint32_t lx, ly;
                    pixel_coords(pid & PID_MASK, pb.tiles_x, lx, ly);
                    const size_t lp = (size_t) fw * pb.frame_px + ly * cam.width + lx;
                    const bool has = hit.inst >= 0;
                    fr.aov_t[lp] = has ? hit.t : RT_INF;
                    fr.aov_inst[lp] = has ? hit.inst : -1;
                    fr.aov_prim[lp] = has ? (int) ((ref_type(hit.leaf) << 28) | sc.leaf_prims8[ref_pos(hit.leaf)]) : -1;
                }
                const bool alive = shade_scatter<true>(sc, CAMS ? pb.frame_cams[fw] : cam, ray, hit, thr, rng, fw * pb.inst_stride);
                depth++;
                if (!(alive && depth < cam.depth)) {
                    // path over: rayColor returns the throughput (Q6); next sample or next pixel
                    acc.x += thr.x; acc.y += thr.y; acc.z += thr.z;
                    acc_dirty = true;
                    sample++;
                    if (sample == cam.sqrtSamples * cam.sqrtSamples) {
                        uint32_t lx, ly;
                        pixel_coords(pid & PID_MASK, pb.tiles_x, lx, ly);
                        const size_t lp = (size_t) fw * pb.frame_px + ly * cam.width + lx;
                        const float w = cam.rcpSqrtSamples * cam.rcpSqrtSamples;
                        const float cr = acc.x * w, cg = acc.y * w, cb = acc.z * w;
                        fr.radiance[3 * lp + 0] = cr;
                        fr.radiance[3 * lp + 1] = cg;
                        fr.radiance[3 * lp + 2] = cb;
                        const size_t op = fr.rgba8_is_full_frame ? (size_t) fw * pb.full_px + (size_t) global_row(fr, ly) * cam.width + lx : lp;
                        reinterpret_cast<uchar4 *>(fr.rgba8)[op] = make_uchar4(to_byte(cr), to_byte(cg), to_byte(cb), 255);
                        if (fw == 0u) atomicAdd(&pb.tile_cost[(pid & PID_MASK) >> 5], pix_rays);      // the next launch hands the dear tiles out first
                        if (pb.frame_done) {
                            // frame completion: the pixel's stores, then its count; the lane that counts the frame's last pixel
                            // publishes the frame (system-wide fence, then the flags the host / rank 0 wait on)
                            if (pb.frame_rays) atomicAdd(&pb.frame_rays[fw], (unsigned long long) pix_rays);
                            __threadfence();
                            if (atomicAdd(&pb.frame_done[fw], 1u) + 1u == pb.valid_pixels) {
                                __threadfence_system();
                                if (pb.peer_flags) pb.peer_flags[fw] = pb.ticket;
                                pb.host_flags[fw] = pb.ticket;
                            }
                        }
                        need_pixel = true;
                    } else {
                        need_primary = true;
                    }
                }
            }
            // lanes that finished a pixel (or have none yet) take the next unstarted pixel; ids outside the image
            // (edge tiles) are skipped
            for (;;) {
                const uint32_t any = __ballot_sync(0xffffffffu, need_pixel);
                if (any == 0u) break;
                const uint32_t cand = warp_append(pb.next_pixel, need_pixel);
                if (need_pixel) {
                    const uint32_t cf = cand / pb.n_pixels;          // frame of the window (its queue follows the previous frame's)
                    if (cf >= pb.n_frames) {
                        state = ST_DEAD; need_pixel = false;
                    } else {
                        const uint32_t cp = cand - cf * pb.n_pixels;
                        const uint32_t tile = __ldg(&pb.tile_order[cp >> 5]);
                        const uint32_t gp = (tile << 5) | (cp & 31u);
                        uint32_t lx, ly;
                        pixel_coords(gp, pb.tiles_x, lx, ly);
                        if (lx < cam.width && ly < fr.local_rows && global_row(fr, ly) < cam.height) {
                            pid = gp | (cf << 28); need_pixel = false; need_primary = true;
                            rng = fr.rng[(size_t) cf * pb.frame_px + ly * cam.width + lx];
                            acc = make_float4(0.f, 0.f, 0.f, __uint_as_float(pid));
                            acc_dirty = true;
                            sample = 0;
                            pix_rays = 0;
                        }
                    }
                }
            }
            if (shading && state != ST_DEAD) {
                if (need_primary) {
                    uint32_t lx, ly;
                    pixel_coords(pid & PID_MASK, pb.tiles_x, lx, ly);
                    const uint32_t si = sample / cam.sqrtSamples, sj = sample % cam.sqrtSamples;
                    ray = primary_ray(CAMS ? pb.frame_cams[pid >> 28] : cam, lx, global_row(fr, ly), si, sj, rng);
                    thr = mk3(1.f, 1.f, 1.f);
                    depth = 0;
                }
                if (acc_dirty) pb.acc[slot] = acc;
                // every lane that is still alive here starts a ray: the path's next segment or a new primary ray
                pb.ray_o[slot] = make_float4(ray.o.x, ray.o.y, ray.o.z, __uint_as_float(pix_rays + 1u));
                pb.ray_d[slot] = make_float4(ray.d.x, ray.d.y, ray.d.z, __uint_as_float(pid >> 28));
                pb.thr[slot] = make_float4(thr.x, thr.y, thr.z, __uint_as_float((sample << 16) | depth));
                pb.rng_a[slot] = make_uint4(rng.d, rng.v0, rng.v1, rng.v2);
                pb.rng_b[slot] = make_uint2(rng.v3, rng.v4);
                hit.t = RT_INF; hit.tc = RT_INF; hit.u = 0.f; hit.v = 0.f; hit.inst = -1; hit.leaf = 0;
                sp = 0; inst_sp = 0;
                if (SINGLE) {
                    // straight into the one instance (Instance::hit: origin as point, direction as vector, not re-normalised)
                    float inv[12];
                    int root8;
                    load_instance_wide(sc, 0, inv, root8, prim_type, (pid >> 28) * pb.inst_stride);
                    if (COUNT) stats.insts++;
                    r.ray.o = xf_point(inv, ray.o);
                    r.ray.d = xf_vector(inv, ray.d);
                    wide_setup_ray(r);
                    inst = 0;
                    ng = make_uint2((uint32_t) root8, 0x80000000u);
                } else {
                    r.ray = ray;
                    wide_setup_ray(r);
                    inst = -1 - (int) (pid >> 28);
                    ng = make_uint2(0u, 0x80000000u);       // TLAS root: one pending inner hit at relative index 0
                }
                tg = make_uint2(0u, 0u);
                state = ST_TRACE;
                rays++;
            }
        }
    }

#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        rays += __shfl_xor_sync(0xffffffffu, rays, off);
        overflow |= __shfl_xor_sync(0xffffffffu, overflow, off);
    }
    if (lane_id() == 0) {
        if (rays) atomicAdd(fr.ray_counter, (unsigned long long) rays);
        if (overflow) atomicOr(fr.overflow_flag, 1u);
    }
    if (COUNT) {
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            stats.nodes += __shfl_xor_sync(0xffffffffu, stats.nodes, off);
            stats.prims += __shfl_xor_sync(0xffffffffu, stats.prims, off);
            stats.insts += __shfl_xor_sync(0xffffffffu, stats.insts, off);
        }
        if (lane_id() == 0) {
            atomicAdd(fr.ray_counter + 2, (unsigned long long) stats.nodes);
            atomicAdd(fr.ray_counter + 3, (unsigned long long) stats.prims);
            atomicAdd(fr.ray_counter + 4, (unsigned long long) stats.insts);
        }
    }
}

// ---- hand-out order of the tiles -------------------------------------------------------------------------------
// The frame ends when the LAST pixel does, and a pixel is sequential work: handing the dear tiles out first and the
// cheap ones (sky: one ray per sample) last keeps the end of the frame short (longest-processing-time-first).  The
// cost of a tile is what it took in the previous frame of this context (rays traced, accumulated by the path kernel);
// a frame without history is handed out in scan order.  Key = cost class (64 classes over 1..depth rays per sample,
// dear first); the radix sort is stable, so tiles of one class stay in scan order and the warps in flight keep
// sharing BVH nodes.
__global__ void __launch_bounds__(256) tile_keys_kernel(uint32_t *__restrict__ tile_cost, uint32_t n_tiles, uint32_t base, uint32_t span,
                                                        uint64_t *__restrict__ keys, uint32_t *__restrict__ vals) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n_tiles) return;
    const uint32_t c = tile_cost[t];
    tile_cost[t] = 0u;                                            // this frame accumulates afresh
    uint32_t cls = 0u;
    if (c > base) {
        const unsigned long long x = (unsigned long long) (c - base) * 64ull / span;
        cls = x > 63ull ? 63u : (uint32_t) x;
    }
    keys[t] = 63u - cls;
    vals[t] = t;
}

} // namespace

void PersistRenderer::render_window(const SceneD &sc, const CameraD &cam, const FrameD &fr, const PersistBuffers *win, cudaStream_t stream, int sm_count) {
    if (blocks_per_sm_[0] == 0) {
        CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks_per_sm_[0], persist8_kernel<false>, PT, 0));
        CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks_per_sm_[1], persist8_kernel<true>, PT, 0));
    }
    const bool count = fr.count_traversal != 0;
    int bps = std::max(1, blocks_per_sm_[count ? 1 : 0]);
    if (blocks_per_sm_override > 0) bps = std::min(bps, blocks_per_sm_override);
    PersistBuffers pb{};
    pb.tiles_x = (cam.width + 7u) / 8u;
    pb.n_tiles = pb.tiles_x * ((fr.local_rows + 3u) / 4u);
    pb.n_pixels = pb.n_tiles * 32u;
    // no more threads than pixels (a warp = one 8x4 tile at a time)
    uint32_t want_blocks = (pb.n_pixels + PT - 1) / PT;
    uint32_t blocks = std::min<uint32_t>((uint32_t) (sm_count * bps), std::max(1u, want_blocks));
    size_t n_threads = (size_t) blocks * PT;
    if (split && !(win && win->frame_cams)) {       // one path-state entry per path of a CTA's pool
        if (split_blocks_per_sm_[count ? 1 : 0] == 0) split_blocks_per_sm_[count ? 1 : 0] = split_blocks_per_sm(count);
        int sb = split_blocks_per_sm_[count ? 1 : 0];
        if (blocks_per_sm_override > 0) sb = std::min(sb, blocks_per_sm_override);
        const uint32_t slots = (uint32_t) split_slots_per_block();
        const size_t total_px = (size_t) pb.n_pixels * (win ? win->n_frames : 1u);
        want_blocks = (uint32_t) std::min<size_t>((total_px + slots - 1) / slots, 1u << 20);
        blocks = std::min<uint32_t>((uint32_t) (sm_count * sb), std::max(1u, want_blocks));
        n_threads = (size_t) blocks * slots;
        hit_.reserve(n_threads); hit_inst_.reserve(n_threads);
    }
    ray_o_.reserve(n_threads); ray_d_.reserve(n_threads); thr_.reserve(n_threads); acc_.reserve(n_threads);
    rng_a_.reserve(n_threads); rng_b_.reserve(n_threads);
    counter_.reserve(1);
    tile_cost_.reserve(pb.n_tiles); keys_a_.reserve(pb.n_tiles); keys_b_.reserve(pb.n_tiles); vals_a_.reserve(pb.n_tiles); vals_b_.reserve(pb.n_tiles);
    // the history belongs to one frame geometry (resolution, shard, samples): anything else starts afresh
    const uint64_t sig = ((uint64_t) cam.width << 40) ^ ((uint64_t) fr.local_rows << 24) ^ ((uint64_t) fr.shard_rank << 16) ^ ((uint64_t) fr.shard_world << 8)
                         ^ ((uint64_t) cam.sqrtSamples << 4) ^ (uint64_t) cam.depth ^ ((uint64_t) fr.tile_rows << 52);
    if (sig != history_sig_ || tile_order == 0) {
        CUDA_CHECK(cudaMemsetAsync(tile_cost_.get(), 0, pb.n_tiles * sizeof(uint32_t), stream));
        history_sig_ = sig;
    }
    uint64_t *keys = keys_a_.get(), *keys_alt = keys_b_.get();
    uint32_t *vals = vals_a_.get(), *vals_alt = vals_b_.get();
    const uint32_t spp = cam.sqrtSamples * cam.sqrtSamples;
    const uint32_t base = 32u * spp, span = std::max(1u, base * (std::max(2u, cam.depth) - 1u));
    tile_keys_kernel<<<div_up(pb.n_tiles, 256), 256, 0, stream>>>(tile_cost_.get(), pb.n_tiles, base, span, keys, vals);
    uint32_t sort_launches = 0;
    sorter_.sort(keys, keys_alt, vals, vals_alt, (int) pb.n_tiles, 6, stream, sm_count, sort_launches);
    launches = 2 + sort_launches;
    pb.ray_o = ray_o_.get(); pb.ray_d = ray_d_.get(); pb.thr = thr_.get(); pb.acc = acc_.get();
    pb.rng_a = rng_a_.get(); pb.rng_b = rng_b_.get(); pb.next_pixel = counter_.get();
    pb.hit = hit_.get(); pb.hit_inst = hit_inst_.get();
    pb.tile_order = vals; pb.tile_cost = tile_cost_.get();
    pb.n_blas_nodes = n_blas_nodes;
    pb.n_frames = 1u;
    pb.frame_px = cam.width * fr.local_rows;
    if (win) {
        pb.n_frames = win->n_frames; pb.frame_px = win->frame_px; pb.full_px = win->full_px;
        pb.tlas_node_stride = win->tlas_node_stride; pb.tlas_leaf_stride = win->tlas_leaf_stride; pb.inst_stride = win->inst_stride;
        pb.valid_pixels = win->valid_pixels; pb.ticket = win->ticket;
        pb.frame_rays = win->frame_rays; pb.frame_done = win->frame_done; pb.host_flags = win->host_flags; pb.peer_flags = win->peer_flags;
        pb.frame_cams = win->frame_cams;
    }
    CUDA_CHECK(cudaMemsetAsync(counter_.get(), 0, sizeof(uint32_t), stream));
    const bool single = single_instance && !split;
    if (pb.frame_cams) {                                    // per-frame cameras (never with counting / split)
        if (single) persist8_kernel<false, true, true><<<blocks, PT, 0, stream>>>(sc, cam, fr, pb, shade_min, node_min);
        else persist8_kernel<false, true><<<blocks, PT, 0, stream>>>(sc, cam, fr, pb, shade_min, node_min);
    } else if (split) launch_persist_split(sc, cam, fr, pb, blocks, node_min, count, stream);
    else if (count) {
        if (single) persist8_kernel<true, false, true><<<blocks, PT, 0, stream>>>(sc, cam, fr, pb, shade_min, node_min);
        else persist8_kernel<true><<<blocks, PT, 0, stream>>>(sc, cam, fr, pb, shade_min, node_min);
    } else {
        if (single) persist8_kernel<false, false, true><<<blocks, PT, 0, stream>>>(sc, cam, fr, pb, shade_min, node_min);
        else persist8_kernel<false><<<blocks, PT, 0, stream>>>(sc, cam, fr, pb, shade_min, node_min);
    }
    CUDA_CHECK(cudaGetLastError());
}
