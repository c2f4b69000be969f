// kernels_wave.cu — wavefront path tracer: per wave one SHADE/GENERATE kernel (thread per slot) and one persistent
// TRACE kernel (warps pull rays from the wave's queue, lanes are refilled as their rays finish).
//
//   wave_shade : consumes the hit of the slot's previous ray — hit record, Rough/Metal::scatter, rayColor's
//                bookkeeping (reference src/Global/Kernel.cu:64-100), sample accumulation and stratified primary-ray
//                generation (Kernel.cu:119-143), pixel write-out (Color3::castToUchar4) — then queues the slot's next
//                ray.  The per-pixel XORWOW stream is advanced exactly in the reference's order.
//   wave_trace : closest hit (the job of TLAS::hit, src/AS/TLAS.cu:131-201) for every queued ray.
#include "wave.h"

#include "rt_pixel.h"
#include "wave_device.h"

#include <algorithm>
#include <cstring>
#include <cstdio>
#include <cstdlib>

namespace {

constexpr int SHADE_THREADS = 256;
template<bool WIDE>
__global__ void __launch_bounds__(SHADE_THREADS) wave_shade_kernel(const __grid_constant__ SceneD sc, const __grid_constant__ CameraD cam,
                                                                   const __grid_constant__ FrameD fr, const __grid_constant__ WaveBuffers wb,
                                                                   uint32_t wave) {
    const uint32_t slot = blockIdx.x * blockDim.x + threadIdx.x;
    const bool in_range = slot < wb.n_slots;
    uint32_t pid = SLOT_IDLE;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    if (in_range) { acc = wb.acc[slot]; pid = __float_as_uint(acc.w); }
    const bool fresh = in_range && pid == SLOT_FRESH;
    bool live = in_range && pid != SLOT_IDLE;

    RngState rng{};
    RayD ray{};
    f3 thr = mk3(1.f, 1.f, 1.f);
    uint32_t sample = 0, depth = 0;
    bool need_pixel = fresh;              // slot has no pixel yet
    bool need_primary = false;
    bool emit = false;

    if (live && !fresh) {
        const float4 h = wb.hit[slot];
        const float4 o = wb.ray_o[slot], d = wb.ray_d[slot], t4 = wb.thr[slot];
        const uint4 ra = wb.rng_a[slot];
        const uint2 rb = wb.rng_b[slot];
        rng.d = ra.x; rng.v0 = ra.y; rng.v1 = ra.z; rng.v2 = ra.w; rng.v3 = rb.x; rng.v4 = rb.y;
        ray.o = mk3(o.x, o.y, o.z); ray.d = mk3(d.x, d.y, d.z);
        thr = mk3(t4.x, t4.y, t4.z);
        const uint32_t sd = __float_as_uint(t4.w);
        sample = sd >> 16; depth = sd & 0xFFFFu;
        HitD hit;
        hit.t = h.x; hit.tc = h.x; hit.u = h.y; hit.v = h.z; hit.leaf = (int) __float_as_uint(h.w); hit.inst = wb.hit_inst[slot];
        if (fr.aov_t && sample == 0u && depth == 0u) {       // primary hit of the pixel's first sample
            uint32_t lx, ly;
            pixel_coords(pid, wb.tiles_x, lx, ly);
            const uint32_t lp = ly * cam.width + lx;
            const bool has = hit.inst >= 0;
            fr.aov_t[lp] = has ? hit.t : RT_INF;
            fr.aov_inst[lp] = has ? hit.inst : -1;
            fr.aov_prim[lp] = has ? (int) ((ref_type(hit.leaf) << 28) | (WIDE ? sc.leaf_prims8[ref_pos(hit.leaf)] : sc.leaf_prims[ref_pos(hit.leaf)])) : -1;
        }
        const bool alive = shade_scatter<WIDE>(sc, cam, ray, hit, thr, rng);
        depth++;
        if (alive && depth < cam.depth) {
            emit = true;
        } else {
            // path over: rayColor returns the throughput (Q6); next sample or next pixel
            acc.x += thr.x; acc.y += thr.y; acc.z += thr.z;
            sample++;
            if (sample == cam.sqrtSamples * cam.sqrtSamples) {
                uint32_t lx, ly;
                pixel_coords(pid, wb.tiles_x, lx, ly);
                const uint32_t lp = ly * cam.width + lx;
                const float w = cam.rcpSqrtSamples * cam.rcpSqrtSamples;
                const float r = acc.x * w, g = acc.y * w, b = acc.z * w;
                fr.radiance[3 * (size_t) lp + 0] = r;
                fr.radiance[3 * (size_t) lp + 1] = g;
                fr.radiance[3 * (size_t) lp + 2] = b;
                const size_t op = fr.rgba8_is_full_frame ? (size_t) global_row(fr, ly) * cam.width + lx : (size_t) lp;
                reinterpret_cast<uchar4 *>(fr.rgba8)[op] = make_uchar4(to_byte(r), to_byte(g), to_byte(b), 255);
                need_pixel = true;
            } else {
                need_primary = true;
            }
        }
    }

    // slots that finished a pixel (or have none yet) take the next unstarted pixel; ids outside the image
    // (edge tiles) are skipped
    for (;;) {
        const uint32_t any = __ballot_sync(0xffffffffu, need_pixel);
        if (any == 0u) break;
        const uint32_t cand = warp_append(wb.next_pixel, need_pixel);
        if (need_pixel) {
            if (cand >= wb.n_pixels) {
                pid = SLOT_IDLE; live = false; need_pixel = false;
                wb.acc[slot] = make_float4(0.f, 0.f, 0.f, __uint_as_float(SLOT_IDLE));
            } else {
                // group-local id -> tile-major id of the shard: the group's k-th tile is tile k * n_groups + group
                const uint32_t gp = ((((cand >> 5) * wb.n_groups) + wb.group) << 5) | (cand & 31u);
                uint32_t lx, ly;
                pixel_coords(gp, wb.tiles_x, lx, ly);
                if (lx < cam.width && ly < fr.local_rows && global_row(fr, ly) < cam.height) {
                    pid = gp; need_pixel = false; need_primary = true;
                    const RngState r0 = fr.rng[ly * cam.width + lx];
                    rng = r0;
                    acc = make_float4(0.f, 0.f, 0.f, __uint_as_float(pid));
                    sample = 0;
                }
            }
        }
    }

    if (live && need_primary) {
        uint32_t lx, ly;
        pixel_coords(pid, wb.tiles_x, lx, ly);
        const uint32_t si = sample / cam.sqrtSamples, sj = sample % cam.sqrtSamples;
        ray = primary_ray(cam, lx, global_row(fr, ly), si, sj, rng);
        thr = mk3(1.f, 1.f, 1.f);
        depth = 0;
        emit = true;
        wb.acc[slot] = acc;
    }
    if (live && emit) {
        wb.ray_o[slot] = make_float4(ray.o.x, ray.o.y, ray.o.z, 0.f);
        wb.ray_d[slot] = make_float4(ray.d.x, ray.d.y, ray.d.z, 0.f);
        wb.thr[slot] = make_float4(thr.x, thr.y, thr.z, __uint_as_float((sample << 16) | depth));
        wb.rng_a[slot] = make_uint4(rng.d, rng.v0, rng.v1, rng.v2);
        wb.rng_b[slot] = make_uint2(rng.v3, rng.v4);
    }
    const uint32_t qi = warp_append(wb.counts + wave, live && emit);
    if (live && emit) wb.queue[qi] = slot;
}

// ---------------------------------------------------------------------------------------------------------------
// TRACE, binary LBVH.  Persistent warps; per lane one ray in flight.  Each outer iteration:
//   1. refill: lanes without a ray take the next queue entries (one atomic per warp) when enough of them are idle;
//   2. node phase: lanes standing on an inner node step down the tree; the phase ends when fewer than NODE_MIN
//      lanes still have an inner node (the others are parked on a leaf / instance / finished);
//   3. leaf phase: parked lanes do their primitive test, instance entry or instance exit together.
// This keeps the expensive reference-exact primitive tests (three IEEE divisions each) from running at one or two
// active lanes per issue, which is what made the one-thread-per-pixel kernel spend 80 % of its issue slots idle
// (profiles/r01a: 6.35 of 32 threads active per instruction).
constexpr int TRACE_THREADS = 128;

template<bool COUNT>
__global__ void __launch_bounds__(TRACE_THREADS, 4) wave_trace2_kernel(const __grid_constant__ SceneD sc, const __grid_constant__ WaveBuffers wb,
                                                                       uint32_t wave, int refill_min, int node_min, unsigned long long *stats_out) {
    const uint32_t n_rays = wb.counts[wave];
    uint32_t *cursor = wb.cursors + wave;
    int stack[RT_STACK_SIZE];
    int sp = 0;
    uint32_t slot = SLOT_IDLE;
    bool exhausted = false;
    int cur = 0, inst = -1;
    RayD ray{}, wray{};
    RaySlabs slabs{};
    HitD hit{};
    uint32_t overflow = 0;
    TravStats stats{0, 0, 0};

    for (;;) {
        // ---- 1. refill ------------------------------------------------------------------------------------------
        const uint32_t idle = __ballot_sync(0xffffffffu, slot == SLOT_IDLE);
        if (idle != 0u && !exhausted && (__popc(idle) >= refill_min || idle == 0xffffffffu)) {
            const uint32_t leader = __ffs(idle) - 1u;
            uint32_t base = 0;
            if (lane_id() == leader) base = atomicAdd(cursor, (uint32_t) __popc(idle));
            base = __shfl_sync(0xffffffffu, base, leader);
            if (slot == SLOT_IDLE) {
                const uint32_t k = base + (uint32_t) __popc(idle & ((1u << lane_id()) - 1u));
                if (k < n_rays) {
                    slot = __ldg(&wb.queue[k]);
                    const float4 o = wb.ray_o[slot], d = wb.ray_d[slot];
                    wray.o = mk3(o.x, o.y, o.z); wray.d = mk3(d.x, d.y, d.z);
                    ray = wray;
                    slabs = make_slabs(ray);
                    hit.t = RT_INF; hit.tc = RT_INF; hit.u = 0.f; hit.v = 0.f; hit.inst = -1; hit.leaf = 0;
                    sp = 0; inst = -1; cur = sc.tlas_root_ref;
                }
            }
            if (base + (uint32_t) __popc(idle) >= n_rays) exhausted = true;
        }
        if (__ballot_sync(0xffffffffu, slot != SLOT_IDLE) == 0u) {
            if (exhausted) break;
            continue;
        }
        // ---- 2. node phase ----------------------------------------------------------------------------------------
        for (;;) {
            const bool on_node = slot != SLOT_IDLE && cur >= 0;
            if (on_node) {
                const BvhNode2 n = load_node2(inst < 0 ? sc.tlas_nodes : sc.blas_nodes, cur);
                if (COUNT) stats.nodes++;
                float tn0, tf0, tn1, tf1;
                node2_test(n, slabs, RT_TMIN, hit.tc, tn0, tf0, tn1, tf1);
                const bool h0 = tn0 <= tf0, h1 = tn1 <= tf1;
                if (h0 && h1) {
                    const bool swap = tn1 < tn0;
                    if (sp < RT_STACK_SIZE) stack[sp++] = swap ? n.refs.x : n.refs.y; else overflow = 1u;
                    cur = swap ? n.refs.y : n.refs.x;
                } else if (h0) cur = n.refs.x;
                else if (h1) cur = n.refs.y;
                else if (sp > 0) cur = stack[--sp];
                else cur = RT_REF_NONE;                                  // traversal finished
            }
            const uint32_t still = __ballot_sync(0xffffffffu, slot != SLOT_IDLE && cur >= 0);
            if (__popc(still) < node_min) break;
        }
        // ---- 3. leaf phase ----------------------------------------------------------------------------------------
        if (slot != SLOT_IDLE && cur < 0 && cur != RT_REF_NONE) {
            if (cur == RT_REF_SENTINEL) {
                inst = -1;
                ray = wray;
                slabs = make_slabs(ray);
            } else if (inst < 0) {
                inst = (int) ref_pos(cur);
                if (COUNT) stats.insts++;
                float inv[12];
                int root;
                load_instance_inv(sc, inst, inv, root);
                ray.o = xf_point(inv, wray.o);
                ray.d = xf_vector(inv, wray.d);
                slabs = make_slabs(ray);
                if (sp < RT_STACK_SIZE) stack[sp++] = RT_REF_SENTINEL; else overflow = 1u;
                cur = root;
                goto parked;                                             // root may be a node or a leaf: next iteration
            } else {
                if (COUNT) stats.prims++;
                leaf_test(sc, cur, ray, RT_TMIN, inst, hit);
            }
            cur = sp > 0 ? stack[--sp] : RT_REF_NONE;
        }
parked:
        if (slot != SLOT_IDLE && cur == RT_REF_NONE) {
            wb.hit[slot] = make_float4(hit.t, hit.u, hit.v, __uint_as_float((uint32_t) hit.leaf));
            wb.hit_inst[slot] = hit.inst;
            slot = SLOT_IDLE;
        }
    }
    if (overflow) atomicOr(reinterpret_cast<unsigned int *>(stats_out + 1), 1u);
    if (COUNT) {
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            stats.nodes += __shfl_xor_sync(0xffffffffu, stats.nodes, off);
            stats.prims += __shfl_xor_sync(0xffffffffu, stats.prims, off);
            stats.insts += __shfl_xor_sync(0xffffffffu, stats.insts, off);
        }
        if (lane_id() == 0) {
            atomicAdd(stats_out + 2, (unsigned long long) stats.nodes);
            atomicAdd(stats_out + 3, (unsigned long long) stats.prims);
            atomicAdd(stats_out + 4, (unsigned long long) stats.insts);
        }
    }
}


// ---------------------------------------------------------------------------------------------------------------
// TRACE, compressed 8-wide BVH (the default).  Same persistent-warp / lane-refill scheme; every lane runs the
// per-ray state machine of trace_closest8 (rt_wide_trace.h) step for step, so a ray sees exactly the node loads
// and primitive tests, in exactly the order, of the serial definition — the warp only decides WHEN a lane may
// take its next step, so that lanes take the expensive steps together:
//   NODE  : pop the nearest pending child of the node group, fetch its 80-byte node, test the 8 child boxes
//   LEAF  : one primitive test (BLAS) or one instance entry (TLAS) of the primitive group
//   cheap : stack pop / instance exit / ray finished (done by whoever needs it before every vote)
// Traversal stacks (24 x 8 B per lane) live in shared memory, interleaved by lane (no bank conflicts).
constexpr uint32_t NG_INNER = 0x00FFFFFFu;      // node group has pending inner hits iff .y > NG_INNER
#ifndef RT_TRACE8_MIN_BLOCKS
#define RT_TRACE8_MIN_BLOCKS 8
#endif
#ifndef RT_PREFETCH
#define RT_PREFETCH 0
#endif
__device__ __forceinline__ void prefetch_l1(const void *p) { asm volatile("prefetch.global.L1 [%0];" :: "l"(p)); }

template<bool COUNT>
__global__ void __launch_bounds__(TRACE_THREADS, RT_TRACE8_MIN_BLOCKS) wave_trace8_kernel(const __grid_constant__ SceneD sc, const __grid_constant__ WaveBuffers wb,
                                                                       uint32_t wave, int refill_min, int node_min, unsigned long long *stats_out) {
    __shared__ uint2 stack_mem[RT_WIDE_STACK * TRACE_THREADS];
    uint2 *const stk = stack_mem + threadIdx.x;               // entry k at stk[k * TRACE_THREADS]
    const uint32_t n_rays = wb.counts[wave];
    uint32_t *cursor = wb.cursors + wave;
    uint32_t slot = SLOT_IDLE;
    bool exhausted = false;
    WideRay r{};
    uint2 ng = make_uint2(0u, 0u), tg = make_uint2(0u, 0u);
    int sp = 0, inst = -1, inst_sp = 0, prim_type = 0;
    HitD hit{};
    uint32_t overflow = 0;
    TravStats stats{0, 0, 0};

    for (;;) {
        // ---- refill -------------------------------------------------------------------------------------------------
        const uint32_t idle = __ballot_sync(0xffffffffu, slot == SLOT_IDLE);
        if (idle != 0u && !exhausted && (__popc(idle) >= refill_min || idle == 0xffffffffu)) {
            const uint32_t leader = __ffs(idle) - 1u;
            uint32_t base = 0;
            if (lane_id() == leader) base = atomicAdd(cursor, (uint32_t) __popc(idle));
            base = __shfl_sync(0xffffffffu, base, leader);
            if (slot == SLOT_IDLE) {
                const uint32_t k = base + (uint32_t) __popc(idle & ((1u << lane_id()) - 1u));
                if (k < n_rays) {
                    slot = __ldg(&wb.queue[k]);
                    const float4 o = wb.ray_o[slot], d = wb.ray_d[slot];
                    r.ray.o = mk3(o.x, o.y, o.z); r.ray.d = mk3(d.x, d.y, d.z);
                    wide_setup_ray(r);
                    hit.t = RT_INF; hit.tc = RT_INF; hit.u = 0.f; hit.v = 0.f; hit.inst = -1; hit.leaf = 0;
                    sp = 0; inst = -1; inst_sp = 0;
                    ng = make_uint2(0u, 0x80000000u);           // TLAS root: one pending inner hit at relative index 0
                    tg = make_uint2(0u, 0u);
                }
            }
            if (base + (uint32_t) __popc(idle) >= n_rays) exhausted = true;
        }
        if (__ballot_sync(0xffffffffu, slot != SLOT_IDLE) == 0u) {
            if (exhausted) break;
            continue;
        }
        // ---- node phase ---------------------------------------------------------------------------------------------
        for (;;) {
            // cheap transitions: nothing pending in registers -> leave instance / finish / pop
            if (slot != SLOT_IDLE && tg.y == 0u && ng.y <= NG_INNER) {
                if (inst >= 0 && sp == inst_sp && sp != 0) {     // BLAS exhausted, TLAS work left: back to world space
                    inst = -1;
                    const float4 o = wb.ray_o[slot], d = wb.ray_d[slot];
                    r.ray.o = mk3(o.x, o.y, o.z); r.ray.d = mk3(d.x, d.y, d.z);
                    wide_setup_ray(r);
                }
                if (sp == 0) {
                    wb.hit[slot] = make_float4(hit.t, hit.u, hit.v, __uint_as_float((uint32_t) hit.leaf));
                    wb.hit_inst[slot] = hit.inst;
                    slot = SLOT_IDLE;
                    ng = make_uint2(0u, 0u);
                } else {
                    sp--;
                    const uint2 e = stk[sp * TRACE_THREADS];
                    if (e.y > NG_INNER) ng = e; else { tg = e; ng = make_uint2(0u, 0u); }
                }
            }
            const bool on_node = slot != SLOT_IDLE && tg.y == 0u && ng.y > NG_INNER;
            const uint32_t node_lanes = __ballot_sync(0xffffffffu, on_node);
            const uint32_t leaf_lanes = __ballot_sync(0xffffffffu, slot != SLOT_IDLE && tg.y != 0u);
            if (node_lanes == 0u || (__popc(node_lanes) < node_min && leaf_lanes != 0u)) break;
            if (on_node) {
                const uint32_t hits_imask = ng.y;
                const int bit = rt_bfind(hits_imask);
                const uint32_t base = ng.x;
                ng.y &= ~(1u << bit);
                if (ng.y > NG_INNER) {
                    if (sp < RT_WIDE_STACK) { stk[sp * TRACE_THREADS] = ng; sp++; } else overflow = 1u;
                }
                const uint32_t sl = ((uint32_t) bit - 24u) ^ (r.oct_inv4 & 0xFFu);
                const uint32_t rel = (uint32_t) __popc(hits_imask & ~(0xFFFFFFFFu << sl));
                const BvhNode8 n = load_node8(inst < 0 ? sc.tlas_nodes8 : sc.blas_nodes8, base + rel);
                if (COUNT) stats.nodes++;
                const uint32_t hm = node8_intersect(n, r, RT_TMIN, hit.tc);
                ng.x = n.n1.x;
                ng.y = hm & 0xFF000000u;
                if (ng.y) ng.y |= __float_as_uint(n.n0.w) >> 24;
                tg.x = n.n1.y;
                tg.y = hm & 0x00FFFFFFu;
#if RT_PREFETCH
                // the next record this lane will need is known now: start its trip through L2 while the lane waits for its phase
                if (tg.y) {
                    if (inst >= 0 && prim_type == 2) prefetch_l1(sc.leaf_tris8 + (tg.x + (uint32_t) rt_bfind(tg.y)));
                } else if (ng.y) {
                    const uint32_t sl2 = ((uint32_t) rt_bfind(ng.y) - 24u) ^ (r.oct_inv4 & 0xFFu);
                    const BvhNode8 *nx = (inst < 0 ? sc.tlas_nodes8 : sc.blas_nodes8) + (ng.x + (uint32_t) __popc(ng.y & ~(0xFFFFFFFFu << sl2)));
                    prefetch_l1(nx);
                    prefetch_l1(reinterpret_cast<const char *>(nx) + 64);
                }
#endif
            }
        }
        // ---- leaf phase: one primitive test or one instance entry per lane -------------------------------------------
        if (slot != SLOT_IDLE && tg.y != 0u) {
            const int idx = rt_bfind(tg.y);
            tg.y &= ~(1u << idx);
            if (inst >= 0) {
                if (COUNT) stats.prims++;
                leaf_test8(sc, (uint32_t) prim_type, tg.x + (uint32_t) idx, r.ray, RT_TMIN, inst, hit);
            } else {
                // TLAS leaf: enter one instance, keep the rest of the group and the node group for later
                if (tg.y) { if (sp < RT_WIDE_STACK) { stk[sp * TRACE_THREADS] = tg; sp++; } else overflow = 1u; }
                if (ng.y > NG_INNER) { if (sp < RT_WIDE_STACK) { stk[sp * TRACE_THREADS] = ng; sp++; } else overflow = 1u; }
                inst = (int) __ldg(&sc.tlas_leaf_inst[tg.x + (uint32_t) idx]);
                if (COUNT) stats.insts++;
                float inv[12];
                int root8;
                load_instance_wide(sc, inst, inv, root8, prim_type);
                const f3 wo = r.ray.o, wd = r.ray.d;
                r.ray.o = xf_point(inv, wo);                     // Instance::hit: origin as point, direction as vector,
                r.ray.d = xf_vector(inv, wd);                    // not re-normalised (reference src/AS/Instance.cu:26-28)
                wide_setup_ray(r);
                inst_sp = sp;
                ng = make_uint2((uint32_t) root8, 0x80000000u);
                tg = make_uint2(0u, 0u);
            }
        }
    }
    if (overflow) atomicOr(reinterpret_cast<unsigned int *>(stats_out + 1), 1u);
    if (COUNT) {
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            stats.nodes += __shfl_xor_sync(0xffffffffu, stats.nodes, off);
            stats.prims += __shfl_xor_sync(0xffffffffu, stats.prims, off);
            stats.insts += __shfl_xor_sync(0xffffffffu, stats.insts, off);
        }
        if (lane_id() == 0) {
            atomicAdd(stats_out + 2, (unsigned long long) stats.nodes);
            atomicAdd(stats_out + 3, (unsigned long long) stats.prims);
            atomicAdd(stats_out + 4, (unsigned long long) stats.insts);
        }
    }
}


// ---------------------------------------------------------------------------------------------------------------
// FINISH.  The last third of a frame's waves carries well under 1 % of its rays (a few pixels with long paths), but
// every wave costs a fixed launch + longest-ray latency.  Once a group's active slots fall below a threshold, this
// kernel takes the slots of its last queue and runs each one to the end of its pixel in ONE thread — shade, trace,
// shade, ... exactly the per-slot sequence the waves would have run (same RNG stream, same results).
template<bool COUNT, bool WIDE>
__global__ void __launch_bounds__(128) wave_finish_kernel(const __grid_constant__ SceneD sc, const __grid_constant__ CameraD cam,
                                                          const __grid_constant__ FrameD fr, const __grid_constant__ WaveBuffers wb, uint32_t wave) {
    __shared__ uint2 stack_mem[WIDE ? RT_WIDE_STACK * 128 : 1];
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= wb.counts[wave]) return;
    const uint32_t slot = wb.queue[i];
    float4 acc = wb.acc[slot];
    uint32_t pid = __float_as_uint(acc.w);
    const float4 h = wb.hit[slot];
    const float4 o = wb.ray_o[slot], d = wb.ray_d[slot], t4 = wb.thr[slot];
    const uint4 ra = wb.rng_a[slot];
    const uint2 rb = wb.rng_b[slot];
    RngState rng;
    rng.d = ra.x; rng.v0 = ra.y; rng.v1 = ra.z; rng.v2 = ra.w; rng.v3 = rb.x; rng.v4 = rb.y;
    RayD ray;
    ray.o = mk3(o.x, o.y, o.z); ray.d = mk3(d.x, d.y, d.z);
    f3 thr = mk3(t4.x, t4.y, t4.z);
    uint32_t sample = __float_as_uint(t4.w) >> 16, depth = __float_as_uint(t4.w) & 0xFFFFu;
    HitD hit;
    hit.t = h.x; hit.tc = h.x; hit.u = h.y; hit.v = h.z; hit.leaf = (int) __float_as_uint(h.w); hit.inst = wb.hit_inst[slot];
    uint32_t rays = 0, overflow = 0;
    TravStats stats{0, 0, 0};
    const uint32_t spp = cam.sqrtSamples * cam.sqrtSamples;
    for (;;) {
        uint32_t lx, ly;
        pixel_coords(pid, wb.tiles_x, lx, ly);
        if (fr.aov_t && sample == 0u && depth == 0u) {
            const uint32_t lp = ly * cam.width + lx;
            const bool has = hit.inst >= 0;
            fr.aov_t[lp] = has ? hit.t : RT_INF;
            fr.aov_inst[lp] = has ? hit.inst : -1;
            fr.aov_prim[lp] = has ? (int) ((ref_type(hit.leaf) << 28) | (WIDE ? sc.leaf_prims8[ref_pos(hit.leaf)] : sc.leaf_prims[ref_pos(hit.leaf)])) : -1;
        }
        const bool alive = shade_scatter<WIDE>(sc, cam, ray, hit, thr, rng);
        depth++;
        if (!(alive && depth < cam.depth)) {
            acc.x += thr.x; acc.y += thr.y; acc.z += thr.z;
            sample++;
            if (sample == spp) {
                const uint32_t lp = ly * cam.width + lx;
                const float w = cam.rcpSqrtSamples * cam.rcpSqrtSamples;
                const float r = acc.x * w, g = acc.y * w, b = acc.z * w;
                fr.radiance[3 * (size_t) lp + 0] = r;
                fr.radiance[3 * (size_t) lp + 1] = g;
                fr.radiance[3 * (size_t) lp + 2] = b;
                const size_t op = fr.rgba8_is_full_frame ? (size_t) global_row(fr, ly) * cam.width + lx : (size_t) lp;
                reinterpret_cast<uchar4 *>(fr.rgba8)[op] = make_uchar4(to_byte(r), to_byte(g), to_byte(b), 255);
                // next unstarted pixel of the group, if any (only when there are fewer slots than pixels)
                bool got = false;
                for (;;) {
                    const uint32_t cand = atomicAdd(wb.next_pixel, 1u);
                    if (cand >= wb.n_pixels) break;
                    const uint32_t gp = ((((cand >> 5) * wb.n_groups) + wb.group) << 5) | (cand & 31u);
                    pixel_coords(gp, wb.tiles_x, lx, ly);
                    if (lx < cam.width && ly < fr.local_rows && global_row(fr, ly) < cam.height) { pid = gp; got = true; break; }
                }
                if (!got) break;
                rng = fr.rng[ly * cam.width + lx];
                acc = make_float4(0.f, 0.f, 0.f, __uint_as_float(pid));
                sample = 0;
            }
            ray = primary_ray(cam, lx, global_row(fr, ly), sample / cam.sqrtSamples, sample % cam.sqrtSamples, rng);
            thr = mk3(1.f, 1.f, 1.f);
            depth = 0;
        }
        if (WIDE) trace_closest8<COUNT>(sc, ray, RT_TMIN, hit, overflow, stats, stack_mem + threadIdx.x, 128);
        else trace_closest<COUNT>(sc, ray, RT_TMIN, hit, overflow, stats);
        rays++;
    }
    wb.acc[slot] = make_float4(0.f, 0.f, 0.f, __uint_as_float(SLOT_IDLE));
    atomicAdd(fr.ray_counter, (unsigned long long) rays);
    if (overflow) atomicOr(fr.overflow_flag, 1u);
    if (COUNT) {
        atomicAdd(fr.ray_counter + 2, (unsigned long long) stats.nodes);
        atomicAdd(fr.ray_counter + 3, (unsigned long long) stats.prims);
        atomicAdd(fr.ray_counter + 4, (unsigned long long) stats.insts);
    }
}

} // namespace

// ---------------------------------------------------------------------------------------------------------------
void WaveRenderer::reserve(uint32_t n_slots) {
    ray_o_.reserve(n_slots); ray_d_.reserve(n_slots); hit_.reserve(n_slots); thr_.reserve(n_slots); acc_.reserve(n_slots);
    hit_inst_.reserve(n_slots); rng_a_.reserve(n_slots); rng_b_.reserve(n_slots); queue_.reserve(n_slots);
}

WaveRenderer::~WaveRenderer() {
    for (int g = 0; g < MAX_GROUPS; g++) {
        if (g_exec_[g]) cudaGraphExecDestroy(g_exec_[g]);
        if (g_event_[g]) cudaEventDestroy(g_event_[g]);
        if (g_stream_[g]) cudaStreamDestroy(g_stream_[g]);
    }
    if (fork_event_) cudaEventDestroy(fork_event_);
}

// The slots are split into n_groups groups (8x4-pixel tiles dealt round-robin) whose wave sequences are independent
// (a wave only orders the rays of ONE slot) and run concurrently, each on its own stream with 1/n_groups of the
// resident trace warps: while one group's trace kernel drains its last long rays, the other groups keep the SMs
// busy, and the fixed per-wave latency (launches, queue hand-over, the longest ray) overlaps with useful work —
// which is what limits a frame once it is sharded over several GPUs and the waves get small.
void WaveRenderer::render(const SceneD &sc, const CameraD &cam, const FrameD &fr, cudaStream_t stream, int sm_count, WaveStats &stats) {
    constexpr int BATCH = 32;
    constexpr int CW = 2 * BATCH + 4;                      // counter words per group
    const uint32_t tiles_x = (cam.width + 7u) / 8u, tiles_y = (fr.local_rows + 3u) / 4u;
    const uint32_t n_tiles = tiles_x * tiles_y;
    int G = std::max(1, std::min(n_groups, (int) MAX_GROUPS));
    if (n_tiles < 64u * (uint32_t) G) G = 1;              // tiny frames: one group
    // every pixel owns a slot for the whole frame (no re-assignment: the queue order stays pixel-coherent and the frame
    // takes max-path-length waves; +10 % over 2^20 slots at 1080p).  slots_override caps the slots of the whole frame.
    // (Binning the queue by direction octant was tried and lost 2.5 %.)
    std::vector<WaveBuffers> wbs(G);
    std::vector<uint32_t> slot_base(G + 1, 0u);
    for (int g = 0; g < G; g++) {
        const uint32_t gt = (n_tiles - (uint32_t) g + (uint32_t) G - 1u) / (uint32_t) G;
        uint32_t ns = gt * 32u;
        if (slots_override) ns = std::min(ns, std::max(32u, (slots_override / (uint32_t) G + 31u) & ~31u));
        wbs[g].n_pixels = gt * 32u;
        wbs[g].n_slots = ns;
        slot_base[g + 1] = slot_base[g] + ns;
    }
    reserve(slot_base[G]);
    counters_.reserve((size_t) G * CW);
    h_counts_.reserve((size_t) G * BATCH);
    if (!fork_event_) CUDA_CHECK(cudaEventCreateWithFlags(&fork_event_, cudaEventDisableTiming));
    for (int g = 0; g < G; g++) {
        if (!g_stream_[g]) CUDA_CHECK(cudaStreamCreateWithFlags(&g_stream_[g], cudaStreamNonBlocking));
        if (!g_event_[g]) CUDA_CHECK(cudaEventCreateWithFlags(&g_event_[g], cudaEventDisableTiming));
        WaveBuffers &wb = wbs[g];
        const uint32_t b = slot_base[g];
        wb.ray_o = ray_o_.get() + b; wb.ray_d = ray_d_.get() + b; wb.hit = hit_.get() + b; wb.hit_inst = hit_inst_.get() + b;
        wb.thr = thr_.get() + b; wb.acc = acc_.get() + b; wb.rng_a = rng_a_.get() + b; wb.rng_b = rng_b_.get() + b;
        wb.queue = queue_.get() + b;
        uint32_t *c = counters_.get() + (size_t) g * CW;
        wb.counts = c; wb.cursors = c + BATCH; wb.next_pixel = c + 2 * BATCH;
        wb.tiles_x = tiles_x; wb.group = (uint32_t) g; wb.n_groups = (uint32_t) G;
    }

    if (trace_blocks_per_sm_[0] == 0) {
        CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&trace_blocks_per_sm_[0], wave_trace2_kernel<false>, TRACE_THREADS, 0));
        CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&trace_blocks_per_sm_[1], wave_trace2_kernel<true>, TRACE_THREADS, 0));
        CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&trace_blocks_per_sm_[2], wave_trace8_kernel<false>, TRACE_THREADS, 0));
        CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&trace_blocks_per_sm_[3], wave_trace8_kernel<true>, TRACE_THREADS, 0));
    }
    const bool count = fr.count_traversal != 0;
    const bool wide = mode == 1;
    int bps = std::max(1, trace_blocks_per_sm_[(wide ? 2 : 0) + (count ? 1 : 0)]);
    if (blocks_per_sm_override > 0) bps = std::min(blocks_per_sm_override, bps);
    // the resident trace warps are shared out among the groups (at least one block per SM each)
    const int trace_blocks = sm_count * std::max(1, (bps + G - 1) / G);

    stats = WaveStats{};
    CUDA_CHECK(cudaMemsetAsync(counters_.get(), 0, (size_t) G * CW * sizeof(uint32_t), stream));
    CUDA_CHECK(cudaMemsetAsync(acc_.get(), 0xFE, (size_t) slot_base[G] * sizeof(float4), stream));     // SLOT_FRESH in every .w
    CUDA_CHECK(cudaEventRecord(fork_event_, stream));
    std::vector<char> first(G, 1), pending(G, 0), done(G, 0);
    std::vector<uint32_t> waves(G, 0u);
    // One batch = [zero the batch's counters] + BATCH x (shade, trace) + [counts -> pinned host].  Every batch of a group
    // is the same work with the same arguments, so it is captured ONCE into a CUDA graph and replayed: a frame is a few
    // hundred graph launches instead of several thousand kernel launches (the host launch rate bounded small shards).
    struct BatchKey { SceneD sc; CameraD cam; FrameD fr; WaveBuffers wb; int trace_blocks, refill_min, node_min, wide, count; uint32_t *h; };
    auto launch_waves = [&](int g, cudaStream_t s) {
        const WaveBuffers &wb = wbs[g];
        const unsigned shade_blocks = div_up(wb.n_slots, SHADE_THREADS);
        CUDA_CHECK(cudaMemsetAsync(wb.counts, 0, 2 * BATCH * sizeof(uint32_t), s));
        for (int w = 0; w < BATCH; w++) {
            if (wide) wave_shade_kernel<true><<<shade_blocks, SHADE_THREADS, 0, s>>>(sc, cam, fr, wb, (uint32_t) w);
            else wave_shade_kernel<false><<<shade_blocks, SHADE_THREADS, 0, s>>>(sc, cam, fr, wb, (uint32_t) w);
            if (wide) {
                if (count) wave_trace8_kernel<true><<<trace_blocks, TRACE_THREADS, 0, s>>>(sc, wb, (uint32_t) w, refill_min, node_min, fr.ray_counter);
                else wave_trace8_kernel<false><<<trace_blocks, TRACE_THREADS, 0, s>>>(sc, wb, (uint32_t) w, refill_min, node_min, fr.ray_counter);
            } else {
                if (count) wave_trace2_kernel<true><<<trace_blocks, TRACE_THREADS, 0, s>>>(sc, wb, (uint32_t) w, refill_min, node_min, fr.ray_counter);
                else wave_trace2_kernel<false><<<trace_blocks, TRACE_THREADS, 0, s>>>(sc, wb, (uint32_t) w, refill_min, node_min, fr.ray_counter);
            }
        }
        CUDA_CHECK(cudaMemcpyAsync(h_counts_.get() + (size_t) g * BATCH, wb.counts, BATCH * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
    };
    auto enqueue_batch = [&](int g) {
        cudaStream_t s = g_stream_[g];
        if (first[g]) {
            CUDA_CHECK(cudaStreamWaitEvent(s, fork_event_, 0));
            first[g] = 0;
            BatchKey key;
            memset(&key, 0, sizeof(key));
            key.sc = sc; key.cam = cam; key.fr = fr; key.wb = wbs[g];
            key.trace_blocks = trace_blocks; key.refill_min = refill_min; key.node_min = node_min; key.wide = wide; key.count = count;
            key.h = h_counts_.get();
            if (g_key_[g].size() != sizeof(key) || memcmp(g_key_[g].data(), &key, sizeof(key)) != 0) {
                if (g_exec_[g]) { cudaGraphExecDestroy(g_exec_[g]); g_exec_[g] = nullptr; }
                cudaGraph_t graph = nullptr;
                CUDA_CHECK(cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
                launch_waves(g, s);
                CUDA_CHECK(cudaStreamEndCapture(s, &graph));
                CUDA_CHECK(cudaGraphInstantiate(&g_exec_[g], graph, 0));
                cudaGraphDestroy(graph);
                g_key_[g].assign(reinterpret_cast<const char *>(&key), reinterpret_cast<const char *>(&key) + sizeof(key));
            }
        }
        if (use_graphs) CUDA_CHECK(cudaGraphLaunch(g_exec_[g], s));
        else launch_waves(g, s);
        stats.launches += 2 * BATCH;
        CUDA_CHECK(cudaGetLastError());
        CUDA_CHECK(cudaEventRecord(g_event_[g], s));
        pending[g] = 1;
    };
    static const bool dump_waves = getenv("RTB200_DUMP_WAVES") != nullptr;     // debugging aid: rays per wave of group 0
    int n_done = 0;
    while (n_done < G) {
        for (int g = 0; g < G; g++) if (!done[g] && !pending[g]) enqueue_batch(g);
        // take whichever group finishes its batch first, so that no group waits for another one's host round trip
        int ready = -1;
        while (ready < 0) {
            for (int g = 0; g < G && ready < 0; g++) {
                if (!pending[g]) continue;
                const cudaError_t q = cudaEventQuery(g_event_[g]);
                if (q == cudaSuccess) ready = g;
                else if (q != cudaErrorNotReady) CUDA_CHECK(q);
            }
        }
        pending[ready] = 0;
        const uint32_t *hc = h_counts_.get() + (size_t) ready * BATCH;
        for (int w = 0; w < BATCH; w++) {
            stats.rays += hc[w];
            if (dump_waves && ready == 0) fprintf(stderr, "wave %u rays %u\n", waves[ready], hc[w]);
            if (hc[w] == 0u) { done[ready] = 1; n_done++; break; }
            waves[ready]++;
        }
        if (!done[ready] && finish_divisor > 0 && hc[BATCH - 1] <= wbs[ready].n_slots / (uint32_t) finish_divisor) {
            // few slots left: one thread per remaining slot runs it to the end of its pixel (rays counted on the device)
            const WaveBuffers &wb = wbs[ready];
            const unsigned blocks = div_up(hc[BATCH - 1], 128);
            cudaStream_t s = g_stream_[ready];
            if (wide) {
                if (count) wave_finish_kernel<true, true><<<blocks, 128, 0, s>>>(sc, cam, fr, wb, (uint32_t) (BATCH - 1));
                else wave_finish_kernel<false, true><<<blocks, 128, 0, s>>>(sc, cam, fr, wb, (uint32_t) (BATCH - 1));
            } else {
                if (count) wave_finish_kernel<true, false><<<blocks, 128, 0, s>>>(sc, cam, fr, wb, (uint32_t) (BATCH - 1));
                else wave_finish_kernel<false, false><<<blocks, 128, 0, s>>>(sc, cam, fr, wb, (uint32_t) (BATCH - 1));
            }
            CUDA_CHECK(cudaGetLastError());
            CUDA_CHECK(cudaEventRecord(g_event_[ready], s));
            stats.launches++;
            done[ready] = 1;
            n_done++;
        }
    }
    for (int g = 0; g < G; g++) {
        stats.waves = std::max(stats.waves, waves[g]);
        CUDA_CHECK(cudaStreamWaitEvent(stream, g_event_[g], 0));    // join: the caller's stream continues after every group
    }
}
