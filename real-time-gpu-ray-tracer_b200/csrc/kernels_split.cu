// kernels_split.cu — the path kernel with the SHADE step moved to a warp of its own (A/B variant of kernels_persist.cu).
//
// In persist8_kernel a lane whose ray has finished waits until eight lanes of its warp want to shade, and the shade step then
// runs at 4-5 of 32 lanes (29 % of the warp instructions, profiles/r02a_persist8_c2_hotlines.txt) while the lanes that wait
// thin out the node and leaf steps.  Here a CTA is seven TRACE warps and one SHADE warp over a pool of more paths than trace
// lanes (320 for 224): a trace lane that finishes a ray writes the hit to the path's state, pushes the path onto a
// shared-memory queue and at once takes the next READY path from a second queue; the shade warp pops finished paths 32 at a
// time — hit record, scatter, sample / pixel bookkeeping, next ray — and pushes them back as ready.  A path still consumes its
// pixel's XORWOW stream strictly in order (it is in exactly one place at a time), so every pixel is bit-identical to the other
// kernels'; only the schedule differs.
#include "wave.h"

#include "rt_pixel.h"
#include "wave_device.h"

namespace {

constexpr int SPLIT_THREADS = 256, TRACE_THREADS = 224, SPLIT_SLOTS = 320, QN = 512;
constexpr uint32_t NONE = 0xFFFFFFFFu, QMASK = QN - 1;
constexpr uint32_t PID_MASK = 0x0FFFFFFFu, NG_INNER = 0x00FFFFFFu;

struct SplitQueues {
    unsigned short ready_q[QN], shade_q[QN];     // path ids of the CTA; shade_q holds id + 1 (0 = not written yet)
    uint32_t ready_head, ready_tail, shade_tail, done;
};

// ---- the shade warp ------------------------------------------------------------------------------------------------------
template<bool COUNT>
__device__ __forceinline__ void split_shade_warp(const SceneD &sc, const CameraD &cam, const FrameD &fr, const PersistBuffers &pb, SplitQueues &q) {
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t gbase = blockIdx.x * (uint32_t) SPLIT_SLOTS;
    volatile uint32_t *v_ready_head = &q.ready_head, *v_ready_tail = &q.ready_tail, *v_shade_tail = &q.shade_tail;
    volatile unsigned short *v_shade_q = q.shade_q;
    uint32_t shade_head = 0, ready_tail = 0, fresh_next = 0, dead = 0, rays = 0;
    for (;;) {
        uint32_t slot = NONE;
        bool fresh = false;
        if (fresh_next < (uint32_t) SPLIT_SLOTS) {                    // the pool starts empty: every path takes its first pixel
            if (fresh_next + lane < (uint32_t) SPLIT_SLOTS) { slot = fresh_next + lane; fresh = true; }
            fresh_next += 32u;
        } else {
            const uint32_t avail = *v_shade_tail - shade_head;
            const bool starving = (int) (*v_ready_head - *v_ready_tail) > 0;     // a trace lane holds a ticket nobody has served yet
            if (avail >= 32u || (avail > 0u && starving)) {
                const uint32_t n = avail < 32u ? avail : 32u;
                if (lane < n) {
                    const uint32_t idx = (shade_head + lane) & QMASK;
                    uint32_t e;
                    while ((e = v_shade_q[idx]) == 0u) { }            // reserved by its producer, written a moment later
                    v_shade_q[idx] = 0;
                    slot = e - 1u;
                }
                shade_head += n;
            } else {
                if (dead == (uint32_t) SPLIT_SLOTS) break;            // every path has run out of pixels
                __nanosleep(40);
                continue;
            }
        }
        // ---- the per-path work of persist8_kernel's SHADE step ------------------------------------------------------------
        const bool active = slot != NONE;
        const uint32_t g = gbase + (active ? slot : 0u);
        uint32_t pid = SLOT_IDLE;
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        RngState rng{};
        RayD ray{};
        f3 thr = mk3(1.f, 1.f, 1.f);
        uint32_t sample = 0, depth = 0, pix_rays = 0;
        bool need_pixel = active && fresh, need_primary = false, acc_dirty = false, alive_path = active;
        if (active && !fresh) {
            acc = __ldcg(&pb.acc[g]);
            pid = __float_as_uint(acc.w);
            const float4 o = __ldcg(&pb.ray_o[g]), d = __ldcg(&pb.ray_d[g]), t4 = __ldcg(&pb.thr[g]), h4 = __ldcg(&pb.hit[g]);
            const uint4 ra = __ldcg(&pb.rng_a[g]);
            const uint2 rb = __ldcg(&pb.rng_b[g]);
            HitD hit;
            hit.t = h4.x; hit.u = h4.y; hit.v = h4.z; hit.tc = h4.x; hit.leaf = (int) __float_as_uint(h4.w); hit.inst = __ldcg(&pb.hit_inst[g]);
            rng.d = ra.x; rng.v0 = ra.y; rng.v1 = ra.z; rng.v2 = ra.w; rng.v3 = rb.x; rng.v4 = rb.y;
            ray.o = mk3(o.x, o.y, o.z); ray.d = mk3(d.x, d.y, d.z);
            thr = mk3(t4.x, t4.y, t4.z);
            const uint32_t sd = __float_as_uint(t4.w);
            sample = sd >> 16; depth = sd & 0xFFFFu;
            pix_rays = __float_as_uint(o.w);
            const uint32_t fw = pid >> 28;
            if (fr.aov_t && sample == 0u && depth == 0u) {
                uint32_t lx, ly;
                pixel_coords(pid & PID_MASK, pb.tiles_x, lx, ly);
                const size_t lp = (size_t) fw * pb.frame_px + ly * cam.width + lx;
                const bool has = hit.inst >= 0;
                fr.aov_t[lp] = has ? hit.t : RT_INF;
                fr.aov_inst[lp] = has ? hit.inst : -1;
                fr.aov_prim[lp] = has ? (int) ((ref_type(hit.leaf) << 28) | sc.leaf_prims8[ref_pos(hit.leaf)]) : -1;
            }
            const bool alive = shade_scatter<true>(sc, cam, ray, hit, thr, rng, fw * pb.inst_stride);
            depth++;
            if (!(alive && depth < cam.depth)) {
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
                    if (fw == 0u) atomicAdd(&pb.tile_cost[(pid & PID_MASK) >> 5], pix_rays);
                    if (pb.frame_done) {
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
        for (;;) {
            const uint32_t any = __ballot_sync(0xffffffffu, need_pixel);
            if (any == 0u) break;
            const uint32_t cand = warp_append(pb.next_pixel, need_pixel);
            if (need_pixel) {
                const uint32_t cf = cand / pb.n_pixels;
                if (cf >= pb.n_frames) {
                    alive_path = false; need_pixel = false;
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
        const bool go = active && alive_path;
        if (go) {
            if (need_primary) {
                uint32_t lx, ly;
                pixel_coords(pid & PID_MASK, pb.tiles_x, lx, ly);
                const uint32_t si = sample / cam.sqrtSamples, sj = sample % cam.sqrtSamples;
                ray = primary_ray(cam, lx, global_row(fr, ly), si, sj, rng);
                thr = mk3(1.f, 1.f, 1.f);
                depth = 0;
            }
            if (acc_dirty) __stcg(&pb.acc[g], acc);
            __stcg(&pb.ray_o[g], make_float4(ray.o.x, ray.o.y, ray.o.z, __uint_as_float(pix_rays + 1u)));
            __stcg(&pb.ray_d[g], make_float4(ray.d.x, ray.d.y, ray.d.z, __uint_as_float(pid >> 28)));
            __stcg(&pb.thr[g], make_float4(thr.x, thr.y, thr.z, __uint_as_float((sample << 16) | depth)));
            __stcg(&pb.rng_a[g], make_uint4(rng.d, rng.v0, rng.v1, rng.v2));
            __stcg(&pb.rng_b[g], make_uint2(rng.v3, rng.v4));
            rays++;
        }
        // ---- back to the trace warps: ready paths in ring order, then the tail moves (this warp is the only producer) -------
        const uint32_t live = __ballot_sync(0xffffffffu, go);
        dead += (uint32_t) __popc(__ballot_sync(0xffffffffu, active && !alive_path));
        if (live) {
            if (go) q.ready_q[(ready_tail + (uint32_t) __popc(live & ((1u << lane) - 1u))) & QMASK] = (unsigned short) slot;
            __threadfence_block();
            __syncwarp();
            ready_tail += (uint32_t) __popc(live);
            if (lane == 0u) *v_ready_tail = ready_tail;
        }
    }
    __syncwarp();
    if (lane == 0u) { __threadfence_block(); *(volatile uint32_t *) &q.done = 1u; }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) rays += __shfl_xor_sync(0xffffffffu, rays, off);
    if (lane == 0u && rays) atomicAdd(fr.ray_counter, (unsigned long long) rays);
}

// ---- a trace warp: persist8_kernel's NODE and LEAF steps; a lane without a path takes the next ready one ---------------------
template<bool COUNT>
__device__ __forceinline__ void split_trace_warp(const SceneD &sc, const FrameD &fr, const PersistBuffers &pb, SplitQueues &q, uint2 *stack_mem, int node_min) {
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t gbase = blockIdx.x * (uint32_t) SPLIT_SLOTS;
    uint2 *const stk = stack_mem + threadIdx.x;               // entry k at stk[k * TRACE_THREADS]
    volatile uint32_t *v_ready_tail = &q.ready_tail, *v_done = &q.done;
    volatile unsigned short *v_ready_q = q.ready_q, *v_shade_q = q.shade_q;
    uint32_t my = NONE, ticket = NONE;
    WideRay r{};
    uint2 ng = make_uint2(0u, 0u), tg = make_uint2(0u, 0u);
    int sp = 0, inst = -1, inst_sp = 0, prim_type = 0;
    HitD hit{};
    uint32_t overflow = 0;
    TravStats stats{0, 0, 0};
    for (;;) {
        // ---- a lane without a path: take a ticket, then the path the shade warp publishes under it ---------------------------
        const bool needs = my == NONE && ticket == NONE;
        const uint32_t want = __ballot_sync(0xffffffffu, needs);
        if (want) {
            const uint32_t leader = __ffs(want) - 1u;
            uint32_t base = 0;
            if (lane == leader) base = atomicAdd(&q.ready_head, (uint32_t) __popc(want));
            base = __shfl_sync(0xffffffffu, base, leader);
            if (needs) ticket = base + (uint32_t) __popc(want & ((1u << lane) - 1u));
        }
        if (my == NONE && (int) (*v_ready_tail - ticket) > 0) {
            my = v_ready_q[ticket & QMASK];
            ticket = NONE;
            const float4 o = __ldcg(&pb.ray_o[gbase + my]), d = __ldcg(&pb.ray_d[gbase + my]);
            r.ray.o = mk3(o.x, o.y, o.z); r.ray.d = mk3(d.x, d.y, d.z);
            wide_setup_ray(r);
            hit.t = RT_INF; hit.tc = RT_INF; hit.u = 0.f; hit.v = 0.f; hit.inst = -1; hit.leaf = 0;
            sp = 0; inst = -1 - (int) __float_as_uint(d.w); inst_sp = 0;
            ng = make_uint2(0u, 0x80000000u);                 // TLAS root: one pending inner hit at relative index 0
            tg = make_uint2(0u, 0u);
        }
        // ---- cheap transitions: leave instance / pop / ray finished ----------------------------------------------------------
        if (my != NONE && tg.y == 0u && ng.y <= NG_INNER) {
            if (inst >= 0 && sp == inst_sp && sp != 0) {
                const float4 o = __ldcg(&pb.ray_o[gbase + my]), d = __ldcg(&pb.ray_d[gbase + my]);
                inst = -1 - (int) __float_as_uint(d.w);
                r.ray.o = mk3(o.x, o.y, o.z); r.ray.d = mk3(d.x, d.y, d.z);
                wide_setup_ray(r);
            }
            if (sp == 0) {
                // finished: the hit goes to the path's state, the path to the shade warp
                __stcg(&pb.hit[gbase + my], make_float4(hit.t, hit.u, hit.v, __uint_as_float((uint32_t) hit.leaf)));
                __stcg(&pb.hit_inst[gbase + my], hit.inst);
                __threadfence_block();
                const uint32_t idx = atomicAdd(&q.shade_tail, 1u);
                v_shade_q[idx & QMASK] = (unsigned short) (my + 1u);
                my = NONE;
                ng = make_uint2(0u, 0u);
            } else {
                sp--;
                const uint2 e = stk[sp * TRACE_THREADS];
                if (e.y > NG_INNER) ng = e; else { tg = e; ng = make_uint2(0u, 0u); }
            }
        }
        const bool on_node = my != NONE && tg.y == 0u && ng.y > NG_INNER;
        const bool on_leaf = my != NONE && tg.y != 0u;
        const uint32_t node_lanes = __ballot_sync(0xffffffffu, on_node);
        const uint32_t leaf_lanes = __ballot_sync(0xffffffffu, on_leaf);
        if ((node_lanes | leaf_lanes) == 0u) {
            const uint32_t busy = __ballot_sync(0xffffffffu, my != NONE);        // lanes that took a path or popped a group this round
            if (busy) continue;
            if (*v_done) break;
            __nanosleep(32);
            continue;
        }
        if (__popc(node_lanes) >= node_min || leaf_lanes == 0u) {
            // ---- NODE ---------------------------------------------------------------------------------------------------
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
                const BvhNode8 n = load_node8(inst < 0 ? sc.tlas_nodes8 + (size_t) (uint32_t) (-1 - inst) * pb.tlas_node_stride : sc.blas_nodes8, base + rel);
                if (COUNT) stats.nodes++;
                const uint32_t hm = node8_intersect(n, r, RT_TMIN, hit.tc);
                ng.x = n.n1.x;
                ng.y = hm & 0xFF000000u;
                if (ng.y) ng.y |= __float_as_uint(n.n0.w) >> 24;
                tg.x = n.n1.y;
                tg.y = hm & 0x00FFFFFFu;
            }
        } else {
            // ---- LEAF: one primitive test or one instance entry per lane ------------------------------------------------------
            if (on_leaf) {
                const int idx = rt_bfind(tg.y);
                tg.y &= ~(1u << idx);
                if (inst >= 0) {
                    if (COUNT) stats.prims++;
                    leaf_test8(sc, (uint32_t) prim_type, tg.x + (uint32_t) idx, r.ray, RT_TMIN, inst, hit);
                } else {
                    if (tg.y) { if (sp < RT_WIDE_STACK) { stk[sp * TRACE_THREADS] = tg; sp++; } else overflow = 1u; }
                    if (ng.y > NG_INNER) { if (sp < RT_WIDE_STACK) { stk[sp * TRACE_THREADS] = ng; sp++; } else overflow = 1u; }
                    const uint32_t fw = (uint32_t) (-1 - inst);
                    inst = (int) __ldg(&sc.tlas_leaf_inst[(size_t) fw * pb.tlas_leaf_stride + tg.x + (uint32_t) idx]);
                    if (COUNT) stats.insts++;
                    float inv[12];
                    int root8;
                    load_instance_wide(sc, inst, inv, root8, prim_type, fw * pb.inst_stride);
                    const f3 wo = r.ray.o, wd = r.ray.d;
                    r.ray.o = xf_point(inv, wo);
                    r.ray.d = xf_vector(inv, wd);
                    wide_setup_ray(r);
                    inst_sp = sp;
                    ng = make_uint2((uint32_t) root8, 0x80000000u);
                    tg = make_uint2(0u, 0u);
                }
            }
        }
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) overflow |= __shfl_xor_sync(0xffffffffu, overflow, off);
    if (lane == 0u && overflow) atomicOr(fr.overflow_flag, 1u);
    if (COUNT) {
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            stats.nodes += __shfl_xor_sync(0xffffffffu, stats.nodes, off);
            stats.prims += __shfl_xor_sync(0xffffffffu, stats.prims, off);
            stats.insts += __shfl_xor_sync(0xffffffffu, stats.insts, off);
        }
        if (lane == 0u) {
            atomicAdd(fr.ray_counter + 2, (unsigned long long) stats.nodes);
            atomicAdd(fr.ray_counter + 3, (unsigned long long) stats.prims);
            atomicAdd(fr.ray_counter + 4, (unsigned long long) stats.insts);
        }
    }
}

template<bool COUNT>
__global__ void __launch_bounds__(SPLIT_THREADS, 3) persist8_split_kernel(const __grid_constant__ SceneD sc, const __grid_constant__ CameraD cam,
                                                                         const __grid_constant__ FrameD fr, const __grid_constant__ PersistBuffers pb,
                                                                         int node_min) {
    __shared__ uint2 stack_mem[RT_WIDE_STACK * TRACE_THREADS];
    __shared__ SplitQueues q;
    if (threadIdx.x == 0) { q.ready_head = 0u; q.ready_tail = 0u; q.shade_tail = 0u; q.done = 0u; }
    for (int i = threadIdx.x; i < QN; i += SPLIT_THREADS) { q.shade_q[i] = 0; q.ready_q[i] = 0; }
    __syncthreads();
    if (threadIdx.x >= TRACE_THREADS) split_shade_warp<COUNT>(sc, cam, fr, pb, q);
    else split_trace_warp<COUNT>(sc, fr, pb, q, stack_mem, node_min);
}

} // namespace

int split_slots_per_block() { return SPLIT_SLOTS; }

int split_blocks_per_sm(bool count) {
    int b = 0;
    if (count) CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, persist8_split_kernel<true>, SPLIT_THREADS, 0));
    else CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, persist8_split_kernel<false>, SPLIT_THREADS, 0));
    return b < 1 ? 1 : b;
}

void launch_persist_split(const SceneD &sc, const CameraD &cam, const FrameD &fr, const PersistBuffers &pb, uint32_t blocks, int node_min, bool count,
                          cudaStream_t stream) {
    if (count) persist8_split_kernel<true><<<blocks, SPLIT_THREADS, 0, stream>>>(sc, cam, fr, pb, node_min);
    else persist8_split_kernel<false><<<blocks, SPLIT_THREADS, 0, stream>>>(sc, cam, fr, pb, node_min);
    CUDA_CHECK(cudaGetLastError());
}
