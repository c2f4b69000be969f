// kernels_wide.cu — collapse of the binary LBVH into the compressed 8-wide BVH (BvhNode8), on the device.
//
// One persistent launch.  Work item i = wide node i; the thread that encodes node i allocates consecutive node
// slots for its inner children and consecutive leaf-order positions for its leaf primitives (two atomic
// counters) and publishes the children's binary roots into queue[child_base ..].  Items are CLAIMED in increasing
// order from a ticket counter (one warp-aggregated atomic per 32 items) and a thread polls queue[i] of the item it
// claimed until the producer has written it.  The producer of item i is an item j < i, and every ticket below a
// claimed one was claimed earlier, by a thread that is running or has finished — so a wait always ends, whatever
// part of the grid is resident (no co-residency assumption: other tenants of the GPU, concurrent builds of other
// contexts and partially scheduled grids only make it slower).  Termination: finished == allocated.
// The node contents are a pure function of the binary tree (rt_wide_core.h); only the memory layout depends on
// the order of the atomics, and the parity tests compare the trees canonically against the serial twin.
#include "build.h"

#include <cstdlib>
#include <cstring>

namespace {

constexpr int CT = 128;
constexpr int QUEUE_EMPTY = (int) 0x80808080u;     // cudaMemsetAsync(0x80) pattern; never a node index or ~position

__device__ __forceinline__ int find_segment_w(const uint32_t *__restrict__ seg_offset, int n_segs, uint32_t j) {
    int lo = 0, hi = n_segs - 1;
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (__ldg(&seg_offset[mid]) <= j) lo = mid; else hi = mid - 1;
    }
    return lo;
}

__global__ void __launch_bounds__(256) collapse_init_kernel(const uint32_t *__restrict__ seg_offset, const int *__restrict__ seg_root, int n_segs,
                                                            int *queue, uint32_t *counters) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s == 0) {
        counters[0] = (uint32_t) n_segs; counters[1] = 0u; counters[2] = 0u; counters[3] = 0u;
        counters[4] = 0u; counters[5] = (uint32_t) n_segs; counters[6] = 0u;      // level launches: begin, end, blocks done
    }
    if (s >= n_segs) return;
    const uint32_t first = seg_offset[s], cnt = seg_offset[s + 1] - first;
    queue[s] = cnt == 1u ? ~(int) first : seg_root[s];
}

// one work item = one wide node: expand, allocate, publish the children, encode, write
__device__ __forceinline__ void collapse_item(const BinaryTreeView &view, int root, int i, int leaf_max, volatile int *vq, uint32_t *counters,
                                              BvhNode8 *__restrict__ nodes8, uint32_t *__restrict__ leaf_pos8) {
    WideChildren w;
    wide_expand(view, root, leaf_max, w);
    const int k = wide_inner_count(w), m = wide_prim_count(w);
    const uint32_t child_base = k ? atomicAdd(&counters[0], (uint32_t) k) : 0u;
    const uint32_t prim_base = m ? atomicAdd(&counters[1], (uint32_t) m) : 0u;
    Aabb box;
    int child_in_slot[8];
    wide_layout(w, box, child_in_slot);
    // the children first: the collapse's dependency chain runs through these stores, not through the record
    uint32_t j = 0;
    for (int sl = 0; sl < 8; sl++) {
        const int c = child_in_slot[sl];
        if (c >= 0 && w.ref[c] >= 0) { vq[child_base + j] = w.ref[c]; j++; }
    }
    BvhNode8 node;
    wide_encode_layout(w, box, child_in_slot, child_base, prim_base, node);
    nodes8[i] = node;
    uint32_t off = 0;
    for (int sl = 0; sl < 8; sl++) {
        const int c = child_in_slot[sl];
        if (c < 0 || w.ref[c] >= 0) continue;
        for (int q = 0; q < w.count[c]; q++) leaf_pos8[prim_base + off + q] = (uint32_t) (w.first[c] + q);
        off += (uint32_t) w.count[c];
    }
}

// ---- the first levels, one launch per level --------------------------------------------------------------------------
// Items are numbered in allocation order, so the nodes of depth L occupy one contiguous index range, known once depth
// L - 1 is finished.  A launch that covers exactly that range has every item ready when it starts: no polling, every SM full
// of warps, the DRAM latency of the dependent binary-node loads hidden by occupancy instead of sitting on a chain of waiting
// threads (the ticket kernel alone: 8 warps per SM, issue slots 10 % busy, 375 us for 1 M triangles).  The host does not
// know the ranges: each launch is sized for the largest range its level can have and reads [begin, end) from `levels`, which
// the last block of the previous level wrote.  Deep or degenerate trees simply leave their remaining levels to the ticket
// kernel, which runs afterwards in any case and starts at the first index no level launch covered.
//   levels[0] = current begin, levels[1] = current end, levels[2] = blocks finished in this launch
__global__ void __launch_bounds__(256) collapse_level_kernel(BinaryTreeView view, int leaf_max, int *queue, uint32_t *counters, uint32_t *levels,
                                                             BvhNode8 *__restrict__ nodes8, uint32_t *__restrict__ leaf_pos8) {
    volatile int *vq = queue;
    const uint32_t begin = ((volatile uint32_t *) levels)[0], end = ((volatile uint32_t *) levels)[1];
    const uint32_t i = begin + blockIdx.x * blockDim.x + threadIdx.x;
    if (i < end) {
        collapse_item(view, vq[i], (int) i, leaf_max, vq, counters, nodes8, leaf_pos8);
        atomicAdd(&counters[2], 1u);
    }
    __shared__ bool last;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) last = atomicAdd(&levels[2], 1u) + 1u == gridDim.x;
    __syncthreads();
    if (last && threadIdx.x == 0) {
        __threadfence();
        const uint32_t alloc = ((volatile uint32_t *) counters)[0];       // every item of this level has allocated its children
        levels[0] = end;
        levels[1] = alloc;
        levels[2] = 0u;
        counters[3] = end;                                                // where the ticket kernel would have to start
        __threadfence();
    }
}

__global__ void __launch_bounds__(CT) collapse_kernel(BinaryTreeView view, int leaf_max, int capacity, int *queue, uint32_t *counters,
                                                      BvhNode8 *__restrict__ nodes8, uint32_t *__restrict__ leaf_pos8) {
    volatile int *vq = queue;
    volatile uint32_t *vc = counters;
    const uint32_t lane = threadIdx.x & 31u;
    for (;;) {
        // ---- claim the next 32 items for this warp (tickets only grow: everything below them is in someone's hands) ----
        uint32_t base = 0;
        if (lane == 0u) base = atomicAdd(&counters[3], 32u);
        base = __shfl_sync(0xffffffffu, base, 0);
        const int i = (int) (base + lane);
        bool finished_all = false;
        uint32_t backoff = 32u;
        bool allocated = false;                    // item i is known to be allocated (so it will be published)
        for (;;) {
            if (i < capacity) {
                const int root = vq[i];
                if (root != QUEUE_EMPTY) {
                    collapse_item(view, root, i, leaf_max, vq, counters, nodes8, leaf_pos8);
                    // no fence: nothing inside this launch reads what the item wrote except the queue entries (single words, polled
                    // through volatile loads); `finished` only ends the launch, and the end of the launch publishes everything
                    atomicAdd(&counters[2], 1u);
                    break;
                }
            }
            // Not published yet.  An index below the allocation count WILL be published (its parent is in someone's hands):
            // poll only the queue.  Only threads whose index is not allocated (yet) look at the termination condition — the
            // counters are one hot line that every polling thread used to hammer on every iteration.
            if (!allocated) allocated = i < capacity && (uint32_t) i < vc[0];       // the count only grows: ask once
            if (allocated) {
                __nanosleep(backoff);
                if (backoff < 256u) backoff <<= 1;
                continue;
            }
            const uint32_t fin = vc[2];
            __threadfence();
            const uint32_t alloc = vc[0];
            if (fin == alloc) { finished_all = true; break; }
            __nanosleep(backoff);                  // waiting lanes must not starve the producers of L2 bandwidth
            if (backoff < 2048u) backoff <<= 1;
        }
        // a lane saw finished == allocated: every node is written, and the lanes of this warp that held real items are done
        if (__any_sync(0xffffffffu, finished_all)) break;
    }
}

// ---- the same collapse with EIGHT LANES PER ITEM ("octets", four items per warp) --------------------------------------
// The one-thread-per-item kernel above keeps the eight children of a node in dynamically indexed local arrays (528 bytes
// of stack): every argmax / slot / quantisation step is a chain of local-memory round trips, about 10 us per item, and the
// collapse is a dependency chain of items (seven levels for 1 M triangles, then several passes over the widest level).
// Here lane c of an octet holds child c in registers; argmax, union, prefix counts and the OR of the quantised bytes are
// 3-step shuffles inside the octet.  The arithmetic per child is that of rt_wide_core.h (wide_set_child, wide_assign_slots,
// wide_quant_*), executed by the child's lane, so the nodes are the ones the serial twin makes (tests compare them).
struct OctChild {
    int ref, first, count, kl, kr;      // as WideChildren::ref/first/count/kl/kr of one child
    Aabb box;
    bool valid;
};
__device__ __forceinline__ void oct_set_child(const BinaryTreeView &t, int child, int leaf_max, OctChild &c) {
    c.valid = true;
    if (child >= 0) {
        const float4 a = t.node_lo[child], b = t.node_hi[child];
        c.box.lo = mk3(a.x, a.y, a.z);
        c.box.hi = mk3(b.x, b.y, b.z);
        c.kl = (int) f2u(a.w);
        c.kr = (int) f2u(b.w);
        int first = 0, cnt = leaf_max + 1;
        if (leaf_max > 1) {
            const int2 rg = t.range[child];
            first = rg.x;
            cnt = rg.y - rg.x + 1;
        }
        c.first = first;
        c.count = cnt;
        c.ref = cnt <= leaf_max ? -1 : child;
    } else {
        const int pos = ~child;
        c.box = view_box(t.prim_lo, t.prim_hi, (int) t.vals[pos]);
        c.first = pos;
        c.count = 1;
        c.ref = -1;
        c.kl = 0; c.kr = 0;
    }
}
// argmax over the octet: largest value, ties to the lower index (the serial loops take the first maximum)
__device__ __forceinline__ void oct_argmax(uint32_t om, float &v, int &i) {
#pragma unroll
    for (int off = 1; off < 8; off <<= 1) {
        const float ov = __shfl_xor_sync(om, v, off, 8);
        const int oi = __shfl_xor_sync(om, i, off, 8);
        if (ov > v || (ov == v && oi < i)) { v = ov; i = oi; }
    }
}
__device__ __forceinline__ uint32_t oct_or(uint32_t om, uint32_t v) {
#pragma unroll
    for (int off = 1; off < 8; off <<= 1) v |= __shfl_xor_sync(om, v, off, 8);
    return v;
}
__device__ __forceinline__ float oct_min(uint32_t om, float v) {
#pragma unroll
    for (int off = 1; off < 8; off <<= 1) v = fminf(v, __shfl_xor_sync(om, v, off, 8));
    return v;
}
__device__ __forceinline__ float oct_max(uint32_t om, float v) {
#pragma unroll
    for (int off = 1; off < 8; off <<= 1) v = fmaxf(v, __shfl_xor_sync(om, v, off, 8));
    return v;
}

// one item by one octet; `l` = lane within the octet (child index), `om` = the octet's lane mask
__device__ __forceinline__ void oct_collapse_item(const BinaryTreeView &view, int root, int item, int leaf_max, uint32_t om, int l, volatile int *vq,
                                                  uint32_t *counters, BvhNode8 *__restrict__ nodes8, uint32_t *__restrict__ leaf_pos8) {
    // ---- wide_expand: children of the node, child c in lane c -------------------------------------------------------------
    OctChild c{};
    c.valid = false; c.ref = -1; c.count = 0;
    int n = 1;
    if (l == 0) oct_set_child(view, root, leaf_max, c);
    if (root >= 0) {
        const int ref0 = __shfl_sync(om, c.ref, 0, 8);
        if (ref0 >= 0) {
            const int kl = __shfl_sync(om, c.kl, 0, 8), kr = __shfl_sync(om, c.kr, 0, 8);
            n = 2;
            if (l == 0) oct_set_child(view, kl, leaf_max, c);
            if (l == 1) oct_set_child(view, kr, leaf_max, c);
            while (n < 8) {
                float a = (c.valid && c.ref >= 0) ? aabb_half_area(c.box) : -1.0f;
                int best = l;
                oct_argmax(om, a, best);
                if (!(a > -1.0f)) break;                      // nothing left to open
                const int bl = __shfl_sync(om, c.kl, best, 8), br = __shfl_sync(om, c.kr, best, 8);
                if (l == best) oct_set_child(view, bl, leaf_max, c);
                if (l == n) oct_set_child(view, br, leaf_max, c);
                n++;
            }
        }
    }
    // ---- counts and allocation ----------------------------------------------------------------------------------------
    const bool inner = c.valid && c.ref >= 0, leaf = c.valid && c.ref < 0;
    const uint32_t inner_lanes = (__ballot_sync(om, inner) >> (__ffs(om) - 1)) & 0xFFu;
    const int k = __popc(inner_lanes);
    int m = leaf ? c.count : 0;
#pragma unroll
    for (int off = 1; off < 8; off <<= 1) m += __shfl_xor_sync(om, m, off, 8);
    uint32_t child_base = 0u, prim_base = 0u;
    if (l == 0) {
        if (k) child_base = atomicAdd(&counters[0], (uint32_t) k);
        if (m) prim_base = atomicAdd(&counters[1], (uint32_t) m);
    }
    child_base = __shfl_sync(om, child_base, 0, 8);
    prim_base = __shfl_sync(om, prim_base, 0, 8);
    // ---- wide_layout: node box, octant-ordered slots (wide_assign_slots) ----------------------------------------------------
    Aabb node;
    node.lo = mk3(oct_min(om, c.valid ? c.box.lo.x : RT_INF), oct_min(om, c.valid ? c.box.lo.y : RT_INF), oct_min(om, c.valid ? c.box.lo.z : RT_INF));
    node.hi = mk3(oct_max(om, c.valid ? c.box.hi.x : -RT_INF), oct_max(om, c.valid ? c.box.hi.y : -RT_INF), oct_max(om, c.valid ? c.box.hi.z : -RT_INF));
    const f3 nc = mk3(0.5f * (node.lo.x + node.hi.x), 0.5f * (node.lo.y + node.hi.y), 0.5f * (node.lo.z + node.hi.z));
    f3 d = mk3(0.0f, 0.0f, 0.0f);
    float prio = -1.0f;
    if (c.valid) {
        d = mk3(0.5f * (c.box.lo.x + c.box.hi.x) - nc.x, 0.5f * (c.box.lo.y + c.box.hi.y) - nc.y, 0.5f * (c.box.lo.z + c.box.hi.z) - nc.z);
        prio = (fabsf(d.x) + fabsf(d.y)) + fabsf(d.z);
    }
    int slot = -1;
    uint32_t slot_free = 0xFFu;
    bool pending = c.valid;
    for (int round = 0; round < n; round++) {
        float bp = pending ? prio : -3.0f;
        int bc = l;
        oct_argmax(om, bp, bc);
        int bs = -1;
        if (l == bc) {
            float best = -RT_INF;
            for (int s = 0; s < 8; s++) {
                if (!(slot_free & (1u << s))) continue;
                const float cost = (((s & 4) ? d.x : -d.x) + ((s & 2) ? d.y : -d.y)) + ((s & 1) ? d.z : -d.z);
                if (cost > best) { best = cost; bs = s; }
            }
            slot = bs;
            pending = false;
        }
        bs = __shfl_sync(om, bs, bc, 8);
        slot_free &= ~(1u << bs);
    }
    // ---- positions among the siblings: inner children are numbered, leaf primitives laid out, in slot order -------------------
    uint32_t j = 0, off = 0;
#pragma unroll
    for (int o = 0; o < 8; o++) {
        const int os = __shfl_sync(om, slot, o, 8);
        const int oc = __shfl_sync(om, leaf ? c.count : 0, o, 8);
        const int oi = __shfl_sync(om, inner ? 1 : 0, o, 8);
        if (os >= 0 && os < slot) { j += (uint32_t) oi; off += (uint32_t) oc; }
    }
    // the children first: the collapse's dependency chain runs through these stores, not through the record
    if (inner) vq[child_base + j] = c.ref;
    // ---- wide_encode_layout: each lane quantises its own child; the words are OR-ed over the octet --------------------------
    const uint32_t ex = wide_grid_exponent(node.hi.x - node.lo.x), ey = wide_grid_exponent(node.hi.y - node.lo.y), ez = wide_grid_exponent(node.hi.z - node.lo.z);
    uint32_t imask = 0, meta[2] = {0, 0}, q[6][2] = {{0, 0}, {0, 0}, {0, 0}, {0, 0}, {0, 0}, {0, 0}};
    if (c.valid) {
        const float sx = wide_grid_step(ex), sy = wide_grid_step(ey), sz = wide_grid_step(ez);
        const float ix = wide_grid_inv_step(ex), iy = wide_grid_inv_step(ey), iz = wide_grid_inv_step(ez);
        uint32_t mm;
        if (inner) { imask = 1u << slot; mm = (1u << 5) | (24u + (uint32_t) slot); }
        else { const uint32_t unary = (1u << c.count) - 1u; mm = (unary << 5) | off; }
        const bool hi_half = slot >= 4;                   // slots 0..3 -> word 0, 4..7 -> word 1 (no dynamic indexing: registers)
        const int sh = 8 * (slot & 3);
        const uint32_t v[7] = {mm << sh,
                               wide_quant_lo(c.box.lo.x, node.lo.x, sx, ix) << sh, wide_quant_lo(c.box.lo.y, node.lo.y, sy, iy) << sh,
                               wide_quant_lo(c.box.lo.z, node.lo.z, sz, iz) << sh, wide_quant_hi(c.box.hi.x, node.lo.x, sx, ix) << sh,
                               wide_quant_hi(c.box.hi.y, node.lo.y, sy, iy) << sh, wide_quant_hi(c.box.hi.z, node.lo.z, sz, iz) << sh};
        meta[0] = hi_half ? 0u : v[0]; meta[1] = hi_half ? v[0] : 0u;
#pragma unroll
        for (int a = 0; a < 6; a++) { q[a][0] = hi_half ? 0u : v[a + 1]; q[a][1] = hi_half ? v[a + 1] : 0u; }
    }
    imask = oct_or(om, imask);
    meta[0] = oct_or(om, meta[0]); meta[1] = oct_or(om, meta[1]);
#pragma unroll
    for (int a = 0; a < 6; a++) { q[a][0] = oct_or(om, q[a][0]); q[a][1] = oct_or(om, q[a][1]); }
    uint4 *dst = reinterpret_cast<uint4 *>(nodes8 + item);
    if (l == 0) dst[0] = make_uint4(f2u(node.lo.x), f2u(node.lo.y), f2u(node.lo.z), ex | (ey << 8) | (ez << 16) | (imask << 24));
    if (l == 1) dst[1] = make_uint4(child_base, prim_base, meta[0], meta[1]);
    if (l == 2) dst[2] = make_uint4(q[0][0], q[0][1], q[1][0], q[1][1]);
    if (l == 3) dst[3] = make_uint4(q[2][0], q[2][1], q[3][0], q[3][1]);
    if (l == 4) dst[4] = make_uint4(q[4][0], q[4][1], q[5][0], q[5][1]);
    if (leaf) for (int t = 0; t < c.count; t++) leaf_pos8[prim_base + off + (uint32_t) t] = (uint32_t) (c.first + t);
    if (l == 0) atomicAdd(&counters[2], 1u);           // no fence (see collapse_kernel)
}

__global__ void __launch_bounds__(CT) collapse_octet_kernel(BinaryTreeView view, int leaf_max, int capacity, int *queue, uint32_t *counters,
                                                            BvhNode8 *__restrict__ nodes8, uint32_t *__restrict__ leaf_pos8) {
    volatile int *vq = queue;
    volatile uint32_t *vc = counters;
    const uint32_t lane = threadIdx.x & 31u;
    const int l = (int) (lane & 7u), octet = (int) (lane >> 3);
    const uint32_t om = 0xFFu << (8 * octet);
    for (;;) {
        // ---- claim the next 4 items for this warp, one per octet (tickets only grow: everything below them is in someone's hands)
        uint32_t base = 0;
        if (lane == 0u) base = atomicAdd(&counters[3], 4u);
        base = __shfl_sync(0xffffffffu, base, 0);
        const int i = (int) base + octet;
        bool finished_all = false, allocated = false;
        uint32_t backoff = 32u;
        for (;;) {
            int root = (int) 0x80808080u;                 // QUEUE_EMPTY
            if (i < capacity) root = vq[i];
            if (root != QUEUE_EMPTY) {
                oct_collapse_item(view, root, i, leaf_max, om, l, vq, counters, nodes8, leaf_pos8);
                break;
            }
            if (!allocated) allocated = i < capacity && (uint32_t) i < vc[0];       // the count only grows: ask once
            if (!allocated) {
                const uint32_t fin = vc[2];
                __threadfence();
                const uint32_t alloc = vc[0];
                if (fin == alloc) { finished_all = true; break; }
            }
            __nanosleep(backoff);
            if (backoff < (allocated ? 128u : 1024u)) backoff <<= 1;
        }
        if (__any_sync(0xffffffffu, finished_all)) break;
    }
}

// leaf-order records of the wide BVH: original primitive index (TLAS: instance index) and, for triangles, the
// 48-byte intersection record (e1/e2 as the reference's constructor, include/Geometry/Triangle.cuh:29-30)
__global__ void __launch_bounds__(256) emit_leaves8_kernel(GeometryD geo, const uint32_t *__restrict__ leaf_pos8, const uint32_t *__restrict__ vals,
                                                           const uint32_t *__restrict__ seg_offset, const uint32_t *__restrict__ seg_type,
                                                           const uint32_t *__restrict__ seg_first, int n_segs, int n, int leaf_mode,
                                                           uint32_t *__restrict__ leaf_prims8, TriIsect *__restrict__ leaf_tris8) {
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= n) return;
    const uint32_t pos = leaf_pos8[q];
    const uint32_t v = vals[pos];
    if (leaf_mode == 1) { leaf_prims8[q] = v; return; }
    const int seg = n_segs > 1 ? find_segment_w(seg_offset, n_segs, pos) : 0;
    const uint32_t prim = __ldg(&seg_first[seg]) + (v - __ldg(&seg_offset[seg]));
    leaf_prims8[q] = prim;
    if (__ldg(&seg_type[seg]) == 2u) {
        const float *s = geo.tri_verts + (size_t) prim * 9;
        const f3 v0 = mk3(__ldg(s), __ldg(s + 1), __ldg(s + 2));
        const f3 v1 = mk3(__ldg(s + 3), __ldg(s + 4), __ldg(s + 5));
        const f3 v2 = mk3(__ldg(s + 6), __ldg(s + 7), __ldg(s + 8));
        const f3 e1 = v1 - v0, e2 = v2 - v0;
        TriIsect t;
        t.v0_id = make_float4(v0.x, v0.y, v0.z, __uint_as_float(prim));
        t.e1 = make_float4(e1.x, e1.y, e1.z, 0.0f);
        t.e2 = make_float4(e2.x, e2.y, e2.z, 0.0f);
        leaf_tris8[q] = t;
    }
}

// ---- small TLAS: the whole per-frame rebuild in ONE block, ONE launch -------------------------------------------------------
// The TLAS of every config but C5 has at most a few hundred instances.  As a chain of general kernels (Morton keys, four
// radix passes, Karras, refit, collapse, leaf emission: 16 launches, each of them a few microseconds of work behind a launch
// gap) it cost 0.22 ms per frame of config 3.  Here one block of 512 threads does the same steps between __syncthreads():
// keys and the sort in shared memory (bitonic on (key, instance index): the order a stable radix sort gives), Karras from the
// shared keys, the atomic refit with its arrival flags in shared memory, the collapse level by level (the items of one level
// are ready when the previous level's barrier has passed: no polling), the leaf-order instance indices.  Every array the long
// path leaves behind (sorted keys, binary topology, boxes, counters) is written too, so introspection sees the same tree;
// the node contents are those of the serial twin (same per-element functions), which the parity tests compare.
struct TlasSmallArgs {
    const float4 *prim_lo, *prim_hi;
    const uint32_t *bounds;
    int n, bits;
    uint64_t *keys;
    uint32_t *vals;
    int *left, *right, *parent, *leaf_parent, *seg_root;
    int2 *range;
    float4 *node_lo, *node_hi;
    uint32_t *seg_offset, *seg_type;
    int *queue;
    uint32_t *counters;
    BvhNode8 *nodes8;
    uint32_t *leaf_pos8, *leaf_prims8;
    int octets;             // collapse with eight lanes per item (oct_collapse_item): a level's latency instead of its throughput
};
constexpr int TS_THREADS = 512;

__global__ void __launch_bounds__(TS_THREADS) tlas_small_kernel(TlasSmallArgs a) {
    __shared__ uint64_t skey[RT_TLAS_SMALL_MAX];
    __shared__ uint32_t sval[RT_TLAS_SMALL_MAX];
    __shared__ int sflag[RT_TLAS_SMALL_MAX];
    const int tid = threadIdx.x, n = a.n;
    int N = 2;
    while (N < n) N <<= 1;
    // ---- Morton keys over the instance centres' bounds (morton_kernel), padding sorts last ----
    Aabb scene;
    scene.lo = mk3(ordered_to_float(a.bounds[0]), ordered_to_float(a.bounds[1]), ordered_to_float(a.bounds[2]));
    scene.hi = mk3(ordered_to_float(a.bounds[3]), ordered_to_float(a.bounds[4]), ordered_to_float(a.bounds[5]));
    for (int j = tid; j < N; j += TS_THREADS) {
        if (j < n) {
            skey[j] = morton_code(view_box(a.prim_lo, a.prim_hi, j), scene, a.bits);
            sval[j] = (uint32_t) j;
        } else {
            skey[j] = ~0ull;
            sval[j] = 0xFFFFFFFFu;
        }
        sflag[j] = 0;
    }
    if (tid == 0) {
        a.seg_offset[0] = 0u; a.seg_offset[1] = (uint32_t) n; a.seg_type[0] = 0u;
        a.counters[0] = 1u; a.counters[1] = 0u; a.counters[2] = 0u; a.counters[3] = 0u;
        a.counters[4] = 0u; a.counters[5] = 1u; a.counters[6] = 0u;
    }
    __syncthreads();
    // ---- bitonic sort by (key, instance index) ----
    for (int k = 2; k <= N; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int p = tid; p < (N >> 1); p += TS_THREADS) {
                const int i = 2 * p - (p & (j - 1)), q = i + j;
                const uint64_t ki = skey[i], kq = skey[q];
                const uint32_t vi = sval[i], vq2 = sval[q];
                const bool greater = ki > kq || (ki == kq && vi > vq2);
                if (greater == ((i & k) == 0)) { skey[i] = kq; skey[q] = ki; sval[i] = vq2; sval[q] = vi; }
            }
            __syncthreads();
        }
    }
    for (int j = tid; j < n; j += TS_THREADS) { a.keys[j] = skey[j]; a.vals[j] = sval[j]; }
    // ---- Karras hierarchy (karras_kernel) ----
    for (int i = tid; i < n - 1; i += TS_THREADS) {
        const KarrasNode k = karras_node(skey, n, i);
        a.left[i] = k.left;
        a.right[i] = k.right;
        a.range[i] = make_int2(k.first, k.last);
        if (k.left >= 0) a.parent[k.left] = i; else a.leaf_parent[~k.left] = i;
        if (k.right >= 0) a.parent[k.right] = i; else a.leaf_parent[~k.right] = i;
        if (i == 0) a.parent[0] = -1;
    }
    volatile int *vq = a.queue;
    if (tid == 0) {
        a.seg_root[0] = n > 1 ? 0 : make_leaf_ref(0u, sval[0]);
        vq[0] = n > 1 ? 0 : ~0;
    }
    __syncthreads();
    // ---- bottom-up refit (refit_kernel): the second child to arrive computes the node ----
    if (n > 1) {
        for (int pos = tid; pos < n; pos += TS_THREADS) {
            int cur = a.leaf_parent[pos];
            while (cur >= 0) {
                __threadfence_block();
                if (atomicAdd(&sflag[cur], 1) == 0) break;
                __threadfence_block();
                const int l = a.left[cur], r = a.right[cur];
                float4 llo, lhi, rlo, rhi;
                if (l >= 0) { llo = __ldcg(&a.node_lo[l]); lhi = __ldcg(&a.node_hi[l]); } else { const uint32_t p = sval[~l]; llo = a.prim_lo[p]; lhi = a.prim_hi[p]; }
                if (r >= 0) { rlo = __ldcg(&a.node_lo[r]); rhi = __ldcg(&a.node_hi[r]); } else { const uint32_t p = sval[~r]; rlo = a.prim_lo[p]; rhi = a.prim_hi[p]; }
                __stcg(&a.node_lo[cur], make_float4(fminf(llo.x, rlo.x), fminf(llo.y, rlo.y), fminf(llo.z, rlo.z), __int_as_float(l)));
                __stcg(&a.node_hi[cur], make_float4(fmaxf(lhi.x, rhi.x), fmaxf(lhi.y, rhi.y), fmaxf(lhi.z, rhi.z), __int_as_float(r)));
                cur = a.parent[cur];
            }
        }
    }
    __syncthreads();
    // ---- collapse into 8-wide nodes, one level per round (collapse_item; items [begin, end) were published by the previous round) ----
    const BinaryTreeView view{a.range, a.node_lo, a.node_hi, a.prim_lo, a.prim_hi, a.vals};
    uint32_t begin = 0u, end = 1u;
    while (begin < end) {
        if (a.octets) {
            const int l = tid & 7;
            const uint32_t om = 0xFFu << (8 * ((tid & 31) >> 3));
            for (uint32_t i = begin + (uint32_t) (tid >> 3); i < end; i += TS_THREADS / 8)
                oct_collapse_item(view, vq[i], (int) i, 1, om, l, vq, a.counters, a.nodes8, a.leaf_pos8);
        } else {
            for (uint32_t i = begin + (uint32_t) tid; i < end; i += TS_THREADS) collapse_item(view, vq[i], (int) i, 1, vq, a.counters, a.nodes8, a.leaf_pos8);
        }
        __syncthreads();
        begin = end;
        end = ((volatile uint32_t *) a.counters)[0];
        __syncthreads();                       // everyone has read the count before the next round's allocations move it
    }
    if (tid == 0) { a.counters[2] = end; a.counters[3] = end; }
    // ---- leaf-order instance indices (emit_leaves8_kernel, leaf_mode 1) ----
    for (int q = tid; q < n; q += TS_THREADS) a.leaf_prims8[q] = sval[a.leaf_pos8[q]];
}

} // namespace

// The collapse inside the one-block TLAS kernel: eight lanes per item (default) or one thread per item (RTB200_TLAS_OCTETS=0).
// For the big builds the octets lost (throughput: three times the instructions); here a level is a handful of items and what
// counts is the latency of ONE item — children in registers instead of dynamically indexed local arrays: 0.152 -> 0.129 ms for
// the TLAS step of a config-3 frame (505 instances; profiles/r02_tlas_octets_ab.txt), same trees (twin equality tests).
static int tlas_small_octets() {
    static const int v = getenv("RTB200_TLAS_OCTETS") ? atoi(getenv("RTB200_TLAS_OCTETS")) : 1;
    return v;
}

bool LbvhBuilder::tlas_small(int n, int morton_bits, cudaStream_t stream) {
    // RTB200_TLAS_FUSED=0: the general kernels for every size (A/B runs)
    static const bool enabled = !(getenv("RTB200_TLAS_FUSED") && atoi(getenv("RTB200_TLAS_FUSED")) == 0);
    if (!enabled || n < 1 || n > RT_TLAS_SMALL_MAX || emit_bvh2 || sah_rotations || ploc) return false;
    keys_ = keys_a_.get(); keys_alt_ = keys_b_.get(); vals_ = vals_a_.get(); vals_alt_ = vals_b_.get();
    TlasSmallArgs a;
    a.prim_lo = prim_lo_.get(); a.prim_hi = prim_hi_.get(); a.bounds = seg_bounds_.get();
    a.n = n; a.bits = morton_bits;
    a.keys = keys_; a.vals = vals_;
    a.left = left_.get(); a.right = right_.get(); a.parent = parent_.get(); a.leaf_parent = leaf_parent_.get(); a.seg_root = seg_root_.get();
    a.range = range_.get();
    a.node_lo = node_lo_.get(); a.node_hi = node_hi_.get();
    a.seg_offset = seg_offset_.get(); a.seg_type = seg_type_.get();
    a.queue = queue_.get(); a.counters = wide_counters_.get();
    a.nodes8 = nodes8_.get(); a.leaf_pos8 = leaf_pos8_.get(); a.leaf_prims8 = leaf_prims8_.get();
    a.octets = tlas_small_octets();
    tlas_small_kernel<<<1, TS_THREADS, 0, stream>>>(a);
    CUDA_CHECK(cudaGetLastError());
    launches++;
    return true;
}

void LbvhBuilder::collapse_wide(int n, int n_segs, int leaf_mode, const GeometryD *geo, cudaStream_t stream, int sm_count) {
    if (collapse_blocks_per_sm_ == 0) {
        CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&collapse_blocks_per_sm_, collapse_kernel, CT, 0));
        if (collapse_blocks_per_sm_ < 1) collapse_blocks_per_sm_ = 1;
    }
    const int capacity = n + n_segs;
    CUDA_CHECK(cudaMemsetAsync(queue_.get(), 0x80, (size_t) capacity * sizeof(int), stream));
    collapse_init_kernel<<<div_up(n_segs, 256), 256, 0, stream>>>(seg_offset_.get(), seg_root_.get(), n_segs, queue_.get(), wide_counters_.get());
    const BinaryTreeView view{range_.get(), node_lo_.get(), node_hi_.get(), prim_lo_.get(), prim_hi_.get(), vals_};
    // items are claimed by ticket (see above): the grid need not be resident as a whole; 3 CTAs per SM measured best
    // tuning (measured): one 1 M-triangle BLAS 0.76 / 0.74 / 0.77 ms at 2 / 3 / 4 CTAs per SM, 0.97 at 8 (profiles/r02j_build_ab.txt); a batch
    // of 504 particle BLASes 0.62 / 0.58 / 0.62 / 0.67 / 0.86 at 1 / 2 / 3 / 4 / 6 (profiles/r02_c3_build_ab.txt, r02_bps_ab.txt): many roots
    // are ready at once there and the waiting threads of a larger grid only take L2 bandwidth from the working ones
    static const int collapse_bps_env = getenv("RTB200_COLLAPSE_BLOCKS_PER_SM") ? atoi(getenv("RTB200_COLLAPSE_BLOCKS_PER_SM")) : 0;
    const int collapse_bps = collapse_bps_env > 0 ? collapse_bps_env : (n_segs >= 64 ? 2 : 3);
    const int max_blocks = sm_count * std::min(collapse_blocks_per_sm_, collapse_bps);
    const int want_blocks = (int) div_up((size_t) capacity, CT);
    const int blocks = want_blocks < max_blocks ? want_blocks : max_blocks;
    static const int leaf_max_env = getenv("RTB200_LEAF_MAX") ? atoi(getenv("RTB200_LEAF_MAX")) : 0;   // tuning experiments only
    const int leaf_max = leaf_mode == 1 ? 1 : (leaf_max_env >= 1 && leaf_max_env <= 3 ? leaf_max_env : RT_WIDE_LEAF_MAX);
    // the first levels one launch each (collapse_level_kernel); small trees are quicker in the single ticket launch
    // Measured (profiles/r02i_build_ab.txt, r02i_build_launches.csv): NOT faster — 0.83 ms with 6 level launches, 0.97 with 12,
    // against 0.76 ms for the ticket kernel alone on the 1 M-triangle BLAS: one item alone takes about 20 us (the first three
    // level launches are 22-27 us each for 1, 8 and 64 items), the chain of levels is what costs, and each level launch adds its
    // own start-up on top.  Off by default (0); RTB200_COLLAPSE_LEVELS=n turns n level launches on for A/B runs.
    static const int level_launches_env = getenv("RTB200_COLLAPSE_LEVELS") ? atoi(getenv("RTB200_COLLAPSE_LEVELS")) : 0;
    if (capacity >= (1 << 15) && level_launches_env > 0) {
        size_t bound = (size_t) n_segs;                       // a level of depth L has at most n_segs * 8^L items, and never more than there are
        for (int level = 0; level < level_launches_env; level++) {
            const size_t items = std::min(bound, (size_t) capacity);
            collapse_level_kernel<<<div_up(items, 256), 256, 0, stream>>>(view, leaf_max, queue_.get(), wide_counters_.get(), wide_counters_.get() + 4,
                                                                         nodes8_.get(), leaf_pos8_.get());
            launches++;
            if (bound < (size_t) capacity) bound *= 8;
        }
    }
    // A/B: RTB200_COLLAPSE=octet runs the eight-lanes-per-item kernel.  Measured (profiles/r02f_build_ab.txt, r02g ncu): correct
    // (same trees as the twin) but slower, 0.88 vs 0.77 ms for the 1 M-triangle BLAS — the collapse is bound by the DRAM
    // latency of the dependent binary-node loads (long scoreboard 42-50 % of the stalls, L2 hit 56 %), not by the local-memory
    // arithmetic the octets remove, and eight fences per item cost more than they save.  One thread per item stays the default.
    static const bool octets = getenv("RTB200_COLLAPSE") && strcmp(getenv("RTB200_COLLAPSE"), "octet") == 0;
    if (octets) {
        if (collapse_octet_blocks_per_sm_ == 0) {
            CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&collapse_octet_blocks_per_sm_, collapse_octet_kernel, CT, 0));
            if (collapse_octet_blocks_per_sm_ < 1) collapse_octet_blocks_per_sm_ = 1;
        }
        static const int bps_env = getenv("RTB200_COLLAPSE_BLOCKS_PER_SM") ? atoi(getenv("RTB200_COLLAPSE_BLOCKS_PER_SM")) : 8;
        const int max_b = sm_count * std::min(collapse_octet_blocks_per_sm_, std::max(1, bps_env));
        const int want_b = (int) div_up((size_t) capacity * 8, CT);             // eight lanes per item
        collapse_octet_kernel<<<want_b < max_b ? want_b : max_b, CT, 0, stream>>>(view, leaf_max, capacity, queue_.get(), wide_counters_.get(), nodes8_.get(),
                                                                                   leaf_pos8_.get());
    } else
    collapse_kernel<<<blocks, CT, 0, stream>>>(view, leaf_max, capacity, queue_.get(), wide_counters_.get(), nodes8_.get(), leaf_pos8_.get());
    const GeometryD none{nullptr, nullptr, nullptr, nullptr};
    emit_leaves8_kernel<<<div_up(n, 256), 256, 0, stream>>>(geo ? *geo : none, leaf_pos8_.get(), vals_, seg_offset_.get(), seg_type_.get(),
                                                           seg_first_.get(), n_segs, n, leaf_mode, leaf_prims8_.get(), leaf_tris8_.get());
    launches += 3;
    CUDA_CHECK(cudaGetLastError());
}

#include "../../include/rtb200.h"

#include <cstdio>
#include <mutex>

// ---- rt_build_tlas_no_alloc: the one-block TLAS build for callers that cannot allocate ---------------------------------------
// No counterpart in the reference (include/rtb200.h).  The kernel is
// tlas_small_kernel above, unchanged; what differs is where its scratch lives: every intermediate array (leaf boxes as float4,
// Morton keys, sorted ids, Karras topology, node boxes, collapse queue and counters, leaf positions) is a field of ONE file-scope
// device array of RT_TLAS_NO_ALLOC_SCRATCH_BYTES, which every GPU that loads the library holds until the process exits.  The
// caller passes the input (instance boxes) and the outputs (wide nodes, leaf-order instance indices, node count), all device
// memory of the current device.  Nothing here calls cudaMalloc / cudaFree or an allocating routine, and nothing synchronises: two
// kernel launches and one 4-byte device-to-device copy on the caller's stream.
//
// In host code `g_tlas_no_alloc_scratch` is a host-side stand-in, not a device address: the address is taken once per device
// with cudaGetSymbolAddress (that device current) and that pointer is what the kernels receive.
// Every call on a device shares the array and finds the previous call's contents in it.  Nothing has to start at zero between
// calls: the prep kernel rewrites the boxes and the bounds, tlas_small_kernel resets its counters and its arrival flags (shared
// memory) and reads only queue / topology entries it wrote in the same launch.  Calls that OVERLAP on the GPU (two streams, two
// host threads, two processes' contexts do not share it — two streams of one process do) race on the array: serialise them.
namespace {

struct TlasNoAllocScratch {
    float4 prim_lo[RT_TLAS_SMALL_MAX], prim_hi[RT_TLAS_SMALL_MAX], node_lo[RT_TLAS_SMALL_MAX], node_hi[RT_TLAS_SMALL_MAX];
    uint64_t keys[RT_TLAS_SMALL_MAX];
    int2 range[RT_TLAS_SMALL_MAX];
    uint32_t vals[RT_TLAS_SMALL_MAX], leaf_pos8[RT_TLAS_SMALL_MAX];
    int left[RT_TLAS_SMALL_MAX], right[RT_TLAS_SMALL_MAX], parent[RT_TLAS_SMALL_MAX], leaf_parent[RT_TLAS_SMALL_MAX];
    int queue[RT_TLAS_SMALL_MAX + 8];
    uint32_t bounds[8], counters[8], seg_offset[4], seg_type[4];
    int seg_root[4];
};
static_assert(RT_TLAS_NO_ALLOC_MAX_INSTANCES == RT_TLAS_SMALL_MAX, "the no-alloc entry point serves what the one-block kernel serves");
static_assert(sizeof(TlasNoAllocScratch) <= RT_TLAS_NO_ALLOC_SCRATCH_BYTES, "scratch layout exceeds the stated size");

static __device__ __align__(256) unsigned char g_tlas_no_alloc_scratch[RT_TLAS_NO_ALLOC_SCRATCH_BYTES];

constexpr int NA_MAX_DEVICES = 64;
std::mutex g_na_mutex;
TlasNoAllocScratch *g_na_scratch[NA_MAX_DEVICES];     // device address of the array on each device, taken on first use

// caller's boxes (6 floats: lo xyz, hi xyz) -> the builder's float4 leaf boxes + the bounds of the box centres (ordered uints):
// what instance_bounds_kernel leaves for tlas_small_kernel in the frame path
__global__ void __launch_bounds__(TS_THREADS) tlas_no_alloc_prep_kernel(const float *__restrict__ boxes, int n, float4 *__restrict__ prim_lo,
                                                                        float4 *__restrict__ prim_hi, uint32_t *__restrict__ bounds) {
    __shared__ uint32_t sb[6];
    if (threadIdx.x < 6) sb[threadIdx.x] = threadIdx.x < 3 ? 0xFFFFFFFFu : 0u;
    __syncthreads();
    for (int j = threadIdx.x; j < n; j += TS_THREADS) {
        const float *b = boxes + 6 * (size_t) j;
        Aabb box;
        box.lo = mk3(b[0], b[1], b[2]);
        box.hi = mk3(b[3], b[4], b[5]);
        prim_lo[j] = make_float4(box.lo.x, box.lo.y, box.lo.z, 0.0f);
        prim_hi[j] = make_float4(box.hi.x, box.hi.y, box.hi.z, 0.0f);
        const Aabb c = aabb_centre_box(box);
        atomicMin(&sb[0], float_to_ordered(c.lo.x)); atomicMin(&sb[1], float_to_ordered(c.lo.y)); atomicMin(&sb[2], float_to_ordered(c.lo.z));
        atomicMax(&sb[3], float_to_ordered(c.hi.x)); atomicMax(&sb[4], float_to_ordered(c.hi.y)); atomicMax(&sb[5], float_to_ordered(c.hi.z));
    }
    __syncthreads();
    if (threadIdx.x < 6) bounds[threadIdx.x] = sb[threadIdx.x];
}

int na_fail(char *error_text, size_t error_cap, int code, const char *msg, const char *detail = nullptr) {
    if (error_text && error_cap) snprintf(error_text, error_cap, "rt_build_tlas_no_alloc: %s%s%s", msg, detail ? ": " : "", detail ? detail : "");
    return code;
}

} // namespace

extern "C" int rt_build_tlas_no_alloc(const float *boxes_dev, uint32_t n_instances, void *nodes_out_dev, uint32_t *leaf_instances_out_dev,
                                      uint32_t *n_nodes_out_dev, void *cuda_stream, char *error_text, size_t error_cap) {
    if (error_text && error_cap) error_text[0] = 0;
    if (n_instances < 1u) return na_fail(error_text, error_cap, RT_ERR_ARGUMENT, "n_instances must be at least 1");
    if (n_instances > (uint32_t) RT_TLAS_NO_ALLOC_MAX_INSTANCES) {
        char d[96];
        snprintf(d, sizeof(d), "%u instances, the static scratch array holds at most %d", n_instances, RT_TLAS_NO_ALLOC_MAX_INSTANCES);
        return na_fail(error_text, error_cap, RT_ERR_ARGUMENT, "too many instances", d);
    }
    if (!boxes_dev || !nodes_out_dev || !leaf_instances_out_dev) return na_fail(error_text, error_cap, RT_ERR_ARGUMENT, "null input or output pointer");
    if (((uintptr_t) boxes_dev & 3u) || ((uintptr_t) leaf_instances_out_dev & 3u) || ((uintptr_t) n_nodes_out_dev & 3u))
        return na_fail(error_text, error_cap, RT_ERR_ARGUMENT, "boxes / leaf_instances_out / n_nodes_out must be 4-byte aligned");
    if ((uintptr_t) nodes_out_dev & 15u) return na_fail(error_text, error_cap, RT_ERR_ARGUMENT, "nodes_out must be 16-byte aligned (80-byte records written as 128-bit words)");
    int dev = -1;
    cudaError_t rc = cudaGetDevice(&dev);
    if (rc != cudaSuccess) return na_fail(error_text, error_cap, RT_ERR_CUDA, "cudaGetDevice", cudaGetErrorString(rc));
    if (dev < 0 || dev >= NA_MAX_DEVICES) return na_fail(error_text, error_cap, RT_ERR_ARGUMENT, "device ordinal beyond the table of scratch addresses (64)");
    TlasNoAllocScratch *sc = nullptr;
    {
        // set-up, once per device, with that device current: the array's name is a host-side stand-in here, its device address
        // comes from cudaGetSymbolAddress and is what the kernels get
        std::lock_guard<std::mutex> lock(g_na_mutex);
        if (!g_na_scratch[dev]) {
            void *p = nullptr;
            rc = cudaGetSymbolAddress(&p, g_tlas_no_alloc_scratch);
            if (rc != cudaSuccess || !p) return na_fail(error_text, error_cap, RT_ERR_CUDA, "cudaGetSymbolAddress(scratch)", cudaGetErrorString(rc));
            g_na_scratch[dev] = static_cast<TlasNoAllocScratch *>(p);
        }
        sc = g_na_scratch[dev];
    }
    cudaStream_t stream = static_cast<cudaStream_t>(cuda_stream);
    const int n = (int) n_instances;
    tlas_no_alloc_prep_kernel<<<1, TS_THREADS, 0, stream>>>(boxes_dev, n, sc->prim_lo, sc->prim_hi, sc->bounds);
    TlasSmallArgs a;
    a.prim_lo = sc->prim_lo; a.prim_hi = sc->prim_hi; a.bounds = sc->bounds;
    a.n = n; a.bits = 10;                                  // tlas_morton_bits(n) for every n <= 1024 (host_scene.cpp)
    a.keys = sc->keys; a.vals = sc->vals;
    a.left = sc->left; a.right = sc->right; a.parent = sc->parent; a.leaf_parent = sc->leaf_parent; a.seg_root = sc->seg_root;
    a.range = sc->range;
    a.node_lo = sc->node_lo; a.node_hi = sc->node_hi;
    a.seg_offset = sc->seg_offset; a.seg_type = sc->seg_type;
    a.queue = sc->queue; a.counters = sc->counters;
    a.nodes8 = static_cast<BvhNode8 *>(nodes_out_dev); a.leaf_pos8 = sc->leaf_pos8; a.leaf_prims8 = leaf_instances_out_dev;
    a.octets = tlas_small_octets();
    tlas_small_kernel<<<1, TS_THREADS, 0, stream>>>(a);
    rc = cudaGetLastError();
    if (rc != cudaSuccess) return na_fail(error_text, error_cap, RT_ERR_CUDA, "kernel launch", cudaGetErrorString(rc));
    if (n_nodes_out_dev) {
        rc = cudaMemcpyAsync(n_nodes_out_dev, &sc->counters[0], sizeof(uint32_t), cudaMemcpyDeviceToDevice, stream);
        if (rc != cudaSuccess) return na_fail(error_text, error_cap, RT_ERR_CUDA, "cudaMemcpyAsync(node count)", cudaGetErrorString(rc));
    }
    return RT_OK;
}
