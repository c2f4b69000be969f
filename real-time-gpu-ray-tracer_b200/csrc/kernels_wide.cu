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
    if (s == 0) { counters[0] = (uint32_t) n_segs; counters[1] = 0u; counters[2] = 0u; counters[3] = 0u; }
    if (s >= n_segs) return;
    const uint32_t first = seg_offset[s], cnt = seg_offset[s + 1] - first;
    queue[s] = cnt == 1u ? ~(int) first : seg_root[s];
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
                    __threadfence();
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

} // namespace

void LbvhBuilder::collapse_wide(int n, int n_segs, int leaf_mode, const GeometryD *geo, cudaStream_t stream, int sm_count) {
    if (collapse_blocks_per_sm_ == 0) {
        CUDA_CHECK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&collapse_blocks_per_sm_, collapse_kernel, CT, 0));
        if (collapse_blocks_per_sm_ < 1) collapse_blocks_per_sm_ = 1;
    }
    const int capacity = n + n_segs;
    CUDA_CHECK(cudaMemsetAsync(queue_.get(), 0x80, (size_t) capacity * sizeof(int), stream));
    collapse_init_kernel<<<div_up(n_segs, 256), 256, 0, stream>>>(seg_offset_.get(), seg_root_.get(), n_segs, queue_.get(), wide_counters_.get());
    const BinaryTreeView view{range_.get(), node_lo_.get(), node_hi_.get(), prim_lo_.get(), prim_hi_.get(), vals_};
    // items are claimed by ticket (see above): the grid need not be resident as a whole; 2 CTAs per SM measured best
    static const int collapse_bps_env = getenv("RTB200_COLLAPSE_BLOCKS_PER_SM") ? atoi(getenv("RTB200_COLLAPSE_BLOCKS_PER_SM")) : 2;   // tuning (measured: 2 beats 3, 4, 8 — fewer pollers)
    const int max_blocks = sm_count * std::min(collapse_blocks_per_sm_, std::max(1, collapse_bps_env));
    const int want_blocks = (int) div_up((size_t) capacity, CT);
    const int blocks = want_blocks < max_blocks ? want_blocks : max_blocks;
    static const int leaf_max_env = getenv("RTB200_LEAF_MAX") ? atoi(getenv("RTB200_LEAF_MAX")) : 0;   // tuning experiments only
    const int leaf_max = leaf_mode == 1 ? 1 : (leaf_max_env >= 1 && leaf_max_env <= 3 ? leaf_max_env : RT_WIDE_LEAF_MAX);
    collapse_kernel<<<blocks, CT, 0, stream>>>(view, leaf_max, capacity, queue_.get(), wide_counters_.get(), nodes8_.get(), leaf_pos8_.get());
    const GeometryD none{nullptr, nullptr, nullptr, nullptr};
    emit_leaves8_kernel<<<div_up(n, 256), 256, 0, stream>>>(geo ? *geo : none, leaf_pos8_.get(), vals_, seg_offset_.get(), seg_type_.get(),
                                                           seg_first_.get(), n_segs, n, leaf_mode, leaf_prims8_.get(), leaf_tris8_.get());
    launches += 3;
    CUDA_CHECK(cudaGetLastError());
}
