// kernels_ploc.cu — PLOC build of the binary tree over the Morton-sorted primitives (rt_build_core.h: ploc_nearest /
// ploc_role), the alternative to Karras + refit for BLASes that are built once.  Replaces, like the LBVH, the
// reference's host builder BLAS::constructBLAS (src/AS/BLAS.cu:4-117: random axis, median split).
//
// One round = four flat passes over the current cluster array (m clusters):
//   nearest : nn[i] = best merge partner within RT_PLOC_RADIUS positions, same segment
//   count   : role of every cluster (stay / lead a merge / absorbed) -> per-block counts of survivors and of leaders
//   scan    : exclusive scan of the block counts (one block), new cluster count, node-id base of the round
//   scatter : survivors move to their new position (order kept); a leader writes the new inner node (box, children,
//             parents) and takes its id
// Node ids are handed out bottom-up (the root of the last segment to finish gets the highest), so, unlike the Karras
// tree, node 0 is not the root: consumers take the roots from seg_root.
#include "build.h"

namespace {

constexpr int PB = 256;

__global__ void __launch_bounds__(PB) ploc_init_kernel(const float4 *__restrict__ prim_lo, const float4 *__restrict__ prim_hi, const uint32_t *__restrict__ vals,
                                                       const uint32_t *__restrict__ seg_offset, int n_segs, int n, float4 *__restrict__ clo,
                                                       float4 *__restrict__ chi) {
    const int pos = blockIdx.x * blockDim.x + threadIdx.x;
    if (pos >= n) return;
    int lo = 0, hi = n_segs - 1;
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (__ldg(&seg_offset[mid]) <= (uint32_t) pos) lo = mid; else hi = mid - 1;
    }
    const uint32_t p = vals[pos];
    const float4 a = prim_lo[p], b = prim_hi[p];
    clo[pos] = make_float4(a.x, a.y, a.z, __int_as_float(~pos));
    chi[pos] = make_float4(b.x, b.y, b.z, __int_as_float(lo));
}

__global__ void __launch_bounds__(PB) ploc_nearest_kernel(const float4 *__restrict__ clo, const float4 *__restrict__ chi, int m, int *__restrict__ nn) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) return;
    nn[i] = ploc_nearest(clo, chi, m, i, RT_PLOC_RADIUS);
}

// per-block counts: x = survivors (roles 0 and 1), y = leaders (role 1)
__global__ void __launch_bounds__(PB) ploc_count_kernel(const int *__restrict__ nn, int m, uint2 *__restrict__ block_counts) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int role = i < m ? ploc_role(nn, i) : 2;
    const uint32_t keep = __ballot_sync(0xffffffffu, role != 2), lead = __ballot_sync(0xffffffffu, role == 1);
    __shared__ uint2 warp_counts[PB / 32];
    if ((threadIdx.x & 31) == 0) warp_counts[threadIdx.x >> 5] = make_uint2(__popc(keep), __popc(lead));
    __syncthreads();
    if (threadIdx.x == 0) {
        uint2 t = make_uint2(0u, 0u);
        for (int w = 0; w < PB / 32; w++) { t.x += warp_counts[w].x; t.y += warp_counts[w].y; }
        block_counts[blockIdx.x] = t;
    }
}

// one block: exclusive scan of the block counts in place; state = {clusters, next node id}: clusters <- survivors,
// next node id advances by the leaders of the round (state[2] keeps the id base the scatter pass uses)
__global__ void __launch_bounds__(1024) ploc_scan_kernel(uint2 *__restrict__ block_counts, int n_blocks, int *__restrict__ state) {
    __shared__ uint2 carry;
    __shared__ uint2 warp_tot[32];
    if (threadIdx.x == 0) carry = make_uint2(0u, 0u);
    __syncthreads();
    for (int base = 0; base < n_blocks; base += 1024) {
        const int b = base + (int) threadIdx.x;
        const uint2 v = b < n_blocks ? block_counts[b] : make_uint2(0u, 0u);
        uint2 inc = v;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            const uint32_t x = __shfl_up_sync(0xffffffffu, inc.x, off), y = __shfl_up_sync(0xffffffffu, inc.y, off);
            if ((threadIdx.x & 31) >= (unsigned) off) { inc.x += x; inc.y += y; }
        }
        if ((threadIdx.x & 31) == 31) warp_tot[threadIdx.x >> 5] = inc;
        __syncthreads();
        uint2 wbase = make_uint2(0u, 0u);
        for (int w = 0; w < (int) (threadIdx.x >> 5); w++) { wbase.x += warp_tot[w].x; wbase.y += warp_tot[w].y; }
        const uint2 c = carry;
        if (b < n_blocks) block_counts[b] = make_uint2(c.x + wbase.x + inc.x - v.x, c.y + wbase.y + inc.y - v.y);
        __syncthreads();
        if (threadIdx.x == 1023) carry = make_uint2(c.x + wbase.x + inc.x, c.y + wbase.y + inc.y);
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        state[2] = state[1];                       // node-id base of this round
        state[0] = (int) carry.x;                  // clusters after the round
        state[1] = state[1] + (int) carry.y;       // next free node id
    }
}

__global__ void __launch_bounds__(PB) ploc_scatter_kernel(const float4 *__restrict__ clo, const float4 *__restrict__ chi, const int *__restrict__ nn, int m,
                                                          const uint2 *__restrict__ block_offsets, const int *__restrict__ state,
                                                          float4 *__restrict__ out_lo, float4 *__restrict__ out_hi,
                                                          float4 *__restrict__ node_lo, float4 *__restrict__ node_hi,
                                                          int *__restrict__ left, int *__restrict__ right, int *__restrict__ parent, int *__restrict__ leaf_parent) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const int role = i < m ? ploc_role(nn, i) : 2;
    const uint32_t keep = __ballot_sync(0xffffffffu, role != 2), lead = __ballot_sync(0xffffffffu, role == 1);
    __shared__ uint2 warp_counts[PB / 32];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) warp_counts[warp] = make_uint2(__popc(keep), __popc(lead));
    __syncthreads();
    if (role == 2) return;
    uint2 off = block_offsets[blockIdx.x];
    for (int w = 0; w < warp; w++) { off.x += warp_counts[w].x; off.y += warp_counts[w].y; }
    const uint32_t below = (1u << lane) - 1u;
    const int dst = (int) (off.x + __popc(keep & below));
    float4 a = clo[i], b = chi[i];
    if (role == 1) {
        const int j = nn[i];
        const float4 c = clo[j], d = chi[j];
        const int id = state[2] + (int) (off.y + __popc(lead & below));
        const int l = __float_as_int(a.w), r = __float_as_int(c.w);
        a = make_float4(fminf(a.x, c.x), fminf(a.y, c.y), fminf(a.z, c.z), __int_as_float(id));
        b = make_float4(fmaxf(b.x, d.x), fmaxf(b.y, d.y), fmaxf(b.z, d.z), b.w);
        // .w of the node record carries its children (what the wide collapse reads, see refit_store in kernels_build.cu)
        node_lo[id] = make_float4(a.x, a.y, a.z, __int_as_float(l));
        node_hi[id] = make_float4(b.x, b.y, b.z, __int_as_float(r));
        left[id] = l;
        right[id] = r;
        if (l >= 0) parent[l] = id; else leaf_parent[~l] = id;
        if (r >= 0) parent[r] = id; else leaf_parent[~r] = id;
    }
    out_lo[dst] = a;
    out_hi[dst] = b;
}

// the clusters left at the end are the roots of their segments
__global__ void __launch_bounds__(PB) ploc_roots_kernel(const float4 *__restrict__ clo, const float4 *__restrict__ chi, int m, int *__restrict__ seg_root) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) return;
    const int ref = __float_as_int(clo[i].w);
    if (ref >= 0) seg_root[__float_as_int(chi[i].w)] = ref;
}

} // namespace

void LbvhBuilder::run_ploc(int n, int n_segs, cudaStream_t stream) {
    const size_t max_blocks = div_up((size_t) n, PB);
    ploc_lo_a_.reserve(n); ploc_hi_a_.reserve(n); ploc_lo_b_.reserve(n); ploc_hi_b_.reserve(n);
    ploc_nn_.reserve(n);
    ploc_blocks_.reserve(max_blocks);
    ploc_state_.reserve(4);
    ploc_state_host_.reserve(4);
    float4 *lo = ploc_lo_a_.get(), *hi = ploc_hi_a_.get(), *lo2 = ploc_lo_b_.get(), *hi2 = ploc_hi_b_.get();
    ploc_init_kernel<<<div_up(n, PB), PB, 0, stream>>>(prim_lo_.get(), prim_hi_.get(), vals_, seg_offset_.get(), n_segs, n, lo, hi);
    CUDA_CHECK(cudaMemsetAsync(parent_.get(), 0xFF, (size_t) (n - 1) * sizeof(int), stream));      // roots keep parent -1
    // a batch of S segments makes n - S inner nodes; the S - 1 unused records of the n - 1 read as (leaf 0, leaf 0)
    CUDA_CHECK(cudaMemsetAsync(left_.get(), 0xFF, (size_t) (n - 1) * sizeof(int), stream));
    CUDA_CHECK(cudaMemsetAsync(right_.get(), 0xFF, (size_t) (n - 1) * sizeof(int), stream));
    CUDA_CHECK(cudaMemsetAsync(node_lo_.get(), 0, (size_t) (n - 1) * sizeof(float4), stream));
    CUDA_CHECK(cudaMemsetAsync(node_hi_.get(), 0, (size_t) (n - 1) * sizeof(float4), stream));
    int *h = ploc_state_host_.get();
    h[0] = n; h[1] = 0; h[2] = 0; h[3] = 0;
    CUDA_CHECK(cudaMemcpyAsync(ploc_state_.get(), h, 4 * sizeof(int), cudaMemcpyHostToDevice, stream));
    launches += 1;
    int m = n;
    ploc_rounds = 0;
    // every round merges at least one pair per segment that still has two clusters: it ends
    while (m > n_segs) {
        const int blocks = (int) div_up((size_t) m, PB);
        ploc_nearest_kernel<<<blocks, PB, 0, stream>>>(lo, hi, m, ploc_nn_.get());
        ploc_count_kernel<<<blocks, PB, 0, stream>>>(ploc_nn_.get(), m, ploc_blocks_.get());
        ploc_scan_kernel<<<1, 1024, 0, stream>>>(ploc_blocks_.get(), blocks, ploc_state_.get());
        ploc_scatter_kernel<<<blocks, PB, 0, stream>>>(lo, hi, ploc_nn_.get(), m, ploc_blocks_.get(), ploc_state_.get(), lo2, hi2, node_lo_.get(), node_hi_.get(),
                                                      left_.get(), right_.get(), parent_.get(), leaf_parent_.get());
        CUDA_CHECK(cudaMemcpyAsync(h, ploc_state_.get(), 4 * sizeof(int), cudaMemcpyDeviceToHost, stream));
        CUDA_CHECK(cudaStreamSynchronize(stream));
        launches += 4;
        ploc_rounds++;
        if (h[0] >= m) throw CudaError("PLOC build made no progress (non-finite primitive bounds?)");
        m = h[0];
        std::swap(lo, lo2);
        std::swap(hi, hi2);
    }
    ploc_roots_kernel<<<div_up(m, PB), PB, 0, stream>>>(lo, hi, m, seg_root_.get());
    launches += 1;
    CUDA_CHECK(cudaGetLastError());
}
