// build.h — host-side drivers of the device LBVH pipeline (replaces the reference's single-threaded host
// builders BLAS::constructBLAS / TLAS::constructTLAS, src/AS/BLAS.cu:4-117, src/AS/TLAS.cu:4-129).
#pragma once

#include "device_util.h"
#include "rt_build_core.h"
#include "rt_prims.h"
#include "rt_wide_core.h"

#include <algorithm>
#include <cstdint>
#include <vector>

// a TLAS of at most this many instances is rebuilt by one block in one launch (kernels_wide.cu tlas_small_kernel)
#define RT_TLAS_SMALL_MAX 1024

class RadixSorter {
public:
    // Sorts n (key, value) pairs by the low `key_bits` bits of the key.  On return `keys`/`vals` point at the
    // sorted arrays and `keys_alt`/`vals_alt` at the scratch copies (the pairs ping-pong between them).
    void sort(uint64_t *&keys, uint64_t *&keys_alt, uint32_t *&vals, uint32_t *&vals_alt, int n, int key_bits,
              cudaStream_t stream, int sm_count, uint32_t &launches);
private:
    DeviceBuffer<uint32_t> scratch_;
};

// One batched LBVH build: all segments (BLASes) of the batch share one sort and one Karras pass; segment s
// owns leaf positions [seg_offset[s], seg_offset[s+1]) and the subtree rooted at seg_root[s].
class LbvhBuilder {
public:
    // ---- BLAS batch.  seg_* are host arrays of n_segs (+1 for offsets) entries. -------------------------
    void build_blas(const GeometryD &geo, const std::vector<uint32_t> &seg_offset, const std::vector<uint32_t> &seg_type,
                    const std::vector<uint32_t> &seg_first, int morton_bits, cudaStream_t stream, int sm_count);
    // ---- TLAS: one segment over instance world boxes already written to prim_lo()/prim_hi() and bounds() --
    void prepare_blas(const std::vector<uint32_t> &seg_offset, const std::vector<uint32_t> &seg_type);   // allocations only
    void prepare_tlas(int n_instances);
    bool sah_rotations = false;              // optional SAH refinement (RT_FLAG_SAH_REFINE): one tree rotation per node inside the refit (rt_build_core.h best_rotation)
    bool ploc = false;                       // build the binary tree with PLOC (kernels_ploc.cu) instead of Karras + refit: BLASes built once
    int ploc_rounds = 0;                     // rounds the last PLOC build took
    bool emit_bvh2 = false;                  // also emit the 64-byte binary nodes + their leaf-order records (RT_FLAG_TRAVERSE_BVH2)
    void build_tlas(int n_instances, int morton_bits, cudaStream_t stream, int sm_count);

    // results (device): the compressed 8-wide BVH the render kernels traverse (root of segment s = node s) ...
    const BvhNode8 *nodes8() const { return nodes8_.get(); }
    const TriIsect *leaf_tris8() const { return leaf_tris8_.get(); }
    const uint32_t *leaf_prims8() const { return leaf_prims8_.get(); }   // BLAS: original primitive index; TLAS: instance index
    const uint32_t *wide_counters() const { return wide_counters_.get(); } // [0] wide nodes, [1] leaf primitives, [2] finished
    size_t node8_capacity() const { return (size_t) n_ + (size_t) n_segs_; }
    // ... and the binary LBVH it was collapsed from (kept for introspection and the BVH2 comparison kernel)
    const BvhNode2 *nodes() const { return nodes_.get(); }
    const TriIsect *leaf_tris() const { return leaf_tris_.get(); }
    const uint32_t *leaf_prims() const { return leaf_prims_.get(); }
    const int *seg_root_dev() const { return seg_root_.get(); }
    std::vector<int> seg_root_host;          // filled by build_blas (synchronises the stream once)
    int root_ref_host = 0;                   // TLAS root (filled by build_tlas via *_dev below, no sync)
    // scratch the instance-update kernel writes into (TLAS mode)
    float4 *prim_lo() const { return prim_lo_.get(); }
    float4 *prim_hi() const { return prim_hi_.get(); }
    uint32_t *bounds() const { return seg_bounds_.get(); }   // 6 ordered-uint words per segment
    // introspection for parity tests (valid after a build)
    int n_prims() const { return n_; }
    const uint64_t *sorted_keys() const { return keys_; }
    const uint32_t *sorted_vals() const { return vals_; }
    const int *left() const { return left_.get(); }
    const int *right() const { return right_.get(); }
    const int *parent() const { return parent_.get(); }
    const float4 *node_lo() const { return node_lo_.get(); }
    const float4 *node_hi() const { return node_hi_.get(); }
    uint32_t launches = 0;
    cudaEvent_t wait_event = nullptr;        // blocking-sync event for the host wait in build_blas (nullptr: cudaStreamSynchronize)
    size_t node_bytes() const { return (size_t) (n_ > 1 ? n_ - 1 : 0) * sizeof(BvhNode2); }

private:
    void reserve(int n, int n_segs, bool tris);
    void run_tree(int n, int n_segs, int key_bits, int leaf_mode, const GeometryD *geo, cudaStream_t stream, int sm_count);
    bool tlas_small(int n, int morton_bits, cudaStream_t stream);                                                      // kernels_wide.cu: one block, one launch
    void run_ploc(int n, int n_segs, cudaStream_t stream);                                                             // kernels_ploc.cu
    void collapse_wide(int n, int n_segs, int leaf_mode, const GeometryD *geo, cudaStream_t stream, int sm_count);   // kernels_wide.cu
    int n_ = 0, n_segs_ = 0;
    RadixSorter sorter_;
    DeviceBuffer<float4> prim_lo_, prim_hi_, node_lo_, node_hi_;
    DeviceBuffer<uint32_t> seg_bounds_, seg_offset_, seg_type_, seg_first_;
    DeviceBuffer<uint64_t> keys_a_, keys_b_;
    DeviceBuffer<uint32_t> vals_a_, vals_b_;
    DeviceBuffer<int> left_, right_, parent_, leaf_parent_, flags_, seg_root_;
    DeviceBuffer<BvhNode2> nodes_;
    DeviceBuffer<int2> range_;
    DeviceBuffer<int> queue_;
    DeviceBuffer<uint32_t> wide_counters_, leaf_pos8_, leaf_prims8_;
    DeviceBuffer<BvhNode8> nodes8_;
    DeviceBuffer<TriIsect> leaf_tris8_;
    int collapse_blocks_per_sm_ = 0, collapse_octet_blocks_per_sm_ = 0;
    DeviceBuffer<float4> ploc_lo_a_, ploc_hi_a_, ploc_lo_b_, ploc_hi_b_;
    DeviceBuffer<int> ploc_nn_, ploc_state_;
    DeviceBuffer<uint2> ploc_blocks_;
    PinnedBuffer<int> ploc_state_host_;
    DeviceBuffer<TriIsect> leaf_tris_;
    DeviceBuffer<uint32_t> leaf_prims_;
    uint64_t *keys_ = nullptr, *keys_alt_ = nullptr;
    uint32_t *vals_ = nullptr, *vals_alt_ = nullptr;
};

// per-frame instance update (device restatement of Instance::updateTransformArguments); writes InstanceD
// records plus the TLAS builder's prim boxes and scene bounds.
// triangle strips of a particle frame -> triangle vertex + shading records (config 3 input stage; kernels_build.cu)
void launch_expand_strips(const float *points, const float *normals, const int *ids, const uint32_t *strip_first_tri, const uint32_t *strip_first_id,
                          int n_strips, uint32_t n_tris, uint32_t material_type, uint32_t material_index, float *tri_verts, TriShade *tri_shade,
                          cudaStream_t stream);
void launch_instance_update(const InstanceXformIn *in, int n, InstanceD *out, float4 *wlo, float4 *whi, float4 *wcentroid,
                            uint32_t *bounds, const float *tri_verts, cudaStream_t stream);
