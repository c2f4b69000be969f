// kernels_build.cu — device LBVH pipeline: primitive bounds -> Morton keys -> onesweep radix sort ->
// Karras hierarchy -> atomic bottom-up refit -> node emission (+ leaf-order triangle records).
// One thread per element everywhere; grids are sized to the element count (>= 148 SMs * resident CTAs for
// the configs that matter).  All of it is HBM-bound streaming work: algorithmic bytes per primitive are
// listed in DESIGN.md ("LBVH build roofline").
#include "build.h"

namespace {

constexpr int BT = 256;   // threads per block for the element-wise kernels

__device__ __forceinline__ int find_segment(const uint32_t *__restrict__ seg_offset, int n_segs, uint32_t j) {
    // largest s with seg_offset[s] <= j
    int lo = 0, hi = n_segs - 1;
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (__ldg(&seg_offset[mid]) <= j) lo = mid; else hi = mid - 1;
    }
    return lo;
}

__global__ void __launch_bounds__(BT) init_bounds_kernel(uint32_t *bounds, int n_segs) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_segs * 6) bounds[i] = (i % 6) < 3 ? 0xFFFFFFFFu : 0u;   // lo = +max, hi = -max in ordered space
}

// TLAS: one segment [0, n) — written by a kernel so the per-frame rebuild needs no host->device copy
__global__ void set_single_segment_kernel(uint32_t *seg_offset, uint32_t *seg_type, uint32_t n) {
    seg_offset[0] = 0u;
    seg_offset[1] = n;
    seg_type[0] = 0u;
}

__device__ __forceinline__ void atomic_bounds(uint32_t *b, const Aabb &box) {
    atomicMin(&b[0], float_to_ordered(box.lo.x)); atomicMin(&b[1], float_to_ordered(box.lo.y)); atomicMin(&b[2], float_to_ordered(box.lo.z));
    atomicMax(&b[3], float_to_ordered(box.hi.x)); atomicMax(&b[4], float_to_ordered(box.hi.y)); atomicMax(&b[5], float_to_ordered(box.hi.z));
}

// K1: per build-primitive box (padded) + per-segment bounds.  Warps whose lanes all belong to one segment
// reduce with shuffles and issue one atomic set per warp.
__global__ void __launch_bounds__(BT) prim_bounds_kernel(GeometryD geo, const uint32_t *__restrict__ seg_offset, const uint32_t *__restrict__ seg_type,
                                                         const uint32_t *__restrict__ seg_first, int n_segs, int n,
                                                         float4 *__restrict__ prim_lo, float4 *__restrict__ prim_hi, uint32_t *bounds) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    const bool valid = j < n;
    int seg = -1;
    Aabb box;
    box.lo = mk3(RT_INF, RT_INF, RT_INF);
    box.hi = mk3(-RT_INF, -RT_INF, -RT_INF);
    if (valid) {
        seg = find_segment(seg_offset, n_segs, (uint32_t) j);
        const uint32_t type = __ldg(&seg_type[seg]);
        const uint32_t prim = __ldg(&seg_first[seg]) + ((uint32_t) j - __ldg(&seg_offset[seg]));
        box = prim_bounds(geo, type, prim);                      // the primitive table (rt_prims.h)
        box = aabb_pad(box);
        prim_lo[j] = make_float4(box.lo.x, box.lo.y, box.lo.z, 0.0f);
        prim_hi[j] = make_float4(box.hi.x, box.hi.y, box.hi.z, 0.0f);
    }
    // reduce per warp, then per block when the whole block lies in one segment (the common case: one atomic set per
    // 256 primitives — per-warp atomics on the same six words serialised in L2 and dominated this kernel)
    __shared__ float sh_box[BT / 32][6];
    __shared__ int sh_seg[BT / 32];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    int seg0 = __shfl_sync(0xffffffffu, seg, 0);
    const bool uniform = __all_sync(0xffffffffu, seg == seg0 || !valid) && seg0 >= 0;
    if (uniform) {
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            box.lo.x = fminf(box.lo.x, __shfl_xor_sync(0xffffffffu, box.lo.x, off));
            box.lo.y = fminf(box.lo.y, __shfl_xor_sync(0xffffffffu, box.lo.y, off));
            box.lo.z = fminf(box.lo.z, __shfl_xor_sync(0xffffffffu, box.lo.z, off));
            box.hi.x = fmaxf(box.hi.x, __shfl_xor_sync(0xffffffffu, box.hi.x, off));
            box.hi.y = fmaxf(box.hi.y, __shfl_xor_sync(0xffffffffu, box.hi.y, off));
            box.hi.z = fmaxf(box.hi.z, __shfl_xor_sync(0xffffffffu, box.hi.z, off));
        }
    } else if (valid) {
        atomic_bounds(bounds + 6 * seg, box);
    }
    if (lane == 0) {
        sh_seg[warp] = uniform ? seg0 : -2;          // -2: this warp already did its own atomics; -1: warp past the end
        sh_box[warp][0] = box.lo.x; sh_box[warp][1] = box.lo.y; sh_box[warp][2] = box.lo.z;
        sh_box[warp][3] = box.hi.x; sh_box[warp][4] = box.hi.y; sh_box[warp][5] = box.hi.z;
    }
    __syncthreads();
    if (warp == 0) {
        const int s = lane < BT / 32 ? sh_seg[lane] : -1;
        const int sfirst = __shfl_sync(0xffffffffu, s, 0);
        const bool block_uniform = __all_sync(0xffffffffu, s == sfirst || s == -1) && sfirst >= 0;
        if (block_uniform) {
            if (lane == 0) {
                Aabb b;
                b.lo = mk3(sh_box[0][0], sh_box[0][1], sh_box[0][2]);
                b.hi = mk3(sh_box[0][3], sh_box[0][4], sh_box[0][5]);
                for (int w = 1; w < BT / 32; w++) {
                    if (sh_seg[w] < 0) continue;
                    b.lo = mk3(fminf(b.lo.x, sh_box[w][0]), fminf(b.lo.y, sh_box[w][1]), fminf(b.lo.z, sh_box[w][2]));
                    b.hi = mk3(fmaxf(b.hi.x, sh_box[w][3]), fmaxf(b.hi.y, sh_box[w][4]), fmaxf(b.hi.z, sh_box[w][5]));
                }
                atomic_bounds(bounds + 6 * sfirst, b);
            }
        } else if (lane < BT / 32 && s >= 0) {
            Aabb b;
            b.lo = mk3(sh_box[lane][0], sh_box[lane][1], sh_box[lane][2]);
            b.hi = mk3(sh_box[lane][3], sh_box[lane][4], sh_box[lane][5]);
            atomic_bounds(bounds + 6 * s, b);
        }
    }
}

// K2: key = (segment << 3*bits) | morton(box centre within segment bounds); payload = build-primitive index
__global__ void __launch_bounds__(BT) morton_kernel(const float4 *__restrict__ prim_lo, const float4 *__restrict__ prim_hi,
                                                    const uint32_t *__restrict__ seg_offset, int n_segs, const uint32_t *__restrict__ bounds,
                                                    int n, int bits, uint64_t *__restrict__ keys, uint32_t *__restrict__ vals) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const int seg = n_segs > 1 ? find_segment(seg_offset, n_segs, (uint32_t) j) : 0;
    const uint32_t *b = bounds + 6 * seg;
    Aabb scene;
    scene.lo = mk3(ordered_to_float(__ldg(b + 0)), ordered_to_float(__ldg(b + 1)), ordered_to_float(__ldg(b + 2)));
    scene.hi = mk3(ordered_to_float(__ldg(b + 3)), ordered_to_float(__ldg(b + 4)), ordered_to_float(__ldg(b + 5)));
    const float4 lo = prim_lo[j], hi = prim_hi[j];
    Aabb box;
    box.lo = mk3(lo.x, lo.y, lo.z);
    box.hi = mk3(hi.x, hi.y, hi.z);
    keys[j] = ((uint64_t) seg << (3 * bits)) | morton_code(box, scene, bits);
    vals[j] = (uint32_t) j;
}

// K3: Karras hierarchy over the sorted keys
__global__ void __launch_bounds__(BT) karras_kernel(const uint64_t *__restrict__ keys, int n, const uint32_t *__restrict__ seg_offset, int n_segs,
                                                    int *__restrict__ left, int *__restrict__ right, int *__restrict__ parent,
                                                    int *__restrict__ leaf_parent, int *__restrict__ seg_root, int2 *__restrict__ range) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    const KarrasNode k = karras_node(keys, n, i);
    left[i] = k.left;
    right[i] = k.right;
    range[i] = make_int2(k.first, k.last);
    if (k.left >= 0) parent[k.left] = i; else leaf_parent[~k.left] = i;
    if (k.right >= 0) parent[k.right] = i; else leaf_parent[~k.right] = i;
    if (i == 0) parent[0] = -1;
    // subtree that covers exactly one segment = that BLAS's root
    const int seg = n_segs > 1 ? find_segment(seg_offset, n_segs, (uint32_t) k.first) : 0;
    if ((uint32_t) k.first == __ldg(&seg_offset[seg]) && (uint32_t) k.last + 1u == __ldg(&seg_offset[seg + 1])) seg_root[seg] = i;
}

// segments with a single primitive have no inner node: their root is the leaf itself
__global__ void __launch_bounds__(BT) single_prim_roots_kernel(const uint32_t *__restrict__ seg_offset, const uint32_t *__restrict__ seg_type,
                                                               int n_segs, int leaf_mode, const uint32_t *__restrict__ vals, int *__restrict__ seg_root) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_segs) return;
    if (seg_offset[s + 1] - seg_offset[s] == 1u) {
        const uint32_t pos = seg_offset[s];
        seg_root[s] = leaf_mode == 0 ? make_leaf_ref(seg_type[s], pos) : make_leaf_ref(0u, vals[pos]);
    }
}

// K4: bottom-up refit; the second child to arrive at a node computes its box.  refit_rotate_kernel (RT_FLAG_SAH_REFINE)
// first tries one SAH tree rotation at the node (best_rotation, rt_build_core.h).  The arriving thread is the only one active in
// the node's subtree (everything below is finished, nothing above has started), so rewriting the node's children and
// one child's record is race-free.  Nodes that span several segments (the junk above the BLAS roots) never rotate.
struct RefitChild { Aabb box; int kl, kr; };
__device__ __forceinline__ RefitChild refit_load(int c, const float4 *prim_lo, const float4 *prim_hi, const uint32_t *vals,
                                                 const float4 *node_lo, const float4 *node_hi) {
    RefitChild r;
    float4 lo, hi;
    if (c >= 0) { lo = __ldcg(&node_lo[c]); hi = __ldcg(&node_hi[c]); r.kl = __float_as_int(lo.w); r.kr = __float_as_int(hi.w); }
    else { const uint32_t p = vals[~c]; lo = prim_lo[p]; hi = prim_hi[p]; r.kl = 0; r.kr = 0; }
    r.box.lo = mk3(lo.x, lo.y, lo.z);
    r.box.hi = mk3(hi.x, hi.y, hi.z);
    return r;
}
__device__ __forceinline__ void refit_store(int node, const Aabb &b, int l, int r, float4 *node_lo, float4 *node_hi, int *left, int *right) {
    // .w carries the node's children so the wide collapse gets box + children in one pair of loads
    __stcg(&node_lo[node], make_float4(b.lo.x, b.lo.y, b.lo.z, __int_as_float(l)));
    __stcg(&node_hi[node], make_float4(b.hi.x, b.hi.y, b.hi.z, __int_as_float(r)));
    left[node] = l;
    right[node] = r;
}
__device__ __forceinline__ void refit_set_parent(int child, int p, int *parent, int *leaf_parent) {
    if (child >= 0) parent[child] = p; else leaf_parent[~child] = p;
}

// The plain refit (the default): no rotation, children as Karras made them.
__global__ void __launch_bounds__(BT) refit_kernel(const float4 *__restrict__ prim_lo, const float4 *__restrict__ prim_hi, const uint32_t *__restrict__ vals,
                                                   const int *__restrict__ left, const int *__restrict__ right, const int *__restrict__ parent,
                                                   const int *__restrict__ leaf_parent, int n, int *flags, float4 *node_lo, float4 *node_hi) {
    const int pos = blockIdx.x * blockDim.x + threadIdx.x;
    if (pos >= n) return;
    int cur = leaf_parent[pos];
    while (cur >= 0) {
        __threadfence();
        if (atomicAdd(&flags[cur], 1) == 0) return;
        __threadfence();
        const int l = left[cur], r = right[cur];
        float4 llo, lhi, rlo, rhi;
        if (l >= 0) { llo = __ldcg(&node_lo[l]); lhi = __ldcg(&node_hi[l]); } else { const uint32_t p = vals[~l]; llo = prim_lo[p]; lhi = prim_hi[p]; }
        if (r >= 0) { rlo = __ldcg(&node_lo[r]); rhi = __ldcg(&node_hi[r]); } else { const uint32_t p = vals[~r]; rlo = prim_lo[p]; rhi = prim_hi[p]; }
        // .w carries the node's Karras children so the wide collapse gets box + children in one pair of loads
        __stcg(&node_lo[cur], make_float4(fminf(llo.x, rlo.x), fminf(llo.y, rlo.y), fminf(llo.z, rlo.z), __int_as_float(l)));
        __stcg(&node_hi[cur], make_float4(fmaxf(lhi.x, rhi.x), fmaxf(lhi.y, rhi.y), fmaxf(lhi.z, rhi.z), __int_as_float(r)));
        cur = parent[cur];
    }
}

__global__ void __launch_bounds__(BT) refit_rotate_kernel(const float4 *__restrict__ prim_lo, const float4 *__restrict__ prim_hi, const uint32_t *__restrict__ vals,
                                                   int *left, int *right, int *parent, int *leaf_parent, int n, int *flags,
                                                   float4 *node_lo, float4 *node_hi, const int2 *__restrict__ range,
                                                   const uint32_t *__restrict__ seg_offset, int n_segs, int rotate) {
    const int pos = blockIdx.x * blockDim.x + threadIdx.x;
    if (pos >= n) return;
    int cur = leaf_parent[pos];
    while (cur >= 0) {
        __threadfence();
        if (atomicAdd(&flags[cur], 1) == 0) return;
        __threadfence();
        int l = __ldcg(&left[cur]), r = __ldcg(&right[cur]);
        RefitChild L = refit_load(l, prim_lo, prim_hi, vals, node_lo, node_hi);
        RefitChild R = refit_load(r, prim_lo, prim_hi, vals, node_lo, node_hi);
        const int up = __ldcg(&parent[cur]);
        bool may = rotate != 0 && (l >= 0 || r >= 0);
        if (may && n_segs > 1) {
            const int2 rg = range[cur];
            may = find_segment(seg_offset, n_segs, (uint32_t) rg.x) == find_segment(seg_offset, n_segs, (uint32_t) rg.y);
        }
        if (may) {
            RefitChild LL{}, LR{}, RL{}, RR{};
            if (l >= 0) { LL = refit_load(L.kl, prim_lo, prim_hi, vals, node_lo, node_hi); LR = refit_load(L.kr, prim_lo, prim_hi, vals, node_lo, node_hi); }
            if (r >= 0) { RL = refit_load(R.kl, prim_lo, prim_hi, vals, node_lo, node_hi); RR = refit_load(R.kr, prim_lo, prim_hi, vals, node_lo, node_hi); }
            Aabb nb;
            const int rot = best_rotation(L.box, R.box, l >= 0, LL.box, LR.box, r >= 0, RL.box, RR.box, nb);
            if (rot == 1) {            // R <-> LL : L = (R, LR), N = (L, LL)
                refit_store(l, nb, r, L.kr, node_lo, node_hi, left, right);
                refit_set_parent(r, l, parent, leaf_parent); refit_set_parent(L.kl, cur, parent, leaf_parent);
                r = L.kl; L.box = nb; R.box = LL.box;
            } else if (rot == 2) {     // R <-> LR : L = (LL, R), N = (L, LR)
                refit_store(l, nb, L.kl, r, node_lo, node_hi, left, right);
                refit_set_parent(r, l, parent, leaf_parent); refit_set_parent(L.kr, cur, parent, leaf_parent);
                r = L.kr; L.box = nb; R.box = LR.box;
            } else if (rot == 3) {     // L <-> RL : R = (L, RR), N = (RL, R)
                refit_store(r, nb, l, R.kr, node_lo, node_hi, left, right);
                refit_set_parent(l, r, parent, leaf_parent); refit_set_parent(R.kl, cur, parent, leaf_parent);
                l = R.kl; R.box = nb; L.box = RL.box;
            } else if (rot == 4) {     // L <-> RR : R = (RL, L), N = (RR, R)
                refit_store(r, nb, R.kl, l, node_lo, node_hi, left, right);
                refit_set_parent(l, r, parent, leaf_parent); refit_set_parent(R.kr, cur, parent, leaf_parent);
                l = R.kr; R.box = nb; L.box = RR.box;
            }
        }
        refit_store(cur, aabb_union(L.box, R.box), l, r, node_lo, node_hi, left, right);
        cur = up;
    }
}

__device__ __forceinline__ Aabb load_box(const float4 *lo, const float4 *hi, int i) {
    const float4 a = lo[i], b = hi[i];
    Aabb r;
    r.lo = mk3(a.x, a.y, a.z);
    r.hi = mk3(b.x, b.y, b.z);
    return r;
}

// K5a: traversal nodes (both child boxes in the parent)
__global__ void __launch_bounds__(BT) emit_nodes_kernel(const float4 *__restrict__ prim_lo, const float4 *__restrict__ prim_hi, const uint32_t *__restrict__ vals,
                                                        const int *__restrict__ left, const int *__restrict__ right,
                                                        const float4 *__restrict__ node_lo, const float4 *__restrict__ node_hi,
                                                        const uint32_t *__restrict__ seg_offset, const uint32_t *__restrict__ seg_type, int n_segs,
                                                        int n, int leaf_mode, BvhNode2 *__restrict__ nodes) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    const int l = left[i], r = right[i];
    Aabb b[2];
    int ref[2];
    const int child[2] = {l, r};
#pragma unroll
    for (int c = 0; c < 2; c++) {
        if (child[c] >= 0) {
            b[c] = load_box(node_lo, node_hi, child[c]);
            ref[c] = child[c];
        } else {
            const uint32_t pos = (uint32_t) ~child[c];
            const uint32_t p = vals[pos];
            b[c] = load_box(prim_lo, prim_hi, (int) p);
            if (leaf_mode == 0) {
                const int seg = n_segs > 1 ? find_segment(seg_offset, n_segs, pos) : 0;
                ref[c] = make_leaf_ref(__ldg(&seg_type[seg]), pos);
            } else {
                ref[c] = make_leaf_ref(0u, p);
            }
        }
    }
    nodes[i] = make_node2(b[0], ref[0], b[1], ref[1]);
}

// K5b: leaf-order primitive ids and triangle intersection records (e1/e2 as the reference's constructor)
__global__ void __launch_bounds__(BT) emit_leaves_kernel(GeometryD geo, const uint32_t *__restrict__ vals,
                                                         const uint32_t *__restrict__ seg_offset, const uint32_t *__restrict__ seg_type,
                                                         const uint32_t *__restrict__ seg_first, int n_segs, int n,
                                                         uint32_t *__restrict__ leaf_prims, TriIsect *__restrict__ leaf_tris) {
    const int pos = blockIdx.x * blockDim.x + threadIdx.x;
    if (pos >= n) return;
    const int seg = n_segs > 1 ? find_segment(seg_offset, n_segs, (uint32_t) pos) : 0;
    const uint32_t prim = __ldg(&seg_first[seg]) + (vals[pos] - __ldg(&seg_offset[seg]));
    leaf_prims[pos] = prim;
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
        leaf_tris[pos] = t;
    }
}

__global__ void __launch_bounds__(BT) instance_update_kernel(const InstanceXformIn *__restrict__ in, int n, InstanceD *__restrict__ out,
                                                             float4 *__restrict__ wlo, float4 *__restrict__ whi, float4 *__restrict__ wcentroid) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    InstanceD o;
    Aabb box;
    f3 c;
    instance_update(in[i], o, box, c);
    out[i] = o;
    wlo[i] = make_float4(box.lo.x, box.lo.y, box.lo.z, 0.0f);      // the reference's transformedBoundingBox; finished below
    whi[i] = make_float4(box.hi.x, box.hi.y, box.hi.z, 0.0f);
    wcentroid[i] = make_float4(c.x, c.y, c.z, 0.0f);
}

// One warp per instance: tighten the world box from the transformed vertices where asked (instance_tight_box), pad it
// for the TLAS and fold it into the scene bounds.
__global__ void __launch_bounds__(BT) instance_bounds_kernel(const InstanceXformIn *__restrict__ in, int n, const InstanceD *__restrict__ inst,
                                                             const float *__restrict__ tri_verts, float4 *__restrict__ wlo, float4 *__restrict__ whi,
                                                             uint32_t *bounds) {
    const int i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (i >= n) return;
    const int cnt = in[i].tri_count;
    const float4 a = wlo[i], b = whi[i];
    Aabb box;
    box.lo = mk3(a.x, a.y, a.z);
    box.hi = mk3(b.x, b.y, b.z);
    if (cnt > 0) {
        float fwd[12];
#pragma unroll
        for (int k = 0; k < 12; k++) fwd[k] = inst[i].fwd[k];
        Aabb v;
        v.lo = mk3(RT_INF, RT_INF, RT_INF);
        v.hi = mk3(-RT_INF, -RT_INF, -RT_INF);
        const float *p = tri_verts + (size_t) in[i].tri_first * 9;
        for (int k = lane; k < cnt * 3; k += 32) aabb_grow(v, xf_point(fwd, mk3(__ldg(p + 3 * k), __ldg(p + 3 * k + 1), __ldg(p + 3 * k + 2))));
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            v.lo.x = fminf(v.lo.x, __shfl_xor_sync(0xffffffffu, v.lo.x, off)); v.lo.y = fminf(v.lo.y, __shfl_xor_sync(0xffffffffu, v.lo.y, off));
            v.lo.z = fminf(v.lo.z, __shfl_xor_sync(0xffffffffu, v.lo.z, off)); v.hi.x = fmaxf(v.hi.x, __shfl_xor_sync(0xffffffffu, v.hi.x, off));
            v.hi.y = fmaxf(v.hi.y, __shfl_xor_sync(0xffffffffu, v.hi.y, off)); v.hi.z = fmaxf(v.hi.z, __shfl_xor_sync(0xffffffffu, v.hi.z, off));
        }
        box = instance_tight_box(box, v);
    }
    if (lane == 0) {
        const Aabb padded = aabb_pad(box);
        wlo[i] = make_float4(padded.lo.x, padded.lo.y, padded.lo.z, 0.0f);
        whi[i] = make_float4(padded.hi.x, padded.hi.y, padded.hi.z, 0.0f);
        atomic_bounds(bounds, aabb_centre_box(padded));      // the TLAS grid spans the instance centres (rt_build_core.h)
    }
}

} // namespace

void launch_instance_update(const InstanceXformIn *in, int n, InstanceD *out, float4 *wlo, float4 *whi, float4 *wcentroid,
                            uint32_t *bounds, const float *tri_verts, cudaStream_t stream) {
    init_bounds_kernel<<<1, BT, 0, stream>>>(bounds, 1);
    instance_update_kernel<<<div_up(n, 64), 64, 0, stream>>>(in, n, out, wlo, whi, wcentroid);
    instance_bounds_kernel<<<div_up((size_t) n * 32, BT), BT, 0, stream>>>(in, n, out, tri_verts, wlo, whi, bounds);
    CUDA_CHECK(cudaGetLastError());
}

void LbvhBuilder::reserve(int n, int n_segs, bool tris) {
    prim_lo_.reserve(n); prim_hi_.reserve(n);
    seg_bounds_.reserve((size_t) n_segs * 6);
    seg_offset_.reserve(n_segs + 1); seg_type_.reserve(n_segs); seg_first_.reserve(n_segs);
    keys_a_.reserve(n); keys_b_.reserve(n); vals_a_.reserve(n); vals_b_.reserve(n);
    const int m = n > 1 ? n - 1 : 1;
    left_.reserve(m); right_.reserve(m); parent_.reserve(m); flags_.reserve(m);
    node_lo_.reserve(m); node_hi_.reserve(m); nodes_.reserve(m);
    leaf_parent_.reserve(n); seg_root_.reserve(n_segs);
    leaf_prims_.reserve(n);
    if (tris) leaf_tris_.reserve(n);
    range_.reserve(m);
    queue_.reserve((size_t) n + n_segs);
    nodes8_.reserve((size_t) n + n_segs);
    wide_counters_.reserve(8);
    leaf_pos8_.reserve(n); leaf_prims8_.reserve(n);
    if (tris) leaf_tris8_.reserve(n);
}

void LbvhBuilder::run_tree(int n, int n_segs, int key_bits, int leaf_mode, const GeometryD *geo, cudaStream_t stream, int sm_count) {
    keys_ = keys_a_.get(); keys_alt_ = keys_b_.get(); vals_ = vals_a_.get(); vals_alt_ = vals_b_.get();
    if (n > 1) {
        sorter_.sort(keys_, keys_alt_, vals_, vals_alt_, n, key_bits, stream, sm_count, launches);
        if (ploc) {
            run_ploc(n, n_segs, stream);
        } else {
        CUDA_CHECK(cudaMemsetAsync(flags_.get(), 0, (size_t) (n - 1) * sizeof(int), stream));
        karras_kernel<<<div_up(n - 1, BT), BT, 0, stream>>>(keys_, n, seg_offset_.get(), n_segs, left_.get(), right_.get(), parent_.get(),
                                                           leaf_parent_.get(), seg_root_.get(), range_.get());
        if (sah_rotations)
            refit_rotate_kernel<<<div_up(n, BT), BT, 0, stream>>>(prim_lo_.get(), prim_hi_.get(), vals_, left_.get(), right_.get(), parent_.get(),
                                                                 leaf_parent_.get(), n, flags_.get(), node_lo_.get(), node_hi_.get(), range_.get(),
                                                                 seg_offset_.get(), n_segs, 1);
        else
            refit_kernel<<<div_up(n, BT), BT, 0, stream>>>(prim_lo_.get(), prim_hi_.get(), vals_, left_.get(), right_.get(), parent_.get(),
                                                          leaf_parent_.get(), n, flags_.get(), node_lo_.get(), node_hi_.get());
        launches += 2;
        }
        if (emit_bvh2) {
            emit_nodes_kernel<<<div_up(n - 1, BT), BT, 0, stream>>>(prim_lo_.get(), prim_hi_.get(), vals_, left_.get(), right_.get(),
                                                                   node_lo_.get(), node_hi_.get(), seg_offset_.get(), seg_type_.get(), n_segs, n, leaf_mode,
                                                                   nodes_.get());
            launches++;
        }
    }
    single_prim_roots_kernel<<<div_up(n_segs, BT), BT, 0, stream>>>(seg_offset_.get(), seg_type_.get(), n_segs, leaf_mode, vals_, seg_root_.get());
    launches++;
    if (leaf_mode == 0 && emit_bvh2) {
        emit_leaves_kernel<<<div_up(n, BT), BT, 0, stream>>>(*geo, vals_, seg_offset_.get(), seg_type_.get(), seg_first_.get(), n_segs, n,
                                                            leaf_prims_.get(), leaf_tris_.get());
        launches++;
    }
    CUDA_CHECK(cudaGetLastError());
    collapse_wide(n, n_segs, leaf_mode, geo, stream, sm_count);
}

void LbvhBuilder::build_blas(const GeometryD &geo, const std::vector<uint32_t> &seg_offset, const std::vector<uint32_t> &seg_type,
                             const std::vector<uint32_t> &seg_first, int morton_bits, cudaStream_t stream, int sm_count) {
    const int n_segs = (int) seg_type.size();
    const int n = (int) seg_offset.back();
    n_ = n; n_segs_ = n_segs; launches = 0;
    prepare_blas(seg_offset, seg_type);
    seg_offset_.upload(seg_offset.data(), seg_offset.size(), stream);
    seg_type_.upload(seg_type.data(), seg_type.size(), stream);
    seg_first_.upload(seg_first.data(), seg_first.size(), stream);
    int seg_bits = 0;
    while ((1 << seg_bits) < n_segs) seg_bits++;
    if (3 * morton_bits + seg_bits > 64) morton_bits = (64 - seg_bits) / 3;
    init_bounds_kernel<<<div_up((size_t) n_segs * 6, BT), BT, 0, stream>>>(seg_bounds_.get(), n_segs);
    prim_bounds_kernel<<<div_up(n, BT), BT, 0, stream>>>(geo, seg_offset_.get(), seg_type_.get(), seg_first_.get(), n_segs, n,
                                                        prim_lo_.get(), prim_hi_.get(), seg_bounds_.get());
    morton_kernel<<<div_up(n, BT), BT, 0, stream>>>(prim_lo_.get(), prim_hi_.get(), seg_offset_.get(), n_segs, seg_bounds_.get(), n, morton_bits,
                                                   keys_a_.get(), vals_a_.get());
    launches += 3;
    run_tree(n, n_segs, 3 * morton_bits + seg_bits, 0, &geo, stream, sm_count);
    seg_root_host.resize(n_segs);
    CUDA_CHECK(cudaMemcpyAsync(seg_root_host.data(), seg_root_.get(), n_segs * sizeof(int), cudaMemcpyDeviceToHost, stream));
    if (wait_event) {                    // sleep instead of spinning (animation lanes, see rt_ctx::wait in api.cu)
        CUDA_CHECK(cudaEventRecord(wait_event, stream));
        CUDA_CHECK(cudaEventSynchronize(wait_event));
    } else {
        CUDA_CHECK(cudaStreamSynchronize(stream));
    }
}

void LbvhBuilder::prepare_blas(const std::vector<uint32_t> &seg_offset, const std::vector<uint32_t> &seg_type) {
    bool tris = false;
    for (uint32_t t: seg_type) tris |= (t == 2u);
    reserve((int) seg_offset.back(), (int) seg_type.size(), tris);
}

void LbvhBuilder::prepare_tlas(int n_instances) {
    reserve(n_instances, 1, false);
}

void LbvhBuilder::build_tlas(int n_instances, int morton_bits, cudaStream_t stream, int sm_count) {
    const int n = n_instances;
    n_ = n; n_segs_ = 1; launches = 0;
    if (tlas_small(n, morton_bits, stream)) return;        // up to RT_TLAS_SMALL_MAX instances: the whole rebuild is one launch
    set_single_segment_kernel<<<1, 1, 0, stream>>>(seg_offset_.get(), seg_type_.get(), (uint32_t) n);
    launches++;
    morton_kernel<<<div_up(n, BT), BT, 0, stream>>>(prim_lo_.get(), prim_hi_.get(), seg_offset_.get(), 1, seg_bounds_.get(), n, morton_bits,
                                                   keys_a_.get(), vals_a_.get());
    launches++;
    run_tree(n, 1, 3 * morton_bits, 1, nullptr, stream, sm_count);
}


// ---- config 3 input stage: triangle strips -> triangle records, on the device (one thread per triangle of the frame;
// the per-triangle work is expand_strip_triangle, rt_build_core.h) ---------------------------------------------------
namespace {
__global__ void __launch_bounds__(256) expand_strips_kernel(const float *__restrict__ points, const float *__restrict__ normals, const int *__restrict__ ids,
                                                            const uint32_t *__restrict__ strip_first_tri, const uint32_t *__restrict__ strip_first_id,
                                                            int n_strips, uint32_t n_tris, uint32_t material_type, uint32_t material_index,
                                                            float *__restrict__ tri_verts, TriShade *__restrict__ tri_shade) {
    const uint32_t q = blockIdx.x * blockDim.x + threadIdx.x;
    if (q >= n_tris) return;
    expand_strip_triangle(q, points, normals, ids, strip_first_tri, strip_first_id, n_strips, material_type, material_index, tri_verts, tri_shade);
}
} // namespace

void launch_expand_strips(const float *points, const float *normals, const int *ids, const uint32_t *strip_first_tri, const uint32_t *strip_first_id,
                          int n_strips, uint32_t n_tris, uint32_t material_type, uint32_t material_index, float *tri_verts, TriShade *tri_shade,
                          cudaStream_t stream) {
    if (n_tris == 0 || n_strips == 0) return;
    expand_strips_kernel<<<div_up((size_t) n_tris, 256), 256, 0, stream>>>(points, normals, ids, strip_first_tri, strip_first_id, n_strips, n_tris,
                                                                         material_type, material_index, tri_verts, tri_shade);
    CUDA_CHECK(cudaGetLastError());
}
