// rt_wide_core.h — compressed 8-wide BVH: collapse of the binary LBVH into BvhNode8 records (build side) and the
// stack-based closest-hit traversal over them (render side).  Pure per-element / per-ray functions, __host__
// __device__, so the CUDA kernels and the g++ twin in tests/emul run the identical definition.
//
// Replaces the reference's binary BLAS/TLAS node arrays (include/AS/BLAS.cuh:20-34, include/AS/TLAS.cuh:24-39:
// 40-byte nodes, each box tested twice) and BLAS::hit / TLAS::hit (src/AS/BLAS.cu:119-206, src/AS/TLAS.cu:131-201).
#pragma once

#include "rt_build_core.h"

// ---- small bit helpers ---------------------------------------------------------------------------------------
RT_HD int rt_popc(uint32_t x) {
#if defined(__CUDA_ARCH__)
    return __popc(x);
#else
    return __builtin_popcount(x);
#endif
}
RT_HD int rt_bfind(uint32_t x) {   // index of the highest set bit, x != 0
#if defined(__CUDA_ARCH__)
    return 31 - __clz((int) x);
#else
    return 31 - __builtin_clz(x);
#endif
}
RT_HD uint32_t rt_byte(uint32_t v, int i) { return (v >> (8 * i)) & 0xFFu; }

#ifndef RT_COLLAPSE_PREFETCH
#define RT_COLLAPSE_PREFETCH 0      /* measured: 0.775 vs 0.780 ms for the 1 M-triangle BLAS (profiles/r02h_build_ab.txt): no effect, off */
#endif

// ---- collapse: children of one wide node --------------------------------------------------------------------
// ref >= 0 : binary inner node that becomes a wide node of its own
// ref <  0 : leaf child holding the Morton-order positions [first, first + count)
struct WideChildren {
    int n;
    int ref[8];
    int first[8], count[8];
    int kl[8], kr[8];                   // Karras children of ref[k] (valid when ref[k] >= 0)
    Aabb box[8];
};

// The binary tree as the refit leaves it: node_lo[i].w / node_hi[i].w carry the bits of the Karras children of
// node i (>= 0 inner, < 0 leaf ~position), so ONE pair of 16-byte loads gives a node's box and its children and
// opening a node is a single memory round trip.
struct BinaryTreeView {
    const int2 *range;                  // per inner node: first/last Morton position
    const float4 *node_lo, *node_hi;    // per inner node box (+ children in .w)
    const float4 *prim_lo, *prim_hi;    // per build primitive box (padded)
    const uint32_t *vals;               // Morton position -> build primitive
};

RT_HD Aabb view_box(const float4 *lo, const float4 *hi, int i) {
    const float4 a = lo[i], b = hi[i];
    Aabb r;
    r.lo = mk3(a.x, a.y, a.z);
    r.hi = mk3(b.x, b.y, b.z);
    return r;
}
// The collapse is a chain of dependent loads: opening a child needs the records of ITS children.  Whichever child is
// opened next, those records are requested the moment the child is added (L2 prefetch, no registers held), so that the
// later load finds them on chip instead of paying a DRAM round trip on the critical path.
RT_HD void wide_prefetch_node(const BinaryTreeView &t, int node) {
#if defined(__CUDA_ARCH__) && RT_COLLAPSE_PREFETCH
    if (node >= 0) {
        asm volatile("prefetch.global.L2 [%0];" :: "l"(t.node_lo + node));
        asm volatile("prefetch.global.L2 [%0];" :: "l"(t.node_hi + node));
    }
#else
    (void) t; (void) node;
#endif
}
RT_HD void wide_set_child(const BinaryTreeView &t, int child, int leaf_max, WideChildren &w, int k) {
    if (child >= 0) {
        const float4 a = t.node_lo[child], b = t.node_hi[child];
        w.box[k].lo = mk3(a.x, a.y, a.z);
        w.box[k].hi = mk3(b.x, b.y, b.z);
        w.kl[k] = (int) f2u(a.w);
        w.kr[k] = (int) f2u(b.w);
        wide_prefetch_node(t, w.kl[k]);
        wide_prefetch_node(t, w.kr[k]);
        int first = 0, cnt = leaf_max + 1;
        if (leaf_max > 1) {                                  // small subtrees become one leaf child
            const int2 rg = t.range[child];
            first = rg.x;
            cnt = rg.y - rg.x + 1;
        }
        w.first[k] = first;
        w.count[k] = cnt;
        w.ref[k] = cnt <= leaf_max ? -1 : child;
    } else {
        const int pos = ~child;
        w.box[k] = view_box(t.prim_lo, t.prim_hi, (int) t.vals[pos]);
        w.first[k] = pos;
        w.count[k] = 1;
        w.ref[k] = -1;
        w.kl[k] = 0; w.kr[k] = 0;
    }
}
// Greedy top-down collapse: start from the two children of `root`, repeatedly open the inner child with the
// largest surface area until eight children exist or nothing can be opened.  root < 0: single-leaf node.
RT_HD void wide_expand(const BinaryTreeView &t, int root, int leaf_max, WideChildren &w) {
    if (root < 0) {
        w.n = 1;
        wide_set_child(t, root, leaf_max, w, 0);
        return;
    }
    wide_set_child(t, root, leaf_max, w, 0);
    if (w.ref[0] < 0) {                               // whole BLAS fits one leaf child
        w.n = 1;
        return;
    }
    {
        const int l = w.kl[0], r = w.kr[0];
        w.n = 2;
        wide_set_child(t, l, leaf_max, w, 0);
        wide_set_child(t, r, leaf_max, w, 1);
    }
    while (w.n < 8) {
        int best = -1;
        float best_area = -1.0f;
        for (int k = 0; k < w.n; k++) {
            if (w.ref[k] < 0) continue;
            const float a = aabb_half_area(w.box[k]);
            if (a > best_area) { best_area = a; best = k; }
        }
        if (best < 0) break;
        const int l = w.kl[best], r = w.kr[best];
        wide_set_child(t, l, leaf_max, w, best);
        wide_set_child(t, r, leaf_max, w, w.n);
        w.n++;
    }
}

// ---- octant-ordered slot assignment --------------------------------------------------------------------------
// Slot s (bit 2 = +x, bit 1 = +y, bit 0 = +z) should hold the child lying furthest towards that corner of the
// node, so that for a ray of octant `oct` visiting slots in descending (s ^ (7 - oct)) order is front-to-back.
// cost(child, slot) = dot(child centre - node centre, corner direction of the slot).
RT_HD void wide_assign_slots(const WideChildren &w, const Aabb &node, int *slot_of /* [8] child -> slot */) {
    // Children in descending order of the best cost they can reach (|dx| + |dy| + |dz|, ties: lower index first);
    // each takes the free slot with its highest cost (ties: lower slot).  O(8 * 8) — the assignment sits on the
    // collapse's dependency chain (a node's children cannot start before it is done), so it has to be short.
    const f3 nc = mk3(0.5f * (node.lo.x + node.hi.x), 0.5f * (node.lo.y + node.hi.y), 0.5f * (node.lo.z + node.hi.z));
    f3 d[8];
    float prio[8];
    for (int c = 0; c < 8; c++) {
        slot_of[c] = -1;
        if (c < w.n) {
            d[c] = mk3(0.5f * (w.box[c].lo.x + w.box[c].hi.x) - nc.x, 0.5f * (w.box[c].lo.y + w.box[c].hi.y) - nc.y,
                       0.5f * (w.box[c].lo.z + w.box[c].hi.z) - nc.z);
            prio[c] = (fabsf(d[c].x) + fabsf(d[c].y)) + fabsf(d[c].z);
        } else {
            d[c] = mk3(0.0f, 0.0f, 0.0f);
            prio[c] = -1.0f;
        }
    }
    uint32_t child_free = (1u << w.n) - 1u, slot_free = 0xFFu;
    for (int round = 0; round < w.n; round++) {
        int bc = -1;
        float bp = -2.0f;
        for (int c = 0; c < 8; c++)
            if ((child_free & (1u << c)) && prio[c] > bp) { bp = prio[c]; bc = c; }
        int bs = -1;
        float best = -RT_INF;
        for (int s = 0; s < 8; s++) {
            if (!(slot_free & (1u << s))) continue;
            const float cost = (((s & 4) ? d[bc].x : -d[bc].x) + ((s & 2) ? d[bc].y : -d[bc].y)) + ((s & 1) ? d[bc].z : -d[bc].z);
            if (cost > best) { best = cost; bs = s; }
        }
        slot_of[bc] = bs;
        child_free &= ~(1u << bc);
        slot_free &= ~(1u << bs);
    }
}

// ---- quantisation grid ---------------------------------------------------------------------------------------
// Biased exponent byte e with step = 2^(e-127) such that extent / step <= 253 (one spare unit at each end).
RT_HD uint32_t wide_grid_exponent(float extent) {
    if (!(extent > 0.0f)) return 27u;                       // degenerate axis: every q is 0, any finite step works
    const float f = extent / 253.0f;
    uint32_t e = (f2u(f) >> 23) & 0xFFu;
    if (f2u(f) & 0x7FFFFFu) e++;                            // round the step up to a power of two
    if (e < 1u) e = 1u;
    if (e > 253u) e = 253u;
    return e;
}
RT_HD float wide_grid_step(uint32_t e) { return u2f(e << 23); }
// step is a power of two, so scaling by its reciprocal `inv` is exact (no division on the collapse's critical path)
RT_HD float wide_grid_inv_step(uint32_t e) { return u2f((254u - e) << 23); }
RT_HD uint32_t wide_quant_lo(float v, float p, float step, float inv) {
    float q = floorf((v - p) * inv);
    if (q < 0.0f) q = 0.0f;
    if (q > 255.0f) q = 255.0f;
    uint32_t qi = (uint32_t) q;
    while (qi > 0u && p + (float) qi * step > v) qi--;      // decoded plane never inside the box
    return qi;
}
RT_HD uint32_t wide_quant_hi(float v, float p, float step, float inv) {
    float q = ceilf((v - p) * inv);
    if (q < 0.0f) q = 0.0f;
    if (q > 255.0f) q = 255.0f;
    uint32_t qi = (uint32_t) q;
    while (qi < 255u && p + (float) qi * step < v) qi++;
    return qi;
}

// ---- encode one wide node ------------------------------------------------------------------------------------
// child_base: index of the first inner child (inner children are numbered in slot order from it),
// prim_base : leaf-order position of the node's first leaf primitive (leaf children in slot order).
// Outputs the node, and for each slot the child placed there (child_in_slot[s], -1 = empty) so the caller can
// enqueue inner children / write leaf records in the same order.
// Split in two so that the collapse can publish a node's children (which needs only the slot order) before it spends
// time on the quantisation: wide_layout = node box + slot assignment, wide_encode_layout = the record.
RT_HD void wide_layout(const WideChildren &w, Aabb &node, int *child_in_slot /* [8] */) {
    node = w.box[0];
    for (int c = 1; c < w.n; c++) node = aabb_union(node, w.box[c]);
    int slot_of[8];
    wide_assign_slots(w, node, slot_of);
    for (int s = 0; s < 8; s++) child_in_slot[s] = -1;
    for (int c = 0; c < w.n; c++) child_in_slot[slot_of[c]] = c;
}
RT_HD void wide_encode_layout(const WideChildren &w, const Aabb &node, const int *child_in_slot, uint32_t child_base, uint32_t prim_base, BvhNode8 &out) {
    const uint32_t ex = wide_grid_exponent(node.hi.x - node.lo.x), ey = wide_grid_exponent(node.hi.y - node.lo.y),
                   ez = wide_grid_exponent(node.hi.z - node.lo.z);
    const float sx = wide_grid_step(ex), sy = wide_grid_step(ey), sz = wide_grid_step(ez);
    const float ix = wide_grid_inv_step(ex), iy = wide_grid_inv_step(ey), iz = wide_grid_inv_step(ez);
    uint32_t imask = 0, meta[2] = {0, 0}, q[6][2] = {{0, 0}, {0, 0}, {0, 0}, {0, 0}, {0, 0}, {0, 0}};
    uint32_t prim_off = 0;
    for (int s = 0; s < 8; s++) {
        const int c = child_in_slot[s];
        if (c < 0) continue;
        uint32_t m;
        if (w.ref[c] >= 0) {
            imask |= 1u << s;
            m = (1u << 5) | (24u + (uint32_t) s);
        } else {
            const uint32_t unary = (1u << w.count[c]) - 1u;   // 1 -> 001, 2 -> 011, 3 -> 111
            m = (unary << 5) | prim_off;
            prim_off += (uint32_t) w.count[c];
        }
        const int h = s >> 2, sh = 8 * (s & 3);
        meta[h] |= m << sh;
        const Aabb &b = w.box[c];
        q[0][h] |= wide_quant_lo(b.lo.x, node.lo.x, sx, ix) << sh;
        q[1][h] |= wide_quant_lo(b.lo.y, node.lo.y, sy, iy) << sh;
        q[2][h] |= wide_quant_lo(b.lo.z, node.lo.z, sz, iz) << sh;
        q[3][h] |= wide_quant_hi(b.hi.x, node.lo.x, sx, ix) << sh;
        q[4][h] |= wide_quant_hi(b.hi.y, node.lo.y, sy, iy) << sh;
        q[5][h] |= wide_quant_hi(b.hi.z, node.lo.z, sz, iz) << sh;
    }
    out.n0 = make_float4(node.lo.x, node.lo.y, node.lo.z, u2f(ex | (ey << 8) | (ez << 16) | (imask << 24)));
    out.n1 = make_uint4(child_base, prim_base, meta[0], meta[1]);
    out.n2 = make_uint4(q[0][0], q[0][1], q[1][0], q[1][1]);
    out.n3 = make_uint4(q[2][0], q[2][1], q[3][0], q[3][1]);
    out.n4 = make_uint4(q[4][0], q[4][1], q[5][0], q[5][1]);
}
RT_HD void wide_encode(const WideChildren &w, uint32_t child_base, uint32_t prim_base, BvhNode8 &out, int *child_in_slot /* [8] */) {
    Aabb node;
    wide_layout(w, node, child_in_slot);
    wide_encode_layout(w, node, child_in_slot, child_base, prim_base, out);
}
RT_HD int wide_inner_count(const WideChildren &w) { int k = 0; for (int c = 0; c < w.n; c++) k += w.ref[c] >= 0; return k; }
RT_HD int wide_prim_count(const WideChildren &w) { int k = 0; for (int c = 0; c < w.n; c++) if (w.ref[c] < 0) k += w.count[c]; return k; }
