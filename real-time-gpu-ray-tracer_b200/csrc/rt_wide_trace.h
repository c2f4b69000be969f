// rt_wide_trace.h — closest-hit traversal of the two-level compressed 8-wide BVH (see rt_types.h BvhNode8).
// One thread = one ray.  Traversal state is a node group (base index | 8 hit bits | imask), a primitive group
// (base | 24 hit bits) and a stack of 8-byte groups, after Ylitie et al. 2017; instances are the "primitives"
// of the TLAS and are entered by transforming the ray (Instance::hit, reference src/AS/Instance.cu:19-50).
// Does the job of TLAS::hit -> Instance::hit -> BLAS::hit (reference src/AS/TLAS.cu:131-201, src/AS/BLAS.cu:119-206);
// acceptance of a primitive candidate keeps the reference's epsilon-closed rule (rt_intersect.h).
#pragma once

#include "rt_trace_core.h"
#include "rt_wide_core.h"


#ifndef RT_NODE_FFMA2
#define RT_NODE_FFMA2 0
#endif

struct WideRay {
    RayD ray;
    f3 idir;
    uint32_t oct_inv4;
};
RT_HD void wide_setup_ray(WideRay &r) {
    r.idir = mk3(safe_rcp(r.ray.d.x), safe_rcp(r.ray.d.y), safe_rcp(r.ray.d.z));
    const uint32_t oct = (r.ray.d.x < 0.0f ? 4u : 0u) | (r.ray.d.y < 0.0f ? 2u : 0u) | (r.ray.d.z < 0.0f ? 1u : 0u);
    r.oct_inv4 = (7u - oct) * 0x01010101u;
}

RT_HD BvhNode8 load_node8(const BvhNode8 *nodes, uint32_t idx) {
    BvhNode8 n;
#if defined(__CUDA_ARCH__)
    const float4 *p = reinterpret_cast<const float4 *>(nodes + idx);
    n.n0 = __ldg(p);
    n.n1 = __ldg(reinterpret_cast<const uint4 *>(p + 1));
    n.n2 = __ldg(reinterpret_cast<const uint4 *>(p + 2));
    n.n3 = __ldg(reinterpret_cast<const uint4 *>(p + 3));
    n.n4 = __ldg(reinterpret_cast<const uint4 *>(p + 4));
#else
    n = nodes[idx];
#endif
    return n;
}

// 0x80 in byte j -> 0xFF in byte j
RT_HD uint32_t sign_extend_s8x4(uint32_t v) {
#if defined(__CUDA_ARCH__)
    return __vcmpgeu4(v, 0x80808080u);
#else
    uint32_t r = 0;
    for (int j = 0; j < 4; j++) if (v & (0x80u << (8 * j))) r |= 0xFFu << (8 * j);
    return r;
#endif
}

// Byte j of w as the float 1 + q * 2^-15 (exact): the byte is dropped into mantissa bits 8..15 of 1.0f by ONE byte
// permute, instead of an int->float conversion (the conversion pipe was the busiest unit of the first wide kernel,
// profiles/r01c: 48 conversions per node).
RT_HD float q_to_unit(uint32_t w, uint32_t one_bits, int j) {
#if defined(__CUDA_ARCH__)
    return __uint_as_float(__byte_perm(w, one_bits, 0x7604u | ((uint32_t) j << 4)));
#else
    return u2f(one_bits | (((w >> (8 * j)) & 0xFFu) << 8));
#endif
}
// 0x3F800000 held in a register the compiler cannot see through: PRMT takes only one immediate, and it must be the
// selector — with the constant as the immediate every permute needed an extra move for its selector.
#if defined(__CUDACC__)
static __constant__ uint32_t c_one_bits = 0x3F800000u;
#endif
RT_HD uint32_t opaque_one_bits() {
#if defined(__CUDA_ARCH__)
    return c_one_bits;
#else
    return 0x3F800000u;
#endif
}

// Ray vs the eight quantised child boxes of a node.  Returns the hit mask: bits 24..31 = inner children in
// traversal priority order (slot ^ octant), bits 0..23 = leaf primitives (offset from prim_base).
//
// Per axis the slab distance of grid plane q is  t(q) = q * a + o,  a = step * idir,  o = (p - origin) * idir.  It is
// evaluated as fma(m, A, O) with m = 1 + q * 2^-15 (q_to_unit), A = 2^15 * a, O = o - A, i.e. one fma per plane and
// no conversion.  The result is made conservative by one absolute bound per node and axis,
//   E = 2^-21 |o| + 2^-22 |A|     (>= the 2^-22-relative error of a and o carried over from the first version, plus the
//                                   roundings of O = o - A, O -/+ E and of the fma; 2^-22 |A| = 1/128 grid step),
// subtracted from every entry distance and added to every exit distance: a child the exact arithmetic hits is never culled.
RT_HD uint32_t node8_intersect(const BvhNode8 &n, const WideRay &r, float tmin, float tbest) {
    const uint32_t e = f2u(n.n0.w);
    const float Ax = u2f(((e & 0xFFu) + 15u) << 23) * r.idir.x, Ay = u2f((((e >> 8) & 0xFFu) + 15u) << 23) * r.idir.y,
                Az = u2f((((e >> 16) & 0xFFu) + 15u) << 23) * r.idir.z;
    const float ox = (n.n0.x - r.ray.o.x) * r.idir.x, oy = (n.n0.y - r.ray.o.y) * r.idir.y, oz = (n.n0.z - r.ray.o.z) * r.idir.z;
    const float Ox = ox - Ax, Oy = oy - Ay, Oz = oz - Az;
    const float Ex = rt_fma(4.76837158e-7f, fabsf(ox), 2.38418579e-7f * fabsf(Ax));
    const float Ey = rt_fma(4.76837158e-7f, fabsf(oy), 2.38418579e-7f * fabsf(Ay));
    const float Ez = rt_fma(4.76837158e-7f, fabsf(oz), 2.38418579e-7f * fabsf(Az));
    const float oxl = Ox - Ex, oxh = Ox + Ex, oyl = Oy - Ey, oyh = Oy + Ey, ozl = Oz - Ez, ozh = Oz + Ez;
    const bool nx = r.ray.d.x < 0.0f, ny = r.ray.d.y < 0.0f, nz = r.ray.d.z < 0.0f;
    const uint32_t one = opaque_one_bits();
    uint32_t hitmask = 0;
#pragma unroll
    for (int h = 0; h < 2; h++) {
        const uint32_t meta4 = h ? n.n1.w : n.n1.z;
        const uint32_t is_inner4 = (meta4 & (meta4 << 1)) & 0x10101010u;
        const uint32_t inner_mask4 = sign_extend_s8x4(is_inner4 << 3);
        const uint32_t bit_index4 = (meta4 ^ (r.oct_inv4 & inner_mask4)) & 0x1F1F1F1Fu;
        const uint32_t child_bits4 = (meta4 >> 5) & 0x07070707u;
        const uint32_t qlx = h ? n.n2.y : n.n2.x, qly = h ? n.n2.w : n.n2.z, qlz = h ? n.n3.y : n.n3.x;
        const uint32_t qhx = h ? n.n3.w : n.n3.z, qhy = h ? n.n4.y : n.n4.x, qhz = h ? n.n4.w : n.n4.z;
        const uint32_t x0 = nx ? qhx : qlx, x1 = nx ? qlx : qhx;
        const uint32_t y0 = ny ? qhy : qly, y1 = ny ? qly : qhy;
        const uint32_t z0 = nz ? qhz : qlz, z1 = nz ? qlz : qhz;
#pragma unroll
        for (int j = 0; j < 4; j++) {
#if RT_NODE_FFMA2 && defined(__CUDA_ARCH__)
            // variant (A/B): the near and far plane of an axis in ONE packed fp32 FMA (sm_100 FFMA2: two independent IEEE fmas,
            // same results) — 24 instead of 48 FMA-pipe issue slots per node
            const float2 tx = __ffma2_rn(make_float2(q_to_unit(x0, one, j), q_to_unit(x1, one, j)), make_float2(Ax, Ax), make_float2(oxl, oxh));
            const float2 ty = __ffma2_rn(make_float2(q_to_unit(y0, one, j), q_to_unit(y1, one, j)), make_float2(Ay, Ay), make_float2(oyl, oyh));
            const float2 tz = __ffma2_rn(make_float2(q_to_unit(z0, one, j), q_to_unit(z1, one, j)), make_float2(Az, Az), make_float2(ozl, ozh));
            const float tnx = tx.x, tfx = tx.y, tny = ty.x, tfy = ty.y, tnz = tz.x, tfz = tz.y;
#else
            const float tnx = rt_fma(q_to_unit(x0, one, j), Ax, oxl), tfx = rt_fma(q_to_unit(x1, one, j), Ax, oxh);
            const float tny = rt_fma(q_to_unit(y0, one, j), Ay, oyl), tfy = rt_fma(q_to_unit(y1, one, j), Ay, oyh);
            const float tnz = rt_fma(q_to_unit(z0, one, j), Az, ozl), tfz = rt_fma(q_to_unit(z1, one, j), Az, ozh);
#endif
            const float tn = rt_max(rt_max(tnx, tny), rt_max(tnz, tmin));
            const float tf = rt_min(rt_min(tfx, tfy), rt_min(tfz, tbest));
            if (tn <= tf) hitmask |= rt_byte(child_bits4, j) << rt_byte(bit_index4, j);
        }
    }
    return hitmask;
}

RT_HD void load_instance_wide(const SceneD &sc, int inst, float *inv, int &root8, int &prim_type, uint32_t inst_base = 0u) {
#if defined(__CUDA_ARCH__)
    const float4 *p = reinterpret_cast<const float4 *>(sc.instances + inst_base + (uint32_t) inst);
    const float4 r0 = __ldg(p), r1 = __ldg(p + 1), r2 = __ldg(p + 2);
    const int4 m = __ldg(reinterpret_cast<const int4 *>(p + 6));
    inv[0] = r0.x; inv[1] = r0.y; inv[2] = r0.z; inv[3] = r0.w;
    inv[4] = r1.x; inv[5] = r1.y; inv[6] = r1.z; inv[7] = r1.w;
    inv[8] = r2.x; inv[9] = r2.y; inv[10] = r2.z; inv[11] = r2.w;
    root8 = m.z; prim_type = m.w;
#else
    for (int i = 0; i < 12; i++) inv[i] = sc.instances[inst].inv[i];
    root8 = sc.instances[inst].root8;
    prim_type = sc.instances[inst].prim_type;
#endif
}

// One leaf primitive (wide leaf order) against the LOCAL ray: dispatched through the primitive table (rt_prims.h)
RT_HD void leaf_test8(const SceneD &sc, uint32_t type, uint32_t pos, const RayD &lray, float tmin, int inst, HitD &best) {
    prim_hit<true>(sc, type, pos, lray, tmin, inst, best);
}

// Stack of 8-byte groups; `stride` lets the CUDA kernels interleave the stacks of a block in shared memory.
struct WideStack {
    uint2 *base;
    int stride;
    int sp;
};
RT_HD void wide_push(WideStack &s, const uint2 &v, uint32_t &overflow) {
    if (s.sp < RT_WIDE_STACK) { s.base[s.sp * s.stride] = v; s.sp++; } else overflow = 1u;
}
RT_HD uint2 wide_pop(WideStack &s) { s.sp--; return s.base[s.sp * s.stride]; }

// Closest hit of a world ray over [tmin, +inf) on the wide two-level BVH.
template<bool COUNT>
RT_HD void trace_closest8(const SceneD &sc, const RayD &wray, float tmin, HitD &hit, uint32_t &overflow, TravStats &stats,
                          uint2 *stack_mem, int stack_stride) {
    hit.t = RT_INF; hit.tc = RT_INF; hit.u = 0.0f; hit.v = 0.0f; hit.inst = -1; hit.leaf = 0;
    WideStack st;
    st.base = stack_mem; st.stride = stack_stride; st.sp = 0;
    WideRay r;
    r.ray = wray;
    wide_setup_ray(r);
    const BvhNode8 *nodes = sc.tlas_nodes8;
    int inst = -1, inst_sp = 0, prim_type = 0;
    uint2 ng = make_uint2(0u, 0x80000000u);      // root: one pending inner hit, empty imask -> relative index 0
    uint2 tg = make_uint2(0u, 0u);
    for (;;) {
        if (ng.y > 0x00FFFFFFu) {
            const uint32_t hits_imask = ng.y;
            const int bit = rt_bfind(hits_imask);
            const uint32_t base = ng.x;
            ng.y &= ~(1u << bit);
            if (ng.y > 0x00FFFFFFu) wide_push(st, ng, overflow);
            const uint32_t slot = ((uint32_t) bit - 24u) ^ (r.oct_inv4 & 0xFFu);
            const uint32_t rel = (uint32_t) rt_popc(hits_imask & ~(0xFFFFFFFFu << slot));
            const BvhNode8 n = load_node8(nodes, base + rel);
            if (COUNT) stats.nodes++;
            const uint32_t hm = node8_intersect(n, r, tmin, hit.tc);
            ng.x = n.n1.x;
            ng.y = (hm & 0xFF000000u) | (f2u(n.n0.w) >> 24);
            tg.x = n.n1.y;
            tg.y = hm & 0x00FFFFFFu;
        } else {
            tg = ng;                               // a postponed primitive group popped from the stack
            ng = make_uint2(0u, 0u);
        }
        if (inst < 0) {
            if (tg.y) {
                // TLAS leaf: enter one instance, keep the rest of the group and the node group for later
                const int idx = rt_bfind(tg.y);
                tg.y &= ~(1u << idx);
                if (tg.y) wide_push(st, tg, overflow);
                if (ng.y > 0x00FFFFFFu) wide_push(st, ng, overflow);
#if defined(__CUDA_ARCH__)
                inst = (int) __ldg(&sc.tlas_leaf_inst[tg.x + (uint32_t) idx]);
#else
                inst = (int) sc.tlas_leaf_inst[tg.x + (uint32_t) idx];
#endif
                if (COUNT) stats.insts++;
                float inv[12];
                int root8;
                load_instance_wide(sc, inst, inv, root8, prim_type);
                r.ray.o = xf_point(inv, wray.o);            // Instance::hit: origin as point, direction as vector,
                r.ray.d = xf_vector(inv, wray.d);           // not re-normalised (reference src/AS/Instance.cu:26-28)
                wide_setup_ray(r);
                nodes = sc.blas_nodes8;
                inst_sp = st.sp;
                ng = make_uint2((uint32_t) root8, 0x80000000u);
                tg = make_uint2(0u, 0u);
                continue;
            }
        } else {
            while (tg.y) {
                const int idx = rt_bfind(tg.y);
                tg.y &= ~(1u << idx);
                if (COUNT) stats.prims++;
                leaf_test8(sc, (uint32_t) prim_type, tg.x + (uint32_t) idx, r.ray, tmin, inst, hit);
            }
        }
        if (ng.y <= 0x00FFFFFFu) {
            if (inst >= 0 && st.sp == inst_sp) {            // BLAS exhausted: back to world space
                inst = -1;
                r.ray = wray;
                wide_setup_ray(r);
                nodes = sc.tlas_nodes8;
            }
            if (st.sp == 0) break;
            ng = wide_pop(st);
        }
    }
}
