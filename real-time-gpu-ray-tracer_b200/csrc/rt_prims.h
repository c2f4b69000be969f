// rt_prims.h — THE primitive table: the one place a geometry type is defined.
//
// The reference documents a ten-step procedure touching ten files to add a geometry type (files/AddGeometryType.txt:1-35:
// new class, enum member, BLAS build switch, BLAS::hit switch, Instance / TLAS signatures, two structs, six allocation /
// copy functions, rayColor's argument list, every scene initialiser).  Here a type is ONE row of RT_PRIMITIVE_TABLE and
// ONE specialisation of PrimOps<type>:
//
//     bounds (geo, prim)                                   true box of primitive `prim` (the builder pads it)
//     hit    (sc, pos, local ray, tmin, inst, best)        candidate at leaf-order position `pos`: accept with the reference's
//                                                          epsilon-closed rule and update `best` (best.leaf = make_leaf_ref(type, pos))
//     surface(sc, prim, local ray, local point, hit, ...)  the hit-record fields Prim::hit writes: outward normal turned
//                                                          against the ray, hitFrontFace, material type + index
//     ref_box / centroid (host scene, prim)                instance-local box and centroid with the reference's semantics
//     count  (host scene)                                  primitives of the type that are committed
//
// Everything else — the batched LBVH / PLOC build, the wide collapse, the TLAS, the persistent path kernel, the wave and
// one-thread-per-pixel comparison kernels, the CPU twin — dispatches through prim_bounds / prim_hit / prim_surface below
// and needs no change: a BLAS holds one primitive type (InstanceD::prim_type), leaf references carry it in two bits.
//
// A fourth type plugs in at compile time: -DRT_EXT_PRIMITIVE='"header.h"' names a header that defines ExtPrimD (its
// 16-byte-aligned device record), RT_EXT_PRIMITIVE_NAME and PrimOps<3>; its records arrive through
// rt_commit_extension_primitives.  tests/ext/rt_prim_disc.h is such a type (a disc), built only into the test variant
// librtb200_disc.so and checked against the oracle port's independent disc (tests/test_ext_primitive.py).
#pragma once

#include "rt_build_core.h"
#include "rt_intersect.h"

struct HitD {
    float t, u, v;
    float tc;       // culling / acceptance bound: == t, except after a bug-compatible parallelogram hit (rt_intersect.h quad_hit_t)
    int inst;       // instance index, -1 = miss
    int leaf;       // leaf ref of the winning primitive (type + leaf-order position)
};

// Geometry sources of a BLAS batch (device pointers, original order)
struct GeometryD {
    const float *tri_verts;     // 9 floats per triangle
    const SphereD *spheres;
    const QuadD *quads;
    const void *ext;            // records of the extension type (RT_EXT_PRIMITIVE), else null
};

template<int TYPE> struct PrimOps;

RT_HD f3 ld3(const float4 &a) { return mk3(a.x, a.y, a.z); }

// ---- SPHERE (reference src/Geometry/Sphere.cu, include/Geometry/Sphere.cuh) ------------------------------------------
template<> struct PrimOps<0> {
    RT_HD static Aabb bounds(const GeometryD &g, uint32_t prim) { return sphere_aabb(g.spheres[prim]); }
    template<bool WIDE> RT_HD static void hit(const SceneD &sc, uint32_t pos, const RayD &lray, float tmin, int inst, HitD &best) {
        const SphereD s = sc.spheres[WIDE ? sc.leaf_prims8[pos] : sc.leaf_prims[pos]];
        float t;
        if (sphere_hit_t(mk3(s.c_r.x, s.c_r.y, s.c_r.z), s.c_r.w, lray, tmin, best.tc, t)) {
            best.t = t; best.tc = t; best.u = 0.0f; best.v = 0.0f; best.inst = inst; best.leaf = make_leaf_ref(0u, pos);
        }
    }
    RT_HD static void surface(const SceneD &sc, uint32_t prim, const RayD &lray, const f3 &lp, const HitD &, f3 &ln, bool &front, uint32_t &mt, uint32_t &mi) {
        const SphereD s = sc.spheres[prim];
        const f3 outward = unit3(lp - ld3(s.c_r));
        front = dot3(lray.d, outward) < 0.0f;
        ln = front ? outward : -outward;
        mt = s.mat_type; mi = s.mat_index;
    }
};

// ---- PARALLELOGRAM (reference src/Geometry/Parallelogram.cu) -----------------------------------------------------------
template<> struct PrimOps<1> {
    RT_HD static Aabb bounds(const GeometryD &g, uint32_t prim) { return quad_aabb(g.quads[prim]); }
    template<bool WIDE> RT_HD static void hit(const SceneD &sc, uint32_t pos, const RayD &lray, float tmin, int inst, HitD &best) {
        const QuadD q = sc.quads[WIDE ? sc.leaf_prims8[pos] : sc.leaf_prims[pos]];
        float t, tc;
        if (quad_hit_t(q, lray, tmin, best.tc, t, tc)) {
            best.t = t; best.tc = tc; best.u = 0.0f; best.v = 0.0f; best.inst = inst; best.leaf = make_leaf_ref(1u, pos);
        }
    }
    RT_HD static void surface(const SceneD &sc, uint32_t prim, const RayD &lray, const f3 &, const HitD &, f3 &ln, bool &front, uint32_t &mt, uint32_t &mi) {
        const QuadD q = sc.quads[prim];
        const f3 n = ld3(q.n_flag);
        front = dot3(lray.d, n) < 0.0;
        ln = front ? n : -n;
        mt = q.mat_type; mi = q.mat_index;
    }
};

// ---- TRIANGLE (reference src/Geometry/Triangle.cu): the one type with leaf-ordered intersection records ------------------
template<> struct PrimOps<2> {
    RT_HD static Aabb bounds(const GeometryD &g, uint32_t prim) {
        float v9[9];
        const float *src = g.tri_verts + (size_t) prim * 9;
#if defined(__CUDA_ARCH__)
#pragma unroll
        for (int k = 0; k < 9; k++) v9[k] = __ldg(src + k);
#else
        for (int k = 0; k < 9; k++) v9[k] = src[k];
#endif
        return tri_aabb(v9);
    }
    template<bool WIDE> RT_HD static void hit(const SceneD &sc, uint32_t pos, const RayD &lray, float tmin, int inst, HitD &best) {
        const TriIsect *rec = (WIDE ? sc.leaf_tris8 : sc.leaf_tris) + pos;
#if defined(__CUDA_ARCH__)
        const float4 *p = reinterpret_cast<const float4 *>(rec);
        const float4 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2);
#else
        const float4 a = rec->v0_id, b = rec->e1, c = rec->e2;
#endif
        float t, u, v;
        if (tri_hit_t(mk3(a.x, a.y, a.z), mk3(b.x, b.y, b.z), mk3(c.x, c.y, c.z), lray, tmin, best.tc, t, u, v)) {
            best.t = t; best.tc = t; best.u = u; best.v = v; best.inst = inst; best.leaf = make_leaf_ref(2u, pos);
        }
    }
    RT_HD static void surface(const SceneD &sc, uint32_t prim, const RayD &lray, const f3 &, const HitD &hit, f3 &ln, bool &front, uint32_t &mt, uint32_t &mi) {
        const TriShade ts = sc.tri_shade[prim];
        const f3 n = unit3((ld3(ts.n0_mt) * (1.0f - hit.u - hit.v) + ld3(ts.n1_mi) * hit.u) + ld3(ts.n2) * hit.v);
        front = dot3(lray.d, n) < 0.0f;
        ln = front ? n : -n;
        mt = f2u(ts.n0_mt.w); mi = f2u(ts.n1_mi.w);
    }
};

#if defined(RT_EXT_PRIMITIVE)
#include RT_EXT_PRIMITIVE
#define RT_PRIMITIVE_TYPE_COUNT 4
#define RT_PRIMITIVE_TABLE_EXT(X) X(3)
#else
#define RT_PRIMITIVE_TYPE_COUNT 3
#define RT_PRIMITIVE_TABLE_EXT(X)
#endif
// the table: dispatch order = row order (triangles first: they are the bulk of every large scene); the last row is the
// chain's final else, so a scene of the three reference types pays no test for a type it cannot contain
#define RT_PRIMITIVE_TABLE(X, LAST) X(2) X(0) RT_PRIMITIVE_TABLE_EXT(X) LAST(1)

// ---- the three dispatchers every kernel goes through ----------------------------------------------------------------
RT_HD Aabb prim_bounds(const GeometryD &g, uint32_t type, uint32_t prim) {
#define RT_ROW(T) if (type == T##u) { return PrimOps<T>::bounds(g, prim); } else
#define RT_LAST(T) { return PrimOps<T>::bounds(g, prim); }
    RT_PRIMITIVE_TABLE(RT_ROW, RT_LAST)
#undef RT_ROW
#undef RT_LAST
}
// One leaf primitive against the LOCAL ray; acceptance rule of BLAS::hit's leaf loop (reference src/AS/BLAS.cu:150-175):
// in_range(t, tmin, best) is epsilon-closed, an accepted candidate becomes best.
template<bool WIDE>
RT_HD void prim_hit(const SceneD &sc, uint32_t type, uint32_t pos, const RayD &lray, float tmin, int inst, HitD &best) {
#define RT_ROW(T) if (type == T##u) { PrimOps<T>::template hit<WIDE>(sc, pos, lray, tmin, inst, best); } else
#define RT_LAST(T) { PrimOps<T>::template hit<WIDE>(sc, pos, lray, tmin, inst, best); }
    RT_PRIMITIVE_TABLE(RT_ROW, RT_LAST)
#undef RT_ROW
#undef RT_LAST
}
RT_HD void prim_surface(const SceneD &sc, uint32_t type, uint32_t prim, const RayD &lray, const f3 &lp, const HitD &hit, f3 &ln, bool &front, uint32_t &mt,
                        uint32_t &mi) {
#define RT_ROW(T) if (type == T##u) { PrimOps<T>::surface(sc, prim, lray, lp, hit, ln, front, mt, mi); } else
#define RT_LAST(T) { PrimOps<T>::surface(sc, prim, lray, lp, hit, ln, front, mt, mi); }
    RT_PRIMITIVE_TABLE(RT_ROW, RT_LAST)
#undef RT_ROW
#undef RT_LAST
}
