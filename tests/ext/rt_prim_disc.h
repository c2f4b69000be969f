// tests/ext/rt_prim_disc.h — TEST-ONLY fourth primitive type: a disc (centre, unit normal, radius).
//
// Proof that a geometry type is ONE table row + ONE PrimOps specialisation (csrc/rt_prims.h; the reference needs ten steps in
// ten files, files/AddGeometryType.txt:1-35).  Compiled only into the test variants (librtb200_disc.so, libemul_disc.so:
// -DRT_EXT_PRIMITIVE='"../../tests/ext/rt_prim_disc.h"'); the product build does not know it.  Checked against the oracle
// port's own, independently written disc (oracle_port.cpp, ORACLE_EXT_DISC) in tests/test_ext_primitive.py.
// Conventions follow the reference's Parallelogram::hit (src/Geometry/Parallelogram.cu:4-46): plane hit by
// t = (d - n.o) / n.dir with the absolute-epsilon parallel test, epsilon-closed range, then the inside test.
#pragma once

struct alignas(16) ExtPrimD {
    float4 c_r;       // centre.xyz, radius
    float4 n_d;       // unit normal.xyz, plane constant d = n . centre
    uint32_t mat_type, mat_index, _p0, _p1;
};
#define RT_EXT_PRIMITIVE_NAME "disc"

template<> struct PrimOps<3> {
    RT_HD static const ExtPrimD *records(const void *p) { return static_cast<const ExtPrimD *>(p); }
    // a disc's extent along axis a is r * sqrt(1 - n_a^2)
    RT_HD static void extent(const ExtPrimD &q, f3 &lo, f3 &hi) {
        const float ex = q.c_r.w * sqrtf(rt_max(0.0f, 1.0f - q.n_d.x * q.n_d.x));
        const float ey = q.c_r.w * sqrtf(rt_max(0.0f, 1.0f - q.n_d.y * q.n_d.y));
        const float ez = q.c_r.w * sqrtf(rt_max(0.0f, 1.0f - q.n_d.z * q.n_d.z));
        lo = mk3(q.c_r.x - ex, q.c_r.y - ey, q.c_r.z - ez);
        hi = mk3(q.c_r.x + ex, q.c_r.y + ey, q.c_r.z + ez);
    }
    RT_HD static Aabb bounds(const GeometryD &g, uint32_t prim) {
        Aabb b;
        extent(records(g.ext)[prim], b.lo, b.hi);
        return b;
    }
    template<bool WIDE> RT_HD static void hit(const SceneD &sc, uint32_t pos, const RayD &lray, float tmin, int inst, HitD &best) {
        const ExtPrimD q = records(sc.ext_prims)[WIDE ? sc.leaf_prims8[pos] : sc.leaf_prims[pos]];
        const f3 n = mk3(q.n_d.x, q.n_d.y, q.n_d.z);
        const float nDotD = dot3(n, lray.d);
        if (near_zero(nDotD)) return;
        const float nDotP = dot3(n, lray.o);
        const float t = (q.n_d.w - nDotP) / nDotD;
        if (!in_range(t, tmin, best.tc)) return;
        const f3 pv = ray_at(lray, t) - mk3(q.c_r.x, q.c_r.y, q.c_r.z);
        if (len2(pv) > q.c_r.w * q.c_r.w) return;
        best.t = t; best.tc = t; best.u = 0.0f; best.v = 0.0f; best.inst = inst; best.leaf = make_leaf_ref(3u, pos);
    }
    RT_HD static void surface(const SceneD &sc, uint32_t prim, const RayD &lray, const f3 &, const HitD &, f3 &ln, bool &front, uint32_t &mt, uint32_t &mi) {
        const ExtPrimD q = records(sc.ext_prims)[prim];
        const f3 n = mk3(q.n_d.x, q.n_d.y, q.n_d.z);
        front = dot3(lray.d, n) < 0.0f;
        ln = front ? n : -n;
        mt = q.mat_type; mi = q.mat_index;
    }
    // host side: instance-local box (before BoundingBox's ensureVolume) and centroid, the material for the commit-time checks
    static void ref_extent(const ExtPrimD &q, f3 &lo, f3 &hi) { extent(q, lo, hi); }
    static f3 centroid(const ExtPrimD &q) { return mk3(q.c_r.x, q.c_r.y, q.c_r.z); }
    static void material(const ExtPrimD &q, uint32_t &mt, uint32_t &mi) { mt = q.mat_type; mi = q.mat_index; }
};
