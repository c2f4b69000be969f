// rt_intersect.h — ray/primitive tests with the reference's exact acceptance semantics and operation order.
// Traversal uses the *_t variants (acceptance + t only); shading re-derives the full hit record.
#pragma once

#include "rt_types.h"

struct RayD { f3 o, d; };
RT_HD f3 ray_at(const RayD &r, float t) { return r.o + r.d * t; }   // Ray::at, reference include/Basic/Ray.cuh:18-20

// ---- triangle: Moller-Trumbore with absolute epsilon and three true divisions -----------------------
// reference src/Geometry/Triangle.cu:4-33 (quirk Q3).  Evaluated in instance-local space with the
// un-normalised local direction, so t is world-parametric.
RT_HD bool tri_hit_t(const f3 &v0, const f3 &e1, const f3 &e2, const RayD &ray, float tlo, float thi,
                     float &t, float &u, float &v) {
    const f3 h = cross3(ray.d, e2);
    const float det = dot3(e1, h);
    if (near_zero(det)) return false;
    const f3 s = ray.o - v0;
    u = dot3(s, h) / det;
    if (!in_range(u, 0.0f, 1.0f)) return false;
    const f3 q = cross3(s, e1);
    v = dot3(ray.d, q) / det;
    if (!in_range(v, 0.0f, 1.0f) || u + v > 1.0f) return false;
    t = dot3(e2, q) / det;
    return in_range(t, tlo, thi);
}

// ---- sphere: un-normalised quadratic, nearer root first (reference src/Geometry/Sphere.cu:4-27, Q4) ---
RT_HD bool sphere_hit_t(const f3 &c, float r, const RayD &ray, float tlo, float thi, float &t) {
    const f3 cq = c - ray.o;
    const float a = dot3(ray.d, ray.d);
    const float b = -2.0f * dot3(cq, ray.d);
    const float cc = dot3(cq, cq) - r * r;
    float delta = b * b - 4.0f * a * cc;
    if (delta < 0.0f) return false;
    delta = sqrtf(delta);
    const float root1 = (-b - delta) / (a * 2.0f);
    const float root2 = (-b + delta) / (a * 2.0f);
    if (in_range(root1, tlo, thi)) { t = root1; return true; }
    if (in_range(root2, tlo, thi)) { t = root2; return true; }
    return false;
}

// ---- reference slab test (reference src/AS/BoundingBox.cu:34-72, Q2): only used as the bug-compatible
//      filter of parallelograms (Q10); BVH culling uses the fast conservative test in rt_trace_core.h -----
RT_HD bool ref_box_hit(const f3 &lo, const f3 &hi, const RayD &ray, float curlo, float curhi, float &t_enter) {
    for (int axis = 0; axis < 3; axis++) {
        const float blo = comp3(lo, axis), bhi = comp3(hi, axis);
        const float q = comp3(ray.o, axis), d = comp3(ray.d, axis);
        if (fabsf(d) < RT_EPS) {
            if (q < blo || q > bhi) return false;
            continue;
        }
        const float t1 = (blo - q) / d;
        const float t2 = (bhi - q) / d;
        if (t1 < t2) {
            if (t1 > curlo) curlo = t1;
            if (t2 < curhi) curhi = t2;
        } else {
            if (t2 > curlo) curlo = t2;
            if (t1 < curhi) curhi = t1;
        }
        if (curlo >= curhi) return false;
    }
    t_enter = curlo;
    return true;
}

// ---- parallelogram (reference src/Geometry/Parallelogram.cu:4-35) -----------------------------------
// thi is the caller's current acceptance bound; on a hit, tbound is the bound to continue with.
RT_HD bool quad_hit_t(const QuadD &p, const RayD &ray, float tlo, float thi, float &t, float &tbound) {
    const f3 n = mk3(p.n_flag.x, p.n_flag.y, p.n_flag.z);
    float t_enter = -RT_INF;
    if (p.n_flag.w != 0.0f) {
        // bug-compatible mode (Q10): the reference only reaches the primitive when the ray passes the slab test of
        // Parallelogram::constructBoundingBox (src/Geometry/Parallelogram.cu:48-50) over [tlo, best-so-far], and it
        // visits boxes near-first (src/AS/TLAS.cu:186-197).  A hit X found BEFORE the wrong box is entered
        // (t_X < t_enter) therefore culls the parallelogram even when the parallelogram's own hit is nearer.
        // To make that independent of OUR traversal order, a parallelogram hit keeps the acceptance bound at
        // max(t, t_enter): a later candidate inside (t, t_enter] replaces it, exactly as if it had been found first.
        if (!ref_box_hit(mk3(p.rb_lo.x, p.rb_lo.y, p.rb_lo.z), mk3(p.rb_hi.x, p.rb_hi.y, p.rb_hi.z), ray, tlo, thi, t_enter)) return false;
    }
    const float nDotD = dot3(n, ray.d);
    if (near_zero(nDotD)) return false;
    const float nDotP = dot3(n, ray.o);
    t = (p.q_d.w - nDotP) / nDotD;
    if (!in_range(t, tlo, thi)) return false;
    const f3 x = ray_at(ray, t);
    const f3 q = mk3(p.q_d.x, p.q_d.y, p.q_d.z), u = mk3(p.u.x, p.u.y, p.u.z), v = mk3(p.v.x, p.v.y, p.v.z);
    const f3 pv = x - q;
    const f3 normal = cross3(u, v);
    const float denom = len2(normal);
    if (near_zero(denom)) return false;
    const float alpha = dot3(cross3(pv, v), normal) / denom;
    const float beta = dot3(cross3(u, pv), normal) / denom;
    if (!(in_range(alpha, 0.0f, 1.0f) && in_range(beta, 0.0f, 1.0f))) return false;
    tbound = t_enter > t ? t_enter : t;
    return true;
}
