// rt_trace_core.h — closest-hit traversal of the two-level BVH and the reference's material/path logic.
// Pure per-ray / per-path functions (no warp intrinsics) so the identical code runs inside the CUDA
// kernels and, compiled by g++, inside the CPU emulation used by the `not gpu` tests.
#pragma once

#include "rt_prims.h"
#include "rt_rng.h"

struct RaySlabs { f3 idir, ood; };

// record fetches of one path/pixel, for the roofline's algorithmic-bytes figure (instrumented variant only)
struct TravStats { uint32_t nodes, prims, insts; };

RT_HD float safe_rcp(float d) {
    // BVH culling only (not parity-critical): keep 1/d finite so 0 * inf never appears in the slab test
    const float a = fabsf(d) < 1e-20f ? (d < 0.0f ? -1e-20f : 1e-20f) : d;
    return 1.0f / a;
}
RT_HD RaySlabs make_slabs(const RayD &r) {
    RaySlabs s;
    s.idir = mk3(safe_rcp(r.d.x), safe_rcp(r.d.y), safe_rcp(r.d.z));
    s.ood = mk3(-(r.o.x * s.idir.x), -(r.o.y * s.idir.y), -(r.o.z * s.idir.z));
    return s;
}

// Both children of a binary node against the ray; returns entry distances, hit iff tn <= tf.
RT_HD void node2_test(const BvhNode2 &n, const RaySlabs &s, float tmin, float tbest, float &tn0, float &tf0, float &tn1, float &tf1) {
    const float a0 = rt_fma(n.xy0.x, s.idir.x, s.ood.x), b0 = rt_fma(n.xy0.y, s.idir.x, s.ood.x);
    const float c0 = rt_fma(n.xy0.z, s.idir.y, s.ood.y), d0 = rt_fma(n.xy0.w, s.idir.y, s.ood.y);
    const float e0 = rt_fma(n.z01.x, s.idir.z, s.ood.z), f0 = rt_fma(n.z01.y, s.idir.z, s.ood.z);
    tn0 = rt_max(rt_max(rt_min(a0, b0), rt_min(c0, d0)), rt_max(rt_min(e0, f0), tmin));
    tf0 = rt_min(rt_min(rt_max(a0, b0), rt_max(c0, d0)), rt_min(rt_max(e0, f0), tbest));
    const float a1 = rt_fma(n.xy1.x, s.idir.x, s.ood.x), b1 = rt_fma(n.xy1.y, s.idir.x, s.ood.x);
    const float c1 = rt_fma(n.xy1.z, s.idir.y, s.ood.y), d1 = rt_fma(n.xy1.w, s.idir.y, s.ood.y);
    const float e1 = rt_fma(n.z01.z, s.idir.z, s.ood.z), f1 = rt_fma(n.z01.w, s.idir.z, s.ood.z);
    tn1 = rt_max(rt_max(rt_min(a1, b1), rt_min(c1, d1)), rt_max(rt_min(e1, f1), tmin));
    tf1 = rt_min(rt_min(rt_max(a1, b1), rt_max(c1, d1)), rt_min(rt_max(e1, f1), tbest));
}

RT_HD BvhNode2 load_node2(const BvhNode2 *nodes, int idx) {
    BvhNode2 n;
#if defined(__CUDA_ARCH__)
    const float4 *p = reinterpret_cast<const float4 *>(nodes + idx);
    n.xy0 = __ldg(p); n.xy1 = __ldg(p + 1); n.z01 = __ldg(p + 2);
    n.refs = __ldg(reinterpret_cast<const int4 *>(p + 3));
#else
    n = nodes[idx];
#endif
    return n;
}

// One leaf primitive (binary-LBVH leaf order) against the LOCAL ray: dispatched through the primitive table (rt_prims.h)
RT_HD void leaf_test(const SceneD &sc, int ref, const RayD &lray, float tmin, int inst, HitD &best) {
    prim_hit<false>(sc, ref_type(ref), ref_pos(ref), lray, tmin, inst, best);
}

RT_HD RayD to_local(const InstanceD &ins, const RayD &w) {
    // Instance::hit, reference src/AS/Instance.cu:26-28: origin as point, direction as vector, NOT re-normalised
    RayD l;
    l.o = xf_point(ins.inv, w.o);
    l.d = xf_vector(ins.inv, w.d);
    return l;
}

RT_HD void load_instance_inv(const SceneD &sc, int inst, float *inv, int &root) {
#if defined(__CUDA_ARCH__)
    const float4 *p = reinterpret_cast<const float4 *>(sc.instances + inst);
    const float4 r0 = __ldg(p), r1 = __ldg(p + 1), r2 = __ldg(p + 2);
    inv[0] = r0.x; inv[1] = r0.y; inv[2] = r0.z; inv[3] = r0.w;
    inv[4] = r1.x; inv[5] = r1.y; inv[6] = r1.z; inv[7] = r1.w;
    inv[8] = r2.x; inv[9] = r2.y; inv[10] = r2.z; inv[11] = r2.w;
    root = __ldg(&sc.instances[inst].root_ref);
#else
    for (int i = 0; i < 12; i++) inv[i] = sc.instances[inst].inv[i];
    root = sc.instances[inst].root_ref;
#endif
}

// Closest hit of a world ray over [tmin, +inf): the job of TLAS::hit -> Instance::hit -> BLAS::hit
// (reference src/AS/TLAS.cu:131-201, src/AS/Instance.cu:19-50, src/AS/BLAS.cu:119-206) on our own BVH.
template<bool COUNT>
RT_HD void trace_closest(const SceneD &sc, const RayD &wray, float tmin, HitD &hit, uint32_t &overflow, TravStats &stats) {
    hit.t = RT_INF; hit.tc = RT_INF; hit.u = 0.0f; hit.v = 0.0f; hit.inst = -1; hit.leaf = 0;
    int stack[RT_STACK_SIZE];
    int sp = 0;
    int cur = sc.tlas_root_ref;
    int inst = -1;
    RayD ray = wray;
    RaySlabs slabs = make_slabs(ray);
    for (;;) {
        if (cur >= 0) {
            const BvhNode2 n = load_node2(inst < 0 ? sc.tlas_nodes : sc.blas_nodes, cur);
            if (COUNT) stats.nodes++;
            float tn0, tf0, tn1, tf1;
            node2_test(n, slabs, tmin, hit.tc, tn0, tf0, tn1, tf1);
            const bool h0 = tn0 <= tf0, h1 = tn1 <= tf1;
            if (h0 && h1) {
                const bool swap = tn1 < tn0;
                const int nearRef = swap ? n.refs.y : n.refs.x;
                const int farRef = swap ? n.refs.x : n.refs.y;
                if (sp < RT_STACK_SIZE) stack[sp++] = farRef; else overflow = 1u;
                cur = nearRef;
                continue;
            }
            if (h0) { cur = n.refs.x; continue; }
            if (h1) { cur = n.refs.y; continue; }
        } else if (cur == RT_REF_SENTINEL) {
            inst = -1;
            ray = wray;
            slabs = make_slabs(ray);
        } else if (inst < 0) {
            // TLAS leaf: enter the instance
            inst = (int) ref_pos(cur);
            if (COUNT) stats.insts++;
            float inv[12];
            int root;
            load_instance_inv(sc, inst, inv, root);
            ray.o = xf_point(inv, wray.o);
            ray.d = xf_vector(inv, wray.d);
            slabs = make_slabs(ray);
            if (sp < RT_STACK_SIZE) stack[sp++] = RT_REF_SENTINEL; else overflow = 1u;
            cur = root;
            continue;
        } else {
            if (COUNT) stats.prims++;
            leaf_test(sc, cur, ray, tmin, inst, hit);
        }
        if (sp == 0) break;
        cur = stack[--sp];
    }
}

// ---- camera ray: reference src/Global/Kernel.cu:121-135 ------------------------------------------------
RT_HD RayD primary_ray(const CameraD &cam, uint32_t x, uint32_t y, uint32_t si, uint32_t sj, RngState &rng) {
    const float ox = (((float) sj + rng_uniform(rng)) * cam.rcpSqrtSamples) - 0.5f;
    const float oy = (((float) si + rng_uniform(rng)) * cam.rcpSqrtSamples) - 0.5f;
    const f3 sample = (cam.pixelOrigin + cam.pixelDx * ((float) x + ox)) + cam.pixelDy * ((float) y + oy);
    f3 origin = cam.center;
    if (cam.focusRadius > 0.0f) {
        float dx, dy;
        rng_plane_vector(rng, cam.focusRadius, dx, dy);
        origin = (cam.center + cam.camU * dx) + cam.camV * dy;
    }
    RayD r;
    r.o = origin;
    r.d = unit3(sample - origin);
    return r;
}

// Hit record + scatter: Prim::hit's record fields (reference src/Geometry/*.cu), Instance::hit's back-transform
// (src/AS/Instance.cu:41-45), Rough/Metal::scatter (include/Material/*.cuh) and rayColor's bookkeeping
// (src/Global/Kernel.cu:76-98).  Returns false when the path ends; `throughput` is rayColor's `result`.
template<bool WIDE>
RT_HD bool shade_scatter(const SceneD &sc, const CameraD &cam, RayD &ray, const HitD &hit, f3 &throughput, RngState &rng, uint32_t inst_base = 0u) {
    if (hit.inst < 0) {
        throughput.x *= cam.background.x; throughput.y *= cam.background.y; throughput.z *= cam.background.z;
        return false;
    }
    const InstanceD &ins = sc.instances[inst_base + (uint32_t) hit.inst];   // inst_base: this frame's copy of the records (frames in flight)
    const RayD lray = to_local(ins, ray);
    const uint32_t type = ref_type(hit.leaf), pos = ref_pos(hit.leaf);
    const uint32_t prim = WIDE ? sc.leaf_prims8[pos] : sc.leaf_prims[pos];
    const f3 lp = ray_at(lray, hit.t);
    f3 ln;
    uint32_t mt, mi;
    bool front;                 // Prim::hit's hitFrontFace: the outward normal faces the (local) ray
    prim_surface(sc, type, prim, lray, lp, hit, ln, front, mt, mi);       // the primitive table (rt_prims.h)
    const f3 wp = xf_point(ins.fwd, lp);
    const f3 wn = unit3(xf_normal(ins.inv, ln));
    // Rough::scatter and Metal::scatter (fuzz > 0) both draw ONE random unit vector; it is drawn at a single place so
    // that the lanes of a warp that shade different materials run the rejection loop together (the stream position
    // and every value are the same as drawing it inside each branch: nothing else consumes the stream in between).
    const bool rough = mt == 0u, metal = mt == 1u, glass = mt == 2u;
    float4 m = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    if (rough) m = sc.roughs[mi]; else if (metal) m = sc.metals[mi]; else if (glass) m = sc.dielectrics[mi];
    f3 rv = mk3(0.0f, 0.0f, 0.0f);
    if (rough || (metal && m.w > 0.0f)) rv = rng_space_vector(rng);
    if (rough) {                                                     // Rough::scatter
        f3 dir = wn + rv;
        if (feq(len2(dir), RT_EPS * RT_EPS)) dir = wn;
        ray.o = wp; ray.d = dir;                                     // not normalised (Q5)
        throughput.x *= m.x; throughput.y *= m.y; throughput.z *= m.z;
        return true;
    }
    if (metal) {                                                     // Metal::scatter
        const f3 v = ray.d;
        f3 refl = unit3(v - wn * (2 * dot3(v, wn)));
        if (m.w > 0.0f) refl = refl + rv * m.w;
        ray.o = wp; ray.d = refl;
        if (!(dot3(refl, wn) > 0.0f)) return false;                  // rayColor returns before attenuating (Q6)
        throughput.x *= m.x; throughput.y *= m.y; throughput.z *= m.z;
        return true;
    }
    if (glass) {
        // EXTENSION — PARITY UNPINNED: the reference has no dielectric (include/Basic/BasicTypes.cuh:14-16).  Snell refraction
        // with Schlick reflectance in Metal::scatter's conventions: incoming direction normalised first (Q5 leaves ray
        // directions un-normalised), `wn` already faces the ray, the side of the surface is Prim::hit's hitFrontFace, one
        // uniform (always drawn) decides reflect / refract, outgoing direction not re-normalised.  m.w = index of refraction.
        const f3 v = unit3(ray.d);
        const float ri = front ? 1.0f / m.w : m.w;
        float cosT = -dot3(v, wn);
        if (cosT > 1.0f) cosT = 1.0f;
        const float sin2 = 1.0f - cosT * cosT;
        const float sinT = sqrtf(sin2 > 0.0f ? sin2 : 0.0f);
        float r0 = (1.0f - ri) / (1.0f + ri);
        r0 = r0 * r0;
        const float c1 = 1.0f - cosT;
        const float c5 = ((c1 * c1) * (c1 * c1)) * c1;
        const float reflectance = r0 + (1.0f - r0) * c5;
        const float xi = rng_uniform(rng);
        f3 dir;
        if (ri * sinT > 1.0f || reflectance > xi) {
            dir = v - wn * (2 * dot3(v, wn));
        } else {
            const f3 perp = (v + wn * cosT) * ri;
            const float k = 1.0f - len2(perp);
            dir = perp + wn * (-sqrtf(fabsf(k)));
        }
        ray.o = wp; ray.d = dir;
        throughput.x *= m.x; throughput.y *= m.y; throughput.z *= m.z;
        return true;
    }
    // MaterialType outside {ROUGH, METAL, DIELECTRIC} is rejected at commit time (host_scene.cpp); unreachable.
    return false;
}

// Color3::castToUchar4 (reference include/Basic/Color3.cuh:99-114): gamma 2, clamp [0, 0.999], x256, truncate
RT_HD uint8_t to_byte(float c) {
    float g = sqrtf(c);
    if (g > 0.999f) g = 0.999f; else if (g < 0.0f) g = 0.0f;
    return (uint8_t) (256.0f * g);
}

