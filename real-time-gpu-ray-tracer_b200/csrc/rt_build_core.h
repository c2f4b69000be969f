// rt_build_core.h — per-element logic of the LBVH build (Morton codes, Karras 2012 hierarchy, refit, node
// emission) and of the per-instance transform update.  The CUDA kernels in kernels_build.cu call these once
// per thread; the CPU twin used by the `not gpu` tests calls them in serial loops, so device topology can be
// compared bit-for-bit against a serial execution of the same definition.
#pragma once

#include "rt_types.h"

// ---- boxes ------------------------------------------------------------------------------------------
struct Aabb { f3 lo, hi; };
RT_HD Aabb aabb_union(const Aabb &a, const Aabb &b) {
    Aabb r;
    r.lo = mk3(rt_min(a.lo.x, b.lo.x), rt_min(a.lo.y, b.lo.y), rt_min(a.lo.z, b.lo.z));
    r.hi = mk3(rt_max(a.hi.x, b.hi.x), rt_max(a.hi.y, b.hi.y), rt_max(a.hi.z, b.hi.z));
    return r;
}
RT_HD float aabb_half_area(const Aabb &b) {
    const float dx = b.hi.x - b.lo.x, dy = b.hi.y - b.lo.y, dz = b.hi.z - b.lo.z;
    return dx * dy + dy * dz + dz * dx;
}

// ---- SAH refinement of the binary LBVH: one tree rotation per node, applied bottom-up inside the refit ---------
// (Kensler 2008 / Kopta et al. 2012.)  At node N = (L, R) with finished subtrees, a grandchild may trade places with
// its uncle: N = ((LL, LR), R) -> ((R, LR), LL) etc.  N's box and leaf set stay, only the box of the inner child
// changes, and with it the surface-area cost of the tree.  The rotation with the largest area reduction is taken
// (first candidate wins ties).  The decision depends only on N's subtree, which is final when the refit reaches N,
// so the result is the same for any bottom-up order — the atomic GPU refit and the serial twin agree bit for bit.
// Returns 0: none, 1: R <-> LL, 2: R <-> LR, 3: L <-> RL, 4: L <-> RR; new_inner = the changed child's new box.
RT_HD int best_rotation(const Aabb &L, const Aabb &R, bool l_inner, const Aabb &LL, const Aabb &LR,
                        bool r_inner, const Aabb &RL, const Aabb &RR, Aabb &new_inner) {
    int best = 0;
    float gain = 0.0f;
    if (l_inner) {
        const float cur = aabb_half_area(L);
        const Aabb a = aabb_union(R, LR), b = aabb_union(R, LL);
        const float ga = cur - aabb_half_area(a), gb = cur - aabb_half_area(b);
        if (ga > gain) { gain = ga; best = 1; new_inner = a; }
        if (gb > gain) { gain = gb; best = 2; new_inner = b; }
    }
    if (r_inner) {
        const float cur = aabb_half_area(R);
        const Aabb a = aabb_union(L, RR), b = aabb_union(L, RL);
        const float ga = cur - aabb_half_area(a), gb = cur - aabb_half_area(b);
        if (ga > gain) { gain = ga; best = 3; new_inner = a; }
        if (gb > gain) { gain = gb; best = 4; new_inner = b; }
    }
    return best;
}

// ---- PLOC: parallel locally-ordered clustering (Meister & Bittner 2018) — the builder for BLASes that are built once ---
// Bottom-up agglomeration over the Morton-sorted primitives: in every round each cluster looks at the clusters within
// RT_PLOC_RADIUS positions of it in the (Morton-ordered) cluster array and picks the one whose union with it has the
// smallest surface area; two clusters that pick each other merge into a new inner node, the rest wait.  The merged array
// keeps the Morton order, so the next round sees ever larger neighbourhoods.  The result approaches a top-down SAH build
// while every step stays a flat data-parallel pass.  Determinism: the distance is symmetric, ties go to the lower
// position, node ids come from a prefix sum over the merge leaders — the device rounds and the serial twin give the
// same tree bit for bit.  Clusters of different segments (BLASes of one batch) never merge.
// A cluster is two float4: lo.xyz | bits(ref) and hi.xyz | bits(segment); ref >= 0 inner node, ref < 0 leaf ~position.
#ifndef RT_PLOC_RADIUS
#define RT_PLOC_RADIUS 16
#endif
RT_HD int ploc_nearest(const float4 *clo, const float4 *chi, int m, int i, int radius) {
    const float4 a = clo[i], b = chi[i];
    int best = -1;
    float best_area = RT_INF;
    const int j0 = i - radius < 0 ? 0 : i - radius, j1 = i + radius > m - 1 ? m - 1 : i + radius;
    for (int j = j0; j <= j1; j++) {
        if (j == i) continue;
        const float4 c = clo[j], d = chi[j];
        if (f2u(d.w) != f2u(b.w)) continue;                      // another BLAS of the batch
        const float dx = rt_max(b.x, d.x) - rt_min(a.x, c.x), dy = rt_max(b.y, d.y) - rt_min(a.y, c.y), dz = rt_max(b.z, d.z) - rt_min(a.z, c.z);
        const float area = dx * dy + dy * dz + dz * dx;
        if (area < best_area) { best_area = area; best = j; }    // strict: the lower position wins ties
    }
    return best;
}
// what cluster i does in this round: 0 stays, 1 merges with nn[i] and leads (i < nn[i]), 2 is absorbed by nn[i]
RT_HD int ploc_role(const int *nn, int i) {
    const int j = nn[i];
    if (j < 0 || nn[j] != i) return 0;
    return i < j ? 1 : 2;
}

// Conservative padding so that the fma-based slab test can never cull a true hit on the box surface.
// The reference pads only degenerate axes by 1e-6 (ensureVolume, include/AS/BoundingBox.cuh:24-28).
RT_HD Aabb aabb_pad(const Aabb &a) {
    Aabb r;
    r.lo = mk3(a.lo.x - (fabsf(a.lo.x) * 4e-7f + 1e-6f), a.lo.y - (fabsf(a.lo.y) * 4e-7f + 1e-6f), a.lo.z - (fabsf(a.lo.z) * 4e-7f + 1e-6f));
    r.hi = mk3(a.hi.x + (fabsf(a.hi.x) * 4e-7f + 1e-6f), a.hi.y + (fabsf(a.hi.y) * 4e-7f + 1e-6f), a.hi.z + (fabsf(a.hi.z) * 4e-7f + 1e-6f));
    return r;
}
// A triangle with v1 == v0 or v2 == v0 has e1 = 0 or e2 = 0 exactly, so Moller-Trumbore's det is exactly 0 for every
// ray and tri_hit_t rejects it (near_zero): it can never be hit, here or in the reference.  The triangle strips of the
// particle meshes are 41 % such triangles (they keep their primitive index, so they stay in the BVH); their box is
// shrunk to the point v0 so that rays stop paying node visits and triangle tests for them.
// (v1 == v2 gives e1 == e2 and a det that is only NEARLY zero — those keep their full box.)
RT_HD Aabb tri_aabb(const float *v9) {
    Aabb r;
    if ((v9[3] == v9[0] && v9[4] == v9[1] && v9[5] == v9[2]) || (v9[6] == v9[0] && v9[7] == v9[1] && v9[8] == v9[2])) {
        r.lo = mk3(v9[0], v9[1], v9[2]);
        r.hi = r.lo;
        return r;
    }
    r.lo = mk3(rt_min(rt_min(v9[0], v9[3]), v9[6]), rt_min(rt_min(v9[1], v9[4]), v9[7]), rt_min(rt_min(v9[2], v9[5]), v9[8]));
    r.hi = mk3(rt_max(rt_max(v9[0], v9[3]), v9[6]), rt_max(rt_max(v9[1], v9[4]), v9[7]), rt_max(rt_max(v9[2], v9[5]), v9[8]));
    return r;
}
RT_HD Aabb sphere_aabb(const SphereD &s) {
    Aabb r;
    r.lo = mk3(s.c_r.x - s.c_r.w, s.c_r.y - s.c_r.w, s.c_r.z - s.c_r.w);
    r.hi = mk3(s.c_r.x + s.c_r.w, s.c_r.y + s.c_r.w, s.c_r.z + s.c_r.w);
    return r;
}
RT_HD Aabb quad_aabb(const QuadD &p) {   // TRUE box of the four corners (the reference's box is a filter, Q10)
    const f3 q = mk3(p.q_d.x, p.q_d.y, p.q_d.z), u = mk3(p.u.x, p.u.y, p.u.z), v = mk3(p.v.x, p.v.y, p.v.z);
    const f3 c1 = q + u, c2 = q + v, c3 = (q + u) + v;
    Aabb r;
    r.lo = mk3(rt_min(rt_min(q.x, c1.x), rt_min(c2.x, c3.x)), rt_min(rt_min(q.y, c1.y), rt_min(c2.y, c3.y)), rt_min(rt_min(q.z, c1.z), rt_min(c2.z, c3.z)));
    r.hi = mk3(rt_max(rt_max(q.x, c1.x), rt_max(c2.x, c3.x)), rt_max(rt_max(q.y, c1.y), rt_max(c2.y, c3.y)), rt_max(rt_max(q.z, c1.z), rt_max(c2.z, c3.z)));
    return r;
}

// order-preserving float <-> uint mapping for atomicMin/atomicMax on bounds
RT_HD uint32_t float_to_ordered(float f) { const uint32_t u = f2u(f); return (u & 0x80000000u) ? ~u : (u | 0x80000000u); }
RT_HD float ordered_to_float(uint32_t o) { return u2f((o & 0x80000000u) ? (o & 0x7FFFFFFFu) : ~o); }

// ---- Morton keys ---------------------------------------------------------------------------------------
// 3*bits-bit Morton code of the box centre, normalised to the segment (BLAS) bounds; bits <= 21.
RT_HD uint64_t expand_bits21(uint64_t v) {
    v &= 0x1FFFFFull;
    v = (v | (v << 32)) & 0x1F00000000FFFFull;
    v = (v | (v << 16)) & 0x1F0000FF0000FFull;
    v = (v | (v << 8)) & 0x100F00F00F00F00Full;
    v = (v | (v << 4)) & 0x10C30C30C30C30C3ull;
    v = (v | (v << 2)) & 0x1249249249249249ull;
    return v;
}
// The grid is CUBIC: every axis is scaled by the largest extent of the segment, so a split always halves a cell along
// its longest side.  (Scaling each axis by its own extent makes a flat mesh — a height field, a particle shell —
// spend every third split on its thin axis: children that overlap completely in the other two.)  Along a short axis
// the top bits are simply equal for all primitives and the Karras hierarchy skips them.
#ifndef RT_MORTON_PER_AXIS
#define RT_MORTON_PER_AXIS 0     /* 1: the first version's per-axis scaling, kept for A/B measurements */
#endif
// The box centre morton_code quantises, as a degenerate box.  The TLAS takes its Morton grid from the union of the
// instances' CENTRES, not of their boxes: one huge instance (the demo scene's r = 1000 sky sphere around particles a
// few units across) would otherwise put every other instance into the same few cells and leave their order — and
// the TLAS topology — to the index tie-break.
RT_HD Aabb aabb_centre_box(const Aabb &b) {
    Aabb c;
    c.lo = mk3(0.5f * (b.lo.x + b.hi.x), 0.5f * (b.lo.y + b.hi.y), 0.5f * (b.lo.z + b.hi.z));
    c.hi = c.lo;
    return c;
}
RT_HD uint64_t morton_code(const Aabb &prim, const Aabb &scene, int bits) {
    const float scale = (float) (1u << bits);
    const float ext_max = rt_max(rt_max(scene.hi.x - scene.lo.x, scene.hi.y - scene.lo.y), scene.hi.z - scene.lo.z);
    uint64_t q[3];
    for (int a = 0; a < 3; a++) {
        const float lo = comp3(scene.lo, a), hi = comp3(scene.hi, a);
        const float c = 0.5f * (comp3(prim.lo, a) + comp3(prim.hi, a));
        const float ext = RT_MORTON_PER_AXIS ? hi - lo : ext_max;
        (void) hi;
        float n = ext > 0.0f ? (c - lo) / ext : 0.0f;
        n = rt_min(rt_max(n * scale, 0.0f), scale - 1.0f);
        q[a] = (uint64_t) n;
    }
    const uint64_t code21 = (expand_bits21(q[0]) << 2) | (expand_bits21(q[1]) << 1) | expand_bits21(q[2]);
    // q has `bits` significant bits per axis -> interleaved code has 3*bits significant bits
    return code21;
}

// ---- Karras 2012: internal node i of a radix tree over n sorted keys (ties broken by position) ------------
RT_HD int clz64(uint64_t x) {
#if defined(__CUDA_ARCH__)
    return __clzll((long long) x);
#else
    return x == 0 ? 64 : __builtin_clzll(x);
#endif
}
RT_HD int clz32(uint32_t x) {
#if defined(__CUDA_ARCH__)
    return __clz((int) x);
#else
    return x == 0 ? 32 : __builtin_clz(x);
#endif
}
RT_HD int karras_delta(const uint64_t *keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    const uint64_t a = keys[i], b = keys[j];
    if (a == b) return 64 + clz32((uint32_t) i ^ (uint32_t) j);
    return clz64(a ^ b);
}
struct KarrasNode { int left, right, first, last; };   // children: >= 0 inner node, < 0 leaf ~position
RT_HD KarrasNode karras_node(const uint64_t *keys, int n, int i) {
    const int d = (karras_delta(keys, n, i, i + 1) - karras_delta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
    const int dmin = karras_delta(keys, n, i, i - d);
    int lmax = 2;
    while (karras_delta(keys, n, i, i + lmax * d) > dmin) lmax <<= 1;
    int l = 0;
    for (int t = lmax >> 1; t >= 1; t >>= 1)
        if (karras_delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
    const int j = i + l * d;
    const int dnode = karras_delta(keys, n, i, j);
    int s = 0;
    for (int div = 2, t = (l + 1) / 2;; div <<= 1, t = (l + div - 1) / div) {
        if (karras_delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
        if (t <= 1) break;
    }
    const int gamma = i + s * d + (d < 0 ? -1 : 0);
    KarrasNode k;
    k.first = i < j ? i : j;
    k.last = i < j ? j : i;
    k.left = (k.first == gamma) ? ~gamma : gamma;
    k.right = (k.last == gamma + 1) ? ~(gamma + 1) : gamma + 1;
    return k;
}

// ---- node emission ----------------------------------------------------------------------------------------
RT_HD BvhNode2 make_node2(const Aabb &b0, int ref0, const Aabb &b1, int ref1) {
    BvhNode2 n;
    n.xy0 = make_float4(b0.lo.x, b0.hi.x, b0.lo.y, b0.hi.y);
    n.xy1 = make_float4(b1.lo.x, b1.hi.x, b1.lo.y, b1.hi.y);
    n.z01 = make_float4(b0.lo.z, b0.hi.z, b1.lo.z, b1.hi.z);
    n.refs = make_int4(ref0, ref1, 0, 0);
    return n;
}

// ---- instance transform update ---------------------------------------------------------------------------
// Instance::updateTransformArguments (reference src/AS/Instance.cu:4-17) with the reference's generic
// Matrix arithmetic restated on plain 4x4 arrays: M = T * (Rx * Ry * Rz) * S  (src/Util/Matrix.cu:71-86,
// 183-250), inverse by Gauss-Jordan with partial pivoting on a 4x8 scratch (Matrix.cu:4-69,101-130; singular
// input is returned unchanged, Q14), world box from the 8 corners (src/AS/BoundingBox.cu:4-32) and the
// centroid.  sin/cos of the three angles are evaluated on the host (glibc) and passed in `trig`.
struct InstanceXformIn {
    float shift[3], scale[3];
    float trig[6];          // cos x, sin x, cos y, sin y, cos z, sin z
    float box_lo[3], box_hi[3], centroid[3];
    int root_ref, blas;
    int root8, prim_type;
    int tri_first, tri_count;   // triangle BLAS whose world box is tightened from its vertices (tri_count = 0: keep the 8-corner box)
};
#define RT_TIGHT_BOUNDS_MAX_TRIS 4096   /* larger meshes keep the 8-corner box (one warp walks the vertices every frame) */
RT_HD void mat4_mul(const float a[4][4], const float b[4][4], float r[4][4]) {
    for (int i = 0; i < 4; i++)
        for (int j = 0; j < 4; j++) {
            float sum = 0.0f;
            for (int k = 0; k < 4; k++) sum += a[i][k] * b[k][j];
            r[i][j] = sum;
        }
}
RT_HD void mat4_identity(float m[4][4]) {
    for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) m[i][j] = i == j ? 1.0f : 0.0f;
}
RT_HD bool mat4_inverse(const float a[4][4], float out[4][4]) {
    float w[4][8];
    for (int i = 0; i < 4; i++) for (int j = 0; j < 8; j++) w[i][j] = j < 4 ? a[i][j] : (j - 4 == i ? 1.0f : 0.0f);
    for (int i = 0; i < 4; i++) {
        float pivot = fabsf(w[i][i]);
        int maxRow = i;
        for (int p = i + 1; p < 4; p++) if (fabsf(w[p][i]) > pivot) { pivot = fabsf(w[p][i]); maxRow = p; }
        if (near_zero(pivot)) return false;
        if (maxRow != i) for (int j = 0; j < 8; j++) { const float t = w[maxRow][j]; w[maxRow][j] = w[i][j]; w[i][j] = t; }
        for (int j = i + 1; j < 4; j++) {
            const float factor = w[j][i] / w[i][i];
            for (int k = i; k < 8; k++) w[j][k] -= factor * w[i][k];
        }
    }
    for (int i = 3; i >= 0; i--) {
        if (near_zero(w[i][i])) return false;
        float factor = 1.0f / w[i][i];
        for (int p = i; p < 8; p++) w[i][p] *= factor;
        for (int j = i - 1; j >= 0; j--) {
            factor = w[j][i];
            for (int k = j; k < 8; k++) w[j][k] -= factor * w[i][k];
        }
    }
    for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) out[i][j] = w[i][4 + j];
    return true;
}
// Range::length / BoundingBox::ensureVolume on one axis (reference include/Util/Range.cuh:59-65, BoundingBox.cuh:24-28)
RT_HD void ensure_volume(float &lo, float &hi) {
    const float len = (lo >= hi || feq(lo, hi)) ? 0.0f : hi - lo;
    if (len < RT_EPS) { lo -= RT_EPS; hi += RT_EPS; }
}
RT_HD void instance_update(const InstanceXformIn &in, InstanceD &out, Aabb &wbox, f3 &wcentroid) {
    float T[4][4], S[4][4], Rx[4][4], Ry[4][4], Rz[4][4], Rxy[4][4], R[4][4], TR[4][4], M[4][4], Minv[4][4];
    mat4_identity(T); T[0][3] = in.shift[0]; T[1][3] = in.shift[1]; T[2][3] = in.shift[2];
    mat4_identity(S); S[0][0] = in.scale[0]; S[1][1] = in.scale[1]; S[2][2] = in.scale[2];
    mat4_identity(Rx); Rx[1][1] = in.trig[0]; Rx[1][2] = -in.trig[1]; Rx[2][1] = in.trig[1]; Rx[2][2] = in.trig[0];
    mat4_identity(Ry); Ry[0][0] = in.trig[2]; Ry[0][2] = in.trig[3]; Ry[2][0] = -in.trig[3]; Ry[2][2] = in.trig[2];
    mat4_identity(Rz); Rz[0][0] = in.trig[4]; Rz[0][1] = -in.trig[5]; Rz[1][0] = in.trig[5]; Rz[1][1] = in.trig[4];
    mat4_mul(Rx, Ry, Rxy);
    mat4_mul(Rxy, Rz, R);
    mat4_mul(T, R, TR);
    mat4_mul(TR, S, M);
    if (!mat4_inverse(M, Minv)) for (int i = 0; i < 4; i++) for (int j = 0; j < 4; j++) Minv[i][j] = M[i][j];
    for (int i = 0; i < 3; i++) for (int j = 0; j < 4; j++) { out.fwd[4 * i + j] = M[i][j]; out.inv[4 * i + j] = Minv[i][j]; }
    out.root_ref = in.root_ref; out.blas = in.blas; out.root8 = in.root8; out.prim_type = in.prim_type;
    // 8-corner world box; the point transform includes the reference's 4th-row terms only through rows 1..3
    f3 lo = mk3(RT_INF, RT_INF, RT_INF), hi = mk3(-RT_INF, -RT_INF, -RT_INF);
    for (int i = 0; i < 2; i++) for (int j = 0; j < 2; j++) for (int k = 0; k < 2; k++) {
        const float x = i * in.box_hi[0] + (1.0f - i) * in.box_lo[0];
        const float y = j * in.box_hi[1] + (1.0f - j) * in.box_lo[1];
        const float z = k * in.box_hi[2] + (1.0f - k) * in.box_lo[2];
        const f3 p = xf_point(out.fwd, mk3(x, y, z));
        lo = mk3(rt_min(lo.x, p.x), rt_min(lo.y, p.y), rt_min(lo.z, p.z));
        hi = mk3(rt_max(hi.x, p.x), rt_max(hi.y, p.y), rt_max(hi.z, p.z));
    }
    ensure_volume(lo.x, hi.x); ensure_volume(lo.y, hi.y); ensure_volume(lo.z, hi.z);
    wbox.lo = lo; wbox.hi = hi;
    wcentroid = xf_point(out.fwd, mk3(in.centroid[0], in.centroid[1], in.centroid[2]));
}
// The 8-corner box of a ROTATED mesh is loose (a unit sphere's grows to side 2*sqrt(3): 3x the surface, so 3x the rays
// that enter the instance for nothing).  For small triangle BLASes the TLAS therefore culls with the exact world box
// of the transformed vertices, clipped to the 8-corner box; min/max are exact, so the walking order does not matter.
RT_HD void aabb_grow(Aabb &b, const f3 &p) {
    b.lo = mk3(rt_min(b.lo.x, p.x), rt_min(b.lo.y, p.y), rt_min(b.lo.z, p.z));
    b.hi = mk3(rt_max(b.hi.x, p.x), rt_max(b.hi.y, p.y), rt_max(b.hi.z, p.z));
}
RT_HD Aabb instance_tight_box(const Aabb &corner_box, const Aabb &vertex_box) {
    Aabb r;
    r.lo = mk3(rt_max(corner_box.lo.x, vertex_box.lo.x), rt_max(corner_box.lo.y, vertex_box.lo.y), rt_max(corner_box.lo.z, vertex_box.lo.z));
    r.hi = mk3(rt_min(corner_box.hi.x, vertex_box.hi.x), rt_min(corner_box.hi.y, vertex_box.hi.y), rt_min(corner_box.hi.z, vertex_box.hi.z));
    ensure_volume(r.lo.x, r.hi.x); ensure_volume(r.lo.y, r.hi.y); ensure_volume(r.lo.z, r.hi.z);
    return r;
}


// ---- config 3 input stage: triangle q of a particle frame from the strips (kernels_build.cu expand_strips_kernel; the
// CPU twin runs the same function).  Strip k with strip_first_tri[k] <= q (binary search), j = q - first; corners
// ids[j], ids[j+1], ids[j+2], the last two swapped for odd j (reference src/Global/VTKReader.cu:184-195); writes the
// 9 vertex floats and the 48-byte shading record (point normals + material) exactly as HostScene::convert_triangles
// lays them out for the same rt_triangle.
RT_HD void expand_strip_triangle(uint32_t q, const float *points, const float *normals, const int *ids, const uint32_t *strip_first_tri,
                                 const uint32_t *strip_first_id, int n_strips, uint32_t material_type, uint32_t material_index,
                                 float *tri_verts, TriShade *tri_shade) {
    int lo = 0, hi = n_strips - 1;
    while (lo < hi) {
        const int mid = (lo + hi + 1) >> 1;
        if (strip_first_tri[mid] <= q) lo = mid; else hi = mid - 1;
    }
    const uint32_t j = q - strip_first_tri[lo];
    const int *s = ids + strip_first_id[lo] + j;
    const int i0 = s[0], a = s[1], b = s[2];
    const int c[3] = {i0, (j & 1u) ? b : a, (j & 1u) ? a : b};
    float *v = tri_verts + (size_t) q * 9;
    float n[9];
    for (int k = 0; k < 3; k++)
        for (int x = 0; x < 3; x++) {
            v[3 * k + x] = points[3 * (size_t) c[k] + x];
            n[3 * k + x] = normals[3 * (size_t) c[k] + x];
        }
    TriShade t;
    t.n0_mt = make_float4(n[0], n[1], n[2], u2f(material_type));
    t.n1_mi = make_float4(n[3], n[4], n[5], u2f(material_index));
    t.n2 = make_float4(n[6], n[7], n[8], 0.0f);
    tri_shade[q] = t;
}
