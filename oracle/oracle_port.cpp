/*
 * oracle_port.cpp — TEST INFRASTRUCTURE ONLY.  Never linked, loaded or called by the product.
 *
 * From-scratch CPU restatement of the reference's per-frame ray-tracing hot path
 * (3169651074/real-time-gpu-ray-tracer), written against the reference sources and PINNED against
 * them: tests/test_oracle_cpu.py checks every export bit-for-bit against oracle/_ref/libref_oracle.so
 * (the reference's own .cu files compiled as C++), and tests/golden/ holds vectors generated from that
 * library by tests/golden/make_golden.py.  Same operation order as the reference everywhere; this file
 * must be compiled WITHOUT fp contraction / fast-math (plain g++ -O2 on x86-64).
 *
 * cuRAND XORWOW (curand_init / curand / curand_uniform) is third-party arithmetic shipped with the
 * CUDA toolkit; it is used here through its own header, made host-callable by overriding QUALIFIERS.
 *
 * Reference citations are file:line relative to the reference root.
 */
#define QUALIFIERS static inline
#include <cuda_runtime.h>
#include <curand_kernel.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <limits>
#include <map>
#include <random>
#include <thread>
#include <vector>

#define ORACLE_PREFIX port_
#include "oracle_api.h"

namespace {

constexpr float EPS = 1e-6f;                                  // FLOAT_ZERO_VALUE  include/Global/Global.cuh:147
constexpr float INF = std::numeric_limits<float>::infinity(); // include/Global/Global.cuh:148
constexpr float PI_F = (float) M_PI;                          // include/Global/Global.cuh:149

struct V3 { float x, y, z; };
inline float comp(const V3 &v, int i) { return i == 0 ? v.x : (i == 1 ? v.y : v.z); }
inline float &comp(V3 &v, int i) { return i == 0 ? v.x : (i == 1 ? v.y : v.z); }

// include/Basic/Vec3.cuh:60-138 — component loops, sums start at 0.0f
inline V3 add(const V3 &a, const V3 &b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline V3 sub(const V3 &a, const V3 &b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V3 mul(const V3 &a, float s) { return {a.x * s, a.y * s, a.z * s}; }
inline V3 neg(const V3 &a) { return {-a.x, -a.y, -a.z}; }
inline float dot(const V3 &a, const V3 &b) {
    float sum = 0.0f;
    sum += a.x * b.x; sum += a.y * b.y; sum += a.z * b.z;
    return sum;
}
inline float len2(const V3 &a) { return dot(a, a); }
inline V3 cross(const V3 &a, const V3 &b) {
    return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
inline V3 unit(const V3 &a) {                                 // Vec3.cuh:127-135: factor = 1/length, then scale
    const float factor = 1.0f / std::sqrt(len2(a));
    return {a.x * factor, a.y * factor, a.z * factor};
}
// Point3::constructVector(from, to) = to - from   include/Basic/Point3.cuh:67-73

// ---- MathHelper / Range ---------------------------------------------------------------------
inline bool nearZero(float v) { return std::fabs(v) < EPS; }             // Global.cuh:172-174
inline bool fEquals(float a, float b) { return std::fabs(a - b) < EPS; } // Global.cuh:180-182

struct Rng { float lo, hi; };
inline bool inRange(const Rng &r, float v) {                  // include/Util/Range.cuh:33-43 (both ends closed)
    if (fEquals(v, r.lo)) return true;
    if (fEquals(v, r.hi)) return true;
    return v > r.lo && v < r.hi;
}
inline float rangeLength(const Rng &r) {                      // Range.cuh:59-65
    if (r.lo >= r.hi || fEquals(r.lo, r.hi)) return 0.0f;
    return r.hi - r.lo;
}

// ---- BoundingBox ----------------------------------------------------------------------------
struct Box { Rng r[3]; };
inline void ensureVolume(Box &b) {                            // include/AS/BoundingBox.cuh:24-28
    for (auto &i: b.r) if (rangeLength(i) < EPS) { i.lo -= EPS; i.hi += EPS; }
}
inline Box boxFromPoints(const V3 &p1, const V3 &p2) {        // BoundingBox.cuh:40-46
    Box b;
    for (int i = 0; i < 3; i++) {
        const float a = comp(p1, i), c = comp(p2, i);
        b.r[i] = a < c ? Rng{a, c} : Rng{c, a};
    }
    ensureVolume(b);
    return b;
}
inline Box boxFromRanges(Rng x, Rng y, Rng z) {               // BoundingBox.cuh:34-37
    Box b{{x, y, z}};
    ensureVolume(b);
    return b;
}
inline Box boxMerge(const Box &a, const Box &b) {             // BoundingBox.cuh:49-54, Range.cuh:24-26
    Box o;
    for (int i = 0; i < 3; i++)
        o.r[i] = {a.r[i].lo < b.r[i].lo ? a.r[i].lo : b.r[i].lo, a.r[i].hi > b.r[i].hi ? a.r[i].hi : b.r[i].hi};
    return o;
}

struct RayF { V3 o, d; };
inline V3 rayAt(const RayF &r, float t) { return add(r.o, mul(r.d, t)); } // include/Basic/Ray.cuh:18-20

// src/AS/BoundingBox.cu:34-72 — slab test with true divisions, shrinking a copy of the range
inline bool boxHit(const Box &b, const RayF &ray, Rng cur, float &tOut) {
    for (int axis = 0; axis < 3; axis++) {
        const Rng &ar = b.r[axis];
        const float q = comp(ray.o, axis), d = comp(ray.d, axis);
        if (std::fabs(d) < EPS) {
            if (q < ar.lo || q > ar.hi) return false;
            continue;
        }
        const float t1 = (ar.lo - q) / d;
        const float t2 = (ar.hi - q) / d;
        if (t1 < t2) {
            if (t1 > cur.lo) cur.lo = t1;
            if (t2 < cur.hi) cur.hi = t2;
        } else {
            if (t2 > cur.lo) cur.lo = t2;
            if (t1 < cur.hi) cur.hi = t1;
        }
        if (cur.lo >= cur.hi) return false;
    }
    tOut = cur.lo;
    return true;
}

// ---- Matrix (include/Util/Matrix.cuh:22-92, src/Util/Matrix.cu) -------------------------------
struct Mat { float d[5][5]; size_t row, col; };
inline Mat matZero(size_t r, size_t c) { Mat m; memset(&m, 0, sizeof(m)); m.row = r; m.col = c; return m; }
inline Mat colPoint(const V3 &p, float w) {                   // Matrix.cuh:28-52 (w=0 vector, w=1 point)
    Mat m = matZero(4, 1);
    m.d[1][1] = p.x; m.d[2][1] = p.y; m.d[3][1] = p.z; m.d[4][1] = w;
    return m;
}
inline V3 colToV3(const Mat &m) { return {m.d[1][1], m.d[2][1], m.d[3][1]}; }
Mat matMul(const Mat &a, const Mat &b) {                      // Matrix.cu:71-86
    Mat r = matZero(a.row, b.col);
    for (size_t i = 1; i <= r.row; i++)
        for (size_t j = 1; j <= r.col; j++) {
            float sum = 0.0f;
            for (size_t n = 1; n <= a.col; n++) sum += a.d[i][n] * b.d[n][j];
            r.d[i][j] = sum;
        }
    return r;
}
Mat matTranspose(const Mat &a) {                              // Matrix.cu:89-98
    Mat r = matZero(a.col, a.row);
    for (size_t i = 1; i <= a.row; i++) for (size_t j = 1; j <= a.col; j++) r.d[j][i] = a.d[i][j];
    return r;
}
Mat matInverse(const Mat &a) {                                // Matrix.cu:4-69,101-130 (Gauss-Jordan, partial pivoting)
    float w[5][9];
    memset(w, 0, sizeof(w));
    for (size_t i = 1; i < 5; i++) {
        for (size_t j = 1; j < 5; j++) w[i][j] = a.d[i][j];
        w[i][4 + i] = 1.0f;
    }
    // forward elimination (Matrix.cu:5-44)
    for (size_t i = 1; i < 5; i++) {
        float pivot = std::fabs(w[i][i]);
        size_t maxRow = i;
        for (size_t p = i + 1; p < 5; p++) if (std::fabs(w[p][i]) > pivot) { pivot = std::fabs(w[p][i]); maxRow = p; }
        if (nearZero(pivot)) return a;                        // singular: input returned unchanged (Q14)
        if (maxRow != i) for (size_t j = 1; j < 9; j++) std::swap(w[maxRow][j], w[i][j]);
        for (size_t j = i + 1; j < 5; j++) {
            const float factor = w[j][i] / w[i][i];
            for (size_t k = i; k < 9; k++) w[j][k] -= factor * w[i][k];
        }
    }
    // back substitution (Matrix.cu:45-68)
    for (size_t i = 4; i >= 1; i--) {
        if (nearZero(w[i][i])) return a;
        float factor = 1.0f / w[i][i];
        for (size_t p = i; p < 9; p++) w[i][p] *= factor;
        for (size_t j = i - 1; j >= 1; j--) {
            factor = w[j][i];
            for (size_t k = j; k < 9; k++) w[j][k] -= factor * w[i][k];
        }
    }
    Mat r = matZero(4, 4);
    for (size_t i = 1; i < 5; i++) for (size_t j = 1; j < 5; j++) r.d[i][j] = w[i][4 + j];
    return r;
}
Mat matIdentity() { Mat m = matZero(4, 4); for (int i = 1; i < 5; i++) m.d[i][i] = 1.0f; return m; }
Mat matShift(const V3 &s) { Mat m = matIdentity(); m.d[1][4] = s.x; m.d[2][4] = s.y; m.d[3][4] = s.z; return m; } // Matrix.cu:183-194
Mat matScale(const V3 &s) { Mat m = matIdentity(); m.d[1][1] = s.x; m.d[2][2] = s.y; m.d[3][3] = s.z; return m; } // Matrix.cu:196-207
Mat matRotateAxis(float degree, int axis) {                   // Matrix.cu:209-243
    const float theta = degree * PI_F / 180.0f;               // Global.cuh:156-158
    const float c = std::cos(theta), s = std::sin(theta);
    Mat m = matIdentity();
    if (axis == 0) { m.d[2][2] = c; m.d[2][3] = -s; m.d[3][2] = s; m.d[3][3] = c; }
    else if (axis == 1) { m.d[1][1] = c; m.d[1][3] = s; m.d[3][1] = -s; m.d[3][3] = c; }
    else { m.d[1][1] = c; m.d[1][2] = -s; m.d[2][1] = s; m.d[2][2] = c; }
    return m;
}
Mat matRotate(const V3 &deg) {                                // Matrix.cu:245-250: Rx * Ry * Rz
    return matMul(matMul(matRotateAxis(deg.x, 0), matRotateAxis(deg.y, 1)), matRotateAxis(deg.z, 2));
}
inline V3 xformPoint(const Mat &m, const V3 &p) { return colToV3(matMul(m, colPoint(p, 1.0f))); }
inline V3 xformVector(const Mat &m, const V3 &v) { return colToV3(matMul(m, colPoint(v, 0.0f))); }

Box boxTransform(const Box &b, const Mat &m) {                // src/AS/BoundingBox.cu:4-32 (8 corners)
    V3 lo{INF, INF, INF}, hi{-INF, -INF, -INF};
    for (int i = 0; i < 2; i++) for (int j = 0; j < 2; j++) for (int k = 0; k < 2; k++) {
        const float x = i * b.r[0].hi + (1.0f - i) * b.r[0].lo;
        const float y = j * b.r[1].hi + (1.0f - j) * b.r[1].lo;
        const float z = k * b.r[2].hi + (1.0f - k) * b.r[2].lo;
        const V3 p = xformPoint(m, {x, y, z});
        for (int l = 0; l < 3; l++) { comp(lo, l) = std::min(comp(lo, l), comp(p, l)); comp(hi, l) = std::max(comp(hi, l), comp(p, l)); }
    }
    return boxFromPoints(lo, hi);
}

// ---- primitives -------------------------------------------------------------------------------
struct Hit {                                                  // include/Basic/BasicTypes.cuh:19-31 (+ ids)
    V3 point, normal;
    float t;
    bool front;
    bool primFront;     // Prim::hit's own hitFrontFace (outward normal against the local ray), before Instance::hit overwrites
                        // it with a test against the already flipped normal — kept for the dielectric extension only
    uint32_t matType, matIndex;
    int32_t prim;       // (type << 28) | index — not in the reference record
    int32_t instance;   // not in the reference record
};

struct SphereP { V3 c; float r; uint32_t mt, mi; };
struct QuadP { V3 q, u, v; float area; V3 n; float d; uint32_t mt, mi; };
struct TriP { V3 v[3], n[3], e1, e2; uint32_t mt, mi; };
struct DiscP { V3 c; float r; V3 n; float d; uint32_t mt, mi; };      // test-only fourth type (ORACLE_EXT_DISC)

QuadP makeQuad(const rt_parallelogram &g) {                   // include/Geometry/Parallelogram.cuh:26-39
    QuadP p{{g.q.x, g.q.y, g.q.z}, {g.u.x, g.u.y, g.u.z}, {g.v.x, g.v.y, g.v.z}, 0, {}, 0, g.material_type, g.material_index};
    p.n = cross(p.u, p.v);
    p.area = std::sqrt(len2(p.n));
    p.n = unit(p.n);
    float sum = 0.0f;
    sum += p.n.x * p.q.x; sum += p.n.y * p.q.y; sum += p.n.z * p.q.z;
    p.d = sum;
    return p;
}
TriP makeTri(const rt_triangle &g) {                          // include/Geometry/Triangle.cuh:26-46
    TriP t;
    for (int i = 0; i < 3; i++) t.v[i] = {g.v[i].x, g.v[i].y, g.v[i].z};
    t.e1 = sub(t.v[1], t.v[0]);
    t.e2 = sub(t.v[2], t.v[0]);
    if (g.has_normals) for (int i = 0; i < 3; i++) t.n[i] = {g.n[i].x, g.n[i].y, g.n[i].z};
    else for (int i = 0; i < 3; i++) t.n[i] = unit(cross(t.e1, t.e2));
    t.mt = g.material_type; t.mi = g.material_index;
    return t;
}

bool sphereHit(const SphereP &s, const RayF &ray, const Rng &range, Hit &rec) { // src/Geometry/Sphere.cu:4-49
    const V3 cq = sub(s.c, ray.o);
    const V3 dir = ray.d;
    const float a = dot(dir, dir);
    const float b = -2.0f * dot(cq, dir);
    const float c = dot(cq, cq) - s.r * s.r;
    float delta = b * b - 4.0f * a * c;
    if (delta < 0.0f) return false;
    delta = std::sqrt(delta);
    const float root1 = (-b - delta) / (a * 2.0f);
    const float root2 = (-b + delta) / (a * 2.0f);
    float root;
    if (inRange(range, root1)) root = root1;
    else if (inRange(range, root2)) root = root2;
    else return false;
    rec.t = root;
    rec.point = rayAt(ray, root);
    rec.matType = s.mt; rec.matIndex = s.mi;
    const V3 outward = unit(sub(rec.point, s.c));
    rec.front = dot(ray.d, outward) < 0.0f;
    rec.primFront = rec.front;
    rec.normal = rec.front ? outward : neg(outward);
    return true;                                              // uv (Sphere.cu:41-47) is never consumed
}
Box sphereBox(const SphereP &s) { const V3 e{s.r, s.r, s.r}; return boxFromPoints(sub(s.c, e), add(s.c, e)); } // Sphere.cu:51-55

bool quadHit(const QuadP &p, const RayF &ray, const Rng &range, Hit &rec) { // src/Geometry/Parallelogram.cu:4-46
    const float nDotD = dot(p.n, ray.d);
    if (nearZero(nDotD)) return false;
    float nDotP = 0.0f;
    nDotP += p.n.x * ray.o.x; nDotP += p.n.y * ray.o.y; nDotP += p.n.z * ray.o.z;
    const float t = (p.d - nDotP) / nDotD;
    if (!inRange(range, t)) return false;
    const V3 x = rayAt(ray, t);
    const V3 pv = sub(x, p.q);
    const V3 normal = cross(p.u, p.v);
    const float denom = len2(normal);
    if (nearZero(denom)) return false;
    const float alpha = dot(cross(pv, p.v), normal) / denom;
    const float beta = dot(cross(p.u, pv), normal) / denom;
    const Rng unitRange{0.0f, 1.0f};
    if (!inRange(unitRange, alpha) || !inRange(unitRange, beta)) return false;
    rec.t = t;
    rec.point = x;
    rec.matType = p.mt; rec.matIndex = p.mi;
    rec.front = dot(ray.d, p.n) < 0.0;
    rec.primFront = rec.front;
    rec.normal = rec.front ? p.n : neg(p.n);
    return true;
}
Box quadBox(const QuadP &p) {                                 // Parallelogram.cu:48-50 — the reference's (wrong) box, Q10
    const V3 h = mul(add(p.u, p.v), 0.5f);
    return boxFromPoints(add(p.q, h), sub(p.q, h));
}
V3 quadCentroid(const QuadP &p) { return add(add(p.q, mul(p.u, 0.5f)), mul(p.v, 0.5f)); } // Parallelogram.cuh:45-47

bool triHit(const TriP &tr, const RayF &ray, const Rng &range, Hit &rec) { // src/Geometry/Triangle.cu:4-44
    const V3 h = cross(ray.d, tr.e2);
    const float det = dot(tr.e1, h);
    if (nearZero(det)) return false;
    const V3 s = sub(ray.o, tr.v[0]);
    const Rng unitRange{0.0f, 1.0f};
    const float u = dot(s, h) / det;
    if (!inRange(unitRange, u)) return false;
    const V3 q = cross(s, tr.e1);
    const float v = dot(ray.d, q) / det;
    if (!inRange(unitRange, v) || u + v > 1.0f) return false;
    const float t = dot(tr.e2, q) / det;
    if (!inRange(range, t)) return false;
    rec.t = t;
    rec.point = rayAt(ray, t);
    rec.matType = tr.mt; rec.matIndex = tr.mi;
    const V3 n = unit(add(add(mul(tr.n[0], 1.0f - u - v), mul(tr.n[1], u)), mul(tr.n[2], v)));
    rec.front = dot(ray.d, n) < 0.0f;
    rec.primFront = rec.front;
    rec.normal = rec.front ? n : neg(n);
    return true;
}
Box triBox(const TriP &t) {                                   // Triangle.cu:46-62
    const V3 lo{std::min({t.v[0].x, t.v[1].x, t.v[2].x}), std::min({t.v[0].y, t.v[1].y, t.v[2].y}), std::min({t.v[0].z, t.v[1].z, t.v[2].z})};
    const V3 hi{std::max({t.v[0].x, t.v[1].x, t.v[2].x}), std::max({t.v[0].y, t.v[1].y, t.v[2].y}), std::max({t.v[0].z, t.v[1].z, t.v[2].z})};
    return boxFromPoints(lo, hi);
}
V3 triCentroid(const TriP &t) {                               // Triangle.cuh:52-60
    V3 c;
    for (int i = 0; i < 3; i++) { comp(c, i) = comp(t.v[0], i) + comp(t.v[1], i) + comp(t.v[2], i); comp(c, i) /= 3.0f; }
    return c;
}

// ---- BVH nodes (include/AS/BLAS.cuh:20-34, include/AS/TLAS.cuh:24-39) --------------------------
struct Node { Box box; size_t count, index; };
struct PrimRef { uint32_t type; size_t index; };
struct Bvh { std::vector<Node> nodes; std::vector<PrimRef> refs; };

struct InstanceP {                                            // include/AS/Instance.cuh:24-65
    uint32_t type; size_t first, count, blas;
    Mat M, Minv, Mnorm;
    Box box; V3 centroid;
    Box wbox; V3 wcentroid;
};

struct Scene {
    std::vector<SphereP> spheres;
    std::vector<QuadP> quads;
    std::vector<TriP> tris;
    std::vector<DiscP> discs;
    std::vector<rt_rough> roughs;
    std::vector<rt_metal> metals;
    std::vector<rt_dielectric> dielectrics;                   // extension, no reference counterpart
    std::vector<InstanceP> instances;
    std::vector<Bvh> blas;
    Bvh tlas;                                                 // refs[].index = instance id
    // camera (include/Global/RendererImpl.cuh:32-61)
    int W, H, pitch;
    V3 background, center, target, up;
    float fov, focusRadius, focusDist;
    V3 camU, camV, camW, pixelDx, pixelDy, pixelOrigin;
    size_t sampleCount, sqrtSamples, depth;
    float rcpSqrtSamples;
};

struct BuildItem { Box box; V3 centroid; uint32_t type; size_t index; };

// Shared shape of BLAS::constructBLAS (src/AS/BLAS.cu:32-116) and TLAS::constructTLAS
// (src/AS/TLAS.cu:12-128): explicit stack, random axis, std::sort by centroid, box = merge over the
// whole range, split at count/2, right pushed first.
Bvh buildMedianSplit(std::vector<BuildItem> &items, size_t leafSize, std::mt19937 &gen) {
    Bvh out;
    const size_t n = items.size();
    out.nodes.resize(2 * n - 1);
    out.refs.reserve(n);
    struct Task { size_t start, count, node; };
    std::vector<Task> stack;
    size_t nodeCount = 0;
    stack.push_back({0, n, 0});
    nodeCount++;
    auto mergeRange = [&](size_t s, size_t e) {               // BLAS.cu:208-216
        Box b = items[s].box;
        for (size_t i = s + 1; i < e; i++) b = boxMerge(b, items[i].box);
        return b;
    };
    while (!stack.empty()) {
        const Task task = stack.back();
        stack.pop_back();
        Node &node = out.nodes[task.node];
        if (task.count <= leafSize) {
            node.count = task.count;
            node.index = out.refs.size();
            node.box = mergeRange(task.start, task.start + task.count);
            for (size_t i = 0; i < task.count; i++) out.refs.push_back({items[task.start + i].type, items[task.start + i].index});
        } else {
            const size_t left = nodeCount++;
            const size_t right = nodeCount++;
            (void) right;
            std::uniform_int_distribution<int> pick(0, 2);    // RandomGenerator::randomInteger  Global.cuh:214-217
            const int axis = pick(gen);
            std::sort(items.begin() + (int) task.start, items.begin() + (int) (task.start + task.count),
                      [axis](const BuildItem &a, const BuildItem &b) { return comp(a.centroid, axis) < comp(b.centroid, axis); });
            node.box = mergeRange(task.start, task.start + task.count);
            node.count = 0;
            node.index = left;
            const size_t mid = task.count / 2;
            stack.push_back({task.start + mid, task.count - mid, left + 1});
            stack.push_back({task.start, mid, left});
        }
    }
    out.nodes.resize(nodeCount);
    return out;
}

#ifdef ORACLE_EXT_DISC
// TEST-ONLY fourth primitive type (tests/ext/): a disc, written here independently of the product's PrimOps<3> — plane hit
// as Parallelogram::hit (src/Geometry/Parallelogram.cu:4-46), then |x - c|^2 <= r^2.  Record layout = the product's ExtPrimD.
struct DiscRec { float c[3], r, n[3], d; uint32_t mt, mi, p0, p1; };
bool discHit(const DiscP &q, const RayF &ray, const Rng &range, Hit &rec) {
    const float nDotD = dot(q.n, ray.d);
    if (nearZero(nDotD)) return false;
    const float nDotP = dot(q.n, ray.o);
    const float t = (q.d - nDotP) / nDotD;
    if (!inRange(range, t)) return false;
    const V3 x = rayAt(ray, t);
    const V3 pv = sub(x, q.c);
    if (len2(pv) > q.r * q.r) return false;
    rec.t = t;
    rec.point = x;
    rec.matType = q.mt; rec.matIndex = q.mi;
    rec.front = dot(ray.d, q.n) < 0.0f;
    rec.primFront = rec.front;
    rec.normal = rec.front ? q.n : neg(q.n);
    return true;
}
Box discBox(const DiscP &q) {
    const float ex = q.r * std::sqrt(std::max(0.0f, 1.0f - q.n.x * q.n.x)), ey = q.r * std::sqrt(std::max(0.0f, 1.0f - q.n.y * q.n.y)),
                ez = q.r * std::sqrt(std::max(0.0f, 1.0f - q.n.z * q.n.z));
    return boxFromPoints({q.c.x - ex, q.c.y - ey, q.c.z - ez}, {q.c.x + ex, q.c.y + ey, q.c.z + ez});
}
#endif

bool primHit(const Scene &s, const PrimRef &ref, const RayF &ray, const Rng &range, Hit &rec) {
    switch (ref.type) {
#ifdef ORACLE_EXT_DISC
        case 3u: return discHit(s.discs[ref.index], ray, range, rec);
#endif
        case RT_SPHERE: return sphereHit(s.spheres[ref.index], ray, range, rec);
        case RT_PARALLELOGRAM: return quadHit(s.quads[ref.index], ray, range, rec);
        case RT_TRIANGLE: return triHit(s.tris[ref.index], ray, range, rec);
    }
    return false;
}

// src/AS/BLAS.cu:119-206
bool blasHit(const Scene &s, const Bvh &bvh, const RayF &ray, const Rng &range, Hit &record) {
    size_t stack[64];
    size_t sp = 0;
    stack[sp++] = 0;
    Hit temp;
    bool isHit = false;
    Rng cur = range;
    while (sp > 0) {
        const size_t idx = stack[--sp];
        float t;
        if (!boxHit(bvh.nodes[idx].box, ray, cur, t)) continue;   // node box re-tested on pop (Q2)
        const Node &node = bvh.nodes[idx];
        if (node.count > 0) {
            for (size_t i = 0; i < node.count; i++) {
                const PrimRef &ref = bvh.refs[node.index + i];
                if (primHit(s, ref, ray, cur, temp)) {
                    isHit = true;
                    cur.hi = temp.t;                              // ε-closed: later candidate within ε replaces (Q1)
                    record = temp;
                    record.prim = (int32_t) ((ref.type << 28) | (uint32_t) ref.index);
                }
            }
        } else {
            const size_t l = node.index, r = l + 1;
            float tl, tr;
            const bool hl = boxHit(bvh.nodes[l].box, ray, cur, tl);
            const bool hr = boxHit(bvh.nodes[r].box, ray, cur, tr);
            if (hl && hr) {
                if (tl > tr) { stack[sp++] = l; stack[sp++] = r; }
                else { stack[sp++] = r; stack[sp++] = l; }
            } else if (hl) stack[sp++] = l;
            else if (hr) stack[sp++] = r;
        }
    }
    return isHit;
}

// src/AS/Instance.cu:19-50
bool instanceHit(const Scene &s, const InstanceP &ins, const RayF &ray, const Rng &range, Hit &record) {
    const RayF local{xformPoint(ins.Minv, ray.o), xformVector(ins.Minv, ray.d)};   // un-normalised direction: t stays world-parametric
    if (!blasHit(s, s.blas[ins.blas], local, range, record)) return false;
    record.point = xformPoint(ins.M, record.point);
    record.normal = unit(xformVector(ins.Mnorm, record.normal));
    record.front = dot(ray.d, record.normal) < 0.0f;
    return true;
}

// src/AS/TLAS.cu:131-201
bool tlasHit(const Scene &s, const RayF &ray, const Rng &range, Hit &record) {
    const Bvh &bvh = s.tlas;
    size_t stack[64];
    size_t sp = 0;
    stack[sp++] = 0;
    Hit temp;
    bool isHit = false;
    Rng cur = range;
    while (sp > 0) {
        const size_t idx = stack[--sp];
        float t;
        if (!boxHit(bvh.nodes[idx].box, ray, cur, t)) continue;
        const Node &node = bvh.nodes[idx];
        if (node.count > 0) {
            for (size_t i = 0; i < node.count; i++) {
                const size_t instanceIndex = bvh.refs[node.index + i].index;
                if (instanceHit(s, s.instances[instanceIndex], ray, cur, temp)) {
                    isHit = true;
                    cur.hi = temp.t;
                    record = temp;
                    record.instance = (int32_t) instanceIndex;
                }
            }
        } else {
            const size_t l = node.index, r = l + 1;
            float tl, tr;
            const bool hl = boxHit(bvh.nodes[l].box, ray, cur, tl);
            const bool hr = boxHit(bvh.nodes[r].box, ray, cur, tr);
            if (hl && hr) {
                if (tl > tr) { stack[sp++] = l; stack[sp++] = r; }
                else { stack[sp++] = r; stack[sp++] = l; }
            } else if (hl) stack[sp++] = l;
            else if (hr) stack[sp++] = r;
        }
    }
    return isHit;
}

// ---- RNG-driven vectors (include/Global/Global.cuh:198-210, src/Basic/Vec3.cu:29-36,54-65) ------
inline float uniform01(curandState *st) { return curand_uniform(st); }                 // (0, 1]
inline float uniformRange(curandState *st, float lo, float hi) { return lo + (hi - lo) * uniform01(st); }
V3 randomSpaceVector(curandState *st, float length) {
    V3 v;
    float l2;
    do {
        v.x = uniformRange(st, -1.0f, 1.0f); v.y = uniformRange(st, -1.0f, 1.0f); v.z = uniformRange(st, -1.0f, 1.0f);
        l2 = len2(v);
    } while (l2 < EPS * EPS);
    v = unit(v);                                               // normalised CUBE sample (Q5)
    return mul(v, length);
}
V3 randomPlaneVector(curandState *st, float maxLength) {
    float x, y;
    do { x = uniformRange(st, -1.0f, 1.0f); y = uniformRange(st, -1.0f, 1.0f); } while (x * x + y * y > maxLength * maxLength);
    return {x, y, 0.0f};
}

// include/Material/Rough.cuh:14-29
void roughScatter(const rt_rough &m, curandState *st, const Hit &rec, V3 &attenuation, RayF &out) {
    V3 dir = add(rec.normal, randomSpaceVector(st, 1.0f));
    if (fEquals(len2(dir), EPS * EPS)) dir = rec.normal;
    out = {rec.point, dir};                                    // not normalised
    attenuation = {m.albedo.x, m.albedo.y, m.albedo.z};
}
// include/Material/Metal.cuh:15-32
bool metalScatter(const rt_metal &m, curandState *st, const RayF &in, const Hit &rec, V3 &attenuation, RayF &out) {
    const V3 v = in.d, n = rec.normal;
    V3 refl = unit(sub(v, mul(n, 2 * dot(v, n))));
    if (m.fuzz > 0.0f) refl = add(refl, mul(randomSpaceVector(st, 1.0f), m.fuzz));
    out = {rec.point, refl};
    attenuation = {m.albedo.x, m.albedo.y, m.albedo.z};
    return dot(out.d, rec.normal) > 0.0f;
}

// EXTENSION — PARITY UNPINNED: the reference has no dielectric (include/Basic/BasicTypes.cuh:14-16 lists ROUGH, METAL).
// Restated from the standard formulation (Snell refraction, Schlick reflectance, total internal reflection) in the
// conventions of Metal::scatter (include/Material/Metal.cuh:15-32): the incoming direction is normalised first (ray
// directions are not unit after a rough bounce, Q5), `normal` is the record's normal (already turned against the ray),
// the side of the surface comes from Prim::hit's own hitFrontFace, ONE uniform decides reflect / refract and is always
// drawn, (1 - cos)^5 is five multiplications (no pow), the outgoing direction is not re-normalised, always scatters.
bool dielectricScatter(const rt_dielectric &m, curandState *st, const RayF &in, const Hit &rec, V3 &attenuation, RayF &out) {
    const V3 v = unit(in.d), n = rec.normal;
    const float ri = rec.primFront ? 1.0f / m.ior : m.ior;
    float cosT = -dot(v, n);
    if (cosT > 1.0f) cosT = 1.0f;
    const float sin2 = 1.0f - cosT * cosT;
    const float sinT = std::sqrt(sin2 > 0.0f ? sin2 : 0.0f);
    float r0 = (1.0f - ri) / (1.0f + ri);
    r0 = r0 * r0;
    const float c1 = 1.0f - cosT;
    const float c5 = ((c1 * c1) * (c1 * c1)) * c1;
    const float reflectance = r0 + (1.0f - r0) * c5;
    const float xi = uniform01(st);
    V3 dir;
    if (ri * sinT > 1.0f || reflectance > xi) {
        dir = sub(v, mul(n, 2 * dot(v, n)));
    } else {
        const V3 perp = mul(add(v, mul(n, cosT)), ri);
        const float k = 1.0f - len2(perp);
        const V3 par = mul(n, -std::sqrt(std::fabs(k)));
        dir = add(perp, par);
    }
    out = {rec.point, dir};
    attenuation = {m.albedo.x, m.albedo.y, m.albedo.z};
    return true;
}

// src/Global/Kernel.cu:64-102 (live branch of rayColor)
V3 rayColor(const Scene &s, const RayF &ray, curandState *st, uint64_t &segments) {
    Hit rec{};
    RayF cur = ray;
    V3 result{1.0f, 1.0f, 1.0f};
    for (size_t depth = 0; depth < s.depth; depth++) {
        const Rng range{0.001f, INF};
        segments++;
        if (tlasHit(s, cur, range, rec)) {
            RayF out{};
            V3 att{};
            if (rec.matType == RT_ROUGH) roughScatter(s.roughs[rec.matIndex], st, rec, att, out);
            else if (rec.matType == RT_METAL) { if (!metalScatter(s.metals[rec.matIndex], st, cur, rec, att, out)) return result; }
            else if (rec.matType == RT_DIELECTRIC) dielectricScatter(s.dielectrics[rec.matIndex], st, cur, rec, att, out);   // extension
            cur = out;
            result.x *= att.x; result.y *= att.y; result.z *= att.z;
        } else {
            result.x *= s.background.x; result.y *= s.background.y; result.z *= s.background.z;
            break;
        }
    }
    return result;                                             // depth exhausted: throughput as is (Q6)
}

// src/Global/Kernel.cu:121-135
RayF primaryRay(const Scene &s, unsigned x, unsigned y, size_t si, size_t sj, curandState *st) {
    const float ox = ((static_cast<float>(sj) + uniform01(st)) * s.rcpSqrtSamples) - 0.5f;
    const float oy = ((static_cast<float>(si) + uniform01(st)) * s.rcpSqrtSamples) - 0.5f;
    const V3 sample = add(add(s.pixelOrigin, mul(s.pixelDx, x + ox)), mul(s.pixelDy, y + oy));
    V3 origin = s.center;
    if (s.focusRadius > 0.0f) {
        const V3 dv = randomPlaneVector(st, s.focusRadius);
        origin = add(add(s.center, mul(s.camU, dv.x)), mul(s.camV, dv.y));
    }
    return {origin, unit(sub(sample, origin))};
}

// include/Basic/Color3.cuh:99-114 (gamma 2, clamp [0, 0.999], x256, truncate)
inline uint8_t toByte(float c) {
    float g = std::pow(c, 1.0f / 2.0f);
    if (g > 0.999f) g = 0.999f; else if (g < 0.0f) g = 0.0f;
    return static_cast<uint8_t>(256.0f * g);
}

double nowMs() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

template<class F>
void parallelRows(int y0, int y1, int threads, F f) {
    if (threads < 1) threads = 1;
    std::atomic<int> next{y0};
    std::vector<std::thread> pool;
    for (int k = 0; k < threads; k++)
        pool.emplace_back([&]() { for (;;) { const int y = next.fetch_add(1); if (y >= y1) break; f(y); } });
    for (auto &t: pool) t.join();
}

inline void initPixelRng(const Scene &s, uint64_t seed, int x, int y, curandState *st) {
    const uint32_t pixelIndex = (uint32_t) s.pitch * (uint32_t) y + (uint32_t) x;   // Kernel.cu:109 (padded pitch)
    curand_init((unsigned long long) ((long long) pixelIndex ^ (long long) seed), pixelIndex, 0, st); // Kernel.cu:114, clock64 := seed
}

} // namespace

extern "C" {

void *port_create(const oracle_scene *sc, oracle_timing *timing) {
    auto *s = new Scene();
    for (size_t i = 0; i < sc->sphere_count; i++) {
        const auto &g = sc->spheres[i];
        s->spheres.push_back({{g.center.x, g.center.y, g.center.z}, g.radius, g.material_type, g.material_index});
    }
    for (size_t i = 0; i < sc->parallelogram_count; i++) s->quads.push_back(makeQuad(sc->parallelograms[i]));
    s->tris.reserve(sc->triangle_count);
    for (size_t i = 0; i < sc->triangle_count; i++) s->tris.push_back(makeTri(sc->triangles[i]));
    s->roughs.assign(sc->roughs, sc->roughs + sc->rough_count);
    s->metals.assign(sc->metals, sc->metals + sc->metal_count);
    if (sc->dielectric_count) s->dielectrics.assign(sc->dielectrics, sc->dielectrics + sc->dielectric_count);

#ifdef ORACLE_EXT_DISC
    for (size_t i = 0; i < sc->ext_count && sc->ext_record_bytes == sizeof(DiscRec); i++) {
        const DiscRec &r = static_cast<const DiscRec *>(sc->ext_records)[i];
        s->discs.push_back({{r.c[0], r.c[1], r.c[2]}, r.r, {r.n[0], r.n[1], r.n[2]}, r.d, r.mt, r.mi});
    }
#endif
    auto primBox = [&](uint32_t type, size_t i) {
#ifdef ORACLE_EXT_DISC
        if (type == 3u) return discBox(s->discs[i]);
#endif
        return type == RT_SPHERE ? sphereBox(s->spheres[i]) : type == RT_PARALLELOGRAM ? quadBox(s->quads[i]) : triBox(s->tris[i]);
    };
    auto primCentroid = [&](uint32_t type, size_t i) {
#ifdef ORACLE_EXT_DISC
        if (type == 3u) return s->discs[i].c;
#endif
        return type == RT_SPHERE ? s->spheres[i].c : type == RT_PARALLELOGRAM ? quadCentroid(s->quads[i]) : triCentroid(s->tris[i]);
    };

    // instance fill-in: src/Global/RenderPin.cu:124-140 (single primitive) / src/Global/VTKReader.cu:203-209 (groups)
    s->instances.resize(sc->instance_count);
    for (size_t i = 0; i < sc->instance_count; i++) {
        const auto &d = sc->instances[i];
        InstanceP &ins = s->instances[i];
        ins.type = d.type; ins.first = d.first_prim;
        if (d.prim_count <= 1) {
            ins.count = 1;
            ins.box = primBox(d.type, d.first_prim);
            ins.centroid = primCentroid(d.type, d.first_prim);
        } else {
            ins.count = d.prim_count;
            if (d.has_bounds) {
                ins.box = boxFromRanges({d.bounds_min[0], d.bounds_max[0]}, {d.bounds_min[1], d.bounds_max[1]}, {d.bounds_min[2], d.bounds_max[2]});
                ins.centroid = {d.centroid[0], d.centroid[1], d.centroid[2]};
            } else {
                Box b = primBox(d.type, d.first_prim);
                for (size_t k = 1; k < d.prim_count; k++) b = boxMerge(b, primBox(d.type, d.first_prim + k));
                ins.box = b;
                ins.centroid = {0.5f * (b.r[0].lo + b.r[0].hi), 0.5f * (b.r[1].lo + b.r[1].hi), 0.5f * (b.r[2].lo + b.r[2].hi)};
            }
        }
    }

    // BLAS per distinct (type, first): src/Global/RenderPin.cu:143-172 with the Q11 value bug fixed
    std::mt19937 gen((uint32_t) sc->build_seed);
    std::map<std::pair<uint32_t, size_t>, size_t> blasOf;
    double tBlas = 0.0;
    uint64_t blasNodes = 0;
    for (auto &ins: s->instances) {
        const auto key = std::make_pair(ins.type, ins.first);
        auto it = blasOf.find(key);
        if (it != blasOf.end()) { ins.blas = it->second; continue; }
        ins.blas = s->blas.size();
        blasOf[key] = ins.blas;
        const double t0 = nowMs();
        std::vector<BuildItem> items;                          // BLAS.cu:12-30
        items.reserve(ins.count);
        for (size_t k = 0; k < ins.count; k++) items.push_back({primBox(ins.type, ins.first + k), primCentroid(ins.type, ins.first + k), ins.type, ins.first + k});
        s->blas.push_back(buildMedianSplit(items, 4, gen));    // PRIMITIVE_COUNT_PER_LEAF_NODE  BLAS.cuh:17
        tBlas += nowMs() - t0;
        blasNodes += s->blas.back().nodes.size();
    }

    // Instance::updateTransformArguments (src/AS/Instance.cu:4-17)
    double t0 = nowMs();
    for (size_t i = 0; i < sc->instance_count; i++) {
        const auto &x = sc->xforms[i];
        InstanceP &ins = s->instances[i];
        ins.M = matMul(matMul(matShift({x.shift.x, x.shift.y, x.shift.z}), matRotate({x.rotate_deg.x, x.rotate_deg.y, x.rotate_deg.z})),
                       matScale({x.scale.x, x.scale.y, x.scale.z}));
        ins.Minv = matInverse(ins.M);
        ins.Mnorm = matTranspose(ins.Minv);
        ins.wbox = boxTransform(ins.box, ins.M);
        ins.wcentroid = xformPoint(ins.M, ins.centroid);
    }
    const double tX = nowMs() - t0;

    // TLAS::constructTLAS (src/AS/TLAS.cu:4-129), INSTANCE_COUNT_PER_LEAF_NODE = 2 (TLAS.cuh:22)
    t0 = nowMs();
    {
        std::vector<BuildItem> items;
        for (size_t i = 0; i < s->instances.size(); i++) items.push_back({s->instances[i].wbox, s->instances[i].wcentroid, 0, i});
        s->tlas = buildMedianSplit(items, 2, gen);
    }
    const double tTlas = nowMs() - t0;
    if (timing) { timing->blas_build_ms = tBlas; timing->xform_ms = tX; timing->tlas_build_ms = tTlas; timing->blas_nodes = blasNodes; timing->tlas_nodes = s->tlas.nodes.size(); }

    // RendererImpl::calculateCameraProperties (src/Global/RenderPin.cu:73-95)
    s->W = (int) sc->window.width; s->H = (int) sc->window.height;
    s->pitch = ((s->W + 15) / 16) * 16;
    s->background = {sc->camera.background.x, sc->camera.background.y, sc->camera.background.z};
    s->center = {sc->camera.center.x, sc->camera.center.y, sc->camera.center.z};
    s->target = {sc->camera.target.x, sc->camera.target.y, sc->camera.target.z};
    s->up = {sc->camera.up.x, sc->camera.up.y, sc->camera.up.z};
    s->fov = sc->camera.fov; s->focusRadius = sc->camera.focus_disk_radius;
    s->sampleCount = sc->camera.sample_count; s->depth = sc->camera.ray_trace_depth;
    {
        float d2 = 0.0;                                        // Point3::distanceSquared  Point3.cuh:78-84
        d2 += (s->center.x - s->target.x) * (s->center.x - s->target.x);
        d2 += (s->center.y - s->target.y) * (s->center.y - s->target.y);
        d2 += (s->center.z - s->target.z) * (s->center.z - s->target.z);
        s->focusDist = std::sqrt(d2);
        const float theta = s->fov * PI_F / 180.0f;
        const float vW = 2.0f * std::tan(theta / 2.0f) * s->focusDist;
        const float vH = vW / (static_cast<float>(s->W) * 1.0f / static_cast<float>(s->H));
        s->camW = unit(sub(s->target, s->center));
        s->camU = unit(cross(s->camW, s->up));
        s->camV = unit(cross(s->camU, s->camW));
        const V3 vpX = mul(s->camU, vW), vpY = mul(s->camV, vH);
        s->pixelDx = {vpX.x / static_cast<float>(s->W), vpX.y / static_cast<float>(s->W), vpX.z / static_cast<float>(s->W)};
        s->pixelDy = {vpY.x / static_cast<float>(s->H), vpY.y / static_cast<float>(s->H), vpY.z / static_cast<float>(s->H)};
        const V3 vpOrigin = sub(sub(add(s->center, mul(s->camW, s->focusDist)), mul(vpX, 0.5f)), mul(vpY, 0.5f));
        s->pixelOrigin = add(add(vpOrigin, mul(s->pixelDx, 0.5f)), mul(s->pixelDy, 0.5f));
        s->sqrtSamples = static_cast<size_t>(std::sqrt(s->sampleCount));
        s->rcpSqrtSamples = 1.0f / static_cast<float>(s->sqrtSamples);
    }
    return s;
}

void port_destroy(void *h) { delete (Scene *) h; }

void port_render(void *h, uint64_t seed, int y0, int y1, int threads, uint8_t *rgba8, float *radiance, uint64_t *rays) {
    const Scene &s = *(Scene *) h;
    std::atomic<uint64_t> total{0};
    parallelRows(y0, y1, threads, [&](int y) {
        uint64_t segs = 0;
        for (int x = 0; x < s.W; x++) {
            curandState st;
            initPixelRng(s, seed, x, y, &st);
            V3 result{0, 0, 0};
            for (size_t si = 0; si < s.sqrtSamples; si++)
                for (size_t sj = 0; sj < s.sqrtSamples; sj++) {
                    const RayF ray = primaryRay(s, (unsigned) x, (unsigned) y, si, sj, &st);
                    const V3 c = rayColor(s, ray, &st, segs);
                    result.x += c.x; result.y += c.y; result.z += c.z;
                }
            const float w = s.rcpSqrtSamples * s.rcpSqrtSamples;   // Kernel.cu:143
            result = mul(result, w);
            const size_t o = (size_t) y * s.W + x;
            if (radiance) { radiance[3 * o] = result.x; radiance[3 * o + 1] = result.y; radiance[3 * o + 2] = result.z; }
            if (rgba8) { rgba8[4 * o] = toByte(result.x); rgba8[4 * o + 1] = toByte(result.y); rgba8[4 * o + 2] = toByte(result.z); rgba8[4 * o + 3] = 255; }
        }
        total += segs;
    });
    if (rays) *rays += total.load();
}

void port_primary(void *h, uint64_t seed, int y0, int y1, int threads, float *t, int32_t *instance, int32_t *prim) {
    const Scene &s = *(Scene *) h;
    parallelRows(y0, y1, threads, [&](int y) {
        for (int x = 0; x < s.W; x++) {
            curandState st;
            initPixelRng(s, seed, x, y, &st);
            const RayF ray = primaryRay(s, (unsigned) x, (unsigned) y, 0, 0, &st);
            Hit rec{};
            rec.t = INF; rec.instance = -1; rec.prim = -1;
            if (!tlasHit(s, ray, {0.001f, INF}, rec)) { rec.t = INF; rec.instance = -1; rec.prim = -1; }
            const size_t o = (size_t) y * s.W + x;
            if (t) t[o] = rec.t;
            if (instance) instance[o] = rec.instance;
            if (prim) prim[o] = rec.prim;
        }
    });
}

void port_trace(void *h, const float *rays, size_t n, float tmin, float *t, int32_t *instance, int32_t *prim, float *normal) {
    const Scene &s = *(Scene *) h;
    for (size_t i = 0; i < n; i++) {
        const RayF ray{{rays[6 * i], rays[6 * i + 1], rays[6 * i + 2]}, {rays[6 * i + 3], rays[6 * i + 4], rays[6 * i + 5]}};
        Hit rec{};
        const bool hit = tlasHit(s, ray, {tmin, INF}, rec);
        if (t) t[i] = hit ? rec.t : INF;
        if (instance) instance[i] = hit ? rec.instance : -1;
        if (prim) prim[i] = hit ? rec.prim : -1;
        if (normal) { normal[3 * i] = hit ? rec.normal.x : 0; normal[3 * i + 1] = hit ? rec.normal.y : 0; normal[3 * i + 2] = hit ? rec.normal.z : 0; }
    }
}

void port_camera(void *h, float *o) {
    const Scene &s = *(Scene *) h;
    const float v[20] = {s.pixelOrigin.x, s.pixelOrigin.y, s.pixelOrigin.z, s.pixelDx.x, s.pixelDx.y, s.pixelDx.z, s.pixelDy.x, s.pixelDy.y, s.pixelDy.z,
                         s.camU.x, s.camU.y, s.camU.z, s.camV.x, s.camV.y, s.camV.z, s.camW.x, s.camW.y, s.camW.z, s.rcpSqrtSamples, (float) s.sqrtSamples};
    memcpy(o, v, sizeof(v));
}

void port_instances(void *h, float *o) {
    const Scene &s = *(Scene *) h;
    for (size_t i = 0; i < s.instances.size(); i++) {
        const InstanceP &ins = s.instances[i];
        float *p = o + 33 * i;
        for (int r = 0; r < 3; r++) for (int c = 0; c < 4; c++) { p[4 * r + c] = ins.M.d[r + 1][c + 1]; p[12 + 4 * r + c] = ins.Minv.d[r + 1][c + 1]; }
        for (int a = 0; a < 3; a++) { p[24 + a] = ins.wbox.r[a].lo; p[27 + a] = ins.wbox.r[a].hi; }
        p[30] = ins.wcentroid.x; p[31] = ins.wcentroid.y; p[32] = ins.wcentroid.z;
    }
}

void port_uniforms(uint64_t seed, uint64_t subsequence, size_t n, float *out) {
    curandState st;
    curand_init(seed, subsequence, 0, &st);
    for (size_t i = 0; i < n; i++) out[i] = uniform01(&st);
}

const char *port_describe(void) { return "CPU restatement (oracle/oracle_port.cpp), pinned against oracle/_ref and tests/golden"; }

} // extern "C"
