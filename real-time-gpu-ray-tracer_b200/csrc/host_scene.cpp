#include "host_scene.h"
#include "host_parallel.h"

#include <algorithm>
#include <cmath>
#include <map>
#include <tuple>

namespace {

constexpr float PI_F = (float) M_PI;   // reference include/Global/Global.cuh:149

inline f3 V(const rt_vec3 &v) { return mk3(v.x, v.y, v.z); }

// BoundingBox(Point3, Point3) (reference include/AS/BoundingBox.cuh:40-46): per-axis ordering + ensureVolume
HostBox ref_box_from_points(const f3 &a, const f3 &b) {
    HostBox r;
    for (int i = 0; i < 3; i++) {
        const float x = comp3(a, i), y = comp3(b, i);
        r.lo[i] = x < y ? x : y;
        r.hi[i] = x < y ? y : x;
        ensure_volume(r.lo[i], r.hi[i]);
    }
    return r;
}
HostBox ref_box_merge(const HostBox &a, const HostBox &b) {   // BoundingBox(b1, b2), BoundingBox.cuh:49-54
    HostBox r;
    for (int i = 0; i < 3; i++) { r.lo[i] = a.lo[i] < b.lo[i] ? a.lo[i] : b.lo[i]; r.hi[i] = a.hi[i] > b.hi[i] ? a.hi[i] : b.hi[i]; }
    return r;
}

struct BlasKey {
    uint32_t type; uint64_t first, count;
    bool operator<(const BlasKey &o) const { return std::tie(type, first, count) < std::tie(o.type, o.first, o.count); }
};

} // namespace

int morton_bits_for(size_t n) {
    // bits per axis: fine enough that cells hold O(1) primitives, coarse enough to keep the sort short
    if (n <= (1u << 12)) return 8;
    if (n <= (1u << 18)) return 10;
    if (n <= (1u << 22)) return 13;
    return 16;
}

int tlas_morton_bits(size_t n_instances) {
    // instances are few and differ wildly in size: never fewer than 10 bits per axis (one more sort pass on ~1e3 keys)
    return std::max(10, morton_bits_for(n_instances));
}

std::string HostScene::commit_geometry(const rt_sphere *s, size_t ns, const rt_parallelogram *p, size_t np, const rt_triangle *t, size_t nt) {
    if ((ns && !s) || (np && !p) || (nt && !t)) return "rt_commit_geometry: null array with non-zero count";
    if (nt > RT_REF_POS_MASK || ns > RT_REF_POS_MASK || np > RT_REF_POS_MASK) return "rt_commit_geometry: more than 2^29 primitives of one type";
    long long mm[3] = {-1, -1, -1};
    auto note = [&mm](uint32_t type, uint32_t index) { if (type <= RT_DIELECTRIC) { if ((long long) index > mm[type]) mm[type] = index; return true; } return false; };
    for (size_t i = 0; i < ns; i++) if (!note(s[i].material_type, s[i].material_index)) return "rt_commit_geometry: sphere with an unknown material type";
    for (size_t i = 0; i < np; i++) if (!note(p[i].material_type, p[i].material_index)) return "rt_commit_geometry: parallelogram with an unknown material type";
    for (size_t i = 0; i < nt; i++) if (!note(t[i].material_type, t[i].material_index)) return "rt_commit_geometry: triangle with an unknown material type";
    for (int k = 0; k < 3; k++) max_mat_index[k] = mm[k];
    spheres.assign(s, s + ns);
    quads.assign(p, p + np);
    n_tris = nt;
    sphere_rec.resize(ns);
    for (size_t i = 0; i < ns; i++) {
        sphere_rec[i].c_r = make_float4(s[i].center.x, s[i].center.y, s[i].center.z, s[i].radius);
        sphere_rec[i].mat_type = s[i].material_type; sphere_rec[i].mat_index = s[i].material_index;
        sphere_rec[i]._p0 = sphere_rec[i]._p1 = 0;
    }
    quad_rec.resize(np);
    for (size_t i = 0; i < np; i++) {
        // Parallelogram constructor (reference include/Geometry/Parallelogram.cuh:26-39)
        const rt_parallelogram &g = p[i];
        const f3 n = unit3(cross3(V(g.u), V(g.v)));
        float d = 0.0f;
        d += n.x * g.q.x; d += n.y * g.q.y; d += n.z * g.q.z;
        const bool compat = !(flags & RT_FLAG_FIX_PARALLELOGRAM_BBOX);
        const f3 h = (V(g.u) + V(g.v)) * 0.5f;                                    // Parallelogram.cu:48-50 (Q10)
        const HostBox rb = ref_box_from_points(V(g.q) + h, V(g.q) - h);
        QuadD &o = quad_rec[i];
        o.q_d = make_float4(g.q.x, g.q.y, g.q.z, d);
        o.u = make_float4(g.u.x, g.u.y, g.u.z, 0.0f);
        o.v = make_float4(g.v.x, g.v.y, g.v.z, 0.0f);
        o.n_flag = make_float4(n.x, n.y, n.z, compat ? 1.0f : 0.0f);
        o.rb_lo = make_float4(rb.lo[0], rb.lo[1], rb.lo[2], 0.0f);
        o.rb_hi = make_float4(rb.hi[0], rb.hi[1], rb.hi[2], 0.0f);
        o.mat_type = g.material_type; o.mat_index = g.material_index; o._p0 = o._p1 = 0;
    }
    tri_verts.resize(nt * 9);
    tri_shade.resize(nt);
    convert_triangles(t, nt, 0);
    return "";
}

size_t HostScene::prim_count_of(uint32_t type) const {
    if (type == RT_SPHERE) return spheres.size();
    if (type == RT_PARALLELOGRAM) return quads.size();
    if (type == RT_TRIANGLE) return n_tris;
#if defined(RT_EXT_PRIMITIVE)
    if (type == 3u) return n_ext;
#endif
    return 0;
}

std::string HostScene::commit_extension(const void *records, size_t record_bytes, size_t count) {
#if defined(RT_EXT_PRIMITIVE)
    if (record_bytes != sizeof(ExtPrimD)) return std::string("rt_commit_extension_primitives: a ") + RT_EXT_PRIMITIVE_NAME + " record is " + std::to_string(sizeof(ExtPrimD)) + " bytes";
    if (count && !records) return "rt_commit_extension_primitives: null array with non-zero count";
    if (count > RT_REF_POS_MASK) return "rt_commit_extension_primitives: more than 2^29 primitives of one type";
    const ExtPrimD *r = static_cast<const ExtPrimD *>(records);
    for (size_t i = 0; i < count; i++) {
        uint32_t mt, mi;
        PrimOps<3>::material(r[i], mt, mi);
        if (mt > RT_DIELECTRIC) return "rt_commit_extension_primitives: unknown material type";
        note_material(mt, mi);
    }
    ext_rec.assign(reinterpret_cast<const unsigned char *>(records), reinterpret_cast<const unsigned char *>(records) + count * sizeof(ExtPrimD));
    n_ext = count;
    return "";
#else
    (void) records; (void) record_bytes; (void) count;
    return "rt_commit_extension_primitives: this build of librtb200 has no extension primitive type (compile with -DRT_EXT_PRIMITIVE, csrc/rt_prims.h)";
#endif
}

std::string HostScene::commit_triangle_tail(size_t n_head, const rt_triangle *t, size_t nt) {
    if (nt && !t) return "commit_triangle_tail: null array with non-zero count";
    if (n_head + nt > RT_REF_POS_MASK) return "rt_commit_geometry: more than 2^29 primitives of one type";
    for (size_t i = 0; i < nt; i++) {
        if (t[i].material_type > RT_DIELECTRIC) return "rt_commit_geometry: triangle with an unknown material type";
        note_material(t[i].material_type, t[i].material_index);
    }
    n_tris = n_head + nt;
    tri_verts.resize(n_tris * 9);
    tri_shade.resize(n_tris);
    convert_triangles(t, nt, n_head);
    return "";
}

void HostScene::convert_triangles(const rt_triangle *t, size_t nt, size_t at) {
    parallel_for(nt, true, [&](size_t q) {
        // Triangle constructors (reference include/Geometry/Triangle.cuh:26-46)
        const size_t i = at + q;
        const rt_triangle &g = t[q];
        float *v = &tri_verts[i * 9];
        for (int k = 0; k < 3; k++) { v[3 * k] = g.v[k].x; v[3 * k + 1] = g.v[k].y; v[3 * k + 2] = g.v[k].z; }
        f3 n0, n1, n2;
        if (g.has_normals) { n0 = V(g.n[0]); n1 = V(g.n[1]); n2 = V(g.n[2]); }
        else {
            const f3 e1 = V(g.v[1]) - V(g.v[0]), e2 = V(g.v[2]) - V(g.v[0]);
            n0 = n1 = n2 = unit3(cross3(e1, e2));
        }
        tri_shade[i].n0_mt = make_float4(n0.x, n0.y, n0.z, u2f(g.material_type));
        tri_shade[i].n1_mi = make_float4(n1.x, n1.y, n1.z, u2f(g.material_index));
        tri_shade[i].n2 = make_float4(n2.x, n2.y, n2.z, 0.0f);
    }, 65536);
}

std::string HostScene::commit_materials(const rt_rough *r, size_t nr, const rt_metal *m, size_t nm) {
    if ((nr && !r) || (nm && !m)) return "rt_commit_materials: null array with non-zero count";
    roughs.resize(nr);
    metals.resize(nm);
    for (size_t i = 0; i < nr; i++) roughs[i] = make_float4(r[i].albedo.x, r[i].albedo.y, r[i].albedo.z, 0.0f);
    for (size_t i = 0; i < nm; i++) metals[i] = make_float4(m[i].albedo.x, m[i].albedo.y, m[i].albedo.z, m[i].fuzz);
    return "";
}

std::string HostScene::commit_dielectrics(const rt_dielectric *d, size_t n) {
    if (n && !d) return "rt_commit_dielectrics: null array";
    dielectrics.resize(n);
    for (size_t i = 0; i < n; i++) dielectrics[i] = make_float4(d[i].albedo.x, d[i].albedo.y, d[i].albedo.z, d[i].ior);
    return "";
}

std::string HostScene::check_materials() const {
    const size_t have[3] = {roughs.size(), metals.size(), dielectrics.size()};
    static const char *name[3] = {"rough", "metal", "dielectric"};
    for (int k = 0; k < 3; k++)
        if (max_mat_index[k] >= (long long) have[k])
            return std::string("geometry refers to ") + name[k] + " material #" + std::to_string(max_mat_index[k]) + " but only " +
                   std::to_string(have[k]) + " are committed";
    return "";
}

std::string HostScene::configure_instances(const rt_instance_desc *d, size_t n, const rt_instance_xform *x) {
    if (!d || n == 0) return "rt_configure_instances: no instances";
    if (n > RT_REF_POS_MASK) return "rt_configure_instances: too many instances";
    for (size_t i = 0; i < n; i++) {
        const uint64_t cnt = d[i].prim_count <= 1 ? 1 : d[i].prim_count;
        const uint64_t limit = prim_count_of(d[i].type);
        if (d[i].type >= (uint32_t) RT_PRIMITIVE_TYPE_COUNT || d[i].first_prim + cnt > limit) return "rt_configure_instances: primitive range out of bounds";
    }
    inst_desc.assign(d, d + n);
    xforms.resize(n);
    for (size_t i = 0; i < n; i++) xforms[i] = x ? x[i] : rt_instance_xform{{0, 0, 0}, {0, 0, 0}, {1, 1, 1}};
    return "";
}

std::string HostScene::plan_blas() {
    // one BLAS per distinct primitive range (the reference's intent, src/Global/RenderPin.cu:143-172; Q11 not replicated)
    const size_t n_inst = inst_desc.size();
    std::map<BlasKey, int> blas_of;
    seg_offset.assign(1, 0u);
    seg_type.clear();
    seg_first.clear();
    inst_blas.resize(n_inst);
    inst_box.resize(n_inst);
    inst_centroid.resize(n_inst);
    for (size_t i = 0; i < n_inst; i++) {
        const rt_instance_desc &d = inst_desc[i];
        const uint64_t cnt = d.prim_count <= 1 ? 1 : d.prim_count;
        const BlasKey key{d.type, d.first_prim, cnt};
        auto it = blas_of.find(key);
        if (it == blas_of.end()) {
            it = blas_of.emplace(key, (int) seg_type.size()).first;
            seg_type.push_back(d.type);
            seg_first.push_back((uint32_t) d.first_prim);
            if ((uint64_t) seg_offset.back() + cnt > RT_REF_POS_MASK) return "BLAS batch exceeds 2^29 primitives";
            seg_offset.push_back(seg_offset.back() + (uint32_t) cnt);
        }
        inst_blas[i] = it->second;

        // local box + centroid, reference semantics (src/Global/RenderPin.cu:124-140 for single primitives;
        // caller-provided bounds as src/Global/VTKReader.cu:206-207; else merged boxes + box centre)
        auto prim_box = [&](uint64_t k) -> HostBox {
            if (d.type == RT_SPHERE) {
                const rt_sphere &s = spheres[k];
                const f3 e = mk3(s.radius, s.radius, s.radius);                   // Sphere.cu:51-55
                return ref_box_from_points(V(s.center) - e, V(s.center) + e);
            }
            if (d.type == RT_PARALLELOGRAM) {
                const rt_parallelogram &p = quads[k];
                if (flags & RT_FLAG_FIX_PARALLELOGRAM_BBOX) {
                    const Aabb b = quad_aabb(quad_rec[k]);
                    return ref_box_from_points(b.lo, b.hi);
                }
                const f3 h = (V(p.u) + V(p.v)) * 0.5f;
                return ref_box_from_points(V(p.q) + h, V(p.q) - h);
            }
#if defined(RT_EXT_PRIMITIVE)
            if (d.type == 3u) {
                f3 lo, hi;
                PrimOps<3>::ref_extent(reinterpret_cast<const ExtPrimD *>(ext_rec.data())[k], lo, hi);
                return ref_box_from_points(lo, hi);
            }
#endif
            const float *v = &tri_verts[k * 9];                                   // Triangle.cu:46-62
            const f3 lo = mk3(std::min({v[0], v[3], v[6]}), std::min({v[1], v[4], v[7]}), std::min({v[2], v[5], v[8]}));
            const f3 hi = mk3(std::max({v[0], v[3], v[6]}), std::max({v[1], v[4], v[7]}), std::max({v[2], v[5], v[8]}));
            return ref_box_from_points(lo, hi);
        };
        HostBox &box = inst_box[i];
        f3 &centroid = inst_centroid[i];
        if (d.prim_count <= 1) {
            box = prim_box(d.first_prim);
            if (d.type == RT_SPHERE) centroid = V(spheres[d.first_prim].center);
            else if (d.type == RT_PARALLELOGRAM) {
                const rt_parallelogram &p = quads[d.first_prim];
                centroid = (V(p.q) + V(p.u) * 0.5f) + V(p.v) * 0.5f;              // Parallelogram.cuh:45-47
#if defined(RT_EXT_PRIMITIVE)
            } else if (d.type == 3u) {
                centroid = PrimOps<3>::centroid(reinterpret_cast<const ExtPrimD *>(ext_rec.data())[d.first_prim]);
#endif
            } else {
                const float *v = &tri_verts[d.first_prim * 9];                    // Triangle.cuh:52-60
                float cc[3];
                for (int a = 0; a < 3; a++) { cc[a] = v[a] + v[3 + a] + v[6 + a]; cc[a] /= 3.0f; }
                centroid = mk3(cc[0], cc[1], cc[2]);
            }
        } else if (d.has_bounds) {
            for (int a = 0; a < 3; a++) { box.lo[a] = d.bounds_min[a]; box.hi[a] = d.bounds_max[a]; ensure_volume(box.lo[a], box.hi[a]); }
            centroid = mk3(d.centroid[0], d.centroid[1], d.centroid[2]);
        } else {
            box = prim_box(d.first_prim);
            for (uint64_t k = 1; k < d.prim_count; k++) box = ref_box_merge(box, prim_box(d.first_prim + k));
            centroid = mk3(0.5f * (box.lo[0] + box.hi[0]), 0.5f * (box.lo[1] + box.hi[1]), 0.5f * (box.lo[2] + box.hi[2]));
        }
    }
    return "";
}

// RendererImpl::calculateCameraProperties (reference src/Global/RenderPin.cu:73-95), same operation order
std::string HostScene::configure_camera(const rt_window &w, const rt_camera_input &in) {
    CameraD c{};
    const std::string e = make_camera(w, in, c);
    if (!e.empty()) return e;
    window = w;
    cam = c;
    return "";
}

std::string HostScene::make_camera(const rt_window &w, const rt_camera_input &in, CameraD &cam) {
    if (w.width == 0 || w.height == 0) return "rt_configure_camera: bad window";
    if (in.sample_count == 0) return "rt_configure_camera: sample_count must be >= 1";
    // the path kernels carry (sample index, depth) of a path in one 32-bit word, 16 bits each
    if (in.sample_count >= 256ull * 256ull) return "rt_configure_camera: sample_count must be below 65536 (255 x 255 strata)";
    if (in.ray_trace_depth > 65535ull) return "rt_configure_camera: ray_trace_depth must be at most 65535";
    const f3 center = V(in.center), target = V(in.target), up = V(in.up);
    float d2 = 0.0f;
    d2 += (center.x - target.x) * (center.x - target.x);
    d2 += (center.y - target.y) * (center.y - target.y);
    d2 += (center.z - target.z) * (center.z - target.z);
    const float focusDistance = std::sqrt(d2);
    const float theta = in.fov * PI_F / 180.0f;
    const float vW = 2.0f * std::tan(theta / 2.0f) * focusDistance;
    const float vH = vW / (static_cast<float>(w.width) * 1.0f / static_cast<float>(w.height));
    const f3 camW = unit3(target - center);
    const f3 camU = unit3(cross3(camW, up));
    const f3 camV = unit3(cross3(camU, camW));
    const f3 vpX = camU * vW, vpY = camV * vH;
    const float fw = static_cast<float>(w.width), fh = static_cast<float>(w.height);
    cam.pixelDx = mk3(vpX.x / fw, vpX.y / fw, vpX.z / fw);
    cam.pixelDy = mk3(vpY.x / fh, vpY.y / fh, vpY.z / fh);
    const f3 vpOrigin = ((center + camW * focusDistance) - vpX * 0.5f) - vpY * 0.5f;
    cam.pixelOrigin = (vpOrigin + cam.pixelDx * 0.5f) + cam.pixelDy * 0.5f;
    cam.background = V(in.background);
    cam.center = center;
    cam.camU = camU;
    cam.camV = camV;
    cam.focusRadius = in.focus_disk_radius;
    cam.sqrtSamples = (uint32_t) static_cast<size_t>(std::sqrt((double) in.sample_count));
    cam.rcpSqrtSamples = 1.0f / static_cast<float>(cam.sqrtSamples);
    cam.depth = (uint32_t) in.ray_trace_depth;
    cam.width = w.width;
    cam.height = w.height;
    cam.pitch = ((w.width + 15u) / 16u) * 16u;   // gridDim.x * blockDim.x of the reference launch (src/Global/Kernel.cu:109)
    return "";
}

void HostScene::fill_xform_in(InstanceXformIn *out, const std::vector<int> &seg_root) const {
    const size_t n = inst_desc.size();
    for (size_t i = 0; i < n; i++) {
        const rt_instance_xform &x = xforms[i];
        InstanceXformIn &o = out[i];
        o.shift[0] = x.shift.x; o.shift[1] = x.shift.y; o.shift[2] = x.shift.z;
        o.scale[0] = x.scale.x; o.scale[1] = x.scale.y; o.scale[2] = x.scale.z;
        // MathHelper::degreeToRadian + cos/sin as Matrix::constructRotateMatrix (reference src/Util/Matrix.cu:209-243)
        const float tx = x.rotate_deg.x * PI_F / 180.0f, ty = x.rotate_deg.y * PI_F / 180.0f, tz = x.rotate_deg.z * PI_F / 180.0f;
        o.trig[0] = std::cos(tx); o.trig[1] = std::sin(tx);
        o.trig[2] = std::cos(ty); o.trig[3] = std::sin(ty);
        o.trig[4] = std::cos(tz); o.trig[5] = std::sin(tz);
        for (int a = 0; a < 3; a++) { o.box_lo[a] = inst_box[i].lo[a]; o.box_hi[a] = inst_box[i].hi[a]; }
        o.centroid[0] = inst_centroid[i].x; o.centroid[1] = inst_centroid[i].y; o.centroid[2] = inst_centroid[i].z;
        o.root_ref = seg_root[inst_blas[i]];
        o.blas = inst_blas[i];
        o.root8 = inst_blas[i];                 // the wide root of BLAS s is wide node s (kernels_wide.cu)
        o.prim_type = (int) inst_desc[i].type;
        const bool rotated = o.trig[1] != 0.0f || o.trig[3] != 0.0f || o.trig[5] != 0.0f;      // unrotated: the 8-corner box is exact
        const bool tight = inst_desc[i].type == RT_TRIANGLE && rotated && inst_desc[i].prim_count >= 2 && inst_desc[i].prim_count <= RT_TIGHT_BOUNDS_MAX_TRIS;
        o.tri_first = tight ? (int) inst_desc[i].first_prim : 0;
        o.tri_count = tight ? (int) inst_desc[i].prim_count : 0;
    }
}

size_t HostScene::prim_bytes() const {
    size_t b = 0;
    for (size_t s = 0; s < seg_type.size(); s++) {
        const size_t cnt = seg_offset[s + 1] - seg_offset[s];
        size_t rec = seg_type[s] == RT_TRIANGLE ? sizeof(TriIsect) : seg_type[s] == RT_SPHERE ? sizeof(SphereD) : sizeof(QuadD);
#if defined(RT_EXT_PRIMITIVE)
        if (seg_type[s] == 3u) rec = sizeof(ExtPrimD);
#endif
        b += cnt * rec;
    }
    return b;
}


// ---- scripted camera input ------------------------------------------------------------------------------------------
namespace {
// Vec3::rotate (reference include/Basic/Vec3.cuh:140-160): Rodrigues, axis normalised first
f3 rotate_about(const f3 &v, const f3 &axis, float angle) {
    const f3 k = unit3(axis);
    const float c = std::cos(angle), s = std::sin(angle);
    const f3 part1 = v * c;
    const f3 part2 = cross3(k, v) * s;
    const f3 part3 = (k * dot3(k, v)) * (1.0f - c);
    return (part1 + part2) + part3;
}
}

bool camera_step(const rt_camera_input &cur, const rt_camera_motion &m, rt_camera_input &next) {
    next = cur;
    const f3 center = V(cur.center), target = V(cur.target), up = V(cur.up);
    // camera frame as RendererImpl::calculateCameraProperties builds it (src/Global/RenderPin.cu:83-85)
    const f3 camW = unit3(target - center);
    const f3 camU = unit3(cross3(camW, up));
    const f3 camV = unit3(cross3(camU, camW));
    f3 new_center = center, new_target = target;
    bool moved = false;
    if (m.dx != 0 || m.dy != 0) {                              // mouse: only the look-at point changes (:193-236)
        moved = true;
        const f3 view = target - center;
        f3 W = unit3(camW);
        const f3 U = unit3(camU), Vv = unit3(camV);
        const float yaw = -static_cast<float>(m.dx) * m.mouse_sensitivity;
        W = rotate_about(W, Vv, yaw);
        const float pitch = -static_cast<float>(m.dy) * m.mouse_sensitivity;
        W = rotate_about(W, U, pitch);
        float new_pitch = std::asin(W.y);
        bool fix = false;
        if (new_pitch > m.pitch_limit) { new_pitch = m.pitch_limit; fix = true; }
        else if (new_pitch < -m.pitch_limit) { new_pitch = -m.pitch_limit; fix = true; }
        if (fix) {
            const f3 horizontal = unit3(mk3(W.x, 0.0f, W.z));
            W = horizontal * std::cos(new_pitch) + mk3(0.0f, std::sin(new_pitch), 0.0f);
        }
        new_target = center + W * std::sqrt(len2(view));
    }
    f3 dir = mk3(0.0f, 0.0f, 0.0f);                            // keys: always in the horizontal plane (:238-254)
    const f3 forward = unit3(mk3(camW.x, 0.0f, camW.z));
    if (m.keys & RT_KEY_W) dir = dir + forward;
    if (m.keys & RT_KEY_S) dir = dir - forward;
    if (m.keys & RT_KEY_D) dir = dir + camU;
    if (m.keys & RT_KEY_A) dir = dir - camU;
    if (m.keys & RT_KEY_SPACE) dir = dir + up;
    if (m.keys & RT_KEY_LSHIFT) dir = dir - up;
    if (len2(dir) > 0.0f) {
        moved = true;
        const f3 t = unit3(dir) * m.move_speed;
        new_center = new_center + t;
        new_target = new_target + t;
    }
    next.center = rt_vec3{new_center.x, new_center.y, new_center.z};
    next.target = rt_vec3{new_target.x, new_target.y, new_target.z};
    return moved;
}
