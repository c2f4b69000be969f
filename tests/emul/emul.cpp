/*
 * emul.cpp — TEST INFRASTRUCTURE ONLY (never part of librtb200.so, never used by the product path).
 *
 * Serial CPU execution of the product's OWN per-element device code: the headers under
 * real-time-gpu-ray-tracer_b200/csrc/ are written __host__ __device__, so g++ compiles the very same
 * Morton/Karras/refit/emit, traversal, intersection, RNG and shading functions that the CUDA kernels call.
 * Purpose (there is no GPU in the development container):
 *   1. `not gpu` tests check this code against the oracle before any GPU time is spent;
 *   2. `gpu` tests use it as the "CPU twin" of the LBVH build: device topology must equal the serial result
 *      bit for bit (same keys, same order, same children, same boxes).
 */
#define QUALIFIERS static inline
#include <cuda_runtime.h>
#include <curand_kernel.h>

#include <algorithm>
#include <atomic>
#include <cstring>
#include <numeric>
#include <thread>
#include <vector>

#include "../../real-time-gpu-ray-tracer_b200/csrc/host_scene.h"
#include "../../real-time-gpu-ray-tracer_b200/csrc/vtk_reader.h"
#include "../../real-time-gpu-ray-tracer_b200/csrc/image_io.h"
#include "../../real-time-gpu-ray-tracer_b200/csrc/rt_pixel.h"
#include "../../oracle/oracle_api.h"

namespace {

struct Lbvh {
    int n = 0;
    std::vector<Aabb> prim_box;
    std::vector<uint64_t> keys;
    std::vector<uint32_t> vals;
    std::vector<int> left, right, parent, leaf_parent;
    std::vector<int2> range;
    std::vector<Aabb> node_box;
    std::vector<BvhNode2> nodes;
    std::vector<int> seg_root;
    // wide (BVH8) collapse of this tree: nodes, and wide leaf order -> Morton position
    std::vector<BvhNode8> nodes8;
    std::vector<uint32_t> leaf_pos8;
};

int find_segment(const std::vector<uint32_t> &off, uint32_t j) {
    int lo = 0, hi = (int) off.size() - 2;
    while (lo < hi) { const int mid = (lo + hi + 1) >> 1; if (off[mid] <= j) lo = mid; else hi = mid - 1; }
    return lo;
}

// serial twin of LbvhBuilder::run_ploc (kernels_ploc.cu): the same rounds, every pass a plain loop
void build_ploc(Lbvh &b, const std::vector<uint32_t> &seg_offset) {
    const int n = b.n, n_segs = (int) seg_offset.size() - 1, m_nodes = n > 1 ? n - 1 : 0;
    std::vector<float4> lo(n), hi(n), lo2(n), hi2(n);
    for (int pos = 0; pos < n; pos++) {
        const Aabb &bx = b.prim_box[b.vals[pos]];
        lo[pos] = make_float4(bx.lo.x, bx.lo.y, bx.lo.z, u2f((uint32_t) ~pos));
        hi[pos] = make_float4(bx.hi.x, bx.hi.y, bx.hi.z, u2f((uint32_t) find_segment(seg_offset, (uint32_t) pos)));
    }
    b.left.assign(m_nodes, -1); b.right.assign(m_nodes, -1); b.parent.assign(m_nodes, -1); b.leaf_parent.assign(n, -1);
    Aabb zero;
    zero.lo = mk3(0.f, 0.f, 0.f); zero.hi = zero.lo;
    b.node_box.assign(m_nodes, zero);
    std::vector<int> nn(n);
    int m = n, next_id = 0;
    while (m > n_segs) {
        for (int i = 0; i < m; i++) nn[i] = ploc_nearest(lo.data(), hi.data(), m, i, RT_PLOC_RADIUS);
        int dst = 0;
        for (int i = 0; i < m; i++) {
            const int role = ploc_role(nn.data(), i);
            if (role == 2) continue;
            float4 a = lo[i], c = hi[i];
            if (role == 1) {
                const int j = nn[i], id = next_id++;
                const int l = (int) f2u(a.w), r = (int) f2u(lo[j].w);
                a = make_float4(fminf(a.x, lo[j].x), fminf(a.y, lo[j].y), fminf(a.z, lo[j].z), u2f((uint32_t) id));
                c = make_float4(fmaxf(c.x, hi[j].x), fmaxf(c.y, hi[j].y), fmaxf(c.z, hi[j].z), c.w);
                b.node_box[id].lo = mk3(a.x, a.y, a.z);
                b.node_box[id].hi = mk3(c.x, c.y, c.z);
                b.left[id] = l; b.right[id] = r;
                if (l >= 0) b.parent[l] = id; else b.leaf_parent[~l] = id;
                if (r >= 0) b.parent[r] = id; else b.leaf_parent[~r] = id;
            }
            lo2[dst] = a; hi2[dst] = c;
            dst++;
        }
        m = dst;
        lo.swap(lo2); hi.swap(hi2);
    }
    for (int i = 0; i < m; i++) {
        const int ref = (int) f2u(lo[i].w);
        if (ref >= 0) b.seg_root[(int) f2u(hi[i].w)] = ref;
    }
}

// serial twin of LbvhBuilder::run_tree
void build_tree(Lbvh &b, const std::vector<uint32_t> &seg_offset, const std::vector<uint32_t> &seg_type, int leaf_mode, bool rotate, bool ploc = false) {
    const int n = b.n;
    const int n_segs = (int) seg_type.size();
    std::vector<uint32_t> order(n);
    std::iota(order.begin(), order.end(), 0u);
    std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t c) { return b.keys[a] < b.keys[c]; });
    std::vector<uint64_t> sk(n);
    for (int i = 0; i < n; i++) sk[i] = b.keys[order[i]];
    b.keys = sk;
    b.vals = order;
    b.seg_root.assign(n_segs, 0);
    const int m = n > 1 ? n - 1 : 0;
    b.left.assign(m, 0); b.right.assign(m, 0); b.parent.assign(m, -1); b.leaf_parent.assign(n, -1);
    b.node_box.resize(m); b.nodes.resize(m); b.range.resize(m);
    if (ploc && m > 0) {
        for (int i = 0; i < m; i++) b.range[i] = make_int2(0, 0);
        build_ploc(b, seg_offset);
    }
    for (int i = 0; i < m && !ploc; i++) {
        const KarrasNode k = karras_node(b.keys.data(), n, i);
        b.range[i] = make_int2(k.first, k.last);
        b.left[i] = k.left; b.right[i] = k.right;
        if (k.left >= 0) b.parent[k.left] = i; else b.leaf_parent[~k.left] = i;
        if (k.right >= 0) b.parent[k.right] = i; else b.leaf_parent[~k.right] = i;
        const int seg = find_segment(seg_offset, (uint32_t) k.first);
        if ((uint32_t) k.first == seg_offset[seg] && (uint32_t) k.last + 1u == seg_offset[seg + 1]) b.seg_root[seg] = i;
    }
    for (int s = 0; s < n_segs; s++)
        if (seg_offset[s + 1] - seg_offset[s] == 1u)
            b.seg_root[s] = leaf_mode == 0 ? make_leaf_ref(seg_type[s], seg_offset[s]) : make_leaf_ref(0u, b.vals[seg_offset[s]]);
    // refit in post-order (explicit stack)
    if (m > 0 && ploc) {
        // PLOC made the boxes while merging; only the traversal records of the binary comparison mode are left to write
        for (int i = 0; i < m; i++) {
            Aabb bx[2];
            int ref[2];
            const int child[2] = {b.left[i], b.right[i]};
            for (int c = 0; c < 2; c++) {
                if (child[c] >= 0) { bx[c] = b.node_box[child[c]]; ref[c] = child[c]; }
                else {
                    const uint32_t pos = (uint32_t) ~child[c];
                    bx[c] = b.prim_box[b.vals[pos]];
                    ref[c] = leaf_mode == 0 ? make_leaf_ref(seg_type[find_segment(seg_offset, pos)], pos) : make_leaf_ref(0u, b.vals[pos]);
                }
            }
            b.nodes[i] = make_node2(bx[0], ref[0], bx[1], ref[1]);
        }
    } else if (m > 0) {
        std::vector<char> done(m, 0);
        std::vector<int> st{0};
        auto child_box = [&](int c) { return c >= 0 ? b.node_box[c] : b.prim_box[b.vals[~c]]; };
        while (!st.empty()) {
            const int i = st.back();
            const int l = b.left[i], r = b.right[i];
            const bool lready = l < 0 || done[l], rready = r < 0 || done[r];
            if (lready && rready) {
                // SAH rotation at node i (same decision function as the device refit, kernels_build.cu refit_kernel)
                int L = l, R = r;
                bool may = rotate && (L >= 0 || R >= 0);
                if (may) may = find_segment(seg_offset, (uint32_t) b.range[i].x) == find_segment(seg_offset, (uint32_t) b.range[i].y);
                if (may) {
                    Aabb z{}, nb{};
                    const Aabb LL = L >= 0 ? child_box(b.left[L]) : z, LR = L >= 0 ? child_box(b.right[L]) : z;
                    const Aabb RL = R >= 0 ? child_box(b.left[R]) : z, RR = R >= 0 ? child_box(b.right[R]) : z;
                    const int rot = best_rotation(child_box(L), child_box(R), L >= 0, LL, LR, R >= 0, RL, RR, nb);
                    auto set_parent = [&](int c, int p) { if (c >= 0) b.parent[c] = p; else b.leaf_parent[~c] = p; };
                    if (rot == 1) { const int ll = b.left[L]; b.left[L] = R; b.node_box[L] = nb; set_parent(R, L); set_parent(ll, i); R = ll; }
                    else if (rot == 2) { const int lr = b.right[L]; b.right[L] = R; b.node_box[L] = nb; set_parent(R, L); set_parent(lr, i); R = lr; }
                    else if (rot == 3) { const int rl = b.left[R]; b.left[R] = L; b.node_box[R] = nb; set_parent(L, R); set_parent(rl, i); L = rl; }
                    else if (rot == 4) { const int rr = b.right[R]; b.right[R] = L; b.node_box[R] = nb; set_parent(L, R); set_parent(rr, i); L = rr; }
                    b.left[i] = L; b.right[i] = R;
                }
                b.node_box[i] = aabb_union(child_box(L), child_box(R));
                done[i] = 1;
                st.pop_back();
            } else {
                if (!lready) st.push_back(l);
                if (!rready) st.push_back(r);
            }
        }
        // nodes that belong to no segment (junk above the BLAS roots) are also reachable from node 0: done.
        for (int i = 0; i < m; i++) {
            Aabb bx[2];
            int ref[2];
            const int child[2] = {b.left[i], b.right[i]};
            for (int c = 0; c < 2; c++) {
                if (child[c] >= 0) { bx[c] = b.node_box[child[c]]; ref[c] = child[c]; }
                else {
                    const uint32_t pos = (uint32_t) ~child[c];
                    bx[c] = b.prim_box[b.vals[pos]];
                    ref[c] = leaf_mode == 0 ? make_leaf_ref(seg_type[find_segment(seg_offset, pos)], pos) : make_leaf_ref(0u, b.vals[pos]);
                }
            }
            b.nodes[i] = make_node2(bx[0], ref[0], bx[1], ref[1]);
        }
    }
}

// serial twin of the device collapse (kernels_wide.cu): queue entries are processed in index order and
// allocations are sequential, so the layout is the canonical one the GPU layout is compared against
void collapse_wide(Lbvh &b, const std::vector<uint32_t> &seg_offset, int leaf_max) {
    const int n = b.n, n_segs = (int) seg_offset.size() - 1;
    std::vector<float4> nlo(b.node_box.size()), nhi(b.node_box.size()), plo(n), phi(n);
    for (size_t i = 0; i < b.node_box.size(); i++) {
        nlo[i] = make_float4(b.node_box[i].lo.x, b.node_box[i].lo.y, b.node_box[i].lo.z, u2f((uint32_t) b.left[i]));
        nhi[i] = make_float4(b.node_box[i].hi.x, b.node_box[i].hi.y, b.node_box[i].hi.z, u2f((uint32_t) b.right[i]));
    }
    for (int i = 0; i < n; i++) {
        plo[i] = make_float4(b.prim_box[i].lo.x, b.prim_box[i].lo.y, b.prim_box[i].lo.z, 0.f);
        phi[i] = make_float4(b.prim_box[i].hi.x, b.prim_box[i].hi.y, b.prim_box[i].hi.z, 0.f);
    }
    const BinaryTreeView view{b.range.data(), nlo.data(), nhi.data(), plo.data(), phi.data(), b.vals.data()};
    std::vector<int> queue;
    for (int s = 0; s < n_segs; s++)
        queue.push_back(seg_offset[s + 1] - seg_offset[s] == 1u ? ~(int) seg_offset[s] : b.seg_root[s]);
    b.nodes8.clear();
    b.leaf_pos8.clear();
    for (size_t i = 0; i < queue.size(); i++) {
        WideChildren w;
        wide_expand(view, queue[i], leaf_max, w);
        const uint32_t child_base = (uint32_t) queue.size(), prim_base = (uint32_t) b.leaf_pos8.size();
        BvhNode8 node;
        int child_in_slot[8];
        wide_encode(w, child_base, prim_base, node, child_in_slot);
        for (int sl = 0; sl < 8; sl++) {
            const int c = child_in_slot[sl];
            if (c < 0) continue;
            if (w.ref[c] >= 0) queue.push_back(w.ref[c]);
            else for (int k = 0; k < w.count[c]; k++) b.leaf_pos8.push_back((uint32_t) (w.first[c] + k));
        }
        b.nodes8.push_back(node);
    }
}

struct Emul {
    HostScene hs;
    Lbvh blas, tlas;
    std::vector<TriIsect> leaf_tris, leaf_tris8;
    std::vector<uint32_t> leaf_prims, leaf_prims8, tlas_leaf_inst;
    bool wide = true;
    std::vector<InstanceD> instances;
    std::vector<Aabb> wbox;
    std::vector<f3> wcentroid;
    SceneD sc{};
    // emul_simt_tokens / emul_simt_run: per pixel (tile-major within rows [tok_y0, tok_y1)) the steps its lane takes in the path kernel
    std::vector<std::vector<uint8_t>> tok;
    int tok_y0 = 0, tok_y1 = 0;
};

template<class F>
void parallel_rows(int y0, int y1, int threads, F f) {
    if (threads < 1) threads = 1;
    std::atomic<int> next{y0};
    std::vector<std::thread> pool;
    for (int k = 0; k < threads; k++) pool.emplace_back([&]() { for (;;) { const int y = next.fetch_add(1); if (y >= y1) break; f(y); } });
    for (auto &t: pool) t.join();
}

// The SINGLE instantiation of the path kernel (kernels_persist.cu) on the CPU: a scene of ONE instance is traced without its
// top level — the world ray goes into the instance's space at once and starts on the BLAS root.  Written here (not in the product
// headers) from the same pieces the kernel uses: load_instance_wide, xf_point / xf_vector, wide_setup_ray, node8_intersect,
// leaf_test8.  emul_single_check runs every ray of rows [y0, y1) through BOTH walks and compares the hit records bit for bit.
template<bool COUNT>
void trace_single8(const SceneD &sc, const RayD &wray, float tmin, HitD &hit, uint32_t &overflow, TravStats &stats, uint2 *stack_mem) {
    hit.t = RT_INF; hit.tc = RT_INF; hit.u = 0.0f; hit.v = 0.0f; hit.inst = -1; hit.leaf = 0;
    WideStack st;
    st.base = stack_mem; st.stride = 1; st.sp = 0;
    float inv[12];
    int root8, prim_type;
    load_instance_wide(sc, 0, inv, root8, prim_type);
    if (COUNT) stats.insts++;
    WideRay r;
    r.ray.o = xf_point(inv, wray.o);
    r.ray.d = xf_vector(inv, wray.d);
    wide_setup_ray(r);
    uint2 ng = make_uint2((uint32_t) root8, 0x80000000u);
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
            const BvhNode8 n = load_node8(sc.blas_nodes8, base + rel);
            if (COUNT) stats.nodes++;
            const uint32_t hm = node8_intersect(n, r, tmin, hit.tc);
            ng.x = n.n1.x;
            ng.y = (hm & 0xFF000000u) | (f2u(n.n0.w) >> 24);
            tg.x = n.n1.y;
            tg.y = hm & 0x00FFFFFFu;
        } else {
            tg = ng;
            ng = make_uint2(0u, 0u);
        }
        while (tg.y) {
            const int idx = rt_bfind(tg.y);
            tg.y &= ~(1u << idx);
            if (COUNT) stats.prims++;
            leaf_test8(sc, (uint32_t) prim_type, tg.x + (uint32_t) idx, r.ray, tmin, 0, hit);
        }
        if (ng.y <= 0x00FFFFFFu) {
            if (st.sp == 0) break;
            ng = wide_pop(st);
        }
    }
}

} // namespace

extern "C" {

void *emul_create(const oracle_scene *s, uint32_t flags) {
    auto *e = new Emul();
    HostScene &hs = e->hs;
    hs.flags = flags;
    if (!hs.commit_geometry(s->spheres, s->sphere_count, s->parallelograms, s->parallelogram_count, s->triangles, s->triangle_count).empty() ||
        !hs.commit_materials(s->roughs, s->rough_count, s->metals, s->metal_count).empty() ||
        !hs.commit_dielectrics(s->dielectrics, s->dielectric_count).empty() ||
        (s->ext_count && !hs.commit_extension(s->ext_records, s->ext_record_bytes, s->ext_count).empty()) || !hs.check_materials().empty() ||
        !hs.configure_instances(s->instances, s->instance_count, s->xforms).empty() || !hs.plan_blas().empty() ||
        !hs.configure_camera(s->window, s->camera).empty()) { delete e; return nullptr; }

    // ---- BLAS batch (twin of LbvhBuilder::build_blas)
    Lbvh &b = e->blas;
    const int n = (int) hs.seg_offset.back(), n_segs = (int) hs.seg_type.size();
    b.n = n;
    b.prim_box.resize(n);
    std::vector<Aabb> seg_bounds(n_segs);
    for (auto &sb: seg_bounds) { sb.lo = mk3(RT_INF, RT_INF, RT_INF); sb.hi = mk3(-RT_INF, -RT_INF, -RT_INF); }
    for (int j = 0; j < n; j++) {
        const int seg = find_segment(hs.seg_offset, (uint32_t) j);
        const uint32_t prim = hs.seg_first[seg] + ((uint32_t) j - hs.seg_offset[seg]);
        const GeometryD geo{hs.tri_verts.data(), hs.sphere_rec.data(), hs.quad_rec.data(), hs.ext_rec.data()};
        Aabb box = prim_bounds(geo, hs.seg_type[seg], prim);             // the primitive table (rt_prims.h), as prim_bounds_kernel
        box = aabb_pad(box);
        b.prim_box[j] = box;
        seg_bounds[seg] = aabb_union(seg_bounds[seg], box);
    }
    int seg_bits = 0;
    while ((1 << seg_bits) < n_segs) seg_bits++;
    int bits = morton_bits_for((size_t) n / std::max(1, n_segs));
    if (3 * bits + seg_bits > 64) bits = (64 - seg_bits) / 3;
    b.keys.resize(n);
    for (int j = 0; j < n; j++) {
        const int seg = find_segment(hs.seg_offset, (uint32_t) j);
        b.keys[j] = ((uint64_t) seg << (3 * bits)) | morton_code(b.prim_box[j], seg_bounds[seg], bits);
    }
    build_tree(b, hs.seg_offset, hs.seg_type, 0, (hs.flags & RT_FLAG_SAH_REFINE) != 0, (hs.flags & RT_FLAG_SAH_BUILD) != 0);
    e->leaf_prims.resize(n);
    e->leaf_tris.resize(n);
    for (int pos = 0; pos < n; pos++) {
        const int seg = find_segment(hs.seg_offset, (uint32_t) pos);
        const uint32_t prim = hs.seg_first[seg] + (b.vals[pos] - hs.seg_offset[seg]);
        e->leaf_prims[pos] = prim;
        if (hs.seg_type[seg] == 2u) {
            const float *v = &hs.tri_verts[(size_t) prim * 9];
            const f3 v0 = mk3(v[0], v[1], v[2]), e1 = mk3(v[3], v[4], v[5]) - v0, e2 = mk3(v[6], v[7], v[8]) - v0;
            e->leaf_tris[pos].v0_id = make_float4(v0.x, v0.y, v0.z, u2f(prim));
            e->leaf_tris[pos].e1 = make_float4(e1.x, e1.y, e1.z, 0.0f);
            e->leaf_tris[pos].e2 = make_float4(e2.x, e2.y, e2.z, 0.0f);
        }
    }

    collapse_wide(b, hs.seg_offset, RT_WIDE_LEAF_MAX);
    e->leaf_prims8.resize(n);
    e->leaf_tris8.resize(n);
    for (int q = 0; q < n; q++) { e->leaf_prims8[q] = e->leaf_prims[b.leaf_pos8[q]]; e->leaf_tris8[q] = e->leaf_tris[b.leaf_pos8[q]]; }

    // ---- instances + TLAS (twin of update_instances_and_tlas)
    const int ni = (int) hs.inst_desc.size();
    std::vector<InstanceXformIn> xin(ni);
    hs.fill_xform_in(xin.data(), b.seg_root);
    for (int i = 0; i < ni; i++) xin[i].root8 = xin[i].blas;      // wide root of BLAS s is wide node s
    e->instances.resize(ni); e->wbox.resize(ni); e->wcentroid.resize(ni);
    Lbvh &t = e->tlas;
    t.n = ni;
    t.prim_box.resize(ni);
    Aabb scene;
    scene.lo = mk3(RT_INF, RT_INF, RT_INF); scene.hi = mk3(-RT_INF, -RT_INF, -RT_INF);
    for (int i = 0; i < ni; i++) {
        instance_update(xin[i], e->instances[i], e->wbox[i], e->wcentroid[i]);
        if (xin[i].tri_count > 0) {          // twin of instance_bounds_kernel: exact world box of the transformed vertices
            Aabb v;
            v.lo = mk3(RT_INF, RT_INF, RT_INF); v.hi = mk3(-RT_INF, -RT_INF, -RT_INF);
            const float *p = hs.tri_verts.data() + (size_t) xin[i].tri_first * 9;
            for (int k = 0; k < xin[i].tri_count * 3; k++) aabb_grow(v, xf_point(e->instances[i].fwd, mk3(p[3 * k], p[3 * k + 1], p[3 * k + 2])));
            t.prim_box[i] = aabb_pad(instance_tight_box(e->wbox[i], v));     // e->wbox stays the reference's transformedBoundingBox
        } else {
            t.prim_box[i] = aabb_pad(e->wbox[i]);
        }
        scene = aabb_union(scene, aabb_centre_box(t.prim_box[i]));        // the grid spans the instance centres
    }
    const int tbits = tlas_morton_bits((size_t) ni);
    t.keys.resize(ni);
    for (int i = 0; i < ni; i++) t.keys[i] = morton_code(t.prim_box[i], scene, tbits);
    const std::vector<uint32_t> toff{0u, (uint32_t) ni}, ttype{0u};
    build_tree(t, toff, ttype, 1, (hs.flags & RT_FLAG_SAH_REFINE) != 0, getenv("EMUL_TLAS_PLOC") != nullptr);
    collapse_wide(t, toff, 1);
    e->tlas_leaf_inst.resize(ni);
    for (int q = 0; q < ni; q++) e->tlas_leaf_inst[q] = t.vals[t.leaf_pos8[q]];

    SceneD &sc = e->sc;
    sc.tlas_nodes = t.nodes.data(); sc.blas_nodes = b.nodes.data(); sc.instances = e->instances.data();
    sc.leaf_tris = e->leaf_tris.data(); sc.leaf_prims = e->leaf_prims.data(); sc.tri_shade = hs.tri_shade.data();
    sc.spheres = hs.sphere_rec.data(); sc.quads = hs.quad_rec.data();
    sc.roughs = hs.roughs.data(); sc.metals = hs.metals.data(); sc.dielectrics = hs.dielectrics.data();
    sc.tlas_root_ref = ni > 1 ? 0 : make_leaf_ref(0u, 0u);
    sc.tlas_nodes8 = t.nodes8.data(); sc.blas_nodes8 = b.nodes8.data(); sc.tlas_leaf_inst = e->tlas_leaf_inst.data();
    sc.leaf_tris8 = e->leaf_tris8.data(); sc.leaf_prims8 = e->leaf_prims8.data();
    sc.ext_prims = hs.ext_rec.data();
    e->wide = !(flags & 0x80000000u);     // test-only switch: traverse the binary LBVH instead of the wide BVH
    return e;
}

void emul_destroy(void *h) { delete (Emul *) h; }

// out: per pixel like the product planes; returns stack-overflow flag
uint32_t emul_render(void *h, uint64_t seed, int y0, int y1, int threads, uint8_t *rgba8, float *radiance, uint64_t *rays,
                     float *aov_t, int32_t *aov_inst, int32_t *aov_prim) {
    Emul &e = *(Emul *) h;
    const CameraD &cam = e.hs.cam;
    std::atomic<uint64_t> total{0};
    std::atomic<uint32_t> ovf{0};
    parallel_rows(y0, y1, threads, [&](int y) {
        uint64_t segs = 0;
        for (uint32_t x = 0; x < cam.width; x++) {
            const uint32_t pixelIndex = cam.pitch * (uint32_t) y + x;
            curandStateXORWOW_t st;
            curand_init((unsigned long long) ((long long) pixelIndex ^ (long long) seed), pixelIndex, 0, &st);
            RngState r{st.d, st.v[0], st.v[1], st.v[2], st.v[3], st.v[4]};
            uint32_t overflow = 0;
            uint2 stack[RT_WIDE_STACK];
            const PixelOut o = e.wide ? render_pixel<false, true>(e.sc, cam, x, (uint32_t) y, r, overflow, stack, 1)
                                      : render_pixel<false, false>(e.sc, cam, x, (uint32_t) y, r, overflow, stack, 1);
            if (overflow) ovf = 1;
            segs += o.rays;
            const size_t p = (size_t) y * cam.width + x;
            if (radiance) { radiance[3 * p] = o.radiance.x; radiance[3 * p + 1] = o.radiance.y; radiance[3 * p + 2] = o.radiance.z; }
            if (rgba8) { rgba8[4 * p] = to_byte(o.radiance.x); rgba8[4 * p + 1] = to_byte(o.radiance.y); rgba8[4 * p + 2] = to_byte(o.radiance.z); rgba8[4 * p + 3] = 255; }
            if (aov_t) { aov_t[p] = o.aov_t; aov_inst[p] = o.aov_inst; aov_prim[p] = o.aov_prim; }
        }
        total += segs;
    });
    if (rays) *rays += total.load();
    return ovf.load();
}

// traversal work of rows [y0, y1): out[0] rays, [1] wide-node fetches, [2] primitive tests, [3] instance entries — the
// counts the device's RT_FLAG_COUNT_TRAVERSAL kernel reports (same functions, COUNT = true), for tree-quality checks
void emul_count(void *h, uint64_t seed, int y0, int y1, int threads, uint64_t *out4) {
    Emul &e = *(Emul *) h;
    const CameraD &cam = e.hs.cam;
    std::atomic<uint64_t> tot[4];
    for (auto &t: tot) t = 0;
    parallel_rows(y0, y1, threads, [&](int y) {
        uint64_t acc[4] = {0, 0, 0, 0};
        for (uint32_t x = 0; x < cam.width; x++) {
            const uint32_t pixelIndex = cam.pitch * (uint32_t) y + x;
            curandStateXORWOW_t st;
            curand_init((unsigned long long) ((long long) pixelIndex ^ (long long) seed), pixelIndex, 0, &st);
            RngState r{st.d, st.v[0], st.v[1], st.v[2], st.v[3], st.v[4]};
            uint32_t overflow = 0;
            uint2 stack[RT_WIDE_STACK];
            const PixelOut o = render_pixel<true, true>(e.sc, cam, x, (uint32_t) y, r, overflow, stack, 1);
            acc[0] += o.rays; acc[1] += o.stats.nodes; acc[2] += o.stats.prims; acc[3] += o.stats.insts;
        }
        for (int k = 0; k < 4; k++) tot[k] += acc[k];
    });
    for (int k = 0; k < 4; k++) out4[k] = tot[k].load();
}

// ---- a trace-driven model of the path kernel's warp votes (profiles/tools/simt_model.py) ----------------------------------------
// Token stream of one ray in the SINGLE instantiation of persist8_kernel: one TOK_NODE per wide-node step, one TOK_LEAF per primitive
// test (the kernel tests ONE primitive per lane and LEAF step), closed by the SHADE step that ends the ray and starts the next.
enum : uint8_t { TOK_NODE = 0, TOK_LEAF = 1, TOK_SHADE_GO_ON = 2, TOK_SHADE_PATH_END = 3 };

void trace_single8_tokens(const SceneD &sc, const RayD &wray, float tmin, HitD &hit, uint2 *stack_mem, std::vector<uint8_t> &tok) {
    hit.t = RT_INF; hit.tc = RT_INF; hit.u = 0.0f; hit.v = 0.0f; hit.inst = -1; hit.leaf = 0;
    WideStack st;
    st.base = stack_mem; st.stride = 1; st.sp = 0;
    uint32_t overflow = 0;
    float inv[12];
    int root8, prim_type;
    load_instance_wide(sc, 0, inv, root8, prim_type);
    WideRay r;
    r.ray.o = xf_point(inv, wray.o);
    r.ray.d = xf_vector(inv, wray.d);
    wide_setup_ray(r);
    uint2 ng = make_uint2((uint32_t) root8, 0x80000000u);
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
            const BvhNode8 n = load_node8(sc.blas_nodes8, base + rel);
            tok.push_back(TOK_NODE);
            const uint32_t hm = node8_intersect(n, r, tmin, hit.tc);
            ng.x = n.n1.x;
            ng.y = (hm & 0xFF000000u) | (f2u(n.n0.w) >> 24);
            tg.x = n.n1.y;
            tg.y = hm & 0x00FFFFFFu;
        } else {
            tg = ng;
            ng = make_uint2(0u, 0u);
        }
        while (tg.y) {
            const int idx = rt_bfind(tg.y);
            tg.y &= ~(1u << idx);
            tok.push_back(TOK_LEAF);
            leaf_test8(sc, (uint32_t) prim_type, tg.x + (uint32_t) idx, r.ray, tmin, 0, hit);
        }
        if (ng.y <= 0x00FFFFFFu) {
            if (st.sp == 0) break;
            ng = wide_pop(st);
        }
    }
}

// out[0] rays, [1] rays whose hit records differ between the two walks (must be 0), [2] wide-node fetches of the two-level walk,
// [3] of the single-level walk, [4] rays that miss the scene's top-level box (the only rays the top level saves work on), [5] primitive
// tests two-level, [6] single-level.  Returns -1 unless the scene is exactly one instance.
int emul_single_check(void *h, uint64_t seed, int y0, int y1, int threads, uint64_t *out7) {
    Emul &e = *(Emul *) h;
    if (e.hs.inst_desc.size() != 1 || !e.wide) return -1;
    const CameraD &cam = e.hs.cam;
    std::atomic<uint64_t> tot[7];
    for (auto &t: tot) t = 0;
    parallel_rows(y0, y1, threads, [&](int y) {
        uint64_t acc[7] = {0, 0, 0, 0, 0, 0, 0};
        for (uint32_t x = 0; x < cam.width; x++) {
            const uint32_t pixelIndex = cam.pitch * (uint32_t) y + x;
            curandStateXORWOW_t cs;
            curand_init((unsigned long long) ((long long) pixelIndex ^ (long long) seed), pixelIndex, 0, &cs);
            RngState rng{cs.d, cs.v[0], cs.v[1], cs.v[2], cs.v[3], cs.v[4]};
            uint2 stack[RT_WIDE_STACK];
            for (uint32_t si = 0; si < cam.sqrtSamples; si++) {
                for (uint32_t sj = 0; sj < cam.sqrtSamples; sj++) {
                    RayD ray = primary_ray(cam, x, (uint32_t) y, si, sj, rng);
                    f3 thr = mk3(1.0f, 1.0f, 1.0f);
                    for (uint32_t depth = 0; depth < cam.depth; depth++) {
                        HitD a, b;
                        TravStats sa{0, 0, 0}, sb{0, 0, 0};
                        uint32_t ovf = 0;
                        trace_closest8<true>(e.sc, ray, RT_TMIN, a, ovf, sa, stack, 1);
                        trace_single8<true>(e.sc, ray, RT_TMIN, b, ovf, sb, stack);
                        acc[0]++;
                        if (std::memcmp(&a.t, &b.t, 4) || std::memcmp(&a.tc, &b.tc, 4) || std::memcmp(&a.u, &b.u, 4) || std::memcmp(&a.v, &b.v, 4) ||
                            a.inst != b.inst || a.leaf != b.leaf) acc[1]++;
                        acc[2] += sa.nodes; acc[3] += sb.nodes; acc[4] += sa.insts == 0; acc[5] += sa.prims; acc[6] += sb.prims;
                        if (!shade_scatter<true>(e.sc, cam, ray, a, thr, rng)) break;
                    }
                }
            }
        }
        for (int k = 0; k < 7; k++) tot[k] += acc[k];
    });
    for (int k = 0; k < 7; k++) out7[k] = tot[k].load();
    return 0;
}

// Token streams of every pixel of rows [y0, y1) (y0, y1 multiples of 4; a one-instance scene), kept in the handle for emul_simt_run.
// NOTE for variant 1 / 2 runs on two_level streams: the instance entry is a LEAF token of its own run (a node token follows it).
// Returns the number of tokens, -1 on a scene that is not one instance.
long long emul_simt_tokens(void *h, uint64_t seed, int y0, int y1, int threads, int two_level) {
    Emul &e = *(Emul *) h;
    if (e.hs.inst_desc.size() != 1 || !e.wide || (y0 & 3) || (y1 & 3)) return -1;
    const CameraD &cam = e.hs.cam;
    const uint32_t tiles_x = (cam.width + 7u) / 8u;
    const int tile_rows = (y1 - y0) / 4;
    e.tok.assign((size_t) tiles_x * tile_rows * 32u, std::vector<uint8_t>());
    e.tok_y0 = y0; e.tok_y1 = y1;
    std::atomic<long long> total{0};
    parallel_rows(y0, y1, threads, [&](int y) {
        long long n = 0;
        for (uint32_t x = 0; x < cam.width; x++) {
            const uint32_t pixelIndex = cam.pitch * (uint32_t) y + x;
            curandStateXORWOW_t cs;
            curand_init((unsigned long long) ((long long) pixelIndex ^ (long long) seed), pixelIndex, 0, &cs);
            RngState rng{cs.d, cs.v[0], cs.v[1], cs.v[2], cs.v[3], cs.v[4]};
            uint2 stack[RT_WIDE_STACK];
            const uint32_t ly = (uint32_t) (y - y0);
            std::vector<uint8_t> &tok = e.tok[((size_t) (ly / 4u) * tiles_x + x / 8u) * 32u + (ly & 3u) * 8u + (x & 7u)];
            for (uint32_t si = 0; si < cam.sqrtSamples; si++) {
                for (uint32_t sj = 0; sj < cam.sqrtSamples; sj++) {
                    RayD ray = primary_ray(cam, x, (uint32_t) y, si, sj, rng);
                    f3 thr = mk3(1.0f, 1.0f, 1.0f);
                    for (uint32_t depth = 0; depth < cam.depth; depth++) {
                        HitD hit;
                        // two_level: the general kernel's walk of a one-instance scene — a node step on the TLAS root and the instance
                        // entry (a LEAF step) in front of every ray (approximation: also for the few rays that miss the top box)
                        if (two_level) { tok.push_back(TOK_NODE); tok.push_back(TOK_LEAF); }
                        trace_single8_tokens(e.sc, ray, RT_TMIN, hit, stack, tok);
                        const bool alive = shade_scatter<true>(e.sc, cam, ray, hit, thr, rng);
                        const bool go_on = alive && depth + 1 < cam.depth;
                        tok.push_back(go_on ? TOK_SHADE_GO_ON : TOK_SHADE_PATH_END);
                        if (!alive) break;
                    }
                }
            }
            n += (long long) tok.size();
        }
        total += n;
    });
    return total.load();
}

// The warp votes of persist8_kernel replayed on those token streams: `n_warps` warps of 32 lanes step in turn; every lane carries
// one pixel, takes the next unstarted one from the shared queue (tile order) inside a SHADE step, and the warp runs ONE phase per
// iteration, chosen as the kernel does (kernels_persist.cu: node while >= node_min lanes stand on a node, else shade once
// >= shade_min lanes wait, else leaf, else node, else shade).
// variant 0: the kernel as it is.  variant 1: a LEAF step hands ALL pending primitive tests of the warp out to its 32 lanes
// (ceil(pending / 32) rounds).  variant 2: a lane tests all pending primitives of its own group in one LEAF step (the rejected
// RT_LEAF_LOOP variant, as a cross-check of the model).
// out[0..2] steps of node / leaf / shade phases, out[3..5] the lanes that took part, summed, out[6] leaf ROUNDS (variant 1 / 2: the
// issue cost of a leaf step is its rounds), out[7] lane-rounds of leaf work, out[8] iterations, out[9] tokens consumed.
int emul_simt_run(void *h, int n_warps, int node_min, int shade_min, int variant, uint64_t *out) {
    Emul &e = *(Emul *) h;
    if (e.tok.empty() || n_warps < 1) return -1;
    struct Lane { const std::vector<uint8_t> *t = nullptr; size_t i = 0; };
    std::vector<Lane> lanes((size_t) n_warps * 32);
    size_t next = 0;
    auto take = [&](Lane &l) {
        while (next < e.tok.size() && e.tok[next].empty()) next++;        // ids outside the image (edge tiles)
        if (next < e.tok.size()) { l.t = &e.tok[next++]; l.i = 0; } else l.t = nullptr;
    };
    for (Lane &l: lanes) take(l);
    for (int k = 0; k < 10; k++) out[k] = 0;
    size_t live = 0;
    for (const Lane &l: lanes) live += l.t != nullptr;
    while (live) {
        for (int w = 0; w < n_warps; w++) {
            Lane *L = &lanes[(size_t) w * 32];
            int nn = 0, nl = 0, ns = 0;
            for (int k = 0; k < 32; k++) {
                if (!L[k].t) continue;
                const uint8_t t = (*L[k].t)[L[k].i];
                if (t == TOK_NODE) nn++; else if (t == TOK_LEAF) nl++; else ns++;
            }
            if (nn + nl + ns == 0) continue;
            out[8]++;
            int phase;
            if (nn >= node_min) phase = 0;
            else if (ns >= shade_min) phase = 2;
            else if (nl) phase = 1;
            else if (nn) phase = 0;
            else phase = 2;
            if (phase == 0) {
                out[0]++; out[3] += (uint64_t) nn; out[9] += (uint64_t) nn;
                for (int k = 0; k < 32; k++) if (L[k].t && (*L[k].t)[L[k].i] == TOK_NODE) L[k].i++;
            } else if (phase == 1) {
                out[1]++; out[4] += (uint64_t) nl;
                if (variant == 0) {
                    out[6]++; out[7] += (uint64_t) nl; out[9] += (uint64_t) nl;
                    for (int k = 0; k < 32; k++) if (L[k].t && (*L[k].t)[L[k].i] == TOK_LEAF) L[k].i++;
                } else {
                    uint64_t pending = 0, longest = 0;
                    for (int k = 0; k < 32; k++) {
                        if (!L[k].t) continue;
                        uint64_t run = 0;
                        while ((*L[k].t)[L[k].i] == TOK_LEAF) { L[k].i++; run++; }
                        pending += run;
                        if (run > longest) longest = run;
                    }
                    out[9] += pending; out[7] += pending;
                    out[6] += variant == 1 ? (pending + 31u) / 32u : longest;
                }
            } else {
                out[2]++; out[5] += (uint64_t) ns; out[9] += (uint64_t) ns;
                for (int k = 0; k < 32; k++) {
                    if (!L[k].t || (*L[k].t)[L[k].i] < TOK_SHADE_GO_ON) continue;
                    L[k].i++;
                    if (L[k].i == L[k].t->size()) { take(L[k]); if (!L[k].t) live--; }
                }
            }
        }
    }
    return 0;
}

// rays traced by every pixel of rows [y0, y1) (row-major, width entries per row): the distribution of the sequential work a pixel
// is (profiles/tools/tail_evidence.py: critical path of a frame / of a shard)
void emul_pixel_rays(void *h, uint64_t seed, int y0, int y1, int threads, uint32_t *out) {
    Emul &e = *(Emul *) h;
    const CameraD &cam = e.hs.cam;
    parallel_rows(y0, y1, threads, [&](int y) {
        for (uint32_t x = 0; x < cam.width; x++) {
            const uint32_t pixelIndex = cam.pitch * (uint32_t) y + x;
            curandStateXORWOW_t st;
            curand_init((unsigned long long) ((long long) pixelIndex ^ (long long) seed), pixelIndex, 0, &st);
            RngState r{st.d, st.v[0], st.v[1], st.v[2], st.v[3], st.v[4]};
            uint32_t overflow = 0;
            uint2 stack[RT_WIDE_STACK];
            const PixelOut o = render_pixel<false, true>(e.sc, cam, x, (uint32_t) y, r, overflow, stack, 1);
            out[(size_t) (y - y0) * cam.width + x] = o.rays;
        }
    });
}

// debugging aid for parity work: every path segment of one pixel — rows of 12 floats: sample, depth, ray origin, ray
// direction, hit t (inf on a miss), instance, primitive ((type << 28) | index); returns the number of rows
int emul_pixel_paths(void *h, uint64_t seed, uint32_t x, uint32_t y, float *out, int cap) {
    Emul &e = *(Emul *) h;
    const CameraD &cam = e.hs.cam;
    const uint32_t pixelIndex = cam.pitch * y + x;
    curandStateXORWOW_t st;
    curand_init((unsigned long long) ((long long) pixelIndex ^ (long long) seed), pixelIndex, 0, &st);
    RngState rng{st.d, st.v[0], st.v[1], st.v[2], st.v[3], st.v[4]};
    uint2 stack[RT_WIDE_STACK];
    int rows = 0;
    for (uint32_t si = 0; si < cam.sqrtSamples; si++)
        for (uint32_t sj = 0; sj < cam.sqrtSamples; sj++) {
            RayD ray = primary_ray(cam, x, y, si, sj, rng);
            f3 thr = mk3(1.0f, 1.0f, 1.0f);
            for (uint32_t depth = 0; depth < cam.depth; depth++) {
                HitD hit;
                uint32_t overflow = 0;
                TravStats ts{0, 0, 0};
                trace_closest8<false>(e.sc, ray, RT_TMIN, hit, overflow, ts, stack, 1);
                if (rows < cap) {
                    float *o = out + 12 * rows;
                    o[0] = (float) (si * cam.sqrtSamples + sj); o[1] = (float) depth;
                    o[2] = ray.o.x; o[3] = ray.o.y; o[4] = ray.o.z; o[5] = ray.d.x; o[6] = ray.d.y; o[7] = ray.d.z;
                    o[8] = hit.inst >= 0 ? hit.t : INFINITY; o[9] = (float) hit.inst;
                    o[10] = hit.inst >= 0 ? (float) e.sc.leaf_prims8[ref_pos(hit.leaf)] : -1.0f; o[11] = hit.inst >= 0 ? (float) ref_type(hit.leaf) : -1.0f;
                }
                rows++;
                if (!shade_scatter<true>(e.sc, cam, ray, hit, thr, rng)) break;
            }
        }
    return rows;
}

void emul_instances(void *h, float *o) {
    Emul &e = *(Emul *) h;
    for (size_t i = 0; i < e.instances.size(); i++) {
        float *p = o + 33 * i;
        memcpy(p, e.instances[i].fwd, 48);
        memcpy(p + 12, e.instances[i].inv, 48);
        p[24] = e.wbox[i].lo.x; p[25] = e.wbox[i].lo.y; p[26] = e.wbox[i].lo.z;
        p[27] = e.wbox[i].hi.x; p[28] = e.wbox[i].hi.y; p[29] = e.wbox[i].hi.z;
        p[30] = e.wcentroid[i].x; p[31] = e.wcentroid[i].y; p[32] = e.wcentroid[i].z;
    }
}

void emul_camera(void *h, float *o) {
    const CameraD &c = ((Emul *) h)->hs.cam;
    const float v[20] = {c.pixelOrigin.x, c.pixelOrigin.y, c.pixelOrigin.z, c.pixelDx.x, c.pixelDx.y, c.pixelDx.z, c.pixelDy.x, c.pixelDy.y, c.pixelDy.z,
                         c.camU.x, c.camU.y, c.camU.z, c.camV.x, c.camV.y, c.camV.z, 0, 0, 0, c.rcpSqrtSamples, (float) c.sqrtSamples};
    memcpy(o, v, sizeof(v));
}

// LBVH arrays of the BLAS batch (which = 0) or the TLAS (which = 1); same layout as rt_read_lbvh
void emul_lbvh(void *h, int which, size_t *n_prims, uint64_t *keys, uint32_t *vals, int32_t *left, int32_t *right, int32_t *parent, float *bounds) {
    Emul &e = *(Emul *) h;
    const Lbvh &b = which == 0 ? e.blas : e.tlas;
    if (n_prims) *n_prims = (size_t) b.n;
    const size_t m = b.n > 1 ? b.n - 1 : 0;
    if (keys) memcpy(keys, b.keys.data(), (size_t) b.n * 8);
    if (vals) memcpy(vals, b.vals.data(), (size_t) b.n * 4);
    if (left && m) memcpy(left, b.left.data(), m * 4);
    if (right && m) memcpy(right, b.right.data(), m * 4);
    if (parent && m) memcpy(parent, b.parent.data(), m * 4);
    if (bounds) for (size_t i = 0; i < m; i++) {
        bounds[6 * i] = b.node_box[i].lo.x; bounds[6 * i + 1] = b.node_box[i].lo.y; bounds[6 * i + 2] = b.node_box[i].lo.z;
        bounds[6 * i + 3] = b.node_box[i].hi.x; bounds[6 * i + 4] = b.node_box[i].hi.y; bounds[6 * i + 5] = b.node_box[i].hi.z;
    }
}

// wide BVH of the BLAS batch (which = 0) or the TLAS (which = 1); same layout as rt_read_wide_bvh
void emul_wide(void *h, int which, size_t *n_nodes, size_t *n_leaf_prims, void *nodes, uint32_t *leaf_prims) {
    Emul &e = *(Emul *) h;
    const Lbvh &b = which == 0 ? e.blas : e.tlas;
    const std::vector<uint32_t> &lp = which == 0 ? e.leaf_prims8 : e.tlas_leaf_inst;
    if (n_nodes) *n_nodes = b.nodes8.size();
    if (n_leaf_prims) *n_leaf_prims = lp.size();
    if (nodes) memcpy(nodes, b.nodes8.data(), b.nodes8.size() * sizeof(BvhNode8));
    if (leaf_prims) memcpy(leaf_prims, lp.data(), lp.size() * 4);
}

// ---- config 3 input stage: the C++ reader + both conversions of one .vtk frame, no GPU -----------------------------
// mode 0: vtk_convert (rt_triangle array) -> HostScene::commit_geometry (the host path, RTB200_VTK_EXPAND=host)
// mode 1: vtk_convert_compact + expand_strip_triangle for every triangle (what expand_strips_kernel does on the device)
// Outputs (any may be NULL): tri_verts [9 per triangle], tri_shade [12 floats per triangle], instance descs, transforms.
int emul_vtk_frame(const char *path, int mode, uint64_t *n_tris, uint64_t *n_inst, float *tri_verts, float *tri_shade,
                   rt_instance_desc *inst_out, rt_instance_xform *xf_out) {
    VtkFrame vf;
    std::string err;
    if (!vf.read(path, err)) return -1;
    std::vector<rt_instance_desc> inst;
    std::vector<rt_instance_xform> xf;
    std::vector<float> verts;
    std::vector<TriShade> shade;
    if (mode == 0) {
        std::vector<rt_triangle> tris;
        vtk_convert(vf, tris, inst, xf);
        HostScene hs;
        if (!hs.commit_geometry(nullptr, 0, nullptr, 0, tris.data(), tris.size()).empty()) return -2;
        verts = hs.tri_verts;
        shade = hs.tri_shade;
    } else {
        VtkCompact vc;
        vtk_convert_compact(vf, vc, inst, xf);
        verts.resize(vc.n_tris * 9);
        shade.resize(vc.n_tris);
        for (uint32_t q = 0; q < (uint32_t) vc.n_tris; q++)
            expand_strip_triangle(q, &vc.points.data()->x, &vc.normals.data()->x, vc.ids.data(), vc.strip_first_tri.data(), vc.strip_first_id.data(),
                                  (int) vc.strip_first_id.size(), RT_METAL, 0u, verts.data(), shade.data());
    }
    if (n_tris) *n_tris = shade.size();
    if (n_inst) *n_inst = inst.size();
    if (tri_verts) memcpy(tri_verts, verts.data(), verts.size() * sizeof(float));
    if (tri_shade) memcpy(tri_shade, shade.data(), shade.size() * sizeof(TriShade));
    if (inst_out) memcpy(inst_out, inst.data(), inst.size() * sizeof(rt_instance_desc));
    if (xf_out) memcpy(xf_out, xf.data(), xf.size() * sizeof(rt_instance_xform));
    return 0;
}

// ---- output stage: the product's PNG encoder (csrc/image_io.h), no GPU ---------------------------------------------
// returns the file size; writes at most `cap` bytes to `out`
uint64_t emul_png_encode(const uint8_t *rgba, uint32_t w, uint32_t h, int bottom_up, uint8_t *out, uint64_t cap) {
    std::vector<uint8_t> png;
    png_encode_rgba8(rgba, w, h, bottom_up != 0, png);
    if (out) memcpy(out, png.data(), std::min<uint64_t>(cap, png.size()));
    return png.size();
}

// the .vtk.series index parser: names joined by '\n' into `names` (at most cap bytes), times[k]; returns the entry
// count, or -1 with the error text in `names`
int64_t emul_vtk_series(const char *path, char *names, uint64_t cap, float *times, uint64_t times_cap) {
    VtkSeries s;
    std::string err;
    std::string joined;
    if (!s.load(path, err)) {
        if (names && cap) { strncpy(names, err.c_str(), cap - 1); names[cap - 1] = 0; }
        return -1;
    }
    for (size_t k = 0; k < s.files.size(); k++) {
        joined += s.files[k].first;
        joined += '\n';
        if (times && k < times_cap) times[k] = s.files[k].second;
    }
    if (names && cap) { strncpy(names, joined.c_str(), cap - 1); names[cap - 1] = 0; }
    return (int64_t) s.files.size();
}

uint64_t emul_exr_encode(const float *plane, uint32_t w, uint32_t h, int n_channels, int bottom_up, uint8_t *out, uint64_t cap) {
    std::vector<uint8_t> exr;
    exr_encode_float(plane, w, h, n_channels, bottom_up != 0, exr);
    if (out) memcpy(out, exr.data(), std::min<uint64_t>(cap, exr.size()));
    return exr.size();
}

} // extern "C"
