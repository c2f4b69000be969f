// host_scene.h — host-side (C++) half of the drop-in: validates and converts the committed scene into the
// device record layouts, keeps what the per-frame loop needs, and restates the reference's host arithmetic
// (primitive constructors, instance local bounds, camera) with the same operation order.  No CUDA calls here,
// so the CPU emulation used by the `not gpu` tests shares this exact code with librtb200.
#pragma once

#include "../../include/rtb200.h"
#include "rt_prims.h"

#include <string>
#include <vector>

struct HostBox { float lo[3], hi[3]; };

struct HostScene {
    uint32_t flags = 0;
    // committed inputs
    std::vector<rt_sphere> spheres;
    std::vector<rt_parallelogram> quads;
    std::vector<float> tri_verts;                 // 9 per triangle, original order
    size_t n_tris = 0;
    // device-layout records (original order)
    std::vector<SphereD> sphere_rec;
    std::vector<QuadD> quad_rec;
    std::vector<TriShade> tri_shade;
    std::vector<float4> roughs, metals, dielectrics;
    // records of the compile-time extension primitive type (rt_prims.h RT_EXT_PRIMITIVE), raw ExtPrimD bytes; empty otherwise
    std::vector<unsigned char> ext_rec;
    size_t n_ext = 0;
    std::string commit_extension(const void *records, size_t record_bytes, size_t count);
    size_t prim_count_of(uint32_t type) const;    // committed primitives of a type (0 for an unknown type)
    // highest material index the committed geometry uses, per rt_material_type (-1: type unused); checked against the
    // committed material counts before a frame is rendered (check_materials)
    long long max_mat_index[3] = {-1, -1, -1};
    void note_material(uint32_t type, uint32_t index) { if (type < 3u && (long long) index > max_mat_index[type]) max_mat_index[type] = index; }
    std::string check_materials() const;
    // instances
    std::vector<rt_instance_desc> inst_desc;
    std::vector<rt_instance_xform> xforms;
    std::vector<HostBox> inst_box;                // local boxes, reference semantics
    std::vector<f3> inst_centroid;
    std::vector<int> inst_blas;
    // BLAS batch: one segment per distinct primitive range
    std::vector<uint32_t> seg_offset, seg_type, seg_first;
    // camera
    rt_window window{};
    CameraD cam{};

    // each returns an empty string on success, else the error text (-> RT_ERR_ARGUMENT)
    std::string commit_geometry(const rt_sphere *s, size_t ns, const rt_parallelogram *p, size_t np, const rt_triangle *t, size_t nt);
    // animation frames whose first n_head triangles are produced on the device (strip expansion): only the caller's own
    // triangles, which follow them, are converted here (tri_verts / tri_shade keep stale values in the head)
    std::string commit_triangle_tail(size_t n_head, const rt_triangle *t, size_t nt);
    std::string commit_materials(const rt_rough *r, size_t nr, const rt_metal *m, size_t nm);
    std::string commit_dielectrics(const rt_dielectric *d, size_t n);
    std::string configure_instances(const rt_instance_desc *d, size_t n, const rt_instance_xform *x);
    std::string plan_blas();                      // fills seg_*, inst_blas, inst_box, inst_centroid
    std::string configure_camera(const rt_window &w, const rt_camera_input &c);
    // the derived camera of (window, input) without touching this scene's own (per-frame cameras of a window of frames in flight)
    static std::string make_camera(const rt_window &w, const rt_camera_input &c, CameraD &out);
    // per frame: transform arguments -> kernel input records (trig evaluated here with the host libm)
    void fill_xform_in(InstanceXformIn *out, const std::vector<int> &seg_root) const;
    size_t prim_bytes() const;
    void convert_triangles(const rt_triangle *t, size_t nt, size_t at);      // rt_triangle[k] -> tri_verts / tri_shade[at + k]
};

int morton_bits_for(size_t n);
int tlas_morton_bits(size_t n_instances);
// SDL_OpenGLWindow::calculateNewPosition restated (reference src/Global/SDL_OpenGLWindow.cu:182-256); true if the camera moved
bool camera_step(const rt_camera_input &cur, const rt_camera_motion &m, rt_camera_input &next);
