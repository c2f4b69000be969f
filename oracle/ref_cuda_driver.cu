/*
 * ref_cuda_driver.cu — BASELINE INFRASTRUCTURE ONLY (never part of librtb200.so, never on the product path).
 *
 * Headless driver for the reference's OWN CUDA build: oracle/Makefile target `refcuda` compiles the reference's
 * hot-path translation units UNMODIFIED from /root/reference/src with nvcc for sm_100a (-rdc=true, as its
 * CMakeLists.txt:16-18 does) and links them with this file into oracle/_ref/libref_cuda.so.  It replaces only
 * src/Global/Renderer.cu + Main.cu + SDL_OpenGLWindow.cu (SDL2 / OpenGL / VTK / json: unavailable on this box):
 * the scene is built with the reference's host builders (BLAS::constructBLAS, TLAS::constructTLAS,
 * Instance::updateTransformArguments), uploaded with the reference's RendererImpl::copy*GlobMem functions, and the
 * reference's `render` kernel (src/Global/Kernel.cu:105-147) is launched with its own launch geometry (16x16
 * blocks, SDL_OpenGLWindow.cu:119-122) on a plain cudaArray surface instead of a GL-interop one.
 * The BLAS de-duplication stores the BLAS index (quirk Q11 fixed, as the survey prescribes for the baseline copy).
 */
#include <Global/RendererImpl.cuh>

#define ORACLE_PREFIX refcuda_
#include "oracle_api.h"

#include <chrono>
#include <cstdio>
#include <cstring>
#include <map>
#include <vector>

extern "C" {
// SDL logging stand-ins (src/Global/Global.cu uses them for error reports only)
const char *SDL_GetError(void) { return "SDL unavailable"; }
void SDL_LogError(int, const char *fmt, ...) { fprintf(stderr, "[reference] %s\n", fmt); }
void SDL_Log(const char *fmt, ...) { (void) fmt; }
}

using namespace renderer;

namespace {
Point3 P(const rt_vec3 &v) { return {v.x, v.y, v.z}; }
Vec3 V(const rt_vec3 &v) { return {v.x, v.y, v.z}; }
double now_ms() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
#define CK(x) do { cudaError_t e__ = (x); if (e__ != cudaSuccess) { fprintf(stderr, "[ref_cuda] %s: %s\n", #x, cudaGetErrorString(e__)); return -200; } } while (0)
}

extern "C" {

/* Builds the scene with the reference's host code, renders `warmup + frames` frames with the reference's kernel and
 * returns per-frame kernel time (CUDA events) of the timed frames.  rgba8 (W*H*4, may be NULL) receives the last frame.
 * host_ms[0] = BLAS build, [1] = transforms, [2] = TLAS build (reference host builders, one thread).                */
int refcuda_bench(const oracle_scene *sc, int warmup, int frames, float *frame_ms, double *host_ms, uint8_t *rgba8) {
    std::vector<Sphere> spheres;
    std::vector<Parallelogram> quads;
    std::vector<Triangle> tris;
    std::vector<Rough> roughs;
    std::vector<Metal> metals;
    for (size_t i = 0; i < sc->sphere_count; i++) { const auto &g = sc->spheres[i]; spheres.emplace_back((MaterialType) g.material_type, (size_t) g.material_index, P(g.center), g.radius); }
    for (size_t i = 0; i < sc->parallelogram_count; i++) { const auto &g = sc->parallelograms[i]; quads.emplace_back((MaterialType) g.material_type, (size_t) g.material_index, P(g.q), V(g.u), V(g.v)); }
    tris.reserve(sc->triangle_count);
    for (size_t i = 0; i < sc->triangle_count; i++) {
        const auto &g = sc->triangles[i];
        const std::array<Point3, 3> vs = {P(g.v[0]), P(g.v[1]), P(g.v[2])};
        if (g.has_normals) { const std::array<Vec3, 3> ns = {V(g.n[0]), V(g.n[1]), V(g.n[2])}; tris.emplace_back((MaterialType) g.material_type, (size_t) g.material_index, vs, ns); }
        else tris.emplace_back((MaterialType) g.material_type, (size_t) g.material_index, vs);
    }
    for (size_t i = 0; i < sc->rough_count; i++) roughs.push_back(Rough{{sc->roughs[i].albedo.x, sc->roughs[i].albedo.y, sc->roughs[i].albedo.z}});
    for (size_t i = 0; i < sc->metal_count; i++) metals.push_back(Metal{{sc->metals[i].albedo.x, sc->metals[i].albedo.y, sc->metals[i].albedo.z}, sc->metals[i].fuzz});

    // instances: Renderer::configureInstances + the fill-in of buildBLASPinMem (Renderer.cu:100-109, RenderPin.cu:124-140)
    const size_t n_inst = sc->instance_count;
    std::vector<Instance> instances(n_inst);
    memset((void *) instances.data(), 0, n_inst * sizeof(Instance));
    for (size_t i = 0; i < n_inst; i++) {
        const auto &d = sc->instances[i];
        Instance &ins = instances[i];
        ins.primitiveType = (PrimitiveType) d.type;
        ins.primitiveIndex = d.first_prim;
        ins.primitiveCount = d.prim_count <= 1 ? 1 : d.prim_count;
        if (d.prim_count > 1 && d.has_bounds) {
            ins.boundingBox = BoundingBox(Range{d.bounds_min[0], d.bounds_max[0]}, Range{d.bounds_min[1], d.bounds_max[1]}, Range{d.bounds_min[2], d.bounds_max[2]});
            ins.centroid = Point3{d.centroid[0], d.centroid[1], d.centroid[2]};
        } else {
            BoundingBox box;
            for (size_t k = 0; k < ins.primitiveCount; k++) {
                BoundingBox b;
                switch (ins.primitiveType) {
                    case PrimitiveType::SPHERE: b = spheres[d.first_prim + k].constructBoundingBox(); break;
                    case PrimitiveType::PARALLELOGRAM: b = quads[d.first_prim + k].constructBoundingBox(); break;
                    case PrimitiveType::TRIANGLE: b = tris[d.first_prim + k].constructBoundingBox(); break;
                }
                box = k == 0 ? b : BoundingBox(box, b);
            }
            ins.boundingBox = box;
            if (ins.primitiveCount == 1) {
                switch (ins.primitiveType) {
                    case PrimitiveType::SPHERE: ins.centroid = spheres[d.first_prim].centroid(); break;
                    case PrimitiveType::PARALLELOGRAM: ins.centroid = quads[d.first_prim].centroid(); break;
                    case PrimitiveType::TRIANGLE: ins.centroid = tris[d.first_prim].centroid(); break;
                }
            } else {
                ins.centroid = Point3{0.5f * (box.range[0].min + box.range[0].max), 0.5f * (box.range[1].min + box.range[1].max), 0.5f * (box.range[2].min + box.range[2].max)};
            }
        }
    }
    // BLAS: one BLAS::constructBLAS per distinct (type, first primitive)
    std::map<std::pair<int, size_t>, size_t> blas_of;
    std::vector<BLASBuildResult> blas_results;
    double t0 = now_ms();
    for (size_t i = 0; i < n_inst; i++) {
        Instance &ins = instances[i];
        const auto key = std::make_pair((int) ins.primitiveType, ins.primitiveIndex);
        auto it = blas_of.find(key);
        if (it != blas_of.end()) { ins.asIndex = it->second; continue; }
        ins.asIndex = blas_results.size();
        blas_of[key] = ins.asIndex;
        switch (ins.primitiveType) {
            case PrimitiveType::SPHERE: blas_results.push_back(BLAS::constructBLAS(spheres.data(), ins.primitiveIndex, ins.primitiveCount, quads.data(), 0, 0, tris.data(), 0, 0)); break;
            case PrimitiveType::PARALLELOGRAM: blas_results.push_back(BLAS::constructBLAS(spheres.data(), 0, 0, quads.data(), ins.primitiveIndex, ins.primitiveCount, tris.data(), 0, 0)); break;
            case PrimitiveType::TRIANGLE: blas_results.push_back(BLAS::constructBLAS(spheres.data(), 0, 0, quads.data(), 0, 0, tris.data(), ins.primitiveIndex, ins.primitiveCount)); break;
        }
    }
    if (host_ms) host_ms[0] = now_ms() - t0;
    std::vector<BLASArray> blas_array(blas_results.size());
    for (size_t i = 0; i < blas_results.size(); i++) {
        blas_array[i].first = {blas_results[i].first.data(), blas_results[i].first.size()};
        blas_array[i].second = {blas_results[i].second.data(), blas_results[i].second.size()};
    }
    t0 = now_ms();
    for (size_t i = 0; i < n_inst; i++) {
        const auto &x = sc->xforms[i];
        instances[i].updateTransformArguments({x.shift.x, x.shift.y, x.shift.z}, {x.rotate_deg.x, x.rotate_deg.y, x.rotate_deg.z}, {x.scale.x, x.scale.y, x.scale.z});
    }
    if (host_ms) host_ms[1] = now_ms() - t0;
    t0 = now_ms();
    TLASBuildResult tlas = TLAS::constructTLAS(instances.data(), n_inst);
    if (host_ms) host_ms[2] = now_ms() - t0;

    Camera cam{};
    cam.windowWidth = (int) sc->window.width;
    cam.windowHeight = (int) sc->window.height;
    cam.backgroundColor = {sc->camera.background.x, sc->camera.background.y, sc->camera.background.z};
    cam.cameraCenter = P(sc->camera.center);
    cam.cameraTarget = P(sc->camera.target);
    cam.fov = sc->camera.fov;
    cam.upDirection = V(sc->camera.up);
    cam.focusDiskRadius = sc->camera.focus_disk_radius;
    cam.sampleRange = sc->camera.sample_range;
    cam.sampleCount = sc->camera.sample_count;
    cam.rayTraceDepth = sc->camera.ray_trace_depth;
    RendererImpl::calculateCameraProperties(&cam);

    // ---- upload with the reference's own copy functions (src/Global/RenderGlob.cu) ----
    cudaStream_t stream;
    CK(cudaStreamCreate(&stream));
    SceneGeometryData geo_pin{spheres.data(), spheres.size(), quads.data(), quads.size(), tris.data(), tris.size()};
    SceneMaterialData mat_pin{roughs.data(), roughs.size(), metals.data(), metals.size()};
    SceneGeometryData geo_dev = RendererImpl::allocGeoGlobMem(stream, geo_pin);
    RendererImpl::copyGeoToGlobMem(stream, geo_dev, geo_pin);
    SceneMaterialData mat_dev = RendererImpl::allocMatGlobMem(stream, mat_pin);
    RendererImpl::copyMatToGlobMem(stream, mat_dev, mat_pin);
    Instance *dev_inst = RendererImpl::allocInstGlobMem(stream, n_inst);
    RendererImpl::copyInstToGlobMem(stream, dev_inst, instances.data(), n_inst);
    Pair<BLASArray *, size_t> blas_pin{blas_array.data(), blas_array.size()};
    Pair<BLASArray *, size_t> blas_dev = RendererImpl::copyBLASToGlobMem(stream, blas_pin);
    TLASArray tlas_pin{{tlas.first.data(), tlas.first.size()}, {tlas.second.data(), tlas.second.size()}};
    TLASArray tlas_dev = RendererImpl::copyTLASGlobMem(stream, tlas_pin);
    RendererImpl::copyCamToConstMem(stream, &cam);
    const TraverseData td = RendererImpl::traverseDevPtr(geo_dev, mat_dev, dev_inst, blas_dev, tlas_dev);
    TraverseData *dev_td = nullptr;
    CK(cudaMalloc(&dev_td, sizeof(TraverseData)));
    CK(cudaMemcpyAsync((void *) dev_td, (const void *) &td, sizeof(TraverseData), cudaMemcpyHostToDevice, stream));

    // ---- headless surface (replaces the GL-interop texture of SDL_OpenGLWindow.cu:113-128,258-270) ----
    const int W = cam.windowWidth, H = cam.windowHeight;
    cudaArray_t arr;
    const cudaChannelFormatDesc desc = cudaCreateChannelDesc<uchar4>();
    CK(cudaMallocArray(&arr, &desc, W, H, cudaArraySurfaceLoadStore));
    cudaResourceDesc rd{};
    rd.resType = cudaResourceTypeArray;
    rd.res.array.array = arr;
    cudaSurfaceObject_t surf;
    CK(cudaCreateSurfaceObject(&surf, &rd));
    const dim3 threads(16, 16), blocks((W + 15) / 16, (H + 15) / 16);
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    for (int f = 0; f < warmup + frames; f++) {
        CK(cudaEventRecord(e0, stream));
        render<<<blocks, threads, 0, stream>>>(dev_td, surf);
        CK(cudaEventRecord(e1, stream));
        CK(cudaStreamSynchronize(stream));
        CK(cudaGetLastError());
        if (f >= warmup) CK(cudaEventElapsedTime(&frame_ms[f - warmup], e0, e1));
    }
    if (rgba8) CK(cudaMemcpy2DFromArray(rgba8, (size_t) W * 4, arr, 0, 0, (size_t) W * 4, H, cudaMemcpyDeviceToHost));
    cudaDestroySurfaceObject(surf);
    cudaFreeArray(arr);
    cudaFree(dev_td);
    RendererImpl::freeTLASGlobMem(stream, tlas_dev);
    RendererImpl::freeBLASGlobMem(stream, blas_dev);
    RendererImpl::freeInstGlobMem(stream, dev_inst);
    RendererImpl::freeMatGlobMem(stream, mat_dev);
    RendererImpl::freeGeoGlobMem(stream, geo_dev);
    cudaStreamSynchronize(stream);
    cudaStreamDestroy(stream);
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    return 0;
}

const char *refcuda_describe(void) { return "reference CUDA build (unmodified /root/reference sources, nvcc sm_100a, headless surface)"; }

}
