/*
 * oracle_api.h — TEST INFRASTRUCTURE ONLY.
 *
 * Common C interface of the two CPU checkers:
 *   - oracle/_ref/libref_oracle.so : the reference's OWN sources (/root/reference/src/...)
 *     compiled as plain C++ (oracle/Makefile, recipe SURVEY.md Appendix A) behind ref_driver.cpp;
 *   - oracle/liboracle_port.so     : oracle_port.cpp, a from-scratch CPU restatement of the same
 *     algorithm (every function cites the reference file:line it follows).
 * Both export the same symbols with prefix `ref_` / `port_` (macro ORACLE_FN).
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load these libraries.  The product (librtb200.so) never links, loads or calls them.
 */
#ifndef ORACLE_API_H
#define ORACLE_API_H

#include "../include/rtb200.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct oracle_scene {
    const rt_sphere *spheres;               size_t sphere_count;
    const rt_parallelogram *parallelograms; size_t parallelogram_count;
    const rt_triangle *triangles;           size_t triangle_count;
    const rt_rough *roughs;                 size_t rough_count;
    const rt_metal *metals;                 size_t metal_count;
    const rt_instance_desc *instances;
    const rt_instance_xform *xforms;        /* one per instance */
    size_t instance_count;
    rt_window window;
    rt_camera_input camera;
    uint64_t build_seed;                    /* seeds the (otherwise unseeded) axis RNG of the BVH build, Q13 */
    /* EXTENSION, parity unpinned (no reference counterpart): used by the port only; the compiled reference ignores it */
    const rt_dielectric *dielectrics;       size_t dielectric_count;
    /* records of a compile-time extension primitive type (value 3 in rt_instance_desc.type; csrc/rt_prims.h): only the
     * test builds that know the type read them (tests/ext/) */
    const void *ext_records;                size_t ext_record_bytes, ext_count;
} oracle_scene;

typedef struct oracle_timing {
    double blas_build_ms;   /* sum over BLAS::constructBLAS calls       */
    double xform_ms;        /* Instance::updateTransformArguments loop  */
    double tlas_build_ms;   /* TLAS::constructTLAS                      */
    uint64_t blas_nodes, tlas_nodes;
} oracle_timing;

#ifndef ORACLE_PREFIX
#define ORACLE_PREFIX port_
#endif
#define ORACLE_CAT2(a, b) a##b
#define ORACLE_CAT(a, b) ORACLE_CAT2(a, b)
#define ORACLE_FN(name) ORACLE_CAT(ORACLE_PREFIX, name)

/* Build BLAS (one per distinct (type, first_prim)), instance transforms, TLAS, camera — in the
 * steady-state order BLAS -> transforms -> TLAS (avoids quirk Q12).  Returns an opaque handle.    */
void *ORACLE_FN(create)(const oracle_scene *scene, oracle_timing *timing);
void  ORACLE_FN(destroy)(void *handle);

/* Render rows [y0, y1) with `threads` host threads.  Per pixel: curand_init(seed ^ pixelIndex,
 * pixelIndex, 0) with pixelIndex = pitch*y + x, pitch = 16*ceil(W/16) (src/Global/Kernel.cu:107-114).
 * Outputs are full-frame W*H arrays (rows outside [y0,y1) untouched); any may be NULL.
 *   rgba8    : 4 bytes/pixel, castToUchar4 (include/Basic/Color3.cuh:99-114)
 *   radiance : 3 floats/pixel, mean over samples (before gamma)
 *   rays     : *rays += number of TLAS::hit calls
 * Row 0 is the bottom scan-line.                                                                   */
void ORACLE_FN(render)(void *handle, uint64_t seed, int y0, int y1, int threads,
                       uint8_t *rgba8, float *radiance, uint64_t *rays);

/* Primary hit of each pixel's FIRST sample (same RNG stream as render): t (+inf on miss),
 * instance index (-1), primitive (type << 28 | index, -1).                                          */
void ORACLE_FN(primary)(void *handle, uint64_t seed, int y0, int y1, int threads,
                        float *t, int32_t *instance, int32_t *prim);

/* Closest hit of arbitrary world rays (n rays, 6 floats each: origin, direction) over
 * range [tmin, +inf): unit test of TLAS::hit.  normal: 3 floats per ray (world, unit) or NULL.      */
void ORACLE_FN(trace)(void *handle, const float *rays, size_t n, float tmin,
                      float *t, int32_t *instance, int32_t *prim, float *normal);

/* Derived camera (the 224-byte reference Camera) as floats, for KATs:
 * out[0..2]=pixelOrigin out[3..5]=viewPortPixelDx out[6..8]=viewPortPixelDy out[9..11]=cameraU
 * out[12..14]=cameraV out[15..17]=cameraW out[18]=reciprocalSqrtSampleCount out[19]=sqrtSampleCount */
void ORACLE_FN(camera)(void *handle, float *out20);

/* Instance matrices for KATs: per instance 12 floats M rows 1..3 (4 cols), 12 floats inverse,
 * 6 floats world box (min xyz, max xyz), 3 floats world centroid  = 33 floats per instance.        */
void ORACLE_FN(instances)(void *handle, float *out33_per_instance);

/* The first `n` curand_uniform draws of curand_init(seed, subsequence, 0).                           */
void ORACLE_FN(uniforms)(uint64_t seed, uint64_t subsequence, size_t n, float *out);

const char *ORACLE_FN(describe)(void);

#ifdef __cplusplus
}
#endif
#endif
