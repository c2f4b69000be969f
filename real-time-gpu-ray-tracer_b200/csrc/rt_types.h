// rt_types.h — HBM-resident record layouts (all 16-byte aligned, loaded as float4/int4) and kernel parameters.
#pragma once

#include "rt_math.h"

// ---- node references ----------------------------------------------------------------------------
// ref >= 0            : inner node index (TLAS node array in TLAS mode, BLAS node pool otherwise)
// ref  < 0            : leaf.  bits 30..29 = primitive type (BLAS) or 0 (TLAS), bits 28..0 = position
//                       (BLAS: position in the Morton-sorted leaf order; TLAS: instance index)
// RT_REF_SENTINEL     : "leave the instance" marker pushed when a TLAS leaf is entered
#define RT_REF_LEAF_BIT 0x80000000u
#define RT_REF_POS_MASK 0x1FFFFFFFu
#define RT_REF_SENTINEL ((int) 0xE0000000u)
#define RT_REF_NONE ((int) 0xFFFFFFFFu)   /* type bits 3 + all-ones position: empty child slot */
RT_HD int make_leaf_ref(uint32_t type, uint32_t pos) { return (int) (RT_REF_LEAF_BIT | (type << 29) | pos); }
RT_HD uint32_t ref_type(int ref) { return ((uint32_t) ref >> 29) & 3u; }
RT_HD uint32_t ref_pos(int ref) { return (uint32_t) ref & RT_REF_POS_MASK; }

#define RT_STACK_SIZE 64   /* reference uses 64-entry stacks without overflow check (Q15); we check */

// ---- binary node with both child boxes (64 B, four 16-byte loads) ----------------------------------
struct alignas(16) BvhNode2 {
    float4 xy0;   // child0: lo.x hi.x lo.y hi.y
    float4 xy1;   // child1: lo.x hi.x lo.y hi.y
    float4 z01;   // child0 lo.z hi.z, child1 lo.z hi.z
    int4 refs;    // ref0, ref1, -, -
};

// ---- compressed 8-wide node (80 B, five 16-byte loads) ------------------------------------------------
// Layout after Ylitie, Karras, Laine 2017 ("Efficient Incoherent Ray Traversal on GPUs Through Compressed Wide
// BVHs"): child boxes are 8-bit offsets on a per-node power-of-two grid anchored at the node's lower corner;
// children sit in octant-ordered slots so that traversal order is a XOR of the slot with the ray octant.
//   n0 : p.x p.y p.z | bytes {ex, ey, ez, imask}   (e* = biased float exponents of the grid step, imask bit s = slot s is an inner node)
//   n1 : child_base | prim_base | meta[0..3] | meta[4..7]
//        meta: 0 = empty; inner = 0b001_11sss (sss = slot); leaf = unary(prim count 1..3) << 5 | offset from prim_base (0..23)
//   n2 : qlo.x[0..3] qlo.x[4..7] qlo.y[0..3] qlo.y[4..7]
//   n3 : qlo.z[0..3] qlo.z[4..7] qhi.x[0..3] qhi.x[4..7]
//   n4 : qhi.y[0..3] qhi.y[4..7] qhi.z[0..3] qhi.z[4..7]
// Inner children of a node are consecutive nodes from child_base in slot order; its leaf primitives are consecutive
// leaf-order records from prim_base (TLAS: positions in the leaf->instance table).
struct alignas(16) BvhNode8 {
    float4 n0;
    uint4 n1, n2, n3, n4;
};
#define RT_WIDE_LEAF_MAX 1      /* primitives per leaf child (format allows 3; 1 measured fastest: the exact triangle test is dearer than a box) */
#ifndef RT_WIDE_STACK
#define RT_WIDE_STACK 24        /* entries (8 B each) of the per-thread traversal stack */
#endif

// ---- primitives -----------------------------------------------------------------------------------
// Triangle intersection record in LEAF ORDER (Morton-sorted), 48 B: v0|primIndex, e1, e2
// (e1, e2 precomputed at commit exactly as the reference's constructor, include/Geometry/Triangle.cuh:29-30,44-45)
struct alignas(16) TriIsect {
    float4 v0_id;   // v0.xyz, w = bits of the original triangle index
    float4 e1;      // e1.xyz, w unused
    float4 e2;      // e2.xyz, w unused
};
// Triangle shading record in ORIGINAL order, 48 B: vertex normals + material
struct alignas(16) TriShade {
    float4 n0_mt;   // n0.xyz, w = bits of material type
    float4 n1_mi;   // n1.xyz, w = bits of material index
    float4 n2;      // n2.xyz
};
struct alignas(16) SphereD {
    float4 c_r;     // center.xyz, radius
    uint32_t mat_type, mat_index, _p0, _p1;
};
struct alignas(16) QuadD {
    float4 q_d;     // q.xyz, plane constant d
    float4 u;       // u.xyz
    float4 v;       // v.xyz
    float4 n_flag;  // unit normal xyz, w = 1.0 if the reference's (wrong) box must be applied as a filter (Q10)
    float4 rb_lo;   // reference box lo.xyz (Parallelogram::constructBoundingBox after ensureVolume)
    float4 rb_hi;   // reference box hi.xyz
    uint32_t mat_type, mat_index, _p0, _p1;
};
struct alignas(16) InstanceD {
    float inv[12];  // rows 1..3 of transformInverse (world -> local)
    float fwd[12];  // rows 1..3 of transformMatrix  (local -> world)
    int root_ref;   // BLAS root (inner node index or leaf ref)
    int blas;
    int root8;      // BLAS root in the wide node pool
    int prim_type;  // rt_primitive_type of the BLAS (a BLAS holds one primitive type)
};

// ---- camera (the derived fields of the reference's Camera, include/Global/RendererImpl.cuh:32-61) -----
struct CameraD {
    f3 background, center, camU, camV;
    f3 pixelOrigin, pixelDx, pixelDy;
    float focusRadius;
    float rcpSqrtSamples;
    uint32_t sqrtSamples;
    uint32_t depth;
    uint32_t width, height, pitch;   // pitch = 16 * ceil(width / 16): the reference's padded pixelIndex stride
};

struct SceneD {
    const BvhNode2 *tlas_nodes;
    const BvhNode2 *blas_nodes;
    const InstanceD *instances;
    const TriIsect *leaf_tris;       // leaf order
    const uint32_t *leaf_prims;      // leaf order: original primitive index (all types)
    const TriShade *tri_shade;       // original order
    const SphereD *spheres;
    const QuadD *quads;
    const float4 *roughs;            // albedo.xyz
    const float4 *metals;            // albedo.xyz, fuzz
    const float4 *dielectrics;       // albedo.xyz, ior (extension)
    int tlas_root_ref;
    int _pad;
    // wide (BVH8) acceleration structure: the one the render kernels traverse
    const BvhNode8 *tlas_nodes8;     // root = node 0
    const BvhNode8 *blas_nodes8;     // roots = InstanceD::root8
    const uint32_t *tlas_leaf_inst;  // TLAS leaf order -> instance index
    const TriIsect *leaf_tris8;      // wide leaf order
    const uint32_t *leaf_prims8;     // wide leaf order: original primitive index (all types)
    const void *ext_prims;           // records of the compile-time extension primitive type (rt_prims.h), original order; else null
};

// XORWOW state (cuRAND curandStateXORWOW_t without the Box-Muller fields): 24 B per pixel
struct RngState { uint32_t d, v0, v1, v2, v3, v4; };

struct FrameD {
    uint8_t *rgba8;        // uchar4 per pixel (packed shard rows when sharded), or peer-mapped full frame
    float *radiance;       // 3 floats per pixel
    float *aov_t;          // primary-hit AOVs (may be null)
    int *aov_inst;
    int *aov_prim;
    const RngState *rng;   // per local pixel
    unsigned long long *ray_counter;   // [0] rays [1] overflow flag [2] node fetches [3] primitive fetches [4] instance fetches
    unsigned int *overflow_flag;
    uint32_t count_traversal;          // 1: run the instrumented variant (fills counters [2..4])
    // tile sharding: local row r -> global row
    uint32_t shard_rank, shard_world, tile_rows, local_rows;
    uint32_t rgba8_is_full_frame;   // 1: rgba8 indexed by GLOBAL pixel (peer store), 0: by local pixel
    uint32_t traverse_bvh2;         // 1: traverse the binary LBVH nodes (comparison mode) instead of the wide BVH
};
