// rt_pixel.h — one pixel of the reference's `render` kernel (all samples, all bounces) as a single serial function.
// Used by the megakernel variant of the product and by the g++ twin in tests/emul; the wavefront kernels
// (kernels_wave.cu) run the same steps split across kernels.
#pragma once

#include "rt_wide_trace.h"

struct PixelOut { f3 radiance; uint32_t rays; float aov_t; int aov_inst, aov_prim; TravStats stats; };

// One pixel of the reference's `render` kernel (src/Global/Kernel.cu:116-143): all samples, all bounces.
// WIDE selects the acceleration structure: the compressed 8-wide BVH (product default) or the binary LBVH nodes.
template<bool COUNT, bool WIDE>
RT_HD PixelOut render_pixel(const SceneD &sc, const CameraD &cam, uint32_t x, uint32_t y, RngState rng, uint32_t &overflow,
                            uint2 *stack_mem, int stack_stride) {
    PixelOut out;
    out.stats.nodes = 0; out.stats.prims = 0; out.stats.insts = 0;
    out.radiance = mk3(0.0f, 0.0f, 0.0f);
    out.rays = 0; out.aov_t = RT_INF; out.aov_inst = -1; out.aov_prim = -1;
    for (uint32_t si = 0; si < cam.sqrtSamples; si++) {
        for (uint32_t sj = 0; sj < cam.sqrtSamples; sj++) {
            RayD ray = primary_ray(cam, x, y, si, sj, rng);
            f3 thr = mk3(1.0f, 1.0f, 1.0f);
            for (uint32_t depth = 0; depth < cam.depth; depth++) {
                HitD hit;
                if (WIDE) trace_closest8<COUNT>(sc, ray, RT_TMIN, hit, overflow, out.stats, stack_mem, stack_stride);
                else trace_closest<COUNT>(sc, ray, RT_TMIN, hit, overflow, out.stats);
                out.rays++;
                if (si == 0 && sj == 0 && depth == 0 && hit.inst >= 0) {
                    out.aov_t = hit.t; out.aov_inst = hit.inst;
                    out.aov_prim = (int) ((ref_type(hit.leaf) << 28) | (WIDE ? sc.leaf_prims8[ref_pos(hit.leaf)] : sc.leaf_prims[ref_pos(hit.leaf)]));
                }
                if (!shade_scatter<WIDE>(sc, cam, ray, hit, thr, rng)) break;
            }
            out.radiance = out.radiance + thr;
        }
    }
    const float w = cam.rcpSqrtSamples * cam.rcpSqrtSamples;
    out.radiance = out.radiance * w;
    return out;
}
