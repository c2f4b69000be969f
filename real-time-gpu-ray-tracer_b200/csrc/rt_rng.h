// rt_rng.h — per-pixel XORWOW stream, bit-identical to cuRAND's curandStateXORWOW_t as the reference uses it:
//   curand_init(pixelIndex ^ seed, pixelIndex, 0, &state)      reference src/Global/Kernel.cu:114
//   RandomGenerator::randomDouble(state) = curand_uniform      reference include/Global/Global.cuh:198-200
// The generator step and the uint -> (0,1] conversion are restated from the toolkit headers
// (curand_kernel.h `curand(curandStateXORWOW_t*)`, curand_uniform.h `_curand_uniform`) so that they are
// evaluated WITHOUT fma contraction, like the reference compiled for the host.  Initialisation (seed
// scrambling + sequence skip-ahead over cuRAND's precomputed matrices) lives in kernels_render.cu.
#pragma once

#include "rt_types.h"

RT_HD uint32_t xorwow_next(RngState &s) {
    const uint32_t t = s.v0 ^ (s.v0 >> 2);
    s.v0 = s.v1; s.v1 = s.v2; s.v2 = s.v3; s.v3 = s.v4;
    s.v4 = (s.v4 ^ (s.v4 << 4)) ^ (t ^ (t << 1));
    s.d += 362437u;
    return s.v4 + s.d;
}

// (0, 1]: x * 2^-32 + 2^-33, two roundings
RT_HD float rng_uniform(RngState &s) {
    const float x = (float) xorwow_next(s);
    const float m = x * 2.3283064e-10f;
    return m + (2.3283064e-10f / 2.0f);
}
// RandomGenerator::randomDouble(state, min, max) = min + (max - min) * u   (Global.cuh:208-210)
RT_HD float rng_range(RngState &s, float lo, float hi) { return lo + (hi - lo) * rng_uniform(s); }

// Vec3::randomSpaceVector(state, 1.0f): normalised CUBE sample (reference src/Basic/Vec3.cu:54-65, quirk Q5)
RT_HD f3 rng_space_vector(RngState &s) {
    f3 v;
    float l2;
    do {
        v.x = rng_range(s, -1.0f, 1.0f);
        v.y = rng_range(s, -1.0f, 1.0f);
        v.z = rng_range(s, -1.0f, 1.0f);
        l2 = len2(v);
    } while (l2 < RT_EPS * RT_EPS);
    v = unit3(v);
    return v * 1.0f;
}
// Vec3::randomPlaneVector(state, maxLength) (reference src/Basic/Vec3.cu:29-36)
RT_HD void rng_plane_vector(RngState &s, float maxLength, float &x, float &y) {
    do {
        x = rng_range(s, -1.0f, 1.0f);
        y = rng_range(s, -1.0f, 1.0f);
    } while (x * x + y * y > maxLength * maxLength);
}
