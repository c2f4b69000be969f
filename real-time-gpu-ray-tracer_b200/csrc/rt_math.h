// rt_math.h — scalar/vector helpers shared by device kernels and the host-side C++ of librtb200.
//
// PARITY RULE: everything here is plain IEEE fp32 with the reference's operation order.  The CUDA
// translation units are compiled with -fmad=false, so `a * b + c` is two roundings exactly like the
// reference compiled for x86 (no contraction); the only fused operations are the explicit rt_fma()
// calls in the (non parity-critical) box tests.
#pragma once

#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define RT_HD __host__ __device__ __forceinline__
#define RT_D __device__ __forceinline__
#else
#define RT_HD inline
#define RT_D inline
#endif

#define RT_EPS 1e-6f            /* FLOAT_ZERO_VALUE, reference include/Global/Global.cuh:147 */
#define RT_TMIN 0.001f          /* rayColor's range.min, reference src/Global/Kernel.cu:66   */
#if defined(__CUDA_ARCH__)
#define RT_INF __int_as_float(0x7f800000)
#else
#define RT_INF INFINITY
#endif

struct f3 { float x, y, z; };

RT_HD f3 mk3(float x, float y, float z) { f3 r; r.x = x; r.y = y; r.z = z; return r; }
RT_HD f3 operator+(const f3 &a, const f3 &b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
RT_HD f3 operator-(const f3 &a, const f3 &b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
RT_HD f3 operator*(const f3 &a, float s) { return mk3(a.x * s, a.y * s, a.z * s); }
RT_HD f3 operator-(const f3 &a) { return mk3(-a.x, -a.y, -a.z); }
// Vec3::dot / lengthSquared (reference include/Basic/Vec3.cuh:103-123): running sum from 0.0f, x then y then z
RT_HD float dot3(const f3 &a, const f3 &b) { float s = a.x * b.x; s = s + a.y * b.y; s = s + a.z * b.z; return s; }
RT_HD float len2(const f3 &a) { return dot3(a, a); }
RT_HD f3 cross3(const f3 &a, const f3 &b) { return mk3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
// Vec3::unitize (Vec3.cuh:127-135): factor = 1 / sqrt(len2), then scale
RT_HD f3 unit3(const f3 &a) { const float f = 1.0f / sqrtf(len2(a)); return mk3(a.x * f, a.y * f, a.z * f); }
RT_HD float comp3(const f3 &a, int i) { return i == 0 ? a.x : (i == 1 ? a.y : a.z); }

RT_HD float rt_fma(float a, float b, float c) { return fmaf(a, b, c); }
RT_HD float rt_min(float a, float b) { return fminf(a, b); }
RT_HD float rt_max(float a, float b) { return fmaxf(a, b); }

RT_HD uint32_t f2u(float f) {
#if defined(__CUDA_ARCH__)
    return __float_as_uint(f);
#else
    union { float f; uint32_t u; } c; c.f = f; return c.u;
#endif
}
RT_HD float u2f(uint32_t u) {
#if defined(__CUDA_ARCH__)
    return __uint_as_float(u);
#else
    union { float f; uint32_t u; } c; c.u = u; return c.f;
#endif
}

// MathHelper::floatValueNearZero / floatValueEquals (reference include/Global/Global.cuh:172-182)
RT_HD bool near_zero(float v) { return fabsf(v) < RT_EPS; }
RT_HD bool feq(float a, float b) { return fabsf(a - b) < RT_EPS; }
// Range::inRange with both ends closed (reference include/Util/Range.cuh:33-43): ε-closed interval (quirk Q1)
RT_HD bool in_range(float v, float lo, float hi) {
    if (feq(v, lo)) return true;
    if (feq(v, hi)) return true;
    return v > lo && v < hi;
}

// affine 3x4 (rows of the reference's 4x4 Matrix, 1-indexed rows 1..3): p' = ((m0*x + m1*y) + m2*z) + m3
// — the same sums as Matrix::operator* on a w=1 column (reference src/Util/Matrix.cu:71-86; the leading
// 0.0f + and the trailing + m*0 terms do not change the value).
RT_HD f3 xf_point(const float *m, const f3 &p) {
    return mk3(((m[0] * p.x + m[1] * p.y) + m[2] * p.z) + m[3],
               ((m[4] * p.x + m[5] * p.y) + m[6] * p.z) + m[7],
               ((m[8] * p.x + m[9] * p.y) + m[10] * p.z) + m[11]);
}
RT_HD f3 xf_vector(const float *m, const f3 &v) {
    return mk3((m[0] * v.x + m[1] * v.y) + m[2] * v.z,
               (m[4] * v.x + m[5] * v.y) + m[6] * v.z,
               (m[8] * v.x + m[9] * v.y) + m[10] * v.z);
}
// (M^-1)^T applied to a vector: rows of the transpose are columns of the inverse
RT_HD f3 xf_normal(const float *inv, const f3 &n) {
    return mk3((inv[0] * n.x + inv[4] * n.y) + inv[8] * n.z,
               (inv[1] * n.x + inv[5] * n.y) + inv[9] * n.z,
               (inv[2] * n.x + inv[6] * n.y) + inv[10] * n.z);
}
