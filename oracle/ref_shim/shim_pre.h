/*
 * shim_pre.h — forced-include (-include) header that lets the reference's .cu sources compile as
 * plain C++ with g++ (recipe: SURVEY.md Appendix A).  TEST INFRASTRUCTURE ONLY.
 * Nothing here re-implements reference arithmetic; it only supplies the CUDA built-ins the
 * sources expect (launch indices, clock64, surf2Dwrite, pinned allocation) as host stand-ins.
 */
#pragma once
#define QUALIFIERS static inline            /* makes the cuRAND device API host-callable */
#include <cuda_runtime.h>
#include <device_launch_parameters.h>
#include <curand_kernel.h>                  /* must precede the macros below */
#include <cstdint>
#include <cstdlib>
extern thread_local uint3 shim_threadIdx, shim_blockIdx;
extern thread_local dim3 shim_blockDim, shim_gridDim;
#define threadIdx shim_threadIdx
#define blockIdx  shim_blockIdx
#define blockDim  shim_blockDim
#define gridDim   shim_gridDim
extern thread_local long long g_clock64;    /* fixed frame seed replaces the clock (quirk Q8) */
static inline long long clock64() { return g_clock64; }
extern uchar4 *g_surface;
extern int g_surface_pitch_px;
template<class T> static inline void surf2Dwrite(T v, unsigned long long, int xbytes, int y) {
    g_surface[(size_t)y * g_surface_pitch_px + xbytes / (int)sizeof(T)] = v;
}
#define cudaHostAlloc(p, sz, flags) ((*(void**)(p) = malloc(sz)), cudaSuccess)
#define cudaFreeHost(p) (free((void*)(p)), cudaSuccess)
