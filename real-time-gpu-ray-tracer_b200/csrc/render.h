// render.h — host-side launchers of the per-frame render kernels.
#pragma once

#include "device_util.h"
#include "rt_types.h"

void launch_rng_init(RngState *states, unsigned long long seed, const CameraD &cam, const FrameD &fr, cudaStream_t stream);
void launch_render(const SceneD &sc, const CameraD &cam, const FrameD &fr, cudaStream_t stream);
// measurement aid: read bandwidth of a buffer of `bytes` streamed `repeats` times (GB/s)
double probe_read_bandwidth(size_t bytes, int repeats, int sm_count, cudaStream_t stream);
