// render.h — host-side launchers of the per-frame render kernels.
#pragma once

#include "device_util.h"
#include "rt_types.h"

void launch_rng_init(RngState *states, unsigned long long seed, const CameraD &cam, const FrameD &fr, cudaStream_t stream);
void launch_render(const SceneD &sc, const CameraD &cam, const FrameD &fr, cudaStream_t stream);
