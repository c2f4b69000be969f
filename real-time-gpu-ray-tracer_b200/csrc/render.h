// render.h — host-side launchers of the per-frame render kernels.
#pragma once

#include "device_util.h"
#include "rt_types.h"

// tables: RNG_TABLE_WORDS device words owned by the context (jump matrices per hex digit, filled on first use)
constexpr size_t RNG_HEX_TABLE_WORDS = 8 * 15 * 1280;                      // T[h][d]: 160 rows x 5 words, padded to 8 (128-bit loads)
constexpr size_t RNG_LOW_TABLE_WORDS = 160 * 5 * 32;                         // lane-interleaved table of the low 5 index bits
constexpr size_t RNG_WIDE_TABLE_WORDS = 2 * 255 * 1280;                       // byte-digit tables of index bits 5..12 and 13..20
constexpr size_t RNG_TABLE_WORDS = RNG_HEX_TABLE_WORDS + RNG_LOW_TABLE_WORDS + RNG_WIDE_TABLE_WORDS;
void launch_rng_init(RngState *states, unsigned long long seed, const CameraD &cam, const FrameD &fr, cudaStream_t stream, uint32_t *tables,
                     bool &tables_ready);
// builds the jump tables once (on `stream`; every later user must be ordered behind it)
void launch_rng_tables(uint32_t *tables, bool &tables_ready, cudaStream_t stream);
void launch_render(const SceneD &sc, const CameraD &cam, const FrameD &fr, cudaStream_t stream);
// measurement aid: read bandwidth of a buffer of `bytes` streamed `repeats` times (GB/s)
double probe_read_bandwidth(size_t bytes, int repeats, int sm_count, cudaStream_t stream);
