// kernels_render.cu — per-frame render kernels: per-pixel XORWOW initialisation and the path-tracing kernel.
#include "render.h"

#include <curand_kernel.h>

#include "rt_pixel.h"

namespace {

// curand_init(pixelIndex ^ seed, pixelIndex, 0) (reference src/Global/Kernel.cu:114 with clock64() := seed).
// Kept in its own coherent kernel: the sequence skip-ahead is ~10^4 table reads per pixel and would
// otherwise sit divergently inside the trace kernel.
__global__ void __launch_bounds__(256) rng_init_kernel(RngState *__restrict__ states, unsigned long long seed, uint32_t width, uint32_t pitch,
                                                       uint32_t local_rows, uint32_t rank, uint32_t world, uint32_t tile_rows) {
    const uint32_t lp = blockIdx.x * blockDim.x + threadIdx.x;
    if (lp >= width * local_rows) return;
    const uint32_t lx = lp % width, ly = lp / width;
    const uint32_t gy = ((ly / tile_rows) * world + rank) * tile_rows + (ly % tile_rows);
    const uint32_t pixelIndex = pitch * gy + lx;
    curandStateXORWOW_t st;
    curand_init((unsigned long long) ((long long) pixelIndex ^ (long long) seed), pixelIndex, 0, &st);
    RngState r;
    r.d = st.d; r.v0 = st.v[0]; r.v1 = st.v[1]; r.v2 = st.v[2]; r.v3 = st.v[3]; r.v4 = st.v[4];
    states[lp] = r;
}

// v1: one thread per pixel, 8x4-pixel warps, all samples and bounces of the pixel in the thread
template<bool COUNT, bool WIDE>
__global__ void __launch_bounds__(128) render_pixels_kernel(const __grid_constant__ SceneD sc, const __grid_constant__ CameraD cam,
                                                            const __grid_constant__ FrameD fr) {
    // block = 4 warps, each warp an 8x4 pixel tile; blocks tile the (local) image in 16x8 pixel groups
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t tiles_x = (cam.width + 15) / 16;
    const uint32_t bx = blockIdx.x % tiles_x, by = blockIdx.x / tiles_x;
    const uint32_t lx = bx * 16 + (warp & 1) * 8 + (lane & 7);
    const uint32_t ly = by * 8 + (warp >> 1) * 4 + (lane >> 3);
    uint32_t rays = 0, overflow = 0;
    TravStats stats{0, 0, 0};
    // per-thread traversal stacks interleaved in shared memory: entry k of thread t at [k * 128 + t]
    __shared__ uint2 stack_mem[WIDE ? RT_WIDE_STACK * 128 : 1];
    if (lx < cam.width && ly < fr.local_rows) {
        const uint32_t gy = ((ly / fr.tile_rows) * fr.shard_world + fr.shard_rank) * fr.tile_rows + (ly % fr.tile_rows);
        if (gy < cam.height) {
            const uint32_t lp = ly * cam.width + lx;
            const PixelOut o = render_pixel<COUNT, WIDE>(sc, cam, lx, gy, fr.rng[lp], overflow, stack_mem + (WIDE ? threadIdx.x : 0), 128);
            rays = o.rays;
            stats = o.stats;
            fr.radiance[3 * (size_t) lp + 0] = o.radiance.x;
            fr.radiance[3 * (size_t) lp + 1] = o.radiance.y;
            fr.radiance[3 * (size_t) lp + 2] = o.radiance.z;
            const uchar4 c = make_uchar4(to_byte(o.radiance.x), to_byte(o.radiance.y), to_byte(o.radiance.z), 255);
            const size_t op = fr.rgba8_is_full_frame ? (size_t) gy * cam.width + lx : (size_t) lp;
            reinterpret_cast<uchar4 *>(fr.rgba8)[op] = c;
            if (fr.aov_t) { fr.aov_t[lp] = o.aov_t; fr.aov_inst[lp] = o.aov_inst; fr.aov_prim[lp] = o.aov_prim; }
        }
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        rays += __shfl_xor_sync(0xffffffffu, rays, off);
        overflow |= __shfl_xor_sync(0xffffffffu, overflow, off);
    }
    if (COUNT) {
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            stats.nodes += __shfl_xor_sync(0xffffffffu, stats.nodes, off);
            stats.prims += __shfl_xor_sync(0xffffffffu, stats.prims, off);
            stats.insts += __shfl_xor_sync(0xffffffffu, stats.insts, off);
        }
    }
    if (lane == 0) {
        if (rays) atomicAdd(fr.ray_counter, (unsigned long long) rays);
        if (overflow) atomicOr(fr.overflow_flag, 1u);
        if (COUNT) {
            atomicAdd(fr.ray_counter + 2, (unsigned long long) stats.nodes);
            atomicAdd(fr.ray_counter + 3, (unsigned long long) stats.prims);
            atomicAdd(fr.ray_counter + 4, (unsigned long long) stats.insts);
        }
    }
}

// measurement aid: every thread streams 128-bit words of the buffer (grid-stride) and folds them into one word
__global__ void __launch_bounds__(512) probe_read_kernel(const uint4 *__restrict__ buf, size_t n_words, int repeats, uint32_t *sink) {
    uint32_t acc = 0;
    const size_t stride = (size_t) gridDim.x * blockDim.x;
    for (int r = 0; r < repeats; r++) {
        for (size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x; i < n_words; i += 4 * stride) {
            uint4 a = make_uint4(0, 0, 0, 0), b = a, c = a, d = a;
            a = __ldcg(buf + i);
            if (i + stride < n_words) b = __ldcg(buf + i + stride);
            if (i + 2 * stride < n_words) c = __ldcg(buf + i + 2 * stride);
            if (i + 3 * stride < n_words) d = __ldcg(buf + i + 3 * stride);
            acc ^= a.x ^ a.y ^ a.z ^ a.w ^ b.x ^ b.y ^ b.z ^ b.w ^ c.x ^ c.y ^ c.z ^ c.w ^ d.x ^ d.y ^ d.z ^ d.w;
        }
    }
    if (acc == 0x9E3779B9u) *sink = acc;       // never true for the zero-filled buffer; keeps the loads alive
}

} // namespace

double probe_read_bandwidth(size_t bytes, int repeats, int sm_count, cudaStream_t stream) {
    DeviceBuffer<uint4> buf;
    DeviceBuffer<uint32_t> sink;
    const size_t n_words = bytes / sizeof(uint4);
    buf.reserve(n_words);
    sink.reserve(1);
    CUDA_CHECK(cudaMemsetAsync(buf.get(), 0, n_words * sizeof(uint4), stream));
    cudaEvent_t e0, e1;
    CUDA_CHECK(cudaEventCreate(&e0));
    CUDA_CHECK(cudaEventCreate(&e1));
    probe_read_kernel<<<sm_count * 4, 512, 0, stream>>>(buf.get(), n_words, 2, sink.get());       // warm-up (fills L2)
    CUDA_CHECK(cudaEventRecord(e0, stream));
    probe_read_kernel<<<sm_count * 4, 512, 0, stream>>>(buf.get(), n_words, repeats, sink.get());
    CUDA_CHECK(cudaEventRecord(e1, stream));
    CUDA_CHECK(cudaEventSynchronize(e1));
    float ms = 0.f;
    CUDA_CHECK(cudaEventElapsedTime(&ms, e0, e1));
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    return (double) n_words * sizeof(uint4) * repeats / (ms * 1e-3) / 1e9;
}

void launch_rng_init(RngState *states, unsigned long long seed, const CameraD &cam, const FrameD &fr, cudaStream_t stream) {
    const size_t n = (size_t) cam.width * fr.local_rows;
    rng_init_kernel<<<div_up(n, 256), 256, 0, stream>>>(states, seed, cam.width, cam.pitch, fr.local_rows, fr.shard_rank, fr.shard_world, fr.tile_rows);
    CUDA_CHECK(cudaGetLastError());
}

void launch_render(const SceneD &sc, const CameraD &cam, const FrameD &fr, cudaStream_t stream) {
    const uint32_t tiles_x = (cam.width + 15) / 16, tiles_y = (fr.local_rows + 7) / 8;
    const unsigned grid = tiles_x * tiles_y;
    if (fr.traverse_bvh2) {
        if (fr.count_traversal) render_pixels_kernel<true, false><<<grid, 128, 0, stream>>>(sc, cam, fr);
        else render_pixels_kernel<false, false><<<grid, 128, 0, stream>>>(sc, cam, fr);
    } else {
        if (fr.count_traversal) render_pixels_kernel<true, true><<<grid, 128, 0, stream>>>(sc, cam, fr);
        else render_pixels_kernel<false, true><<<grid, 128, 0, stream>>>(sc, cam, fr);
    }
    CUDA_CHECK(cudaGetLastError());
}
