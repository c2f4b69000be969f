// kernels_render.cu — per-frame render kernels: per-pixel XORWOW initialisation and the path-tracing kernel.
#include "render.h"

#include <curand_kernel.h>

#include <cstdlib>
#include <cstring>

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

// ---- the same states, 2.7x faster ---------------------------------------------------------------------------------
// curand_init's cost is skipahead_sequence(pixelIndex): the jump by pixelIndex * 2^67 draws is a GF(2)-linear map of the
// 160-bit state, applied by cuRAND as one 160x160 matrix-vector product per unit of every base-4 digit of pixelIndex
// (precalc_xorwow_matrix[k] = one-sequence jump ^ (4^k); ~16 products for a 21-bit index, each with a data-dependent
// branch per bit).  Here the products of those matrices are tabulated once per context for every value of every
// HEX digit (T[h][d] = M[2h]^(d & 3) * M[2h+1]^(d >> 2), 8 x 15 matrices, 384 KB), so a pixel needs one branch-free
// product per non-zero hex digit (~5).  Powers of one matrix commute, so the result is the same state bit for bit —
// the parity tests against the oracle (host curand_init) cover it; RTB200_RNG_INIT=curand selects the kernel above.
constexpr int RNG_HEX_DIGITS = 8, RNG_ROWS = 160, RNG_ROW_WORDS = 8, RNG_MAT_WORDS = RNG_ROWS * RNG_ROW_WORDS;

// v <- M v over GF(2); row r of M = image of basis vector e_r, 5 words.  STRIDE 5: cuRAND's own precalc matrices;
// STRIDE 8: the tables built here, rows padded to 32 bytes so that a row is one 128-bit load + one word.
template<int STRIDE>
__device__ __forceinline__ void xorwow_matvec_s(const uint32_t *__restrict__ m, uint32_t v[5]) {
    uint32_t r0 = 0, r1 = 0, r2 = 0, r3 = 0, r4 = 0;
#pragma unroll
    for (int i = 0; i < 5; i++) {
        const uint32_t w = v[i];
#pragma unroll 8
        for (int j = 0; j < 32; j++) {
            const uint32_t mask = 0u - ((w >> j) & 1u);
            const uint32_t *row = m + STRIDE * (32 * i + j);
            if (STRIDE == 8) {
                const uint4 a = __ldg(reinterpret_cast<const uint4 *>(row));
                r0 ^= a.x & mask; r1 ^= a.y & mask; r2 ^= a.z & mask; r3 ^= a.w & mask; r4 ^= __ldg(row + 4) & mask;
            } else {
                r0 ^= __ldg(row) & mask; r1 ^= __ldg(row + 1) & mask; r2 ^= __ldg(row + 2) & mask; r3 ^= __ldg(row + 3) & mask; r4 ^= __ldg(row + 4) & mask;
            }
        }
    }
    v[0] = r0; v[1] = r1; v[2] = r2; v[3] = r3; v[4] = r4;
}
__device__ __forceinline__ void xorwow_matvec(const uint32_t *__restrict__ m, uint32_t v[5]) { xorwow_matvec_s<RNG_ROW_WORDS>(m, v); }

// block (h, d - 1), thread = row r: row r of T[h][d] = image of basis vector e_r under M[2h]^(d & 3) M[2h+1]^(d >> 2)
__global__ void __launch_bounds__(RNG_ROWS) rng_tables_kernel(uint32_t *__restrict__ tables) {
    const int h = blockIdx.x / 15, d = blockIdx.x % 15 + 1, r = threadIdx.x;
    uint32_t v[5] = {0, 0, 0, 0, 0};
    v[r >> 5] = 1u << (r & 31);
    for (int t = 0; t < (d & 3); t++) xorwow_matvec_s<5>(precalc_xorwow_matrix[2 * h], v);
    for (int t = 0; t < (d >> 2); t++) xorwow_matvec_s<5>(precalc_xorwow_matrix[2 * h + 1], v);
    uint32_t *out = tables + ((size_t) (h * 15 + d - 1) * RNG_ROWS + r) * RNG_ROW_WORDS;
    for (int k = 0; k < RNG_ROW_WORDS; k++) out[k] = k < 5 ? v[k] : 0u;
}

// The 5 index bits that differ between the lanes of a warp, as ONE product per lane: LOW[row][word][c] = word of the
// image of e_row under the jump by c sequences, c = 0..31.  For a fixed (row, word) the 32 values are one 128-byte
// line, so lanes with c = 0..31 (in any order) load their own matrix with one coalesced transaction per (row, word)
// — the lane-specific matrices that were a 16-way L1 gather in row-major layout.
// block c, thread = row r
__global__ void __launch_bounds__(RNG_ROWS) rng_low_table_kernel(uint32_t *__restrict__ tables) {
    const int c = blockIdx.x, r = threadIdx.x;
    uint32_t v[5] = {0, 0, 0, 0, 0};
    v[r >> 5] = 1u << (r & 31);
    if (c & 15) xorwow_matvec(tables + (size_t) ((c & 15) - 1) * RNG_MAT_WORDS, v);            // T[0][c & 15]
    if (c >> 4) xorwow_matvec(tables + (size_t) (15 + 1 - 1) * RNG_MAT_WORDS, v);              // T[1][1]: bit 4
    uint32_t *low = tables + RNG_HEX_TABLE_WORDS;
    for (int k = 0; k < 5; k++) low[(size_t) (r * 5 + k) * 32 + c] = v[k];
}

// Byte-digit tables for the index bits every lane of a warp shares: WIDE[0][d] = jump by d << 5 sequences, WIDE[1][d] =
// jump by d << 13, d = 1..255 (products of the hex tables) — two products cover bits 5..20 (a 1080p frame has 21 index
// bits) instead of up to five hex-digit products.  block (which, d - 1), thread = row r
__global__ void __launch_bounds__(RNG_ROWS) rng_wide_tables_kernel(uint32_t *__restrict__ tables) {
    const int which = blockIdx.x / 255, d = blockIdx.x % 255 + 1, r = threadIdx.x;
    uint32_t v[5] = {0, 0, 0, 0, 0};
    v[r >> 5] = 1u << (r & 31);
    const uint32_t val = (uint32_t) d << (which ? 13 : 5);
    for (int h = 0; h < RNG_HEX_DIGITS; h++) {
        const uint32_t dd = (val >> (4 * h)) & 15u;
        if (dd) xorwow_matvec(tables + (size_t) (h * 15 + (int) dd - 1) * RNG_MAT_WORDS, v);
    }
    uint32_t *out = tables + RNG_HEX_TABLE_WORDS + RNG_LOW_TABLE_WORDS + ((size_t) (which * 255 + d - 1) * RNG_ROWS + r) * RNG_ROW_WORDS;
    for (int k = 0; k < RNG_ROW_WORDS; k++) out[k] = k < 5 ? v[k] : 0u;
}

__device__ __forceinline__ void xorwow_matvec_low(const uint32_t *__restrict__ low, uint32_t c, uint32_t v[5]) {
    uint32_t r0 = 0, r1 = 0, r2 = 0, r3 = 0, r4 = 0;
    const uint32_t *base = low + c;
#pragma unroll
    for (int i = 0; i < 5; i++) {
        const uint32_t w = v[i];
#pragma unroll 8
        for (int j = 0; j < 32; j++) {
            const uint32_t mask = 0u - ((w >> j) & 1u);
            const uint32_t *row = base + (size_t) (5 * (32 * i + j)) * 32;
            r0 ^= __ldg(row) & mask; r1 ^= __ldg(row + 32) & mask; r2 ^= __ldg(row + 64) & mask; r3 ^= __ldg(row + 96) & mask; r4 ^= __ldg(row + 128) & mask;
        }
    }
    v[0] = r0; v[1] = r1; v[2] = r2; v[3] = r3; v[4] = r4;
}

// pixelIndex of local pixel lp (tile-interleaved shard rows, cam.pitch)
__device__ __forceinline__ uint32_t rng_pixel_index(uint32_t lp, uint32_t width, uint32_t pitch, uint32_t rank, uint32_t world, uint32_t tile_rows) {
    const uint32_t lx = lp % width, ly = lp / width;
    const uint32_t gy = ((ly / tile_rows) * world + rank) * tile_rows + (ly % tile_rows);
    return pitch * gy + lx;
}
// _curand_init_scratch (curand_kernel.h): seed scrambling; returns the state's d word
__device__ __forceinline__ uint32_t rng_seed_vector(uint32_t pixelIndex, unsigned long long seed, uint32_t v[5]) {
    const unsigned long long sd = (unsigned long long) ((long long) pixelIndex ^ (long long) seed);
    const uint32_t s0 = ((uint32_t) sd) ^ 0xaad26b49u, s1 = (uint32_t) (sd >> 32) ^ 0xf7dcefddu;
    const uint32_t t0 = 1099087573u * s0, t1 = 2591861531u * s1;
    v[0] = 123456789u + t0; v[1] = 362436069u ^ t0; v[2] = 521288629u + t1; v[3] = 88675123u ^ t1; v[4] = 5783321u + t0;
    return 6615241u + t1 + t0;
}
// The sequence jump by pixelIndex for ANY assignment of pixels to lanes (all 32 lanes call it together).
// Lanes of a warp hold neighbouring pixels: their indices differ in a few low bits and agree above.  A product whose
// matrix differs from lane to lane is a 16-way gather in L1 (it bound the first version of this kernel), so the
// bits that vary across the warp are applied one at a time with the warp-uniform single-bit matrix T[b / 4][1 << b % 4]
// (every lane computes, lanes whose bit is clear keep their vector), and only the bits all lanes share go through
// the hex-digit tables.
__device__ __forceinline__ void rng_jump_generic(const uint32_t *__restrict__ tables, uint32_t pixelIndex, uint32_t v[5]) {
    // the low 5 bits: one product with the lane's own matrix from the lane-interleaved table (coalesced when the warp's
    // pixels are consecutive; correct in any case)
    xorwow_matvec_low(tables + RNG_HEX_TABLE_WORDS, pixelIndex & 31u, v);
    const uint32_t vary = __reduce_or_sync(0xffffffffu, pixelIndex ^ __shfl_sync(0xffffffffu, pixelIndex, 0)) & ~31u;
    for (uint32_t rest = vary; rest != 0u; rest &= rest - 1u) {
        const int b = __ffs(rest) - 1;
        uint32_t w[5] = {v[0], v[1], v[2], v[3], v[4]};
        xorwow_matvec(tables + (size_t) ((b >> 2) * 15 + (1 << (b & 3)) - 1) * RNG_MAT_WORDS, w);
        if ((pixelIndex >> b) & 1u) { v[0] = w[0]; v[1] = w[1]; v[2] = w[2]; v[3] = w[3]; v[4] = w[4]; }
    }
    uint32_t p = pixelIndex & ~(vary | 31u);
    const uint32_t *wide = tables + RNG_HEX_TABLE_WORDS + RNG_LOW_TABLE_WORDS;
    const uint32_t da = (p >> 5) & 255u, db = (p >> 13) & 255u;
    if (da) xorwow_matvec(wide + (size_t) (da - 1u) * RNG_MAT_WORDS, v);
    if (db) xorwow_matvec(wide + (size_t) (255u + db - 1u) * RNG_MAT_WORDS, v);
    p &= ~0x1FFFFFu;                                   // bits 21 and up (frames beyond 2 M pixels): hex digits
    for (int h = 0; p != 0u; h++, p >>= 4) {
        const uint32_t d = p & 15u;
        if (d) xorwow_matvec(tables + (size_t) (h * 15 + (int) d - 1) * RNG_MAT_WORDS, v);
    }
}

// One pixel per thread (RTB200_RNG_INIT=single; the kernel of round 1)
__global__ void __launch_bounds__(256) rng_init_tables_kernel(RngState *__restrict__ states, const uint32_t *__restrict__ tables, unsigned long long seed,
                                                              uint32_t width, uint32_t pitch, uint32_t local_rows, uint32_t rank, uint32_t world,
                                                              uint32_t tile_rows) {
    const uint32_t lp0 = blockIdx.x * blockDim.x + threadIdx.x;
    const bool valid = lp0 < width * local_rows;
    const uint32_t lp = valid ? lp0 : width * local_rows - 1u;          // lanes past the end shadow the last pixel (warp collectives below)
    const uint32_t pixelIndex = rng_pixel_index(lp, width, pitch, rank, world, tile_rows);
    uint32_t v[5];
    RngState r;
    r.d = rng_seed_vector(pixelIndex, seed, v);
    rng_jump_generic(tables, pixelIndex, v);
    r.v0 = v[0]; r.v1 = v[1]; r.v2 = v[2]; r.v3 = v[3]; r.v4 = v[4];
    if (valid) states[lp0] = r;
}

// ---- four pixels per thread: the rows of a matrix are loaded once for four vectors ---------------------------------------
// The one-pixel kernel issues 9 loads per matrix row index and pixel (5 for the lane's own low-bits matrix, 2 + 2 for the two
// byte-digit matrices) against 15 logic operations: it is bound as much by L1 wavefronts (1 440 per warp) as by issue slots.
// A thread here takes the pixels lp, lp + 32, lp + 64, lp + 96 of its warp's 128 consecutive pixels: same lane, hence the SAME
// low-bits matrix, and — 128 aligned pixels never straddle an 8 192-pixel block — the same high byte digit; only the middle
// digit differs (consecutive values).  Loads per row index and pixel: 5/4 + 2 + 2/4 = 3.75.  The bit of the vector that selects
// a row is taken from the top of a bit-reversed copy that shifts left once per row (sign-extending shift = the mask: two
// instructions instead of four).  Same products, same order per pixel as the one-pixel kernel: the states are bit-identical
// (every parity test depends on them).  Warps whose 128 pixels are not that regular (shard boundaries of odd geometries, the
// end of the frame, indices beyond 2^21) take the generic path pixel by pixel.
constexpr int RNG_K = 4;
template<int K, bool LOW>
__device__ __forceinline__ void xorwow_matvec_k(const uint32_t *__restrict__ m, uint32_t v[K][5]) {
    uint32_t r[K][5];
#pragma unroll
    for (int k = 0; k < K; k++) { r[k][0] = 0u; r[k][1] = 0u; r[k][2] = 0u; r[k][3] = 0u; r[k][4] = 0u; }
#pragma unroll
    for (int i = 0; i < 5; i++) {
        uint32_t wr[K];
#pragma unroll
        for (int k = 0; k < K; k++) wr[k] = __brev(v[k][i]);
#pragma unroll 4
        for (int j = 0; j < 32; j++) {
            uint32_t a0, a1, a2, a3, a4;
            if (LOW) {                      // m = lane table + lane: word w of row at [(5 * row + w) * 32]
                const uint32_t *row = m + (size_t) (5 * (32 * i + j)) * 32;
                a0 = __ldg(row); a1 = __ldg(row + 32); a2 = __ldg(row + 64); a3 = __ldg(row + 96); a4 = __ldg(row + 128);
            } else {
                const uint32_t *row = m + RNG_ROW_WORDS * (32 * i + j);
                const uint4 a = __ldg(reinterpret_cast<const uint4 *>(row));
                a0 = a.x; a1 = a.y; a2 = a.z; a3 = a.w; a4 = __ldg(row + 4);
            }
#pragma unroll
            for (int k = 0; k < K; k++) {
                const uint32_t mask = (uint32_t) ((int32_t) wr[k] >> 31);
                wr[k] <<= 1;
                r[k][0] ^= a0 & mask; r[k][1] ^= a1 & mask; r[k][2] ^= a2 & mask; r[k][3] ^= a3 & mask; r[k][4] ^= a4 & mask;
            }
        }
    }
#pragma unroll
    for (int k = 0; k < K; k++) { v[k][0] = r[k][0]; v[k][1] = r[k][1]; v[k][2] = r[k][2]; v[k][3] = r[k][3]; v[k][4] = r[k][4]; }
}

__global__ void __launch_bounds__(256) rng_init_tables4_kernel(RngState *__restrict__ states, const uint32_t *__restrict__ tables, unsigned long long seed,
                                                               uint32_t width, uint32_t pitch, uint32_t local_rows, uint32_t rank, uint32_t world,
                                                               uint32_t tile_rows) {
    const uint32_t n = width * local_rows;
    const uint32_t warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31u;
    const uint32_t base = warp * (32u * RNG_K);
    if (base >= n) return;                                             // the whole warp
    uint32_t pix[RNG_K], d[RNG_K], v[RNG_K][5];
    bool regular = true;
#pragma unroll
    for (int k = 0; k < RNG_K; k++) {
        const uint32_t lp0 = base + 32u * k + lane;
        pix[k] = rng_pixel_index(lp0 < n ? lp0 : n - 1u, width, pitch, rank, world, tile_rows);      // lanes past the end shadow the last pixel
        d[k] = rng_seed_vector(pix[k], seed, v[k]);
    }
    const uint32_t p0 = __shfl_sync(0xffffffffu, pix[0], 0);
#pragma unroll
    for (int k = 0; k < RNG_K; k++) regular = regular && pix[k] == p0 + 32u * k + lane;
    const uint32_t last = p0 + 32u * RNG_K - 1u;
    regular = __all_sync(0xffffffffu, regular) && (p0 & 31u) == 0u && (p0 >> 13) == (last >> 13) && last < (1u << 21);
    if (regular) {
        const uint32_t *wide = tables + RNG_HEX_TABLE_WORDS + RNG_LOW_TABLE_WORDS;
        xorwow_matvec_k<RNG_K, true>(tables + RNG_HEX_TABLE_WORDS + lane, v);                       // bits 0..4: the lane's matrix, four vectors
#pragma unroll
        for (int k = 0; k < RNG_K; k++) {                                                    // bits 5..12: one matrix per 32 pixels
            const uint32_t da = ((p0 >> 5) + (uint32_t) k) & 255u;
            if (da) xorwow_matvec_k<1, false>(wide + (size_t) (da - 1u) * RNG_MAT_WORDS, &v[k]);
        }
        const uint32_t db = (p0 >> 13) & 255u;                                               // bits 13..20: one matrix, four vectors
        if (db) xorwow_matvec_k<RNG_K, false>(wide + (size_t) (255u + db - 1u) * RNG_MAT_WORDS, v);
    } else {
#pragma unroll
        for (int k = 0; k < RNG_K; k++) rng_jump_generic(tables, pix[k], v[k]);
    }
#pragma unroll
    for (int k = 0; k < RNG_K; k++) {
        const uint32_t lp0 = base + 32u * k + lane;
        if (lp0 < n) {
            RngState r;
            r.d = d[k]; r.v0 = v[k][0]; r.v1 = v[k][1]; r.v2 = v[k][2]; r.v3 = v[k][3]; r.v4 = v[k][4];
            states[lp0] = r;
        }
    }
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

void launch_rng_tables(uint32_t *tables, bool &tables_ready, cudaStream_t stream) {
    if (tables_ready || !tables) return;
    rng_tables_kernel<<<RNG_HEX_DIGITS * 15, RNG_ROWS, 0, stream>>>(tables);
    rng_low_table_kernel<<<32, RNG_ROWS, 0, stream>>>(tables);
    rng_wide_tables_kernel<<<2 * 255, RNG_ROWS, 0, stream>>>(tables);
    CUDA_CHECK(cudaGetLastError());
    tables_ready = true;
}

void launch_rng_init(RngState *states, unsigned long long seed, const CameraD &cam, const FrameD &fr, cudaStream_t stream, uint32_t *tables,
                     bool &tables_ready) {
    const size_t n = (size_t) cam.width * fr.local_rows;
    if (n == 0) return;
    static const bool use_curand = getenv("RTB200_RNG_INIT") && strcmp(getenv("RTB200_RNG_INIT"), "curand") == 0;
    if (tables && !use_curand) {
        launch_rng_tables(tables, tables_ready, stream);
        static const bool single = getenv("RTB200_RNG_INIT") && strcmp(getenv("RTB200_RNG_INIT"), "single") == 0;       // A/B: one pixel per thread
        if (single)
            rng_init_tables_kernel<<<div_up(n, 256), 256, 0, stream>>>(states, tables, seed, cam.width, cam.pitch, fr.local_rows, fr.shard_rank, fr.shard_world,
                                                                      fr.tile_rows);
        else
            rng_init_tables4_kernel<<<div_up(div_up(n, (size_t) 32 * RNG_K) * 32, 256), 256, 0, stream>>>(states, tables, seed, cam.width, cam.pitch, fr.local_rows,
                                                                                                        fr.shard_rank, fr.shard_world, fr.tile_rows);
        CUDA_CHECK(cudaGetLastError());
        return;
    }
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
