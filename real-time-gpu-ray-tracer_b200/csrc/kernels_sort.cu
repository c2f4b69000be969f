// kernels_sort.cu — least-significant-digit radix sort of (64-bit Morton key, 32-bit payload) pairs in the
// "onesweep" formulation: ONE histogram kernel for all digit positions, then one kernel per 8-bit digit whose
// tiles obtain their global offsets by decoupled look-back over single-word {flag|count} tile descriptors
// (no separate scan kernel, each key is read once and written once per digit).  Stable.
//
// HBM roofline per pass: n * (8 + 4) bytes read + n * 12 bytes written; histogram: n * 8 bytes read.
#include "build.h"

namespace {

constexpr int RADIX = 256;
constexpr int SORT_THREADS = 256;
constexpr int SORT_WARPS = SORT_THREADS / 32;
constexpr int SORT_ITEMS = 12;
constexpr int SORT_TILE = SORT_THREADS * SORT_ITEMS;
constexpr int MAX_PASSES = 8;

constexpr uint32_t FLAG_AGGREGATE = 1u << 30;
constexpr uint32_t FLAG_PREFIX = 2u << 30;
constexpr uint32_t FLAG_MASK = 3u << 30;
constexpr uint32_t VALUE_MASK = (1u << 30) - 1u;

// hist[p][d] = number of keys whose p-th digit (from begin_bit) equals d
__global__ void __launch_bounds__(256) radix_histogram_kernel(const uint64_t *__restrict__ keys, int n, int passes, int begin_bit,
                                                              uint32_t *__restrict__ hist) {
    __shared__ uint32_t sh[MAX_PASSES * RADIX];
    for (int i = threadIdx.x; i < passes * RADIX; i += blockDim.x) sh[i] = 0;
    __syncthreads();
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint64_t k = keys[i] >> begin_bit;
        for (int p = 0; p < passes; p++) atomicAdd(&sh[p * RADIX + (int) ((k >> (8 * p)) & 255u)], 1u);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < passes * RADIX; i += blockDim.x)
        if (sh[i]) atomicAdd(&hist[i], sh[i]);
}

// in-place exclusive scan of each pass's 256 bins (one block per pass)
__global__ void __launch_bounds__(RADIX) radix_scan_histogram_kernel(uint32_t *__restrict__ hist) {
    __shared__ uint32_t sh[RADIX];
    uint32_t *h = hist + blockIdx.x * RADIX;
    const int d = threadIdx.x;
    const uint32_t mine = h[d];
    sh[d] = mine;
    __syncthreads();
    for (int off = 1; off < RADIX; off <<= 1) {
        const uint32_t add = d >= off ? sh[d - off] : 0u;
        __syncthreads();
        sh[d] += add;
        __syncthreads();
    }
    h[d] = sh[d] - mine;
}

__global__ void __launch_bounds__(SORT_THREADS) onesweep_pass_kernel(
        const uint64_t *__restrict__ keys_in, uint64_t *__restrict__ keys_out,
        const uint32_t *__restrict__ vals_in, uint32_t *__restrict__ vals_out,
        int n, int shift, const uint32_t *__restrict__ digit_base, volatile uint32_t *status, uint32_t *ticket) {
    __shared__ uint32_t warp_cnt[SORT_WARPS][RADIX];
    __shared__ uint32_t base[RADIX];
    __shared__ uint32_t s_tile;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    // tiles are numbered in the order blocks START, so every predecessor of a tile is resident or finished
    if (tid == 0) s_tile = atomicAdd(ticket, 1u);
    for (int i = tid; i < SORT_WARPS * RADIX; i += SORT_THREADS) (&warp_cnt[0][0])[i] = 0;
    __syncthreads();
    const uint32_t tile = s_tile;
    const long long warp_start = (long long) tile * SORT_TILE + (long long) warp * 32 * SORT_ITEMS;

    uint64_t k[SORT_ITEMS];
    uint32_t v[SORT_ITEMS];
    uint32_t rank[SORT_ITEMS];
#pragma unroll
    for (int r = 0; r < SORT_ITEMS; r++) {
        const long long idx = warp_start + r * 32 + lane;
        const bool valid = idx < n;
        k[r] = valid ? keys_in[idx] : ~0ull;
        v[r] = valid ? vals_in[idx] : 0u;
    }
    const uint32_t lt_mask = (1u << lane) - 1u;
#pragma unroll
    for (int r = 0; r < SORT_ITEMS; r++) {
        const long long idx = warp_start + r * 32 + lane;
        const bool valid = idx < n;
        const int digit = valid ? (int) ((k[r] >> shift) & 255u) : RADIX;   // RADIX = "no digit"
        const uint32_t peers = __match_any_sync(0xffffffffu, digit);
        uint32_t prev = 0;
        if (valid) {
            prev = warp_cnt[warp][digit];
            rank[r] = prev + __popc(peers & lt_mask);
        } else {
            rank[r] = 0;
        }
        __syncwarp();
        if (valid && lane == __ffs(peers) - 1) warp_cnt[warp][digit] = prev + __popc(peers);
        __syncwarp();
    }
    __syncthreads();

    {   // thread d owns digit d: warp-exclusive offsets, tile count, look-back
        const int d = tid;
        uint32_t acc = 0;
#pragma unroll
        for (int w = 0; w < SORT_WARPS; w++) {
            const uint32_t c = warp_cnt[w][d];
            warp_cnt[w][d] = acc;
            acc += c;
        }
        const uint32_t count = acc;
        uint32_t excl = 0;
        if (tile == 0) {
            status[d] = FLAG_PREFIX | count;
        } else {
            status[(size_t) tile * RADIX + d] = FLAG_AGGREGATE | count;
            int p = (int) tile - 1;
            for (;;) {
                const uint32_t s = status[(size_t) p * RADIX + d];
                const uint32_t f = s & FLAG_MASK;
                if (f == 0u) continue;
                excl += s & VALUE_MASK;
                if (f == FLAG_PREFIX) break;
                p--;
            }
            status[(size_t) tile * RADIX + d] = FLAG_PREFIX | (excl + count);
        }
        base[d] = digit_base[d] + excl;
    }
    __syncthreads();

#pragma unroll
    for (int r = 0; r < SORT_ITEMS; r++) {
        const long long idx = warp_start + r * 32 + lane;
        if (idx < n) {
            const int digit = (int) ((k[r] >> shift) & 255u);
            const uint32_t dst = base[digit] + warp_cnt[warp][digit] + rank[r];
            keys_out[dst] = k[r];
            vals_out[dst] = v[r];
        }
    }
}

} // namespace

void RadixSorter::sort(uint64_t *&keys, uint64_t *&keys_alt, uint32_t *&vals, uint32_t *&vals_alt, int n, int key_bits,
                       cudaStream_t stream, int sm_count, uint32_t &launches) {
    if (n <= 1 || key_bits <= 0) return;
    const int passes = (key_bits + 7) / 8;
    if (passes > MAX_PASSES) throw CudaError("RadixSorter: more than 64 key bits");
    const int tiles = (n + SORT_TILE - 1) / SORT_TILE;
    const size_t status_words = (size_t) passes * tiles * RADIX;
    // one scratch allocation: [hist passes*256][ticket passes (padded to 256)][status]
    const size_t total = (size_t) passes * RADIX + RADIX + status_words;
    scratch_.reserve(total);
    uint32_t *hist = scratch_.get();
    uint32_t *ticket = hist + (size_t) passes * RADIX;
    uint32_t *status = ticket + RADIX;
    CUDA_CHECK(cudaMemsetAsync(hist, 0, total * sizeof(uint32_t), stream));
    const int hist_blocks = std::min(div_up(n, 256 * 8), (unsigned) (sm_count * 8));
    radix_histogram_kernel<<<hist_blocks, 256, 0, stream>>>(keys, n, passes, 0, hist);
    radix_scan_histogram_kernel<<<passes, RADIX, 0, stream>>>(hist);
    launches += 2;
    for (int p = 0; p < passes; p++) {
        onesweep_pass_kernel<<<tiles, SORT_THREADS, 0, stream>>>(keys, keys_alt, vals, vals_alt, n, 8 * p, hist + (size_t) p * RADIX,
                                                               status + (size_t) p * tiles * RADIX, ticket + p);
        launches++;
        std::swap(keys, keys_alt);
        std::swap(vals, vals_alt);
    }
    CUDA_CHECK(cudaGetLastError());
}
