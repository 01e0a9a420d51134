// Device-wide exclusive scan of up to 4 Mi 32-bit values in three launches (block sums, scan of the
// sums, apply). Hand-written (no CUB): used for histogram prefixes and bucket offsets.
#pragma once
#include "common.cuh"

namespace uda {
namespace scan4 {

constexpr int THREADS = 1024;
constexpr int ITEMS = 4;
constexpr int PER_BLOCK = THREADS * ITEMS;
constexpr int MAX_ELEMS = PER_BLOCK * THREADS;  // 4 Mi

__device__ __forceinline__ uint32_t block_excl_scan(uint32_t v, uint32_t* ws /* [32] shared */, uint32_t* total) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    uint32_t x = v;
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
        if (lane >= o) x += y;
    }
    if (lane == 31) ws[w] = x;
    __syncthreads();
    if (w == 0) {
        uint32_t t = ws[lane], z = t;
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t y = __shfl_up_sync(0xffffffffu, z, o);
            if (lane >= o) z += y;
        }
        ws[lane] = z - t;
        if (lane == 31 && total) *total = z;
    }
    __syncthreads();
    return ws[w] + x - v;
}

static __global__ void __launch_bounds__(THREADS) sums_kernel(const uint32_t* __restrict__ v, int m, uint32_t* __restrict__ sums) {
    __shared__ uint32_t ws[32];
    __shared__ uint32_t tot;
    const int base = blockIdx.x * PER_BLOCK + threadIdx.x * ITEMS;
    uint32_t s = 0;
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) s += base + i < m ? v[base + i] : 0u;
    block_excl_scan(s, ws, &tot);
    if (threadIdx.x == 0) sums[blockIdx.x] = tot;
}

static __global__ void __launch_bounds__(THREADS) scan_sums_kernel(uint32_t* __restrict__ sums, int nb) {
    __shared__ uint32_t ws[32];
    uint32_t s = (int)threadIdx.x < nb ? sums[threadIdx.x] : 0u;
    uint32_t e = block_excl_scan(s, ws, nullptr);
    if ((int)threadIdx.x < nb) sums[threadIdx.x] = e;
}

// out[i] = exclusive prefix (out may alias v); out[m] = grand total
static __global__ void __launch_bounds__(THREADS) apply_kernel(const uint32_t* __restrict__ v, int m, const uint32_t* __restrict__ sums,
                                                        uint32_t* __restrict__ out) {
    __shared__ uint32_t ws[32];
    const int base = blockIdx.x * PER_BLOCK + threadIdx.x * ITEMS;
    uint32_t x[ITEMS], s = 0;
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) { x[i] = base + i < m ? v[base + i] : 0u; s += x[i]; }
    uint32_t e = sums[blockIdx.x] + block_excl_scan(s, ws, nullptr);
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
        if (base + i < m) out[base + i] = e;
        e += x[i];
        if (base + i == m - 1) out[m] = e;
    }
}

// sums: scratch of THREADS u32
static inline int exclusive(const uint32_t* v, int m, uint32_t* out, uint32_t* sums, cudaStream_t stream) {
    if (m <= 0) return 0;
    if (m > MAX_ELEMS) return -2;
    const int nb = (m + PER_BLOCK - 1) / PER_BLOCK;
    sums_kernel<<<nb, THREADS, 0, stream>>>(v, m, sums);
    scan_sums_kernel<<<1, THREADS, 0, stream>>>(sums, nb);
    apply_kernel<<<nb, THREADS, 0, stream>>>(v, m, sums, out);
    return 0;
}

// Any length: the block sums are scanned recursively. scratch must hold scratch_elems(m) u32.
static inline long long scratch_elems(long long m) {
    long long tot = 0;
    while (m > PER_BLOCK) { m = (m + PER_BLOCK - 1) / PER_BLOCK; tot += m + 1; }
    return tot + THREADS + 1;
}
static inline int exclusive_any(const uint32_t* v, long long m, uint32_t* out, uint32_t* scratch, cudaStream_t stream) {
    if (m <= 0) return 0;
    if (m > (1LL << 31) - 1) return -2;
    if (m <= MAX_ELEMS) return exclusive(v, (int)m, out, scratch, stream);
    const long long nb = (m + PER_BLOCK - 1) / PER_BLOCK;
    uint32_t* sums = scratch;                 // nb + 1 entries, then the next level's scratch
    sums_kernel<<<(unsigned)nb, THREADS, 0, stream>>>(v, (int)m, sums);
    int rc = exclusive_any(sums, nb, sums, sums + nb + 1, stream);
    if (rc) return rc;
    apply_kernel<<<(unsigned)nb, THREADS, 0, stream>>>(v, (int)m, sums, out);
    return 0;
}

}  // namespace scan4
}  // namespace uda
