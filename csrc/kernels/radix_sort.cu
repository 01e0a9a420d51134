// K2 — map-side LSD radix sort of 16-byte sort entries (hand-written; no CUB/Thrust).
//
// What it stands in for: Hadoop's MapTask sorts each map output before UDA ever sees it; the
// reference only consumes the sorted MOFs (src/Merger/MergeQueue.h:299-321 assumes every segment is
// sorted). The new engine owns the map side, so the sort is a kernel of this library.
//
// Design: 8-bit digits over the 80-bit key prefix, least significant byte first. Per pass:
//   histogram (per-block digit counts, digit-major) -> single-block exclusive scan -> stable scatter.
// The scatter ranks each 32-element warp round with __match_any_sync and per-warp shared histograms,
// so there are no shared-memory atomics on the ranking path and the order (warp, round, lane) equals
// memory order, which keeps the sort stable. The low 48 bits of w1 (run, index) are never sorted on:
// stability already orders equal keys by original position.
#include "common.cuh"
#include "../../include/uda_b200.h"

namespace uda {

constexpr int RS_THREADS = 256;
constexpr int RS_IPT = 8;
constexpr int RS_TILE = RS_THREADS * RS_IPT;  // 2048 entries = 32 KiB
constexpr int RS_MAX_BLOCKS = kNumSMs * 4;

// byte j of the 80-bit prefix, j = 0 is the most significant key byte
__device__ __forceinline__ uint32_t key_byte(const Entry& e, int j) {
    return j < 8 ? (uint32_t)(e.w0 >> (8 * (7 - j))) & 0xFFu : (uint32_t)(e.w1 >> (48 + 8 * (9 - j))) & 0xFFu;
}

__global__ void __launch_bounds__(RS_THREADS) rs_hist_kernel(const Entry* __restrict__ in, long long n, long long chunk,
                                                             int nblk, int byte_j, uint32_t* __restrict__ counts) {
    __shared__ uint32_t h[256];
    h[threadIdx.x] = 0;
    __syncthreads();
    long long beg = (long long)blockIdx.x * chunk;
    long long end = beg + chunk < n ? beg + chunk : n;
    for (long long i = beg + threadIdx.x; i < end; i += RS_THREADS) {
        Entry e = in[i];
        atomicAdd(&h[key_byte(e, byte_j)], 1u);
    }
    __syncthreads();
    counts[(long long)threadIdx.x * nblk + blockIdx.x] = h[threadIdx.x];
}

// exclusive scan of m values in place, one block
__global__ void __launch_bounds__(1024) rs_scan_kernel(uint32_t* __restrict__ v, int m) {
    __shared__ uint32_t warp_sums[32];
    __shared__ uint32_t carry;
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const int per = (m + 1023) / 1024;
    const int beg = tid * per, end = beg + per < m ? beg + per : m;
    uint32_t s = 0;
    for (int i = beg; i < end; ++i) s += v[i];
    uint32_t x = s;
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
        if (lane >= o) x += y;
    }
    if (lane == 31) warp_sums[w] = x;
    __syncthreads();
    if (w == 0) {
        uint32_t t = warp_sums[lane], z = t;
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t y = __shfl_up_sync(0xffffffffu, z, o);
            if (lane >= o) z += y;
        }
        warp_sums[lane] = z - t;
        if (lane == 31) carry = z;
    }
    __syncthreads();
    uint32_t run = warp_sums[w] + (x - s);
    for (int i = beg; i < end; ++i) {
        uint32_t t = v[i];
        v[i] = run;
        run += t;
    }
}

__global__ void __launch_bounds__(RS_THREADS) rs_scatter_kernel(const Entry* __restrict__ in, Entry* __restrict__ out,
                                                                long long n, long long chunk, int nblk, int byte_j,
                                                                const uint32_t* __restrict__ offsets) {
    __shared__ uint32_t whist[RS_THREADS / 32][256];
    __shared__ uint32_t running[256];
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    const uint32_t lt_mask = (1u << lane) - 1u;
    running[tid] = offsets[(long long)tid * nblk + blockIdx.x];
    const long long beg = (long long)blockIdx.x * chunk;
    const long long end = beg + chunk < n ? beg + chunk : n;
    for (long long tile = beg; tile < end; tile += RS_TILE) {
#pragma unroll
        for (int i = 0; i < RS_THREADS / 32; ++i) whist[i][tid] = 0;
        __syncthreads();
        Entry e[RS_IPT];
        uint32_t d[RS_IPT], pre[RS_IPT];
#pragma unroll
        for (int j = 0; j < RS_IPT; ++j) {
            long long idx = tile + (long long)w * (32 * RS_IPT) + j * 32 + lane;
            bool valid = idx < end;
            if (valid) e[j] = in[idx];
            d[j] = valid ? key_byte(e[j], byte_j) : 256u;
        }
#pragma unroll
        for (int j = 0; j < RS_IPT; ++j) {
            uint32_t mask = __match_any_sync(0xffffffffu, d[j]);
            int leader = __ffs(mask) - 1;
            uint32_t base = 0;
            if (lane == leader && d[j] < 256u) {
                base = whist[w][d[j]];
                whist[w][d[j]] = base + __popc(mask);
            }
            base = __shfl_sync(0xffffffffu, base, leader);
            pre[j] = base + __popc(mask & lt_mask);
            __syncwarp();
        }
        __syncthreads();
        {
            uint32_t acc = running[tid];
#pragma unroll
            for (int ww = 0; ww < RS_THREADS / 32; ++ww) {
                uint32_t t = whist[ww][tid];
                whist[ww][tid] = acc;
                acc += t;
            }
            running[tid] = acc;
        }
        __syncthreads();
#pragma unroll
        for (int j = 0; j < RS_IPT; ++j)
            if (d[j] < 256u) out[whist[w][d[j]] + pre[j]] = e[j];
        __syncthreads();
    }
}

static inline void rs_plan(long long n, int* nblk, long long* chunk) {
    long long tiles = (n + RS_TILE - 1) / RS_TILE;
    long long b = tiles < RS_MAX_BLOCKS ? tiles : RS_MAX_BLOCKS;
    if (b < 1) b = 1;
    long long tiles_per = (tiles + b - 1) / b;
    *chunk = tiles_per * RS_TILE;
    *nblk = (int)((n + *chunk - 1) / *chunk);
    if (*nblk < 1) *nblk = 1;
}

}  // namespace uda

extern "C" long long uda_radix_sort_scratch_bytes(long long n) {
    (void)n;
    return 256LL * uda::RS_MAX_BLOCKS * (long long)sizeof(uint32_t);
}

extern "C" int uda_radix_sort_entries(void* a, void* b, long long n, int key_bytes, void* scratch, cudaStream_t stream) {
    using namespace uda;
    if (n <= 1) return 0;
    if (key_bytes < 1 || key_bytes > 10 || n >= (1LL << 32)) return UDA_E_INVALID;
    int nblk;
    long long chunk;
    rs_plan(n, &nblk, &chunk);
    Entry* src = (Entry*)a;
    Entry* dst = (Entry*)b;
    uint32_t* counts = (uint32_t*)scratch;
    for (int j = key_bytes - 1; j >= 0; --j) {
        rs_hist_kernel<<<nblk, RS_THREADS, 0, stream>>>(src, n, chunk, nblk, j, counts);
        rs_scan_kernel<<<1, 1024, 0, stream>>>(counts, 256 * nblk);
        rs_scatter_kernel<<<nblk, RS_THREADS, 0, stream>>>(src, dst, n, chunk, nblk, j, counts);
        Entry* t = src; src = dst; dst = t;
    }
    UDA_CHECK_LAUNCH();
    if (src != (Entry*)a) UDA_CUDA(cudaMemcpyAsync(a, src, (size_t)n * sizeof(Entry), cudaMemcpyDeviceToDevice, stream));
    return 0;
}
