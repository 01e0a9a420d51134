// K1 (second mode) — fused key partition + record scatter for map outputs that are NOT yet grouped by
// reducer: every record's destination is computed on the fly (Hadoop HashPartitioner or a total-order
// range partitioner) and the record is stored straight into the destination rank's receive arena
// (a peer pointer over NVLink), staged per destination in shared memory so that the stores are coalesced.
//
// Reference context: the reference never partitions — Hadoop's MapTask does, and UDA then fetches partition r
// of every map (src/MOFServer/IndexInfo.cc:237-335, src/DataNet/RDMAServer.cc:537-631). Here the map side is
// on the GPU too, so partitioning fuses with the transfer. Skew (Zipf keys) shows up as very different
// per-destination counts; uda_partition_count gives exact counts first so that arena regions can be sized
// before the scatter (no overflow handling on the data path).
#include "common.cuh"
#include "scan.cuh"
#include "../../include/uda_b200.h"

namespace uda {

constexpr int PS_THREADS = 256;
constexpr int PS_IPT = 8;
constexpr int PS_TILE = PS_THREADS * PS_IPT;  // records per block tile
constexpr int PS_MAXP = 256;                  // destinations

// WritableComparator.hashBytes over the key body, then (h & MAX_INT) % P  (Hadoop HashPartitioner)
__device__ __forceinline__ int hash_partition(const uint8_t* k, int len, int P) {
    int h = 1;
    for (int i = 0; i < len; ++i) h = 31 * h + (int)(int8_t)k[i];
    return (int)((unsigned)(h & 0x7FFFFFFF) % (unsigned)P);
}
// index of the first splitter strictly greater than the key prefix (TotalOrderPartitioner); splitters = P-1
// big-endian 8-byte prefixes
__device__ __forceinline__ int range_partition(unsigned long long w0, const unsigned long long* __restrict__ spl, int P) {
    int lo = 0, hi = P - 1;
    while (lo < hi) {
        int mid = (lo + hi) >> 1;
        if (spl[mid] <= w0) lo = mid + 1; else hi = mid;
    }
    return lo;
}

__device__ __forceinline__ int record_dest(const uint8_t* rec, int key_off, int key_len, int skip, int mode, int P,
                                           const unsigned long long* __restrict__ spl) {
    const uint8_t* k = rec + key_off;
    if (mode == 0) return hash_partition(k + skip, key_len - skip, P);
    unsigned long long w0;
    uint32_t t;
    key_prefix(k + skip, key_len - skip, &w0, &t);
    return range_partition(w0, spl, P);
}

__global__ void __launch_bounds__(PS_THREADS) partition_count_kernel(const uint8_t* __restrict__ recs, long long n, int stride,
                                                                     int key_off, int key_len, int skip, int mode, int P,
                                                                     const unsigned long long* __restrict__ spl,
                                                                     unsigned long long* __restrict__ counts) {
    __shared__ unsigned h[PS_MAXP];
    for (int i = threadIdx.x; i < P; i += PS_THREADS) h[i] = 0;
    __syncthreads();
    for (long long i = (long long)blockIdx.x * PS_THREADS + threadIdx.x; i < n; i += (long long)gridDim.x * PS_THREADS)
        atomicAdd(&h[record_dest(recs + i * stride, key_off, key_len, skip, mode, P, spl)], 1u);
    __syncthreads();
    for (int i = threadIdx.x; i < P; i += PS_THREADS)
        if (h[i]) atomicAdd(&counts[i], (unsigned long long)h[i]);
}

// dst[p] = base of this source's region in destination p (peer pointer); cursor[p] = records already placed
template <int STRIDE>
__global__ void __launch_bounds__(PS_THREADS) partition_scatter_kernel(const uint8_t* __restrict__ recs, long long n, int stride_rt,
                                                                       int key_off, int key_len, int skip, int mode, int P,
                                                                       const unsigned long long* __restrict__ spl,
                                                                       const unsigned long long* __restrict__ dst,
                                                                       unsigned long long* __restrict__ cursor) {
    const int stride = STRIDE > 0 ? STRIDE : stride_rt;
    extern __shared__ __align__(16) uint8_t ps_smem[];  // PS_TILE * stride staged bytes
    __shared__ unsigned cnt[PS_MAXP], excl[PS_MAXP + 1];
    __shared__ unsigned long long gbase[PS_MAXP];
    const long long ntiles = (n + PS_TILE - 1) / PS_TILE;
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const long long t0 = tile * PS_TILE;
        const int m = (int)(n - t0 < PS_TILE ? n - t0 : PS_TILE);
        for (int i = threadIdx.x; i < P; i += PS_THREADS) cnt[i] = 0;
        __syncthreads();
        int dest[PS_IPT];
        unsigned rank[PS_IPT];
#pragma unroll
        for (int j = 0; j < PS_IPT; ++j) {
            const int i = j * PS_THREADS + threadIdx.x;
            dest[j] = -1;
            if (i < m) {
                dest[j] = record_dest(recs + (t0 + i) * stride, key_off, key_len, skip, mode, P, spl);
                rank[j] = atomicAdd(&cnt[dest[j]], 1u);
            }
        }
        __syncthreads();
        if (threadIdx.x == 0) {  // P <= 256: a serial scan is cheaper than another barrier ladder
            unsigned acc = 0;
            for (int p = 0; p < P; ++p) { excl[p] = acc; acc += cnt[p]; }
            excl[P] = acc;
        }
        for (int p = threadIdx.x; p < P; p += PS_THREADS)
            gbase[p] = cnt[p] ? atomicAdd(&cursor[p], (unsigned long long)cnt[p]) : 0ull;  // one reservation per (tile, destination)
        __syncthreads();
        // stage the tile grouped by destination
#pragma unroll
        for (int j = 0; j < PS_IPT; ++j) {
            const int i = j * PS_THREADS + threadIdx.x;
            if (i < m) {
                const uint8_t* s = recs + (t0 + i) * stride;
                uint8_t* d = ps_smem + (size_t)(excl[dest[j]] + rank[j]) * stride;
                if (STRIDE == 16) *reinterpret_cast<int4*>(d) = __ldg(reinterpret_cast<const int4*>(s));
                else for (int b = 0; b < stride; b += 2) *reinterpret_cast<uint16_t*>(d + b) = *reinterpret_cast<const uint16_t*>(s + b);
            }
        }
        __syncthreads();
        // consecutive threads now hold consecutive slots of one destination: coalesced (peer) stores
        for (int i = threadIdx.x; i < m; i += PS_THREADS) {
            int lo = 0, hi = P - 1;
            while (lo < hi) {
                int mid = (lo + hi + 1) >> 1;
                if (excl[mid] <= (unsigned)i) lo = mid; else hi = mid - 1;
            }
            uint8_t* d = (uint8_t*)(uintptr_t)dst[lo] + (gbase[lo] + (unsigned)i - excl[lo]) * (unsigned long long)stride;
            const uint8_t* s = ps_smem + (size_t)i * stride;
            if (STRIDE == 16) st_stream(reinterpret_cast<int4*>(d), *reinterpret_cast<const int4*>(s));
            else for (int b = 0; b < stride; b += 2) *reinterpret_cast<uint16_t*>(d + b) = *reinterpret_cast<const uint16_t*>(s + b);
        }
        __syncthreads();
    }
}

// ---- combiner: sum the 4-byte big-endian values of equal adjacent keys in a sorted fixed-stride stream -------
// pass 1: head[i] = 1 when record i starts a new key group; pass 2 (after an exclusive scan of head): group
// leaders walk their group (groups are short except for the few Zipf heads, which are split across warps by a
// second-level atomic add) and emit one record per key.
__global__ void combine_heads_kernel(const uint8_t* __restrict__ recs, long long n, int stride, int key_off, int key_len,
                                     uint32_t* __restrict__ head) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        bool h = i == 0;
        if (!h) {
            const uint8_t* a = recs + i * stride + key_off;
            const uint8_t* b = a - stride;
            for (int j = 0; j < key_len; ++j) if (a[j] != b[j]) { h = true; break; }
        }
        head[i] = h ? 1u : 0u;
    }
}

__global__ void combine_emit_kernel(const uint8_t* __restrict__ recs, long long n, int stride, int key_off, int key_len,
                                    int val_off, const uint32_t* __restrict__ gid /* exclusive scan of head, n+1 entries */,
                                    uint8_t* __restrict__ out, unsigned long long* __restrict__ sums) {
    const int lane = threadIdx.x & 31;
    const long long nthreads = (long long)gridDim.x * blockDim.x;
    const long long t0 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    for (long long wi = t0 - lane; wi < n; wi += nthreads) {  // warp-uniform: the warp-level reduction needs all lanes
        const long long i = wi + lane;
        const bool valid = i < n;
        uint32_t g = 0xFFFFFFFFu, x = 0;
        bool is_head = false;
        const uint8_t* r = recs + (valid ? i : 0) * stride;
        if (valid) {
            // head[i] = gid[i+1] - gid[i]; the group of record i is gid[i+1] - 1
            is_head = gid[i + 1] != gid[i];
            g = gid[i + 1] - 1u;
            const uint8_t* v = r + val_off;
            x = ((uint32_t)v[0] << 24) | ((uint32_t)v[1] << 16) | ((uint32_t)v[2] << 8) | v[3];
        }
        // records are sorted, so a warp sees few groups: one atomic per (warp, group) instead of one per record —
        // this is what keeps the Zipf head words (millions of equal keys) from serialising on one address
        const uint32_t mask = __match_any_sync(0xffffffffu, g);
        const uint32_t lo16 = __reduce_add_sync(mask, x & 0xFFFFu), hi16 = __reduce_add_sync(mask, x >> 16);
        if (valid && lane == __ffs(mask) - 1) atomicAdd(&sums[g], (unsigned long long)lo16 + ((unsigned long long)hi16 << 16));
        if (is_head)  // the group leader copies header + key; the value is patched in by combine_finish
            for (int b = 0; b < stride; ++b) out[(long long)g * stride + b] = r[b];
    }
}

__global__ void combine_finish_kernel(uint8_t* __restrict__ out, long long groups, int stride, int val_off,
                                      const unsigned long long* __restrict__ sums) {
    for (long long g = (long long)blockIdx.x * blockDim.x + threadIdx.x; g < groups; g += (long long)gridDim.x * blockDim.x) {
        const unsigned long long s = sums[g];
        uint8_t* v = out + g * stride + val_off;
        v[0] = (uint8_t)(s >> 24); v[1] = (uint8_t)(s >> 16); v[2] = (uint8_t)(s >> 8); v[3] = (uint8_t)s;
    }
}

}  // namespace uda

extern "C" int uda_partition_count(const void* recs, long long n, int stride, int key_off, int key_len, int key_skip, int mode,
                                   int P, const unsigned long long* splitters_dev, unsigned long long* counts_dev,
                                   cudaStream_t stream) {
    using namespace uda;
    if (P < 1 || P > PS_MAXP || stride < 2 || (stride & 1)) return UDA_E_INVALID;
    UDA_CUDA(cudaMemsetAsync(counts_dev, 0, (size_t)P * 8, stream));
    if (n <= 0) return 0;
    long long b = (n + PS_THREADS - 1) / PS_THREADS;
    if (b > kNumSMs * 8) b = kNumSMs * 8;
    partition_count_kernel<<<(int)b, PS_THREADS, 0, stream>>>((const uint8_t*)recs, n, stride, key_off, key_len, key_skip, mode, P,
                                                               splitters_dev, counts_dev);
    UDA_CHECK_LAUNCH();
    return 0;
}

// cursor_dev[P] must be zeroed by the caller before the first call of a job (calls may be chained per map).
extern "C" int uda_partition_scatter(const void* recs, long long n, int stride, int key_off, int key_len, int key_skip, int mode,
                                     int P, const unsigned long long* splitters_dev, const unsigned long long* dst_dev,
                                     unsigned long long* cursor_dev, cudaStream_t stream) {
    using namespace uda;
    if (P < 1 || P > PS_MAXP || stride < 2 || (stride & 1) || stride > 104) return UDA_E_INVALID;
    if (n <= 0) return 0;
    long long tiles = (n + PS_TILE - 1) / PS_TILE;
    long long b = tiles < kNumSMs * 2 ? tiles : kNumSMs * 2;
    const size_t smem = (size_t)PS_TILE * stride;
    if (stride == 16) {
        partition_scatter_kernel<16><<<(int)b, PS_THREADS, smem, stream>>>((const uint8_t*)recs, n, stride, key_off, key_len, key_skip,
                                                                            mode, P, splitters_dev, dst_dev, cursor_dev);
    } else {
        UDA_CUDA(ensure_dynamic_smem((const void*)partition_scatter_kernel<0>, smem));
        partition_scatter_kernel<0><<<(int)b, PS_THREADS, smem, stream>>>((const uint8_t*)recs, n, stride, key_off, key_len, key_skip,
                                                                           mode, P, splitters_dev, dst_dev, cursor_dev);
    }
    UDA_CHECK_LAUNCH();
    return 0;
}

extern "C" int uda_combine_heads(const void* recs, long long n, int stride, int key_off, int key_len, unsigned* head_dev,
                                 cudaStream_t stream) {
    if (n <= 0) return 0;
    int b = (int)((n + 255) / 256 < 148LL * 16 ? (n + 255) / 256 : 148LL * 16);
    uda::combine_heads_kernel<<<b, 256, 0, stream>>>((const uint8_t*)recs, n, stride, key_off, key_len, head_dev);
    UDA_CHECK_LAUNCH();
    return 0;
}

// gid_dev = exclusive scan of head (so gid of a record = index of its key group); sums_dev[groups] zeroed here.
extern "C" int uda_combine_emit(const void* recs, long long n, long long groups, int stride, int key_off, int key_len, int val_off,
                                const unsigned* gid_dev, void* out, unsigned long long* sums_dev, cudaStream_t stream) {
    if (n <= 0) return 0;
    UDA_CUDA(cudaMemsetAsync(sums_dev, 0, (size_t)groups * 8, stream));
    int b = (int)((n + 255) / 256 < 148LL * 16 ? (n + 255) / 256 : 148LL * 16);
    uda::combine_emit_kernel<<<b, 256, 0, stream>>>((const uint8_t*)recs, n, stride, key_off, key_len, val_off, gid_dev, (uint8_t*)out,
                                                    sums_dev);
    int b2 = (int)((groups + 255) / 256 < 148LL * 16 ? (groups + 255) / 256 : 148LL * 16);
    if (b2 < 1) b2 = 1;
    uda::combine_finish_kernel<<<b2, 256, 0, stream>>>((uint8_t*)out, groups, stride, val_off, sums_dev);
    UDA_CHECK_LAUNCH();
    return 0;
}

extern "C" long long uda_scan_u32_scratch_bytes(long long m) { return uda::scan4::scratch_elems(m) * 4; }

extern "C" int uda_scan_u32(const unsigned* v, long long m, unsigned* out, unsigned* sums_scratch, cudaStream_t stream) {
    int rc = uda::scan4::exclusive_any(v, m, out, sums_scratch, stream);
    if (rc) return UDA_E_TOO_LARGE;
    UDA_CHECK_LAUNCH();
    return 0;
}
