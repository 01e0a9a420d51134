// General (variable-length record) reduce path: exact tie-break of equal-prefix groups and the 64-bit output
// offsets of the merged stream.
//
// The sort entry carries the first 10 compared key bytes (common.cuh). Keys that are longer, or shorter and zero
// padded, can tie on that prefix; the radix sort / k-way merge then leaves them in (run, index) order, which is
// not the order of the exact comparator. This file re-orders exactly those groups.
//
// Reference behaviour being replaced: the byte comparators of src/Merger/CompareFunc.cc:70-91 (Text: skip the
// leading VInt, BytesWritable: skip the 4-byte length, fixed-width Writables: raw bytes; memcmp on the common
// length, then shorter-key-first) called per heap sift from src/Merger/StreamRW.h:163. Equal keys keep the order
// of their segments, as a stable heap of segments produces (src/Merger/MergeQueue.h:250-269).
//
// Steps, all full-width kernels except the last:
//   1. head flag per merged position (prefix differs from the previous entry)
//   2. exclusive scan of the flags = group id; start[] table per group
//   3. every non-head position compares itself with its predecessor using the exact comparator; an inversion
//      claims its group once (high bit of start[]) and appends it to a work list
//   4. one CTA per listed group: small groups by a stable rank sort (rank = number of group members that sort
//      before the element) into a temporary, large ones by a bitonic network in place.
// Groups of equal keys (Zipf heads) never enter the work list because (run, index) order is already right, so
// step 4 only sees groups of distinct keys sharing 10 bytes; the largest such group is reported.
#include "common.cuh"
#include "scan.cuh"
#include "../../include/uda_b200.h"

namespace uda {
namespace {

struct KeyRef {
    const uint8_t* p;
    int len;
};

__device__ __forceinline__ KeyRef key_of(const Entry& e, const unsigned long long* __restrict__ run_base,
                                         const unsigned long long* __restrict__ run_recoff, int key_kind) {
    const uint32_t run = (uint32_t)(e.w1 >> 32) & 0xFFFFu, idx = (uint32_t)e.w1;
    const long long* ro = (const long long*)(uintptr_t)run_recoff[run];
    const uint8_t* p = (const uint8_t*)(uintptr_t)run_base[run] + ro[idx];
    long long klen, vlen;
    int n1 = vint_decode(p, &klen);
    int n2 = vint_decode(p + n1, &vlen);
    KeyRef k{p + n1 + n2, (int)klen};
    if (key_kind == 1 && k.len > 0) {
        int s = vint_size_from_first((int8_t)k.p[0]);
        k.p += s; k.len -= s;
    } else if (key_kind == 2) {
        k.p += 4; k.len -= 4;
    }
    if (k.len < 0) k.len = 0;
    return k;
}

__device__ __forceinline__ int cmp_keys(const KeyRef& a, const KeyRef& b) {
    const int m = a.len < b.len ? a.len : b.len;
    for (int i = 0; i < m; ++i) {
        int d = (int)a.p[i] - (int)b.p[i];
        if (d) return d;
    }
    return a.len - b.len;
}

__device__ __forceinline__ bool same_prefix(const Entry& a, const Entry& b) {
    return a.w0 == b.w0 && (a.w1 >> 48) == (b.w1 >> 48);
}

__global__ void tb_heads_kernel(const Entry* __restrict__ S, long long n, uint32_t* __restrict__ head) {
    for (long long j = (long long)blockIdx.x * blockDim.x + threadIdx.x; j < n; j += (long long)gridDim.x * blockDim.x)
        head[j] = (j == 0 || !same_prefix(S[j], S[j - 1])) ? 1u : 0u;
}

// gidx = exclusive scan of the head flags (n+1 entries): group of j is gidx[j+1]-1, j is a head iff gidx[j+1] != gidx[j]
__global__ void tb_starts_kernel(const uint32_t* __restrict__ gidx, long long n, uint32_t* __restrict__ start) {
    for (long long j = (long long)blockIdx.x * blockDim.x + threadIdx.x; j < n; j += (long long)gridDim.x * blockDim.x) {
        if (gidx[j + 1] != gidx[j]) start[gidx[j]] = (uint32_t)j;
        if (j == n - 1) start[gidx[n]] = (uint32_t)n;
    }
}

__global__ void tb_claim_kernel(const Entry* __restrict__ S, long long n, const uint32_t* __restrict__ gidx,
                                const unsigned long long* __restrict__ run_base, const unsigned long long* __restrict__ run_recoff,
                                int key_kind, uint32_t* __restrict__ start, uint32_t* __restrict__ work,
                                unsigned long long* __restrict__ stats /* [0]=groups listed [1]=largest [2]=records moved */) {
    for (long long j = (long long)blockIdx.x * blockDim.x + threadIdx.x; j < n; j += (long long)gridDim.x * blockDim.x) {
        if (j == 0 || gidx[j + 1] != gidx[j]) continue;          // heads have no predecessor inside their group
        KeyRef a = key_of(S[j - 1], run_base, run_recoff, key_kind), b = key_of(S[j], run_base, run_recoff, key_kind);
        if (cmp_keys(a, b) <= 0) continue;
        const uint32_t g = gidx[j + 1] - 1;
        if (!(atomicOr(&start[g], 0x80000000u) >> 31)) work[atomicAdd(&stats[0], 1ull)] = g;
    }
}

constexpr int TB_THREADS = 1024;
constexpr int TB_RANK_MAX = 64;   // groups up to this size: rank sort; larger: bitonic network

__global__ void __launch_bounds__(TB_THREADS) tb_sort_kernel(Entry* __restrict__ S, Entry* __restrict__ tmp,
                                                             const unsigned long long* __restrict__ run_base,
                                                             const unsigned long long* __restrict__ run_recoff, int key_kind,
                                                             const uint32_t* __restrict__ start, const uint32_t* __restrict__ work,
                                                             unsigned long long* __restrict__ stats) {
    const unsigned long long nwork = stats[0];
    for (unsigned long long w = blockIdx.x; w < nwork; w += gridDim.x) {
        const uint32_t g = work[w];
        const long long s = start[g] & 0x7FFFFFFFu, e = start[g + 1] & 0x7FFFFFFFu;
        const int size = (int)(e - s);
        if (threadIdx.x == 0) {
            atomicMax(&stats[1], (unsigned long long)size);
            atomicAdd(&stats[2], (unsigned long long)size);
        }
        if (size > TB_RANK_MAX) {
            // bitonic network over the group in global memory (all comparisons ascending, indices >= size act as +inf):
            // O(g log^2 g) exact compares instead of the O(g^2) of the rank sort. Equal keys are ordered by
            // (run, index), which is the order a stable sort would keep.
            Entry* G = S + s;
            for (unsigned k = 2; (k >> 1) < (unsigned)size; k <<= 1) {
                for (unsigned j = k >> 1; j > 0; j >>= 1) {
                    const unsigned mask = (j == (k >> 1)) ? (k - 1) : j;   // first step of a merge flips, the rest shuffle
                    for (unsigned i = threadIdx.x; i < (unsigned)size; i += TB_THREADS) {
                        const unsigned l = i ^ mask;
                        if (l > i && l < (unsigned)size) {
                            const Entry a = G[i], b = G[l];
                            int c = cmp_keys(key_of(b, run_base, run_recoff, key_kind), key_of(a, run_base, run_recoff, key_kind));
                            if (c == 0) c = (b.w1 & 0xFFFFFFFFFFFFull) < (a.w1 & 0xFFFFFFFFFFFFull) ? -1 : 1;
                            if (c < 0) { G[i] = b; G[l] = a; }
                        }
                    }
                    __syncthreads();
                }
            }
            continue;
        }
        for (int i = threadIdx.x; i < size; i += TB_THREADS) {
            const Entry ei = S[s + i];
            const KeyRef ki = key_of(ei, run_base, run_recoff, key_kind);
            int rank = 0;
            for (int j = 0; j < size; ++j) {
                if (j == i) continue;
                const int c = cmp_keys(key_of(S[s + j], run_base, run_recoff, key_kind), ki);
                rank += (c < 0 || (c == 0 && j < i)) ? 1 : 0;
            }
            tmp[s + rank] = ei;
        }
        __syncthreads();
        for (int i = threadIdx.x; i < size; i += TB_THREADS) S[s + i] = tmp[s + i];
        __syncthreads();
    }
}

// ---- u32 lengths -> u64 exclusive offsets ---------------------------------------------------------------------
constexpr int S64_THREADS = 1024;
constexpr int S64_ITEMS = 4;
constexpr int S64_PER_BLOCK = S64_THREADS * S64_ITEMS;

__global__ void __launch_bounds__(S64_THREADS) s64_sums_kernel(const uint32_t* __restrict__ v, long long m,
                                                               unsigned long long* __restrict__ sums) {
    __shared__ unsigned long long ws[32];
    const long long base = (long long)blockIdx.x * S64_PER_BLOCK + threadIdx.x * S64_ITEMS;
    unsigned long long s = 0;
#pragma unroll
    for (int i = 0; i < S64_ITEMS; ++i) s += base + i < m ? v[base + i] : 0u;
    for (int o = 16; o; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x < 32) {
        s = ws[threadIdx.x];
        for (int o = 16; o; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o);
        if (threadIdx.x == 0) sums[blockIdx.x] = s;
    }
}

// one CTA walks all block sums (at most 2^31/4096 = 512 Ki of them) with a running carry
__global__ void __launch_bounds__(S64_THREADS) s64_scan_sums_kernel(unsigned long long* __restrict__ sums, long long nb) {
    __shared__ unsigned long long ws[32];
    __shared__ unsigned long long carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    for (long long b0 = 0; b0 < nb; b0 += S64_THREADS) {
        const long long i = b0 + threadIdx.x;
        const unsigned long long v = i < nb ? sums[i] : 0ull;
        unsigned long long x = v;
        for (int o = 1; o < 32; o <<= 1) {
            unsigned long long y = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= o) x += y;
        }
        if (lane == 31) ws[w] = x;
        __syncthreads();
        if (w == 0) {
            unsigned long long t = ws[lane], z = t;
            for (int o = 1; o < 32; o <<= 1) {
                unsigned long long y = __shfl_up_sync(0xffffffffu, z, o);
                if (lane >= o) z += y;
            }
            ws[lane] = z - t;
        }
        __syncthreads();
        const unsigned long long excl = carry + ws[w] + x - v;
        if (i < nb) sums[i] = excl;
        __syncthreads();
        if (threadIdx.x == S64_THREADS - 1) carry = excl + v;
        __syncthreads();
    }
    if (threadIdx.x == 0) sums[nb] = carry;
}

__global__ void __launch_bounds__(S64_THREADS) s64_apply_kernel(const uint32_t* __restrict__ v, long long m,
                                                                const unsigned long long* __restrict__ sums,
                                                                unsigned long long* __restrict__ out) {
    const long long base = (long long)blockIdx.x * S64_PER_BLOCK + threadIdx.x * S64_ITEMS;
    uint32_t x[S64_ITEMS];
    unsigned long long wide = 0;
#pragma unroll
    for (int i = 0; i < S64_ITEMS; ++i) { x[i] = base + i < m ? v[base + i] : 0u; wide += x[i]; }
    // 4096 record lengths can exceed 2^32, so the in-block scan is 64-bit as well
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    __shared__ unsigned long long ws64[32];
    unsigned long long xs = wide;
    for (int o = 1; o < 32; o <<= 1) {
        unsigned long long y = __shfl_up_sync(0xffffffffu, xs, o);
        if (lane >= o) xs += y;
    }
    if (lane == 31) ws64[w] = xs;
    __syncthreads();
    if (w == 0) {
        unsigned long long t = ws64[lane], z = t;
        for (int o = 1; o < 32; o <<= 1) {
            unsigned long long y = __shfl_up_sync(0xffffffffu, z, o);
            if (lane >= o) z += y;
        }
        ws64[lane] = z - t;
    }
    __syncthreads();
    unsigned long long e = sums[blockIdx.x] + ws64[w] + xs - wide;
#pragma unroll
    for (int i = 0; i < S64_ITEMS; ++i) {
        if (base + i < m) out[base + i] = e;
        e += x[i];
        if (base + i == m - 1) out[m] = e;
    }
}

__global__ void eof_marker_kernel(uint8_t* dst_base, const unsigned long long* total) {
    uint8_t* d = dst_base + (total ? *total : 0ull);
    d[0] = 0xFF;
    d[1] = 0xFF;
}

}  // namespace
}  // namespace uda

extern "C" long long uda_scan_u32_to_u64_scratch_bytes(long long m) {
    if (m < 0) return UDA_E_INVALID;
    return ((m + uda::S64_PER_BLOCK - 1) / uda::S64_PER_BLOCK + 2) * 8;
}

extern "C" int uda_scan_u32_to_u64(const unsigned* v, long long m, unsigned long long* out, void* scratch, cudaStream_t stream) {
    using namespace uda;
    if (m < 0) return UDA_E_INVALID;
    if (m == 0) return cudaMemsetAsync(out, 0, 8, stream) == cudaSuccess ? 0 : (int)cudaGetLastError();
    if (m > (1LL << 31) - 1) return UDA_E_TOO_LARGE;
    const long long nb = (m + S64_PER_BLOCK - 1) / S64_PER_BLOCK;
    unsigned long long* sums = (unsigned long long*)scratch;
    s64_sums_kernel<<<(unsigned)nb, S64_THREADS, 0, stream>>>(v, m, sums);
    s64_scan_sums_kernel<<<1, S64_THREADS, 0, stream>>>(sums, nb);
    s64_apply_kernel<<<(unsigned)nb, S64_THREADS, 0, stream>>>(v, m, sums, out);
    UDA_CHECK_LAUNCH();
    return 0;
}

extern "C" long long uda_tiebreak_scratch_bytes(long long n) {
    if (n < 0) return UDA_E_INVALID;
    // gidx u32[n+1] | start u32[n+1] | work u32[n/2+1] | stats u64[4] | scan scratch
    long long words = (n + 1) + (n + 1) + (n / 2 + 1);
    words = (words + 3) / 4 * 4;
    return words * 4 + 32 + uda::scan4::scratch_elems(n) * 4 + 64;
}

extern "C" int uda_tiebreak_var(void* sorted_entries, long long n, const unsigned long long* run_base_dev,
                                const unsigned long long* run_recoff_dev, int key_kind, void* tmp_entries, void* scratch,
                                cudaStream_t stream) {
    using namespace uda;
    if (n < 0 || key_kind < 0 || key_kind > 2) return UDA_E_INVALID;
    if (n >= (1LL << 31) - 1) return UDA_E_TOO_LARGE;
    long long words = (n + 1) + (n + 1) + (n / 2 + 1);
    words = (words + 3) / 4 * 4;
    uint32_t* gidx = (uint32_t*)scratch;
    uint32_t* start = gidx + (n + 1);
    uint32_t* work = start + (n + 1);
    unsigned long long* stats = (unsigned long long*)((uint32_t*)scratch + words);
    uint32_t* scan_scratch = (uint32_t*)(stats + 4);
    if (cudaMemsetAsync(stats, 0, 32, stream) != cudaSuccess) return (int)cudaGetLastError();
    if (n < 2) return 0;
    Entry* S = (Entry*)sorted_entries;
    int blocks = (int)((n + 255) / 256 < kNumSMs * 16LL ? (n + 255) / 256 : kNumSMs * 16LL);
    tb_heads_kernel<<<blocks, 256, 0, stream>>>(S, n, gidx);
    if (scan4::exclusive_any(gidx, n, gidx, scan_scratch, stream)) return UDA_E_TOO_LARGE;
    tb_starts_kernel<<<blocks, 256, 0, stream>>>(gidx, n, start);
    tb_claim_kernel<<<blocks, 256, 0, stream>>>(S, n, gidx, run_base_dev, run_recoff_dev, key_kind, start, work, stats);
    tb_sort_kernel<<<kNumSMs * 2, TB_THREADS, 0, stream>>>(S, (Entry*)tmp_entries, run_base_dev, run_recoff_dev, key_kind, start, work,
                                                          stats);
    UDA_CHECK_LAUNCH();
    return 0;
}

extern "C" int uda_tiebreak_stats(const void* scratch, long long n, unsigned long long* out3_host, cudaStream_t stream) {
    if (n < 0 || !out3_host) return UDA_E_INVALID;
    long long words = (n + 1) + (n + 1) + (n / 2 + 1);
    words = (words + 3) / 4 * 4;
    const unsigned long long* stats = (const unsigned long long*)((const uint32_t*)scratch + words);
    cudaError_t e = cudaMemcpyAsync(out3_host, stats, 24, cudaMemcpyDeviceToHost, stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
    return e == cudaSuccess ? 0 : (int)e;
}

extern "C" int uda_write_eof_at(void* dst, const unsigned long long* total_dev, cudaStream_t stream) {
    uda::eof_marker_kernel<<<1, 1, 0, stream>>>((uint8_t*)dst, total_dev);
    UDA_CHECK_LAUNCH();
    return 0;
}
