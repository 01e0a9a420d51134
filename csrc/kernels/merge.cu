// K4 — reducer-side streaming k-way merge of sorted runs, as sm_100a kernels.
//
// Reference behaviour being replaced (one CPU thread per reducer):
//   src/Merger/MergeQueue.h:126-270  binary min-heap of segments (put/top/pop/adjustTop)
//   src/Merger/MergeQueue.h:299-321  MergeQueue::next(): advance min segment, expose new top
//   src/Merger/StreamRW.cc:151-225   write_kv_to_stream(): the per-record hot loop
// A heap is inherently serial, so the device algorithm is different:
//
//   1. merge_range      : key range of all runs -> bin shift (2^20 bins over [kmin,kmax])
//   2. merge_hist       : histogram of entries over the bins (one pass over the 16 B entries)
//   3. scan             : exclusive prefix of the histogram
//   4. merge_partition  : k-way "merge path": bucket q owns output ranks [q*TARGET,(q+1)*TARGET) rounded
//                         to bin edges; per run the split position is a binary search. Bucket sizes
//                         are < TARGET + max bin count, independent of the key distribution shape.
//   5. merge_tiles      : per bucket, TMA bulk copies (cp.async.bulk + mbarrier complete_tx) pull the
//                         k run slices into shared memory, a pairwise merge-path tree merges them in
//                         shared memory, and the sorted tile is written out. Buckets larger than the
//                         tile (heavy duplicate keys) are split on the fly by 128-bit value bisection.
//
// Output order is (key, run, index-in-run): stable by run, one of the orders the reference heap can
// produce. The sorted entries drive the record gather in gather.cu.
#include "common.cuh"
#include "merge_tree.cuh"
#include "../../include/uda_b200.h"

namespace uda {

constexpr int MG_THREADS = 256;
constexpr int MG_VT = 8;
constexpr int MG_TILE = MG_THREADS * MG_VT;  // 2048 entries (32 KiB) per shared buffer
constexpr int MG_TARGET = 1536;              // target bucket size (slack absorbs bin granularity)
constexpr int MG_NB_LOG = 20;
constexpr int MG_NB = 1 << MG_NB_LOG;
constexpr int MG_KMAX = 4096;

struct MergeParams {
    unsigned long long kmin;
    unsigned long long kmax;
    int shift;
    int pad;
    unsigned long long fallback_buckets;  // statistics
    unsigned long long fallback_iters;
};

// ------------------------------------------------------------------------------------------- 1
__global__ void merge_range_kernel(const Entry* __restrict__ E, const long long* __restrict__ run_off, int k,
                                   MergeParams* __restrict__ prm) {
    __shared__ unsigned long long smin, smax;
    if (threadIdx.x == 0) { smin = ~0ull; smax = 0ull; }
    __syncthreads();
    for (int r = threadIdx.x; r < k; r += blockDim.x) {
        long long b = run_off[r], e = run_off[r + 1];
        if (e > b) {
            atomicMin(&smin, E[b].w0);
            atomicMax(&smax, E[e - 1].w0);
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long lo = smin, hi = smax;
        if (lo > hi) { lo = 0; hi = 0; }
        unsigned long long range = hi - lo;
        int shift = 0;
        while (shift < 64 && (range >> shift) >= (unsigned long long)MG_NB) ++shift;
        prm->kmin = lo;
        prm->kmax = hi;
        prm->shift = shift;
        prm->fallback_buckets = 0;
        prm->fallback_iters = 0;
    }
}

// ------------------------------------------------------------------------------------------- 2
__global__ void merge_hist_kernel(const Entry* __restrict__ E, long long n, const MergeParams* __restrict__ prm,
                                  uint32_t* __restrict__ hist) {
    const unsigned long long kmin = prm->kmin;
    const int shift = prm->shift;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        unsigned long long w0 = __ldg(&E[i].w0);
        atomicAdd(&hist[(w0 - kmin) >> shift], 1u);
    }
}

// ------------------------------------------------------------------------------------- 1+2 fused
// Fixed-stride front end: the key range comes straight from the first/last record of every run, so the
// histogram can be accumulated by the same pass that normalises the keys (saves one 16 B/entry pass).
__global__ void merge_range_records_kernel(const UdaRunDesc* __restrict__ runs, int k, int stride, int key_off, int key_len,
                                           MergeParams* __restrict__ prm) {
    __shared__ unsigned long long smin, smax;
    if (threadIdx.x == 0) { smin = ~0ull; smax = 0ull; }
    __syncthreads();
    for (int r = threadIdx.x; r < k; r += blockDim.x) {
        const UdaRunDesc rd = runs[r];
        if (rd.nrec > 0) {
            const uint8_t* base = (const uint8_t*)rd.base;
            unsigned long long w0;
            uint32_t t;
            key_prefix(base + key_off, key_len, &w0, &t);
            atomicMin(&smin, w0);
            key_prefix(base + (rd.nrec - 1) * stride + key_off, key_len, &w0, &t);
            atomicMax(&smax, w0);
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long lo = smin, hi = smax;
        if (lo > hi) { lo = 0; hi = 0; }
        unsigned long long range = hi - lo;
        int shift = 0;
        while (shift < 64 && (range >> shift) >= (unsigned long long)MG_NB) ++shift;
        prm->kmin = lo;
        prm->kmax = hi;
        prm->shift = shift;
        prm->fallback_buckets = 0;
        prm->fallback_iters = 0;
    }
}

__global__ void extract_hist_fixed_kernel(const UdaRunDesc* __restrict__ runs, int stride, int key_off, int key_len,
                                          const MergeParams* __restrict__ prm, Entry* __restrict__ out,
                                          uint32_t* __restrict__ hist) {
    const UdaRunDesc rd = runs[blockIdx.y];
    const uint8_t* base = (const uint8_t*)rd.base;
    Entry* o = out + rd.out_off;
    const unsigned long long hi_bits = ((unsigned long long)rd.run_id & 0xFFFFull) << 32;
    const unsigned long long kmin = prm->kmin;
    const int shift = prm->shift;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < rd.nrec; i += (long long)gridDim.x * blockDim.x) {
        unsigned long long w0;
        uint32_t t;
        key_prefix(base + i * stride + key_off, key_len, &w0, &t);
        Entry e;
        e.w0 = w0;
        e.w1 = ((unsigned long long)t << 48) | hi_bits | (unsigned long long)(uint32_t)(rd.idx0 + i);
        o[i] = e;
        atomicAdd(&hist[(w0 - kmin) >> shift], 1u);
    }
}

// ------------------------------------------------------------------------------------------- 3
// device-wide exclusive scan of m (<= 1024*1024) u32 values, three launches
constexpr int SC_THREADS = 1024;
__global__ void __launch_bounds__(SC_THREADS) scan_block_sums_kernel(const uint32_t* __restrict__ v, int m,
                                                                     uint32_t* __restrict__ sums) {
    __shared__ uint32_t ws[32];
    int i = blockIdx.x * SC_THREADS + threadIdx.x;
    uint32_t x = i < m ? v[i] : 0u;
    for (int o = 16; o; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
    if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = x;
    __syncthreads();
    if (threadIdx.x < 32) {
        uint32_t y = ws[threadIdx.x];
        for (int o = 16; o; o >>= 1) y += __shfl_xor_sync(0xffffffffu, y, o);
        if (threadIdx.x == 0) sums[blockIdx.x] = y;
    }
}
// single block: exclusive scan of nb (<= 1024) block sums in place
__global__ void __launch_bounds__(SC_THREADS) scan_sums_kernel(uint32_t* __restrict__ sums, int nb) {
    __shared__ uint32_t ws[32];
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    uint32_t s = tid < nb ? sums[tid] : 0u, x = s;
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
    }
    __syncthreads();
    if (tid < nb) sums[tid] = ws[w] + x - s;
}
// out[i] = exclusive prefix (out may alias v); out[m] = total is written by the last block
__global__ void __launch_bounds__(SC_THREADS) scan_apply_kernel(const uint32_t* __restrict__ v, int m,
                                                                const uint32_t* __restrict__ sums,
                                                                uint32_t* __restrict__ out) {
    __shared__ uint32_t ws[32];
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    int i = blockIdx.x * SC_THREADS + tid;
    uint32_t s = i < m ? v[i] : 0u, x = s;
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
    }
    __syncthreads();
    uint32_t excl = sums[blockIdx.x] + ws[w] + x - s;
    if (i < m) out[i] = excl;
    if (i == m - 1) out[m] = excl + s;
}

static int scan_u32_exclusive(const uint32_t* v, int m, uint32_t* out, uint32_t* sums, cudaStream_t stream) {
    int nb = (m + SC_THREADS - 1) / SC_THREADS;
    if (nb > SC_THREADS) return UDA_E_TOO_LARGE;
    scan_block_sums_kernel<<<nb, SC_THREADS, 0, stream>>>(v, m, sums);
    scan_sums_kernel<<<1, SC_THREADS, 0, stream>>>(sums, nb);
    scan_apply_kernel<<<nb, SC_THREADS, 0, stream>>>(v, m, sums, out);
    return 0;
}

// ------------------------------------------------------------------------------------------- 4
__device__ __forceinline__ uint32_t lower_bound_entries(const Entry* __restrict__ E, uint32_t lo, uint32_t hi,
                                                        unsigned long long w0, unsigned long long w1) {
    while (lo < hi) {
        uint32_t mid = lo + ((hi - lo) >> 1);
        unsigned long long m0 = __ldg(&E[mid].w0);
        bool less = m0 < w0;
        if (m0 == w0) less = __ldg(&E[mid].w1) < w1;
        if (less) lo = mid + 1; else hi = mid;
    }
    return lo;
}

// P[q*k + r] = first entry of run r that belongs to bucket >= q ; outoff[q] = output rank of bucket q
__global__ void merge_partition_kernel(const Entry* __restrict__ E, const long long* __restrict__ run_off, int k, int G,
                                       const uint32_t* __restrict__ cum, const MergeParams* __restrict__ prm,
                                       uint32_t* __restrict__ P, uint32_t* __restrict__ outoff) {
    const long long total = (long long)(G + 1) * k;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < total; t += (long long)gridDim.x * blockDim.x) {
        int q = (int)(t / k), r = (int)(t - (long long)q * k);
        // first bin whose exclusive prefix reaches q*TARGET
        unsigned long long want = (unsigned long long)q * MG_TARGET;
        int lo = 0, hi = MG_NB;  // cum has MG_NB+1 entries
        while (lo < hi) {
            int mid = (lo + hi) >> 1;
            if ((unsigned long long)__ldg(&cum[mid]) < want) lo = mid + 1; else hi = mid;
        }
        const int s = lo;
        uint32_t rb = (uint32_t)run_off[r], re = (uint32_t)run_off[r + 1];
        uint32_t pos;
        if (q == 0) pos = rb;
        else if (s >= MG_NB || q == G) pos = re;
        else {
            // boundary key; bins above the populated range saturate
            unsigned long long off = (unsigned long long)s << prm->shift;
            unsigned long long key = prm->kmin + off;
            bool overflow = (prm->shift > 0 && (off >> prm->shift) != (unsigned long long)s) || key < prm->kmin;
            pos = overflow ? re : lower_bound_entries(E, rb, re, key, 0ull);
        }
        P[t] = pos;
        if (r == 0) outoff[q] = (q == G || s >= MG_NB) ? cum[MG_NB] : (q == 0 ? 0u : cum[s]);
    }
}

// ------------------------------------------------------------------------------------------- 5
struct MergeSmem {
    Entry* bufA;
    Entry* bufB;
    uint32_t* lo;    // [k]   slice begin per run (absolute entry index)
    uint32_t* hi;    // [k]   slice end per run
    uint32_t* bend;  // [k]   bucket end per run (fallback only)
    uint32_t* offA;  // [k+1] list offsets in the current source buffer
    uint32_t* offB;  // [k+1]
};

// load slices [lo[r],hi[r]) of every run with TMA, merge them, write `total` sorted entries to out
__device__ __forceinline__ void merge_one_tile(const Entry* __restrict__ E, const MergeSmem& sm, int k, uint32_t total,
                                               Entry* __restrict__ out, uint64_t* mbar, uint32_t& phase) {
    const int rpt = (k + MG_THREADS - 1) / MG_THREADS;
    const int r0 = threadIdx.x * rpt, r1 = min(k, r0 + rpt);
    uint32_t mine = 0;
    for (int r = r0; r < r1; ++r) mine += sm.hi[r] - sm.lo[r];
    uint32_t tot;
    uint32_t base = block_exclusive_scan<MG_THREADS>(mine, &tot);
    for (int r = r0; r < r1; ++r) { sm.offA[r] = base; base += sm.hi[r] - sm.lo[r]; }
    if (threadIdx.x == 0) {
        sm.offA[k] = total;
        mbar_expect_tx(mbar, total * (uint32_t)sizeof(Entry));
    }
    fence_proxy_async_smem();  // order the previous tile's generic-proxy smem traffic before the TMA writes
    __syncthreads();
    for (int r = r0; r < r1; ++r) {
        uint32_t len = sm.hi[r] - sm.lo[r];
        if (len) tma_load_1d(sm.bufA + sm.offA[r], E + sm.lo[r], len * (uint32_t)sizeof(Entry), mbar);
    }
    mbar_wait(mbar, phase);
    phase ^= 1u;

    Entry* src = sm.bufA; Entry* dst = sm.bufB;
    uint32_t* soff = sm.offA; uint32_t* doff = sm.offB;
    int L = k;
    while (L > 1) {
        merge_level<MG_THREADS, MG_VT>(src, soff, L, dst, doff, total);
        __syncthreads();
        Entry* t = src; src = dst; dst = t;
        uint32_t* u = soff; soff = doff; doff = u;
        L = (L + 1) >> 1;
    }
    for (uint32_t i = threadIdx.x; i < total; i += MG_THREADS) out[i] = src[i];
    __syncthreads();  // buffers are reused by the next tile
}

__global__ void __launch_bounds__(MG_THREADS, 3)
merge_tiles_kernel(const Entry* __restrict__ E, int k, int G, const uint32_t* __restrict__ P,
                   const uint32_t* __restrict__ outoff, MergeParams* __restrict__ prm, Entry* __restrict__ out) {
    extern __shared__ __align__(128) uint8_t smem_raw[];
    __shared__ __align__(8) uint64_t mbar;
    __shared__ uint32_t s_cnt;
    MergeSmem sm;
    sm.bufA = reinterpret_cast<Entry*>(smem_raw);
    sm.bufB = sm.bufA + MG_TILE;
    sm.lo = reinterpret_cast<uint32_t*>(sm.bufB + MG_TILE);
    sm.hi = sm.lo + k;
    sm.bend = sm.hi + k;
    sm.offA = sm.bend + k;
    sm.offB = sm.offA + (k + 1);
    if (threadIdx.x == 0) { mbar_init(&mbar, 1); mbar_fence_init(); }
    __syncthreads();
    uint32_t phase = 0;

    for (int q = blockIdx.x; q < G; q += gridDim.x) {
        const uint32_t obase = outoff[q];
        const uint32_t total = outoff[q + 1] - obase;
        if (total == 0) continue;
        for (int r = threadIdx.x; r < k; r += MG_THREADS) {
            sm.lo[r] = P[(long long)q * k + r];
            sm.bend[r] = P[(long long)(q + 1) * k + r];
            sm.hi[r] = sm.bend[r];
        }
        __syncthreads();
        if (total <= MG_TILE) {
            merge_one_tile(E, sm, k, total, out + obase, &mbar, phase);
            continue;
        }
        // ---- oversize bucket (heavy duplicate keys): split by bisection over the 128-bit entry value
        if (threadIdx.x == 0) atomicAdd(&prm->fallback_buckets, 1ull);
        uint32_t remaining = total, done = 0;
        unsigned __int128 vlo = 0;  // count(vlo) == 0 w.r.t. the current lo[]
        {
            // lower bound of the bucket in value space: smallest head entry
            // (computed lazily: start from 0, which also has count 0)
        }
        while (remaining > 0) {
            uint32_t cnt;
            if (remaining <= (uint32_t)MG_TILE) {
                for (int r = threadIdx.x; r < k; r += MG_THREADS) sm.hi[r] = sm.bend[r];
                cnt = remaining;
                __syncthreads();
            } else {
                unsigned __int128 lo128 = vlo, hi128 = ~(unsigned __int128)0;
                unsigned __int128 best = vlo;
                uint32_t bestc = 0;
                int iters = 0;
                while (hi128 - lo128 > 1) {
                    unsigned __int128 mid = lo128 + ((hi128 - lo128) >> 1);
                    unsigned long long m0 = (unsigned long long)(mid >> 64), m1 = (unsigned long long)mid;
                    uint32_t c = 0;
                    for (int r = threadIdx.x; r < k; r += MG_THREADS)
                        c += lower_bound_entries(E, sm.lo[r], sm.bend[r], m0, m1) - sm.lo[r];
                    // block reduce
                    for (int o = 16; o; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
                    __syncthreads();
                    if (threadIdx.x == 0) s_cnt = 0;
                    __syncthreads();
                    if ((threadIdx.x & 31) == 0) atomicAdd(&s_cnt, c);
                    __syncthreads();
                    c = s_cnt;
                    ++iters;
                    if (c > (uint32_t)MG_TILE) hi128 = mid;
                    else {
                        lo128 = mid; best = mid; bestc = c;
                        if (c >= (uint32_t)MG_TILE / 2) break;
                    }
                }
                if (threadIdx.x == 0) atomicAdd(&prm->fallback_iters, (unsigned long long)iters);
                unsigned long long b0 = (unsigned long long)(best >> 64), b1 = (unsigned long long)best;
                for (int r = threadIdx.x; r < k; r += MG_THREADS)
                    sm.hi[r] = lower_bound_entries(E, sm.lo[r], sm.bend[r], b0, b1);
                cnt = bestc;
                vlo = best;
                __syncthreads();
            }
            if (cnt) merge_one_tile(E, sm, k, cnt, out + obase + done, &mbar, phase);
            done += cnt;
            remaining -= cnt;
            for (int r = threadIdx.x; r < k; r += MG_THREADS) sm.lo[r] = sm.hi[r];
            __syncthreads();
        }
    }
}

static inline int merge_num_buckets(long long n) { return (int)((n + MG_TARGET - 1) / MG_TARGET) + 1; }
static inline size_t align256(size_t x) { return (x + 255) & ~(size_t)255; }

struct MergeScratch {
    MergeParams* prm;
    uint32_t* hist;   // [NB+1] histogram then exclusive prefix in place
    uint32_t* sums;   // [1024]
    uint32_t* outoff; // [G+1]
    uint32_t* P;      // [(G+1)*k]
    size_t bytes;
};
static MergeScratch merge_scratch_layout(void* base, long long n, int k) {
    MergeScratch s;
    int G = merge_num_buckets(n);
    size_t off = 0;
    uint8_t* b = (uint8_t*)base;
    s.prm = (MergeParams*)(b + off); off += align256(sizeof(MergeParams));
    s.hist = (uint32_t*)(b + off); off += align256((size_t)(MG_NB + 1) * 4);
    s.sums = (uint32_t*)(b + off); off += align256(1024 * 4);
    s.outoff = (uint32_t*)(b + off); off += align256((size_t)(G + 2) * 4);
    s.P = (uint32_t*)(b + off); off += align256((size_t)(G + 1) * k * 4);
    s.bytes = off;
    return s;
}

}  // namespace uda

extern "C" long long uda_merge_scratch_bytes(long long n, int nruns) {
    if (n < 0 || nruns < 1) return UDA_E_INVALID;
    return (long long)uda::merge_scratch_layout(nullptr, n, nruns).bytes;
}

static int merge_back(const uda::Entry* E, const long long* run_off_dev, int nruns, long long n, void* sorted_out,
                      const uda::MergeScratch& s, cudaStream_t stream) {
    using namespace uda;
    const int G = merge_num_buckets(n);
    int rc = scan_u32_exclusive(s.hist, MG_NB, s.hist, s.sums, stream);
    if (rc) return rc;
    {
        long long t = (long long)(G + 1) * nruns;
        long long b = (t + 255) / 256;
        if (b > kNumSMs * 32) b = kNumSMs * 32;
        merge_partition_kernel<<<(int)b, 256, 0, stream>>>(E, run_off_dev, nruns, G, s.hist, s.prm, s.P, s.outoff);
    }
    {
        size_t smem = (size_t)2 * MG_TILE * sizeof(Entry) + ((size_t)5 * nruns + 2) * sizeof(uint32_t);
        UDA_CUDA(ensure_dynamic_smem((const void*)merge_tiles_kernel, smem));
        int per_sm = 0;
        UDA_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, merge_tiles_kernel, MG_THREADS, smem));
        if (per_sm < 1) return UDA_E_TOO_LARGE;
        int grid = kNumSMs * per_sm;
        if (grid > G) grid = G;
        merge_tiles_kernel<<<grid, MG_THREADS, smem, stream>>>(E, nruns, G, s.P, s.outoff, s.prm, (Entry*)sorted_out);
    }
    UDA_CHECK_LAUNCH();
    return 0;
}

extern "C" int uda_kway_merge_entries(const void* entries, const long long* run_off_dev, int nruns, long long n,
                                      void* sorted_out, void* scratch, cudaStream_t stream) {
    using namespace uda;
    if (nruns < 1 || nruns > MG_KMAX || n < 0 || n >= (1LL << 32) - MG_TILE) return UDA_E_INVALID;
    if (n == 0) return 0;
    const Entry* E = (const Entry*)entries;
    MergeScratch s = merge_scratch_layout(scratch, n, nruns);
    merge_range_kernel<<<1, 256, 0, stream>>>(E, run_off_dev, nruns, s.prm);
    UDA_CUDA(cudaMemsetAsync(s.hist, 0, (size_t)(MG_NB + 1) * 4, stream));
    {
        long long b = (n + 255) / 256;
        if (b > kNumSMs * 16) b = kNumSMs * 16;
        merge_hist_kernel<<<(int)b, 256, 0, stream>>>(E, n, s.prm, s.hist);
    }
    return merge_back(E, run_off_dev, nruns, n, sorted_out, s, stream);
}

// Fused fixed-stride pipeline: key normalise + histogram in one pass over the landed records, then the
// merge proper. entries_tmp and sorted_out hold n entries each.
extern "C" int uda_merge_runs_fixed(const UdaRunDesc* runs_dev, const long long* run_off_dev, int nruns, long long n,
                                    long long max_nrec, int stride, int key_off, int key_len, void* entries_tmp,
                                    void* sorted_out, void* scratch, cudaStream_t stream) {
    using namespace uda;
    if (nruns < 1 || nruns > MG_KMAX || n < 0 || n >= (1LL << 32) - MG_TILE) return UDA_E_INVALID;
    if (n == 0) return 0;
    Entry* E = (Entry*)entries_tmp;
    MergeScratch s = merge_scratch_layout(scratch, n, nruns);
    merge_range_records_kernel<<<1, 256, 0, stream>>>(runs_dev, nruns, stride, key_off, key_len, s.prm);
    UDA_CUDA(cudaMemsetAsync(s.hist, 0, (size_t)(MG_NB + 1) * 4, stream));
    {
        long long bx = (max_nrec + 255) / 256;
        long long cap = (kNumSMs * 16 + nruns - 1) / nruns;
        if (cap < 1) cap = 1;
        if (bx > cap) bx = cap;
        if (bx < 1) bx = 1;
        dim3 grid((unsigned)bx, (unsigned)nruns);
        extract_hist_fixed_kernel<<<grid, 256, 0, stream>>>(runs_dev, stride, key_off, key_len, s.prm, E, s.hist);
    }
    return merge_back(E, run_off_dev, nruns, n, sorted_out, s, stream);
}

extern "C" int uda_merge_stats(const void* scratch, unsigned long long* out2_host, cudaStream_t stream) {
    const uda::MergeParams* p = (const uda::MergeParams*)scratch;
    UDA_CUDA(cudaMemcpyAsync(out2_host, &p->fallback_buckets, 2 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, stream));
    UDA_CUDA(cudaStreamSynchronize(stream));
    return 0;
}
