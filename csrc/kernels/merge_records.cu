// K4 (fixed-stride fast path) — fused k-way merge that moves every record exactly once.
//
// Reference hot loop being replaced: write_kv_to_stream (src/Merger/StreamRW.cc:151-225) pops the heap
// (src/Merger/MergeQueue.h:299-321,390-419) and serialises one record per iteration on one CPU thread.
//
// The entry-based pipeline in merge.cu reads every record three times (key normalise, gather) and
// writes 2 x 16 B of entries per record. For fixed-stride records this kernel does the whole merge with
// one read and one write of each record:
//
//   1. fr_sample_hist  : every FR_SAMPLE-th key of every run -> 2^22-bin histogram (sector-sparse read)
//   2. scan            : bucket q starts at the first bin whose sample prefix reaches q*TARGET/FR_SAMPLE
//   3. fr_partition    : exact per-run split position of every bucket boundary (binary search on the keys
//                        inside the records) — the k-way merge-path partition
//   4. fr_sizes + scan : exact bucket sizes -> output offsets. Regular sampling bounds the estimate error
//                        by FR_SAMPLE records per run per boundary, so buckets fit the tile by construction
//                        for small k and with overwhelming probability for large k; oversize buckets take the
//                        bisection path below.
//   5. fr_merge        : per bucket, TMA bulk copies (cp.async.bulk, mbarrier expect_tx/complete_tx) land the
//                        k record slices in shared memory; 16-byte entries (key prefix, slot) are built from
//                        the staged records; the pairwise merge-path tree orders them; records leave shared
//                        memory in merged order as coalesced stores straight to their final position.
//
// Output order is (key, run, index-in-run), identical to merge.cu + gather.cu.
#include "common.cuh"
#include "merge_tree.cuh"
#include "scan.cuh"
#include "../../include/uda_b200.h"

namespace uda {

// Tuning (measured on B200, 125 M records, k = 8; ms per TeraSort step incl. the 4.4 ms push):
//   VT=3 / 2 CTAs per SM 13.95   VT=2 / 3 CTAs 13.53   VT=4 / 1 CTA 17.76   VT=2 / 3 CTAs, 128 merge threads 13.91
#ifndef UDA_FR_VT
#define UDA_FR_VT 2
#endif
#ifndef UDA_FR_OUT32
#define UDA_FR_OUT32 0   // measured: 9.23 ms vs 8.95 ms for the 16-byte path on the 125 M record merge
#endif
#ifndef UDA_FR_MERGE_THREADS
#define UDA_FR_MERGE_THREADS UDA_FR_THREADS
#endif
#ifndef UDA_FR_CTAS
#define UDA_FR_CTAS 3
#endif
#ifndef UDA_FR_THREADS
#define UDA_FR_THREADS 256
#endif
constexpr int FR_THREADS = UDA_FR_THREADS;
constexpr int FR_VT = UDA_FR_VT;
constexpr int FR_TILE = FR_THREADS * FR_VT;  // record slots per tile (512 by default)
constexpr int FR_MERGE_THREADS = UDA_FR_MERGE_THREADS;   // threads that run the merge tree (fewer = more items per thread)
constexpr int FR_MERGE_VT = FR_TILE / FR_MERGE_THREADS;
constexpr int FR_TARGET = FR_TILE * 3 / 4;   // target bucket size: a quarter of the tile absorbs the sampling error
constexpr int FR_SAMPLE = 8;
constexpr int FR_NB_LOG = 22;
constexpr int FR_NB = 1 << FR_NB_LOG;
constexpr int FR_KMAX = 1024;
constexpr int FR_SLICE_PAD = 48;             // bytes of shared memory reserved per run for 16-byte alignment slop

struct FrParams {
    unsigned long long kmin, kmax;
    int shift, pad;
    unsigned long long fallback_buckets, fallback_iters;
    unsigned int next_bucket, pad2;   // dynamic bucket scheduler of fr_merge_kernel (reset by fr_range_kernel)
};

__device__ __forceinline__ void load_key80(const uint8_t* p, int key_len, unsigned long long* w0, uint32_t* t) {
    key_prefix(p, key_len, w0, t);
}

// first record index in [lo,hi) of the run whose key is >= (w0,t)
__device__ __forceinline__ uint32_t lb_records(const uint8_t* __restrict__ base, int stride, int key_off, int key_len,
                                               uint32_t lo, uint32_t hi, unsigned long long w0, uint32_t t) {
    while (lo < hi) {
        uint32_t mid = lo + ((hi - lo) >> 1);
        unsigned long long m0;
        uint32_t mt;
        load_key80(base + (unsigned long long)mid * stride + key_off, key_len, &m0, &mt);
        bool less = m0 < w0 || (m0 == w0 && mt < t);
        if (less) lo = mid + 1; else hi = mid;
    }
    return lo;
}
// first record index whose key is > (w0,t)
__device__ __forceinline__ uint32_t ub_records(const uint8_t* __restrict__ base, int stride, int key_off, int key_len,
                                               uint32_t lo, uint32_t hi, unsigned long long w0, uint32_t t) {
    while (lo < hi) {
        uint32_t mid = lo + ((hi - lo) >> 1);
        unsigned long long m0;
        uint32_t mt;
        load_key80(base + (unsigned long long)mid * stride + key_off, key_len, &m0, &mt);
        bool le = m0 < w0 || (m0 == w0 && mt <= t);
        if (le) lo = mid + 1; else hi = mid;
    }
    return lo;
}

// ------------------------------------------------------------------------------------------- 0
__global__ void fr_range_kernel(const UdaRunDesc* __restrict__ runs, int k, int stride, int key_off, int key_len,
                                FrParams* __restrict__ prm) {
    __shared__ unsigned long long smin, smax;
    if (threadIdx.x == 0) { smin = ~0ull; smax = 0ull; }
    __syncthreads();
    for (int r = threadIdx.x; r < k; r += blockDim.x) {
        const UdaRunDesc rd = runs[r];
        if (rd.nrec > 0) {
            const uint8_t* base = (const uint8_t*)rd.base;
            unsigned long long w0;
            uint32_t t;
            load_key80(base + key_off, key_len, &w0, &t);
            atomicMin(&smin, w0);
            load_key80(base + (rd.nrec - 1) * stride + key_off, key_len, &w0, &t);
            atomicMax(&smax, w0);
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long lo = smin, hi = smax;
        if (lo > hi) { lo = 0; hi = 0; }
        unsigned long long range = hi - lo;
        int shift = 0;
        while (shift < 64 && (range >> shift) >= (unsigned long long)FR_NB) ++shift;
        prm->kmin = lo; prm->kmax = hi; prm->shift = shift;
        prm->fallback_buckets = 0; prm->fallback_iters = 0;
        prm->next_bucket = 0;
    }
}

// ------------------------------------------------------------------------------------------- 1
// samples at record positions FR_SAMPLE-1, 2*FR_SAMPLE-1, ...: c samples below x  =>  between
// c*FR_SAMPLE and c*FR_SAMPLE + FR_SAMPLE - 1 records of that run below x
__global__ void fr_sample_hist_kernel(const UdaRunDesc* __restrict__ runs, int stride, int key_off, int key_len,
                                      const FrParams* __restrict__ prm, uint32_t* __restrict__ hist) {
    const UdaRunDesc rd = runs[blockIdx.y];
    const uint8_t* base = (const uint8_t*)rd.base;
    const unsigned long long kmin = prm->kmin;
    const int shift = prm->shift;
    const long long ns = rd.nrec / FR_SAMPLE;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < ns; i += (long long)gridDim.x * blockDim.x) {
        unsigned long long w0;
        uint32_t t;
        load_key80(base + (i * FR_SAMPLE + FR_SAMPLE - 1) * stride + key_off, key_len, &w0, &t);
        atomicAdd(&hist[(w0 - kmin) >> shift], 1u);
    }
}

// ------------------------------------------------------------------------------------------- 3
// boundary key of every bucket: bkey[q] = first key of bucket q (w0 only, tail 0); kind[q]: 0 = start of the
// run (q == 0), 1 = search for bkey[q], 2 = end of the run (past the last sample / q == G)
__global__ void fr_bkeys_kernel(int G, const uint32_t* __restrict__ cum, const FrParams* __restrict__ prm,
                                unsigned long long* __restrict__ bkey, uint8_t* __restrict__ kind) {
    for (int q = blockIdx.x * blockDim.x + threadIdx.x; q <= G; q += gridDim.x * blockDim.x) {
        const unsigned long long want = (unsigned long long)q * (FR_TARGET / FR_SAMPLE);
        int lo = 0, hi = FR_NB;
        while (lo < hi) {
            int mid = (lo + hi) >> 1;
            if ((unsigned long long)__ldg(&cum[mid]) < want) lo = mid + 1; else hi = mid;
        }
        uint8_t kd = 1;
        unsigned long long key = 0;
        if (q == 0) kd = 0;
        else if (lo >= FR_NB || q == G) kd = 2;
        else {
            key = prm->kmin + ((unsigned long long)lo << prm->shift);
            if (key < prm->kmin) kd = 2;  // wrapped past the top of the key space
        }
        bkey[q] = key;
        kind[q] = kd;
    }
}

// Per-boundary binary search: (G+1)*k independent searches over the strided keys. Cheapest for few runs
// (k <= 16: 0.4 ms at k = 8), where a boundary falls only every ~50 records of a run.
__global__ void fr_partition_search_kernel(const UdaRunDesc* __restrict__ runs, int k, int G, int stride, int key_off,
                                           int key_len, const unsigned long long* __restrict__ bkey,
                                           const uint8_t* __restrict__ kind, uint32_t* __restrict__ P) {
    const long long total = (long long)(G + 1) * k;
    for (long long x = (long long)blockIdx.x * blockDim.x + threadIdx.x; x < total; x += (long long)gridDim.x * blockDim.x) {
        const int q = (int)(x / k), r = (int)(x - (long long)q * k);
        const UdaRunDesc rd = runs[r];
        const uint8_t kd = kind[q];
        uint32_t pos = kd == 0 ? 0u : (uint32_t)rd.nrec;
        if (kd == 1) pos = lb_records((const uint8_t*)rd.base, stride, key_off, key_len, 0u, (uint32_t)rd.nrec, bkey[q], 0u);
        P[x] = pos;
    }
}

// Exact split position of every bucket boundary in every run: the k-way merge-path partition.
//
// A boundary is a bin edge, so "first record of run r in bucket >= q" can be read off a single pass over
// the run's keys: record i lies in bucket b_i = cum[bin(key_i)] / (TARGET/SAMPLE) (O(1), the prefix array is
// L2 resident), b_i is non-decreasing in i, and P[q][r] = i for every q in (b_{i-1}, b_i]. One coalesced,
// sector-sparse pass (2 ms for 125 M records) replaces (G+1)*k independent binary searches over strided
// records, which for k = 64 dragged the whole data set through DRAM at 2 TB/s (6.3 ms). Runs that are much
// shorter than the bucket count keep the per-boundary binary search (their gaps would serialise one thread).
__global__ void fr_partition_kernel(const UdaRunDesc* __restrict__ runs, int k, int G, int stride, int key_off, int key_len,
                                    const uint32_t* __restrict__ cum, const FrParams* __restrict__ prm,
                                    const unsigned long long* __restrict__ bkey, const uint8_t* __restrict__ kind,
                                    uint32_t* __restrict__ P) {
    const int r = blockIdx.y;
    const UdaRunDesc rd = runs[r];
    const uint8_t* base = (const uint8_t*)rd.base;
    const long long n = rd.nrec;
    const long long nthreads = (long long)gridDim.x * blockDim.x;
    const long long t0 = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (n * 4 < (long long)G) {  // short run: one binary search per boundary
        for (long long q = t0; q <= G; q += nthreads) {
            const uint8_t kd = kind[q];
            uint32_t pos = kd == 0 ? 0u : (uint32_t)n;
            if (kd == 1) pos = lb_records(base, stride, key_off, key_len, 0u, (uint32_t)n, bkey[q], 0u);
            P[q * k + r] = pos;
        }
        return;
    }
    const unsigned long long kmin = prm->kmin;
    const int shift = prm->shift;
    const uint32_t per = FR_TARGET / FR_SAMPLE;
    auto bucket_of = [&](long long i) -> int {
        unsigned long long w0;
        uint32_t t;
        load_key80(base + i * stride + key_off, key_len, &w0, &t);
        const uint32_t bq = __ldg(&cum[(w0 - kmin) >> shift]) / per;
        return (int)(bq < (uint32_t)(G - 1) ? bq : (uint32_t)(G - 1));
    };
    const int lane = threadIdx.x & 31;
    for (long long wi = t0 - lane; wi < n; wi += nthreads) {  // warp-uniform trip count: the shuffle needs every lane
        const long long i = wi + lane;
        const bool valid = i < n;
        const int bi = valid ? bucket_of(i) : G;
        int bp = __shfl_up_sync(0xffffffffu, bi, 1);
        if (lane == 0) bp = wi > 0 ? bucket_of(wi - 1) : -1;
        if (valid) {
            for (int q = bp + 1; q <= bi; ++q) P[(long long)q * k + r] = (uint32_t)i;
            if (i == n - 1)
                for (int q = bi + 1; q <= G; ++q) P[(long long)q * k + r] = (uint32_t)n;
        }
    }
}

// ------------------------------------------------------------------------------------------- 4
__global__ void fr_sizes_kernel(const uint32_t* __restrict__ P, int k, int G, uint32_t* __restrict__ sizes) {
    const int lane = threadIdx.x & 31;
    const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    for (long long q = warp; q < G; q += nwarps) {
        uint32_t s = 0;
        for (int r = lane; r < k; r += 32) s += P[(q + 1) * k + r] - P[q * k + r];
        for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
        if (lane == 0) sizes[q] = s;
    }
}

// ------------------------------------------------------------------------------------------- 5
struct FrSmem {
    uint8_t* rec;        // staged record bytes
    Entry* entA;
    Entry* entB;
    uint32_t* slotaddr;  // [FR_TILE] byte offset of slot j's record inside rec
    uint32_t* lo;        // [k]
    uint32_t* hi;        // [k]
    uint32_t* bend;      // [k]
    uint32_t* sbase;     // [k] byte offset of the first record of run r's slice inside rec
    uint32_t* soff;      // [k+1] slot offsets (list boundaries for the merge tree)
    uint32_t* doff;      // [k+1]
};

__host__ __device__ inline size_t fr_smem_bytes(int k, int stride) {
    size_t rec = ((size_t)FR_TILE * stride + (size_t)FR_SLICE_PAD * k + 64 + 15) & ~(size_t)15;
    return rec + 2 * (size_t)FR_TILE * sizeof(Entry) + (size_t)FR_TILE * 4 + ((size_t)4 * k + 2 * ((size_t)k + 1)) * 4 + 16;
}

__device__ __forceinline__ void mbar_expect_tx_only(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.expect_tx.relaxed.cta.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}

template <bool ALIGN2, int STRIDE>
__device__ __forceinline__ void fr_process_tile(const UdaRunDesc* __restrict__ runs, const FrSmem& sm, int k, int stride_rt,
                                                int key_off, int key_len, uint32_t total, uint8_t* __restrict__ out,
                                                uint64_t* mbar, uint32_t& phase) {
    const int stride = STRIDE > 0 ? STRIDE : stride_rt;
    const int tid = threadIdx.x;
    const int rpt = (k + FR_THREADS - 1) / FR_THREADS;
    const int r0 = tid * rpt, r1 = min(k, r0 + rpt);
    uint32_t mine = 0;
    for (int r = r0; r < r1; ++r) mine += sm.hi[r] - sm.lo[r];
    uint32_t tot;
    uint32_t slot = block_exclusive_scan<FR_THREADS>(mine, &tot);
    fence_proxy_async_smem();  // previous tile's generic-proxy accesses to `rec` are ordered before the TMA writes
    for (int r = r0; r < r1; ++r) {
        const uint32_t len = sm.hi[r] - sm.lo[r];
        sm.soff[r] = slot;
        const uint32_t B = ((slot * (uint32_t)stride + (uint32_t)FR_SLICE_PAD * r) + 15u) & ~15u;
        if (len) {
            const uint8_t* src = (const uint8_t*)runs[r].base + (unsigned long long)sm.lo[r] * stride;
            const uint32_t mis = (uint32_t)((uintptr_t)src & 15);
            const uint32_t nbytes = (mis + len * (uint32_t)stride + 15u) & ~15u;
            sm.sbase[r] = B + mis;
            mbar_expect_tx_only(mbar, nbytes);
            tma_load_1d(sm.rec + B, src - mis, nbytes, mbar);
        } else {
            sm.sbase[r] = B;
        }
        slot += len;
    }
    if (tid == 0) sm.soff[k] = total;
    __syncthreads();               // every expect_tx precedes the single arrival
    if (tid == 0) mbar_arrive(mbar);
    mbar_wait(mbar, phase);
    phase ^= 1u;

    // entries from the staged records
    for (uint32_t j = tid; j < total; j += FR_THREADS) {
        int lo = 0, hi = k - 1;  // run owning slot j: last r with soff[r] <= j
        while (lo < hi) {
            int mid = (lo + hi + 1) >> 1;
            if (sm.soff[mid] <= j) lo = mid; else hi = mid - 1;
        }
        const uint32_t addr = sm.sbase[lo] + (j - sm.soff[lo]) * (uint32_t)stride;
        const uint8_t* kp = sm.rec + addr + key_off;
        unsigned long long w0 = 0;
        uint32_t t = 0;
        if (ALIGN2 && key_len == 10) {
            const uint16_t* h = reinterpret_cast<const uint16_t*>(kp);
            const uint32_t h0 = h[0], h1 = h[1], h2 = h[2], h3 = h[3], h4 = h[4];
            // bytes are big-endian key digits; each little-endian half holds (b0 | b1 << 8)
            auto sw = [](uint32_t x) { return ((x & 0xFFu) << 8) | (x >> 8); };
            w0 = ((unsigned long long)sw(h0) << 48) | ((unsigned long long)sw(h1) << 32) | ((unsigned long long)sw(h2) << 16) | sw(h3);
            t = sw(h4);
        } else {
            key_prefix(kp, key_len, &w0, &t);
        }
        Entry e;
        e.w0 = w0;
        e.w1 = ((unsigned long long)t << 48) | j;
        sm.entA[j] = e;
        sm.slotaddr[j] = addr;
    }
    __syncthreads();

    Entry* src = sm.entA; Entry* dst = sm.entB;
    uint32_t* soff = sm.soff; uint32_t* doff = sm.doff;
    int L = k;
    while (L > 1) {
        if (tid < FR_MERGE_THREADS) merge_level<FR_MERGE_THREADS, FR_MERGE_VT>(src, soff, L, dst, doff, total);
        __syncthreads();
        Entry* te = src; src = dst; dst = te;
        uint32_t* tu = soff; soff = doff; doff = tu;
        L = (L + 1) >> 1;
    }

    // Records leave in merged order. One thread per 16-byte *destination* word (all 32 lanes busy, the
    // per-record work is amortised over the lanes): the word's bytes come from at most two staged records at
    // arbitrary (even) shared-memory offsets, each fetched as aligned 32-bit shared loads + funnel shifts.
    // Shared-space addresses are explicit so the loads are LDS, not generic LD.
    const uint32_t rec_s = smem_u32(sm.rec), ent_s = smem_u32(src), slot_s = smem_u32(sm.slotaddr);
    auto slot_addr = [&](uint32_t p) -> uint32_t {  // shared address of merged record p
        uint32_t j, sa;
        asm volatile("ld.shared.u16 %0, [%1];" : "=r"(j) : "r"(ent_s + p * 16u + 8u));
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(sa) : "r"(slot_s + j * 4u));
        return rec_s + sa;
    };
    auto lds16 = [](uint32_t a) -> uint4 {  // 16 bytes at arbitrary shared address a
        const uint32_t w = a & ~3u, sh = (a & 3u) * 8u;
        uint32_t x0, x1, x2, x3, x4;
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(x0) : "r"(w));
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(x1) : "r"(w + 4u));
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(x2) : "r"(w + 8u));
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(x3) : "r"(w + 12u));
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(x4) : "r"(w + 16u));
        uint4 r;
        r.x = __funnelshift_r(x0, x1, sh); r.y = __funnelshift_r(x1, x2, sh);
        r.z = __funnelshift_r(x2, x3, sh); r.w = __funnelshift_r(x3, x4, sh);
        return r;
    };
    const uint32_t span = total * (uint32_t)stride;
    const uint32_t head = (uint32_t)((16u - ((uintptr_t)out & 15u)) & 15u) < span ? (uint32_t)((16u - ((uintptr_t)out & 15u)) & 15u) : span;
    const uint32_t nfull = (span - head) >> 4;
    auto emit16 = [&](uint32_t w) {
        const uint32_t rel = head + (w << 4);
        const uint32_t p = rel / (uint32_t)stride;
        const uint32_t o = rel - p * (uint32_t)stride;
        const uint32_t len0 = (uint32_t)stride - o < 16u ? (uint32_t)stride - o : 16u;
        uint4 v = lds16(slot_addr(p) + o);
        if (len0 < 16u) {  // the word straddles two records (stride >= 16: never three)
            const uint4 u = lds16(slot_addr(p + 1) - len0);
            // byte k of the word comes from v for k < len0, from u otherwise: one 128-bit mask, 4 LOP3 selects
            const unsigned long long mlo = len0 >= 8u ? ~0ull : ((1ull << (8u * len0)) - 1ull);
            const unsigned long long mhi = len0 <= 8u ? 0ull : ((1ull << (8u * (len0 - 8u))) - 1ull);
            const uint32_t m0 = (uint32_t)mlo, m1 = (uint32_t)(mlo >> 32), m2 = (uint32_t)mhi, m3 = (uint32_t)(mhi >> 32);
            v.x = (v.x & m0) | (u.x & ~m0); v.y = (v.y & m1) | (u.y & ~m1);
            v.z = (v.z & m2) | (u.z & ~m2); v.w = (v.w & m3) | (u.w & ~m3);
        }
        st_stream(reinterpret_cast<int4*>(out + rel), make_int4((int)v.x, (int)v.y, (int)v.z, (int)v.w));
    };
#if UDA_FR_OUT32
    // 32 bytes per thread and iteration (strides >= 40): the record look-up, the division and the straddle handling
    // are paid once per 32 bytes instead of once per 16
    auto lds32 = [](uint32_t a, uint32_t (&r)[8]) {  // 32 bytes at arbitrary shared address a
        const uint32_t w = a & ~3u, sh = (a & 3u) * 8u;
        uint32_t x[9];
#pragma unroll
        for (int i = 0; i < 9; ++i) asm volatile("ld.shared.u32 %0, [%1];" : "=r"(x[i]) : "r"(w + 4u * i));
#pragma unroll
        for (int i = 0; i < 8; ++i) r[i] = __funnelshift_r(x[i], x[i + 1], sh);
    };
    const uint32_t nunits = stride >= 40 ? (nfull >> 1) : 0u;
    for (uint32_t un = tid; un < nunits; un += FR_THREADS) {
        const uint32_t rel = head + (un << 5);
        const uint32_t p = rel / (uint32_t)stride;
        const uint32_t o = rel - p * (uint32_t)stride;
        const uint32_t len0 = (uint32_t)stride - o < 32u ? (uint32_t)stride - o : 32u;
        uint32_t v[8];
        lds32(slot_addr(p) + o, v);
        if (len0 < 32u) {  // the 32 bytes straddle two records (stride >= 40: never three)
            uint32_t u[8];
            lds32(slot_addr(p + 1) - len0, u);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int rem = (int)len0 - 4 * i;   // bytes of word i that still belong to the first record
                const uint32_t m = rem >= 4 ? 0xFFFFFFFFu : rem <= 0 ? 0u : ((1u << (8 * rem)) - 1u);
                v[i] = (v[i] & m) | (u[i] & ~m);
            }
        }
        st_stream(reinterpret_cast<int4*>(out + rel), make_int4((int)v[0], (int)v[1], (int)v[2], (int)v[3]));
        st_stream(reinterpret_cast<int4*>(out + rel + 16), make_int4((int)v[4], (int)v[5], (int)v[6], (int)v[7]));
    }
    for (uint32_t w = (nunits << 1) + tid; w < nfull; w += FR_THREADS) emit16(w);
#else
#pragma unroll 2
    for (uint32_t w = tid; w < nfull; w += FR_THREADS) emit16(w);
#endif
    // unaligned head and tail bytes of the tile's span (< 16 bytes each)
    {
        const uint32_t tail0 = head + (nfull << 4);
        const uint32_t nedge = head + (span - tail0);
        for (uint32_t e = tid; e < nedge; e += FR_THREADS) {
            const uint32_t rel = e < head ? e : tail0 + (e - head);
            const uint32_t p = rel / (uint32_t)stride, o = rel - p * (uint32_t)stride;
            uint32_t v;
            asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(slot_addr(p) + o));
            out[rel] = (uint8_t)v;
        }
    }
    __syncthreads();  // buffers are reused by the next tile
}

template <bool ALIGN2, int STRIDE>
__global__ void __launch_bounds__(FR_THREADS, UDA_FR_CTAS)
fr_merge_kernel(const UdaRunDesc* __restrict__ runs, int k, int G, int stride_rt, int key_off, int key_len,
                const uint32_t* __restrict__ P, const uint32_t* __restrict__ outoff, FrParams* __restrict__ prm,
                uint8_t* __restrict__ out) {
    const int stride = STRIDE > 0 ? STRIDE : stride_rt;
    extern __shared__ __align__(128) uint8_t fr_smem_raw[];
    __shared__ __align__(8) uint64_t mbar;
    __shared__ uint32_t s_cnt;
    FrSmem sm;
    const size_t rec_bytes = ((size_t)FR_TILE * stride + (size_t)FR_SLICE_PAD * k + 64 + 15) & ~(size_t)15;
    sm.rec = fr_smem_raw;
    sm.entA = reinterpret_cast<Entry*>(fr_smem_raw + rec_bytes);
    sm.entB = sm.entA + FR_TILE;
    sm.slotaddr = reinterpret_cast<uint32_t*>(sm.entB + FR_TILE);
    sm.lo = sm.slotaddr + FR_TILE;
    sm.hi = sm.lo + k;
    sm.bend = sm.hi + k;
    sm.sbase = sm.bend + k;
    sm.soff = sm.sbase + k;
    sm.doff = sm.soff + (k + 1);
    if (threadIdx.x == 0) { mbar_init(&mbar, 1); mbar_fence_init(); }
    __syncthreads();
    uint32_t phase = 0;

    // Buckets are handed out through a counter, not by blockIdx stride: when part of the SMs is held by a
    // concurrently running push kernel, the CTAs that do get an SM still share the work evenly.
    __shared__ int s_q;
    for (;;) {
        if (threadIdx.x == 0) s_q = (int)atomicAdd(&prm->next_bucket, 1u);
        __syncthreads();
        const int q = s_q;
        __syncthreads();
        if (q >= G) break;
        const uint32_t obase = outoff[q];
        const uint32_t total = outoff[q + 1] - obase;
        if (total == 0) continue;
        for (int r = threadIdx.x; r < k; r += FR_THREADS) {
            sm.lo[r] = P[(long long)q * k + r];
            sm.bend[r] = P[(long long)(q + 1) * k + r];
            sm.hi[r] = sm.bend[r];
        }
        __syncthreads();
        if (total <= (uint32_t)FR_TILE) {
            fr_process_tile<ALIGN2, STRIDE>(runs, sm, k, stride, key_off, key_len, total, out + (unsigned long long)obase * stride, &mbar, phase);
            continue;
        }
        // ---- oversize bucket: split by bisection over the 128-bit value (key80 | run | index)
        if (threadIdx.x == 0) atomicAdd(&prm->fallback_buckets, 1ull);
        uint32_t remaining = total, done = 0;
        unsigned __int128 vlo = 0;
        while (remaining > 0) {
            uint32_t cnt;
            if (remaining <= (uint32_t)FR_TILE) {
                for (int r = threadIdx.x; r < k; r += FR_THREADS) sm.hi[r] = sm.bend[r];
                cnt = remaining;
                __syncthreads();
            } else {
                unsigned __int128 lo128 = vlo, hi128 = ~(unsigned __int128)0, best = vlo;
                uint32_t bestc = 0;
                int iters = 0;
                auto split_pos = [&](int r, unsigned __int128 v) -> uint32_t {
                    const unsigned long long v0 = (unsigned long long)(v >> 64), v1 = (unsigned long long)v;
                    const uint32_t t = (uint32_t)(v1 >> 48), vrun = (uint32_t)(v1 >> 32) & 0xFFFFu, vidx = (uint32_t)v1;
                    const uint8_t* base = (const uint8_t*)runs[r].base;
                    uint32_t lb = lb_records(base, stride, key_off, key_len, sm.lo[r], sm.bend[r], v0, t);
                    const uint32_t rid = (uint32_t)r;  // slots are laid out run-major, so the run *index* is the tie-break rank
                    if (rid > vrun) return lb;
                    uint32_t ub = ub_records(base, stride, key_off, key_len, lb, sm.bend[r], v0, t);
                    if (rid < vrun) return ub;
                    uint32_t x = vidx < lb ? lb : (vidx > ub ? ub : vidx);
                    return x;
                };
                while (hi128 - lo128 > 1) {
                    unsigned __int128 mid = lo128 + ((hi128 - lo128) >> 1);
                    uint32_t c = 0;
                    for (int r = threadIdx.x; r < k; r += FR_THREADS) c += split_pos(r, mid) - sm.lo[r];
                    for (int o = 16; o; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
                    __syncthreads();
                    if (threadIdx.x == 0) s_cnt = 0;
                    __syncthreads();
                    if ((threadIdx.x & 31) == 0) atomicAdd(&s_cnt, c);
                    __syncthreads();
                    c = s_cnt;
                    ++iters;
                    if (c > (uint32_t)FR_TILE) hi128 = mid;
                    else {
                        lo128 = mid; best = mid; bestc = c;
                        if (c >= (uint32_t)FR_TILE / 2) break;
                    }
                }
                if (threadIdx.x == 0) atomicAdd(&prm->fallback_iters, (unsigned long long)iters);
                for (int r = threadIdx.x; r < k; r += FR_THREADS) sm.hi[r] = split_pos(r, best);
                cnt = bestc;
                vlo = best;
                __syncthreads();
            }
            if (cnt) fr_process_tile<ALIGN2, STRIDE>(runs, sm, k, stride, key_off, key_len, cnt, out + (unsigned long long)(obase + done) * stride, &mbar, phase);
            done += cnt;
            remaining -= cnt;
            for (int r = threadIdx.x; r < k; r += FR_THREADS) sm.lo[r] = sm.hi[r];
            __syncthreads();
        }
    }
}

__global__ void fr_eof_kernel(uint8_t* dst) { dst[0] = 0xFF; dst[1] = 0xFF; }

static inline int fr_num_buckets(long long n) { return (int)(n / FR_TARGET) + 2; }
static inline size_t fr_al(size_t x) { return (x + 255) & ~(size_t)255; }

struct FrScratch {
    FrParams* prm;
    uint32_t* hist;    // [FR_NB+1]
    uint32_t* sums;    // [1024]
    uint32_t* sizes;   // [G+1] then exclusive prefix in place (= outoff)
    uint32_t* P;       // [(G+1)*k]
    unsigned long long* bkey;  // [G+1]
    uint8_t* kind;     // [G+1]
    size_t bytes;
};
static FrScratch fr_layout(void* base, long long n, int k) {
    FrScratch s;
    const int G = fr_num_buckets(n);
    uint8_t* b = (uint8_t*)base;
    size_t off = 0;
    s.prm = (FrParams*)(b + off); off += fr_al(sizeof(FrParams));
    s.hist = (uint32_t*)(b + off); off += fr_al((size_t)(FR_NB + 1) * 4);
    s.sums = (uint32_t*)(b + off); off += fr_al(1024 * 4);
    s.sizes = (uint32_t*)(b + off); off += fr_al((size_t)(G + 2) * 4);
    s.P = (uint32_t*)(b + off); off += fr_al((size_t)(G + 1) * k * 4);
    s.bkey = (unsigned long long*)(b + off); off += fr_al((size_t)(G + 1) * 8);
    s.kind = (uint8_t*)(b + off); off += fr_al((size_t)(G + 1));
    s.bytes = off;
    return s;
}

}  // namespace uda

// cap on resident merge CTAs per SM (0 = as many as fit): lowered when the merge shares the GPU with a
// concurrently running push kernel, so that the push CTAs find registers and thread slots
static int g_fr_ctas_per_sm = 0;
extern "C" int uda_merge_records_set_ctas_per_sm(int n) {
    if (n < 0 || n > 32) return UDA_E_INVALID;
    g_fr_ctas_per_sm = n;
    return 0;
}

extern "C" long long uda_merge_records_scratch_bytes(long long n, int nruns) {
    if (n < 0 || nruns < 1) return UDA_E_INVALID;
    return (long long)uda::fr_layout(nullptr, n, nruns).bytes;
}

extern "C" int uda_merge_records_supported(int nruns, long long n, int stride, int key_len) {
    using namespace uda;
    return nruns >= 1 && nruns <= FR_KMAX && n >= 0 && n < (1LL << 32) - FR_TILE && stride >= 16 && stride <= 128 && key_len >= 1 &&
           key_len <= 10 && n / FR_TARGET + 2 <= scan4::MAX_ELEMS && fr_smem_bytes(nruns, stride) <= 200 * 1024;
}

// Merge nruns sorted fixed-stride runs into `out` (n*stride bytes, + 2 bytes EOF marker if write_eof).
// Runs must not alias `out`; every run needs 32 readable bytes after its last record.
extern "C" int uda_merge_records_fixed(const UdaRunDesc* runs_dev, int nruns, long long n, long long max_nrec, int stride,
                                       int key_off, int key_len, int align2, void* out, int write_eof, void* scratch,
                                       cudaStream_t stream) {
    using namespace uda;
    if (!uda_merge_records_supported(nruns, n, stride, key_len)) return UDA_E_INVALID;
    if (align2 && ((stride | key_off) & 1)) return UDA_E_INVALID;
    uint8_t* o = (uint8_t*)out;
    if (n > 0) {
        FrScratch s = fr_layout(scratch, n, nruns);
        const int G = fr_num_buckets(n);
        fr_range_kernel<<<1, 256, 0, stream>>>(runs_dev, nruns, stride, key_off, key_len, s.prm);
        UDA_CUDA(cudaMemsetAsync(s.hist, 0, (size_t)(FR_NB + 1) * 4, stream));
        {
            long long ns = max_nrec / FR_SAMPLE;
            long long bx = (ns + 255) / 256;
            long long cap = (kNumSMs * 16 + nruns - 1) / nruns;
            if (cap < 1) cap = 1;
            if (bx > cap) bx = cap;
            if (bx < 1) bx = 1;
            dim3 grid((unsigned)bx, (unsigned)nruns);
            fr_sample_hist_kernel<<<grid, 256, 0, stream>>>(runs_dev, stride, key_off, key_len, s.prm, s.hist);
        }
        int rc = scan4::exclusive(s.hist, FR_NB, s.hist, s.sums, stream);
        if (rc) return rc;
        {
            int bq = (G + 1 + 255) / 256;
            if (bq > kNumSMs * 8) bq = kNumSMs * 8;
            fr_bkeys_kernel<<<bq, 256, 0, stream>>>(G, s.hist, s.prm, s.bkey, s.kind);
            if (nruns <= 16) {  // measured: search 0.4 ms vs streaming 2.0 ms at k = 8; search 5.5 ms vs streaming 2.4 ms at k = 64
                long long t = (long long)(G + 1) * nruns;
                long long b = (t + 255) / 256;
                if (b > kNumSMs * 32) b = kNumSMs * 32;
                fr_partition_search_kernel<<<(int)b, 256, 0, stream>>>(runs_dev, nruns, G, stride, key_off, key_len, s.bkey, s.kind, s.P);
            } else {
                long long bx = (max_nrec + 255) / 256;
                long long cap = (kNumSMs * 16 + nruns - 1) / nruns;
                if (cap < 1) cap = 1;
                if (bx > cap) bx = cap;
                if (bx < 1) bx = 1;
                dim3 grid((unsigned)bx, (unsigned)nruns);
                fr_partition_kernel<<<grid, 256, 0, stream>>>(runs_dev, nruns, G, stride, key_off, key_len, s.hist, s.prm, s.bkey, s.kind, s.P);
            }
        }
        {
            long long b = ((long long)G * 32 + 255) / 256;
            if (b > kNumSMs * 16) b = kNumSMs * 16;
            fr_sizes_kernel<<<(int)b, 256, 0, stream>>>(s.P, nruns, G, s.sizes);
        }
        rc = scan4::exclusive(s.sizes, G, s.sizes, s.sums, stream);
        if (rc) return rc;
        {
            const size_t smem = fr_smem_bytes(nruns, stride);
            void (*kern)(const UdaRunDesc*, int, int, int, int, int, const uint32_t*, const uint32_t*, FrParams*, uint8_t*);
            int slot;
            if (align2 && stride == 102) { kern = fr_merge_kernel<true, 102>; slot = 0; }
            else if (align2 && stride == 100) { kern = fr_merge_kernel<true, 100>; slot = 1; }
            else if (align2) { kern = fr_merge_kernel<true, 0>; slot = 2; }
            else { kern = fr_merge_kernel<false, 0>; slot = 3; }
            static size_t configured[4] = {0, 0, 0, 0};
            if (smem > configured[slot]) {
                UDA_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
                configured[slot] = smem;
            }
            int per_sm = 0;
            UDA_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, FR_THREADS, smem));
            if (per_sm < 1) return UDA_E_TOO_LARGE;
            if (g_fr_ctas_per_sm > 0 && per_sm > g_fr_ctas_per_sm) per_sm = g_fr_ctas_per_sm;
            int grid = kNumSMs * per_sm;
            if (grid > G) grid = G;
            kern<<<grid, FR_THREADS, smem, stream>>>(runs_dev, nruns, G, stride, key_off, key_len, s.P, s.sizes, s.prm, o);
        }
    }
    if (write_eof) uda::fr_eof_kernel<<<1, 1, 0, stream>>>(o + n * stride);
    UDA_CHECK_LAUNCH();
    return 0;
}

extern "C" int uda_merge_records_stats(const void* scratch, unsigned long long* out2_host, cudaStream_t stream) {
    const uda::FrParams* p = (const uda::FrParams*)scratch;
    UDA_CUDA(cudaMemcpyAsync(out2_host, &p->fallback_buckets, 2 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, stream));
    UDA_CUDA(cudaStreamSynchronize(stream));
    return 0;
}

// Positions of nkeys probe keys (8-byte big-endian prefixes, tail 0) in one sorted fixed-stride run:
// pos_out[i] = first record whose key is >= keys[i]. Used by the spill path to cut merged runs into key ranges.
namespace uda {
__global__ void fr_lower_bound_kernel(const uint8_t* __restrict__ base, long long n, int stride, int key_off, int key_len,
                                      const unsigned long long* __restrict__ keys, int nkeys, long long* __restrict__ pos_out) {
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < nkeys; i += gridDim.x * blockDim.x)
        pos_out[i] = (long long)lb_records(base, stride, key_off, key_len, 0u, (uint32_t)n, keys[i], 0u);
}
__global__ void fr_sample_keys_kernel(const uint8_t* __restrict__ base, long long n, int stride, int key_off, int key_len,
                                      int nkeys, unsigned long long* __restrict__ keys_out) {
    // keys_out[i] = key prefix of record (i+1)*n/(nkeys+1): nkeys equi-depth splitters of this run
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < nkeys; i += gridDim.x * blockDim.x) {
        unsigned long long w0;
        uint32_t t;
        load_key80(base + ((long long)(i + 1) * n / (nkeys + 1)) * stride + key_off, key_len, &w0, &t);
        keys_out[i] = w0;
    }
}
}  // namespace uda

extern "C" int uda_lower_bound_records(const void* base, long long n, int stride, int key_off, int key_len,
                                       const unsigned long long* keys_dev, int nkeys, long long* pos_out_dev, cudaStream_t stream) {
    if (nkeys <= 0) return 0;
    if (n >= (1LL << 32)) return UDA_E_INVALID;
    uda::fr_lower_bound_kernel<<<(nkeys + 127) / 128, 128, 0, stream>>>((const uint8_t*)base, n, stride, key_off, key_len, keys_dev, nkeys,
                                                                        pos_out_dev);
    UDA_CHECK_LAUNCH();
    return 0;
}

extern "C" int uda_sample_splitters(const void* base, long long n, int stride, int key_off, int key_len, int nkeys,
                                    unsigned long long* keys_out_dev, cudaStream_t stream) {
    if (nkeys <= 0 || n <= 0) return 0;
    uda::fr_sample_keys_kernel<<<(nkeys + 127) / 128, 128, 0, stream>>>((const uint8_t*)base, n, stride, key_off, key_len, nkeys, keys_out_dev);
    UDA_CHECK_LAUNCH();
    return 0;
}
