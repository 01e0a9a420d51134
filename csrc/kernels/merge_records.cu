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
//                        k record slices in shared memory. The tile is then ORDERED IN O(1) PASSES, independent
//                        of k: every staged record is dropped into one of FR_NBIN equal-width key bins of the
//                        bucket's key range (a monotone map, so bins are ordered), a shared-memory counting pass
//                        + scan gives each bin its output range, and the few records that share a bin are ranked
//                        exactly against each other on (80-bit key, slot). Buckets are ~400 records wide, so for
//                        any smooth key distribution a bin holds < 1 record on average. Tiles whose keys pile up
//                        in one bin (duplicate-heavy data) fall back to the pairwise merge-path tree (one
//                        block-wide level per log2 k). Records leave shared memory in merged order as coalesced
//                        16-byte stores straight to their final position.
//
// Output order is (key, run, index-in-run), identical to merge.cu + gather.cu.
#include "common.cuh"
#include "merge_tree.cuh"
#include "scan.cuh"
#include "../../include/uda_b200.h"
#include <cstdlib>

namespace uda {

// Tuning history (B200, 125 M records, k = 8; ms per TeraSort step incl. the 4.4 ms push), tree ordering:
//   VT=3 / 2 CTAs per SM 13.95   VT=2 / 3 CTAs 13.53   VT=4 / 1 CTA 17.76   VT=2 / 3 CTAs, 128 merge threads 13.91
#ifndef UDA_FR_VT
#define UDA_FR_VT 2
#endif
#ifndef UDA_FR_CTAS
#define UDA_FR_CTAS 2
#endif
#ifndef UDA_FR_THREADS
#define UDA_FR_THREADS 256
#endif
#ifndef UDA_FR_MERGE_THREADS
#define UDA_FR_MERGE_THREADS UDA_FR_THREADS
#endif
#ifndef UDA_FR_BIN_LIMIT
#define UDA_FR_BIN_LIMIT 48   // records in one key bin beyond which a tile is ordered by the merge tree instead
#endif
constexpr int FR_THREADS = UDA_FR_THREADS;
constexpr int FR_VT = UDA_FR_VT;
constexpr int FR_TILE = FR_THREADS * FR_VT;  // largest tile capacity in record slots (512); the capacity in use is a launch parameter
constexpr int FR_CAP_MIN = 128;              // smallest capacity the host ever picks (bounds the scratch size)
constexpr int FR_MERGE_THREADS = UDA_FR_MERGE_THREADS;   // threads that run the merge tree (fewer = more items per thread)
constexpr int FR_MERGE_VT = FR_TILE / FR_MERGE_THREADS;
constexpr int FR_NBIN = 4 * FR_THREADS;      // key bins per tile: four per thread in the scan
constexpr int FR_BIN_LIMIT = UDA_FR_BIN_LIMIT;
constexpr int FR_SAMPLE = 8;
constexpr int FR_NB_LOG = 22;
constexpr int FR_NB = 1 << FR_NB_LOG;
constexpr int FR_KMAX = 1024;
constexpr int FR_SLICE_PAD = 48;             // bytes of shared memory reserved per run for 16-byte alignment slop
// target bucket size for a tile capacity: a quarter of the tile absorbs the sampling error
__host__ __device__ inline int fr_target(int cap) { return cap * 3 / 4; }

struct FrParams {
    unsigned long long kmin, kmax;
    int shift, pad;
    unsigned long long fallback_buckets, fallback_iters;
    unsigned long long tree_tiles;    // tiles ordered by the merge tree instead of the key bins
    unsigned int next_bucket, pad2;   // dynamic bucket scheduler of fr_merge_kernel (reset by fr_range_kernel)
};

// Key prefix of a record in global memory. A 10-byte key at an even address (TeraSort; any 2-byte aligned layout) is read
// with two aligned 8-byte loads and shifted into place: the byte-wise path issues ten scattered one-byte loads per key,
// which made the partition pre-passes L1-throughput bound (every lane of a warp touches a different cache line).
// The aligned window starts at most 6 bytes before the key (never before an 8-byte aligned allocation start) and ends at
// most 16 bytes after its first byte (inside the 32 readable bytes every run has after its last record).
__device__ __forceinline__ void load_key80(const uint8_t* p, int key_len, unsigned long long* w0, uint32_t* t) {
    if (key_len == 10 && ((uintptr_t)p & 1u) == 0) {
        const unsigned long long* a = reinterpret_cast<const unsigned long long*>((uintptr_t)p & ~(uintptr_t)7);
        const uint32_t sh = (uint32_t)((uintptr_t)p & 7u) * 8u;   // 0, 16, 32, 48
        const unsigned long long q0 = __ldg(a), q1 = __ldg(a + 1);
        const unsigned long long le = sh ? (q0 >> sh) | (q1 << (64u - sh)) : q0;   // key bytes 0..7, little-endian
        *w0 = bswap64(le);
        const uint32_t tl = (uint32_t)(q1 >> sh) & 0xFFFFu;                          // key bytes 8..9
        *t = ((tl & 0xFFu) << 8) | (tl >> 8);
        return;
    }
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
__global__ void fr_range_kernel(const UdaRunDesc* __restrict__ runs, int k, int stride, int key_off, int key_len, int nb,
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
        while (shift < 64 && (range >> shift) >= (unsigned long long)nb) ++shift;
        prm->kmin = lo; prm->kmax = hi; prm->shift = shift;
        prm->fallback_buckets = 0; prm->fallback_iters = 0; prm->tree_tiles = 0;
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
__global__ void fr_bkeys_kernel(int G, int per, int nb, const uint32_t* __restrict__ cum, const FrParams* __restrict__ prm,
                                unsigned long long* __restrict__ bkey, uint8_t* __restrict__ kind) {
    for (int q = blockIdx.x * blockDim.x + threadIdx.x; q <= G; q += gridDim.x * blockDim.x) {
        const unsigned long long want = (unsigned long long)q * (unsigned long long)per;
        int lo = 0, hi = nb;
        while (lo < hi) {
            int mid = (lo + hi) >> 1;
            if ((unsigned long long)__ldg(&cum[mid]) < want) lo = mid + 1; else hi = mid;
        }
        uint8_t kd = 1;
        unsigned long long key = 0;
        if (q == 0) kd = 0;
        else if (lo >= nb || q == G) kd = 2;
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
// sector-sparse pass (2.2 ms for 125 M records) replaces (G+1)*k independent binary searches over strided
// records, which for k = 64 dragged the whole data set through DRAM at 2 TB/s (6.3 ms). Runs that are much
// shorter than the bucket count keep the per-boundary binary search (their gaps would serialise one thread).
// The pass is DRAM bound, not latency bound: one key per 102-byte record touches 62 % of the 64-byte DRAM atoms, so
// no key pass can take less than ~1.4 ms here. Two walking variants that avoid the prefix-table look-up (a thread per
// 32 boundaries: 4.9 ms; a warp per run with shared-memory key windows: 2.8 ms) measured slower and were dropped.
__global__ void fr_partition_kernel(const UdaRunDesc* __restrict__ runs, int k, int G, int per_i, int stride, int key_off, int key_len,
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
    const uint32_t per = (uint32_t)per_i;
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
// The merge kernel is a producer / consumer pipeline inside every CTA:
//
//   producer warp (1)   claims the next bucket, reads its split positions and key range, lays the k slices out in one
//                       of TWO record buffers and issues the TMA bulk copies for it — one tile ahead of the consumers,
//                       so the bucket's metadata round trips and the DRAM latency of the copies are off the critical
//                       path (in the single-buffered version they were 40 % of the stall samples, profiles/NCU_SUMMARY.md).
//   consumer warps (8)  wait on the buffer's `full` mbarrier, order the tile (key bins, see the file header), write the
//                       records out, and hand the buffer back through its `empty` mbarrier. They synchronise among
//                       themselves with a named barrier; the producer never takes part.
//
// Shared memory of one CTA (`cap` = tile capacity in record slots):
//   rec x 2                  staged record bytes of the k slices, double buffered
//   work  (cap * 36 bytes)   bin ordering: skey[cap] u64 | sts[cap] u32 | bincnt[NBIN] | binbase[NBIN + 4]
//                            tree ordering (fallback): entA[cap] | entB[cap] | slotaddr[cap]      (same bytes)
//   srcaddr[cap]             byte offset inside rec of the record that goes to output position p
//   meta x 3                 per tile in flight: header, sbase[k], soff[k + 1], lo[k], bend[k]
//   hi[k], doff[k + 1]       oversize-bucket / merge-tree scratch (consumers only)
constexpr int FR_CONSUMERS = FR_THREADS;            // threads that order and emit a tile
constexpr int FR_THREADS_ALL = FR_CONSUMERS + 32;   // + the producer warp
constexpr uint32_t FR_META_DONE = 0xFFFFFFFFu;

struct FrTileHdr {
    unsigned long long klo, kspan;   // key bins cover [klo, klo + kspan]
    uint32_t total;                  // records in the tile; FR_META_DONE: no more work
    uint32_t obase;                  // first output record of the tile
    uint32_t big;                    // oversize bucket: lo / bend hold its slice bounds, nothing was staged
    uint32_t pad;
};
struct FrMeta {
    FrTileHdr* hdr;
    uint32_t* sbase;   // [k] byte offset of the first record of run r's slice inside rec
    uint32_t* soff;    // [k+1] slot offsets (list boundaries for the merge tree)
    uint32_t* lo;      // [k]
    uint32_t* bend;    // [k]
};
struct FrSmem {
    uint8_t* rec0;       // two record buffers, rec_bytes apart
    uint8_t* meta0;      // three metadata blocks, meta_bytes apart
    uint32_t rec_bytes, meta_bytes;
    int k;
    __device__ __forceinline__ uint8_t* rec(uint32_t b) const { return rec0 + b * rec_bytes; }
    __device__ __forceinline__ FrMeta meta(uint32_t b) const {   // computed, not indexed: keeps the struct in registers
        uint8_t* mp = meta0 + b * meta_bytes;
        FrMeta m;
        m.hdr = reinterpret_cast<FrTileHdr*>(mp);
        m.sbase = reinterpret_cast<uint32_t*>(mp + sizeof(FrTileHdr));
        m.soff = m.sbase + k;
        m.lo = m.soff + (k + 1);
        m.bend = m.lo + k;
        return m;
    }
    Entry* entA;
    Entry* entB;
    uint32_t* slotaddr;
    unsigned long long* skey;
    uint32_t* sts;
    uint32_t* bincnt;    // all zero between tiles
    uint32_t* binbase;
    uint32_t* srcaddr;
    uint32_t* hi;
    uint32_t* doff;
};

__host__ __device__ inline size_t fr_rec_bytes(int cap, int k, int stride) {
    return ((size_t)cap * stride + (size_t)FR_SLICE_PAD * k + 64 + 15) & ~(size_t)15;
}
__host__ __device__ inline size_t fr_work_bytes(int cap) {
    const size_t tree = (size_t)cap * 36, bins = (size_t)cap * 12 + ((size_t)2 * FR_NBIN + 4) * 4;
    return ((tree > bins ? tree : bins) + 15) & ~(size_t)15;
}
__host__ __device__ inline size_t fr_meta_bytes(int k) { return sizeof(FrTileHdr) + ((size_t)4 * k + 1) * 4 + 4; }   // multiple of 8
__host__ __device__ inline size_t fr_smem_bytes(int cap, int k, int stride) {
    return 2 * fr_rec_bytes(cap, k, stride) + fr_work_bytes(cap) + (size_t)cap * 4 + 3 * fr_meta_bytes(k) + ((size_t)2 * k + 1) * 4 + 16;
}

__device__ __forceinline__ void mbar_expect_tx_only(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.expect_tx.relaxed.cta.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
// barrier among the consumer warps only (the producer warp runs ahead on its own)
__device__ __forceinline__ void fr_cbar() { asm volatile("bar.sync 1, %0;" ::"n"(FR_CONSUMERS) : "memory"); }

// One full warp lays the k slices [lo[r], hi[r]) out in `rec` (slot offsets by a shuffle scan, 32 runs at a time), issues one
// bulk copy per non-empty slice and arrives on `bar` once; the data (and the tables) are visible to whoever waits on it.
__device__ __forceinline__ void fr_stage(const UdaRunDesc* __restrict__ runs, int k, int stride, const uint32_t* lo_a,
                                         const uint32_t* hi_a, uint8_t* rec, uint32_t* sbase, uint32_t* soff, uint32_t total,
                                         uint64_t* bar) {
    const int lane = threadIdx.x & 31;
    fence_proxy_async_smem();  // earlier generic-proxy accesses to `rec` are ordered before the TMA writes
    uint32_t run_slot = 0;
    for (int r0 = 0; r0 < k; r0 += 32) {
        const int r = r0 + lane;
        const uint32_t lo = r < k ? lo_a[r] : 0u;
        const uint32_t len = r < k ? hi_a[r] - lo : 0u;
        uint32_t x = len;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= o) x += y;
        }
        const uint32_t slot = run_slot + x - len;
        if (r < k) {
            soff[r] = slot;
            const uint32_t B = ((slot * (uint32_t)stride + (uint32_t)FR_SLICE_PAD * r) + 15u) & ~15u;
            if (len) {
                const uint8_t* src = (const uint8_t*)runs[r].base + (unsigned long long)lo * stride;
                const uint32_t mis = (uint32_t)((uintptr_t)src & 15);
                const uint32_t nbytes = (mis + len * (uint32_t)stride + 15u) & ~15u;
                sbase[r] = B + mis;
                mbar_expect_tx_only(bar, nbytes);
                tma_load_1d(rec + B, src - mis, nbytes, bar);
            } else {
                sbase[r] = B;
            }
        }
        run_slot += __shfl_sync(0xffffffffu, x, 31);
    }
    if (lane == 0) soff[k] = total;
    __syncwarp();                      // every expect_tx and table store precedes the single arrival
    if (lane == 0) mbar_arrive(bar);
}

// Orders one staged tile and writes its records to `out` (consumer threads only). `release`: mbarrier of the OTHER record
// buffer, handed back to the producer once every consumer has left the previous tile's output phase (first barrier here).
// There is no barrier at the end: the next tile's phase A may overlap this tile's output phase (it touches neither
// srcaddr nor rec of this buffer).
template <bool ALIGN2, int STRIDE>
__device__ __forceinline__ void fr_order_emit(const FrSmem& sm, const uint8_t* rec, const FrMeta& mt, int k, int stride_rt, int key_off,
                                              int key_len, uint32_t total, unsigned long long klo, unsigned long long kspan,
                                              int force_tree, uint8_t* __restrict__ out, uint64_t* release, uint32_t* s_ws,
                                              uint32_t* s_flag, const uint4* s_mask, FrParams* __restrict__ prm) {
    const int stride = STRIDE > 0 ? STRIDE : stride_rt;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    // ---- A: key of every staged record -> bin, position inside the bin by a shared-memory atomic
    if (tid == 0) *s_flag = (uint32_t)force_tree;
    const int bits = 64 - __clzll((long long)kspan);
    const int sh = bits > 32 ? bits - 32 : 0;
    const float scale = (float)FR_NBIN / (__uint2float_rz((uint32_t)(kspan >> sh)) + 1.0f);
    unsigned long long kw0[FR_VT];
    uint32_t kts[FR_VT], kaddr[FR_VT], kbin[FR_VT];
#pragma unroll
    for (int v = 0; v < FR_VT; ++v) {
        const uint32_t j = (uint32_t)tid + (uint32_t)v * FR_CONSUMERS;
        kw0[v] = 0; kts[v] = 0; kaddr[v] = 0; kbin[v] = 0;
        if (j < total) {
            int lo = 0, hi = k - 1;  // run owning slot j: last r with soff[r] <= j
            while (lo < hi) {
                const int mid = (lo + hi + 1) >> 1;
                if (mt.soff[mid] <= j) lo = mid; else hi = mid - 1;
            }
            const uint32_t addr = mt.sbase[lo] + (j - mt.soff[lo]) * (uint32_t)stride;
            const uint8_t* kp = rec + addr + key_off;
            unsigned long long w0 = 0;
            uint32_t t = 0;
            if (ALIGN2 && key_len == 10) {
                const uint16_t* h = reinterpret_cast<const uint16_t*>(kp);
                const uint32_t h0 = h[0], h1 = h[1], h2 = h[2], h3 = h[3], h4 = h[4];
                // bytes are big-endian key digits; each little-endian half holds (b0 | b1 << 8)
                w0 = ((unsigned long long)__byte_perm(h0, h1, 0x0145) << 32) | __byte_perm(h2, h3, 0x0145);
                t = __byte_perm(h4, 0, 0x4401);
            } else {
                key_prefix(kp, key_len, &w0, &t);
            }
            unsigned long long rel = w0 > klo ? w0 - klo : 0ull;
            if (rel > kspan) rel = kspan;
            uint32_t b = __float2uint_rz(__uint2float_rz((uint32_t)(rel >> sh)) * scale);   // monotone in the key
            b = b < (uint32_t)FR_NBIN - 1u ? b : (uint32_t)FR_NBIN - 1u;
            const uint32_t o = atomicAdd(&sm.bincnt[b], 1u);
            kw0[v] = w0; kts[v] = (t << 16) | j; kaddr[v] = addr; kbin[v] = b | (o << 16);
        }
    }
    fr_cbar();
    if (release != nullptr && tid == 0) {   // everybody is out of the previous tile: its record buffer may be refilled
        fence_proxy_async_smem();
        mbar_arrive(release);
    }

    // ---- B: exclusive scan of the bin counts (four bins per thread); the counters go back to zero for the next tile
    {
        const uint4 c = *reinterpret_cast<const uint4*>(&sm.bincnt[4 * tid]);
        *reinterpret_cast<uint4*>(&sm.bincnt[4 * tid]) = make_uint4(0u, 0u, 0u, 0u);
        const uint32_t s4 = c.x + c.y + c.z + c.w;
        uint32_t x = s4;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t y = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= o) x += y;
        }
        if (lane == 31) s_ws[warp] = x;
        const uint32_t mx = max(max(c.x, c.y), max(c.z, c.w));
        if (mx > (uint32_t)FR_BIN_LIMIT) *s_flag = 1u;
        fr_cbar();
        uint32_t wp = lane < warp ? s_ws[lane] : 0u;   // FR_CONSUMERS / 32 <= 32 warps
#pragma unroll
        for (int o = 16; o; o >>= 1) wp += __shfl_xor_sync(0xffffffffu, wp, o);
        const uint32_t e = wp + x - s4;
        *reinterpret_cast<uint4*>(&sm.binbase[4 * tid]) = make_uint4(e, e + c.x, e + c.x + c.y, e + c.x + c.y + c.z);
        if (tid == FR_CONSUMERS - 1) sm.binbase[FR_NBIN] = e + s4;
    }
    fr_cbar();

    if (*s_flag == 0u) {
        // ---- C: a record alone in its bin knows its output position; the others are staged by bin
        uint32_t b0[FR_VT], cn[FR_VT];
#pragma unroll
        for (int v = 0; v < FR_VT; ++v) {
            const uint32_t j = (uint32_t)tid + (uint32_t)v * FR_CONSUMERS;
            b0[v] = 0; cn[v] = 0;
            if (j < total) {
                const uint32_t b = kbin[v] & 0xFFFFu, o = kbin[v] >> 16;
                b0[v] = sm.binbase[b];
                cn[v] = sm.binbase[b + 1] - b0[v];
                if (cn[v] == 1u) sm.srcaddr[b0[v]] = kaddr[v];
                else { sm.skey[b0[v] + o] = kw0[v]; sm.sts[b0[v] + o] = kts[v]; }
            }
        }
        fr_cbar();
        // ---- D: exact rank among the records of the same bin on (key, tail, slot); slots are run-major, so equal
        //         keys come out in (run, index) order
#pragma unroll
        for (int v = 0; v < FR_VT; ++v) {
            if (cn[v] > 1u) {
                uint32_t rank = 0;
                for (uint32_t i = 0; i < cn[v]; ++i) {
                    const unsigned long long a = sm.skey[b0[v] + i];
                    const uint32_t bt = sm.sts[b0[v] + i];
                    rank += (a < kw0[v]) || (a == kw0[v] && bt < kts[v]) ? 1u : 0u;
                }
                sm.srcaddr[b0[v] + rank] = kaddr[v];
            }
        }
        fr_cbar();
    } else {
        // ---- fallback: many records share a key bin (duplicate-heavy keys). Pairwise merge-path tree over 16-byte
        //      entries, one block-wide level per log2 k.
        if (tid == 0) atomicAdd(&prm->tree_tiles, 1ull);
        fr_cbar();               // nobody reads binbase / bincnt any more: the work area becomes entA | entB | slotaddr
#pragma unroll
        for (int v = 0; v < FR_VT; ++v) {
            const uint32_t j = (uint32_t)tid + (uint32_t)v * FR_CONSUMERS;
            if (j < total) {
                Entry e;
                e.w0 = kw0[v];
                e.w1 = ((unsigned long long)(kts[v] >> 16) << 48) | j;
                sm.entA[j] = e;
                sm.slotaddr[j] = kaddr[v];
            }
        }
        fr_cbar();
        Entry* src = sm.entA; Entry* dst = sm.entB;
        uint32_t* soff = mt.soff; uint32_t* doff = sm.doff;
        int L = k;
        while (L > 1) {
            if (tid < FR_MERGE_THREADS) merge_level<FR_MERGE_THREADS, FR_MERGE_VT>(src, soff, L, dst, doff, total);
            fr_cbar();
            Entry* te = src; src = dst; dst = te;
            uint32_t* tu = soff; soff = doff; doff = tu;
            L = (L + 1) >> 1;
        }
        for (uint32_t p = tid; p < total; p += FR_CONSUMERS) sm.srcaddr[p] = sm.slotaddr[(uint32_t)src[p].w1 & 0xFFFFu];
        fr_cbar();
        for (int b = tid; b < FR_NBIN; b += FR_CONSUMERS) sm.bincnt[b] = 0u;   // restore the invariant the bin path relies on
        fr_cbar();
    }

    // ---- E: records leave in merged order, one thread per 16-byte *destination* word (all 32 lanes busy, fully coalesced
    // stores). Words that lie inside one record are a single unaligned 16-byte read of shared memory (aligned 32-bit loads +
    // funnel shifts); the one word per record boundary that takes bytes from two records is handled in a second, dense
    // loop over the boundaries, so the two-source path is not paid by every warp of the first loop.
    // Shared-space addresses are explicit so the loads are LDS, not generic LD.
    const uint32_t rec_s = smem_u32(rec), sa_s = smem_u32(sm.srcaddr), mask_s = smem_u32(s_mask);
    auto slot_addr = [&](uint32_t p) -> uint32_t {  // shared address of merged record p
        uint32_t sa;
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(sa) : "r"(sa_s + p * 4u));
        return rec_s + sa;
    };
    auto lds16 = [](uint32_t a) -> uint4 {  // 16 bytes at arbitrary shared address a
        const uint32_t w = a & ~3u, sh8 = (a & 3u) * 8u;
        uint32_t x0, x1, x2, x3, x4;
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(x0) : "r"(w));
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(x1) : "r"(w + 4u));
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(x2) : "r"(w + 8u));
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(x3) : "r"(w + 12u));
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(x4) : "r"(w + 16u));
        uint4 r;
        r.x = __funnelshift_r(x0, x1, sh8); r.y = __funnelshift_r(x1, x2, sh8);
        r.z = __funnelshift_r(x2, x3, sh8); r.w = __funnelshift_r(x3, x4, sh8);
        return r;
    };
    const uint32_t span = total * (uint32_t)stride;
    const uint32_t head = (uint32_t)((16u - ((uintptr_t)out & 15u)) & 15u) < span ? (uint32_t)((16u - ((uintptr_t)out & 15u)) & 15u) : span;
    const uint32_t nfull = (span - head) >> 4;
#pragma unroll 2
    for (uint32_t w = tid; w < nfull; w += FR_CONSUMERS) {
        const uint32_t rel = head + (w << 4);
        const uint32_t p = rel / (uint32_t)stride;
        const uint32_t o = rel - p * (uint32_t)stride;
        if (o + 16u <= (uint32_t)stride) {   // the word lies inside record p
            const uint4 v = lds16(slot_addr(p) + o);
            st_stream(reinterpret_cast<int4*>(out + rel), make_int4((int)v.x, (int)v.y, (int)v.z, (int)v.w));
        }
    }
    // the word around the boundary between records i-1 and i (stride >= 16: a word never spans three records)
    for (uint32_t i = (uint32_t)tid + 1u; i < total; i += FR_CONSUMERS) {
        const uint32_t rel_b = i * (uint32_t)stride;
        if (rel_b <= head) continue;                 // inside the unaligned head bytes (byte loop below)
        const uint32_t d = rel_b - head;
        if ((d & 15u) == 0u) continue;               // the boundary coincides with a word boundary
        const uint32_t wrel = head + (d & ~15u);
        if (wrel + 16u > head + (nfull << 4)) continue;   // inside the unaligned tail bytes
        const uint32_t len0 = rel_b - wrel;          // bytes of record i-1 in this word (1..15)
        uint4 v = lds16(slot_addr(i - 1u) + (uint32_t)stride - len0);
        const uint4 u = lds16(slot_addr(i) - len0);
        // byte b of the word comes from record i-1 for b < len0: a 16-entry mask table in shared memory (one LDS.128) instead
        // of per-word shift arithmetic, which was 12 % of the kernel's instructions
        uint32_t m0, m1, m2, m3;
        asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(m0), "=r"(m1), "=r"(m2), "=r"(m3) : "r"(mask_s + len0 * 16u));
        v.x = (v.x & m0) | (u.x & ~m0); v.y = (v.y & m1) | (u.y & ~m1);
        v.z = (v.z & m2) | (u.z & ~m2); v.w = (v.w & m3) | (u.w & ~m3);
        st_stream(reinterpret_cast<int4*>(out + wrel), make_int4((int)v.x, (int)v.y, (int)v.z, (int)v.w));
    }
    // unaligned head and tail bytes of the tile's span (< 16 bytes each)
    {
        const uint32_t tail0 = head + (nfull << 4);
        const uint32_t nedge = head + (span - tail0);
        for (uint32_t e = tid; e < nedge; e += FR_CONSUMERS) {
            const uint32_t rel = e < head ? e : tail0 + (e - head);
            const uint32_t p = rel / (uint32_t)stride, o = rel - p * (uint32_t)stride;
            uint32_t v;
            asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(slot_addr(p) + o));
            out[rel] = (uint8_t)v;
        }
    }
}

template <bool ALIGN2, int STRIDE>
__global__ void __launch_bounds__(FR_THREADS_ALL, UDA_FR_CTAS)
fr_merge_kernel(const UdaRunDesc* __restrict__ runs, int k, int G, int cap, int stride_rt, int key_off, int key_len, int force_tree,
                const uint32_t* __restrict__ P, const uint32_t* __restrict__ outoff, const unsigned long long* __restrict__ bkey,
                const uint8_t* __restrict__ kind, FrParams* __restrict__ prm, uint8_t* __restrict__ out) {
    const int stride = STRIDE > 0 ? STRIDE : stride_rt;
    extern __shared__ __align__(128) uint8_t fr_smem_raw[];
    __shared__ __align__(8) uint64_t bar_full[2], bar_empty[2], bar_own;
    __shared__ uint32_t s_cnt, s_flag;
    __shared__ uint32_t s_ws[32];
    __shared__ __align__(16) uint4 s_mask[16];   // s_mask[n]: the low n bytes of a 16-byte word set
    if (threadIdx.x < 16) {
        const int n = threadIdx.x;
        auto w = [n](int wi) -> uint32_t { const int rem = n - 4 * wi; return rem >= 4 ? 0xFFFFFFFFu : (rem <= 0 ? 0u : (0xFFFFFFFFu >> (32 - 8 * rem))); };
        s_mask[n] = make_uint4(w(0), w(1), w(2), w(3));
    }
    FrSmem sm;
    const size_t rec_bytes = fr_rec_bytes(cap, k, stride);
    sm.rec0 = fr_smem_raw;
    sm.rec_bytes = (uint32_t)rec_bytes;
    sm.k = k;
    uint8_t* work = fr_smem_raw + 2 * rec_bytes;
    sm.entA = reinterpret_cast<Entry*>(work);
    sm.entB = sm.entA + cap;
    sm.slotaddr = reinterpret_cast<uint32_t*>(sm.entB + cap);
    sm.skey = reinterpret_cast<unsigned long long*>(work);
    sm.sts = reinterpret_cast<uint32_t*>(sm.skey + cap);
    sm.bincnt = sm.sts + cap;
    sm.binbase = sm.bincnt + FR_NBIN;
    sm.srcaddr = reinterpret_cast<uint32_t*>(work + fr_work_bytes(cap));
    uint8_t* mp = reinterpret_cast<uint8_t*>(sm.srcaddr + cap);
    mp += (8 - ((uintptr_t)mp & 7)) & 7;
    sm.meta0 = mp;
    sm.meta_bytes = (uint32_t)fr_meta_bytes(k);
    mp += 3 * fr_meta_bytes(k);
    sm.hi = reinterpret_cast<uint32_t*>(mp);
    sm.doff = sm.hi + k;
    if (threadIdx.x == 0) {
        mbar_init(&bar_full[0], 1); mbar_init(&bar_full[1], 1);
        mbar_init(&bar_empty[0], 1); mbar_init(&bar_empty[1], 1);
        mbar_init(&bar_own, 1);
        mbar_fence_init();
    }
    for (int b = threadIdx.x; b < FR_NBIN; b += FR_THREADS_ALL) sm.bincnt[b] = 0u;
    __syncthreads();   // the only block-wide barrier: from here on the two roles run on their own
    const int lane = threadIdx.x & 31;

    if (threadIdx.x >= FR_CONSUMERS) {
        // =============================================================== producer warp
        // Buckets are handed out through a counter, not by blockIdx stride: when part of the SMs is held by a
        // concurrently running push kernel, the CTAs that do get an SM still share the work evenly.
        const unsigned long long gmin = prm->kmin, gmax = prm->kmax;
        // Metadata of the next non-empty bucket -> meta block `slot` (three blocks: the consumers may still read the one
        // of the tile they are on while the next tile's is complete and the one after that is being fetched).
        // The bucket counter is read one fetch ahead (qn), and everything that depends only on the bucket number (output
        // range, split positions, key range) is loaded at once: one global round trip per fetch instead of three.
        int qn = 0;
        if (lane == 0) qn = (int)atomicAdd(&prm->next_bucket, 1u);
        qn = __shfl_sync(0xffffffffu, qn, 0);
        auto fetch = [&](uint32_t slot) {
            const FrMeta mt = sm.meta(slot);
            for (;;) {
                const int q = qn;
                if (q >= G) {
                    if (lane == 0) mt.hdr->total = FR_META_DONE;
                    break;
                }
                int q2 = 0;
                if (lane == 0) q2 = (int)atomicAdd(&prm->next_bucket, 1u);   // for the next fetch; in flight with the loads below
                const uint32_t obase = outoff[q];
                const uint32_t total = outoff[q + 1] - obase;
                uint32_t plo[4], phi[4];   // k <= 128 in registers; larger k in a second loop
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const int r = c * 32 + lane;
                    plo[c] = r < k ? P[(long long)q * k + r] : 0u;
                    phi[c] = r < k ? P[(long long)(q + 1) * k + r] : 0u;
                }
                unsigned long long klo = 0, khi = 0;
                if (lane == 0) {   // the bucket's key range: [first key of the bucket, last key before the next bucket]
                    klo = kind[q] == 1 ? bkey[q] : gmin;
                    khi = kind[q + 1] == 1 ? bkey[q + 1] - 1ull : gmax;
                }
                qn = __shfl_sync(0xffffffffu, q2, 0);
                if (total == 0) continue;
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const int r = c * 32 + lane;
                    if (r < k) { mt.lo[r] = plo[c]; mt.bend[r] = phi[c]; }
                }
                for (int r = 128 + lane; r < k; r += 32) {
                    mt.lo[r] = P[(long long)q * k + r];
                    mt.bend[r] = P[(long long)(q + 1) * k + r];
                }
                if (lane == 0) {
                    mt.hdr->klo = klo;
                    mt.hdr->kspan = khi > klo ? khi - klo : 0ull;
                    mt.hdr->total = total;
                    mt.hdr->obase = obase;
                    mt.hdr->big = total > (uint32_t)cap ? 1u : 0u;
                }
                break;
            }
            __syncwarp();
        };
        fetch(0);
        for (uint32_t it = 0;; ++it) {
            const uint32_t b = it & 1u;
            if (it >= 2) mbar_wait(&bar_empty[b], ((it >> 1) & 1u) ^ 1u);   // the consumers have left this buffer's previous tile
            const FrMeta mt = sm.meta(it % 3u);
            const uint32_t total = mt.hdr->total;
            if (total == FR_META_DONE) {
                if (lane == 0) mbar_arrive(&bar_full[b]);
                break;
            }
            if (total > (uint32_t)cap) {
                if (lane == 0) mbar_arrive(&bar_full[b]);   // the consumers split and stage an oversize bucket themselves
            } else {
                fr_stage(runs, k, stride, mt.lo, mt.bend, sm.rec(b), mt.sbase, mt.soff, total, &bar_full[b]);
            }
            fetch((it + 1u) % 3u);   // its global-memory round trips overlap the consumers' work on tile `it`
        }
        return;
    }

    // =================================================================== consumer warps
    uint32_t own_phase = 0;
    for (uint32_t it = 0;; ++it) {
        const uint32_t b = it & 1u;
        mbar_wait(&bar_full[b], (it >> 1) & 1u);     // acquire: tables and records of this tile are visible
        const FrMeta mt = sm.meta(it % 3u);
        const uint32_t total = mt.hdr->total;
        if (total == FR_META_DONE) break;
        const uint32_t obase = mt.hdr->obase;
        const unsigned long long klo = mt.hdr->klo, kspan = mt.hdr->kspan;
        uint64_t* release = it > 0 ? &bar_empty[b ^ 1u] : nullptr;
        if (!mt.hdr->big) {
            fr_order_emit<ALIGN2, STRIDE>(sm, sm.rec(b), mt, k, stride, key_off, key_len, total, klo, kspan, force_tree,
                                          out + (unsigned long long)obase * stride, release, s_ws, &s_flag, s_mask, prm);
            continue;
        }
        // ---- oversize bucket: split by bisection over the 128-bit value (key80 | run | index); every piece is staged by
        //      consumer warp 0 into this buffer and ordered like a regular tile
        fr_cbar();
        if (threadIdx.x == 0) {
            if (release != nullptr) { fence_proxy_async_smem(); mbar_arrive(release); }
            atomicAdd(&prm->fallback_buckets, 1ull);
        }
        uint32_t remaining = total, done = 0;
        unsigned __int128 vlo = 0;
        while (remaining > 0) {
            uint32_t cnt;
            if (remaining <= (uint32_t)cap) {
                for (int r = threadIdx.x; r < k; r += FR_CONSUMERS) sm.hi[r] = mt.bend[r];
                cnt = remaining;
                fr_cbar();
            } else {
                unsigned __int128 lo128 = vlo, hi128 = ~(unsigned __int128)0, best = vlo;
                uint32_t bestc = 0;
                int iters = 0;
                auto split_pos = [&](int r, unsigned __int128 v) -> uint32_t {
                    const unsigned long long v0 = (unsigned long long)(v >> 64), v1 = (unsigned long long)v;
                    const uint32_t t = (uint32_t)(v1 >> 48), vrun = (uint32_t)(v1 >> 32) & 0xFFFFu, vidx = (uint32_t)v1;
                    const uint8_t* base = (const uint8_t*)runs[r].base;
                    uint32_t lb = lb_records(base, stride, key_off, key_len, mt.lo[r], mt.bend[r], v0, t);
                    const uint32_t rid = (uint32_t)r;  // slots are laid out run-major, so the run *index* is the tie-break rank
                    if (rid > vrun) return lb;
                    uint32_t ub = ub_records(base, stride, key_off, key_len, lb, mt.bend[r], v0, t);
                    if (rid < vrun) return ub;
                    uint32_t x = vidx < lb ? lb : (vidx > ub ? ub : vidx);
                    return x;
                };
                while (hi128 - lo128 > 1) {
                    unsigned __int128 mid = lo128 + ((hi128 - lo128) >> 1);
                    uint32_t c = 0;
                    for (int r = threadIdx.x; r < k; r += FR_CONSUMERS) c += split_pos(r, mid) - mt.lo[r];
                    for (int o = 16; o; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
                    fr_cbar();
                    if (threadIdx.x == 0) s_cnt = 0;
                    fr_cbar();
                    if ((threadIdx.x & 31) == 0) atomicAdd(&s_cnt, c);
                    fr_cbar();
                    c = s_cnt;
                    ++iters;
                    if (c > (uint32_t)cap) hi128 = mid;
                    else {
                        lo128 = mid; best = mid; bestc = c;
                        if (c >= (uint32_t)cap / 2) break;
                    }
                }
                if (threadIdx.x == 0) atomicAdd(&prm->fallback_iters, (unsigned long long)iters);
                for (int r = threadIdx.x; r < k; r += FR_CONSUMERS) sm.hi[r] = split_pos(r, best);
                cnt = bestc;
                vlo = best;
                fr_cbar();
            }
            if (cnt) {
                if (threadIdx.x < 32) fr_stage(runs, k, stride, mt.lo, sm.hi, sm.rec(b), mt.sbase, mt.soff, cnt, &bar_own);
                mbar_wait(&bar_own, own_phase);
                own_phase ^= 1u;
                fr_order_emit<ALIGN2, STRIDE>(sm, sm.rec(b), mt, k, stride, key_off, key_len, cnt, klo, kspan, force_tree,
                                              out + (unsigned long long)(obase + done) * stride, nullptr, s_ws, &s_flag, s_mask, prm);
            }
            fr_cbar();   // the piece has left the record buffer and the tables
            done += cnt;
            remaining -= cnt;
            for (int r = threadIdx.x; r < k; r += FR_CONSUMERS) mt.lo[r] = sm.hi[r];
            fr_cbar();
        }
    }
}

__global__ void fr_eof_kernel(uint8_t* dst) { dst[0] = 0xFF; dst[1] = 0xFF; }

static inline int fr_num_buckets(long long n, int cap) { return (int)(n / fr_target(cap)) + 2; }
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
static FrScratch fr_layout(void* base, long long n, int k, int cap) {
    FrScratch s;
    const int G = fr_num_buckets(n, cap);
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

// Tile capacity for a run count and stride: the largest multiple of 32 slots (<= FR_TILE) whose shared memory lets
// `want_ctas` CTAs share an SM (227 KB per SM, 1 KB reserved per CTA + the kernel's static shared memory).
static int fr_pick_cap(int k, int stride, int want_ctas) {
    const size_t budget = (size_t)232448 / (size_t)want_ctas - 1024 - 256;
    int cap = FR_TILE;
    while (cap > FR_CAP_MIN && fr_smem_bytes(cap, k, stride) > budget) cap -= 32;
    return cap;
}
// UDA_FR_CAP (tile capacity) / UDA_FR_WANT_CTAS / UDA_FR_FORCE_TREE: experiment overrides, read once
struct FrEnv { int cap, want_ctas, force_tree; };
static const FrEnv& fr_env() {
    static const FrEnv e = [] {
        FrEnv x{0, 0, 0};
        if (const char* s = getenv("UDA_FR_CAP")) x.cap = atoi(s);
        if (const char* s = getenv("UDA_FR_WANT_CTAS")) x.want_ctas = atoi(s);
        if (const char* s = getenv("UDA_FR_FORCE_TREE")) x.force_tree = atoi(s) != 0;
        return x;
    }();
    return e;
}
static int fr_cap_for(int k, int stride) {
    const FrEnv& e = fr_env();
    if (e.cap >= 32 && e.cap <= FR_TILE) return e.cap & ~31;
    int want = e.want_ctas >= 1 && e.want_ctas <= 8 ? e.want_ctas : UDA_FR_CTAS;
    int cap = fr_pick_cap(k, stride, want);
    while (want > 1 && fr_smem_bytes(cap, k, stride) > (size_t)232448 / (size_t)want - 1024 - 256) cap = fr_pick_cap(k, stride, --want);
    return cap;
}

}  // namespace uda

extern "C" long long uda_merge_records_scratch_bytes(long long n, int nruns) {
    if (n < 0 || nruns < 1) return UDA_E_INVALID;
    return (long long)uda::fr_layout(nullptr, n, nruns, uda::FR_CAP_MIN).bytes;   // upper bound over every tile capacity
}

extern "C" int uda_merge_records_supported(int nruns, long long n, int stride, int key_len) {
    using namespace uda;
    return nruns >= 1 && nruns <= FR_KMAX && n >= 0 && n < (1LL << 32) - FR_TILE && stride >= 16 && stride <= 128 && key_len >= 1 &&
           key_len <= 10 && n / fr_target(FR_CAP_MIN) + 2 <= scan4::MAX_ELEMS && fr_smem_bytes(FR_CAP_MIN, nruns, stride) <= 200 * 1024;
}

// Merge nruns sorted fixed-stride runs into `out` (n*stride bytes, + 2 bytes EOF marker if write_eof).
// Runs must not alias `out`; every run needs 32 readable bytes after its last record.
extern "C" int uda_merge_records_fixed(const UdaRunDesc* runs_dev, int nruns, long long n, long long max_nrec, int stride,
                                       int key_off, int key_len, int align2, void* out, int write_eof, void* scratch,
                                       int ctas_per_sm, cudaStream_t stream) {
    using namespace uda;
    if (!uda_merge_records_supported(nruns, n, stride, key_len)) return UDA_E_INVALID;
    if (align2 && ((stride | key_off) & 1)) return UDA_E_INVALID;
    if (ctas_per_sm < 0 || ctas_per_sm > 32) return UDA_E_INVALID;
    uint8_t* o = (uint8_t*)out;
    if (n > 0) {
        const int cap = fr_cap_for(nruns, stride);
        const int per = fr_target(cap) / FR_SAMPLE;     // samples per bucket
        FrScratch s = fr_layout(scratch, n, nruns, cap);
        const int G = fr_num_buckets(n, cap);
        // histogram bins: ~10 per bucket is plenty; a sub-partition of a pipelined step (10 M records) then clears and scans
        // 2 MB instead of 16 MB
        int nb = 1 << 16;
        while (nb < FR_NB && (long long)nb * 32 < n) nb <<= 1;
        fr_range_kernel<<<1, 256, 0, stream>>>(runs_dev, nruns, stride, key_off, key_len, nb, s.prm);
        UDA_CUDA(cudaMemsetAsync(s.hist, 0, (size_t)(nb + 1) * 4, stream));
        {
            long long ns = max_nrec / FR_SAMPLE;
            long long bx = (ns + 255) / 256;
            long long capb = (kNumSMs * 16 + nruns - 1) / nruns;
            if (capb < 1) capb = 1;
            if (bx > capb) bx = capb;
            if (bx < 1) bx = 1;
            dim3 grid((unsigned)bx, (unsigned)nruns);
            fr_sample_hist_kernel<<<grid, 256, 0, stream>>>(runs_dev, stride, key_off, key_len, s.prm, s.hist);
        }
        int rc = scan4::exclusive(s.hist, nb, s.hist, s.sums, stream);
        if (rc) return rc;
        {
            int bq = (G + 1 + 255) / 256;
            if (bq > kNumSMs * 8) bq = kNumSMs * 8;
            fr_bkeys_kernel<<<bq, 256, 0, stream>>>(G, per, nb, s.hist, s.prm, s.bkey, s.kind);
            if (nruns <= 16) {  // measured: search 0.5 ms at k = 8 (a boundary falls only every ~50 records of a run)
                long long t = (long long)(G + 1) * nruns;
                long long b = (t + 255) / 256;
                if (b > kNumSMs * 32) b = kNumSMs * 32;
                fr_partition_search_kernel<<<(int)b, 256, 0, stream>>>(runs_dev, nruns, G, stride, key_off, key_len, s.bkey, s.kind, s.P);
            } else {
                long long bx = (max_nrec + 255) / 256;
                long long capb = (kNumSMs * 16 + nruns - 1) / nruns;
                if (capb < 1) capb = 1;
                if (bx > capb) bx = capb;
                if (bx < 1) bx = 1;
                dim3 grid((unsigned)bx, (unsigned)nruns);
                fr_partition_kernel<<<grid, 256, 0, stream>>>(runs_dev, nruns, G, per, stride, key_off, key_len, s.hist, s.prm, s.bkey, s.kind, s.P);
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
            const size_t smem = fr_smem_bytes(cap, nruns, stride);
            void (*kern)(const UdaRunDesc*, int, int, int, int, int, int, int, const uint32_t*, const uint32_t*, const unsigned long long*,
                         const uint8_t*, FrParams*, uint8_t*);
            if (align2 && stride == 102) kern = fr_merge_kernel<true, 102>;
            else if (align2 && stride == 100) kern = fr_merge_kernel<true, 100>;
            else if (align2) kern = fr_merge_kernel<true, 0>;
            else kern = fr_merge_kernel<false, 0>;
            UDA_CUDA(ensure_dynamic_smem((const void*)kern, smem));
            int per_sm = 0;
            UDA_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, FR_THREADS_ALL, smem));
            if (per_sm < 1) return UDA_E_TOO_LARGE;
            if (ctas_per_sm > 0 && per_sm > ctas_per_sm) per_sm = ctas_per_sm;
            int grid = kNumSMs * per_sm;
            if (grid > G) grid = G;
            kern<<<grid, FR_THREADS_ALL, smem, stream>>>(runs_dev, nruns, G, cap, stride, key_off, key_len, fr_env().force_tree, s.P, s.sizes,
                                                     s.bkey, s.kind, s.prm, o);
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
// {oversize buckets, bisection iterations, tiles ordered by the merge tree} of the last merge; synchronises the stream
extern "C" int uda_merge_records_stats3(const void* scratch, unsigned long long* out3_host, cudaStream_t stream) {
    const uda::FrParams* p = (const uda::FrParams*)scratch;
    UDA_CUDA(cudaMemcpyAsync(out3_host, &p->fallback_buckets, 3 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, stream));
    UDA_CUDA(cudaStreamSynchronize(stream));
    return 0;
}
// tile capacity (record slots) and shared memory per CTA the merge would use for this shape
extern "C" int uda_merge_records_tile(int nruns, int stride, int* cap_out, long long* smem_out) {
    if (nruns < 1 || stride < 16) return UDA_E_INVALID;
    const int cap = uda::fr_cap_for(nruns, stride);
    if (cap_out) *cap_out = cap;
    if (smem_out) *smem_out = (long long)uda::fr_smem_bytes(cap, nruns, stride);
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
