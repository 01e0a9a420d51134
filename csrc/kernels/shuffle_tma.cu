// K1, copy-engine variant — the shuffle push driven by TMA bulk copies from a handful of SMs.
//
// Same protocol as shuffle.cu (segments stored into the destination rank's arena over NVLink, then a system-scope
// release add on that rank's landed-byte counter), different engine: every CTA is a few independent warps, each owning a small
// ring of 16 KiB shared-memory stages. A stage is filled with `cp.async.bulk.shared.global` (completion on an mbarrier) and
// drained with `cp.async.bulk.global.shared` straight into peer memory (completion through bulk async-groups), so
// the bytes in flight are set by the ring (192 KiB per SM), not by how many threads sit in loads.
//
// Why it exists: in the pipelined step the push overlaps the merge of the previous sub-partition. The LSU push
// kernel needs ~300 CTAs to cover NVLink latency; those CTAs take registers and issue slots from the merge on every
// SM and both kernels run at about half speed (profiles/pipeline_sweep.md). This kernel asks for 192 KiB of shared
// memory per CTA, so each CTA owns an SM and nothing else fits beside it: `num_ctas` SMs do the shuffle, the other
// 148 - num_ctas SMs run the merge undisturbed. All sub-partitions go through ONE launch, in order, so the SMs are
// claimed once at the start of the step.
//
// Reference behaviour being replaced: as in shuffle.cu (src/DataNet/RDMAServer.cc:537-631 RDMA-WRITE + ack,
// src/MOFServer/IndexInfo.cc:237-335 chunked read, src/DataNet/RDMAClient.cc:559-600 request).
#include "common.cuh"
#include "../../include/uda_b200.h"

namespace uda {
namespace {

#ifndef UDA_TP_WARPS
#define UDA_TP_WARPS 4
#endif
#ifndef UDA_TP_CHUNK_KB
#define UDA_TP_CHUNK_KB 16
#endif
#ifndef UDA_TP_STAGES
#define UDA_TP_STAGES 3
#endif
#ifndef UDA_TP_AHEAD
#define UDA_TP_AHEAD 2
#endif
constexpr int TP_WARPS = UDA_TP_WARPS;      // independent copy pipelines per CTA (one warp each): hides the per-chunk bookkeeping
constexpr int TP_CHUNK = UDA_TP_CHUNK_KB * 1024;
constexpr int TP_STAGES = UDA_TP_STAGES;    // per warp; TP_WARPS * TP_STAGES * TP_CHUNK = 192 KiB per CTA
constexpr int TP_AHEAD = UDA_TP_AHEAD;      // loads run this many chunks ahead of the stores: AHEAD chunks of loads and
                                            // STAGES - AHEAD chunks of stores are in flight per warp
constexpr int TP_MAXSUB = 16;
constexpr int TP_CACHE = 2048;   // chunk_prefix entries kept in shared memory
constexpr int TP_DCACHE = 256;   // descriptors kept in shared memory
constexpr int TP_FLUSH_MAX = 64;  // landed-byte counters are advanced once per `flush` (<= this) chunks of a warp ...
constexpr int TP_LAG = 32;        // ... and only for stores older than this many groups: those are waited for in full,
                                  // and a remote write under all-to-all load takes long to be acknowledged
constexpr int TP_HIST = TP_FLUSH_MAX + TP_LAG + 64;  // ring of (counter, bytes) of recent chunks (a batch of 32 is resolved ahead)

static_assert(TP_HIST > TP_FLUSH_MAX + TP_LAG + TP_AHEAD + TP_STAGES + 32, "history ring too small");
static_assert(TP_STAGES - 1 - TP_AHEAD >= 0, "stage ring too small");

struct TpSub {
    unsigned vstart[TP_MAXSUB + 1];  // virtual chunk index space: sub j owns [vstart[j], vstart[j+1])
    unsigned cstart[TP_MAXSUB + 1];  // first real chunk of sub j
    unsigned per[TP_MAXSUB];         // ceil(chunks of sub j / nway)
};

// chunk_prefix[j] = number of TP_CHUNK chunks in segments < j ; then the per-sub-partition tables
__global__ void __launch_bounds__(1024) tp_plan_kernel(const UdaPushDesc* __restrict__ d, int nseg, int nsub, unsigned nway,
                                                       unsigned* __restrict__ chunk_prefix, TpSub* __restrict__ sub) {
    __shared__ unsigned ws[32];
    __shared__ unsigned carry;
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    if (tid == 0) carry = 0;
    __syncthreads();
    for (int base = 0; base < nseg; base += 1024) {
        int j = base + tid;
        unsigned c = j < nseg ? (unsigned)((d[j].bytes + TP_CHUNK - 1) / TP_CHUNK) : 0u, x = c;
        for (int o = 1; o < 32; o <<= 1) {
            unsigned y = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= o) x += y;
        }
        if (lane == 31) ws[w] = x;
        __syncthreads();
        if (w == 0) {
            unsigned t = ws[lane], z = t;
            for (int o = 1; o < 32; o <<= 1) {
                unsigned y = __shfl_up_sync(0xffffffffu, z, o);
                if (lane >= o) z += y;
            }
            ws[lane] = z - t;
        }
        __syncthreads();
        unsigned excl = carry + ws[w] + x - c;
        if (j < nseg) chunk_prefix[j] = excl;
        __syncthreads();
        if (tid == 1023) carry = excl + c;
        __syncthreads();
    }
    if (tid == 0) {
        chunk_prefix[nseg] = carry;
        const int per_sub = nseg / nsub;
        unsigned v = 0;
        for (int j = 0; j < nsub; ++j) {
            const unsigned c0 = chunk_prefix[j * per_sub], c1 = j == nsub - 1 ? carry : chunk_prefix[(j + 1) * per_sub];
            const unsigned per = (c1 - c0 + nway - 1) / nway;
            sub->vstart[j] = v;
            sub->cstart[j] = c0;
            sub->per[j] = per;
            v += per * nway;
        }
        sub->vstart[nsub] = v;
        sub->cstart[nsub] = carry;
    }
}

__device__ __forceinline__ void fence_proxy_async_global() { asm volatile("fence.proxy.async.global;" ::: "memory"); }
__device__ __forceinline__ void red_add_relaxed_sys(unsigned long long* p, unsigned long long v) {
    asm volatile("red.relaxed.sys.global.add.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

struct TpHist {
    unsigned long long* flag[TP_HIST];
    unsigned len[TP_HIST];
};
// Publishes chunks [published, upto): one system-scope release fence covers them all, the counters are then
// advanced with relaxed reductions (a fence per chunk would serialise this one-warp CTA on the fence latency).
// The caller has waited for the bulk groups of those chunks to complete in full (not just .read).
__device__ __forceinline__ void tp_publish(const TpHist& h, int& published, int upto) {
    if (upto <= published) return;
    fence_proxy_async_global();
    __threadfence_system();
    for (int c = published; c < upto; ++c) {
        const int i = c % TP_HIST;
        if (h.len[i] && h.flag[i]) red_add_relaxed_sys(h.flag[i], (unsigned long long)h.len[i]);
    }
    published = upto;
}

struct TpStage {
    uint8_t* dst_body;
    unsigned body, pad;
};
// One resolved chunk: where its 16-byte aligned body comes from and goes to (head / tail bytes were already copied
// by the resolving lane).
struct TpDesc {
    const uint8_t* src_body;
    uint8_t* dst_body;
    unsigned body, pad;
};
constexpr int TP_BATCH = 32;  // chunks resolved at once, one per lane

// Receive ring (credit != nullptr): the destination arenas hold only `slots` sub-partitions. Before a warp touches
// sub-partition j it finishes and publishes everything it has in flight (a reducer may be waiting for exactly those bytes
// before it can merge and hand out the credit this warp is about to wait for), then waits until every reducer has merged
// the sub-partition that used the slot before: credit[p] >= need0 + j, written by reducer p with a system-scope release
// (uda_credit_publish). All ranks push in sub-partition order, so the oldest outstanding sub-partition can always complete.
__global__ void __launch_bounds__(32 * TP_WARPS, 1) push_tma_kernel(const UdaPushDesc* __restrict__ d, int nseg, int nsub, unsigned nway,
                                                         const unsigned* __restrict__ chunk_prefix, const TpSub* __restrict__ subg, int flush,
                                                         const unsigned long long* __restrict__ credit, int world, long long need0,
                                                         unsigned long long max_spins, unsigned* __restrict__ status) {
    extern __shared__ __align__(128) uint8_t tp_smem[];
    __shared__ __align__(8) uint64_t mbar_all[TP_WARPS][TP_STAGES];
    __shared__ TpStage stage_all[TP_WARPS][TP_STAGES];
    __shared__ TpHist hist_all[TP_WARPS];
    __shared__ TpDesc tab_all[TP_WARPS][TP_BATCH];
    __shared__ TpSub sub;
    __shared__ unsigned cp_cache[TP_CACHE];
    __shared__ UdaPushDesc dcache[TP_DCACHE];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, tid = threadIdx.x, nthr = 32 * TP_WARPS;
    uint64_t* mbar = mbar_all[wid];
    TpStage* stage = stage_all[wid];
    TpHist& hist = hist_all[wid];
    TpDesc* tab = tab_all[wid];
    uint8_t* ring = tp_smem + (size_t)wid * TP_STAGES * TP_CHUNK;
    for (int i = tid; i < (int)(sizeof(TpSub) / 4); i += nthr) ((unsigned*)&sub)[i] = ((const unsigned*)subg)[i];
    const bool cached = nseg + 1 <= TP_CACHE;
    if (cached)
        for (int i = tid; i <= nseg; i += nthr) cp_cache[i] = chunk_prefix[i];
    const bool dcached = nseg <= TP_DCACHE;
    if (dcached)
        for (int i = tid; i < nseg * (int)(sizeof(UdaPushDesc) / 8); i += nthr)
            ((unsigned long long*)dcache)[i] = ((const unsigned long long*)d)[i];
    const UdaPushDesc* dd = dcached ? dcache : d;
    if (lane == 0) {
        for (int s = 0; s < TP_STAGES; ++s) mbar_init(&mbar[s], 1);
        mbar_fence_init();
    }
    __syncthreads();
    const unsigned* cp = cached ? cp_cache : chunk_prefix;
    const unsigned vtotal = sub.vstart[nsub];
    const unsigned vfirst = blockIdx.x * TP_WARPS + wid, vstride = gridDim.x * TP_WARPS;
    const int n_my = vfirst < vtotal ? (int)((vtotal - vfirst + vstride - 1) / vstride) : 0;
    int published = 0;  // lane 0 only: chunks [0, published) of this warp have had their counters advanced

    int it_begin = 0;
    for (int js = 0; js < nsub; ++js) {
        // this warp's chunks of sub-partition js: `it` in [it_begin, it_end)
        const unsigned ve = sub.vstart[js + 1];
        const int it_end = ve > vfirst ? (int)((ve - vfirst + vstride - 1) / vstride) : 0;
        if (credit != nullptr) {
            int ok = 1;
            if (lane == 0) {
                tma_store_wait<0>();                       // everything issued so far has landed ...
                tp_publish(hist, published, it_begin);     // ... and is visible to its reducer
                const long long need = need0 + js;
                if (need > 0) {
                    for (int p = 0; p < world && ok; ++p) {
                        unsigned long long spins = 0;
                        while ((long long)ld_acquire_sys(credit + p) < need) {
                            if (++spins > max_spins) { ok = 0; atomicExch(status, 3u); break; }
                            __nanosleep(500);
                        }
                    }
                }
            }
            ok = __shfl_sync(0xffffffffu, ok, 0);
            if (!ok) return;                               // a reducer never freed its slot: give up (status word = 3)
        }
        for (int base = it_begin; base < it_end + TP_AHEAD; base += TP_BATCH) {
            // ---- all lanes: resolve chunk base+lane (which bytes is virtual chunk v?) and copy its unaligned fringes
            const int it_l = base + lane;
            {
                unsigned body = 0, len = 0;
                unsigned long long* flag = nullptr;
                uint8_t* dst_body = nullptr;
                const uint8_t* src_body = nullptr;
                if (it_l < it_end) {
                    const unsigned v = vfirst + (unsigned)it_l * vstride;
                    int j = 0;
                    while (j + 1 < nsub && v >= sub.vstart[j + 1]) ++j;
                    const unsigned li = v - sub.vstart[j];
                    const unsigned rel = (li % nway) * sub.per[j] + li / nway;
                    const unsigned cnt = sub.cstart[j + 1] - sub.cstart[j];
                    if (rel < cnt) {
                        const unsigned c = sub.cstart[j] + rel;
                        int lo = 0, hi = nseg - 1;  // last segment with chunk_prefix[seg] <= c
                        while (lo < hi) {
                            int mid = (lo + hi + 1) >> 1;
                            if (cp[mid] <= c) lo = mid; else hi = mid - 1;
                        }
                        const UdaPushDesc sd = dd[lo];
                        const long long off = (long long)(c - cp[lo]) * TP_CHUNK;
                        len = (unsigned)(sd.bytes - off < TP_CHUNK ? sd.bytes - off : TP_CHUNK);
                        const uint8_t* src = (const uint8_t*)sd.src + off;
                        uint8_t* dst = (uint8_t*)sd.dst + off;
                        flag = (unsigned long long*)sd.flag;
                        if ((((uintptr_t)src ^ (uintptr_t)dst) & 15) != 0) {
                            for (unsigned i = 0; i < len; ++i) dst[i] = src[i];  // incompatible phases (not produced by our layouts)
                        } else {
                            unsigned head = (unsigned)((16 - ((uintptr_t)src & 15)) & 15);
                            if (head > len) head = len;
                            body = (len - head) & ~15u;
                            for (unsigned i = 0; i < head; ++i) dst[i] = src[i];
                            for (unsigned i = head + body; i < len; ++i) dst[i] = src[i];
                            src_body = src + head;
                            dst_body = dst + head;
                        }
                    }
                    hist.flag[it_l % TP_HIST] = flag;
                    hist.len[it_l % TP_HIST] = len;
                }
                tab[lane].src_body = src_body;
                tab[lane].dst_body = dst_body;
                tab[lane].body = body;
            }
            __syncwarp();
            // ---- lane 0: the copy pipeline over this batch
            if (lane == 0) {
                const int end = base + TP_BATCH < it_end + TP_AHEAD ? base + TP_BATCH : it_end + TP_AHEAD;
                for (int it = base; it < end; ++it) {
                    if (it < it_end) {
                        const int s = it % TP_STAGES;
                        // the stage is free again once the store that used it has READ it; landing is tracked separately
                        if (it >= TP_STAGES) tma_store_wait_read<TP_STAGES - 1 - TP_AHEAD>();
                        if (it % flush == 0 && it - TP_AHEAD - TP_LAG > published) {
                            tma_store_wait<TP_LAG>();   // groups committed so far: chunks [0, it - TP_AHEAD)
                            tp_publish(hist, published, it - TP_AHEAD - TP_LAG);
                        }
                        const TpDesc t = tab[it - base];
                        stage[s].body = t.body;
                        stage[s].dst_body = t.dst_body;
                        mbar_expect_tx(&mbar[s], t.body);
                        if (t.body) tma_load_1d(ring + (size_t)s * TP_CHUNK, t.src_body, t.body, &mbar[s]);
                    }
                    const int jt = it - TP_AHEAD;
                    if (jt >= it_begin && jt < it_end) {
                        const int s2 = jt % TP_STAGES;
                        mbar_wait(&mbar[s2], (uint32_t)(jt / TP_STAGES) & 1u);
                        if (stage[s2].body) tma_store_1d(stage[s2].dst_body, ring + (size_t)s2 * TP_CHUNK, stage[s2].body);
                        tma_store_commit();  // one group per chunk, empty or not, so that wait_group counts chunks
                    }
                }
            }
            __syncwarp();
        }
        it_begin = it_end;
    }
    if (lane == 0) {
        tma_store_wait<0>();
        tp_publish(hist, published, n_my);
    }
}

}  // namespace
}  // namespace uda

extern "C" long long uda_push_tma_scratch_bytes(int nseg) {
    if (nseg < 0) return UDA_E_INVALID;
    return ((long long)(nseg + 1) * 4 + 15) / 16 * 16 + (long long)sizeof(uda::TpSub) + 64;
}

static int push_tma_launch(const UdaPushDesc* descs_dev, int nseg, int nsub, int nway, void* scratch, int num_ctas, int flush_chunks,
                           const unsigned long long* credit, int world, long long need0, unsigned long long max_spins, unsigned* status,
                           cudaStream_t stream) {
    using namespace uda;
    if (flush_chunks == 0) flush_chunks = 8;
    if (flush_chunks < 1 || flush_chunks > TP_FLUSH_MAX) return UDA_E_INVALID;
    if (nseg <= 0) return 0;
    if (nsub < 1 || nsub > TP_MAXSUB || nseg % nsub != 0 || nway < 1) return UDA_E_INVALID;
    if (num_ctas <= 0) num_ctas = 32;
    if (num_ctas > kNumSMs) num_ctas = kNumSMs;
    unsigned* chunk_prefix = (unsigned*)scratch;
    TpSub* sub = (TpSub*)((uint8_t*)scratch + ((size_t)(nseg + 1) * 4 + 15) / 16 * 16);
    const int smem = TP_WARPS * TP_STAGES * TP_CHUNK;
    // the opt-in is a per-device attribute of the function
    UDA_CUDA(cudaFuncSetAttribute(push_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    tp_plan_kernel<<<1, 1024, 0, stream>>>(descs_dev, nseg, nsub, (unsigned)nway, chunk_prefix, sub);
    push_tma_kernel<<<num_ctas, 32 * TP_WARPS, smem, stream>>>(descs_dev, nseg, nsub, (unsigned)nway, chunk_prefix, sub, flush_chunks, credit, world,
                                                              need0, max_spins, status);
    UDA_CHECK_LAUNCH();
    return 0;
}

extern "C" int uda_shuffle_push_tma(const UdaPushDesc* descs_dev, int nseg, int nsub, int nway, void* scratch, int num_ctas,
                                    int flush_chunks, cudaStream_t stream) {
    return push_tma_launch(descs_dev, nseg, nsub, nway, scratch, num_ctas, flush_chunks, nullptr, 0, 0, 0, nullptr, stream);
}

extern "C" int uda_shuffle_push_tma_ring(const UdaPushDesc* descs_dev, int nseg, int nsub, int nway, void* scratch, int num_ctas,
                                         int flush_chunks, const unsigned long long* credits_dev, int world, long long need0,
                                         unsigned long long max_spins, unsigned* status_dev, cudaStream_t stream) {
    if (!credits_dev || !status_dev || world < 1 || world > 64) return UDA_E_INVALID;
    return push_tma_launch(descs_dev, nseg, nsub, nway, scratch, num_ctas, flush_chunks, credits_dev, world, need0, max_spins, status_dev, stream);
}

namespace uda {
// reducer side of the ring: one more sub-partition merged -> the new count goes to every sender's credit slot
__global__ void credit_publish_kernel(unsigned long long* const* __restrict__ slots, int world, unsigned long long* __restrict__ counter) {
    __shared__ unsigned long long v;
    if (threadIdx.x == 0) { v = *counter + 1ull; *counter = v; }
    __syncthreads();
    if ((int)threadIdx.x < world) st_release_sys(slots[threadIdx.x], v);
}
}  // namespace uda

extern "C" int uda_credit_publish(unsigned long long* const* credit_slots_dev, int world, unsigned long long* counter_dev, cudaStream_t stream) {
    if (world < 1 || world > 64 || !credit_slots_dev || !counter_dev) return UDA_E_INVALID;
    uda::credit_publish_kernel<<<1, 64, 0, stream>>>(credit_slots_dev, world, counter_dev);
    UDA_CHECK_LAUNCH();
    return 0;
}
