// Shared device helpers for the sm_100a kernels: sort-entry layout, Hadoop VInt codec,
// mbarrier / TMA bulk-copy / system-scope acquire-release PTX wrappers.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <map>
#include <mutex>
#include <utility>

#define UDA_CHECK_LAUNCH()                                  \
    do {                                                    \
        cudaError_t e__ = cudaGetLastError();               \
        if (e__ != cudaSuccess) return (int)e__;            \
    } while (0)

#define UDA_CUDA(call)                                      \
    do {                                                    \
        cudaError_t e__ = (call);                           \
        if (e__ != cudaSuccess) return (int)e__;            \
    } while (0)

namespace uda {

constexpr int kNumSMs = 148;  // B200: 2 dies x 74 SMs

// Opt a kernel in to `smem` bytes of dynamic shared memory. The limit is an attribute of (function, device), and a
// launch fails when it is below the launch's request, so it only ever grows here: two host threads launching different
// shapes of one kernel on one device cannot lower each other's limit between the set and the launch. Set on every call
// (a few hundred ns) so that a device whose context was reset is opted in again.
inline cudaError_t ensure_dynamic_smem(const void* kernel, size_t smem) {
    static std::mutex mu;
    static std::map<std::pair<const void*, int>, size_t> high;
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    std::lock_guard<std::mutex> lock(mu);
    size_t& h = high[std::make_pair(kernel, dev)];
    if (smem > h) h = smem;
    return cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)h);
}

// ---------------------------------------------------------------------------------------------
// Sort entry: 16 bytes, compared as the 128-bit big integer (w0:w1).
//   w0 = key bytes 0..7, big-endian (order-preserving)
//   w1 = key bytes 8..9 << 48 | run id << 32 | record index inside the run
// (run, idx) as the low bits makes every entry unique and the merge stable by run order, which is
// one of the orders the reference heap can produce (src/Merger/MergeQueue.h:250-269).
// ---------------------------------------------------------------------------------------------
struct __align__(16) Entry {
    unsigned long long w0;
    unsigned long long w1;
};

__host__ __device__ __forceinline__ bool entry_less(const Entry& a, const Entry& b) {
    return (a.w0 < b.w0) || (a.w0 == b.w0 && a.w1 < b.w1);
}
__host__ __device__ __forceinline__ uint32_t entry_run(const Entry& e) { return (uint32_t)(e.w1 >> 32) & 0xFFFFu; }
__host__ __device__ __forceinline__ uint32_t entry_idx(const Entry& e) { return (uint32_t)e.w1; }

// ---------------------------------------------------------------------------------------------
// Hadoop VInt (same rules as uda_b200/utils/vint.py; reference: src/CommUtils/IOUtility.cc:167-196,
// :287-333, :389-396).
// ---------------------------------------------------------------------------------------------
__host__ __device__ __forceinline__ int vint_size_from_first(int8_t b) {
    if (b >= -112) return 1;
    return (b < -120) ? (-119 - b) : (-111 - b);
}

__host__ __device__ __forceinline__ int vint_size(long long v) {
    if (v >= -112 && v <= 127) return 1;
    if (v < 0) v = ~v;
    int bits = 0;
    unsigned long long t = (unsigned long long)v;
    while (t) { t >>= 1; ++bits; }
    return (bits + 7) / 8 + 1;
}

// Decodes at p (caller guarantees the bytes are readable). Returns encoded size.
__host__ __device__ __forceinline__ int vint_decode(const uint8_t* p, long long* out) {
    int8_t b = (int8_t)p[0];
    if (b >= -112) { *out = b; return 1; }
    bool neg = b < -120;
    int n = neg ? (-120 - b) : (-112 - b);
    unsigned long long t = 0;
    for (int i = 0; i < n; ++i) t = (t << 8) | p[1 + i];
    *out = neg ? (long long)~t : (long long)t;
    return n + 1;
}

__host__ __device__ __forceinline__ int vint_encode(uint8_t* p, long long v) {
    if (v >= -112 && v <= 127) { p[0] = (uint8_t)v; return 1; }
    int tag = -112;
    if (v < 0) { v = ~v; tag = -120; }
    int n = 0;
    for (unsigned long long t = (unsigned long long)v; t; t >>= 8) ++n;
    p[0] = (uint8_t)(tag - n);
    for (int i = 0; i < n; ++i) p[1 + i] = (uint8_t)((unsigned long long)v >> (8 * (n - 1 - i)));
    return n + 1;
}

#ifdef __CUDACC__
// ---------------------------------------------------------------------------------------------
// PTX wrappers
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t phase) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" ::"r"(smem_u32(bar)),
        "r"(phase)
        : "memory");
}
// 1-D TMA bulk copy global -> shared, completion counted on an mbarrier (SASS: UBLKCP).
// dst, src and bytes must be multiples of 16.
__device__ __forceinline__ void tma_load_1d(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(smem_dst)),
        "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}
// 1-D TMA bulk copy shared -> global (bulk async-group completion).
__device__ __forceinline__ void tma_store_1d(void* gdst, const void* smem_src, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_u32(smem_src)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void tma_store_wait_read() {
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
template <int N>
__device__ __forceinline__ void tma_store_wait() {
    asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}
// make generic-proxy smem writes visible to the async proxy (before a TMA store reads them)
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// streaming 16-byte global accesses
__device__ __forceinline__ int4 ld_stream(const int4* p) {
    int4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}
__device__ __forceinline__ void st_stream(int4* p, const int4& v) {
    asm volatile("st.global.L1::no_allocate.v4.s32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
                 : "memory");
}

// system-scope release/acquire used for the cross-GPU "tile landed" protocol
__device__ __forceinline__ void red_add_release_sys(unsigned long long* p, unsigned long long v) {
    asm volatile("red.release.sys.global.add.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// first 10 compared key bytes -> (w0, 16-bit tail); shorter keys are zero padded
__device__ __forceinline__ void key_prefix(const uint8_t* k, int len, unsigned long long* w0, uint32_t* tail16) {
    unsigned long long a = 0;
    uint32_t t = 0;
#pragma unroll
    for (int i = 0; i < 8; ++i) a = (a << 8) | (i < len ? k[i] : 0u);
#pragma unroll
    for (int i = 8; i < 10; ++i) t = (t << 8) | (i < len ? k[i] : 0u);
    *w0 = a;
    *tail16 = t;
}

__device__ __forceinline__ unsigned long long bswap64(unsigned long long x) {
    uint32_t lo = (uint32_t)x, hi = (uint32_t)(x >> 32);
    return ((unsigned long long)__byte_perm(lo, 0, 0x0123) << 32) | __byte_perm(hi, 0, 0x0123);
}
#endif  // __CUDACC__

}  // namespace uda
