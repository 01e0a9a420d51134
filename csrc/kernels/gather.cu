// Record gather: materialise the merged stream by copying every record to its merged rank.
//
// Reference behaviour being replaced: the serialisation half of write_kv_to_stream()
// (src/Merger/StreamRW.cc:184-201: per record, write VInt(klen) VInt(vlen) key val into the 1 MiB
// staging buffer) plus the final EOF marker (src/Merger/StreamRW.cc:216-218).
//
// Fixed-stride fast path (TeraSort: 102-byte IFile-framed records, header bytes travel with the
// record so no re-encode is needed): one thread per 16-byte *destination* word. The word's bytes come
// from at most two source records at arbitrary byte alignment; each piece is fetched as five aligned
// 32-bit loads that are funnel-shifted into place, so the store side is a fully coalesced, streaming
// 128-bit store and no shared memory or block barrier is involved.
//
// General path (variable-length records): one warp per record, byte-granular copy, destination
// offsets from a prefix sum of record lengths.
#include "common.cuh"
#include "../../include/uda_b200.h"

namespace uda {

// 16 bytes starting at arbitrary address a (reads [a&~3, (a&~3)+20): up to 22 bytes past the record end,
// which is why source buffers need GATHER_SLACK readable bytes after their last record)
__device__ __forceinline__ uint4 load16_unaligned(const uint8_t* a) {
    const uintptr_t ai = (uintptr_t)a;
    const uint32_t* w = (const uint32_t*)(ai & ~(uintptr_t)3);
    const uint32_t sh = (uint32_t)(ai & 3) * 8;
    uint32_t x0 = __ldg(w), x1 = __ldg(w + 1), x2 = __ldg(w + 2), x3 = __ldg(w + 3), x4 = __ldg(w + 4);
    uint4 r;
    r.x = __funnelshift_r(x0, x1, sh);
    r.y = __funnelshift_r(x1, x2, sh);
    r.z = __funnelshift_r(x2, x3, sh);
    r.w = __funnelshift_r(x3, x4, sh);
    return r;
}

// same, but never touches memory below `floor` (4-byte aligned down): words below it read as zero.
// Used for the second piece of a straddling word, whose window starts before the record.
__device__ __forceinline__ uint4 load16_unaligned_floor(const uint8_t* a, const uint8_t* floor) {
    const uintptr_t ai = (uintptr_t)a;
    const uint32_t* w = (const uint32_t*)(ai & ~(uintptr_t)3);
    const uint32_t* f = (const uint32_t*)((uintptr_t)floor & ~(uintptr_t)3);
    const uint32_t sh = (uint32_t)(ai & 3) * 8;
    uint32_t x0 = w >= f ? __ldg(w) : 0u, x1 = w + 1 >= f ? __ldg(w + 1) : 0u, x2 = w + 2 >= f ? __ldg(w + 2) : 0u,
             x3 = w + 3 >= f ? __ldg(w + 3) : 0u, x4 = __ldg(w + 4);
    uint4 r;
    r.x = __funnelshift_r(x0, x1, sh);
    r.y = __funnelshift_r(x1, x2, sh);
    r.z = __funnelshift_r(x2, x3, sh);
    r.w = __funnelshift_r(x3, x4, sh);
    return r;
}

__device__ __forceinline__ uint32_t low_bytes_mask(int nbytes) {  // nbytes in [0,4]
    return nbytes >= 4 ? 0xFFFFFFFFu : (nbytes <= 0 ? 0u : ((1u << (8 * nbytes)) - 1u));
}

__device__ __forceinline__ const uint8_t* record_ptr(const Entry* __restrict__ S, long long j,
                                                     const unsigned long long* __restrict__ run_base, int stride) {
    unsigned long long w1 = __ldg(&S[j].w1);
    uint32_t run = (uint32_t)(w1 >> 32) & 0xFFFFu, idx = (uint32_t)w1;
    return (const uint8_t*)(uintptr_t)__ldg(&run_base[run]) + (unsigned long long)idx * (unsigned)stride;
}

template <int STRIDE>
__global__ void __launch_bounds__(256) gather_fixed_kernel(const Entry* __restrict__ S, long long n,
                                                           const unsigned long long* __restrict__ run_base, int stride_rt,
                                                           uint8_t* __restrict__ dst, long long total_bytes) {
    const int stride = STRIDE > 0 ? STRIDE : stride_rt;
    const long long nwords = total_bytes >> 4;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < nwords; t += (long long)gridDim.x * blockDim.x) {
        const long long byte0 = t << 4;
        const long long j0 = byte0 / stride;
        const int o0 = (int)(byte0 - j0 * stride);
        const int len0 = stride - o0 < 16 ? stride - o0 : 16;
        uint4 v = load16_unaligned(record_ptr(S, j0, run_base, stride) + o0);
        if (len0 < 16) {  // word straddles two records (stride >= 16 so never three)
            const uint8_t* nx = record_ptr(S, j0 + 1, run_base, stride);
            uint4 u = load16_unaligned_floor(nx - len0, nx);
            uint32_t m0 = low_bytes_mask(len0), m1 = low_bytes_mask(len0 - 4), m2 = low_bytes_mask(len0 - 8),
                     m3 = low_bytes_mask(len0 - 12);
            v.x = (v.x & m0) | (u.x & ~m0);
            v.y = (v.y & m1) | (u.y & ~m1);
            v.z = (v.z & m2) | (u.z & ~m2);
            v.w = (v.w & m3) | (u.w & ~m3);
        }
        int4 o = make_int4((int)v.x, (int)v.y, (int)v.z, (int)v.w);
        st_stream(reinterpret_cast<int4*>(dst) + t, o);
    }
    // tail (< 16 bytes) and optional trailer are written by the first thread of block 0
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        for (long long b = nwords << 4; b < total_bytes; ++b) {
            long long j = b / stride;
            int o = (int)(b - j * stride);
            dst[b] = record_ptr(S, j, run_base, stride)[o];
        }
    }
}

// EOF marker VInt(-1) VInt(-1) = 0xFF 0xFF after the last record
__global__ void write_eof_kernel(uint8_t* dst) { dst[0] = 0xFF; dst[1] = 0xFF; }

// ---- general path -------------------------------------------------------------------------------
// rec_len[j] = encoded length of merged record j ; dst_off = exclusive prefix (both by merged rank)
__global__ void gather_var_len_kernel(const Entry* __restrict__ S, long long n,
                                      const unsigned long long* __restrict__ run_recoff /* per run: long long* */,
                                      uint32_t* __restrict__ rec_len) {
    for (long long j = (long long)blockIdx.x * blockDim.x + threadIdx.x; j < n; j += (long long)gridDim.x * blockDim.x) {
        unsigned long long w1 = S[j].w1;
        uint32_t run = (uint32_t)(w1 >> 32) & 0xFFFFu, idx = (uint32_t)w1;
        const long long* ro = (const long long*)(uintptr_t)run_recoff[run];
        rec_len[j] = (uint32_t)(ro[idx + 1] - ro[idx]);
    }
}

__global__ void gather_var_copy_kernel(const Entry* __restrict__ S, long long n,
                                       const unsigned long long* __restrict__ run_base,
                                       const unsigned long long* __restrict__ run_recoff,
                                       const unsigned long long* __restrict__ dst_off, uint8_t* __restrict__ dst) {
    const int lane = threadIdx.x & 31;
    const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
    for (long long j = warp; j < n; j += nwarps) {
        unsigned long long w1 = S[j].w1;
        uint32_t run = (uint32_t)(w1 >> 32) & 0xFFFFu, idx = (uint32_t)w1;
        const long long* ro = (const long long*)(uintptr_t)run_recoff[run];
        const long long s = ro[idx];
        const int len = (int)(ro[idx + 1] - s);
        const uint8_t* src = (const uint8_t*)(uintptr_t)run_base[run] + s;
        uint8_t* d = dst + dst_off[j];
        for (int b = lane; b < len; b += 32) d[b] = src[b];
    }
}

}  // namespace uda

extern "C" int uda_gather_fixed(const void* sorted_entries, long long n, const unsigned long long* run_base_dev, int stride,
                                void* dst, int write_eof, cudaStream_t stream) {
    using namespace uda;
    if (stride < 16) return UDA_E_INVALID;
    if (((uintptr_t)dst & 15) != 0) return UDA_E_INVALID;
    const long long total = n * stride;
    if (n > 0) {
        const Entry* S = (const Entry*)sorted_entries;
        long long words = total >> 4;
        long long b = (words + 255) / 256;
        if (b > kNumSMs * 32) b = kNumSMs * 32;
        if (b < 1) b = 1;
        // (a warp-cooperative shared-memory variant was measured at 6.8 ms vs 6.3 ms for this kernel on the
        //  125 M record TeraSort step: both sit at ~80 % of the copy roofline, see profiles/)
        if (stride == 102) gather_fixed_kernel<102><<<(int)b, 256, 0, stream>>>(S, n, run_base_dev, stride, (uint8_t*)dst, total);
        else if (stride == 100) gather_fixed_kernel<100><<<(int)b, 256, 0, stream>>>(S, n, run_base_dev, stride, (uint8_t*)dst, total);
        else gather_fixed_kernel<0><<<(int)b, 256, 0, stream>>>(S, n, run_base_dev, stride, (uint8_t*)dst, total);
    }
    if (write_eof) write_eof_kernel<<<1, 1, 0, stream>>>((uint8_t*)dst + total);
    UDA_CHECK_LAUNCH();
    return 0;
}

extern "C" int uda_gather_var_lengths(const void* sorted_entries, long long n, const unsigned long long* run_recoff_dev,
                                      unsigned* rec_len_out, cudaStream_t stream) {
    if (n <= 0) return 0;
    int b = (int)((n + 255) / 256 < 148LL * 16 ? (n + 255) / 256 : 148LL * 16);
    uda::gather_var_len_kernel<<<b, 256, 0, stream>>>((const uda::Entry*)sorted_entries, n, run_recoff_dev, rec_len_out);
    UDA_CHECK_LAUNCH();
    return 0;
}

extern "C" int uda_gather_var(const void* sorted_entries, long long n, const unsigned long long* run_base_dev,
                              const unsigned long long* run_recoff_dev, const unsigned long long* dst_off_dev, void* dst,
                              cudaStream_t stream) {
    if (n <= 0) return 0;
    long long warps = n;
    long long b = (warps * 32 + 255) / 256;
    if (b > 148LL * 16) b = 148LL * 16;
    uda::gather_var_copy_kernel<<<(int)b, 256, 0, stream>>>((const uda::Entry*)sorted_entries, n, run_base_dev, run_recoff_dev,
                                                          dst_off_dev, (uint8_t*)dst);
    UDA_CHECK_LAUNCH();
    return 0;
}
