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

// ---- warp-cooperative fast path (stride <= 112) ------------------------------------------------------
// A warp owns 32 consecutive *output* records (32*stride bytes, always a multiple of 16). Eight lanes
// fetch one record as aligned 128-bit vectors (four records per load instruction), deposit the bytes at
// their packed position in a per-warp shared-memory span, and the span then leaves as fully coalesced
// 128-bit streaming stores. Per-record work (entry load, pointer arithmetic) is paid once per record
// instead of once per 16 output bytes, which is what made the word-per-thread kernel issue bound.
constexpr int GW_THREADS = 256;
constexpr int GW_WARPS = GW_THREADS / 32;
constexpr int GW_MAX_STRIDE = 112;  // (15 + stride) <= 128: eight 16-byte vectors always cover a record

template <int STRIDE>
__global__ void __launch_bounds__(GW_THREADS) gather_fixed_warp_kernel(const Entry* __restrict__ S, long long n,
                                                                      const unsigned long long* __restrict__ run_base,
                                                                      int stride_rt, uint8_t* __restrict__ dst) {
    const int stride = STRIDE > 0 ? STRIDE : stride_rt;
    extern __shared__ __align__(16) uint8_t gw_smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint8_t* span = gw_smem + (size_t)warp * (32 * GW_MAX_STRIDE);
    const long long ngroups = (n + 31) >> 5;
    const int sub = lane & 7, recsel = lane >> 3;
    for (long long g = (long long)blockIdx.x * GW_WARPS + warp; g < ngroups; g += (long long)gridDim.x * GW_WARPS) {
        const long long j = (g << 5) + lane;
        unsigned long long ptr = 0;
        if (j < n) ptr = (unsigned long long)(uintptr_t)record_ptr(S, j, run_base, stride);
        const int cnt = (int)((n - (g << 5)) < 32 ? (n - (g << 5)) : 32);
#pragma unroll
        for (int it = 0; it < 8; ++it) {
            const int rec = it * 4 + recsel;
            const unsigned long long a = __shfl_sync(0xffffffffu, ptr, rec);
            if (rec < cnt) {
                const int mis = (int)(a & 15);
                const int q0 = 16 * sub - mis;  // packed byte offset (inside the record) of this lane's vector
                if (q0 < stride) {
                    const int4 v = __ldg(reinterpret_cast<const int4*>(a - mis) + sub);
                    uint8_t* d = span + rec * stride + q0;
                    if (((a | (unsigned)stride) & 1) == 0) {  // 2-byte granular (TeraSort: everything is even)
                        const uint32_t w[4] = {(uint32_t)v.x, (uint32_t)v.y, (uint32_t)v.z, (uint32_t)v.w};
#pragma unroll
                        for (int h = 0; h < 8; ++h) {
                            const int q = q0 + 2 * h;
                            if (q >= 0 && q < stride)
                                *reinterpret_cast<uint16_t*>(d + 2 * h) = (uint16_t)(w[h >> 1] >> (16 * (h & 1)));
                        }
                    } else {
                        const uint32_t w[4] = {(uint32_t)v.x, (uint32_t)v.y, (uint32_t)v.z, (uint32_t)v.w};
#pragma unroll
                        for (int h = 0; h < 16; ++h) {
                            const int q = q0 + h;
                            if (q >= 0 && q < stride) d[h] = (uint8_t)(w[h >> 2] >> (8 * (h & 3)));
                        }
                    }
                }
            }
        }
        __syncwarp();
        const int bytes = cnt * stride;
        uint8_t* out = dst + (g << 5) * stride;
        const int nw = bytes >> 4;
        for (int w = lane; w < nw; w += 32)
            st_stream(reinterpret_cast<int4*>(out) + w, *reinterpret_cast<const int4*>(span + 16 * w));
        for (int b = (nw << 4) + lane; b < bytes; b += 32) out[b] = span[b];
        __syncwarp();
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
        if (stride <= GW_MAX_STRIDE) {
            const size_t smem = (size_t)GW_WARPS * 32 * GW_MAX_STRIDE;
            long long groups = (n + 31) / 32;
            long long b = (groups + GW_WARPS - 1) / GW_WARPS;
            if (b > kNumSMs * 8) b = kNumSMs * 8;
            if (stride == 102) gather_fixed_warp_kernel<102><<<(int)b, GW_THREADS, smem, stream>>>(S, n, run_base_dev, stride, (uint8_t*)dst);
            else if (stride == 100) gather_fixed_warp_kernel<100><<<(int)b, GW_THREADS, smem, stream>>>(S, n, run_base_dev, stride, (uint8_t*)dst);
            else gather_fixed_warp_kernel<0><<<(int)b, GW_THREADS, smem, stream>>>(S, n, run_base_dev, stride, (uint8_t*)dst);
        } else {
            long long words = total >> 4;
            long long b = (words + 255) / 256;
            if (b > kNumSMs * 32) b = kNumSMs * 32;
            if (b < 1) b = 1;
            gather_fixed_kernel<0><<<(int)b, 256, 0, stream>>>(S, n, run_base_dev, stride, (uint8_t*)dst, total);
        }
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
