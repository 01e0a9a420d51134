// Synthetic teragen-shaped records (10 B key + 90 B value) generated on the device, and the
// TeraValidate-equivalent checker (global order + order-independent record fingerprint).
//
// The reference has no generator: its regression drives Hadoop's teragen / teravalidate
// (scripts/regression/sortcountRunner.sh:123-241, scripts/regression/mr-dstatExcel.sh:246-340).
// BASELINE.json asks for synthetic teragen-shaped 100-byte records with no datasets, so both live
// here; uda_b200/models/teragen.py holds the bit-identical NumPy restatement used as oracle.
#include "common.cuh"
#include "../../include/uda_b200.h"

namespace uda {

__host__ __device__ __forceinline__ unsigned long long mix64(unsigned long long z) {
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

// byte j (0..99) of raw record with global row id r
__device__ __forceinline__ uint32_t teragen_byte(unsigned long long a, unsigned long long b, unsigned long long r, int j,
                                              bool tail_zero) {
    if (j < 8) return (uint32_t)(a >> (8 * (7 - j))) & 0xFF;
    if (j < 10) return tail_zero ? 0u : (uint32_t)(b >> (8 * (7 - (j - 8)))) & 0xFF;
    if (j < 18) return (uint32_t)(r >> (8 * (7 - (j - 10)))) & 0xFF;
    return ((uint32_t)(b >> (8 * (j & 7))) + (uint32_t)j) & 0xFF;
}

// one thread per 2-byte unit; framed records carry the IFile header VInt(10)=0x0A, VInt(90)=0x5A
__global__ void teragen_kernel(uint8_t* __restrict__ out, long long n, unsigned long long seed, unsigned long long first_row,
                               int framed, unsigned long long key_mask_hi) {
    const int units = framed ? 51 : 50;
    const bool tz = key_mask_hi != ~0ull;  // reduced key entropy (duplicate-key tests): zero the 2-byte tail too
    long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    long long total = n * units;
    for (; t < total; t += (long long)gridDim.x * blockDim.x) {
        long long i = t / units;
        int u = (int)(t - i * units);
        unsigned long long r = first_row + (unsigned long long)i;
        unsigned long long a = mix64(seed + 2 * r) & key_mask_hi, b = mix64(seed + 2 * r + 1);
        uint32_t lo, hi;
        if (framed) {
            if (u == 0) { lo = 10; hi = 90; }
            else { lo = teragen_byte(a, b, r, 2 * u - 2, tz); hi = teragen_byte(a, b, r, 2 * u - 1, tz); }
        } else {
            lo = teragen_byte(a, b, r, 2 * u, tz); hi = teragen_byte(a, b, r, 2 * u + 1, tz);
        }
        reinterpret_cast<uint16_t*>(out)[t] = (uint16_t)(lo | (hi << 8));
    }
}

// Validation: counts adjacent misordered pairs (10-byte key memcmp order) and accumulates an
// order-independent fingerprint sum_i mix64(first 8 key bytes ^ mix64(row id bytes ^ tail)).
__device__ __forceinline__ unsigned long long load_be(const uint8_t* p, int n) {
    unsigned long long v = 0;
    for (int i = 0; i < n; ++i) v = (v << 8) | p[i];
    return v;
}

__device__ __forceinline__ unsigned long long record_fingerprint(const uint8_t* rec /* raw 100 B */) {
    unsigned long long h = 0;
    for (int j = 0; j < 96; j += 8) h = mix64(h ^ load_be(rec + j, 8));
    return mix64(h ^ load_be(rec + 96, 4));
}

__global__ void validate_kernel(const uint8_t* __restrict__ recs, long long n, int stride, int key_off,
                                unsigned long long* __restrict__ result /* [0]=misordered [1]=fingerprint [2]=count */) {
    unsigned long long bad = 0, fp = 0, cnt = 0;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const uint8_t* r = recs + i * stride + key_off;
        fp += record_fingerprint(r);
        ++cnt;
        if (i + 1 < n) {
            const uint8_t* q = r + stride;
            unsigned long long a0 = load_be(r, 8), b0 = load_be(q, 8);
            unsigned long long a1 = load_be(r + 8, 2), b1 = load_be(q + 8, 2);
            if (a0 > b0 || (a0 == b0 && a1 > b1)) ++bad;
        }
    }
    for (int o = 16; o; o >>= 1) {
        bad += __shfl_xor_sync(0xffffffffu, bad, o);
        fp += __shfl_xor_sync(0xffffffffu, fp, o);
        cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    }
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(&result[0], bad);
        atomicAdd(&result[1], fp);
        atomicAdd(&result[2], cnt);
    }
}

}  // namespace uda

extern "C" int uda_teragen(void* out, long long n, unsigned long long seed, unsigned long long first_row, int framed,
                           int key_bits, cudaStream_t stream) {
    if (n <= 0) return 0;
    unsigned long long mask = key_bits >= 64 || key_bits <= 0 ? ~0ull : ~0ull << (64 - key_bits);
    long long total = n * (framed ? 51 : 50);
    int blocks = (int)((total + 255) / 256 < 148LL * 16 ? (total + 255) / 256 : 148LL * 16);
    uda::teragen_kernel<<<blocks, 256, 0, stream>>>((uint8_t*)out, n, seed, first_row, framed, mask);
    UDA_CHECK_LAUNCH();
    return 0;
}

extern "C" int uda_validate_fixed(const void* recs, long long n, int stride, int key_off, unsigned long long* result3,
                                  cudaStream_t stream) {
    UDA_CUDA(cudaMemsetAsync(result3, 0, 3 * sizeof(unsigned long long), stream));
    if (n <= 0) return 0;
    int blocks = (int)((n + 255) / 256 < 148LL * 8 ? (n + 255) / 256 : 148LL * 8);
    uda::validate_kernel<<<blocks, 256, 0, stream>>>((const uint8_t*)recs, n, stride, key_off, result3);
    UDA_CHECK_LAUNCH();
    return 0;
}
