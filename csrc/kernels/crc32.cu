// K7 — CRC32 (IEEE 802.3, reflected 0xEDB88320, zlib compatible) of a device buffer.
//
// The reference never computes it: it fetches partLength bytes (src/MOFServer/IndexInfo.cc:308-310)
// but stops parsing at the EOF marker (src/Merger/StreamRW.cc:361-367), so Hadoop's trailing 4-byte
// IFile checksum is transferred and ignored. This engine *writes* MOF segments, so it must produce the
// checksum (and can optionally verify fetched segments).
//
// Parallelisation: every thread computes the zero-init remainder of one 2 KiB chunk with slice-by-4
// tables held in shared memory, multiplies it by x^(8*bytes_after_chunk) mod P (GF(2) square and
// multiply), and the contributions are XOR-reduced; CRC is linear, so
//   crc(M) = ~( x^(8*len)*0xFFFFFFFF  ^  XOR_i x^(8*after_i)*rem(chunk_i) ).
#include "common.cuh"
#include "../../include/uda_b200.h"

namespace uda {

constexpr uint32_t CRC_POLY = 0xEDB88320u;
constexpr int CRC_CHUNK = 2048;

// a*b mod P in the reflected representation (bit 31 = x^0)
__host__ __device__ __forceinline__ uint32_t gf2_mulmod(uint32_t a, uint32_t b) {
    uint32_t r = 0;
    for (int i = 0; i < 32; ++i) {
        if (a & 0x80000000u) r ^= b;
        a <<= 1;
        b = (b >> 1) ^ ((b & 1u) ? CRC_POLY : 0u);
    }
    return r;
}
// x^(8*nbytes) mod P
__host__ __device__ __forceinline__ uint32_t gf2_xpow8(unsigned long long nbytes) {
    uint32_t result = 0x80000000u;  // 1
    uint32_t base = 0x00800000u;    // x^8
    while (nbytes) {
        if (nbytes & 1ull) result = gf2_mulmod(result, base);
        base = gf2_mulmod(base, base);
        nbytes >>= 1;
    }
    return result;
}

__global__ void __launch_bounds__(256) crc32_kernel(const uint8_t* __restrict__ data, long long n,
                                                    uint32_t* __restrict__ acc) {
    __shared__ uint32_t T[4][256];
    {
        uint32_t c = threadIdx.x;
        for (int k = 0; k < 8; ++k) c = (c >> 1) ^ ((c & 1u) ? CRC_POLY : 0u);
        T[0][threadIdx.x] = c;
        uint32_t c1 = (c >> 8) ^ 0;  // filled after sync below
        (void)c1;
    }
    __syncthreads();
    {
        uint32_t c = T[0][threadIdx.x];
        for (int t = 1; t < 4; ++t) {
            c = T[0][c & 0xFF] ^ (c >> 8);
            T[t][threadIdx.x] = c;
        }
    }
    __syncthreads();
    const long long nchunks = (n + CRC_CHUNK - 1) / CRC_CHUNK;
    uint32_t mine = 0;
    for (long long ch = (long long)blockIdx.x * blockDim.x + threadIdx.x; ch < nchunks; ch += (long long)gridDim.x * blockDim.x) {
        const long long beg = ch * CRC_CHUNK;
        const long long end = beg + CRC_CHUNK < n ? beg + CRC_CHUNK : n;
        const uint8_t* p = data + beg;
        const uint8_t* e = data + end;
        uint32_t crc = 0;
        while (p < e && ((uintptr_t)p & 3)) crc = T[0][(crc ^ *p++) & 0xFF] ^ (crc >> 8);
        while (p + 4 <= e) {
            crc ^= __ldg(reinterpret_cast<const uint32_t*>(p));
            crc = T[3][crc & 0xFF] ^ T[2][(crc >> 8) & 0xFF] ^ T[1][(crc >> 16) & 0xFF] ^ T[0][crc >> 24];
            p += 4;
        }
        while (p < e) crc = T[0][(crc ^ *p++) & 0xFF] ^ (crc >> 8);
        mine ^= gf2_mulmod(crc, gf2_xpow8((unsigned long long)(n - end)));
    }
    for (int o = 16; o; o >>= 1) mine ^= __shfl_xor_sync(0xffffffffu, mine, o);
    if ((threadIdx.x & 31) == 0 && mine) atomicXor(acc, mine);
}

__global__ void crc32_finish_kernel(uint32_t* acc, long long n, uint32_t* out, uint8_t* out_be4) {
    uint32_t v = ~(*acc ^ gf2_mulmod(0xFFFFFFFFu, gf2_xpow8((unsigned long long)n)));
    if (out) *out = v;
    if (out_be4) { out_be4[0] = (uint8_t)(v >> 24); out_be4[1] = (uint8_t)(v >> 16); out_be4[2] = (uint8_t)(v >> 8); out_be4[3] = (uint8_t)v; }
}

}  // namespace uda

// acc_scratch: 4 bytes of device scratch. out_dev (u32, host order) and/or out_be4_dev (4 bytes
// big-endian, i.e. the IFile trailer) may be NULL.
extern "C" int uda_crc32(const void* data, long long n, unsigned* acc_scratch, unsigned* out_dev, void* out_be4_dev,
                         cudaStream_t stream) {
    using namespace uda;
    if (n < 0) return UDA_E_INVALID;
    UDA_CUDA(cudaMemsetAsync(acc_scratch, 0, 4, stream));
    if (n > 0) {
        long long chunks = (n + CRC_CHUNK - 1) / CRC_CHUNK;
        long long b = (chunks + 255) / 256;
        if (b > kNumSMs * 8) b = kNumSMs * 8;
        crc32_kernel<<<(int)b, 256, 0, stream>>>((const uint8_t*)data, n, acc_scratch);
    }
    crc32_finish_kernel<<<1, 1, 0, stream>>>(acc_scratch, n, out_dev, (uint8_t*)out_be4_dev);
    UDA_CHECK_LAUNCH();
    return 0;
}
