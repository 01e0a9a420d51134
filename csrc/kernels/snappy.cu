// K9 — block decompression of compressed map-output segments on the device (Snappy raw format inside Hadoop's
// block framing `[u32 BE rawLen][u32 BE compLen][compLen bytes]`, repeated).
//
// Reference behaviour being replaced: DecompressorWrapper (src/Merger/DecompressorWrapper.cc:168-197) runs a
// decompress thread per reducer that takes the RDMA landing buffer, decodes one block at a time with a dlopen'ed
// libsnappy / liblzo2 (src/Merger/SnappyDecompressor.cc:89-105, src/Merger/LzoDecompressor.cc:151-167) into a cyclic
// buffer the merge reads from. Here the compressed segment lands in the arena as it is, and two kernels turn it into
// the plain IFile segment the merge kernels expect:
//   1. snappy_scan_kernel   one warp walks the block headers (a serial chain: every header position depends on
//                           the previous compLen) and writes a table {src offset, compLen, rawLen, dst offset}
//   2. snappy_decode_kernel one warp per block: all lanes parse the same element header (uniform control flow,
//                           the bytes come out of L1), then copy the literal / the match cooperatively. A match that
//                           overlaps its own output (offset < length) is a repeating pattern and is written as
//                           out[i] = out[start - offset + i % offset], which needs no byte-serial loop.
// Blocks are independent, so a segment of N blocks keeps N warps busy; the element chain inside a block is serial
// by the format's definition.
#include "block_table.cuh"
#include "common.cuh"
#include "../../include/uda_b200.h"

namespace uda {
namespace {

using SnBlock = CodecBlock;   // block_table.cuh: {src offset, dst offset, compLen, rawLen}

__device__ __forceinline__ unsigned be32(const uint8_t* p) {
    return ((unsigned)p[0] << 24) | ((unsigned)p[1] << 16) | ((unsigned)p[2] << 8) | (unsigned)p[3];
}

// status: 0 ok, 1 table overflow, 2 corrupt header
__global__ void snappy_scan_kernel(const uint8_t* __restrict__ seg, long long seg_len, int max_blocks, SnBlock* __restrict__ tab,
                                   unsigned long long* __restrict__ out2 /* [0]=blocks [1]=total raw */, unsigned* __restrict__ status) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    long long pos = 0;
    unsigned long long dst = 0;
    int n = 0;
    while (pos + 8 <= seg_len) {
        const unsigned raw = be32(seg + pos), comp = be32(seg + pos + 4);
        if (raw == 0 && comp == 0) break;
        if (raw > (1u << 30) || comp > (1u << 30) || pos + 8 + (long long)comp > seg_len) { atomicExch(status, 2u); break; }
        if (n >= max_blocks) { atomicExch(status, 1u); break; }
        tab[n].src_off = (unsigned long long)(pos + 8);
        tab[n].dst_off = dst;
        tab[n].comp = comp;
        tab[n].raw = raw;
        ++n;
        dst += raw;
        pos += 8 + (long long)comp;
    }
    out2[0] = (unsigned long long)n;
    out2[1] = dst;
}

constexpr int SN_WARPS = 4;

__global__ void __launch_bounds__(32 * SN_WARPS) snappy_decode_kernel(const uint8_t* __restrict__ seg, const SnBlock* __restrict__ tab,
                                                                      const unsigned long long* __restrict__ nblocks_dev,
                                                                      uint8_t* __restrict__ out, unsigned* __restrict__ status) {
    const int lane = threadIdx.x & 31;
    const unsigned long long nblocks = *nblocks_dev;
    const unsigned long long warp = (unsigned long long)blockIdx.x * SN_WARPS + (threadIdx.x >> 5);
    const unsigned long long nwarps = (unsigned long long)gridDim.x * SN_WARPS;
    for (unsigned long long b = warp; b < nblocks; b += nwarps) {
        const SnBlock bk = tab[b];
        const uint8_t* in = seg + bk.src_off;
        uint8_t* o = out + bk.dst_off;
        const unsigned n = bk.comp, out_len = bk.raw;
        unsigned ip = 0, op = 0;
        // varint: uncompressed length of this block
        unsigned long long ulen = 0;
        bool bad = false;
        for (int shift = 0;; shift += 7) {
            if (ip >= n || shift > 35) { bad = true; break; }
            const unsigned c = in[ip++];
            ulen |= (unsigned long long)(c & 0x7F) << shift;
            if (!(c & 0x80)) break;
        }
        if (ulen != out_len) bad = true;
        while (!bad && ip < n) {
            const unsigned tag = in[ip++];
            unsigned len, off = 0;
            const unsigned kind = tag & 3;
            if (kind == 0) {
                len = (tag >> 2) + 1;
                if (len > 60) {
                    const unsigned nb = len - 60;
                    if (ip + nb > n) { bad = true; break; }
                    len = 0;
                    for (unsigned i = 0; i < nb; ++i) len |= (unsigned)in[ip + i] << (8 * i);
                    len += 1;
                    ip += nb;
                }
                if (ip + len > n || op + len > out_len) { bad = true; break; }
                for (unsigned i = lane; i < len; i += 32) o[op + i] = in[ip + i];
                ip += len;
            } else {
                if (kind == 1) {
                    if (ip + 1 > n) { bad = true; break; }
                    len = ((tag >> 2) & 7) + 4;
                    off = ((tag >> 5) << 8) | in[ip];
                    ip += 1;
                } else if (kind == 2) {
                    if (ip + 2 > n) { bad = true; break; }
                    len = (tag >> 2) + 1;
                    off = in[ip] | ((unsigned)in[ip + 1] << 8);
                    ip += 2;
                } else {
                    if (ip + 4 > n) { bad = true; break; }
                    len = (tag >> 2) + 1;
                    off = in[ip] | ((unsigned)in[ip + 1] << 8) | ((unsigned)in[ip + 2] << 16) | ((unsigned)in[ip + 3] << 24);
                    ip += 4;
                }
                if (off == 0 || off > op || op + len > out_len) { bad = true; break; }
                const uint8_t* from = o + op - off;
                if (off >= len) {
                    for (unsigned i = lane; i < len; i += 32) o[op + i] = from[i];
                } else {
                    for (unsigned i = lane; i < len; i += 32) o[op + i] = from[i % off];   // self-overlapping match = repeating pattern
                }
            }
            op += len;
            __syncwarp();   // the next element may read what this one wrote
        }
        if (bad || op != out_len) {
            if (lane == 0) atomicExch(status, 3u);
        }
        __syncwarp();
    }
}

}  // namespace
}  // namespace uda

extern "C" long long uda_snappy_table_bytes(int max_blocks) {
    if (max_blocks < 0) return UDA_E_INVALID;
    return (long long)max_blocks * (long long)sizeof(uda::SnBlock) + 64;
}

// out2_dev: {blocks, total plain bytes}; status_dev: 0 ok, 1 more than max_blocks blocks, 2 corrupt block header
extern "C" int uda_snappy_scan_blocks(const void* seg, long long seg_len, int max_blocks, void* table, unsigned long long* out2_dev,
                                      unsigned* status_dev, cudaStream_t stream) {
    if (seg_len < 0 || max_blocks < 0) return UDA_E_INVALID;
    uda::snappy_scan_kernel<<<1, 32, 0, stream>>>((const uint8_t*)seg, seg_len, max_blocks, (uda::SnBlock*)table, out2_dev, status_dev);
    UDA_CHECK_LAUNCH();
    return 0;
}

// Decodes every block of the table into `out` (which must hold the total plain bytes reported by the scan).
// status_dev becomes 3 when a block is corrupt (bad varint, element out of bounds, length mismatch).
extern "C" int uda_snappy_decode_blocks(const void* seg, const void* table, const unsigned long long* nblocks_dev, int max_blocks, void* out,
                                        unsigned* status_dev, cudaStream_t stream) {
    using namespace uda;
    if (max_blocks <= 0) return 0;
    long long ctas = ((long long)max_blocks + SN_WARPS - 1) / SN_WARPS;
    if (ctas > kNumSMs * 8LL) ctas = kNumSMs * 8LL;
    snappy_decode_kernel<<<(int)ctas, 32 * SN_WARPS, 0, stream>>>((const uint8_t*)seg, (const SnBlock*)table, nblocks_dev, (uint8_t*)out, status_dev);
    UDA_CHECK_LAUNCH();
    return 0;
}
