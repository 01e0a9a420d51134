// K9 — LZO1X block decompression of compressed map-output segments on the device.
//
// Reference behaviour being replaced: LzoDecompressor (src/Merger/LzoDecompressor.cc:121-167) hands one block at a
// time to a dlopen'ed liblzo2 on the reducer's decompress thread; LZO is the codec the reference's install guide
// recommends. The block framing is Hadoop's BlockCompressorStream framing, the same as for Snappy, so the header scan
// of snappy.cu (uda_snappy_scan_blocks) builds the block table and this file only adds the per-block decoder:
// one warp per block, every lane runs the parser of csrc/runtime/lzo1x_parse.h over the same bytes (uniform control
// flow, the bytes come out of L1) and takes its lane's share of each literal run / match. The parser and the per-lane
// copy halves are the code the host runtime decodes with (and the CPU tests exercise, lane emulation included); the
// only device-specific part is the __syncwarp() that publishes an instruction's bytes to the next instruction.
#include "block_table.cuh"
#include "common.cuh"
#include "../runtime/lzo1x_parse.h"
#include "../../include/uda_b200.h"

namespace uda {
namespace {

struct WarpCopier {
    unsigned lane;
    __host__ __device__ __forceinline__ void literals(uint8_t* dst, const uint8_t* src, size_t len) {
        lzo1x_lane_literals(dst, src, len, lane, 32u);
#ifdef __CUDA_ARCH__
        __syncwarp();   // the next instruction may read what this one wrote
#endif
    }
    __host__ __device__ __forceinline__ void match(uint8_t* dst, size_t dist, size_t len) {
        lzo1x_lane_match(dst, dist, len, lane, 32u);
#ifdef __CUDA_ARCH__
        __syncwarp();
#endif
    }
};

constexpr int LZ_WARPS = 4;

__global__ void __launch_bounds__(32 * LZ_WARPS) lzo_decode_kernel(const uint8_t* __restrict__ seg, const CodecBlock* __restrict__ tab,
                                                                   const unsigned long long* __restrict__ nblocks_dev, uint8_t* out,
                                                                   unsigned* __restrict__ status) {
    const unsigned lane = threadIdx.x & 31;
    const unsigned long long nblocks = *nblocks_dev;
    const unsigned long long warp = (unsigned long long)blockIdx.x * LZ_WARPS + (threadIdx.x >> 5);
    const unsigned long long nwarps = (unsigned long long)gridDim.x * LZ_WARPS;
    for (unsigned long long b = warp; b < nblocks; b += nwarps) {
        const CodecBlock bk = tab[b];
        WarpCopier copier{lane};
        // every lane computes the same verdict: the parse is a function of the input bytes alone
        const bool ok = lzo1x_decode_block(seg + bk.src_off, (size_t)bk.comp, out + bk.dst_off, (size_t)bk.raw, copier);
        if (!ok && lane == 0) atomicExch(status, 3u);
        __syncwarp();
    }
}

}  // namespace
}  // namespace uda

// Decodes every block of the table (built by uda_snappy_scan_blocks: the framing is codec independent) into `out`,
// which must hold the total plain bytes the scan reported. status_dev becomes 3 when a block is corrupt (instruction
// out of bounds, distance before the block's first byte, length mismatch, missing end marker). A block's matches
// never reach before its own first byte: blocks are compressed independently (BlockCompressorStream).
extern "C" int uda_lzo_decode_blocks(const void* seg, const void* table, const unsigned long long* nblocks_dev, int max_blocks, void* out,
                                     unsigned* status_dev, cudaStream_t stream) {
    using namespace uda;
    if (max_blocks <= 0) return 0;
    long long ctas = ((long long)max_blocks + LZ_WARPS - 1) / LZ_WARPS;
    if (ctas > kNumSMs * 8LL) ctas = kNumSMs * 8LL;
    lzo_decode_kernel<<<(int)ctas, 32 * LZ_WARPS, 0, stream>>>((const uint8_t*)seg, (const CodecBlock*)table, nblocks_dev, (uint8_t*)out, status_dev);
    UDA_CHECK_LAUNCH();
    return 0;
}
