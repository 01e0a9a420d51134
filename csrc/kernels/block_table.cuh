// Block table of a compressed segment in Hadoop's block framing `[u32 BE rawLen][u32 BE compLen][compLen bytes]`,
// written by the header scan (snappy.cu: uda_snappy_scan_blocks — the framing does not depend on the codec) and read by
// the per-codec block decoders (snappy.cu, lzo.cu).
#pragma once
#include <stdint.h>

namespace uda {

struct CodecBlock {
    unsigned long long src_off;  // offset of the compressed bytes inside the segment
    unsigned long long dst_off;  // offset of the plain bytes in the output
    unsigned comp, raw;
};

}  // namespace uda
