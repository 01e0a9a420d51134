// LZO1X block decoder, one source for the host runtime and the device kernel.
//
// The instruction stream of a block is a serial chain (every instruction's position depends on the one before it),
// so decoding splits into a parser that walks the chain and a copier that moves the bytes an instruction names.
// The parser below is plain integer code without any host- or device-only construct: `lzo1x.cc` runs it with a
// memcpy copier, `csrc/kernels/lzo.cu` runs it with all 32 lanes of a warp parsing the same bytes (uniform control
// flow) and copying cooperatively. `lzo1x_lane_*` are the per-lane halves of that cooperative copy; the host can run
// them lane by lane in any order (`Lzo1xLaneEmulation`), which is how the CPU tests check the device copier's logic
// without a GPU.
//
// Reference: src/Merger/LzoDecompressor.cc:121-167 calls liblzo2's lzo1x_decompress_safe per block; stream format in
// lzo1x.cc's header comment.
#pragma once
#include <stddef.h>
#include <stdint.h>
#include <string.h>

#if defined(__CUDACC__)
#define UDA_LZO_HD __host__ __device__ __forceinline__
#else
#define UDA_LZO_HD inline
#endif

namespace uda {

// ---- per-lane copy halves (lane in [0, lanes)) ----------------------------------------------------------------------
UDA_LZO_HD void lzo1x_lane_literals(uint8_t* dst, const uint8_t* src, size_t len, unsigned lane, unsigned lanes) {
    for (size_t i = lane; i < len; i += lanes) dst[i] = src[i];
}
// dst[i] = dst[i - dist] for i in [0, len), defined byte-serially by the format. When dist < len the match overlaps its
// own output and is a repetition of the dist bytes before dst, so byte i equals (dst - dist)[i % dist], which every lane
// can read before any lane has written: no lane depends on another lane's store.
UDA_LZO_HD void lzo1x_lane_match(uint8_t* dst, size_t dist, size_t len, unsigned lane, unsigned lanes) {
    const uint8_t* from = dst - dist;
    if (dist >= len) {
        for (size_t i = lane; i < len; i += lanes) dst[i] = from[i];
    } else {
        for (size_t i = lane; i < len; i += lanes) dst[i] = from[i % dist];
    }
}

// ---- parser ---------------------------------------------------------------------------------------------------------
// Copier concept: void literals(uint8_t* dst, const uint8_t* src, size_t len); void match(uint8_t* dst, size_t dist,
// size_t len). Both must make their bytes visible to the next call before they return.
// Returns true when the block decodes to exactly out_len bytes and consumes exactly n input bytes.
template <class Copier>
UDA_LZO_HD bool lzo1x_decode_block(const uint8_t* in, size_t n, uint8_t* out, size_t out_len, Copier& copier) {
    size_t ip = 0, op = 0;
    unsigned state = 0;

// zero-byte run-length extension of a length field: BASE + 255 per zero byte + the next non-zero byte
#define UDA_LZO_EXTENDED(BASE, LEN)                                              \
    do {                                                                          \
        size_t t__ = 0;                                                           \
        while (ip < n && in[ip] == 0) {                                           \
            t__ += 255;                                                           \
            ++ip;                                                                 \
            if (t__ > (1u << 30)) return false;                                   \
        }                                                                         \
        if (ip >= n) return false;                                                \
        (LEN) = (size_t)(BASE) + t__ + in[ip++];                                  \
    } while (0)
#define UDA_LZO_LITERALS(LEN)                                                     \
    do {                                                                          \
        const size_t l__ = (LEN);                                                 \
        if (ip + l__ > n || op + l__ > out_len) return false;                     \
        copier.literals(out + op, in + ip, l__);                                  \
        ip += l__;                                                                \
        op += l__;                                                                \
    } while (0)

    if (n < 1) return false;
    if (in[0] > 17) {
        const size_t t = (size_t)in[ip++] - 17;
        UDA_LZO_LITERALS(t);
        state = t < 4 ? (unsigned)t : 4u;
    }
    for (;;) {
        if (ip >= n) return false;
        const unsigned t = in[ip++];
        size_t len = 0, dist = 0;
        unsigned next_state = 0;
        if (t < 16) {
            if (state == 0) {
                if (t == 0) UDA_LZO_EXTENDED(15, len);
                else len = t;
                UDA_LZO_LITERALS(len + 3);
                state = 4;
                continue;
            }
            if (ip >= n) return false;
            const unsigned h = in[ip++];
            dist = ((size_t)h << 2) + (t >> 2) + (state == 4 ? 2049 : 1);
            len = state == 4 ? 3 : 2;
            next_state = t & 3;
        } else if (t < 32) {
            if ((t & 7) == 0) UDA_LZO_EXTENDED(7, len);
            else len = t & 7;
            len += 2;
            if (ip + 2 > n) return false;
            const unsigned w = in[ip] | ((unsigned)in[ip + 1] << 8);
            ip += 2;
            dist = 16384 + ((size_t)(t & 8) << 11) + (w >> 2);
            next_state = w & 3;
            if (dist == 16384) return op == out_len && ip == n;   // end-of-stream marker
        } else if (t < 64) {
            if ((t & 31) == 0) UDA_LZO_EXTENDED(31, len);
            else len = t & 31;
            len += 2;
            if (ip + 2 > n) return false;
            const unsigned w = in[ip] | ((unsigned)in[ip + 1] << 8);
            ip += 2;
            dist = (size_t)(w >> 2) + 1;
            next_state = w & 3;
        } else {
            if (ip >= n) return false;
            const unsigned h = in[ip++];
            dist = ((size_t)h << 3) + ((t >> 2) & 7) + 1;
            len = t < 128 ? 3 + ((t >> 5) & 1) : 5 + ((t >> 5) & 3);
            next_state = t & 3;
        }
        if (dist == 0 || dist > op || op + len > out_len) return false;
        copier.match(out + op, dist, len);
        op += len;
        if (next_state) UDA_LZO_LITERALS((size_t)next_state);
        state = next_state;
    }
#undef UDA_LZO_EXTENDED
#undef UDA_LZO_LITERALS
}

// Host copiers ----------------------------------------------------------------------------------------------------------
struct Lzo1xHostCopier {
    void literals(uint8_t* dst, const uint8_t* src, size_t len) { memcpy(dst, src, len); }
    void match(uint8_t* dst, size_t dist, size_t len) {
        for (size_t i = 0; i < len; ++i) dst[i] = dst[i - dist];   // may overlap: byte-serial by definition
    }
};
// Runs the device copier's per-lane halves one lane after the other, highest lane first: equal output proves that no
// lane reads a byte another lane of the same instruction writes.
struct Lzo1xLaneEmulation {
    unsigned lanes;
    void literals(uint8_t* dst, const uint8_t* src, size_t len) {
        for (unsigned l = lanes; l-- > 0;) lzo1x_lane_literals(dst, src, len, l, lanes);
    }
    void match(uint8_t* dst, size_t dist, size_t len) {
        for (unsigned l = lanes; l-- > 0;) lzo1x_lane_match(dst, dist, len, l, lanes);
    }
};

}  // namespace uda
