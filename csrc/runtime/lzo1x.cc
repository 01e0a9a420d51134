// Built-in LZO1X block decompressor (and a small compressor for tests and tools).
//
// The reference decompresses LZO map outputs with a dlopen'ed liblzo2 (src/Merger/LzoDecompressor.cc:66-167; LZO is
// the codec its install guide recommends) and falls back to vanilla Hadoop when the library is missing. liblzo2 is
// still preferred here when it can be loaded; this file removes the hard dependency.
//
// Stream format (as documented for the decompressor in the Linux kernel, Documentation/staging/lzo.rst). Every
// instruction ends by saying how many literal bytes (0..3, "state") follow it before the next instruction:
//   first byte   18..21  copy 0..3 literals (byte - 17); 22..255 copy (byte - 17) literals, state = 4;
//                0..17   is an ordinary instruction
//   0000 LLLL    state 0 (previous instruction left no literals): literal run of 3 + L bytes, L == 0 -> 15 +
//                255 per zero byte + the next non-zero byte; state = 4
//                state 1..3: 2-byte match, distance = (H << 2) + D + 1          [0000 DDSS  H]
//                state 4   : 3-byte match, distance = (H << 2) + D + 2049       [0000 DDSS  H]
//   0001 HLLL    match of 2 + L bytes (L == 0 -> 7 + extension), distance = 16384 + (H << 14) + D, followed by a
//                little-endian 16-bit word DDDDDDDD DDDDDDSS; distance == 16384 ends the stream
//   001L LLLL    match of 2 + L bytes (L == 0 -> 31 + extension), distance = D + 1, same 16-bit word
//   01LD DDSS    match of 3 + L bytes, distance = (H << 3) + D + 1, followed by H
//   1LLD DDSS    match of 5 + L bytes, distance = (H << 3) + D + 1, followed by H
// The decoder itself lives in lzo1x_parse.h, shared with the device kernel (csrc/kernels/lzo.cu).
#include <string.h>

#include <cstdint>
#include <vector>

#include "consumer.h"
#include "lzo1x_parse.h"

namespace uda {

bool lzo1x_decompress_builtin(const uint8_t* in, size_t n, uint8_t* out, size_t out_len) {
    Lzo1xHostCopier copier;
    return lzo1x_decode_block(in, n, out, out_len, copier);
}

// The same parser with the device kernel's cooperative copy run lane by lane on the host (tests: the warp copier's
// logic is checked against the byte-serial one without a GPU).
bool lzo1x_decompress_lanes(const uint8_t* in, size_t n, uint8_t* out, size_t out_len, unsigned lanes) {
    if (lanes < 1) return false;
    Lzo1xLaneEmulation copier{lanes};
    return lzo1x_decode_block(in, n, out, out_len, copier);
}

// ---- compressor (greedy, hash of 4 bytes; matches are emitted in the 001L LLLL form, distance <= 16384) -------------
namespace {
void put_literal_run(const uint8_t* p, size_t len, bool first, std::vector<uint8_t>* out) {
    // precondition: len > 0 and the previous instruction left state 0 (or this is the very first instruction)
    if (first && len <= 238) {
        out->push_back((uint8_t)(17 + len));
    } else if (len <= 3) {
        // a run of 1..3 bytes cannot be expressed as a 0000 LLLL run (its minimum is 4): the caller attaches such runs
        // to the preceding match's state bits instead; reaching here would be a logic error
        out->push_back(0);  // never taken
    } else if (len <= 18) {
        out->push_back((uint8_t)(len - 3));
    } else {
        size_t t = len - 18;
        out->push_back(0);
        while (t > 255) { out->push_back(0); t -= 255; }
        out->push_back((uint8_t)t);
    }
    out->insert(out->end(), p, p + len);
}
void put_match(size_t dist, size_t len, unsigned trailing, std::vector<uint8_t>* out) {
    // 001L LLLL, length = 2 + L (extension when L == 0), 16-bit word = (dist - 1) << 2 | trailing literals
    if (len <= 33) {
        out->push_back((uint8_t)(32 | (len - 2)));
    } else {
        size_t t = len - 33;
        out->push_back(32);
        while (t > 255) { out->push_back(0); t -= 255; }
        out->push_back((uint8_t)t);
    }
    const unsigned w = (unsigned)((dist - 1) << 2) | trailing;
    out->push_back((uint8_t)(w & 0xFF));
    out->push_back((uint8_t)(w >> 8));
}
}  // namespace

size_t lzo1x_compress(const uint8_t* in, size_t n, std::vector<uint8_t>* out) {
    out->clear();
    // pass 1: greedy parse into (literal run, match) pairs
    struct Tok { size_t lit_pos, lit_len, dist, mlen; };
    std::vector<Tok> toks;
    constexpr int kBits = 14;
    std::vector<uint32_t> table(1u << kBits, 0xFFFFFFFFu);
    auto load32 = [&](size_t p) { uint32_t v; memcpy(&v, in + p, 4); return v; };
    size_t ip = 0, lit = 0;
    while (ip + 4 <= n) {
        const uint32_t h = (load32(ip) * 0x1E35A7BDu) >> (32 - kBits);
        const uint32_t cand = table[h];
        table[h] = (uint32_t)ip;
        if (cand != 0xFFFFFFFFu && ip - cand <= 16384 && load32(cand) == load32(ip)) {
            size_t len = 4;
            while (ip + len < n && in[cand + len] == in[ip + len]) ++len;
            toks.push_back({lit, ip - lit, ip - cand, len});
            ip += len;
            lit = ip;
        } else {
            ++ip;
        }
    }
    // pass 2: emit. A literal run of 1..3 bytes between two matches rides in the state bits of the match before it;
    // a short run with no match before it (start of the stream) uses the first-byte form; a short run elsewhere is
    // avoided by shortening the previous match so that the run reaches 4 bytes.
    bool first = true;
    unsigned carried = 0;   // literals already announced by the previous match's state bits
    for (size_t i = 0; i < toks.size(); ++i) {
        Tok& t = toks[i];
        if (t.lit_len) {
            if (carried) {
                out->insert(out->end(), in + t.lit_pos, in + t.lit_pos + t.lit_len);   // announced by the previous match
            } else {
                put_literal_run(in + t.lit_pos, t.lit_len, first, out);
            }
        }
        first = false;
        // how many literals follow this match (before the next match or the end)?
        size_t next_lit = i + 1 < toks.size() ? toks[i + 1].lit_len : n - (t.lit_pos + t.lit_len + t.mlen);
        if (next_lit >= 1 && next_lit <= 3) {
            carried = (unsigned)next_lit;
        } else {
            carried = 0;
        }
        put_match(t.dist, t.mlen, carried, out);
    }
    // tail literals
    const size_t tail_pos = toks.empty() ? 0 : toks.back().lit_pos + toks.back().lit_len + toks.back().mlen;
    const size_t tail = n - tail_pos;
    if (tail) {
        if (carried) out->insert(out->end(), in + tail_pos, in + n);
        else put_literal_run(in + tail_pos, tail, first, out);
    }
    out->push_back(0x11);   // 0001 0001 + distance word 0: end of stream
    out->push_back(0);
    out->push_back(0);
    return out->size();
}

}  // namespace uda
