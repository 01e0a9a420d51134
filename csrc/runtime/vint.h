// Host-side Hadoop VInt codec and a bounds-checked byte cursor.
// Reference behaviour: src/CommUtils/IOUtility.cc:167-196 (encode), :287-333 (decode),
// :228-273 (decode across two buffers), :367-396 (sizes); DataStream src/CommUtils/IOUtility.cc:47-158.
#pragma once
#include <cstddef>
#include <cstdint>
#include <cstring>

namespace uda {

inline int vint_size(int64_t v) {
    if (v >= -112 && v <= 127) return 1;
    if (v < 0) v = ~v;
    int bits = 0;
    for (uint64_t t = (uint64_t)v; t; t >>= 1) ++bits;
    return (bits + 7) / 8 + 1;
}

inline int vint_size_from_first(int8_t b) {
    if (b >= -112) return 1;
    return b < -120 ? -119 - b : -111 - b;
}

inline int vint_encode(uint8_t* p, int64_t v) {
    if (v >= -112 && v <= 127) { p[0] = (uint8_t)v; return 1; }
    int tag = -112;
    if (v < 0) { v = ~v; tag = -120; }
    int n = 0;
    for (uint64_t t = (uint64_t)v; t; t >>= 8) ++n;
    p[0] = (uint8_t)(tag - n);
    for (int i = 0; i < n; ++i) p[1 + i] = (uint8_t)((uint64_t)v >> (8 * (n - 1 - i)));
    return n + 1;
}

// Decode from [p, end). Returns bytes consumed, or 0 when the value is truncated.
inline int vint_decode(const uint8_t* p, const uint8_t* end, int64_t* out) {
    if (p >= end) return 0;
    int8_t b = (int8_t)p[0];
    if (b >= -112) { *out = b; return 1; }
    bool neg = b < -120;
    int n = neg ? -120 - b : -112 - b;
    if (p + 1 + n > end) return 0;
    uint64_t t = 0;
    for (int i = 0; i < n; ++i) t = (t << 8) | p[1 + i];
    *out = neg ? (int64_t)~t : (int64_t)t;
    return n + 1;
}

// A logical byte stream made of two consecutive buffers (the tail of one fetched chunk followed by the
// head of the next): what the reference's "read across two buffers" DataStream overload provides for
// records that straddle a chunk boundary (src/Merger/StreamRW.cc:592-662).
struct SplitCursor {
    const uint8_t* a; size_t na;
    const uint8_t* b; size_t nb;
    size_t pos = 0;
    size_t size() const { return na + nb; }
    size_t remaining() const { return size() - pos; }
    uint8_t at(size_t i) const { return i < na ? a[i] : b[i - na]; }
    bool read(void* dst, size_t n) {
        if (remaining() < n) return false;
        uint8_t* d = (uint8_t*)dst;
        size_t p = pos;
        if (p < na) {
            size_t k = na - p < n ? na - p : n;
            memcpy(d, a + p, k);
            d += k; n -= k; p += k;
        }
        if (n) memcpy(d, b + (p - na), n);
        pos += (size_t)(d - (uint8_t*)dst) + n;
        return true;
    }
    // 0 = truncated
    int read_vint(int64_t* out) {
        uint8_t tmp[9];
        if (remaining() < 1) return 0;
        tmp[0] = at(pos);
        int n = vint_size_from_first((int8_t)tmp[0]);
        if (remaining() < (size_t)n) return 0;
        for (int i = 1; i < n; ++i) tmp[i] = at(pos + i);
        int r = vint_decode(tmp, tmp + n, out);
        if (r) pos += r;
        return r;
    }
};

}  // namespace uda
