// Key comparators selected by the Java map-output key class name.
// Reference behaviour: src/Merger/CompareFunc.cc:29-42 (class tables), :70-91 (compare bodies),
// :95-113 (selection; unknown class throws so that the Java side falls back to the vanilla shuffle).
#pragma once
#include <cstring>
#include <string>

#include "uda_rt.h"
#include "vint.h"

namespace uda {

enum class KeyKind { RAW_BYTES = 0, TEXT = 1, BYTES_WRITABLE = 2 };

inline KeyKind key_kind_for_class(const std::string& java_class) {
    static const char* kText[] = {"org.apache.hadoop.io.Text"};
    static const char* kRaw[] = {"org.apache.hadoop.io.BooleanWritable", "org.apache.hadoop.io.ByteWritable",
                                 "org.apache.hadoop.io.ShortWritable", "org.apache.hadoop.io.IntWritable",
                                 "org.apache.hadoop.io.LongWritable"};
    static const char* kBytes[] = {"org.apache.hadoop.io.BytesWritable", "org.apache.hadoop.hbase.io.ImmutableBytesWritable"};
    for (auto c : kText) if (java_class == c) return KeyKind::TEXT;
    for (auto c : kRaw) if (java_class == c) return KeyKind::RAW_BYTES;
    for (auto c : kBytes) if (java_class == c) return KeyKind::BYTES_WRITABLE;
    UDA_LOG(lsERROR, "using compare function for unsupported type: '%s'", java_class.c_str());
    throw UdaError("using compare function for unsupported type");
}

// The compared part of a serialized key: skips the type-specific length header.
inline void key_body(KeyKind kind, const uint8_t* k, int len, const uint8_t** body, int* blen) {
    int skip = 0;
    if (kind == KeyKind::TEXT) skip = len > 0 ? vint_size_from_first((int8_t)k[0]) : 0;
    else if (kind == KeyKind::BYTES_WRITABLE) skip = 4;
    if (skip > len) skip = len;
    *body = k + skip;
    *blen = len - skip;
}

// memcmp then shorter-first, on the compared part
inline int compare_keys(KeyKind kind, const uint8_t* k1, int l1, const uint8_t* k2, int l2) {
    const uint8_t *b1, *b2;
    int n1, n2;
    key_body(kind, k1, l1, &b1, &n1);
    key_body(kind, k2, l2, &b2, &n2);
    int c = memcmp(b1, b2, (size_t)(n1 < n2 ? n1 : n2));
    return c ? c : n1 - n2;
}

}  // namespace uda
