// A host-only stand-in for libuda_kernels.so, for testing csrc/runtime/device_merge.cc on machines without a GPU.
//
// device_merge.cc reaches the kernel library only through its C ABI, resolved with dlopen (UDA_KERNELS_LIB selects the
// file). This library exports the 23 entry points it resolves and honours their contracts (include/uda_b200.h) with plain
// host code: "device memory" is malloc, the copies are memcpy, the "stream" is the calling thread (everything completes
// before the call returns), the kernels are loops and std::sort. What a test learns from it: the host-side logic of the
// device merge (layout of the uploads, offsets, fixed-shape detection, windowed delivery in blocks of whole records, the
// fall-back rules before / after the first delivered block) -- under ASAN / TSAN too. What it cannot show: stream-order
// mistakes (a missing synchronize is invisible when every call is synchronous) and the kernels themselves, which
// tests/test_device_merge_gpu.py covers on a B200.
//
// Knobs (environment, read at every call so one process can change them between jobs):
//   UDA_MOCK_DEVICES=N            devices reported (default 1)
//   UDA_MOCK_FREE_BYTES=N         free device memory reported (default 8 GiB)
//   UDA_MOCK_NO_FUSED=1           uda_merge_records_supported answers no (forces the general path)
//   UDA_MOCK_FAIL_MALLOC_AT=K     the K-th device allocation of the process fails (1-based)
//   UDA_MOCK_FAIL_D2H_AT=K        the K-th device-to-host copy of the process fails (1-based)
//   UDA_MOCK_FAIL_FN=name         the named kernel entry point (e.g. uda_tiebreak_var) returns an error instead of running
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <atomic>
#include <vector>

#include "../../../include/uda_b200.h"

namespace {

long long env_ll(const char* name, long long dflt) {
    const char* v = getenv(name);
    return v && *v ? atoll(v) : dflt;
}
std::atomic<long long> g_mallocs{0}, g_d2h{0}, g_live{0};
bool fails(const char* fn) {
    const char* v = getenv("UDA_MOCK_FAIL_FN");
    return v && strcmp(v, fn) == 0;
}
#define MOCK_FAIL_POINT() do { if (fails(__func__)) return 700; } while (0)

// Hadoop VInt (csrc/runtime/vint.h restated: the mock must not link the library under test)
int vint_decode(const uint8_t* p, int64_t* out) {
    const int8_t b = (int8_t)p[0];
    if (b >= -112) { *out = b; return 1; }
    const bool neg = b < -120;
    const int n = neg ? -120 - b : -112 - b;
    uint64_t t = 0;
    for (int i = 0; i < n; ++i) t = (t << 8) | p[1 + i];
    *out = neg ? (int64_t)~t : (int64_t)t;
    return n + 1;
}
int vint_size_from_first(int8_t b) { return b >= -112 ? 1 : (b < -120 ? -119 - b : -111 - b); }

struct Key { const uint8_t* p; int len; };
// compared part of a serialized key: raw bytes, Text (own VInt length prefix skipped), BytesWritable (4-byte length skipped)
Key key_of_record(const uint8_t* rec, int kind) {
    int64_t kl = 0, vl = 0;
    const int n1 = vint_decode(rec, &kl);
    const int n2 = vint_decode(rec + n1, &vl);
    const uint8_t* k = rec + n1 + n2;
    int skip = 0;
    if (kind == 1) skip = kl > 0 ? vint_size_from_first((int8_t)k[0]) : 0;
    else if (kind == 2) skip = 4;
    if (skip > kl) skip = (int)kl;
    return Key{k + skip, (int)kl - skip};
}
int compare(const Key& a, const Key& b) {
    const int c = memcmp(a.p, b.p, (size_t)std::min(a.len, b.len));
    return c ? c : a.len - b.len;
}
inline unsigned run_of(const UdaEntry& e) { return (unsigned)(e.w1 >> 32) & 0xFFFFu; }
inline unsigned idx_of(const UdaEntry& e) { return (unsigned)e.w1; }
inline bool entry_less(const UdaEntry& a, const UdaEntry& b) { return a.w0 < b.w0 || (a.w0 == b.w0 && a.w1 < b.w1); }

}  // namespace

extern "C" {

int uda_device_count(int* out) { *out = (int)env_ll("UDA_MOCK_DEVICES", 1); return 0; }
int uda_set_device(int) { return 0; }
int uda_mem_info(long long* free_bytes, long long* total_bytes) {
    *free_bytes = env_ll("UDA_MOCK_FREE_BYTES", 8LL << 30);
    *total_bytes = 180LL << 30;
    return 0;
}
int uda_dev_malloc(void** ptr, long long bytes) {
    const long long k = ++g_mallocs;
    if (k == env_ll("UDA_MOCK_FAIL_MALLOC_AT", -1)) { *ptr = nullptr; return 2; /* cudaErrorMemoryAllocation */ }
    *ptr = malloc((size_t)std::max<long long>(bytes, 1));
    if (!*ptr) return 2;
    ++g_live;
    return 0;
}
int uda_dev_free(void* p) { if (p) { free(p); --g_live; } return 0; }
long long uda_mock_live_allocations(void) { return g_live.load(); }
long long uda_mock_allocation_count(void) { return g_mallocs.load(); }
long long uda_mock_d2h_count(void) { return g_d2h.load(); }
int uda_host_alloc_pinned(void** ptr, long long bytes) { *ptr = malloc((size_t)std::max<long long>(bytes, 1)); return *ptr ? 0 : 2; }
int uda_host_free_pinned(void* p) { free(p); return 0; }
int uda_memcpy_h2d_async(void* dst, const void* src, long long bytes, cudaStream_t) { memcpy(dst, src, (size_t)bytes); return 0; }
int uda_memcpy_d2h_async(void* dst, const void* src, long long bytes, cudaStream_t) {
    if (++g_d2h == env_ll("UDA_MOCK_FAIL_D2H_AT", -1)) return 700; /* cudaErrorIllegalAddress: what a faulted copy reports */
    memcpy(dst, src, (size_t)bytes);
    return 0;
}
int uda_stream_synchronize(cudaStream_t) { return 0; }
int uda_memset_async(void* p, int v, long long bytes, cudaStream_t) { memset(p, v, (size_t)bytes); return 0; }
const char* uda_error_string(int rc) {
    switch (rc) {
        case 0: return "no error";
        case 2: return "out of memory (mock)";
        case 700: return "an illegal memory access was encountered (mock)";
        case UDA_E_INVALID: return "invalid argument";
        case UDA_E_TOO_LARGE: return "size exceeds a documented maximum";
        default: return "mock error";
    }
}

// ---- general path -------------------------------------------------------------------------------------------------
int uda_extract_keys_var(const void* data, const long long* rec_off, long long n, int key_kind, unsigned run_id, unsigned idx0, void* entries_out,
                         unsigned* long_key_flag, cudaStream_t) {
    MOCK_FAIL_POINT();
    if (key_kind < 0 || key_kind > 2 || run_id > 0xFFFFu) return UDA_E_INVALID;
    UdaEntry* e = (UdaEntry*)entries_out;
    for (long long i = 0; i < n; ++i) {
        const Key k = key_of_record((const uint8_t*)data + rec_off[i], key_kind);
        uint8_t pre[10] = {0};
        memcpy(pre, k.p, (size_t)std::min(k.len, 10));
        unsigned long long w0 = 0;
        for (int b = 0; b < 8; ++b) w0 = (w0 << 8) | pre[b];
        e[i].w0 = w0;
        e[i].w1 = ((unsigned long long)pre[8] << 56) | ((unsigned long long)pre[9] << 48) | ((unsigned long long)run_id << 32) | (unsigned)(idx0 + i);
        if (k.len > 10 && long_key_flag) *long_key_flag = 1;
    }
    return 0;
}
long long uda_merge_scratch_bytes(long long n, int nruns) { return n < 0 || nruns < 0 ? UDA_E_INVALID : 256; }
int uda_kway_merge_entries(const void* entries, const long long* run_off, int nruns, long long n, void* sorted_out, void*, cudaStream_t) {
    MOCK_FAIL_POINT();
    if (nruns > 4096) return UDA_E_TOO_LARGE;
    const UdaEntry* e = (const UdaEntry*)entries;
    for (int r = 0; r < nruns; ++r)   // the contract: every run is sorted already
        for (long long i = run_off[r] + 1; i < run_off[r + 1]; ++i)
            if (entry_less(e[i], e[i - 1])) return UDA_E_INVALID;
    UdaEntry* o = (UdaEntry*)sorted_out;
    memcpy(o, e, (size_t)n * sizeof(UdaEntry));
    std::sort(o, o + n, entry_less);
    return 0;
}
long long uda_tiebreak_scratch_bytes(long long n) { return n < 0 ? UDA_E_INVALID : 256; }
int uda_tiebreak_var(void* sorted_entries, long long n, const unsigned long long* run_base, const unsigned long long* run_recoff, int key_kind,
                     void*, void*, cudaStream_t) {
    MOCK_FAIL_POINT();
    UdaEntry* e = (UdaEntry*)sorted_entries;
    auto rec = [&](const UdaEntry& x) {
        const unsigned r = run_of(x);
        const long long* ro = (const long long*)(uintptr_t)run_recoff[r];
        return (const uint8_t*)(uintptr_t)run_base[r] + ro[idx_of(x)];
    };
    long long i = 0;
    while (i < n) {
        long long j = i + 1;
        while (j < n && e[j].w0 == e[i].w0 && (e[j].w1 >> 48) == (e[i].w1 >> 48)) ++j;
        if (j - i > 1)   // equal 10-byte prefixes: the exact comparator decides, equal keys stay in (run, index) order
            std::stable_sort(e + i, e + j, [&](const UdaEntry& a, const UdaEntry& b) {
                const int c = compare(key_of_record(rec(a), key_kind), key_of_record(rec(b), key_kind));
                return c ? c < 0 : (a.w1 & 0xFFFFFFFFFFFFull) < (b.w1 & 0xFFFFFFFFFFFFull);
            });
        i = j;
    }
    return 0;
}
int uda_gather_var_lengths(const void* sorted_entries, long long n, const unsigned long long* run_recoff, unsigned* rec_len_out, cudaStream_t) {
    MOCK_FAIL_POINT();
    const UdaEntry* e = (const UdaEntry*)sorted_entries;
    for (long long i = 0; i < n; ++i) {
        const long long* ro = (const long long*)(uintptr_t)run_recoff[run_of(e[i])];
        rec_len_out[i] = (unsigned)(ro[idx_of(e[i]) + 1] - ro[idx_of(e[i])]);
    }
    return 0;
}
long long uda_scan_u32_to_u64_scratch_bytes(long long m) { return m < 0 ? UDA_E_INVALID : 256; }
int uda_scan_u32_to_u64(const unsigned* v, long long m, unsigned long long* out, void*, cudaStream_t) {
    MOCK_FAIL_POINT();
    unsigned long long run = 0;
    for (long long i = 0; i < m; ++i) { out[i] = run; run += v[i]; }
    out[m] = run;   // m + 1 outputs: the total closes the table
    return 0;
}
int uda_gather_var(const void* sorted_entries, long long n, const unsigned long long* run_base, const unsigned long long* run_recoff,
                   const unsigned long long* dst_off, void* dst, cudaStream_t) {
    MOCK_FAIL_POINT();
    const UdaEntry* e = (const UdaEntry*)sorted_entries;
    for (long long i = 0; i < n; ++i) {
        const unsigned r = run_of(e[i]);
        const long long* ro = (const long long*)(uintptr_t)run_recoff[r];
        const long long a = ro[idx_of(e[i])], b = ro[idx_of(e[i]) + 1];
        memcpy((uint8_t*)dst + dst_off[i], (const uint8_t*)(uintptr_t)run_base[r] + a, (size_t)(b - a));
    }
    return 0;
}
int uda_write_eof_at(void* dst, const unsigned long long* total, cudaStream_t) {
    MOCK_FAIL_POINT();
    uint8_t* p = (uint8_t*)dst + (total ? *total : 0);
    p[0] = 0xFF; p[1] = 0xFF;
    return 0;
}

// ---- fused fixed-stride path ----------------------------------------------------------------------------------------
int uda_merge_records_supported(int nruns, long long n, int stride, int key_len) {
    if (env_ll("UDA_MOCK_NO_FUSED", 0)) return 0;
    return nruns >= 1 && nruns <= 4096 && n >= 0 && n < (1LL << 31) && stride >= 16 && stride <= 128 && key_len >= 1 && key_len <= 10;
}
long long uda_merge_records_scratch_bytes(long long n, int nruns) { return n < 0 || nruns < 1 ? UDA_E_INVALID : 256; }
int uda_merge_records_fixed(const UdaRunDesc* runs, int nruns, long long n, long long max_nrec, int stride, int key_off, int key_len, int align2,
                            void* out, int write_eof, void*, int ctas_per_sm, cudaStream_t) {
    MOCK_FAIL_POINT();
    if (!uda_merge_records_supported(nruns, n, stride, key_len) && !env_ll("UDA_MOCK_NO_FUSED", 0)) return UDA_E_INVALID;
    if (align2 && ((stride | key_off) & 1)) return UDA_E_INVALID;
    if (ctas_per_sm < 0 || ctas_per_sm > 32) return UDA_E_INVALID;
    struct Ref { const uint8_t* p; unsigned run; long long idx; };
    std::vector<Ref> refs;
    refs.reserve((size_t)n);
    long long total = 0;
    for (int r = 0; r < nruns; ++r) {
        if (runs[r].nrec > max_nrec) return UDA_E_INVALID;
        if (align2 && ((uintptr_t)runs[r].base & 1)) return UDA_E_INVALID;
        for (long long i = 0; i < runs[r].nrec; ++i) refs.push_back(Ref{(const uint8_t*)runs[r].base + i * stride, runs[r].run_id, i});
        total += runs[r].nrec;
    }
    if (total != n) return UDA_E_INVALID;
    std::stable_sort(refs.begin(), refs.end(), [&](const Ref& a, const Ref& b) {
        const int c = memcmp(a.p + key_off, b.p + key_off, (size_t)key_len);
        return c ? c < 0 : (a.run != b.run ? a.run < b.run : a.idx < b.idx);
    });
    uint8_t* o = (uint8_t*)out;
    for (long long i = 0; i < n; ++i) memcpy(o + i * stride, refs[(size_t)i].p, (size_t)stride);
    if (write_eof) { o[n * stride] = 0xFF; o[n * stride + 1] = 0xFF; }
    return 0;
}

}  // extern "C"
