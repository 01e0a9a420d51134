#include "device_merge.h"

#include <dlfcn.h>
#include <string.h>

#include <algorithm>
#include <chrono>

#include "uda_rt.h"
#include "vint.h"

namespace uda {

void index_segment(HostSegment* seg) {
    const uint8_t* p = seg->bytes.data();
    const uint8_t* end = p + seg->bytes.size();
    const uint8_t* base = p;
    seg->rec_off.clear();
    for (;;) {
        int64_t kl = 0, vl = 0;
        const int n1 = vint_decode(p, end, &kl);
        const int n2 = n1 ? vint_decode(p + n1, end, &vl) : 0;
        if (!n1 || !n2) throw UdaError("segment ended without EOF marker");
        seg->rec_off.push_back((long long)(p - base));
        if (kl == -1 && vl == -1) return;   // the entry just pushed is the EOF marker position
        // lengths first (Hadoop's are ints), then the bound as a length comparison: a pointer sum could wrap
        if (kl < 0 || vl < 0 || kl > 0x7FFFFFFF || vl > 0x7FFFFFFF || (int64_t)(end - p) - n1 - n2 < kl + vl)
            throw UdaError("corrupt record header in a fetched segment");
        p += n1 + n2 + kl + vl;
    }
}

// ---- the C ABI of libuda_kernels.so, resolved at run time (include/uda_b200.h) ----------------------------------
struct DeviceMerge::Api {
    int (*device_count)(int*);
    int (*mem_info)(long long*, long long*);
    int (*dev_malloc)(void**, long long);
    int (*dev_free)(void*);
    int (*h2d)(void*, const void*, long long, void*);
    int (*d2h)(void*, const void*, long long, void*);
    int (*sync)(void*);
    int (*memset_async)(void*, int, long long, void*);
    int (*extract_keys_var)(const void*, const long long*, long long, int, unsigned, unsigned, void*, unsigned*, void*);
    long long (*merge_scratch_bytes)(long long, int);
    int (*kway_merge_entries)(const void*, const long long*, int, long long, void*, void*, void*);
    long long (*tiebreak_scratch_bytes)(long long);
    int (*tiebreak_var)(void*, long long, const unsigned long long*, const unsigned long long*, int, void*, void*, void*);
    int (*gather_var_lengths)(const void*, long long, const unsigned long long*, unsigned*, void*);
    long long (*scan64_scratch_bytes)(long long);
    int (*scan64)(const unsigned*, long long, unsigned long long*, void*, void*);
    int (*gather_var)(const void*, long long, const unsigned long long*, const unsigned long long*, const unsigned long long*, void*, void*);
    int (*write_eof_at)(void*, const unsigned long long*, void*);
    const char* (*error_string)(int);
    int (*host_alloc_pinned)(void**, long long);
    // fused fixed-stride record merge (one read + one write per record)
    int (*merge_records_supported)(int, long long, int, int);
    long long (*merge_records_scratch_bytes)(long long, int);
    int (*merge_records_fixed)(const void*, int, long long, long long, int, int, int, int, void*, int, void*, int, void*);
};

namespace {
// the kernel library's run descriptor (include/uda_b200.h: UdaRunDesc)
struct RunDesc {
    const void* base;
    long long nrec;
    long long out_off;
    unsigned run_id;
    unsigned idx0;
};
static_assert(sizeof(RunDesc) == 32, "UdaRunDesc layout");

// Do all records of all segments have the same serialized shape (same VInt headers, hence the same key and value lengths),
// with a compared key part of at most 10 bytes? Then the fused record merge orders them exactly on its own (no entry
// pipeline, no tie-break) and every record is `stride` bytes.
struct FixedShape { int stride, key_off, key_len; };
bool fixed_shape(const std::vector<HostSegment>& segs, KeyKind kind, FixedShape* out) {
    const HostSegment* first = nullptr;
    for (const auto& sg : segs) if (sg.records() > 0) { first = &sg; break; }
    if (!first) return false;
    const uint8_t* p = first->bytes.data();
    const uint8_t* end = p + first->bytes.size();
    int64_t kl = 0, vl = 0;
    const int n1 = vint_decode(p, end, &kl);
    const int n2 = n1 ? vint_decode(p + n1, end, &vl) : 0;
    if (!n1 || !n2 || kl < 1 || vl < 0) return false;
    const int hdr = n1 + n2;
    const long long stride = hdr + kl + vl;
    int skip = 0;
    if (kind == KeyKind::TEXT) skip = vint_size_from_first((int8_t)p[hdr]);
    else if (kind == KeyKind::BYTES_WRITABLE) skip = 4;
    const long long key_len = kl - skip;
    if (stride < 16 || stride > 128 || key_len < 1 || key_len > 10 || hdr > 8) return false;
    uint8_t tmpl[9];
    memcpy(tmpl, p, (size_t)hdr);
    const uint8_t kfirst = p[hdr];    // Text: the key's own length prefix must be the same everywhere too
    for (const auto& sg : segs) {
        const long long n = sg.records();
        if (n == 0) continue;
        if (sg.rec_off[(size_t)n] != n * stride) return false;
        const uint8_t* b = sg.bytes.data();
        for (long long i = 0; i < n; ++i) {
            const uint8_t* r = b + i * stride;
            if (memcmp(r, tmpl, (size_t)hdr) != 0) return false;
            if (kind == KeyKind::TEXT && r[hdr] != kfirst) return false;
        }
    }
    out->stride = (int)stride;
    out->key_off = hdr + skip;
    out->key_len = (int)key_len;
    return true;
}
}  // namespace

namespace {
std::string dir_of_this_library() {
    Dl_info info;
    if (dladdr((void*)&index_segment, &info) && info.dli_fname) {
        std::string p = info.dli_fname;
        size_t s = p.rfind('/');
        if (s != std::string::npos) return p.substr(0, s);
    }
    return ".";
}
}  // namespace

static std::atomic<long long> g_fused_merges{0};
long long DeviceMerge::fused_merges() { return g_fused_merges.load(); }

DeviceMerge& DeviceMerge::instance() {
    static DeviceMerge d;
    return d;
}

DeviceMerge::DeviceMerge() {
    const char* env = getenv("UDA_KERNELS_LIB");
    const std::string path = env && *env ? std::string(env) : dir_of_this_library() + "/libuda_kernels.so";
    lib_ = dlopen(path.c_str(), RTLD_NOW | RTLD_LOCAL);
    if (!lib_) {
        const char* err = dlerror();   // a second call would return null: the message is consumed by the first
        why_ = std::string("cannot load ") + path + ": " + (err ? err : "unknown error");
        return;
    }
    api_ = new Api();
    bool all = true;
    auto sym = [&](const char* n) { void* s = dlsym(lib_, n); if (!s) { all = false; why_ = std::string("missing symbol ") + n; } return s; };
#define UDA_SYM(field, name) api_->field = reinterpret_cast<decltype(api_->field)>(sym(name))
    UDA_SYM(device_count, "uda_device_count");
    UDA_SYM(mem_info, "uda_mem_info");
    UDA_SYM(dev_malloc, "uda_dev_malloc");
    UDA_SYM(dev_free, "uda_dev_free");
    UDA_SYM(h2d, "uda_memcpy_h2d_async");
    UDA_SYM(d2h, "uda_memcpy_d2h_async");
    UDA_SYM(sync, "uda_stream_synchronize");
    UDA_SYM(memset_async, "uda_memset_async");
    UDA_SYM(extract_keys_var, "uda_extract_keys_var");
    UDA_SYM(merge_scratch_bytes, "uda_merge_scratch_bytes");
    UDA_SYM(kway_merge_entries, "uda_kway_merge_entries");
    UDA_SYM(tiebreak_scratch_bytes, "uda_tiebreak_scratch_bytes");
    UDA_SYM(tiebreak_var, "uda_tiebreak_var");
    UDA_SYM(gather_var_lengths, "uda_gather_var_lengths");
    UDA_SYM(scan64_scratch_bytes, "uda_scan_u32_to_u64_scratch_bytes");
    UDA_SYM(scan64, "uda_scan_u32_to_u64");
    UDA_SYM(gather_var, "uda_gather_var");
    UDA_SYM(write_eof_at, "uda_write_eof_at");
    UDA_SYM(error_string, "uda_error_string");
    UDA_SYM(host_alloc_pinned, "uda_host_alloc_pinned");
    UDA_SYM(merge_records_supported, "uda_merge_records_supported");
    UDA_SYM(merge_records_scratch_bytes, "uda_merge_records_scratch_bytes");
    UDA_SYM(merge_records_fixed, "uda_merge_records_fixed");
#undef UDA_SYM
    if (!all) return;
    int n = 0;
    if (api_->device_count(&n) != 0 || n < 1) { why_ = "no CUDA device visible"; return; }
    ok_ = true;
    UDA_LOG(lsINFO, "device merge back-end ready: %s, %d device(s)", path.c_str(), n);
}

void DeviceMerge::warm_up() {
    if (!ok_) return;
    long long f = 0, t = 0;
    api_->mem_info(&f, &t);   // the first runtime call that needs a context creates it (about a second on a cold process)
}

namespace {
// every device allocation of one merge, released together
struct DeviceArena {
    DeviceMerge* owner;
    int (*free_fn)(void*);
    std::vector<void*> ptrs;
    ~DeviceArena() { for (void* p : ptrs) if (p) free_fn(p); }
};
inline long long align_up(long long x, long long a) { return (x + a - 1) / a * a; }
}  // namespace

// The merged stream `d_out` (device) with record offsets `off[0..n]` (off[n] = position of the EOF marker) -> `deliver`.
void DeviceMerge::stream_out(const uint8_t* d_out, const std::vector<unsigned long long>& off, size_t n, size_t max_block,
                             const std::function<void(const uint8_t*, size_t)>& deliver) {
    // The merged stream comes back window by window through two pinned staging buffers: the copy of window w+1 is
    // in flight while window w is delivered in blocks of whole records; the EOF marker rides with the last block
    // (or alone when it does not fit). A record larger than a window travels through a one-off pageable buffer.
    const Api& a = *api_;
    void* const stream = nullptr;
    auto ck = [&](int rc, const char* what) { if (rc != 0) throw UdaError(std::string(what) + ": " + a.error_string(rc)); };
    std::lock_guard<std::mutex> stage_lock(stage_mu_);
    if (!stage_[0]) {
        for (int b = 0; b < 2; ++b) {
            void* p = nullptr;
            if (a.host_alloc_pinned(&p, (long long)kStageBytes) != 0 || !p) throw UdaError("pinned staging buffer allocation failed");
            stage_[b] = (uint8_t*)p;
        }
    }
    struct Window { size_t r0, r1; };            // records [r0, r1)
    auto next_window = [&](size_t r0) {
        size_t r1 = r0 + 1;
        // largest r1 with off[r1] - off[r0] (+2 for the EOF marker at the very end) <= kStageBytes, found by bisection
        size_t lo = r0 + 1, hi = n;
        while (lo < hi) {
            size_t mid = lo + (hi - lo + 1) / 2;
            if (off[mid] - off[r0] + (mid == n ? 2 : 0) <= kStageBytes) lo = mid; else hi = mid - 1;
        }
        r1 = std::max(r1, lo);
        return Window{r0, r1};
    };
    std::vector<uint8_t> big[2];                 // only for a record that does not fit a staging buffer (one per slot)
    auto start_copy = [&](const Window& w, int slot) -> uint8_t* {
        const size_t bytes = (size_t)(off[w.r1] - off[w.r0]) + (w.r1 == n ? 2 : 0);
        uint8_t* dst = stage_[slot];
        if (bytes > kStageBytes) { big[slot].resize(bytes); dst = big[slot].data(); }
        ck(a.d2h(dst, d_out + off[w.r0], (long long)bytes, stream), "copy of the merged stream to the host");
        return dst;
    };
    Window cur = next_window(0);
    uint8_t* cur_buf = start_copy(cur, 0);
    int slot = 0;
    while (true) {
        ck(a.sync(stream), "stream synchronize");
        const bool final_window = cur.r1 == n;
        Window nxt{0, 0};
        uint8_t* nxt_buf = nullptr;
        if (!final_window) {
            nxt = next_window(cur.r1);
            nxt_buf = start_copy(nxt, slot ^ 1);
        }
        const unsigned long long base = off[cur.r0];
        size_t i = cur.r0;
        while (i < cur.r1) {
            size_t j = i + 1;
            while (j < cur.r1 && off[j + 1] - off[i] <= max_block) ++j;
            size_t len = (size_t)(off[j] - off[i]);
            const bool last = final_window && j == cur.r1;
            if (last && len + 2 <= max_block) len += 2;
            deliver(cur_buf + (off[i] - base), len);
            if (last && len == (size_t)(off[j] - off[i])) deliver(cur_buf + (off[n] - base), 2);
            i = j;
        }
        if (final_window) break;
        cur = nxt;
        cur_buf = nxt_buf;
        slot ^= 1;
    }
}

// Fixed-shape records: H2D of the segments, the fused record merge (uda_merge_records_fixed: every record is read once and
// written once on the device), D2H window by window. Returns false with *why set when nothing has been delivered yet.
bool DeviceMerge::merge_fixed(const std::vector<HostSegment>& segs, int stride, int key_off, int key_len, long long n, size_t max_block,
                              const std::function<void(const uint8_t*, size_t)>& sink, std::string* why, bool* out_of_memory) {
    *out_of_memory = false;
    const Api& a = *api_;
    const int k = (int)segs.size();
    bool delivered = false;
    const std::function<void(const uint8_t*, size_t)> deliver = [&](const uint8_t* p, size_t len) { delivered = true; sink(p, len); };
    try {
        using Clk = std::chrono::steady_clock;
        const auto t0 = Clk::now();
        auto ms_since = [](Clk::time_point t) { return (long long)std::chrono::duration_cast<std::chrono::milliseconds>(Clk::now() - t).count(); };
        std::vector<long long> seg_pos(k, 0);
        long long data_bytes = 64, max_nrec = 0;
        for (int r = 0; r < k; ++r) {
            seg_pos[r] = data_bytes;
            data_bytes = align_up(data_bytes + segs[r].records() * stride + 64, 16);
            max_nrec = std::max(max_nrec, segs[r].records());
        }
        const long long out_bytes = n * stride + 2;
        const long long scratch_bytes = a.merge_records_scratch_bytes(n, k);
        if (scratch_bytes < 0) { *why = "scratch size query failed"; return false; }
        const long long need = data_bytes + out_bytes + scratch_bytes + (long long)k * 32 + 4096;
        long long free_b = 0, total_b = 0;
        if (a.mem_info(&free_b, &total_b) != 0 || need > (long long)(0.9 * (double)free_b)) {
            *why = "partition needs " + std::to_string(need >> 20) + " MiB of device memory, " + std::to_string(free_b >> 20) + " MiB free";
            *out_of_memory = true;
            return false;
        }
        DeviceArena arena{this, a.dev_free, {}};
        auto dmalloc = [&](long long bytes) -> uint8_t* {
            void* p = nullptr;
            int rc = a.dev_malloc(&p, std::max<long long>(bytes, 16));
            if (rc != 0 || !p) throw UdaError(std::string("device allocation failed: ") + a.error_string(rc));
            arena.ptrs.push_back(p);
            return (uint8_t*)p;
        };
        auto ck = [&](int rc, const char* what) { if (rc != 0) throw UdaError(std::string(what) + ": " + a.error_string(rc)); };
        void* const stream = nullptr;
        uint8_t* d_data = dmalloc(data_bytes + 64);
        uint8_t* d_out = dmalloc(out_bytes + 64);
        uint8_t* d_scr = dmalloc(scratch_bytes);
        uint8_t* d_runs = dmalloc((long long)k * 32);
        std::vector<RunDesc> runs((size_t)k);
        long long at = 0;
        const bool even = (stride % 2) == 0 && (key_off % 2) == 0;
        for (int r = 0; r < k; ++r) {
            const long long nr = segs[r].records();
            if (nr) ck(a.h2d(d_data + seg_pos[r], segs[r].bytes.data(), nr * stride, stream), "copy of a segment to the device");
            runs[(size_t)r] = RunDesc{d_data + seg_pos[r], nr, at, (unsigned)r, 0u};
            at += nr;
        }
        ck(a.h2d(d_runs, runs.data(), (long long)k * 32, stream), "run table upload");
        const long long ms_h2d = ms_since(t0);
        const auto t1 = Clk::now();
        ck(a.merge_records_fixed(d_runs, k, n, max_nrec, stride, key_off, key_len, even ? 1 : 0, d_out, 1, d_scr, 0, stream), "uda_merge_records_fixed");
        ck(a.sync(stream), "stream synchronize");
        const long long ms_merge = ms_since(t1);
        const auto t2 = Clk::now();
        std::vector<unsigned long long> off((size_t)n + 1);
        for (long long i = 0; i <= n; ++i) off[(size_t)i] = (unsigned long long)i * (unsigned long long)stride;
        stream_out(d_out, off, (size_t)n, max_block, deliver);
        ++g_fused_merges;
        UDA_LOG(lsINFO, "device merge (fused fixed-stride path): H2D %lld ms, merge %lld ms, D2H + delivery %lld ms (%lld records of %d bytes, %d runs)",
                ms_h2d, ms_merge, ms_since(t2), n, stride, k);
    } catch (const UdaError& e) {
        if (delivered) throw;
        *why = e.info();
        return false;
    }
    return true;
}

bool DeviceMerge::merge(const std::vector<HostSegment>& segs, KeyKind kind, size_t max_block,
                        const std::function<void(const uint8_t*, size_t)>& sink, long long* records_out, std::string* why,
                        Refusal* refusal) {
    Refusal dummy;
    if (!refusal) refusal = &dummy;
    *refusal = Refusal::NONE;
    if (!ok_) { if (why) *why = why_; *refusal = Refusal::UNAVAILABLE; return false; }
    const Api& a = *api_;
    const int k = (int)segs.size();
    long long n = 0, data_bytes = 64, out_bytes = 2;
    std::vector<long long> run_off(k + 1, 0), seg_pos(k, 0), ro_pos(k, 0);
    long long ro_bytes = 0;
    for (int r = 0; r < k; ++r) {
        run_off[r] = n;
        n += segs[r].records();
        seg_pos[r] = data_bytes;
        data_bytes = align_up(data_bytes + segs[r].raw_len() + 32, 16);
        ro_pos[r] = ro_bytes;
        ro_bytes += (long long)segs[r].rec_off.size() * 8;
        out_bytes += segs[r].raw_len() - 2;
    }
    run_off[k] = n;
    if (records_out) *records_out = n;
    if (k == 0 || n == 0) {
        const uint8_t eof[2] = {0xFF, 0xFF};
        sink(eof, 2);
        return true;
    }
    // limits of the kernel library: uda_kway_merge_entries takes at most 4096 runs, the entry format 2^31 - 2 records
    if (k > 4096 || n >= (1LL << 31) - 1) {
        if (why) *why = "partition too large for the device merge (runs > 4096 or records >= 2^31 - 1)";
        *refusal = Refusal::LIMITS;
        return false;
    }
    // ---- fast path: records of one fixed shape go through the fused record merge (the kernel of the TeraSort headline)
    FixedShape fs{0, 0, 0};
    if (fixed_shape(segs, kind, &fs) && a.merge_records_supported(k, n, fs.stride, fs.key_len)) {
        std::string why_fixed;
        bool oom = false;
        if (merge_fixed(segs, fs.stride, fs.key_off, fs.key_len, n, max_block, sink, &why_fixed, &oom)) return true;
        if (oom) {   // the general path needs more memory than the fused one: no point in trying it
            if (why) *why = why_fixed;
            *refusal = Refusal::MEMORY;
            return false;
        }
        UDA_LOG(lsWARN, "fused record merge not used (%s): general device path", why_fixed.c_str());
    }
    const long long nn = std::max<long long>(n, 1);
    const long long merge_scratch = a.merge_scratch_bytes(nn, k), tb_scratch = a.tiebreak_scratch_bytes(nn), scan_scratch = a.scan64_scratch_bytes(nn);
    if (merge_scratch < 0 || tb_scratch < 0 || scan_scratch < 0) { if (why) *why = "scratch size query failed"; *refusal = Refusal::FAILED; return false; }
    const long long need = data_bytes + ro_bytes + 2 * nn * 16 + merge_scratch + tb_scratch + scan_scratch + nn * 4 + (nn + 1) * 8 + out_bytes + 64 +
                           (long long)k * 24 + 4096;
    long long free_b = 0, total_b = 0;
    if (a.mem_info(&free_b, &total_b) != 0 || need > (long long)(0.9 * (double)free_b)) {
        if (why) *why = "partition needs " + std::to_string(need >> 20) + " MiB of device memory, " + std::to_string(free_b >> 20) + " MiB free";
        *refusal = Refusal::MEMORY;
        return false;
    }
    // Anything that goes wrong before the first block has been handed to the sink (allocation, a copy, a kernel launch, a
    // limit of the kernel library) leaves the caller free to merge on the CPU: report it through `why` and return false.
    // Once output has been delivered the stream cannot be restarted, so the error propagates.
    bool delivered = false;
    const std::function<void(const uint8_t*, size_t)> deliver = [&](const uint8_t* p, size_t len) { delivered = true; sink(p, len); };
    try {
    using Clk = std::chrono::steady_clock;
    auto ms_since = [](Clk::time_point t) { return (long long)std::chrono::duration_cast<std::chrono::milliseconds>(Clk::now() - t).count(); };
    const auto t0 = Clk::now();
    DeviceArena arena{this, a.dev_free, {}};
    auto dmalloc = [&](long long bytes) -> uint8_t* {
        void* p = nullptr;
        int rc = a.dev_malloc(&p, std::max<long long>(bytes, 16));
        if (rc != 0 || !p) throw UdaError(std::string("device allocation failed: ") + a.error_string(rc));
        arena.ptrs.push_back(p);
        return (uint8_t*)p;
    };
    auto ck = [&](int rc, const char* what) { if (rc != 0) throw UdaError(std::string(what) + ": " + a.error_string(rc)); };
    void* const stream = nullptr;   // the legacy default stream: this back-end runs one merge at a time

    uint8_t* d_data = dmalloc(data_bytes + 64);
    uint8_t* d_ro = dmalloc(ro_bytes + 16);
    uint8_t* d_ent = dmalloc(nn * 16);
    uint8_t* d_sorted = dmalloc(nn * 16);
    uint8_t* d_mscr = dmalloc(merge_scratch);
    uint8_t* d_tscr = dmalloc(tb_scratch);
    uint8_t* d_sscr = dmalloc(scan_scratch);
    uint8_t* d_len = dmalloc(nn * 4);
    uint8_t* d_off = dmalloc((nn + 1) * 8);
    uint8_t* d_out = dmalloc(out_bytes + 64);
    uint8_t* d_tab = dmalloc((long long)k * 16 + (long long)(k + 1) * 8 + 16);   // run_base | run_recoff | run_off | flag
    unsigned long long* d_run_base = (unsigned long long*)d_tab;
    unsigned long long* d_run_recoff = d_run_base + k;
    long long* d_run_off = (long long*)(d_run_recoff + k);
    unsigned* d_flag = (unsigned*)(d_run_off + k + 1);

    const long long ms_alloc = ms_since(t0);
    const auto t1 = Clk::now();
    std::vector<unsigned long long> h_base(k), h_recoff(k);
    for (int r = 0; r < k; ++r) {
        ck(a.h2d(d_data + seg_pos[r], segs[r].bytes.data(), segs[r].raw_len(), stream), "copy of a segment to the device");
        ck(a.h2d(d_ro + ro_pos[r], segs[r].rec_off.data(), (long long)segs[r].rec_off.size() * 8, stream), "copy of a record-offset table");
        h_base[r] = (unsigned long long)(uintptr_t)(d_data + seg_pos[r]);
        h_recoff[r] = (unsigned long long)(uintptr_t)(d_ro + ro_pos[r]);
    }
    ck(a.h2d(d_run_base, h_base.data(), (long long)k * 8, stream), "table upload");
    ck(a.h2d(d_run_recoff, h_recoff.data(), (long long)k * 8, stream), "table upload");
    ck(a.h2d(d_run_off, run_off.data(), (long long)(k + 1) * 8, stream), "table upload");
    ck(a.memset_async(d_flag, 0, 4, stream), "memset");
    const int kk = (int)kind;
    for (int r = 0; r < k; ++r)
        if (segs[r].records())
            ck(a.extract_keys_var(d_data + seg_pos[r], (const long long*)(d_ro + ro_pos[r]), segs[r].records(), kk, (unsigned)r, 0u,
                                  d_ent + run_off[r] * 16, d_flag, stream), "uda_extract_keys_var");
    ck(a.kway_merge_entries(d_ent, d_run_off, k, n, d_sorted, d_mscr, stream), "uda_kway_merge_entries");
    ck(a.tiebreak_var(d_sorted, n, d_run_base, d_run_recoff, kk, d_ent, d_tscr, stream), "uda_tiebreak_var");
    ck(a.gather_var_lengths(d_sorted, n, d_run_recoff, (unsigned*)d_len, stream), "uda_gather_var_lengths");
    ck(a.scan64((const unsigned*)d_len, n, (unsigned long long*)d_off, d_sscr, stream), "uda_scan_u32_to_u64");
    ck(a.gather_var(d_sorted, n, d_run_base, d_run_recoff, (const unsigned long long*)d_off, d_out, stream), "uda_gather_var");
    ck(a.write_eof_at(d_out, (const unsigned long long*)d_off + n, stream), "uda_write_eof_at");

    ck(a.sync(stream), "stream synchronize");
    const long long ms_h2d_kernels = ms_since(t1);
    const auto t2 = Clk::now();
    std::vector<unsigned long long> off((size_t)n + 1);
    ck(a.d2h(off.data(), d_off, (n + 1) * 8, stream), "copy of the stream offsets to the host");
    ck(a.sync(stream), "stream synchronize");
    if ((long long)off[n] + 2 != out_bytes) throw UdaError("device merge produced a stream of unexpected length");
    const long long ms_d2h = ms_since(t2);
    const auto t3 = Clk::now();

    stream_out(d_out, off, (size_t)n, max_block, deliver);
    UDA_LOG(lsINFO, "device merge phases: alloc %lld ms, H2D + kernels %lld ms, D2H %lld ms, delivery %lld ms (%lld records, %lld bytes, %d runs)",
            ms_alloc, ms_h2d_kernels, ms_d2h, ms_since(t3), n, out_bytes, k);
    } catch (const UdaError& e) {
        if (delivered) throw;
        if (why) *why = e.info();
        *refusal = Refusal::FAILED;
        return false;
    }
    return true;
}

}  // namespace uda
