// Plain C API of the host runtime (libuda.so). The JNI bridge (csrc/jni/uda_bridge.cc) and the Python
// "fake Hadoop" driver (uda_b200/bridge.py) are both thin layers over these functions, so the same
// role code runs under a JVM and under pytest.
//
// The reference exposes this surface only through JNI (src/UdaBridge.cc:187-330: startNative,
// doCommandNative, reduceExitMsgNative, setLogLevelNative) with one role per process held in globals;
// here a role is a handle, which lets one test process run a provider and several consumers.
#include <set>
#include <list>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

#include "threads.h"
#include "../../include/uda_rt_api.h"
#include "consumer.h"
#include "device_merge.h"
#include "provider.h"
#include "protocol.h"
#include "uda_rt.h"

using namespace uda;

namespace {

struct Role {
    bool is_net_merger = false;
    uda_host_callbacks_t c{};
    HostCallbacks host{};
    Options opt;
    std::unique_ptr<Provider> provider;
    std::unique_ptr<Consumer> consumer;
    std::thread provider_thread;
    FailureLatch latch;
    std::string last_error;
    bool failed = false;
};

void tramp_fetch_over(void* u) { auto* r = (Role*)u; if (r->c.fetch_over_message) r->c.fetch_over_message(r->c.user); }
void tramp_data(void* u, const void* b, int n) { auto* r = (Role*)u; if (r->c.data_from_uda) r->c.data_from_uda(r->c.user, b, n); }
bool tramp_path(void* u, const char* job, const char* map, int red, IndexRecord* out) {
    auto* r = (Role*)u;
    if (!r->c.get_path_uda) return false;
    char path[2048] = {0};
    long long off = 0, raw = 0, part = 0;
    if (!r->c.get_path_uda(r->c.user, job, map, red, &off, &raw, &part, path, (int)sizeof(path))) return false;
    out->offset = off; out->raw_length = raw; out->part_length = part; out->path = path;
    return true;
}
bool tramp_conf(void* u, const char* name, const char* dflt, std::string* out) {
    auto* r = (Role*)u;
    if (!r->c.get_conf_data) return false;
    char buf[1024] = {0};
    if (!r->c.get_conf_data(r->c.user, name, dflt ? dflt : "", buf, (int)sizeof(buf))) return false;
    *out = buf;
    return true;
}
void tramp_log(void* u, const char* m, int s) { auto* r = (Role*)u; if (r->c.log_to_java) r->c.log_to_java(r->c.user, m, s); }
void tramp_fail(void* u) { auto* r = (Role*)u; r->failed = true; if (r->c.failure_in_uda) r->c.failure_in_uda(r->c.user); }

int guarded(Role* r, const std::function<void()>& f) {
    try {
        f();
        return 0;
    } catch (const UdaError& e) {
        r->last_error = e.full_message();
        UDA_LOG(lsERROR, "%s", e.info().c_str());
    } catch (const std::exception& e) {
        r->last_error = std::string("unexpected: ") + e.what();
    } catch (...) {
        r->last_error = "unexpected failure";
    }
    r->failed = true;  // after a failure further commands are ignored (null down-call handler in the reference)
    return -1;
}

}  // namespace

extern "C" {

uda_role_t uda_rt_start(int is_net_merger, int argc, const char* const* argv, int log_level, int log_to_unique_file,
                        const uda_host_callbacks_t* cb) {
    auto* r = new Role();
    r->is_net_merger = is_net_merger != 0;
    if (cb) r->c = *cb;
    r->host.user = r;
    r->host.fetch_over_message = tramp_fetch_over;
    r->host.data_from_uda = tramp_data;
    r->host.get_path_uda = tramp_path;
    r->host.get_conf_data = tramp_conf;
    r->host.log_to_java = tramp_log;
    r->host.failure_in_uda = tramp_fail;
    log_set_threshold(log_level);
    log_set_to_unique_file(log_to_unique_file != 0);
    log_set_host(&r->host);
    std::vector<std::string> args;
    for (int i = 0; i < argc; ++i) args.emplace_back(argv[i] ? argv[i] : "");
    int rc = guarded(r, [&] {
        if (!parse_options(args, &r->opt)) throw UdaError("error in main: bad command line");
        if (r->opt.log_threshold >= 0) log_set_threshold(r->opt.log_threshold);
        if (r->is_net_merger) {
            r->consumer = std::make_unique<Consumer>(r->opt, &r->host);
            r->consumer->start();
        } else {
            r->provider = std::make_unique<Provider>(r->opt, &r->host);
            r->provider->start();
            Provider* p = r->provider.get();
            r->provider_thread = spawn_guarded("MOFSupplierRun", &r->host, &r->latch, [p] { p->run(); });
        }
        UDA_LOG(lsINFO, "main initialization finished ret=0");
    });
    (void)rc;
    return r;
}

int uda_rt_failed(uda_role_t h) { return h && ((Role*)h)->failed ? 1 : 0; }
const char* uda_rt_last_error(uda_role_t h) { return h ? ((Role*)h)->last_error.c_str() : ""; }

int uda_rt_do_command(uda_role_t h, const char* cmd) {
    auto* r = (Role*)h;
    if (!r) return -1;
    if (r->failed) return 0;  // ignored after a failure
    return guarded(r, [&] {
        if (r->is_net_merger) r->consumer->handle_command(cmd ? cmd : "");
        else r->provider->handle_command(cmd ? cmd : "");
    });
}

void uda_rt_reduce_exit(uda_role_t h) {
    auto* r = (Role*)h;
    if (!r || !r->consumer) return;
    try {
        r->consumer->exit_msg();
    } catch (const std::exception& e) {
        UDA_LOG(lsWARN, "failure in reduce_exit_msg_handler: %s", e.what());
    }
}

void uda_rt_set_log_level(int level) { log_set_threshold(level); }

int uda_rt_port(uda_role_t h) {
    auto* r = (Role*)h;
    return r && r->provider ? r->provider->port() : 0;
}

long long uda_rt_stat(uda_role_t h, const char* what) {
    auto* r = (Role*)h;
    if (!r) return -1;
    std::string w = what ? what : "";
    if (r->consumer) {
        if (w == "records_emitted") return (long long)r->consumer->records_emitted();
        if (w == "lpqs_spilled") return r->consumer->lpqs_spilled();
        if (w == "device_merges") return r->consumer->device_merges();
        if (w == "fetch_stall_us") return fetch_stall_us();
        if (w == "fetch_stalls") return fetch_stalls();
        if (w == "device_groups") return r->consumer->device_groups();   // groups of the last partition that was merged group-wise
        if (w == "device_fused_merges") return DeviceMerge::fused_merges();   // process-wide: merges that took the fused record merge
        if (w == "fetch_ms") return r->consumer->fetch_ms();
        if (w == "device_ms") return r->consumer->device_ms();
    }
    if (r->provider && w == "chunks_served") return (long long)r->provider->chunks_served();
    return -1;
}

void uda_rt_destroy(uda_role_t h) {
    auto* r = (Role*)h;
    if (!r) return;
    if (r->consumer) {
        try { r->consumer->exit_msg(); } catch (...) {}
        r->consumer.reset();
    }
    if (r->provider) {
        r->provider->stop();
        if (r->provider_thread.joinable()) r->provider_thread.join();
        r->provider.reset();
    }
    log_set_host(nullptr);
    delete r;
}

// ---- protocol helpers exported for golden tests ---------------------------------------------------------
int uda_rt_vint_encode(long long v, unsigned char* out9) { return vint_encode(out9, v); }
int uda_rt_vint_decode(const unsigned char* p, int n, long long* out) {
    int64_t v = 0;
    int r = vint_decode(p, p + n, &v);
    *out = v;
    return r;
}
int uda_rt_vint_size(long long v) { return vint_size(v); }

// Writes "count\nheader\nparam0\nparam1..." (params separated by '\n'); returns -1 when unparsable.
int uda_rt_parse_cmd(const char* s, char* out, int cap) {
    HadoopCmd c;
    if (!parse_hadoop_cmd(s ? s : "", &c)) return -1;
    std::string r = std::to_string(c.count) + "\n" + std::to_string(c.header);
    for (auto& p : c.params) r += "\n" + p;
    if ((int)r.size() + 1 > cap) return -2;
    memcpy(out, r.c_str(), r.size() + 1);
    return (int)c.params.size();
}

int uda_rt_form_cmd(int cmd, int nparams, const char* const* params, char* out, int cap) {
    std::vector<std::string> p;
    for (int i = 0; i < nparams; ++i) p.emplace_back(params[i]);
    std::string r = form_cmd(cmd, p);
    if ((int)r.size() + 1 > cap) return -2;
    memcpy(out, r.c_str(), r.size() + 1);
    return (int)r.size();
}

// out7 = {wqes, port, approach, mode, buf_size, log_threshold, ok}
int uda_rt_parse_options(int argc, const char* const* argv, long long* out7, char* log_dir, int cap) {
    std::vector<std::string> a;
    for (int i = 0; i < argc; ++i) a.emplace_back(argv[i]);
    Options o;
    bool ok = parse_options(a, &o);
    out7[0] = o.wqes_per_conn; out7[1] = o.data_port; out7[2] = o.merge_approach; out7[3] = o.mode;
    out7[4] = o.buf_size; out7[5] = o.log_threshold; out7[6] = ok;
    snprintf(log_dir, cap, "%s", o.log_dir.c_str());
    return ok ? 0 : 1;
}

int uda_rt_fetch_request_roundtrip(const char* in, char* out, int cap) {
    FetchRequest r;
    if (!parse_fetch_request(in, &r)) return -1;
    try {
        std::string s = format_fetch_request(r);
        if ((int)s.size() + 1 > cap) return -2;
        memcpy(out, s.c_str(), s.size() + 1);
        return (int)s.size();
    } catch (const UdaError&) {
        return -3;
    }
}

int uda_rt_fetch_ack_roundtrip(const char* in, char* out, int cap) {
    FetchAck a;
    if (!parse_fetch_ack(in, &a)) return -1;
    try {
        std::string s = format_fetch_ack(a);
        if ((int)s.size() + 1 > cap) return -2;
        memcpy(out, s.c_str(), s.size() + 1);
        return (int)s.size();
    } catch (const UdaError&) {
        return -3;
    }
}

// kind: 0 raw, 1 Text, 2 BytesWritable; returns <0, 0, >0
int uda_rt_compare_keys(int kind, const unsigned char* a, int la, const unsigned char* b, int lb) {
    return compare_keys((KeyKind)kind, a, la, b, lb);
}
// -1 when the class is unsupported (the condition that makes the Java side fall back)
int uda_rt_key_kind(const char* java_class) {
    try {
        return (int)key_kind_for_class(java_class);
    } catch (const UdaError&) {
        return -1;
    }
}

int uda_rt_snappy_uncompress(const unsigned char* in, long long n, unsigned char* out, long long out_len) {
    return snappy_uncompress_builtin(in, (size_t)n, out, (size_t)out_len) ? 0 : -1;
}
long long uda_rt_snappy_store(const unsigned char* in, long long n, unsigned char* out, long long cap) {
    std::vector<uint8_t> v;
    snappy_compress_literal(in, (size_t)n, &v);
    if ((long long)v.size() > cap) return -1;
    memcpy(out, v.data(), v.size());
    return (long long)v.size();
}

long long uda_rt_snappy_compress(const unsigned char* in, long long n, unsigned char* out, long long cap) {
    std::vector<uint8_t> v;
    snappy_compress(in, (size_t)n, &v);
    if ((long long)v.size() > cap) return -1;
    memcpy(out, v.data(), v.size());
    return (long long)v.size();
}

long long uda_rt_lzo1x_compress(const unsigned char* in, long long n, unsigned char* out, long long cap) {
    std::vector<uint8_t> v;
    lzo1x_compress(in, (size_t)n, &v);
    if ((long long)v.size() > cap) return -1;
    memcpy(out, v.data(), v.size());
    return (long long)v.size();
}
int uda_rt_lzo1x_decompress(const unsigned char* in, long long n, unsigned char* out, long long out_len) {
    return lzo1x_decompress_builtin(in, (size_t)n, out, (size_t)out_len) ? 0 : -1;
}
// test hook: the device kernel's cooperative copy, run lane by lane on the host
int uda_rt_lzo1x_decompress_lanes(const unsigned char* in, long long n, unsigned char* out, long long out_len, int lanes) {
    return lanes > 0 && lzo1x_decompress_lanes(in, (size_t)n, out, (size_t)out_len, (unsigned)lanes) ? 0 : -1;
}

// out4 = {buffer_size, pairs, pair0, pair1}; returns 0, or 1/2 for the two sizing failures
int uda_rt_plan_memory(int num_maps, long long max_buf, long long min_buf, long long shuffle_mem, int compressed, double ratio,
                       long long* out4) {
    InitParams i;
    i.num_maps = num_maps; i.max_buffer_bytes = max_buf; i.min_buffer_bytes = min_buf; i.shuffle_memory = shuffle_mem;
    try {
        MemoryPlan p = plan_memory(i, num_maps, compressed ? Codec::SNAPPY : Codec::NONE, ratio, max_buf);
        out4[0] = (long long)p.buffer_size; out4[1] = p.pairs; out4[2] = (long long)p.pair_size[0]; out4[3] = (long long)p.pair_size[1];
        return 0;
    } catch (const UdaError& e) {
        return e.info().find("Not enough") != std::string::npos ? 1 : 2;
    }
}

// Test hook: `n` bytes handed to the record parser of the merge (Segment) as one map-output segment arriving in blocks of
// `chunk` bytes, every block in a heap buffer of exactly its size (a read past a block is visible to a sanitizer). Returns the
// number of records before the EOF marker, or -1 when the parser rejects the stream. *digest = order-sensitive FNV-1a hash
// over (key length, value length, key bytes, value bytes) of every record, for comparison with an independent parser.
long long uda_rt_selftest_parse_segment(const unsigned char* data, long long n, int chunk, unsigned long long* digest) {
    class Chunked : public ByteSource {
    public:
        Chunked(const uint8_t* p, size_t n, size_t c) : p_(p), n_(n), c_(c ? c : 1) {}
        bool next_block(const uint8_t** d, size_t* len) override {
            if (off_ >= n_) return false;
            const size_t take = std::min(c_, n_ - off_);
            block_.reset(new uint8_t[take]);          // the previous block dies here, like a recycled fetch buffer
            memcpy(block_.get(), p_ + off_, take);
            off_ += take;
            *d = block_.get();
            *len = take;
            return true;
        }
    private:
        const uint8_t* p_;
        size_t n_, c_, off_ = 0;
        std::unique_ptr<uint8_t[]> block_;
    };
    if (n < 0 || chunk < 1) return -1;
    unsigned long long h = 1469598103934665603ull;
    auto mix = [&](const uint8_t* b, size_t len) { for (size_t i = 0; i < len; ++i) { h ^= b[i]; h *= 1099511628211ull; } };
    long long records = 0;
    try {
        Segment seg(std::make_unique<Chunked>(data, (size_t)n, (size_t)chunk), 0);
        while (seg.next()) {
            const Record& r = seg.cur();
            const uint32_t lens[2] = {(uint32_t)r.klen, (uint32_t)r.vlen};
            mix((const uint8_t*)lens, sizeof(lens));
            mix(r.key, (size_t)r.klen);
            mix(r.val, (size_t)r.vlen);
            ++records;
        }
    } catch (const UdaError&) {
        return -1;
    }
    if (digest) *digest = h;
    return records;
}

// Exercises the three queues with real threads; returns 0 when all invariants held.
int uda_rt_selftest_queues(void) {
    ConcurrentQueue<int> q;
    int tmp;
    if (q.try_pop(tmp)) return 1;
    q.push(7);
    if (!q.try_pop(tmp) || tmp != 7) return 2;  // the reference's try_pop never reports success
    std::atomic<long long> sum{0};
    std::vector<std::thread> ts;
    for (int t = 0; t < 4; ++t) ts.emplace_back([&] { int v; while (q.wait_and_pop(v)) sum += v; });
    for (int i = 1; i <= 1000; ++i) q.push(i);
    q.close();
    for (auto& t : ts) t.join();
    if (sum != 500500) return 3;
    QuotaQueue<int> qq(2);
    std::thread prod([&] { for (int i = 0; i < 50; ++i) qq.push(i); });
    long long s2 = 0;
    for (int i = 0; i < 50; ++i) { if (qq.size() > 2) return 4; s2 += qq.wait_and_pop(); }
    prod.join();
    if (s2 != 1225) return 5;
    ExternalQuotaQueue<int> eq(3);
    std::atomic<int> built{0};
    std::atomic<bool> over{false};
    std::thread f([&] { for (int i = 0; i < 20; ++i) { eq.wait_and_reserve(); if (eq.reserved() > 3) over = true; ++built; eq.push_reserved(i); } });
    for (int i = 0; i < 20; ++i) {
        int v = eq.wait_and_pop_without_dereserve();
        if (v != i) return 6;
        if (eq.reserved() > 3) return 7;
        eq.dereserve();
    }
    f.join();
    return over ? 8 : 0;
}

const char* uda_rt_version(void) { return VERSION_UDA; }


// Exercises spawn_guarded / FailureLatch and the shuffle helpers; returns 0 when all invariants held.
// perm_out (n ints, may be null) receives vector_shuffle(0..n-1) for `seed` so that the order can be pinned.
int uda_rt_selftest_threads(unsigned long long seed, int n, int* perm_out) {
    static std::atomic<int> fails{0};
    fails = 0;
    HostCallbacks cb;
    cb.failure_in_uda = [](void*) { ++fails; };
    FailureLatch latch;
    std::thread a = spawn_guarded("selftest-a", &cb, &latch, [] { throw UdaError("boom"); });
    a.join();
    std::thread b = spawn_guarded("selftest-b", &cb, &latch, [] { throw std::runtime_error("again"); });
    b.join();
    std::thread c = spawn_guarded("selftest-c", &cb, &latch, [] {});
    c.join();
    if (fails != 1 || !latch.failed()) return 1;  // exactly one up-call for two failing threads
    ShuffleRng rng(seed);
    std::vector<int> v(n);
    for (int i = 0; i < n; ++i) v[i] = i;
    vector_shuffle(v, rng);
    std::vector<char> seen(n, 0);
    for (int x : v) { if (x < 0 || x >= n || seen[x]) return 2; seen[x] = 1; }
    if (perm_out) for (int i = 0; i < n; ++i) perm_out[i] = v[i];
    std::list<int> l;
    for (int i = 0; i < n; ++i) l.push_back(i);
    auto lv = list_shuffle_in_vector(l, rng);
    if (!l.empty() || (int)lv.size() != n) return 3;
    // 3 hosts with 4, 2 and 1 entries: no host repeats back to back while another still has work
    std::vector<std::pair<std::string, int>> work;
    for (int i = 0; i < 4; ++i) work.push_back({"h0", i});
    for (int i = 0; i < 2; ++i) work.push_back({"h1", i});
    work.push_back({"h2", 0});
    auto out = shuffle_interleave_hosts(work, rng, [](const std::pair<std::string, int>& x) { return x.first; });
    if (out.size() != work.size()) return 4;
    std::set<std::string> first3{out[0].first, out[1].first, out[2].first};
    if (first3.size() != 3) return 5;
    return 0;
}
}  // extern "C"
