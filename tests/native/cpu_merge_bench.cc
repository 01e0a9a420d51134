// The host runtime's own reducer (online / hybrid CPU merge) fed from memory: the counterpart of baseline/ref_harness.cc,
// which feeds the REFERENCE's reducer the same way. Same record shape (teragen: 10-byte key, 90-byte value, IFile framed),
// same generator, same "transport" (a copy thread that memcpy's the next chunk into the consumer's buffer and completes the
// fetch), same timed span (INIT -> last block handed to "Java", a byte counter), so the two CPU mergers can be compared
// without a network, disks or a JVM in the way:
//
//   build/bin/cpu_merge_bench <libuda.so> --maps 8 --records 20000000 --steps 3 --warmup 1 [--approach 1|2|3] [--check]
//     (approach 3 = the device merge: needs a GPU, or UDA_KERNELS_LIB pointing at the host-only mock)
//   baseline/_ref/bin/ref_merger           --maps 8 --records 20000000 --steps 3 --warmup 1 [--check]
//
//   ... --codec snappy|lzo      the map outputs are block-compressed (256 KiB blocks); the reducer decompresses them
//
//   build/bin/cpu_merge_bench <libuda.so> --tcp --maps 4 --reducers 2 --records 10000000    (BASELINE config 1, native)
//     the map outputs are written as Hadoop file.out files, a provider role of this process serves them over TCP
//     loop-back (index look-up, chunk pool, disk reader, fetch protocol with credits) and --reducers consumer roles
//     fetch and merge them at the same time; --records is the total over all reducers
//
// prints one line:  OURS steps=<k> seconds=<s0,...> bytes=<merged stream bytes per step> blocks=<n> check=<ok|skipped|FAILED>
// (the first argument is accepted for symmetry with the other native tests; the program links libuda.so)
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include <algorithm>
#include <condition_variable>
#include <deque>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/uda_rt_api.h"
#include "../../csrc/runtime/consumer.h"
#include "../../csrc/runtime/transport.h"

namespace {

constexpr int KEY = 10, VAL = 90, REC = 102;
std::vector<std::vector<char>> g_maps;       // what the "provider" holds per segment: plain, or codec blocks when --codec is given
std::vector<long long> g_plain_len;          // plain length of every segment (rawLength of its index record)
std::string g_codec;                         // "", "snappy" or "lzo"
constexpr size_t kCodecBlock = 256 * 1024;   // io.compression.codec.{snappy,lzo}.buffersize default

// plain segment -> Hadoop block framing [u32 BE rawLen][u32 BE compLen][bytes]... with this runtime's own compressors
void compress_segment(std::vector<char>* seg) {
    std::vector<char> out;
    std::vector<uint8_t> blk;
    for (size_t off = 0; off < seg->size(); off += kCodecBlock) {
        const size_t n = std::min(kCodecBlock, seg->size() - off);
        if (g_codec == "snappy") uda::snappy_compress((const uint8_t*)seg->data() + off, n, &blk);
        else uda::lzo1x_compress((const uint8_t*)seg->data() + off, n, &blk);
        const uint32_t hdr[2] = {__builtin_bswap32((uint32_t)n), __builtin_bswap32((uint32_t)blk.size())};
        out.insert(out.end(), (const char*)hdr, (const char*)hdr + 8);
        out.insert(out.end(), (const char*)blk.data(), (const char*)blk.data() + blk.size());
    }
    seg->swap(out);
}

inline uint64_t splitmix(uint64_t& s) {
    uint64_t z = (s += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
struct KeyIdx {
    uint64_t hi;
    uint16_t lo;
    uint32_t idx;
    bool operator<(const KeyIdx& o) const { return hi != o.hi ? hi < o.hi : (lo != o.lo ? lo < o.lo : idx < o.idx); }
};
void build_map(int m, long long n, uint64_t seed) {   // identical to baseline/ref_harness.cc: both arms merge the same bytes
    std::vector<KeyIdx> keys((size_t)n);
    uint64_t s = seed * 0x100000001B3ull + (uint64_t)m * 0x9E3779B1ull + 17;
    for (long long i = 0; i < n; ++i) {
        keys[(size_t)i].hi = splitmix(s);
        keys[(size_t)i].lo = (uint16_t)splitmix(s);
        keys[(size_t)i].idx = (uint32_t)i;
    }
    std::sort(keys.begin(), keys.end());
    std::vector<char>& seg = g_maps[(size_t)m];
    seg.resize((size_t)n * REC + 2);
    char* p = seg.data();
    for (long long i = 0; i < n; ++i) {
        const KeyIdx& k = keys[(size_t)i];
        p[0] = (char)KEY; p[1] = (char)VAL;
        for (int b = 0; b < 8; ++b) p[2 + b] = (char)(k.hi >> (56 - 8 * b));
        p[10] = (char)(k.lo >> 8); p[11] = (char)k.lo;
        uint64_t v = ((uint64_t)m << 40) ^ k.idx ^ seed;
        for (int b = 0; b < VAL; b += 8) {
            const uint64_t x = splitmix(v);
            memcpy(p + 12 + b, &x, (size_t)std::min(8, VAL - b));
        }
        p += REC;
    }
    p[0] = (char)0xFF; p[1] = (char)0xFF;
}
bool g_verbose = false;   // UDA_BENCH_VERBOSE=1: per-step busy time of the copy thread on stderr
double now_s() {
    timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

// ---- transport: one copy thread, like the reference harness ------------------------------------------------------
class MemoryInputClient : public uda::InputClient {
public:
    MemoryInputClient() { th_ = std::thread([this] { run(); }); }
    ~MemoryInputClient() override {
        stop();
        if (g_verbose) fprintf(stderr, "copy thread: %.3f s busy in %lld copies\n", copy_busy_, copies_);
    }
    void start() override {}
    void stop() override {
        {
            std::lock_guard<std::mutex> lk(mu_);
            if (stop_) return;
            stop_ = true;
        }
        cv_.notify_all();
        if (th_.joinable()) th_.join();
        for (auto& j : q_) j.done(false, uda::FetchAck());   // fail what was still queued, like a closed connection would
        q_.clear();
    }
    void fetch(const std::string&, uda::FetchRequest req, uint8_t* dest, size_t capacity, Done done) override {
        {
            std::lock_guard<std::mutex> lk(mu_);
            q_.push_back(Job{std::move(req), dest, capacity, std::move(done)});
        }
        cv_.notify_one();
    }

private:
    struct Job { uda::FetchRequest req; uint8_t* dest; size_t cap; Done done; };
    void run() {
        for (;;) {
            Job j;
            {
                std::unique_lock<std::mutex> lk(mu_);
                cv_.wait(lk, [&] { return stop_ || !q_.empty(); });
                if (stop_) return;
                j = std::move(q_.front());
                q_.pop_front();
            }
            const int m = atoi(j.req.map_id.c_str());
            const std::vector<char>& seg = g_maps[(size_t)m];
            const long long total = (long long)seg.size();
            long long n = total - j.req.map_offset;
            if (n > j.req.chunk_size) n = j.req.chunk_size;
            if (n > (long long)j.cap) n = (long long)j.cap;
            if (n < 0) n = 0;
            const double tc = g_verbose ? now_s() : 0;
            memcpy(j.dest, seg.data() + j.req.map_offset, (size_t)n);
            if (g_verbose) { copy_busy_ += now_s() - tc; ++copies_; }
            uda::FetchAck ack;
            ack.raw_length = g_plain_len[(size_t)m];
            ack.part_length = total;
            ack.sent = n;
            ack.file_offset = 0;
            ack.path = "mem";
            j.done(true, ack);
        }
    }
    std::mutex mu_;
    std::condition_variable cv_;
    std::deque<Job> q_;
    bool stop_ = false;
    double copy_busy_ = 0;
    long long copies_ = 0;
    std::thread th_;
};

// ---- "Java" side ---------------------------------------------------------------------------------------------------
struct Sink {
    std::mutex mu;
    std::condition_variable cv;
    long long bytes = 0, blocks = 0, records = 0, expected = 0;
    double t_last = 0;
    bool check = false, check_ok = true, failed = false;
    std::string prev_key;
    void reset(bool chk, long long expect) {
        bytes = blocks = records = 0;
        expected = expect;
        t_last = 0;
        check = chk;
        check_ok = true;
        failed = false;
        prev_key.clear();
    }
};
std::deque<Sink> g_sinks;   // one per reducer (a deque: Sink is not movable)

void on_data(void* user, const void* buf, int len) {
    Sink& g_sink = *(Sink*)user;
    const char* p = (const char*)buf;
    if (g_sink.check) {
        int off = 0;
        while (off + 2 <= len) {
            if ((unsigned char)p[off] == 0xFF && (unsigned char)p[off + 1] == 0xFF) break;
            if (p[off] != KEY || p[off + 1] != VAL || off + REC > len) { g_sink.check_ok = false; break; }
            std::string k(p + off + 2, KEY);
            if (!g_sink.prev_key.empty() && k < g_sink.prev_key) g_sink.check_ok = false;
            g_sink.prev_key.swap(k);
            ++g_sink.records;
            off += REC;
        }
    }
    std::lock_guard<std::mutex> lk(g_sink.mu);
    g_sink.bytes += len;
    ++g_sink.blocks;
    g_sink.t_last = now_s();
    if (g_sink.bytes >= g_sink.expected) g_sink.cv.notify_all();
}
void on_failure(void* user) {
    Sink& g_sink = *(Sink*)user;
    std::lock_guard<std::mutex> lk(g_sink.mu);
    g_sink.failed = true;
    g_sink.cv.notify_all();
}
bool g_inline_decompress = false;   // --inline-decompress: no decompress thread (A/B of the decode-ahead)
int on_conf(void*, const char* name, const char*, char* out, int cap) {
    if (g_inline_decompress && strcmp(name, "mapred.uda.decompress.ahead") == 0) { snprintf(out, (size_t)cap, "false"); return 1; }
    return 0;
}

struct StepResult { double seconds; long long bytes, blocks, records; int check_ok; };

// ---- file-backed map outputs for the --tcp mode -------------------------------------------------------------------------
struct MofIndex { long long start, raw, part; };
std::vector<std::string> g_mof_path;                 // per map
std::vector<std::vector<MofIndex>> g_mof_index;      // [map][reducer]
int g_reducers = 1;

// segment (map m, reducer r) lives in g_maps[m * g_reducers + r]; file.out = the map's segments, each followed by 4 checksum bytes
bool write_mofs(const std::string& dir, int maps) {
    g_mof_path.assign((size_t)maps, "");
    g_mof_index.assign((size_t)maps, std::vector<MofIndex>((size_t)g_reducers));
    for (int m = 0; m < maps; ++m) {
        g_mof_path[(size_t)m] = dir + "/map_" + std::to_string(m) + ".out";
        FILE* f = fopen(g_mof_path[(size_t)m].c_str(), "wb");
        if (!f) return false;
        long long pos = 0;
        for (int r = 0; r < g_reducers; ++r) {
            const std::vector<char>& seg = g_maps[(size_t)(m * g_reducers + r)];
            const char crc[4] = {0, 0, 0, 0};
            if (fwrite(seg.data(), 1, seg.size(), f) != seg.size() || fwrite(crc, 1, 4, f) != 4) { fclose(f); return false; }
            g_mof_index[(size_t)m][(size_t)r] = MofIndex{pos, g_plain_len[(size_t)(m * g_reducers + r)], (long long)seg.size() + 4};
            pos += (long long)seg.size() + 4;
        }
        fclose(f);
    }
    return true;
}
int on_path(void*, const char*, const char* map, int reduce, long long* start, long long* raw, long long* part, char* path_out, int cap) {
    const int m = atoi(map);
    if (m < 0 || m >= (int)g_mof_path.size() || reduce < 0 || reduce >= g_reducers) return 0;
    const MofIndex& ix = g_mof_index[(size_t)m][(size_t)reduce];
    *start = ix.start; *raw = ix.raw; *part = ix.part;
    snprintf(path_out, (size_t)cap, "%s", g_mof_path[(size_t)m].c_str());
    return 1;
}

// One job: every reducer's consumer role is started, initialised and told about every map; the span runs from the first
// INIT to the last block of the slowest reducer. `port` > 0: fetch over TCP from the provider role listening there.
bool run_job(int maps, int approach, bool check, int port, StepResult* out) {
    const int R = g_reducers;
    uda_host_callbacks_t cb;
    memset(&cb, 0, sizeof(cb));
    cb.data_from_uda = on_data;
    cb.failure_in_uda = on_failure;
    cb.get_conf_data = on_conf;
    const double t0 = now_s();
    const std::string a = std::to_string(approach), prt = std::to_string(port > 0 ? port : 9011);
    const char* argv[] = {"-w", "256", "-r", prt.c_str(), "-a", a.c_str(), "-m", "1", "-g", "/tmp/", "-s", "1024"};
    std::vector<uda_role_t> roles;
    bool ok = true;
    for (int r = 0; r < R && ok; ++r) {
        long long expect = 2;
        for (int m = 0; m < maps; ++m) expect += g_plain_len[(size_t)(m * R + r)] - 2;
        g_sinks[(size_t)r].reset(check, expect);
        cb.user = &g_sinks[(size_t)r];
        uda_role_t h = uda_rt_start(1, (int)(sizeof(argv) / sizeof(argv[0])), argv, 0 /* logging off */, 0, &cb);
        if (!h) { ok = false; break; }
        roles.push_back(h);
        char init[512];
        const long long shuffle_mem = (long long)maps * 2 * (1 << 20) + (64LL << 20);
        const char* codec_class = g_codec == "snappy" ? "org.apache.hadoop.io.compress.SnappyCodec"
                                  : g_codec == "lzo" ? "com.hadoop.compression.lzo.LzoCodec" : "null";
        snprintf(init, sizeof(init), "13:7:%d:job_ref_0001:attempt_ref_0001_r_%06d_0:0:%d:%d:org.apache.hadoop.io.ByteWritable:%s:%d:%lld:1:/tmp",
                 maps, r, 1 << 20, 16 << 10, codec_class, g_codec.empty() ? 0 : (int)kCodecBlock, shuffle_mem);
        ok = uda_rt_do_command(h, init) == 0;
        for (int m = 0; ok && m < maps; ++m) {
            char fetch[256];
            // memory-fed: the map id is the segment index; over TCP the provider resolves (map, reduce) through on_path
            snprintf(fetch, sizeof(fetch), "5:4:127.0.0.1:job_ref_0001:%d:%d", port > 0 ? m : m * R + r, r);
            ok = uda_rt_do_command(h, fetch) == 0;
        }
    }
    out->seconds = 0;
    out->bytes = out->blocks = out->records = 0;
    out->check_ok = 1;
    for (size_t r = 0; r < roles.size(); ++r) {
        Sink& sk = g_sinks[r];
        if (ok) {
            std::unique_lock<std::mutex> lk(sk.mu);
            sk.cv.wait(lk, [&] { return sk.failed || sk.bytes >= sk.expected; });
            ok = !sk.failed;
        }
        out->seconds = std::max(out->seconds, sk.t_last - t0);
        out->bytes += sk.bytes;
        out->blocks += sk.blocks;
        out->records += sk.records;
        out->check_ok = out->check_ok && sk.check_ok;
        if (!ok) fprintf(stderr, "reduce task %zu failed: %s\n", r, uda_rt_last_error(roles[r]));
    }
    if (g_verbose && !roles.empty())
        fprintf(stderr, "merge threads: %lld stalls, %.3f s waiting for chunks (cumulative)\n", uda_rt_stat(roles[0], "fetch_stalls"),
                1e-6 * (double)uda_rt_stat(roles[0], "fetch_stall_us"));
    for (uda_role_t h : roles) {   // outside the timed span, like the JVM's exit
        uda_rt_reduce_exit(h);
        uda_rt_destroy(h);
    }
    return ok;
}

}  // namespace

int main(int argc, char** argv) {
    int maps = 8, steps = 3, warmup = 1, approach = 1;
    long long records = 2000000;
    uint64_t seed = 1234;
    double budget_s = 600;
    bool check = false, tcp = false;
    for (int i = 1; i < argc; ++i) {
        std::string a = argv[i];
        auto val = [&]() -> const char* { return i + 1 < argc ? argv[++i] : "0"; };
        if (a == "--maps") maps = atoi(val());
        else if (a == "--reducers") g_reducers = atoi(val());
        else if (a == "--records") records = atoll(val());
        else if (a == "--steps") steps = atoi(val());
        else if (a == "--warmup") warmup = atoi(val());
        else if (a == "--seed") seed = strtoull(val(), nullptr, 10);
        else if (a == "--approach") approach = atoi(val());
        else if (a == "--budget-s") budget_s = atof(val());
        else if (a == "--check") check = true;
        else if (a == "--tcp") tcp = true;
        else if (a == "--codec") g_codec = val();
        else if (a == "--inline-decompress") g_inline_decompress = true;
    }
    if (maps < 1 || records < 1 || g_reducers < 1 || (approach < 1 || approach > 3) || (!g_codec.empty() && g_codec != "snappy" && g_codec != "lzo")) {
        fprintf(stderr, "bad arguments\n");
        return 2;
    }
    g_verbose = getenv("UDA_BENCH_VERBOSE") != nullptr;
    const double t_begin = now_s();
    const int R = g_reducers, nseg = maps * R;
    g_maps.resize((size_t)nseg);
    g_plain_len.assign((size_t)nseg, 0);
    g_sinks.resize((size_t)R);
    {
        // `records` is the total of the job (what bench.py calls records per GPU for one reducer), spread over maps x reducers
        std::vector<std::thread> th;
        for (int sgi = 0; sgi < nseg; ++sgi) {
            const long long n = records / nseg + (sgi < records % nseg ? 1 : 0);
            th.emplace_back([=] {
                build_map(sgi, n, seed);
                g_plain_len[(size_t)sgi] = (long long)g_maps[(size_t)sgi].size();
                if (!g_codec.empty()) compress_segment(&g_maps[(size_t)sgi]);
            });
            if (th.size() == 16) { for (auto& t : th) t.join(); th.clear(); }
        }
        for (auto& t : th) t.join();
    }
    uda_role_t provider = nullptr;
    int port = 0;
    char tmpl[] = "/tmp/uda_cpu_merge_bench_XXXXXX";
    std::string dir;
    if (tcp) {
        if (!mkdtemp(tmpl)) { perror("mkdtemp"); return 1; }
        dir = tmpl;
        if (!write_mofs(dir, maps)) { fprintf(stderr, "cannot write the map outputs under %s\n", dir.c_str()); return 1; }
        uda_host_callbacks_t pcb;
        memset(&pcb, 0, sizeof(pcb));
        pcb.get_path_uda = on_path;
        pcb.get_conf_data = on_conf;
        const char* pargv[] = {"-w", "256", "-r", "0", "-m", "1", "-g", "/tmp/", "-s", "1024"};
        provider = uda_rt_start(0, (int)(sizeof(pargv) / sizeof(pargv[0])), pargv, 0, 0, &pcb);
        port = provider ? uda_rt_port(provider) : 0;
        if (port <= 0) { fprintf(stderr, "the provider role did not start\n"); return 1; }
    } else {
        uda::set_input_client_factory([](int, int) { return std::unique_ptr<uda::InputClient>(new MemoryInputClient()); });
    }
    std::vector<double> secs;
    long long bytes = 0, blocks = 0;
    bool all_ok = true, checked = false;
    double per_step = 0;
    for (int s = 0; s < warmup + steps; ++s) {
        if (s > 0 && per_step > 0 && (now_s() - t_begin) + per_step > budget_s) break;
        StepResult r;
        if (!run_job(maps, approach, check && s == 0, port, &r)) { all_ok = false; break; }
        per_step = r.seconds;
        if (check && s == 0) { checked = true; all_ok = all_ok && r.check_ok && r.records == records; }
        if (s >= warmup) secs.push_back(r.seconds);
        bytes = r.bytes;
        blocks = r.blocks;
    }
    if (provider) {
        uda_rt_do_command(provider, "");   // the empty command is EXIT
        uda_rt_destroy(provider);
        for (const auto& p : g_mof_path) remove(p.c_str());
        remove(dir.c_str());
    }
    printf("OURS steps=%zu seconds=", secs.size());
    for (size_t i = 0; i < secs.size(); ++i) printf("%s%.6f", i ? "," : "", secs[i]);
    printf(" bytes=%lld blocks=%lld check=%s\n", bytes, blocks, !all_ok ? "FAILED" : (checked ? "ok" : "skipped"));
    return all_ok ? 0 : 1;
}
