// Reference arm of bench.py: drives the UNMODIFIED reducer side of chaoyang223/UDA on the CPU.
//
// What is the reference's own code here (compiled as-is from a copy of /root/reference/src, see
// baseline/build_ref.py): NetMergerMain.cc (MergeManager_main), reducer.cc (INIT / FETCH / EXIT handling,
// memory plan), MergeManager.cc (fetching phase, merging phase, merge thread), StreamRW.cc (Segment,
// nextKV, switch_mem/join, write_kv_to_stream), MergeQueue.h (heap), CompareFunc.cc, IOUtility.cc (streams,
// VInt, log), C2JNexus.cc (option + command parser), UdaUtil.cc, the decompressor wrappers.
//
// What this file supplies instead of the parts that cannot exist on this image:
//   * RdmaClient (declared by the reference's own DataNet/RDMAClient.h): libibverbs / librdmacm are absent,
//     so the fetch "transport" is a copy thread that memcpy's the next chunk of the requested map-output
//     partition from provider memory into the reducer's registered buffer and then completes the request
//     through the reference's own MergeManager::update_fetch_req / mark_req_as_ready — exactly the two calls
//     its RDMA completion handler makes (src/DataNet/RDMAClient.cc:529-543). A memcpy is the most
//     favourable transport the reference could have.
//   * UdaBridge_* (src/UdaBridge.cc needs a JVM): dataFromUda counts the delivered bytes (the Java side's
//     copy out of the direct buffer is NOT charged to the reference), getConfData returns the defaults.
//   * the data: `maps` sorted IFile segments of teragen-shaped records (10-byte random key, 90-byte value,
//     102 bytes framed) generated in host memory at set-up.
//
// One reduce task per process, as in Hadoop: every step forks a child that runs MergeManager_main ->
// INIT -> FETCH x maps -> merge -> EXIT and reports the time from INIT to the last delivered block.
//
//   ref_merger --maps 8 --records 125000000 --steps 5 --warmup 1 [--check] [--seed S] [--budget-s 600]
// prints one line:  REF steps=<k> seconds=<s0,s1,...> bytes=<merged stream bytes per step> blocks=<n> check=<ok|skipped|FAILED>
#include <algorithm>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include <sys/wait.h>
#include <time.h>
#include <unistd.h>

#include "DataNet/RDMAClient.h"
#include "Merger/reducer.h"
#include "Merger/MergeManager.h"
#include "UdaBridge.h"
#include "IOUtility.h"

int MergeManager_main(int argc, char* argv[]) throw(UdaException*);

// ------------------------------------------------------------------------------------------ data
namespace {

constexpr int KEY = 10, VAL = 90, REC = 102;   // VInt(10) VInt(90) key value

struct MapSegment {
    std::vector<char> bytes;   // IFile segment: records + EOF marker (-1, -1)
};
std::vector<MapSegment> g_maps;
long long g_expected_stream = 0;    // bytes the merged stream must have (all records + one EOF marker)

inline uint64_t splitmix(uint64_t& s) {
    uint64_t z = (s += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

struct KeyIdx {
    uint64_t hi;     // key bytes 0..7 big-endian
    uint16_t lo;     // key bytes 8..9
    uint32_t idx;
    bool operator<(const KeyIdx& o) const { return hi != o.hi ? hi < o.hi : (lo != o.lo ? lo < o.lo : idx < o.idx); }
};

void build_map(int m, long long n, uint64_t seed) {
    std::vector<KeyIdx> keys((size_t)n);
    uint64_t s = seed * 0x100000001B3ull + (uint64_t)m * 0x9E3779B1ull + 17;
    for (long long i = 0; i < n; ++i) {
        keys[(size_t)i].hi = splitmix(s);
        keys[(size_t)i].lo = (uint16_t)splitmix(s);
        keys[(size_t)i].idx = (uint32_t)i;
    }
    std::sort(keys.begin(), keys.end());
    MapSegment& seg = g_maps[(size_t)m];
    seg.bytes.resize((size_t)n * REC + 2);
    char* p = seg.bytes.data();
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
    p[0] = (char)0xFF; p[1] = (char)0xFF;   // VInt(-1) VInt(-1)
}

double now_s() {
    timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

// ------------------------------------------------------------------------------------------ "Java" side
struct Sink {
    std::mutex mu;
    std::condition_variable cv;
    long long bytes = 0, blocks = 0;
    double t_last = 0;
    bool check = false, check_ok = true;
    std::string prev_key;
    long long records = 0;
} g_sink;

JNIEnv g_env;

// ------------------------------------------------------------------------------------------ transport
struct CopyJob { client_part_req_t* req; char* buff; int len; };
bool g_verbose = false;   // UDA_BENCH_VERBOSE=1: how busy the stand-in transport was (it must not be what limits the reducer)
struct CopyThread {
    double busy_s = 0;
    long long copies = 0;
    std::mutex mu;
    std::condition_variable cv;
    std::deque<CopyJob> q;
    bool stop = false;
    std::thread th;
    RdmaClient* owner = nullptr;

    void run() {
        for (;;) {
            CopyJob job;
            {
                std::unique_lock<std::mutex> lk(mu);
                cv.wait(lk, [&] { return stop || !q.empty(); });
                if (q.empty()) return;
                job = q.front();
                q.pop_front();
            }
            client_part_req_t* req = job.req;
            const int m = atoi(req->info->params[2]);          // map id = index of the segment
            const MapSegment& seg = g_maps[(size_t)m];
            const long long total = (long long)seg.bytes.size();
            const long long off = req->mop->fetched_len_rdma;
            long long n = total - off;
            if (n > job.len) n = job.len;
            if (n < 0) n = 0;
            const double tc = g_verbose ? now_s() : 0;
            memcpy(job.buff, seg.bytes.data() + off, (size_t)n);   // the "RDMA write" into the registered buffer
            if (g_verbose) { busy_s += now_s() - tc; ++copies; }
            // ack text the provider sends: rawLength:partLength:sent:mofOffset:mofPath:  (parsed by update_fetch_req)
            snprintf(req->recvd_msg, sizeof(req->recvd_msg), "%lld:%lld:%lld:%lld:%s:", total, total, n, 0LL, "mem");
            owner->comp_fetch_req(req);
        }
    }
} g_copy;

}  // namespace

// ------------------------------------------------------------------------------------------ RdmaClient (memory fed)
RdmaClient::RdmaClient(int port, reduce_task_t* rt) {
    svc_port = port;
    reduce_task = rt;
    parent = NULL;
    memset(&helper, 0, sizeof(helper));
    memset(&ctx, 0, sizeof(ctx));
    INIT_LIST_HEAD(&register_mems_head);
    g_copy.owner = this;
    g_copy.stop = false;
    g_copy.th = std::thread([] { g_copy.run(); });
}

RdmaClient::~RdmaClient() {
    {
        std::lock_guard<std::mutex> lk(g_copy.mu);
        g_copy.stop = true;
    }
    g_copy.cv.notify_all();
    if (g_copy.th.joinable()) g_copy.th.join();
}

void RdmaClient::start_client() { parent = reduce_task->client; }
void RdmaClient::stop_client() {}
RdmaClient* RdmaClient::getRdmaClient() { return this; }

// the receive pool: `num` pairs of buffers carved out of one page-aligned allocation (the reference registers the
// same region with the HCA; here there is nothing to register)
void RdmaClient::register_mem(struct memory_pool* pool, double_buffer_t buffers) {
    void* mem = NULL;
    if (posix_memalign(&mem, (size_t)getpagesize(), (size_t)pool->total_size) != 0 || !mem) throw new UdaException("pool allocation failed");
    memset(mem, 0, (size_t)pool->total_size);
    pool->mem = (char*)mem;
    if (split_mem_pool_to_pairs(pool, buffers)) throw new UdaException("failure in split_mem_pool_to_pairs");
}

int RdmaClient::split_mem_pool_to_pairs(memory_pool_t* pool, double_buffer_t buffers) {
    pthread_mutex_init(&pool->lock, NULL);
    INIT_LIST_HEAD(&pool->free_descs);
    const int pairs = pool->num;
    pool->desc_arr = new mem_desc_t[2 * pairs];
    pool->pair_desc_arr = new mem_set_desc_t[pairs];
    char* at = pool->mem;
    for (int i = 0; i < pairs; ++i) {
        const int len[NUM_STAGE_MEM] = {buffers.buffer1, buffers.buffer2};
        for (int b = 0; b < NUM_STAGE_MEM; ++b) {
            mem_desc_t* d = &pool->desc_arr[2 * i + b];
            d->init(at, len[b]);
            pool->pair_desc_arr[i].buffer_unit[b] = d;
            at += len[b];
        }
        list_add_tail(&pool->pair_desc_arr[i].list, &pool->free_descs);
    }
    return 0;
}

int RdmaClient::start_fetch_req(client_part_req_t* freq, char* buff, int32_t buf_len) {
    if (buf_len <= 0) throw new UdaException("illegal fetch request size of 0 or less bytes");
    {
        std::lock_guard<std::mutex> lk(g_copy.mu);
        g_copy.q.push_back(CopyJob{freq, buff, (int)buf_len});
    }
    g_copy.cv.notify_one();
    return 0;
}

// same two calls as the reference's completion path (src/DataNet/RDMAClient.cc:529-543)
void RdmaClient::comp_fetch_req(client_part_req_t* req) {
    if (parent == this || parent == NULL) {
        MergeManager* mm = req->mop->task->merge_man;
        mm->update_fetch_req(req);
        mm->mark_req_as_ready(req);
    } else {
        parent->comp_fetch_req(req);
    }
}

netlev_conn_t* RdmaClient::connect(const char*, int) { return NULL; }
netlev_conn_t* RdmaClient::netlev_get_conn(unsigned long, int, netlev_ctx_t*, list_head_t*) { return NULL; }
unsigned long RdmaClient::get_hostip(const char*) { return 0; }

// ------------------------------------------------------------------------------------------ UdaBridge (no JVM)
JNIEnv* UdaBridge_threadGetEnv() { return &g_env; }
JNIEnv* UdaBridge_attachNativeThread() { return &g_env; }
void UdaBridge_detachNativeThread() {}
void UdaBridge_onUnloadCleanup() {}
jobject UdaBridge_registerDirectByteBuffer(JNIEnv*, void* address, long) { return (jobject)address; }
void UdaBridge_invoke_fetchOverMessage_callback(JNIEnv*) {}
void UdaBridge_invoke_logToJava_callback(const char* msg, int severity) {
    if (severity <= (int)lsERROR) fprintf(stderr, "[ref log] %s\n", msg);
}
std::string UdaBridge_invoke_getConfData_callback(const char*, const char* defaultValue) { return std::string(defaultValue); }
index_record* UdaBridge_invoke_getPathUda_callback(JNIEnv*, const char*, const char*, int) { return NULL; }
void UdaBridge_exceptionInNativeThread(JNIEnv*, UdaException* ex) {
    fprintf(stderr, "[ref] exception in native thread: %s\n", ex ? ex->getFullMessage().c_str() : "?");
    _exit(3);
}

// One merged block (<= 1 MiB of whole records) handed to "Java".
void UdaBridge_invoke_dataFromUda_callback(JNIEnv*, jobject jbuf, int len) {
    Sink& s = g_sink;
    if (s.check) {   // untimed validation pass: keys must be non-decreasing across the whole stream
        const unsigned char* p = (const unsigned char*)jbuf;
        int off = 0;
        while (off + 2 <= len) {
            if (p[off] == 0xFF && p[off + 1] == 0xFF) { off += 2; break; }
            if (p[off] != KEY || p[off + 1] != VAL || off + REC > len) { s.check_ok = false; break; }
            std::string k((const char*)p + off + 2, KEY);
            if (!s.prev_key.empty() && k < s.prev_key) s.check_ok = false;
            s.prev_key.swap(k);
            ++s.records;
            off += REC;
        }
        if (off != len) s.check_ok = false;
    }
    std::lock_guard<std::mutex> lk(s.mu);
    s.bytes += len;
    ++s.blocks;
    if (s.bytes >= g_expected_stream) {
        s.t_last = now_s();
        s.cv.notify_all();
    }
}

// ------------------------------------------------------------------------------------------ one reduce task
namespace {

struct StepResult { double seconds; long long bytes, blocks, records; int check_ok; };

StepResult run_reduce_task(int maps, bool check) {
    g_sink.check = check;
    const double t0 = now_s();
    // the consumer command line the Java plugin builds (UdaPluginRT): online merge, 1 MiB buffers, logging off
    const char* argv_c[] = {"NetMerger", "-w", "256", "-r", "9011", "-a", "1", "-m", "1", "-g", "/tmp/", "-t", "0", "-s", "1024"};
    const int argc = (int)(sizeof(argv_c) / sizeof(argv_c[0]));
    std::vector<char*> argv;
    for (int i = 0; i < argc; ++i) argv.push_back(strdup(argv_c[i]));
    argv.push_back(NULL);
    optind = 1;
    MergeManager_main(argc, argv.data());
    // INIT: num_maps, job, reduce task, lpq size, rdma buffer bytes, min buffer bytes, key class, codec, block size,
    //       shuffle memory, #dirs, dir. ("count:header:p0:...": count = header + parameters; the last parameter runs to
    //       the end of the string.) The key class selects the plain byte comparator
    //       (src/Merger/CompareFunc.cc:33-38,75-77): teragen keys are 10 raw bytes.
    char init[512];
    const long long shuffle_mem = (long long)maps * 2 * (1 << 20) + (64LL << 20);
    snprintf(init, sizeof(init), "13:7:%d:job_ref_0001:attempt_ref_0001_r_000000_0:0:%d:%d:org.apache.hadoop.io.ByteWritable:null:0:%lld:1:/tmp",
             maps, 1 << 20, 16 << 10, shuffle_mem);
    reduce_downcall_handler(std::string(init));
    for (int m = 0; m < maps; ++m) {   // FETCH: host, job, map (= segment index), reduce
        char fetch[256];
        snprintf(fetch, sizeof(fetch), "5:4:localhost:job_ref_0001:%d:0", m);
        reduce_downcall_handler(std::string(fetch));
    }
    {
        std::unique_lock<std::mutex> lk(g_sink.mu);
        g_sink.cv.wait(lk, [] { return g_sink.bytes >= g_expected_stream; });
    }
    StepResult r;
    r.seconds = g_sink.t_last - t0;
    r.bytes = g_sink.bytes;
    r.blocks = g_sink.blocks;
    r.records = g_sink.records;
    r.check_ok = g_sink.check_ok ? 1 : 0;
    if (g_verbose)
        fprintf(stderr, "copy thread (stand-in transport): %.3f s busy in %lld copies during a %.3f s reduce task\n", g_copy.busy_s, g_copy.copies,
                r.seconds);
    reduce_exit_msg_handler();   // joins the merge thread, frees the pools (outside the timed span, like the JVM's exit)
    return r;
}

bool run_in_child(int maps, bool check, StepResult* out) {
    int fd[2];
    if (pipe(fd) != 0) return false;
    fflush(NULL);
    const pid_t pid = fork();
    if (pid < 0) return false;
    if (pid == 0) {
        close(fd[0]);
        StepResult r = run_reduce_task(maps, check);
        ssize_t w = write(fd[1], &r, sizeof(r));
        (void)w;
        _exit(0);
    }
    close(fd[1]);
    StepResult r;
    memset(&r, 0, sizeof(r));
    size_t got = 0;
    while (got < sizeof(r)) {
        const ssize_t n = read(fd[0], (char*)&r + got, sizeof(r) - got);
        if (n <= 0) break;
        got += (size_t)n;
    }
    close(fd[0]);
    int status = 0;
    waitpid(pid, &status, 0);
    if (got != sizeof(r) || !WIFEXITED(status) || WEXITSTATUS(status) != 0) return false;
    *out = r;
    return true;
}

}  // namespace

int main(int argc, char** argv) {
    g_verbose = getenv("UDA_BENCH_VERBOSE") != NULL;
    int maps = 8, steps = 5, warmup = 1;
    long long records = 1000000;
    uint64_t seed = 1234;
    bool check = false;
    double budget_s = 600;
    for (int i = 1; i < argc; ++i) {
        std::string a = argv[i];
        auto val = [&]() -> const char* { return i + 1 < argc ? argv[++i] : "0"; };
        if (a == "--maps") maps = atoi(val());
        else if (a == "--records") records = atoll(val());
        else if (a == "--steps") steps = atoi(val());
        else if (a == "--warmup") warmup = atoi(val());
        else if (a == "--seed") seed = strtoull(val(), NULL, 10);
        else if (a == "--budget-s") budget_s = atof(val());
        else if (a == "--check") check = true;
    }
    if (maps < 1 || records < 0 || steps < 1) { fprintf(stderr, "bad arguments\n"); return 2; }
    // ---- set-up (untimed): the map outputs of this reducer's partition, in "provider memory"
    g_maps.resize((size_t)maps);
    {
        std::vector<std::thread> th;
        for (int m = 0; m < maps; ++m) {
            const long long n = records / maps + (m < records % maps ? 1 : 0);
            th.emplace_back([m, n, seed] { build_map(m, n, seed); });
        }
        for (auto& t : th) t.join();
    }
    g_expected_stream = records * REC + 2;

    const double t_begin = now_s();
    StepResult r;
    for (int w = 0; w < warmup; ++w)
        if (!run_in_child(maps, false, &r)) { fprintf(stderr, "reference reduce task failed (warm-up)\n"); return 1; }
    const double per_step = warmup > 0 ? (now_s() - t_begin) / warmup : 0;
    std::vector<double> secs;
    long long bytes = 0, blocks = 0;
    for (int s = 0; s < steps; ++s) {
        if (s > 0 && per_step > 0 && (now_s() - t_begin) + per_step > budget_s) break;   // keep the arm inside its time budget
        if (!run_in_child(maps, false, &r)) { fprintf(stderr, "reference reduce task failed\n"); return 1; }
        secs.push_back(r.seconds);
        bytes = r.bytes;
        blocks = r.blocks;
    }
    const char* chk = "skipped";
    if (check) {
        if (!run_in_child(maps, true, &r)) { fprintf(stderr, "reference reduce task failed (check)\n"); return 1; }
        chk = (r.check_ok && r.records == records && r.bytes == g_expected_stream) ? "ok" : "FAILED";
    }
    printf("REF steps=%zu seconds=", secs.size());
    for (size_t i = 0; i < secs.size(); ++i) printf("%s%.6f", i ? "," : "", secs[i]);
    printf(" bytes=%lld blocks=%lld check=%s\n", bytes, blocks, chk);
    return strcmp(chk, "FAILED") == 0 ? 1 : 0;
}
