// Reader micro-benchmark: asynchronous block readers vs a plain sequential read over a directory of MOF-like files.
//
// Counterpart of the reference's only native test program, src/tests/AIOHandler_test.cc:56-256 (manual benchmark:
// AIO O_DIRECT vs sequential read, prints "BW: ... MBpS"). Usage:
//     reader_bench <dir> [files=8] [MiB per file=64] [chunk KiB=1024] [threads=4] [--direct]
// Creates the files when they do not exist, reads every one of them through
//   1. one thread, sequential pread of `chunk` bytes           (the reference's "sequential" arm)
//   2. the "aio" back-end (batched, 4 KiB-aligned requests)     (AIOHandler's contract)
//   3. the "blocked" back-end (per-disk queues, N threads)      (src/AsyncIO)
// and prints one "BW: <MB/s> MBpS <arm>" line per arm plus a checksum that must agree between the arms.
// With no arguments it runs a 4 x 4 MiB self-test in a temporary directory and exits 0 when the checksums agree.
#include <fcntl.h>
#include <stdlib.h>
#include <string.h>
#include <sys/stat.h>
#include <unistd.h>

#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdio>
#include <mutex>
#include <string>
#include <vector>

#include "reader.h"
#include "uda_rt.h"

using namespace uda;
using Clock = std::chrono::steady_clock;

static uint64_t mix(uint64_t h, const uint8_t* p, size_t n) {
    for (size_t i = 0; i + 8 <= n; i += 512) {  // sampled: the benchmark measures IO, not hashing
        uint64_t v;
        memcpy(&v, p + i, 8);
        h += v * 0x9E3779B97F4A7C15ull + (h << 7);
    }
    return h;
}

struct Arm {
    const char* name;
    double mbps;
    uint64_t sum;
};

int main(int argc, char** argv) {
    bool direct = false;
    std::vector<std::string> pos;
    for (int i = 1; i < argc; ++i) {
        if (!strcmp(argv[i], "--direct")) direct = true; else pos.push_back(argv[i]);
    }
    char tmpl[] = "/tmp/uda_reader_bench_XXXXXX";
    const bool selftest = pos.empty();
    std::string dir = selftest ? std::string(mkdtemp(tmpl)) : pos[0];
    const int nfiles = pos.size() > 1 ? atoi(pos[1].c_str()) : (selftest ? 4 : 8);
    const size_t file_bytes = (size_t)(pos.size() > 2 ? atoi(pos[2].c_str()) : (selftest ? 4 : 64)) << 20;
    const size_t chunk = (size_t)(pos.size() > 3 ? atoi(pos[3].c_str()) : 1024) << 10;
    const int threads = pos.size() > 4 ? atoi(pos[4].c_str()) : 4;
    if (chunk % kAioAlignment || file_bytes % chunk) { fprintf(stderr, "chunk must be a 4 KiB multiple dividing the file size\n"); return 2; }
    log_set_threshold(lsERROR);

    std::vector<std::string> files;
    for (int f = 0; f < nfiles; ++f) {
        std::string p = dir + "/mof_" + std::to_string(f) + ".out";
        struct stat st;
        if (stat(p.c_str(), &st) != 0 || (size_t)st.st_size != file_bytes) {
            int fd = open(p.c_str(), O_CREAT | O_TRUNC | O_WRONLY, 0644);
            if (fd < 0) { perror(p.c_str()); return 2; }
            std::vector<uint8_t> buf(1 << 20);
            uint64_t x = 88172645463325252ull + (uint64_t)f;
            for (size_t off = 0; off < file_bytes; off += buf.size()) {
                for (size_t i = 0; i < buf.size(); i += 8) { x ^= x << 13; x ^= x >> 7; x ^= x << 17; memcpy(&buf[i], &x, 8); }
                if (write(fd, buf.data(), buf.size()) != (ssize_t)buf.size()) { perror("write"); return 2; }
            }
            fsync(fd);
            close(fd);
        }
        files.push_back(p);
    }
    const int oflags = O_RDONLY | (direct ? O_DIRECT : 0);
    const double total_mb = (double)nfiles * (double)file_bytes / 1e6;
    std::vector<Arm> arms;

    {   // ---- arm 1: sequential
        void* raw = nullptr;
        if (posix_memalign(&raw, kAioAlignment, chunk)) return 2;
        uint8_t* buf = (uint8_t*)raw;
        uint64_t sum = 0;
        auto t0 = Clock::now();
        for (auto& p : files) {
            int fd = open(p.c_str(), oflags);
            if (fd < 0) { perror("open"); return 2; }
            uint64_t h = 0;
            for (size_t off = 0; off < file_bytes; off += chunk) {
                ssize_t r = pread(fd, buf, chunk, (off_t)off);
                if (r != (ssize_t)chunk) { fprintf(stderr, "short read\n"); return 2; }
                h = mix(h, buf, chunk);
            }
            sum ^= h;
            close(fd);
        }
        double s = std::chrono::duration<double>(Clock::now() - t0).count();
        arms.push_back({"sequential", total_mb / s, sum});
        free(raw);
    }
    for (const char* kind : {"aio", "blocked"}) {   // ---- arms 2 and 3
        auto reader = BlockReader::create(kind, {dir}, threads);
        reader->start();
        const int depth = threads * 2;            // buffers in flight per file
        std::mutex mu;
        std::condition_variable cv;
        uint64_t sum = 0;
        auto t0 = Clock::now();
        for (auto& p : files) {
            int fd = open(p.c_str(), oflags);
            if (fd < 0) { perror("open"); return 2; }
            const size_t nchunks = file_bytes / chunk;
            std::vector<uint64_t> hs(nchunks, 0);
            std::vector<uint8_t*> bufs(depth);
            for (auto& b : bufs) { void* raw = nullptr; if (posix_memalign(&raw, kAioAlignment, chunk)) return 2; b = (uint8_t*)raw; }
            std::vector<int> free_bufs;
            for (int i = 0; i < depth; ++i) free_bufs.push_back(i);
            size_t done = 0;
            bool bad = false;
            for (size_t c = 0; c < nchunks; ++c) {
                int bi;
                {
                    std::unique_lock<std::mutex> lk(mu);
                    cv.wait(lk, [&] { return !free_bufs.empty(); });
                    bi = free_bufs.back();
                    free_bufs.pop_back();
                }
                ReadRequest r;
                r.fd = fd; r.offset = (int64_t)(c * chunk); r.length = chunk; r.buffer = bufs[bi]; r.disk = dir;
                r.on_done = [&, c, bi](ssize_t got) {
                    uint64_t h = got == (ssize_t)chunk ? mix(0, bufs[bi], chunk) : 0;
                    std::lock_guard<std::mutex> lk(mu);
                    if (got != (ssize_t)chunk) bad = true;
                    hs[c] = h;
                    free_bufs.push_back(bi);
                    ++done;
                    cv.notify_all();
                };
                reader->insert(std::move(r));
                reader->submit();
            }
            {
                std::unique_lock<std::mutex> lk(mu);
                cv.wait(lk, [&] { return done == nchunks; });
            }
            if (bad) { fprintf(stderr, "%s: read failed\n", kind); return 2; }
            // same per-file hash as the sequential arm: fold the chunk hashes in order
            uint64_t h = 0;   // fold the per-chunk hashes in file order (completion order differs between runs)
            for (size_t c = 0; c < nchunks; ++c) h = h * 1099511628211ull + hs[c];
            sum ^= h;
            for (auto b : bufs) free(b);
            close(fd);
        }
        double s = std::chrono::duration<double>(Clock::now() - t0).count();
        reader->stop();
        arms.push_back({kind, total_mb / s, sum});
    }
    for (auto& a : arms) printf("BW: %.1f MBpS %s (checksum %016llx)\n", a.mbps, a.name, (unsigned long long)a.sum);
    const bool ok = arms[1].sum == arms[2].sum;   // both asynchronous arms saw the same bytes in every chunk
    printf("%s\n", ok ? "reader bench: checksums agree" : "reader bench: CHECKSUM MISMATCH");
    if (selftest) {
        for (auto& p : files) unlink(p.c_str());
        rmdir(dir.c_str());
    }
    return ok ? 0 : 1;
}
