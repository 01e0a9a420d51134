#include "reader.h"

#include <errno.h>
#include <sys/stat.h>
#include <unistd.h>

#include <map>
#include <mutex>

#include "uda_rt.h"

namespace uda {
namespace {

ssize_t pread_full(int fd, uint8_t* buf, size_t len, int64_t off) {
    size_t got = 0;
    while (got < len) {
        ssize_t r = ::pread(fd, buf + got, len - got, off + (int64_t)got);
        if (r < 0) {
            if (errno == EINTR) continue;
            return -errno;
        }
        if (r == 0) break;  // EOF
        got += (size_t)r;
    }
    return (ssize_t)got;
}

// ------------------------------------------------------------------------------------------ "aio"
class BatchedAlignedReader : public BlockReader {
public:
    explicit BatchedAlignedReader(int workers) : nworkers_(workers < 1 ? 1 : workers) {}
    ~BatchedAlignedReader() override { stop(); }

    void start() override {
        if (running_.exchange(true)) return;
        for (int i = 0; i < nworkers_; ++i) threads_.emplace_back([this] { run(); });
    }
    void insert(ReadRequest r) override {
        if ((r.offset % kAioAlignment) || (r.length % kAioAlignment) || ((uintptr_t)r.buffer % kAioAlignment)) {
            UDA_LOG(lsERROR, "unaligned aio read: offset=%lld length=%zu buffer=%p", (long long)r.offset, r.length, (void*)r.buffer);
            throw UdaError("aio read must be 4 KiB aligned (offset, size and buffer)");
        }
        std::lock_guard<std::mutex> lk(mu_);
        batch_.push_back(std::move(r));
    }
    void submit() override {
        std::vector<ReadRequest> b;
        {
            std::lock_guard<std::mutex> lk(mu_);
            b.swap(batch_);
        }
        for (auto& r : b) {
            ++in_flight_;
            q_.push(std::move(r));
        }
    }
    void stop() override {
        if (!running_.exchange(false)) return;
        q_.close();
        for (auto& t : threads_) t.join();
        threads_.clear();
    }
    int in_flight() const override { return in_flight_.load(); }

private:
    void run() {
        ReadRequest r;
        while (q_.wait_and_pop(r)) {
            ssize_t got = pread_full(r.fd, r.buffer, r.length, r.offset);
            if (got >= 0 && (size_t)got != r.length) {
                // a short read is legal only at the end of the file (within the last two blocks)
                struct stat st;
                bool at_eof = fstat(r.fd, &st) == 0 && r.offset + got >= st.st_size && r.length - (size_t)got < 2 * (size_t)kAioAlignment;
                if (!at_eof) {
                    UDA_LOG(lsERROR, "aio short read: wanted %zu got %zd at offset %lld", r.length, got, (long long)r.offset);
                    got = -EIO;
                }
            }
            --in_flight_;
            if (r.on_done) r.on_done(got);
        }
    }
    int nworkers_;
    std::atomic<bool> running_{false};
    std::atomic<int> in_flight_{0};
    std::mutex mu_;
    std::vector<ReadRequest> batch_;
    ConcurrentQueue<ReadRequest> q_;
    std::vector<std::thread> threads_;
};

// ------------------------------------------------------------------------------------------ "blocked"
class PerDiskReader : public BlockReader {
public:
    PerDiskReader(const std::vector<std::string>& disks, int threads_per_disk)
        : disks_(disks.empty() ? std::vector<std::string>{""} : disks), tpd_(threads_per_disk < 1 ? 1 : threads_per_disk) {
        for (auto& d : disks_) queues_[d] = std::make_unique<ConcurrentQueue<ReadRequest>>();
    }
    ~PerDiskReader() override { stop(); }
    void start() override {
        if (running_.exchange(true)) return;
        for (auto& kv : queues_)
            for (int i = 0; i < tpd_; ++i) threads_.emplace_back([this, q = kv.second.get()] { run(q); });
    }
    void insert(ReadRequest r) override {
        auto it = queues_.find(r.disk);
        if (it == queues_.end()) {
            // longest configured disk prefix of the request's routing key, else the first disk
            size_t best = 0;
            it = queues_.begin();
            for (auto q = queues_.begin(); q != queues_.end(); ++q)
                if (!q->first.empty() && r.disk.compare(0, q->first.size(), q->first) == 0 && q->first.size() >= best) { best = q->first.size(); it = q; }
        }
        ++in_flight_;
        it->second->push(std::move(r));
    }
    void submit() override {}
    void stop() override {
        if (!running_.exchange(false)) return;
        for (auto& kv : queues_) kv.second->close();
        for (auto& t : threads_) t.join();
        threads_.clear();
    }
    int in_flight() const override { return in_flight_.load(); }

private:
    void run(ConcurrentQueue<ReadRequest>* q) {
        ReadRequest r;
        while (q->wait_and_pop(r)) {
            ssize_t got = pread_full(r.fd, r.buffer, r.length, r.offset);
            --in_flight_;
            if (r.on_done) r.on_done(got);
        }
    }
    std::vector<std::string> disks_;
    int tpd_;
    std::map<std::string, std::unique_ptr<ConcurrentQueue<ReadRequest>>> queues_;
    std::atomic<bool> running_{false};
    std::atomic<int> in_flight_{0};
    std::vector<std::thread> threads_;
};

}  // namespace

std::unique_ptr<BlockReader> BlockReader::create(const std::string& kind, const std::vector<std::string>& disks, int tpd) {
    if (kind == "blocked") return std::make_unique<PerDiskReader>(disks, tpd);
    if (kind == "aio") return std::make_unique<BatchedAlignedReader>(tpd < 2 ? 2 : tpd);
    UDA_LOG(lsERROR, "unknown reader kind '%s'", kind.c_str());
    throw UdaError("unknown reader kind");
}

}  // namespace uda
