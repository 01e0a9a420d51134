#include "provider.h"

#include <fcntl.h>
#include <sys/resource.h>
#include <unistd.h>

#include <cstdlib>
#include <cstring>

namespace uda {

// ------------------------------------------------------------------------------------------ chunk pool
ChunkPool::ChunkPool(int count, size_t buf_size) {
    const size_t cap = buf_size + 2 * (size_t)kAioAlignment;
    total_ = cap * (size_t)count;
    void* p = nullptr;
    if (posix_memalign(&p, kAioAlignment, total_) != 0) throw UdaError("cannot allocate the provider chunk region");
    region_ = (uint8_t*)p;
    chunks_.resize(count);
    for (int i = 0; i < count; ++i) {
        chunks_[i].data = region_ + cap * (size_t)i;
        chunks_[i].capacity = cap;
        free_.push_back(&chunks_[i]);
    }
}
ChunkPool::~ChunkPool() { free(region_); }
Chunk* ChunkPool::occupy() {
    std::unique_lock<std::mutex> lk(mu_);
    cv_.wait(lk, [&] { return !free_.empty() || down_; });  // `while`, not the reference's `if` (IndexInfo.cc:282-284)
    if (free_.empty()) return nullptr;
    Chunk* c = free_.back();
    free_.pop_back();
    return c;
}
void ChunkPool::release(Chunk* c) {
    {
        std::lock_guard<std::mutex> lk(mu_);
        free_.push_back(c);
    }
    cv_.notify_one();
}
void ChunkPool::shutdown() {
    {
        std::lock_guard<std::mutex> lk(mu_);
        down_ = true;
    }
    cv_.notify_all();
}
int ChunkPool::free_count() const {
    std::lock_guard<std::mutex> lk(mu_);
    return (int)free_.size();
}

// ------------------------------------------------------------------------------------------ fd cache
FdCache::~FdCache() {
    for (auto& kv : map_) ::close(kv.second.fd);
}
int FdCache::acquire(const std::string& path) {
    std::lock_guard<std::mutex> lk(mu_);
    auto it = map_.find(path);
    if (it != map_.end()) { ++it->second.refs; return it->second.fd; }
    int fd = -1;
    if (o_direct_) fd = ::open(path.c_str(), O_RDONLY | O_DIRECT);
    if (fd < 0) fd = ::open(path.c_str(), O_RDONLY);  // tmpfs and friends refuse O_DIRECT
    if (fd < 0) {
        UDA_LOG(lsERROR, "cannot open MOF %s: %s", path.c_str(), strerror(errno));
        return -1;
    }
    map_[path] = Ent{fd, 1};
    return fd;
}
void FdCache::release(const std::string& path) {
    std::lock_guard<std::mutex> lk(mu_);
    auto it = map_.find(path);
    if (it == map_.end()) return;
    if (--it->second.refs == 0) {
        ::close(it->second.fd);
        map_.erase(it);
    }
}
size_t FdCache::open_count() const {
    std::lock_guard<std::mutex> lk(mu_);
    return map_.size();
}

// ------------------------------------------------------------------------------------------ provider
Provider::Provider(const Options& opt, const HostCallbacks* host) : opt_(opt), host_(host), fds_(true) {
    if (opt_.buf_size <= 0) opt_.buf_size = 1024 * 1024;
}
Provider::~Provider() { stop(); }

void Provider::start() {
    log_open_mofsupplier(opt_.log_dir);
    /* the next line is parsed by deployment automation: keep its wording */
    UDA_LOG(lsINFO, "The version is %s", VERSION_UDA);
    struct rlimit rl;
    if (getrlimit(RLIMIT_NOFILE, &rl) == 0) UDA_LOG(lsINFO, "RLIMIT_NOFILE: soft=%ld hard=%ld", (long)rl.rlim_cur, (long)rl.rlim_max);
    server_ = std::make_unique<ShuffleServer>(opt_.data_port, [this](const FetchRequest& r, std::shared_ptr<Connection> c) {
        requests_.push(Job{r, std::move(c)});
    });
    server_->start();
    pool_ = std::make_unique<ChunkPool>(kNumChunks, (size_t)opt_.buf_size);
    int tpd = atoi(conf_get(host_, "mapred.uda.provider.blocked.threads.per.disk", "1").c_str());
    std::string kind = conf_get(host_, "mapred.uda.provider.reader", "aio");
    std::vector<std::string> disks;
    {
        std::string dirs = conf_get(host_, "mapred.local.dir", "");
        size_t pos = 0;
        while (pos <= dirs.size() && !dirs.empty()) {
            size_t c = dirs.find(',', pos);
            if (c == std::string::npos) c = dirs.size();
            if (c > pos) disks.push_back(dirs.substr(pos, c - pos));
            pos = c + 1;
        }
    }
    reader_ = BlockReader::create(kind, disks, tpd);
    reader_->start();
    UDA_LOG(lsINFO, "MOFSupplier ready: port=%d chunks=%d chunk_bytes=%zu reader=%s", server_->port(), kNumChunks,
            (size_t)opt_.buf_size + 2 * kAioAlignment, kind.c_str());
}

void Provider::run() {
    Job j;
    while (!stopped_ && requests_.wait_and_pop(j)) {
        try {
            process(std::move(j));
        } catch (const UdaError& e) {
            UDA_LOG(lsERROR, "request failed: %s", e.info().c_str());
        }
    }
    UDA_LOG(lsINFO, "MOFSupplier engine loop finished");
}

void Provider::process(Job j) {
    FetchRequest& rq = j.req;
    IndexRecord rec;
    // A request that cannot be served is answered with a rejecting ack (path = kFetchRejected, nothing sent): the
    // reducer fails that fetch at once instead of waiting for a reply that never comes.
    auto reject = [&](const char* why) {
        UDA_LOG(lsERROR, "fetch rejected (%s): job=%s map=%s reduce=%d offset=%lld chunk=%d", why, rq.job_id.c_str(), rq.map_id.c_str(),
                rq.reduce_id, (long long)rq.map_offset, rq.chunk_size);
        FetchAck nack;
        nack.path = kFetchRejected;
        nack.sent = 0;
        ShuffleServer::reply(j.conn, rq.request_ptr, format_fetch_ack(nack), nullptr, 0);
    };
    // the request text comes off a socket: nothing in it is trusted
    if (rq.chunk_size <= 0 || rq.map_offset < 0) { reject("non-positive chunk size or negative offset"); return; }
    if (rq.path.empty()) {
        // first fetch of this (job, map, reduce): ask the host for the index record
        if (!host_ || !host_->get_path_uda || !host_->get_path_uda(host_->user, rq.job_id.c_str(), rq.map_id.c_str(), rq.reduce_id, &rec)) {
            reject("getPathUda returned nothing");
            return;
        }
    } else {
        rec.offset = rq.file_offset;
        rec.raw_length = rq.raw_length;
        rec.part_length = rq.part_length;
        rec.path = rq.path;
    }
    if (rec.offset < 0 || rec.part_length < 0 || rec.raw_length < 0 || rq.map_offset > rec.part_length) {
        reject("offset beyond the partition or negative index record");
        return;
    }
    if (pool_->free_count() == 0) reader_->submit();  // do not sleep on the pool with reads still un-submitted
    Chunk* chunk = pool_->occupy();
    if (!chunk) return;   // shutting down
    FetchAck ack;
    ack.raw_length = rec.raw_length;
    ack.part_length = rec.part_length;
    ack.file_offset = rec.offset;
    ack.path = rec.path;
    if (rec.path.size() > kMaxMofPath) {
        ack.path = kPathTooLong;
        ack.sent = 0;
        ShuffleServer::reply(j.conn, rq.request_ptr, format_fetch_ack(ack), nullptr, 0);
        pool_->release(chunk);
        return;
    }
    long long want = rec.part_length - rq.map_offset;     // >= 0, checked above
    if (want > rq.chunk_size) want = rq.chunk_size;
    if (want > opt_.buf_size) want = opt_.buf_size;
    if (want == 0) {
        ack.sent = 0;
        ShuffleServer::reply(j.conn, rq.request_ptr, format_fetch_ack(ack), nullptr, 0);
        pool_->release(chunk);
        return;
    }
    const long long file_pos = rec.offset + rq.map_offset;
    const long long aligned = file_pos & ~((long long)kAioAlignment - 1);
    const size_t lead = (size_t)(file_pos - aligned);
    const size_t len = (lead + (size_t)want + kAioAlignment - 1) & ~((size_t)kAioAlignment - 1);
    if (len > chunk->capacity) {   // cannot happen with want <= buf_size; the read must never leave the chunk
        pool_->release(chunk);
        reject("read span exceeds the chunk");
        return;
    }
    int fd = fds_.acquire(rec.path);
    if (fd < 0) {
        pool_->release(chunk);
        reject("cannot open the map output file");
        return;
    }
    ReadRequest rr;
    rr.fd = fd;
    rr.offset = aligned;
    rr.length = len;
    rr.buffer = chunk->data;
    rr.disk = rec.path;
    auto conn = j.conn;
    const uint64_t cookie = rq.request_ptr;
    const std::string path = rec.path;
    rr.on_done = [this, chunk, lead, want, ack, conn, cookie, path](ssize_t got) mutable {
        if (got < 0 || (size_t)got < lead + (size_t)want) {
            UDA_LOG(lsERROR, "read of %s failed or short: rc=%zd", path.c_str(), got);
            FetchAck nack;
            nack.path = kFetchRejected;
            nack.sent = 0;
            ShuffleServer::reply(conn, cookie, format_fetch_ack(nack), nullptr, 0);
        } else {
            ack.sent = want;
            if (ShuffleServer::reply(conn, cookie, format_fetch_ack(ack), chunk->data + lead, (size_t)want)) ++served_;
        }
        fds_.release(path);
        pool_->release(chunk);
    };
    reader_->insert(std::move(rr));
    reader_->submit();
}

void Provider::handle_command(const std::string& s) {
    HadoopCmd cmd;
    if (!parse_hadoop_cmd(s, &cmd)) { UDA_LOG(lsWARN, "Hadoop's command '%s' cannot be parsed", s.c_str()); return; }
    switch (cmd.header) {
        case INIT_MSG: break;  // nothing to do on the provider
        case EXIT_MSG: stop(); break;
        default: UDA_LOG(lsWARN, "unexpected command %d on the provider", cmd.header); break;
    }
}

void Provider::stop() {
    if (stopped_.exchange(true)) return;
    requests_.close();
    if (pool_) pool_->shutdown();
    if (server_) server_->stop();
    if (reader_) reader_->stop();
    UDA_LOG(lsINFO, "MOFSupplier stopped");
    log_close();
}

}  // namespace uda
