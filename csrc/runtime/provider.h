// Provider role ("MOFSupplier"): serves partitions of map-output files to reducers.
//
// Reference behaviour covered:
//   * MOFSupplier_main / MOFSupplierRun / mof_downcall_handler  (src/MOFServer/MOFSupplierMain.cc:37-157)
//   * OutputServer: owns the listener and the incoming-request queue (src/MOFServer/MOFServlet.cc:99-180)
//   * DataEngine: chunk pool of 1000 chunks of (buf + 2*4096) bytes, first-fetch index lookup through the
//     getPathUda up-call, aligned chunk reads on fds cached per MOF path with an in-flight refcount,
//     600-character path limit answered with the "MOF_PATH_SIZE_TOO_LONG" sentinel
//     (src/MOFServer/IndexInfo.cc:97-376, src/MOFServer/IndexInfo.h:39-138)
#pragma once
#include <atomic>
#include <condition_variable>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "protocol.h"
#include "queues.h"
#include "reader.h"
#include "transport.h"
#include "uda_rt.h"

namespace uda {

constexpr int kNumChunks = 1000;  // NETLEV_RDMA_MEM_CHUNKS_NUM (src/include/NetlevComm.h:35)

struct Chunk {
    uint8_t* data = nullptr;  // 4 KiB aligned, capacity = buf_size + 2*4096
    size_t capacity = 0;
};

class ChunkPool {
public:
    ChunkPool(int count, size_t buf_size);
    ~ChunkPool();
    Chunk* occupy();            // blocks until a chunk is free; nullptr after shutdown()
    void release(Chunk* c);
    void shutdown();
    int free_count() const;
    size_t total_bytes() const { return total_; }

private:
    uint8_t* region_ = nullptr;  // the single "registered" region all chunks are carved from
    size_t total_ = 0;
    std::vector<Chunk> chunks_;
    std::vector<Chunk*> free_;
    mutable std::mutex mu_;
    std::condition_variable cv_;
    bool down_ = false;
};

// fd cache keyed by MOF path; an fd is closed when its last in-flight read completes
class FdCache {
public:
    explicit FdCache(bool o_direct) : o_direct_(o_direct) {}
    ~FdCache();
    int acquire(const std::string& path);  // -1 on failure
    void release(const std::string& path);
    size_t open_count() const;

private:
    struct Ent { int fd; int refs; };
    bool o_direct_;
    mutable std::mutex mu_;
    std::map<std::string, Ent> map_;
};

class Provider {
public:
    Provider(const Options& opt, const HostCallbacks* host);
    ~Provider();
    void start();                 // MOFSupplier_main: listener up, chunk pool carved
    void run();                   // MOFSupplierRun: request loop until stop (call on its own thread)
    void stop();                  // EXIT_MSG
    void handle_command(const std::string& cmd);  // mof_downcall_handler
    int port() const { return server_ ? server_->port() : 0; }
    uint64_t chunks_served() const { return served_.load(); }

private:
    struct Job { FetchRequest req; std::shared_ptr<Connection> conn; };
    void process(Job j);
    Options opt_;
    const HostCallbacks* host_;
    std::unique_ptr<ShuffleServer> server_;
    std::unique_ptr<ChunkPool> pool_;
    std::unique_ptr<BlockReader> reader_;
    FdCache fds_;
    ConcurrentQueue<Job> requests_;
    std::atomic<bool> stopped_{false};
    std::atomic<uint64_t> served_{0};
    std::thread engine_;
};

}  // namespace uda
