// Asynchronous block readers for MOF files on local disks.
//
// Reference behaviour covered:
//   * AIOHandler (src/CommUtils/AIOHandler.cc:27-235, src/include/AIOHandler.h:26-105): requests are
//     prepared into a batch, submit() hands the batch to the kernel, a completion thread invokes the
//     callback per request; offset, size and buffer must be 4 KiB aligned; a short read is tolerated
//     only when it ends inside the last two blocks of the file (EOF).
//   * AsyncIO "blocked" reader (src/AsyncIO/AsyncReaderManager.cc:16-152, AsyncReaderThread.cc:36-81):
//     one queue and N threads per local disk, blocking pread, optional prepareRead hook on the
//     subscriber, readCallback on completion. In the reference it is compiled in but never wired to the
//     DataEngine; here either back-end can serve the MofStore.
// There is no libaio on this image, so the "aio" back-end implements the same contract (batching,
// alignment rules, completion thread) over pread().
#pragma once
#include <atomic>
#include <cstdint>
#include <functional>
#include <memory>
#include <string>
#include <thread>
#include <vector>

#include "queues.h"

namespace uda {

constexpr int kAioAlignment = 4096;  // AIO_ALIGNMENT (src/include/AIOHandler.h:26-27)

struct ReadRequest {
    int fd = -1;
    int64_t offset = 0;
    size_t length = 0;
    uint8_t* buffer = nullptr;
    std::string disk;                       // routing key for the per-disk back-end
    std::function<void(ssize_t)> on_done;   // bytes read, or -errno
};

class BlockReader {
public:
    virtual ~BlockReader() = default;
    virtual void start() = 0;
    // queue one request (AIOHandler::prepare_read / AbstractReader::insert)
    virtual void insert(ReadRequest req) = 0;
    // make queued requests eligible to run (AIOHandler::submit); a no-op for back-ends that start at insert
    virtual void submit() = 0;
    virtual void stop() = 0;
    virtual int in_flight() const = 0;
    // "aio": batched + alignment checked, one completion thread pool; "blocked": per-disk queues
    static std::unique_ptr<BlockReader> create(const std::string& kind, const std::vector<std::string>& disks,
                                               int threads_per_disk);
};

}  // namespace uda
