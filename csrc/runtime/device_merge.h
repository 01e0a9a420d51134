// Device merge back-end of the native consumer: merge approach 3.
//
// The reference's consumer merges on one CPU thread per reducer (src/Merger/MergeManager.cc:47-182 online,
// :202-288 hybrid; selected with `-a` / mapred.netmerger.merge.approach). Approach 3 has no counterpart there: the
// consumer fetches every map output of its partition into host memory (same fetch protocol, same buffers), hops
// over the record headers once to build the record-offset tables, and hands the segments to the sm_100a kernels of
// libuda_kernels.so. Records of one fixed shape with a compared key of at most 10 bytes (TeraSort, fixed-width keys) go
// through the fused record merge — the kernel behind the headline benchmark, one read and one write per record; every
// other shape takes key normalise -> k-way merge of entries -> exact-comparator tie-break -> record gather. The
// merged stream comes back to the host and is delivered through dataFromUda in blocks of whole records, exactly
// like the CPU merges deliver it.
//
// libuda.so must keep loading on machines without a GPU (the JVM loads it before any decision is taken), so the
// kernel library is dlopen'ed on first use; when it is missing, when no device is visible or when the partition does
// not fit the HBM budget, `available()` / `merge()` say so and the consumer falls back to its CPU heap merge over
// the segments it already holds.
#pragma once
#include <atomic>
#include <cstdint>
#include <functional>
#include <mutex>
#include <string>
#include <vector>

#include "compare.h"

namespace uda {

struct HostSegment {
    std::vector<uint8_t> bytes;        // IFile records + EOF marker (trailing checksum bytes are ignored)
    std::vector<long long> rec_off;    // offsets of the records, then of the EOF marker (records + 1 entries)
    long long records() const { return rec_off.empty() ? 0 : (long long)rec_off.size() - 1; }
    long long raw_len() const { return rec_off.empty() ? 0 : rec_off.back() + 2; }
};

// Walks the record headers of `seg->bytes` and fills `seg->rec_off`; throws UdaError on a corrupt segment.
void index_segment(HostSegment* seg);

class DeviceMerge {
public:
    // Loads the kernel library (UDA_KERNELS_LIB, else libuda_kernels.so next to libuda.so). Never throws.
    static DeviceMerge& instance();
    bool available() const { return ok_; }
    // Creates the CUDA context now (call it from a helper thread while the fetches are still running)
    void warm_up();
    const std::string& why_not() const { return why_; }
    // merges of this process that took the fused fixed-stride record merge
    static long long fused_merges();

    // Merges the segments (each sorted by `kind`'s comparator) and calls `sink` with consecutive blocks of the merged
    // stream, every block <= max_block bytes and made of whole records (a single larger record travels alone); the
    // last block ends with the EOF marker. Returns false (nothing delivered) when the device cannot take the job:
    // *why says so and the caller merges on the CPU. Throws UdaError on CUDA errors after work has started.
    // *refusal (optional) tells the caller what kind of "no" it was: a partition that only lacks device memory can be
    // merged in groups (Consumer::merge_device_grouped), the other refusals go to the CPU merge.
    enum class Refusal { NONE, UNAVAILABLE, LIMITS, MEMORY, FAILED };
    bool merge(const std::vector<HostSegment>& segs, KeyKind kind, size_t max_block,
               const std::function<void(const uint8_t*, size_t)>& sink, long long* records_out, std::string* why,
               Refusal* refusal = nullptr);

private:
    DeviceMerge();
    bool merge_fixed(const std::vector<HostSegment>& segs, int stride, int key_off, int key_len, long long n, size_t max_block,
                     const std::function<void(const uint8_t*, size_t)>& sink, std::string* why, bool* out_of_memory);
    void stream_out(const uint8_t* d_out, const std::vector<unsigned long long>& off, size_t n, size_t max_block,
                    const std::function<void(const uint8_t*, size_t)>& deliver);
    bool ok_ = false;
    std::string why_;
    void* lib_ = nullptr;
    struct Api;
    Api* api_ = nullptr;
    static constexpr size_t kStageBytes = 32u << 20;   // pinned D2H staging, two buffers, allocated on first use
    std::mutex stage_mu_;                               // one merge at a time streams through the staging buffers
    uint8_t* stage_[2] = {nullptr, nullptr};
};

}  // namespace uda
